"""Dynamic batching of synthesis requests that arrive from several threads.

The reference serves strictly one request at a time per GPU process (`web_api/socket_server.py:125-131`,
`vits_wrap.py:177`: the segments of a request are synthesised one after the other).  `DynamicBatcher` is the library-side
replacement for that loop: worker threads (socket handlers, segment loops) call `submit(z_p, g)` and get a future; one
dispatcher thread groups whatever is queued - up to `max_batch` rows / `max_batch_frames` padded frames, waiting at most
`max_wait_ms` for company - into ONE ragged `FlowDecoder.synthesize` call and hands every caller its own rows.  Row
independence of the ragged plan makes each result bit-identical to the B = 1 call on that request.
"""
from __future__ import annotations

import queue
import threading
import time
from concurrent.futures import Future
from typing import List, Optional, Tuple

import torch

from .engine import FlowDecoder


class DynamicBatcher:
    def __init__(self, eng: FlowDecoder, max_batch: int = 64, max_batch_frames: int = 32768, max_wait_ms: float = 2.0,
                 pcm16: bool = False, volume: float = 1.0):
        self.eng, self.max_batch, self.max_frames, self.max_wait = eng, max_batch, max_batch_frames, max_wait_ms * 1e-3
        self.pcm16, self.volume = pcm16, volume
        self._q: "queue.Queue[Optional[Tuple[torch.Tensor, torch.Tensor, Future]]]" = queue.Queue()
        self.batches: List[int] = []                 # rows per dispatched batch (for tests / monitoring)
        self._thread = threading.Thread(target=self._run, name="vitsb-batcher", daemon=True)
        self._thread.start()

    def submit(self, z_p: torch.Tensor, g: torch.Tensor) -> Future:
        """z_p [1, inter, T] or [inter, T], g [1, gin] or [gin] (any device); the future resolves to the waveform
        [1, 1, T * prod(upsample_rates)] (float32, or int16 PCM with pcm16=True) on the engine's device."""
        z = z_p if z_p.dim() == 3 else z_p.unsqueeze(0)
        gg = g if g.dim() == 2 else g.reshape(1, -1)
        if z.size(0) != 1 or z.size(1) != self.eng.inter_channels or tuple(gg.shape) != (1, self.eng.gin_channels):
            raise ValueError(f"expected z_p [1,{self.eng.inter_channels},T] and g [1,{self.eng.gin_channels}]")
        fut: Future = Future()
        self._q.put((z, gg, fut))
        return fut

    def close(self):
        self._q.put(None)
        self._thread.join()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- dispatcher ------------------------------------------------------------
    def _run(self):
        stop = False
        with torch.cuda.device(self.eng.device):
            stream = torch.cuda.Stream(self.eng.device)
            while not stop:
                item = self._q.get()
                if item is None:
                    break
                batch = [item]
                t_max = item[0].size(2)
                deadline = time.monotonic() + self.max_wait
                while len(batch) < self.max_batch:
                    try:
                        nxt = self._q.get(timeout=max(0.0, deadline - time.monotonic()))
                    except queue.Empty:
                        break
                    if nxt is None:
                        stop = True
                        break
                    if max(t_max, nxt[0].size(2)) * (len(batch) + 1) > self.max_frames:
                        self._q.put(nxt)                    # does not fit this batch: it opens the next one
                        break
                    batch.append(nxt)
                    t_max = max(t_max, nxt[0].size(2))
                self._dispatch(batch, t_max, stream)

    def _dispatch(self, batch, t_max, stream):
        eng, n = self.eng, len(batch)
        try:
            with torch.cuda.stream(stream), torch.no_grad():
                zb = torch.zeros(n, eng.inter_channels, t_max, dtype=torch.float32, device=eng.device)
                gb = torch.empty(n, eng.gin_channels, dtype=torch.float32, device=eng.device)
                lens = []
                for r, (z, g, _) in enumerate(batch):
                    zb[r, :, :z.size(2)] = z[0].to(eng.device, torch.float32)
                    gb[r] = g[0].to(eng.device, torch.float32)
                    lens.append(int(z.size(2)))
                out = eng.synthesize_pcm16(zb, gb, lens, volume=self.volume) if self.pcm16 else eng.synthesize(zb, gb, lens)
                res = [out[r:r + 1, :, :lens[r] * eng.total_upsample].clone() for r in range(n)]
                stream.synchronize()
            self.batches.append(n)
            for (_, _, fut), w in zip(batch, res):
                fut.set_result(w)
        except Exception as e:       # noqa: BLE001 - every waiter must be released
            for _, _, fut in batch:
                if not fut.done():
                    fut.set_exception(e)
