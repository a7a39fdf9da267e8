"""Host-side mirror of the reference's flow+decoder interface over the C ABI.

`FlowDecoder` owns one `vitsb_handle` (packed weights on one GPU).  Method
names and argument meaning follow the reference (paths relative to
emotional-vits/):

  flow_reverse(z_p, g, lengths)   ResidualCouplingBlock.infer / masked forward(reverse=True)  models.py:219-235
  decode(z, g, lengths, ragged)   Generator.forward                                          models.py:306-318
  infer_p2(attn, m_p, s_p, g, noise)  SynthesizerTrn.infer_p2                                models.py:568-575
  synthesize(z_p, g, lengths, ragged) tail of infer_p2 / inference                           models.py:533-534,573-574
  synthesize_segments(segments)   the per-segment loop of VITSWrap.speaking as ragged batches  vits_wrap.py:177-183

PyTorch is plumbing here (device memory, streams); every FLOP of the path runs
in libvitsb200.so.  There is no CPU fallback: constructing a FlowDecoder
without a CUDA device or without the built library raises.
"""
from __future__ import annotations

import ctypes as C
import numpy as np
from typing import Dict, Optional, Sequence

import torch

from . import lib as L

_TORCH_DT = {torch.float32: L.DT_F32, torch.float16: L.DT_F16, torch.bfloat16: L.DT_BF16}


def _hot_items(state_dict):
    for k, v in state_dict.items():
        if k.startswith(("dec.", "flow.", "enc_p.", "dp.", "emb_g.")):
            yield k, v


class FlowDecoder:
    def __init__(self, cfg, state_dict: Dict[str, torch.Tensor], device="cuda:0", precision: str = "fp16",
                 options: Optional[Dict[str, int]] = None):
        """`options`: vitsb_set_option keys (include/vitsb200.h), applied before the weights are packed."""
        if not torch.cuda.is_available():
            raise RuntimeError("vits_b200.FlowDecoder needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = L.load_library()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("vits_b200.FlowDecoder runs on CUDA devices only")
        self.precision = precision
        self._cfg_struct = L.make_config(cfg, precision)
        self.inter_channels = self._cfg_struct.inter_channels
        self.gin_channels = self._cfg_struct.gin_channels
        h = C.c_void_p()
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        rc = self.lib.vitsb_create(C.byref(self._cfg_struct), idx, C.byref(h))
        if rc != 0:
            raise L.VitsbError(rc, self.lib.vitsb_last_error(None).decode())
        self._h = h
        self._ws: Optional[torch.Tensor] = None
        self._ws1: Optional[torch.Tensor] = None
        for k, v in (options or {}).items():
            self.set_option(k, v)
        self.load_state_dict(state_dict)
        self.total_upsample = self.lib.vitsb_total_upsample(self._h)
        self.flop_per_frame = self.lib.vitsb_flop_per_frame(self._h)

    # -- lifecycle -----------------------------------------------------------
    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and getattr(self, "lib", None) is not None:
            self.lib.vitsb_destroy(h)
            self._h = None

    def _check(self, rc: int):
        if rc != 0:
            raise L.VitsbError(rc, self.lib.vitsb_last_error(self._h).decode())

    def load_state_dict(self, state_dict: Dict[str, torch.Tensor]):
        """Accepts the reference checkpoint's `dec.*` / `flow.*` entries, folded
        or weight-normed (utils.py:54-57, export.py:150-153); other keys are ignored."""
        keep, descs = [], []
        for k, v in _hot_items(state_dict):
            t = v.detach().to("cpu")
            if t.dtype not in _TORCH_DT:
                t = t.float()
            t = t.contiguous()
            keep.append((k.encode(), t))
        arr = (L.VitsbTensorDesc * len(keep))()
        for i, (name, t) in enumerate(keep):
            arr[i].name = name
            arr[i].data = t.data_ptr()
            arr[i].dtype = _TORCH_DT[t.dtype]
            arr[i].ndim = t.dim()
            for d in range(t.dim()):
                arr[i].shape[d] = t.shape[d]
        self._check(self.lib.vitsb_load_weights(self._h, arr, len(keep)))

    def set_option(self, key: str, value: int):
        self._check(self.lib.vitsb_set_option(self._h, key.encode(), int(value)))

    def get_option(self, key: str) -> int:
        v = C.c_int()
        if self.lib.vitsb_get_option(self._h, key.encode(), C.byref(v)) != 0:
            raise KeyError(key)
        return int(v.value)

    # -- helpers -------------------------------------------------------------
    def workspace_bytes(self, B: int, T: int) -> int:
        return int(self.lib.vitsb_workspace_bytes(self._h, B, T))

    def launches_per_call(self, B: int, T: int) -> int:
        return int(self.lib.vitsb_launches_per_call(self._h, B, T))

    TAG_NAMES = {0: "pack", 1: "cond_gemv", 2: "unpack", 3: "conv_post", 4: "mrf_sum", 10: "flow.pre", 11: "flow.in_layers",
                 12: "flow.res", 13: "flow.skip", 14: "flow.post", 15: "flow.fused", 20: "conv_pre"}

    def profile(self, enable: bool):
        self._check(self.lib.vitsb_profile(self._h, int(enable)))

    def profile_table(self):
        """[(name, ms, launches, algorithmic flop)] accumulated since profile(True)."""
        cap = 64
        tags, ms = (C.c_int32 * cap)(), (C.c_double * cap)()
        n_l, fl = (C.c_int64 * cap)(), (C.c_double * cap)()
        n = self.lib.vitsb_profile_read(self._h, tags, ms, n_l, fl, cap)
        rows = []
        for i in range(n):
            t = tags[i]
            name = self.TAG_NAMES.get(t) or (f"ups.{t - 30}" if 30 <= t < 40 else
                                             f"resblock.c1.s{t - 40}" if 40 <= t < 50 else
                                             f"resblock.c2.s{t - 50}" if 50 <= t < 60 else f"resblock.fused.s{t - 60}")
            rows.append((name, ms[i], int(n_l[i]), fl[i]))
        return rows

    def _workspace(self, B: int, T: int) -> torch.Tensor:
        need = self.workspace_bytes(B, T)
        if self._ws is None or self._ws.numel() < need:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._ws

    def _prep(self, x: torch.Tensor, g: torch.Tensor, lengths):
        if x.dim() != 3 or x.size(1) != self.inter_channels:
            raise ValueError(f"expected [B,{self.inter_channels},T], got {tuple(x.shape)}")
        B, _, T = x.shape
        if g.dim() == 3:            # tolerate [B,gin,1]
            g = g.squeeze(-1)
        if tuple(g.shape) != (B, self.gin_channels):
            raise ValueError(f"expected g [B,{self.gin_channels}], got {tuple(g.shape)}")
        x = x.to(self.device, torch.float32).contiguous()
        g = g.to(self.device, torch.float32).contiguous()
        lens = None
        if lengths is not None:
            lst = [int(v) for v in (lengths.tolist() if torch.is_tensor(lengths) else lengths)]
            if len(lst) != B:
                raise ValueError("lengths must have B entries")
            lens = (C.c_int32 * B)(*lst)
        return x, g, lens, B, T

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # -- reference-shaped API --------------------------------------------------
    @torch.no_grad()
    def flow_reverse(self, z_p: torch.Tensor, g: torch.Tensor, lengths: Optional[Sequence[int]] = None) -> torch.Tensor:
        z_p, g, lens, B, T = self._prep(z_p, g, lengths)
        with torch.cuda.device(self.device):
            ws = self._workspace(B, T)
            z = torch.empty_like(z_p)
            self._check(self.lib.vitsb_flow_reverse(self._h, z_p.data_ptr(), g.data_ptr(), lens, B, T, z.data_ptr(),
                                                    ws.data_ptr(), ws.numel(), self._stream()))
        return z

    @torch.no_grad()
    def resample(self, wav: torch.Tensor, orig_sr: int, target_sr: int, pcm16: bool = False, volume: float = 1.0) -> torch.Tensor:
        """VITSWrap.speaking's resampling step on the device (vits_wrap.py:186-194): `wav` [B, 1, N] or [B, N] fp32 at
        `orig_sr` -> the same shape at `target_sr` (ceil(N * target / orig) samples), optionally followed by the int16
        conversion of vits_wrap.py:195.  `pitch` is resample(wav, int(sr / pitch), sr); a requested output rate is
        resample(wav, sr, sampling_rate).  Method: scipy.signal.resample_poly's polyphase filter (vits_b200/resample.py)."""
        from . import resample as R
        key = (int(orig_sr), int(target_sr))
        cache = self.__dict__.setdefault("_resample_plans", {})
        if key not in cache:
            p = R.plan(*key)
            cache[key] = (p, torch.from_numpy(np.ascontiguousarray(p.taps)).to(self.device))
        p, taps = cache[key]
        x = wav.to(self.device, torch.float32).contiguous()
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        B, n_in = x2.shape
        n_out = p.n_out(n_in)
        with torch.cuda.device(self.device):
            out = torch.empty(B, n_out, dtype=torch.int16 if pcm16 else torch.float32, device=self.device)
            if p.up == 1 and p.down == 1 and not pcm16:
                out.copy_(x2)
            else:
                self._check(self.lib.vitsb_resample_poly(self._h, x2.data_ptr(), B, n_in, n_in, p.up, p.down, taps.data_ptr(),
                                                         taps.numel(), p.n_pre_remove, n_out,
                                                         None if pcm16 else out.data_ptr(), out.data_ptr() if pcm16 else None,
                                                         float(volume), n_out, self._stream()))
        return out.reshape(*shape[:-1], n_out)

    @torch.no_grad()
    def speaking_output(self, wav: torch.Tensor, volume: float = 1.0, pitch: float = 1.0, sampling_rate: Optional[int] = None,
                        tail_silence: float = 0.0, default_sampling_rate: int = 16000) -> torch.Tensor:
        """The per-segment output steps of VITSWrap.speaking (vits_wrap.py:186-197) on the device, for one waveform
        [N] / [1, N] / [1, 1, N] fp32 at `default_sampling_rate`: `pitch` resampling from int(sr / pitch) to sr, resampling
        to the requested `sampling_rate`, int16 conversion with `volume`, zero padding of int(tail_silence * sampling_rate)
        samples.  Returns int16 [N_out] on the device.  (The optional bandwidth-extension model `self.asv` of line 189 is
        a separate network and is not part of this library; the resampler is the polyphase one of `resample()`.)"""
        sr = int(default_sampling_rate)
        out_sr = sr if sampling_rate is None else int(sampling_rate)
        x = wav.to(self.device, torch.float32).reshape(1, -1)
        steps = []
        if pitch != 1.0:
            steps.append((int(sr / pitch), sr))
        if out_sr != sr:
            steps.append((sr, out_sr))
        for i, (a, b) in enumerate(steps[:-1]):
            x = self.resample(x, a, b)
        if steps:
            pcm = self.resample(x, steps[-1][0], steps[-1][1], pcm16=True, volume=volume)
        else:                                              # no resampling: the conversion alone (vits_wrap.py:195)
            pcm = self.resample(x, sr, sr, pcm16=True, volume=volume)
        pcm = pcm.reshape(-1)
        pad = int(tail_silence * out_sr) if tail_silence > 0 else 0
        return torch.nn.functional.pad(pcm, (0, pad)) if pad else pcm

    @torch.no_grad()
    def decode(self, z: torch.Tensor, g: torch.Tensor, lengths: Optional[Sequence[int]] = None,
               ragged: bool = True) -> torch.Tensor:
        z, g, lens, B, T = self._prep(z, g, lengths)
        with torch.cuda.device(self.device):
            ws = self._workspace(B, T)
            wav = torch.empty(B, 1, T * self.total_upsample, dtype=torch.float32, device=self.device)
            self._check(self.lib.vitsb_decode(self._h, z.data_ptr(), g.data_ptr(), lens, B, T, int(bool(ragged)),
                                              wav.data_ptr(), ws.data_ptr(), ws.numel(), self._stream()))
        return wav

    @torch.no_grad()
    def synthesize(self, z_p: torch.Tensor, g: torch.Tensor, lengths: Optional[Sequence[int]] = None,
                   ragged: bool = True, return_z: bool = False, out: Optional[torch.Tensor] = None):
        z_p, g, lens, B, T = self._prep(z_p, g, lengths)
        with torch.cuda.device(self.device):
            ws = self._workspace(B, T)
            wav = out if out is not None else torch.empty(B, 1, T * self.total_upsample, dtype=torch.float32,
                                                          device=self.device)
            z = torch.empty_like(z_p) if return_z else None
            self._check(self.lib.vitsb_synthesize(self._h, z_p.data_ptr(), g.data_ptr(), lens, B, T, int(bool(ragged)),
                                                  z.data_ptr() if return_z else None, wav.data_ptr(),
                                                  ws.data_ptr(), ws.numel(), self._stream()))
        return (wav, z) if return_z else wav

    @torch.no_grad()
    def synthesize_pcm16(self, z_p: torch.Tensor, g: torch.Tensor, lengths: Optional[Sequence[int]] = None,
                         ragged: bool = True, volume: float = 1.0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """synthesize() + VITSWrap.speaking's output step (vits_wrap.py:195) in the last kernel: int16 [B,1,192*T] =
        clip(wav * volume * 32767, -32768, 32767) truncated toward zero; half the D2H bytes."""
        z_p, g, lens, B, T = self._prep(z_p, g, lengths)
        with torch.cuda.device(self.device):
            ws = self._workspace(B, T)
            pcm = out if out is not None else torch.empty(B, 1, T * self.total_upsample, dtype=torch.int16, device=self.device)
            self._check(self.lib.vitsb_synthesize_pcm16(self._h, z_p.data_ptr(), g.data_ptr(), lens, B, T, int(bool(ragged)),
                                                        float(volume), pcm.data_ptr(), ws.data_ptr(), ws.numel(), self._stream()))
        return pcm

    @torch.no_grad()
    def length_regulate(self, attn, m_p, s_p, noise) -> torch.Tensor:
        """Head of SynthesizerTrn.infer_p2 (models.py:570-572): z_p = attn @ m_p^T + noise * (attn @ s_p^T) as one
        gather/FMA kernel.  attn [B,Ty,Tx], m_p/s_p [B,inter,Tx], noise [B,inter,Ty] -> z_p [B,inter,Ty]."""
        dev = self.device
        attn, m_p, s_p, noise = (t.to(dev, torch.float32).contiguous() for t in (attn, m_p, s_p, noise))
        if attn.dim() != 3 or m_p.dim() != 3 or tuple(s_p.shape) != tuple(m_p.shape):
            raise ValueError("expected attn [B,Ty,Tx] and m_p, s_p [B,inter,Tx]")
        B, Ty, Tx = attn.shape
        if tuple(m_p.shape) != (B, self.inter_channels, Tx) or tuple(noise.shape) != (B, self.inter_channels, Ty):
            raise ValueError(f"shape mismatch: attn {tuple(attn.shape)}, m_p {tuple(m_p.shape)}, noise {tuple(noise.shape)}")
        with torch.cuda.device(dev):
            z_p = torch.empty_like(noise)
            self._check(self.lib.vitsb_length_regulate(self._h, attn.data_ptr(), m_p.data_ptr(), s_p.data_ptr(), noise.data_ptr(),
                                                       B, Ty, Tx, z_p.data_ptr(), self._stream()))
        return z_p

    @torch.no_grad()
    def infer_p1(self, text: torch.Tensor, emo: torch.Tensor, sid, x_lengths: Optional[Sequence[int]] = None):
        """SynthesizerTrn.infer_p1 (models.py:558-566) on the library, fp32: text [B,Tx,text_channels], emo [B,1024],
        sid [B] -> (m_p [B,inter,Tx], s_p, logw [B,1,Tx], g [B,gin]).  B > 1 rows are ragged (x_lengths), each equal to
        its own B = 1 call.  Needs the checkpoint's enc_p.* / dp.* / emb_g.* entries."""
        dev = self.device
        text = text.to(dev, torch.float32).contiguous()
        emo = emo.to(dev, torch.float32).reshape(text.size(0), -1).contiguous()
        if text.dim() != 3:
            raise ValueError("expected text [B,Tx,text_channels]")
        B, Tx, _ = text.shape
        sid_l = [int(v) for v in (sid.tolist() if torch.is_tensor(sid) else sid)]
        if len(sid_l) != B:
            raise ValueError("sid must have B entries")
        sid_a = (C.c_int32 * B)(*sid_l)
        len_a = None
        if x_lengths is not None:
            ll = [int(v) for v in (x_lengths.tolist() if torch.is_tensor(x_lengths) else x_lengths)]
            if len(ll) != B:
                raise ValueError("x_lengths must have B entries")
            len_a = (C.c_int32 * B)(*ll)
        with torch.cuda.device(dev):
            need = int(self.lib.vitsb_p1_workspace_bytes(self._h, B, Tx))
            if need == 0:
                raise L.VitsbError(3, "infer_p1 weights not loaded (state_dict had no enc_p.* / dp.* / emb_g.* entries)")
            if self._ws1 is None or self._ws1.numel() < need:
                self._ws1 = torch.empty(need, dtype=torch.uint8, device=dev)
            m_p = torch.empty(B, self.inter_channels, Tx, dtype=torch.float32, device=dev)
            s_p = torch.empty_like(m_p)
            logw = torch.empty(B, 1, Tx, dtype=torch.float32, device=dev)
            g = torch.empty(B, self.gin_channels, dtype=torch.float32, device=dev)
            self._check(self.lib.vitsb_infer_p1(self._h, text.data_ptr(), emo.data_ptr(), sid_a, len_a, B, Tx, m_p.data_ptr(),
                                                s_p.data_ptr(), logw.data_ptr(), g.data_ptr(), self._ws1.data_ptr(),
                                                self._ws1.numel(), self._stream()))
        return m_p, s_p, logw, g

    @torch.no_grad()
    def synthesize_durations(self, m_p, s_p, logw, g, noise, T_cap: int, x_lengths: Optional[Sequence[int]] = None,
                             length_scale: float = 1.0, compact_noise: bool = False, ragged: bool = True, pcm16: bool = False,
                             volume: float = 1.0, return_z: bool = False):
        """infer_p1's outputs -> waveform with the durations handled on the device (replaces the glue of
        infer.py:169-176 and infer_p2, models.py:568-575): no `.item()` sync, no dense attn.  `noise` is [B,inter,>=T_cap]
        (already scaled, infer.py:69) or, with compact_noise, the reference's flat pool viewed as [inter, y_len] blocks.
        Returns (wav or pcm [B,1,T_cap*up] - zero beyond each row's length, y_len int32 [B] on the device[, z])."""
        dev = self.device
        m_p, s_p, logw, g, noise = (t.to(dev, torch.float32).contiguous() for t in (m_p, s_p, logw, g, noise))
        B, IC, Tx = m_p.shape
        if IC != self.inter_channels or tuple(s_p.shape) != (B, IC, Tx) or logw.numel() != B * Tx or tuple(g.shape) != (B, self.gin_channels):
            raise ValueError("expected m_p, s_p [B,inter,Tx], logw [B,1,Tx], g [B,gin]")
        if compact_noise:
            stride = 0
            if noise.numel() < B * IC * T_cap:
                raise ValueError("compact noise pool must hold B * inter * T_cap values")
        else:
            if noise.dim() != 3 or noise.size(0) != B or noise.size(1) != IC or noise.size(2) < T_cap:
                raise ValueError("expected noise [B,inter,>=T_cap]")
            stride = noise.size(2)
        len_a = None
        if x_lengths is not None:
            ll = [int(v) for v in (x_lengths.tolist() if torch.is_tensor(x_lengths) else x_lengths)]
            len_a = (C.c_int32 * B)(*ll)
        with torch.cuda.device(dev):
            ws = self._workspace(B, T_cap)
            out = torch.empty(B, 1, T_cap * self.total_upsample, dtype=torch.int16 if pcm16 else torch.float32, device=dev)
            y_len = torch.empty(B, dtype=torch.int32, device=dev)
            z = torch.empty(B, IC, T_cap, dtype=torch.float32, device=dev) if return_z else None
            self._check(self.lib.vitsb_synthesize_durations(
                self._h, m_p.data_ptr(), s_p.data_ptr(), logw.data_ptr(), g.data_ptr(), len_a, noise.data_ptr(), stride, B, Tx, int(T_cap),
                float(length_scale), int(bool(ragged)), z.data_ptr() if return_z else None, None if pcm16 else out.data_ptr(),
                out.data_ptr() if pcm16 else None, float(volume), y_len.data_ptr(), ws.data_ptr(), ws.numel(), self._stream()))
        return (out, y_len, z) if return_z else (out, y_len)

    @torch.no_grad()
    def infer(self, text, emo, sid, noise, T_cap: int, x_lengths: Optional[Sequence[int]] = None, length_scale: float = 1.0,
              **kw):
        """SynthesizerTrn.infer (models.py:537-556) / EmoVITS.infer (infer.py:136-184) end to end on the library with
        injected noise: infer_p1 -> device-side durations -> reverse flow -> decoder, no host synchronisation in between."""
        m_p, s_p, logw, g = self.infer_p1(text, emo, sid, x_lengths)
        return self.synthesize_durations(m_p, s_p, logw, g, noise, T_cap, x_lengths, length_scale, **kw)

    @torch.no_grad()
    def infer_p2(self, attn, m_p, s_p, g, noise) -> torch.Tensor:
        """SynthesizerTrn.infer_p2 (models.py:568-575): length regulation, reverse flow, decoder - all on the CUDA path."""
        return self.synthesize(self.length_regulate(attn, m_p, s_p, noise), g)

    @torch.no_grad()
    def synthesize_host(self, z_p_host: torch.Tensor, g_host: torch.Tensor, lengths: Sequence[int],
                        ragged: bool = True, out_host: Optional[torch.Tensor] = None) -> torch.Tensor:
        """End-to-end call with HOST buffers (pinned for async copies): H2D of the
        latents, synthesis, D2H of the waveform, one stream synchronisation."""
        with torch.cuda.device(self.device):
            z_p = z_p_host.to(self.device, non_blocking=True)
            g = g_host.to(self.device, non_blocking=True)
            wav = self.synthesize(z_p, g, lengths, ragged)
            if out_host is None:
                out_host = torch.empty(wav.shape, dtype=torch.float32, pin_memory=True)
            out_host.copy_(wav, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
        return out_host

    @torch.no_grad()
    def synthesize_segments(self, segments: Sequence[Sequence[torch.Tensor]], max_batch_frames: int = 16384,
                            max_batch: int = 64) -> list:
        """Batched form of the serving loop's per-segment calls (`VITSWrap.speaking` runs the segments of a request one
        after the other, vits_wrap.py:177-183).  `segments` is a list of (z_p [1,inter,T_i] or
        [inter,T_i], g [1,gin] or [gin]); they are sorted by length, packed into ragged batches of at most `max_batch`
        rows and `max_batch_frames` padded frames (B * T_max: bounds the workspace), and each batch is ONE synthesize
        call.  Returns the waveforms [1,1,192*T_i] in the order of `segments`; by row independence of the ragged plan
        every waveform is bit-identical to the B=1 call on that segment."""
        items = []
        for i, (z, g) in enumerate(segments):
            z = z if z.dim() == 3 else z.unsqueeze(0)
            g = g if g.dim() == 2 else g.reshape(1, -1)
            if z.size(0) != 1 or z.size(1) != self.inter_channels or g.size(0) != 1:
                raise ValueError(f"segment {i}: expected z_p [1,{self.inter_channels},T] and g [1,{self.gin_channels}]")
            items.append((int(z.size(2)), i, z, g))
        out = [None] * len(items)
        order = sorted(items, key=lambda it: -it[0])
        pos = 0
        while pos < len(order):
            T = order[pos][0]                                  # longest of the batch (descending order)
            n = max(1, min(max_batch, max_batch_frames // max(T, 1), len(order) - pos))
            batch = order[pos:pos + n]
            zb = torch.zeros(n, self.inter_channels, T, dtype=torch.float32, device=self.device)
            gb = torch.empty(n, self.gin_channels, dtype=torch.float32, device=self.device)
            lens = []
            for r, (Ti, _, z, g) in enumerate(batch):
                zb[r, :, :Ti] = z[0].to(self.device, torch.float32)
                gb[r] = g[0].to(self.device, torch.float32)
                lens.append(Ti)
            wav = self.synthesize(zb, gb, lens, ragged=True)
            for r, (Ti, i, _, _) in enumerate(batch):
                out[i] = wav[r:r + 1, :, :Ti * self.total_upsample].clone()
            pos += n
        return out

    def host_pipeline(self, B: int, T_max: int, depth: int = 2, pcm16: bool = False, volume: float = 1.0) -> "HostPipeline":
        """Serving loop for HOST buffers: see HostPipeline."""
        return HostPipeline(self, B, T_max, depth, pcm16, volume)


class HostPipeline:
    """Bounded-depth pipeline over `FlowDecoder.synthesize` for requests that live in pinned host memory.

    submit() enqueues H2D of the latents on a copy stream, the synthesis on the caller's current stream and
    the D2H of the waveform on a second copy stream, and returns a ticket; wait(ticket) blocks until that
    request's waveform is in its host buffer.  With depth 2 the copies of requests i-1 and i+1 overlap the
    synthesis of request i (PCIe moves ~10 MB each way per 16-utterance batch, ~8 % of the synthesis time).
    Every request still pays its own H2D and D2H; nothing is cached between requests."""

    def __init__(self, eng: FlowDecoder, B: int, T_max: int, depth: int = 2, pcm16: bool = False, volume: float = 1.0):
        """pcm16: the last kernel writes int16 PCM (vits_wrap.py:195, `volume` applied) and that is what travels to the
        host: half the device->host bytes of the float waveform."""
        self.eng, self.B, self.T, self.depth = eng, B, T_max, depth
        self.pcm16, self.volume = pcm16, volume
        self.out_dtype = torch.int16 if pcm16 else torch.float32
        dev = eng.device
        with torch.cuda.device(dev):
            self.h2d, self.d2h = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
            self.z = [torch.empty(B, eng.inter_channels, T_max, dtype=torch.float32, device=dev) for _ in range(depth)]
            self.g = [torch.empty(B, eng.gin_channels, dtype=torch.float32, device=dev) for _ in range(depth)]
            self.wav = [torch.empty(B, 1, T_max * eng.total_upsample, dtype=self.out_dtype, device=dev) for _ in range(depth)]
            self.ev_in = [torch.cuda.Event() for _ in range(depth)]      # latents of the slot are on the device
            self.ev_comp = [torch.cuda.Event() for _ in range(depth)]    # synthesis of the slot has finished
            self.ev_out = [torch.cuda.Event() for _ in range(depth)]     # waveform of the slot is in host memory
        self.n = 0

    def submit(self, z_p_host: torch.Tensor, g_host: torch.Tensor, lengths: Sequence[int], out_host: torch.Tensor,
               ragged: bool = True, after: Optional[torch.cuda.Event] = None) -> int:
        if tuple(z_p_host.shape) != tuple(self.z[0].shape) or tuple(g_host.shape) != tuple(self.g[0].shape):
            raise ValueError("HostPipeline: request shape differs from the pipeline's (B, T_max)")
        s = self.n % self.depth
        dev = self.eng.device
        with torch.cuda.device(dev):
            cur = torch.cuda.current_stream(dev)
            if self.n >= self.depth:
                self.h2d.wait_event(self.ev_comp[s])         # the slot's previous synthesis has consumed its inputs
            if after is not None:
                self.h2d.wait_event(after)
            with torch.cuda.stream(self.h2d):
                self.z[s].copy_(z_p_host, non_blocking=True)
                self.g[s].copy_(g_host, non_blocking=True)
                self.ev_in[s].record(self.h2d)
            cur.wait_event(self.ev_in[s])
            if self.n >= self.depth:
                cur.wait_event(self.ev_out[s])               # the slot's previous waveform has left the device
            if self.pcm16:
                self.eng.synthesize_pcm16(self.z[s], self.g[s], lengths, ragged, volume=self.volume, out=self.wav[s])
            else:
                self.eng.synthesize(self.z[s], self.g[s], lengths, ragged, out=self.wav[s])
            self.ev_comp[s].record(cur)
            self.d2h.wait_event(self.ev_comp[s])
            with torch.cuda.stream(self.d2h):
                out_host.copy_(self.wav[s], non_blocking=True)
                self.ev_out[s].record(self.d2h)
        self.n += 1
        return self.n - 1

    def wait(self, ticket: int) -> None:
        # The slot's event may have been re-recorded by a later request that reused the slot; the D2H stream is in
        # order, so that later event also covers `ticket`'s copy.
        self.ev_out[ticket % self.depth].synchronize()

    def join_event(self, ticket: int) -> torch.cuda.Event:
        """Event that fires when `ticket`'s waveform is in host memory (to order other streams after it)."""
        return self.ev_out[ticket % self.depth]
