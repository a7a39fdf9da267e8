"""CPU, world_size 2 over gloo: the utterance sharding (LPT) and the whole-job
aggregation bench.py uses for N>1 (sum of audio over ranks / max of time over ranks).
The synthesis path itself has no collective (replicas only)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import bench


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lens = bench.global_lengths(world)
    mine = bench.lpt_shard(lens, world)[rank]
    audio = sum(lens[i] for i in mine) * bench.FRAME_SEC
    ms = 10.0 + rank                                   # pretend rank 1 is slower
    audio_all, frames_all, ms_max = bench.reduce_job(audio, sum(lens[i] for i in mine), ms, world, torch.device("cpu"))
    q.put((rank, sorted(mine), audio, audio_all, frames_all, ms_max))
    dist.barrier()
    dist.destroy_process_group()


def test_lpt_shard_and_aggregation_world2():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    lens = bench.global_lengths(world)
    assert len(lens) == bench.PER_GPU_BATCH * world
    all_idx = sorted(res[0][1] + res[1][1])
    assert all_idx == list(range(len(lens)))           # every utterance on exactly one rank
    loads = [sum(lens[i] for i in r[1]) for r in res]
    assert abs(loads[0] - loads[1]) <= max(lens)       # LPT balance: within one utterance
    total_audio = sum(lens) * bench.FRAME_SEC
    for r in res:
        assert abs(r[3] - total_audio) < 1e-6          # whole-job audio = sum over ranks
        assert r[4] == sum(lens)
        assert r[5] == 11.0                            # job time = max over ranks


def test_lpt_shard_single_rank_is_identity():
    lens = bench.global_lengths(1)
    assert sorted(bench.lpt_shard(lens, 1)[0]) == list(range(16))
    assert min(lens) >= bench.LEN_LO and max(lens) <= bench.LEN_HI
