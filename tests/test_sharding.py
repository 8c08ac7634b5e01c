"""
The multi-GPU host logic on CPU: voice sharding covers every voice exactly once, and the
partial-mix reduce (torch.distributed, gloo, world_size 2) sums shards onto rank 0 — the same
calls bench.py and VoiceBank.render_sharded make with NCCL on the GPUs.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from synthesizer_b200.voicebank import shard_range


@pytest.mark.parametrize("n,world", [(65536, 8), (8192, 1), (10, 3), (7, 8), (0, 2), (128, 4)])
def test_shard_range_partitions(n, world):
    covered = []
    for r in range(world):
        b, e = shard_range(n, r, world)
        assert 0 <= b <= e <= n
        covered.extend(range(b, e))
    assert covered == list(range(n))
    sizes = [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]
    assert max(sizes) - min(sizes) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nvoices, frames, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, e = shard_range(nvoices, rank, world)
    # stand-in for the per-rank partial mix: voice v contributes v+1 to every frame
    mix = torch.zeros(2 * frames, dtype=torch.float32)
    for v in range(b, e):
        mix += float(v + 1)
    dist.reduce(mix, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        out.put(mix.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_partial_mix_reduce_gloo_world2():
    nvoices, frames, world = 11, 64, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, nvoices, frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert np.all(got == nvoices * (nvoices + 1) / 2)
