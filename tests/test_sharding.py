"""
The multi-GPU host logic on CPU: voice sharding covers every voice exactly once; the product's
own VoiceBank.render_sharded runs under torch.distributed (gloo, world_size 2) with the kernel
launch stubbed out (it is the call bench.py makes with NCCL on the GPUs); and the
destination-range sharding of Sample.mix_at_many cuts the placements into pieces that tile
the originals exactly.
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


def _sharded_worker(rank, world, port, nvoices, frames, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from synthesizer_b200 import _cabi as C
    from synthesizer_b200.voicebank import VoiceBank
    bank = VoiceBank.from_voice_array([C.Voice() for _ in range(nvoices)], 48000)
    calls = []

    def stub_render(nframes, start=0, begin=0, end=None, **kw):
        # stands in for the kernel launch: voice v contributes v+1 to every value of the mix
        calls.append((nframes, start, begin, end))
        mix = torch.zeros(2 * nframes, dtype=torch.float32)
        for v in range(begin, len(bank) if end is None else end):
            mix += float(v + 1)
        return mix

    bank.render = stub_render
    mix = bank.render_sharded(frames, start=7)           # the product function under test
    assert calls == [(frames, 7) + shard_range(nvoices, rank, world)], calls
    if rank == 0:
        out.put(mix.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_render_sharded_gloo_world2():
    """VoiceBank.render_sharded itself: each rank renders exactly its slice of the voices, and
    rank 0 ends up with the sum over all voices."""
    nvoices, frames, world = 11, 64, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sharded_worker, args=(r, world, port, nvoices, frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert got.shape == (2 * frames,) and np.all(got == nvoices * (nvoices + 1) / 2)


@pytest.mark.parametrize("total,world", [(28_800_000, 8), (100_000, 3), (2048, 4), (5, 2), (0, 2)])
def test_mix_at_shard_bounds_partition(total, world):
    from synthesizer_b200.sample import Sample
    edges = [Sample.shard_bounds(total, r, world) for r in range(world)]
    assert edges[0][0] == 0 and edges[-1][1] == total
    for (a, b), (c, d) in zip(edges, edges[1:]):
        assert b == c and a <= b
    assert all(a % 2048 == 0 for a, _ in edges)          # cuts sit on the kernel's tile boundaries


def test_mix_at_clipped_placements_tile_the_originals():
    """Every destination sample of every source lands in exactly one rank's piece, at the same
    source sample, and the pieces keep the list order (mix_at saturates per step, in order)."""
    from synthesizer_b200.sample import Sample
    rng = np.random.default_rng(4)
    total, world, n = 300_000, 4, 500
    lens = rng.integers(0, 9000, n)
    starts = rng.integers(0, total - 9000, n)
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]])
    seen = np.zeros(int(lens.sum()), dtype=np.int64)      # per source sample: destination it was sent to (+1)
    for r in range(world):
        lo, hi = Sample.shard_bounds(total, r, world)
        o, l, d = Sample.clip_placements(offs, lens, starts, lo, hi)
        assert np.all(l > 0) and np.all(d >= 0) and np.all(d + l <= hi - lo)
        src_index = np.searchsorted(offs, o, side="right") - 1
        assert np.all(np.diff(src_index[l > 0]) >= 0)     # original order kept
        for oo, ll, dd in zip(o, l, d):
            assert np.all(seen[oo:oo + ll] == 0)
            seen[oo:oo + ll] = lo + dd + np.arange(ll) + 1
    want = np.concatenate([s + np.arange(l) + 1 for s, l in zip(starts, lens)])
    assert np.array_equal(seen, want)
