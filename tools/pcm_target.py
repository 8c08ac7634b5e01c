"""
One HBM-bound Sample kernel on the sizes bench_extra.py quotes, a few launches: a target for
`ncu --set full -k regex:<kernel>`.   python tools/pcm_target.py <fadeout|ratecv|mix_at|add|mul|tomono>
(mix_at = BASELINE config 4's kernel: 4096 hits of 3000-29000 samples into 10 min @48 kHz mono)
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from synthesizer_b200 import _cabi as C, _device as D      # noqa: E402

op = sys.argv[1]
torch.cuda.set_device(0)
n = 600 * 48000 * 2
a = torch.randint(-20000, 20000, (n,), dtype=torch.int16, device="cuda")
b = torch.randint(-20000, 20000, (n,), dtype=torch.int16, device="cuda")
out = torch.empty_like(a)
if op == "fadeout":
    fn = lambda: C.call("sb_pcm16_fadeout", D.ptr(out), n, 0.1, D.stream())
elif op == "add":
    fn = lambda: C.call("sb_pcm16_add", D.ptr(a), D.ptr(b), n, D.ptr(out), D.stream())
elif op == "mul":
    fn = lambda: C.call("sb_pcm16_mul", D.ptr(a), n, 0.731, D.ptr(out), D.stream())
elif op == "tomono":
    fn = lambda: C.call("sb_pcm16_tomono", D.ptr(a), n // 2, 0.5, 0.5, D.ptr(out), D.stream())
elif op == "ratecv":
    nin = 44100 * 600
    nout = C.lib.sb_pcm16_ratecv_frames(nin, 44100, 48000)
    dst = torch.empty(nout * 2, dtype=torch.int16, device="cuda")
    fn = lambda: C.call("sb_pcm16_ratecv", D.ptr(a), nin, 2, 44100, 48000, D.ptr(dst), D.stream())
elif op == "mix_at":
    rng = np.random.default_rng(0)
    total = 600 * 48000
    lens = rng.integers(3000, 29000, 4096).astype(np.int64)
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    starts = rng.integers(0, total - 30000, 4096).astype(np.int64)
    pool = torch.randint(-8000, 8000, (int(lens.sum()),), dtype=torch.int16, device="cuda")
    idx = torch.tensor(np.stack([offs, lens, starts]), dtype=torch.int64, device="cuda")
    dstbuf = torch.zeros(total, dtype=torch.int16, device="cuda")
    nscr = C.lib.sb_pcm16_mix_at_many_scratch_bytes(total, 4096, int(lens.sum()))
    scr = torch.empty(nscr, dtype=torch.uint8, device="cuda")
    fn = lambda: C.call("sb_pcm16_mix_at_many", D.ptr(dstbuf), total, D.ptr(pool), D.ptr(idx[0]), D.ptr(idx[1]),
                        D.ptr(idx[2]), 4096, D.ptr(scr), nscr, D.stream())
    print("hit samples", int(lens.sum()))
else:
    raise SystemExit("unknown op")
flush = torch.zeros(512 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    flush.add_(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    print("%s: %.4f ms" % (op, e0.elapsed_time(e1)))
