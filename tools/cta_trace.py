"""
Per-warp timeline of the fast kernels of one bank render (debugging build:
  SB_EXTRA=-DSB_CTA_TRACE SB_OUT=build/trace/libsynth_b200.so SB_OBJ=build/trace/obj bash synthesizer_b200/csrc/build.sh
  SYNTH_B200_LIB=build/trace/libsynth_b200.so python tools/cta_trace.py <carrier> <modulator> [voices]
arguments as tools/class_bank.py).  Prints, per kernel part: the kernel's span, the distribution of warp
durations, busy time per SM, and the slowest units.
"""
import ctypes
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv_saved = list(sys.argv)
import synthesizer_b200 as sb            # noqa: E402
from synthesizer_b200 import _cabi       # noqa: E402

lib = _cabi.lib
NW = 1 << 16
buf = np.zeros((3, NW, 4), dtype=np.uint64)
lib.sb_debug_cta_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
exec(open(os.path.join(os.path.dirname(__file__), "class_bank.py")).read().split("vd, _ = b._upload")[0])
vd, _ = b._upload(0, None)
g, sm, e, sc = b.count_classes()
kw = dict(voices_dev=vd, generic_voices=g, smooth_voices=sm, edged_voices=e, scan_voices=sc)
b.render(nfr, **kw)
lib.sb_debug_cta_trace(None, 1)
b.render(nfr, **kw)
lib.sb_debug_cta_trace(buf.ctypes.data, 0)
for part in range(3):
    t = buf[part]
    live = t[:, 1] > 0
    if not live.any():
        continue
    t0, t1 = t[live, 0].astype(np.int64), t[live, 1].astype(np.int64)
    grp, span = (t[live, 2] >> np.uint64(32)).astype(np.int64), (t[live, 2] & np.uint64(0xffffffff)).astype(np.int64)
    nv, smid = (t[live, 3] >> np.uint64(32)).astype(np.int64), (t[live, 3] & np.uint64(0xff)).astype(np.int64)
    fixes = ((t[live, 3] >> np.uint64(20)) & np.uint64(0xfff)).astype(np.int64)      # fix-up calls (divergent groups)
    lows = ((t[live, 3] >> np.uint64(8)) & np.uint64(0xfff)).astype(np.int64)        # lane-runs taken exactly throughout
    d = (t1 - t0) / 1e3
    begin = t0.min()
    print("part %d: %d warps, kernel span %.1f us; warp duration us: min %.1f  median %.1f  p90 %.1f  p99 %.1f  max %.1f"
          % (part, live.sum(), (t1.max() - begin) / 1e3, d.min(), np.median(d), np.percentile(d, 90), np.percentile(d, 99), d.max()))
    busy = np.zeros(smid.max() + 1)
    last = np.zeros(smid.max() + 1)
    np.add.at(busy, smid, d)
    np.maximum.at(last, smid, (t1 - begin) / 1e3)
    print("   per SM: warp-busy us min %.0f median %.0f max %.0f; last warp ends at us: min %.0f median %.0f max %.0f"
          % (busy.min(), np.median(busy), busy.max(), last.min(), np.median(last), last.max()))
    print("   fix-up calls per warp: median %d, p99 %d" % (np.median(fixes), np.percentile(fixes, 99)))
    order = np.argsort(-d)[:12]
    for i in order:
        print("   slow: group %d span %d voices %d sm %d: %.1f us, starts at %.1f; fix-up calls %d, whole runs %d"
              % (grp[i], span[i], nv[i], smid[i], d[i], (t0[i] - begin) / 1e3, fixes[i], lows[i]))
    late = np.argsort(-(t1 - begin))[:6]
    for i in late:
        print("   late: group %d span %d voices %d sm %d: %.1f us, starts at %.1f, ends %.1f" % (grp[i], span[i], nv[i], smid[i], d[i], (t0[i] - begin) / 1e3, (t1[i] - begin) / 1e3))
