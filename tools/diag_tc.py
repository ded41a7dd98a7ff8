"""Development diagnostic (GPU box): run each contraction through one kernel family and report how
it compares with the oracle, without asserting, so one gpurun call shows every failure mode.
usage: python tools/diag_tc.py simt|tc [probe]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from oracle.oracle import Oracle, shape
from mandheling_dsp_training_b200 import NitiB200
from mandheling_dsp_training_b200.lib import PATH_SIMT, PATH_TC, PATH_TMA
from tests.gpu_util import Runner

path = sys.argv[1] if len(sys.argv) > 1 else "tc"
nb = NitiB200(0); run = Runner(nb); orc = Oracle()
nb.set_path({"simt": PATH_SIMT, "tc": PATH_TC, "tma": PATH_TMA, "auto": 0}[path])


def report(tag, got, want):
    if np.array_equal(got, want):
        print(f"  {tag}: OK", flush=True); return
    bad = np.argwhere(got != want)
    print(f"  {tag}: MISMATCH {len(bad)}/{got.size}; first {bad[:6].tolist()} got {got[tuple(bad[0])]} want {want[tuple(bad[0])]}; "
          f"got-nonzero {np.count_nonzero(got)} want-nonzero {np.count_nonzero(want)}", flush=True)
    # which channels / pixels are wrong (NCHW): summarise along axes
    for ax, name in enumerate(["n", "ch", "y", "x"][:got.ndim]):
        other = tuple(i for i in range(got.ndim) if i != ax)
        wrong = np.any(got != want, axis=other)
        print(f"      wrong along {name}: {np.flatnonzero(wrong)[:40].tolist()} ({wrong.sum()}/{wrong.size})", flush=True)


cases = [(2, 64, 8, 8, 64, 3, 1, 1), (2, 128, 8, 8, 128, 3, 1, 1), (2, 64, 9, 9, 64, 3, 1, 1), (2, 128, 9, 9, 64, 3, 2, 1),
         (256, 512, 1, 1, 512, 1, 1, 0), (3, 128, 7, 5, 192, 3, 1, 1),
         (2, 16, 8, 8, 32, 3, 1, 1), (4, 64, 8, 8, 128, 3, 1, 1), (2, 128, 6, 6, 256, 3, 1, 1),
         (3, 32, 9, 7, 48, 3, 1, 1), (2, 32, 10, 10, 64, 3, 2, 1), (1, 256, 4, 4, 512, 3, 1, 1),
         (64, 832, 1, 1, 512, 1, 1, 0)]
for cfg in cases:
    n, c, h, w, oc, k, st, p = cfg
    s = shape(n, c, h, w, oc, k, stride=st, pad=p)
    rng = np.random.default_rng(1)
    x = rng.integers(-127, 128, size=(n, c, h, w), dtype=np.int8)
    wt = rng.integers(-64, 65, size=(oc, c, k, k), dtype=np.int8)
    dy = rng.integers(-50, 51, size=(n, oc, s.oh, s.ow), dtype=np.int8)
    print(cfg, flush=True)
    for name, fn, ofn, args in [("fwd", run.conv_fwd_acc, orc.conv_fwd_acc, (s, x, wt)),
                                ("dx", run.conv_dx_acc, orc.conv_dx_acc, (s, dy, wt)),
                                ("dw", run.conv_dw_acc, orc.conv_dw_acc, (s, x, dy))]:
        try:
            t = time.time(); acc, mx = fn(*args); dt = time.time() - t
            oacc, _ = ofn(*args)
            report(f"{name} [{nb.last_path}] mx={mx} want {int(np.abs(oacc).max())} ({dt*1e3:.1f} ms)", acc, oacc)
        except Exception as ex:
            print(f"  {name}: EXC {ex}", flush=True)

if "probe" in sys.argv:
    for it in (2000, 8000):
        print("i8 tcgen05 issue-rate probe:", it, "iters ->", round(nb.probe_i8_peak(it), 1), "TOPS", flush=True)
print("launches", nb.launches)
