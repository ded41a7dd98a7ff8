"""Run one op a few times (the short command ncu wraps).  usage: python tools/prof_op.py stem|dw|fwd|dx [n c h w oc k stride pad]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mandheling_dsp_training_b200 import NitiB200
from mandheling_dsp_training_b200.lib import desc

def main():
    op = sys.argv[1] if len(sys.argv) > 1 else "stem"
    dims = [int(v) for v in sys.argv[2:10]] if len(sys.argv) >= 10 else [256, 16, 32, 32, 64, 3, 1, 1]
    n, c, h, w, oc, k, st, pad = dims
    nb = NitiB200(0)
    d = desc(n, c, h, w, oc, k, stride=st, pad=pad)
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randint(-127, 128, (n, h, w, c), dtype=torch.int8, device="cuda", generator=g)
    wt = torch.randint(-64, 65, (oc, k, k, c), dtype=torch.int8, device="cuda", generator=g)
    dy = torch.randint(-50, 51, (n, d.oh, d.ow, oc), dtype=torch.int8, device="cuda", generator=g)
    y = torch.empty((n, d.oh, d.ow, oc), dtype=torch.int8, device="cuda")
    e = torch.full((1,), -7, dtype=torch.int8, device="cuda"); ey = torch.zeros(1, dtype=torch.int8, device="cuda")
    acc = torch.empty(max(y.numel(), x.numel(), wt.numel()), dtype=torch.int32, device="cuda")
    mx = torch.zeros(1, dtype=torch.int32, device="cuda")
    trace = torch.zeros(148 * 8, dtype=torch.int64, device="cuda")
    if os.environ.get("TRACE"):
        nb.debug_trace(trace)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(4):
        mx.zero_()
        t0.record()
        if op == "stem":
            nb.conv_fwd_fused(d, x, wt, e, e, y, ey, True, acc, mx)
        elif op == "fwd":
            nb.conv_fwd_acc(d, x, wt, acc, mx)
        elif op == "dx":
            nb.conv_dx_acc(d, dy, wt, acc, mx)
        else:
            nb.conv_dw_acc(d, x, dy, acc, None)
        t1.record(); torch.cuda.synchronize()
        print(f"{op} {dims}: {t0.elapsed_time(t1) * 1e3:.1f} us  path={nb.last_path}", flush=True)

    if os.environ.get("TRACE"):
        t = trace.cpu().numpy().reshape(148, 8)
        t = t[t[:, 0] > 0]
        base = t[:, 0].min()
        names = ["entry", "setup", "tile0 ready", "max pass", "grid barrier", "rescale+store", "exit"]
        print("phase: min / median / max ns since first CTA entry, over", len(t), "CTAs")
        for i, nme in enumerate(names):
            col = t[:, i] - base
            print(f"  {nme:14s} {col.min():7d} {int(np.median(col)):7d} {col.max():7d}")


if __name__ == "__main__":
    main()
