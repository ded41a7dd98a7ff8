"""BASELINE.json config 5: int8 conv fwd / dX / dW and GEMM shape sweep over the VGG-11 (b256) and ResNet-18
(b128) layer shapes of SURVEY.md section 8 — TOP/s of the phase-1 contraction kernel (int8 x int8 -> int32 + max|acc|; columns
fwd / dx / dw) and of the WHOLE fused op the training step launches (contraction + max|acc| + exponent + rule A + int8 store in one
launch, no int32 in HBM; columns fwd_fused / dx_fused) against the measured dense INT8 tcgen05 peak.  ops = 2*M*N*K with the unpadded extents.

Each shape: operands resident in HBM, 3 warm-up launches, then `iters` launches timed with CUDA events on the
launching stream, a 256 MiB write between launches to flush L2 (timed separately and subtracted? no: the event
pair brackets only the contraction call).  Prints one JSON object per shape and a summary line.

usage: python tools/sweep.py [--iters 10] [--out profiles/sweep_rNN.json]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# (name, n, c, h, w, oc, k, stride, pad)
SHAPES = [
    ("vgg.conv2  64->128@16^2 b256", 256, 64, 16, 16, 128, 3, 1, 1),
    ("vgg.conv3 128->256@8^2  b256", 256, 128, 8, 8, 256, 3, 1, 1),
    ("vgg.conv4 256->256@8^2  b256", 256, 256, 8, 8, 256, 3, 1, 1),
    ("vgg.conv5 256->512@4^2  b256", 256, 256, 4, 4, 512, 3, 1, 1),
    ("vgg.conv6 512->512@4^2  b256", 256, 512, 4, 4, 512, 3, 1, 1),
    ("vgg.conv7 512->512@2^2  b256", 256, 512, 2, 2, 512, 3, 1, 1),
    ("vgg.fc    512->512      b256", 256, 512, 1, 1, 512, 1, 1, 0),
    ("res.l1   64->64@56^2   b128", 128, 64, 56, 56, 64, 3, 1, 1),
    ("res.l2a  64->128@56^2 s2 b128", 128, 64, 56, 56, 128, 3, 2, 1),
    ("res.l2  128->128@28^2  b128", 128, 128, 28, 28, 128, 3, 1, 1),
    ("res.l3a 128->256@28^2 s2 b128", 128, 128, 28, 28, 256, 3, 2, 1),
    ("res.l3  256->256@14^2  b128", 128, 256, 14, 14, 256, 3, 1, 1),
    ("res.l4a 256->512@14^2 s2 b128", 128, 256, 14, 14, 512, 3, 2, 1),
    ("res.l4  512->512@7^2   b128", 128, 512, 7, 7, 512, 3, 1, 1),
    ("res.ds  64->128@56^2 1x1 s2 b128", 128, 64, 56, 56, 128, 1, 2, 0),
    ("res.fc  512->1000      b128", 128, 512, 1, 1, 1000, 1, 1, 0),
    ("gemm-like 1x1 1024->1024 M=65536", 64, 1024, 32, 32, 1024, 1, 1, 0),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-flush", action="store_true")
    args = ap.parse_args()
    import numpy as np
    import torch
    from mandheling_dsp_training_b200 import NitiB200
    from mandheling_dsp_training_b200.lib import desc
    assert torch.cuda.is_available(), "the sweep needs a CUDA device (no CPU fallback)"
    nb = NitiB200(0)
    peak = nb.probe_i8_peak(4000)
    flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda")
    g = torch.Generator(device="cuda"); g.manual_seed(0)
    rows = []
    for name, n, c, h, w, oc, k, st, pad in SHAPES:
        d = desc(n, c, h, w, oc, k, stride=st, pad=pad)
        oh, ow = d.oh, d.ow
        x = torch.randint(-127, 128, (n, h, w, c), dtype=torch.int8, device="cuda", generator=g)
        wt = torch.randint(-127, 128, (oc, k, k, c), dtype=torch.int8, device="cuda", generator=g)
        dy = torch.randint(-40, 41, (n, oh, ow, oc), dtype=torch.int8, device="cuda", generator=g)
        ops = 2.0 * n * oh * ow * oc * c * k * k
        acc_y = torch.empty(n * oh * ow * oc, dtype=torch.int32, device="cuda")
        acc_x = torch.empty(n * h * w * c, dtype=torch.int32, device="cuda")
        acc_w = torch.empty(oc * k * k * c, dtype=torch.int32, device="cuda")
        mx = torch.zeros(4, dtype=torch.int32, device="cuda")
        # whole ops as the training step runs them (one launch: contraction + max|acc| + exponent + rule A + int8 store; the
        # exponent prediction has settled after the warm-up launches, so the timed launches are single-sweep hits)
        y8 = torch.empty((n, oh, ow, oc), dtype=torch.int8, device="cuda"); x8 = torch.empty((n, h, w, c), dtype=torch.int8, device="cuda")
        ey = torch.zeros(1, dtype=torch.int8, device="cuda"); e7 = torch.full((1,), -7, dtype=torch.int8, device="cuda")
        wtt = wt.permute(3, 1, 2, 0).contiguous()
        calls = {"fwd": lambda: nb.conv_fwd_acc(d, x, wt, acc_y, mx),
                 "dx": lambda: nb.conv_dx_acc(d, dy, wt, acc_x, mx),
                 "dw": lambda: nb.conv_dw_acc(d, x, dy, acc_w, mx),
                 "fwd_fused": lambda: nb.conv_fwd_fused(d, x, wt, e7, e7, y8, ey, True, acc_y, mx[1:2], dp_slot=1),
                 "dx_fused": lambda: nb.conv_dx_fused(d, dy, wt, wtt, x8, None, acc_x, mx[2:3], dp_slot=2)}
        row = {"shape": name, "M": n * oh * ow, "N": oc, "K": c * k * k, "gop": ops / 1e9}
        for kind, fn in calls.items():
            for _ in range(4):
                mx.zero_(); fn()
            torch.cuda.synchronize()
            ts = []
            for _ in range(args.iters):
                if not args.no_flush:
                    flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda._sleep(1_000_000)      # ~0.5 ms: the host (ctypes call + 2-3 tensor-map encodes) runs ahead of the device
                mx.zero_()
                a.record(); fn(); b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b) * 1e-3)
            t = float(np.median(ts))
            row[kind] = {"us": round(t * 1e6, 2), "TOP/s": round(ops / t / 1e12, 1), "frac": round(ops / t / 1e12 / peak, 4),
                         "path": nb.last_path}
        rows.append(row)
        print(json.dumps(row), flush=True)
        del x, wt, dy, acc_y, acc_x, acc_w, y8, x8, wtt
    best = max(r[k]["TOP/s"] for r in rows for k in ("fwd", "dx", "dw", "fwd_fused", "dx_fused"))
    summary = {"peak_TOPs_measured": round(peak, 1), "best_TOPs": best, "best_frac": round(best / peak, 4),
               "l2": "not flushed" if args.no_flush else "flushed (256 MiB write) before every timed launch",
               "iters": args.iters, "rows": rows}
    print(json.dumps({k: v for k, v in summary.items() if k != "rows"}))
    if args.out:
        with open(os.path.join(ROOT, args.out) if not os.path.isabs(args.out) else args.out, "w") as f:
            json.dump(summary, f, indent=1)


if __name__ == "__main__":
    main()
