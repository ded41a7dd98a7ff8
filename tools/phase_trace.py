"""Per-layer phase breakdown of the fused forward / dX launches of one training step (niti_debug_trace).

usage: python tools/phase_trace.py [vgg11|lenet|resnet18] [batch] > profiles/phase_rN_<workload>.txt
Each fused launch writes %globaltimer stamps per CTA: 0 entry, 1 set-up done (barrier init, TMEM alloc, tensor-map prefetch,
cluster sync, griddepcontrol.wait), 2 first tile accumulated (K loop of the first resident tile), 3 max + predicted-store sweep done (one TMEM sweep, overlapping the
K loops of the CTA's later tiles), 4 grid barrier passed, 5 re-store done (only when the exponent prediction missed; else = 4),
6 next layer released, 7 exit.  Printed per launch: median over CTAs of each phase's
duration in us, the launch's wall time from the first entry to the last exit, and its event-timed duration."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from mandheling_dsp_training_b200 import NitiB200
from mandheling_dsp_training_b200.step import NitiNet, SPECS


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "vgg11"
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    spec_fn, chw = SPECS[wl]
    nb = NitiB200(0)
    net = NitiNet(nb, spec_fn(), batch, chw)
    net.init_weights(seed=1)
    rng = np.random.default_rng(0)
    x = torch.from_numpy(rng.integers(-127, 128, size=(batch, chw[1], chw[2], chw[0]), dtype=np.int8)).cuda()
    lab = torch.from_numpy(rng.integers(0, 10, size=batch).astype(np.int32)).cuda()
    net.set_input(x, -7, lab)
    for _ in range(3):
        net.step()
    torch.cuda.synchronize()
    trace = torch.zeros(24 * 160 * 8, dtype=torch.int64, device="cuda")    # [layer][block][8], chains walk up to 24 layers
    nb.debug_trace(trace)
    rows = []
    convs = iter([])

    def timer(kind, ops, fn):
        trace.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        rows.append((kind, nb.last_path, ops, a.elapsed_time(b) * 1e3, trace.cpu().numpy().copy()))
    net.timer = timer
    net.step()
    torch.cuda.synchronize()
    nb.debug_trace(None)
    # stamps (csrc/conv_chain.cu): per launch [0][block][0] entry, [0][block][1] set-up done, [last][block][7] exit; per layer
    # 2 first tile accumulated, 3 max + predicted-store sweep done, 4 grid barrier (+ peer exchange) passed,
    # 5 re-store + second barrier done (prediction missed; otherwise stamped right after 4), 6 next layer released
    names = ["K loop (first tile)", "max+store sweep", "grid barrier", "re-store (miss)", "release"]
    print(f"# phase trace, {wl} batch {batch}, one eager step (every launch synchronised: cold pipeline, no PDL overlap)")
    print("# per launch: kind path GOP | event us | in-kernel span us (first entry .. last exit) | set-up us")
    print("#   per layer: CTAs | median us per phase: " + " | ".join(names))
    tot = {}
    for kind, path, ops, ev_us, raw in rows:
        grids = [g for g in (148, 146, 144, 132, 130, 128) if False]
        t_all = raw.reshape(-1)
        if not (t_all > 0).any():
            print(f"{kind:4s} {path:6s} {ops / 1e9:7.2f} | {ev_us:7.1f} | (no trace: kernel family without phase stamps)")
            continue
        # the grid size is whatever makes entry stamps [0][b][0] contiguous: count blocks with an entry stamp
        entry = raw.reshape(-1, 8)[:, 0]
        grid = int((entry[:160] > 0).sum())
        t = raw[: (raw.size // (grid * 8)) * grid * 8].reshape(-1, grid, 8)
        nl = int((t[:, :, 3] > 0).any(axis=1).sum())
        t0, t1 = t[0, :, 0].min(), t[nl - 1, :, 7].max()
        setup = float(np.median(t[0, :, 1] - t[0, :, 0])) / 1e3
        print(f"{kind:4s} {path:6s} {ops / 1e9:7.2f} | {ev_us:7.1f} | {(t1 - t0) / 1e3:7.1f} | {setup:5.1f}")
        tot["launch overhead (event - span)"] = tot.get("launch overhead (event - span)", 0.0) + max(0.0, ev_us - (t1 - t0) / 1e3)
        tot["set-up"] = tot.get("set-up", 0.0) + setup
        prev_end = t[0, :, 1].astype(np.float64)
        for l in range(nl):
            act = t[l, :, 2] > 0
            k = np.where(act, t[l, :, 2] - prev_end, np.nan)
            ph = [np.nanmedian(k) / 1e3 if act.any() else 0.0,
                  float(np.median((t[l, :, 3] - np.where(act, t[l, :, 2], prev_end)))) / 1e3,
                  float(np.median(t[l, :, 4] - t[l, :, 3])) / 1e3,
                  float(np.median(t[l, :, 5] - t[l, :, 4])) / 1e3,
                  float(np.median(t[l, :, 6] - t[l, :, 5])) / 1e3 if l + 1 < nl else 0.0]
            print(f"     layer {l:2d}: {int(act.sum()):3d} | " + " | ".join(f"{v:6.1f}" for v in ph))
            for n_, v in zip(names, ph):
                tot[n_] = tot.get(n_, 0.0) + float(v)
            prev_end = (t[l, :, 6] if l + 1 < nl else t[l, :, 5]).astype(np.float64)
    print("# sum over traced launches, us:", {k: round(float(v), 1) for k, v in tot.items()})


if __name__ == "__main__":
    main()
