"""Per-layer phase breakdown of the fused forward / dX launches of one training step (niti_debug_trace).

usage: python tools/phase_trace.py [vgg11|lenet|resnet18] [batch] > profiles/phase_rN_<workload>.txt
Each fused launch writes %globaltimer stamps per CTA: 0 entry, 1 set-up done (barrier init, TMEM alloc, tensor-map prefetch,
cluster sync, griddepcontrol.wait), 2 first tile accumulated (K loop of the first resident tile), 3 max pass done (TMEM sweep 1),
4 grid barrier passed, 5 rescale + store done (TMEM sweep 2), 6 exit.  Printed per launch: median over CTAs of each phase's
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
    trace = torch.zeros(160 * 8, dtype=torch.int64, device="cuda")
    nb.debug_trace(trace)
    rows = []
    convs = iter([])

    def timer(kind, ops, fn):
        trace.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        t = trace.cpu().numpy().reshape(-1, 8)
        t = t[t[:, 0] > 0]
        rows.append((kind, nb.last_path, ops, a.elapsed_time(b) * 1e3, t.copy()))
    net.timer = timer
    net.step()
    torch.cuda.synchronize()
    nb.debug_trace(None)
    names = ["set-up", "first tile (K loop)", "max pass", "grid barrier", "rescale+store", "exit"]
    print(f"# phase trace, {wl} batch {batch}, one eager step (every launch synchronised: cold pipeline, no PDL overlap)")
    print(f"# columns: kind path GOP | event us | CTAs | first-entry..last-exit us | median us per phase: " + " | ".join(names))
    tot = {}
    for kind, path, ops, ev_us, t in rows:
        if len(t) == 0:
            print(f"{kind:4s} {path:6s} {ops / 1e9:7.2f} | {ev_us:7.1f} | (no trace: kernel family without phase stamps)")
            continue
        span = (t[:, 6].max() - t[:, 0].min()) / 1e3
        d = np.diff(t[:, :7], axis=1) / 1e3
        med = np.median(d, axis=0)
        print(f"{kind:4s} {path:6s} {ops / 1e9:7.2f} | {ev_us:7.1f} | {len(t):3d} | {span:7.1f} | " + " | ".join(f"{v:6.1f}" for v in med))
        for n_, v in zip(names, med):
            tot[n_] = tot.get(n_, 0.0) + float(v)
        tot["launch overhead (event - span)"] = tot.get("launch overhead (event - span)", 0.0) + max(0.0, ev_us - span)
    print("# sum over traced launches, us:", {k: round(v, 1) for k, v in tot.items()})


if __name__ == "__main__":
    main()
