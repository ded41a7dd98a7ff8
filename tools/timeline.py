"""Launch timeline of one training step UNDER CUDA-GRAPH REPLAY (niti_debug_timeline): for every instrumented launch the
time of its first CTA entry and last CTA exit (%globaltimer), relative to the first entry of the step.  This is the picture
nsys would give (not installed here): what overlaps what, and where the device idles between launches.

usage: python tools/timeline.py [vgg11|lenet|resnet18] [batch] > profiles/timeline_rN_<workload>.txt
(under torchrun: every rank captures; rank 0 prints)"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from mandheling_dsp_training_b200 import NitiB200, lib
from mandheling_dsp_training_b200.step import NitiNet, SPECS


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "vgg11"
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        lib.dp_env()
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    spec_fn, chw = SPECS[wl]
    nb = NitiB200(local)
    net = NitiNet(nb, spec_fn(), batch, chw, device=local, dist=dist)
    net.init_weights(seed=1)
    rng = np.random.default_rng(rank)
    x = torch.from_numpy(rng.integers(-127, 128, size=(batch, chw[1], chw[2], chw[0]), dtype=np.int8)).cuda()
    lab = torch.from_numpy(rng.integers(0, 10, size=batch).astype(np.int32)).cuda()
    net.set_input(x, -7, lab)
    net.step(); torch.cuda.synchronize()
    cap = 512
    tl = torch.zeros(cap * 2, dtype=torch.int64, device="cuda")
    nb.debug_timeline(tl, cap)
    replay = net.capture()               # warm-up step (records 0..n-1) then the captured step (records n..2n-1)
    names = nb.timeline_names()
    nb.debug_timeline(None, 0)
    n = len(names) // 2
    flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda")
    spans = []
    for rep in range(5):
        flush.zero_()
        t = tl.view(cap, 2)
        t[:, 0] = torch.iinfo(torch.int64).max; t[:, 1] = 0
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        replay()
        torch.cuda.synchronize()
        spans.append(tl.view(cap, 2)[n:2 * n].cpu().numpy().copy())
    net.check()
    if rank != 0:
        if world > 1:
            dist.barrier()
        os._exit(0)
    t = spans[-1].astype(np.float64)
    t0 = t[:, 0].min()
    order = np.argsort(t[:, 0])
    print(f"# launch timeline, {wl} batch {batch}/GPU, world {world}, CUDA-graph replay #5 (L2 flushed before), rank 0")
    print(f"# step span (first entry .. last exit of instrumented launches): {(t[:, 1].max() - t0) / 1e3:.1f} us; "
          f"sum of launch durations {((t[:, 1] - t[:, 0]).sum()) / 1e3:.1f} us; {n} instrumented launches")
    print("#  start us |   end us |  dur us | gap to previous end on the critical chain | kernel")
    prev_end = None
    busy_until = 0.0
    for i in order:
        s_, e_ = (t[i, 0] - t0) / 1e3, (t[i, 1] - t0) / 1e3
        gap = s_ - busy_until if busy_until > 0 else 0.0
        print(f"  {s_:9.1f} | {e_:8.1f} | {e_ - s_:7.1f} | {gap:7.1f} | {names[n + i]}")
        busy_until = max(busy_until, e_)
    durs = {}
    for i in range(n):
        durs.setdefault(names[n + i], []).append((t[i, 1] - t[i, 0]) / 1e3)
    print("# per kernel: launches, total us, mean us")
    for k, v in sorted(durs.items(), key=lambda kv: -sum(kv[1])):
        print(f"#   {k:16s} {len(v):3d} {sum(v):8.1f} {sum(v) / len(v):7.1f}")
    allspans = [(s[:, 1].max() - s[:, 0].min()) / 1e3 for s in spans]
    print("# step span of the 5 replays, us:", [round(float(v), 1) for v in allspans])
    if world > 1:
        dist.barrier()
        os._exit(0)


if __name__ == "__main__":
    main()
