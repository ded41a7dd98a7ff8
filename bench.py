#!/usr/bin/env python
"""bench.py — int8 NITI training throughput on B200 (BASELINE.json metric: int8 train samples/sec).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload vgg11|lenet|resnet18] [--impl reference]

A "step" is one full int8 training step (forward, integer loss gradient, dX, dW, shared-exponent
rescale, int8 SGD update) of the hot path over one synthetic batch.  One process per GPU (torchrun for
N>1, NCCL); prints ONE JSON line on rank 0.  `value` times steps whose inputs are already in HBM;
`e2e` times the same step fed from pinned host memory with the logits read back every step.
`--impl reference` times the unmodified reference (oracle/_ref, built from /root/reference by
oracle/Makefile.ref) on the host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (spec key, default per-GPU batch, description)
    "vgg11": ("vgg11", 256, "VGG-11 int8 NITI, CIFAR-shaped 3x32x32 synthetic, batch 256/GPU"),
    "lenet": ("lenet", 64, "MNIST NITIInt8 LeNet-style CNN, 1x28x28 synthetic, batch 64/GPU"),
    "resnet18": ("resnet18", 128, "ResNet-18 int8 contraction chain, 3x224x224 synthetic, batch 128/GPU"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="vgg11", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the workload's)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ref-batch", type=int, default=0, help="override the reference arm's batch (default: same config as our arm; sets config.reduced)")
    ap.add_argument("--ref-budget", type=float, default=170.0, help="seconds of timed reference steps (reference arm)")
    ap.add_argument("--cpu-sample", type=int, default=32, help="samples per step of the bounded cpu_baseline leg of the default run")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: the workload's batch PER GPU; strong: the workload's batch is the GLOBAL batch, split over the GPUs")
    ap.add_argument("--no-dp-check", action="store_true", help="skip the data-parallel correctness witness (world > 1)")
    ap.add_argument("--no-parity-check", action="store_true", help="skip the reference-fixture parity witness (1 GPU, vgg11)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--via", default="cabi", choices=["cabi", "mnn"],
                    help="mnn (workload lenet only): the reference's own training loop on its CPU backend and on the B200 MNN plug-in "
                         "(BASELINE.json configs 1 and 2), timed by the reference's clock in one torch-free subprocess")
    ap.add_argument("--no-graph", action="store_true", help="launch eagerly instead of replaying a CUDA graph")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------------
# CPU reference leg (cpu_baseline and --impl reference).  The only place bench.py touches oracle/.
# ----------------------------------------------------------------------------------------------------
def _walk_weights(workload):
    import numpy as np
    from mandheling_dsp_training_b200.step import SPECS, niti_xavier_int8
    spec_fn, chw = SPECS[WORKLOADS[workload][0]]
    spec = spec_fn()
    wrng = np.random.default_rng(1)
    ws, c = [], chw[0]
    h, w_ = chw[1], chw[2]
    for it in spec:
        if it[0] == "conv":
            ws.append(niti_xavier_int8(wrng, it[1], c, it[2], it[2]))
            c = it[1]; h = (h + 2 * it[4] - it[2]) // it[3] + 1; w_ = (w_ + 2 * it[4] - it[2]) // it[3] + 1
        elif it[0] == "pool":
            h = (h - it[1]) // it[2] + 1; w_ = (w_ - it[1]) // it[2] + 1
        elif it[0] == "flatten":
            c, h, w_ = c * h * w_, 1, 1
    return spec, chw, ws


def reference_step_fn(workload, batch):
    """Returns (step_callable, kind, cores, note).  With oracle/_ref built (kind "reference") a step is one iteration of
    the UNMODIFIED reference's own training loop — its Module forward, _NITI_LOSS_SUM, OpGrad over the whole graph and
    NITI_SGD::step (ref_net_train in oracle/ref_driver.cpp = demo/MnistUtils.cpp:79-147 with injected data and weights) —
    on its CPU backend with every host thread MNN will take.  Otherwise the C oracle port (kind "port", 1 core)."""
    import numpy as np
    from oracle.oracle import Oracle, Reference, shape
    from oracle.step_twin import oracle_step
    spec, chw, ws = _walk_weights(workload)
    rng = np.random.default_rng(0)
    x = rng.integers(-127, 128, size=(batch,) + chw, dtype=np.int8)
    labels = rng.integers(0, 10, size=batch).astype(np.int32)
    if Reference.available():
        cores = min(os.cpu_count() or 1, 32)            # MNN caps its pool at 32 threads (CPUBackend.cpp:29)
        ref = Reference(threads=cores)
        state = {"ws": [(w.copy(), e) for w, e in ws]}
        classes = [it for it in spec if it[0] == "conv"][-1][1]

        def step():
            out = ref.net_train(spec, x[None], -7, labels[None], state["ws"], tclasses=classes)
            state["ws"] = [(w, e) for w, (_, e) in zip(out["weights"], state["ws"])]
        return step, "reference", cores, (f"unmodified reference training loop (Module forward + OpGrad + NITI_SGD, oracle/_ref) "
                                          f"on its CPU backend, batch {batch}, {cores} threads")
    orc = Oracle()
    state = [(w.copy(), e) for w, e in ws]
    return (lambda: oracle_step(orc, spec, x, -7, labels, state, shape)), "port", 1, \
        f"C oracle port, batch {batch}, 1 thread (oracle/_ref not built)"


class _QuietStdout:
    """the reference printf()s debug lines to stdout (e.g. shape/ShapeNITI_*.cpp); keep our stdout to the
    one JSON line by pointing fd 1 at /dev/null while it runs"""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *a):
        os.dup2(self.saved, 1)
        os.close(self.null); os.close(self.saved)


def time_reference(workload, batch, steps, warmup, budget_s=45.0, warm_batch=None):
    """Times up to `steps` reference steps of `batch` samples (stops once budget_s is spent, at least one step).
    Warm-up: `warmup` steps at warm_batch samples (default: the timed batch)."""
    with _QuietStdout():
        step, kind, cores, note = reference_step_fn(workload, batch)
        if warmup > 0:
            wstep = step if not warm_batch or warm_batch == batch else reference_step_fn(workload, warm_batch)[0]
            for _ in range(warmup):
                wstep()
        t0 = time.perf_counter(); done = 0
        while done < steps and (time.perf_counter() - t0 < budget_s or done == 0):
            step(); done += 1
        dt = time.perf_counter() - t0
    return {"value": batch * done / dt, "unit": "samples/s", "cores": cores, "kind": kind,
            "sample": f"{note}; {done} steps in {dt:.1f}s"}, dt / done, done


def run_reference_arm(args):
    """The reference arm: the reference's own CPU implementation on this box's host cores, SAME config as our arm (the
    workload's batch) unless --ref-batch overrides it (then `reduced` is set).  A batch-256 VGG-11 step is about a minute
    of CPU time, so the timed region is as many whole steps as fit `--ref-budget` seconds (steps_done), after `--warmup`
    warm-up steps on 16 samples each (they page the code in; full-size warm-ups would multiply the run time)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = args.ref_batch or (args.batch or WORKLOADS[args.workload][1])
    reduced = batch != (args.batch or WORKLOADS[args.workload][1])
    warm = max(1, args.warmup)
    warm_batch = min(16, batch)
    cb, s_per_step, done = time_reference(args.workload, batch, args.steps, warm, budget_s=args.ref_budget, warm_batch=warm_batch)
    scaling = {}
    for b in (4, 16, 64):                               # how the CPU path scales with batch (threads over the batch)
        if b < batch:
            r, _, _ = time_reference(args.workload, b, 1, 0, budget_s=1.0)
            scaling[str(b)] = round(r["value"], 3)
    scaling[str(batch)] = round(cb["value"], 3)
    world = max(1, args.gpus)
    # `config` carries the same keys, with the same values, as the B200 arm's line for the same flags.  Under --gpus N (weak)
    # the job's global batch is batch x N; the host then times a BOUNDED sample of it: whole steps of `batch` samples (CPU
    # samples/s does not depend on how the global batch is cut into steps, see ref_batch_scaling), as the cpu_baseline says.
    gbatch = batch * world if args.scaling == "weak" else batch
    line = {"impl": "reference", "metric": "int8_train_samples_per_sec", "value": cb["value"], "unit": "samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "steps_done": done, "warmup": warm, "ms_per_step": s_per_step * 1e3,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "int8", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload][2], "per_gpu_batch": batch if args.scaling == "weak" else batch // world,
                       "global_batch": gbatch, "parallelism": f"dp{world}",
                       "notes": {"launch": "reference CPU backend, one host process", "reduced": reduced, "ref_batch_scaling": scaling,
                                 "timed_step_batch": batch, "warmup_step_batch": warm_batch,
                                 "simd": "oracle/_ref is built without MNN_USE_SSE (oracle/Makefile.ref): the NITI GEMM unit is scalar C either way, the float glue ops (Raster, Cast) run their portable C loops",
                                 "sample": "one host runs whole steps of timed_step_batch samples; value does not grow with --gpus"}},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock + throttle reasons sampled through NVML every 20 ms while the benchmark runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False
        self.ok = False
        self.active = False            # set by the main thread around the timed regions
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.max_mhz = None

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((mhz, reasons, self.active))
            except Exception:
                pass
            time.sleep(0.002)

    def summary(self):
        self.stop_flag = True
        nv = getattr(self, "nv", None)
        busy = [s for s in self.samples if s[2]] or self.samples      # samples taken inside the timed regions
        sm = [s[0] for s in busy]
        names = {}
        if nv is not None:
            for key, label in (("nvmlClocksThrottleReasonHwSlowdown", "hw_slowdown"),
                               ("nvmlClocksThrottleReasonHwThermalSlowdown", "hw_thermal_slowdown"),
                               ("nvmlClocksThrottleReasonSwThermalSlowdown", "sw_thermal_slowdown"),
                               ("nvmlClocksThrottleReasonSwPowerCap", "sw_power_cap")):
                if hasattr(nv, key):
                    names[getattr(nv, key)] = label
        reasons = sorted({label for s in busy for bit, label in names.items() if s[1] & bit})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": reasons, "samples": len(busy), "samples_total": len(self.samples)}


def run_via_mnn(args):
    """configs 1/2: MNIST NITIInt8 model, batch 64, the unmodified reference graph with the B200 plug-in as its backend.  The
    plug-in copies every graph input host->device and every output device->host itself, so this is an end-to-end number."""
    import subprocess
    if args.workload != "lenet":
        raise SystemExit("--via mnn measures the reference's MNIST model: use --workload lenet")
    if int(os.environ.get("RANK", "0")) != 0:
        return
    here = os.path.dirname(os.path.abspath(__file__))
    batch = args.batch or WORKLOADS["lenet"][1]
    steps = max(4, min(args.steps, 50))
    p = subprocess.run([sys.executable, os.path.join(here, "tools", "mnn_lenet_bench.py"), str(steps), str(batch)],
                       capture_output=True, text=True, timeout=1200)
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    if p.returncode != 0 or not lines:
        raise SystemExit(f"mnn_lenet_bench failed ({p.returncode}):\n{p.stdout[-1500:]}\n{p.stderr[-1500:]}")
    r = json.loads(lines[-1])
    value = batch / (r["b200_ms_per_step"] * 1e-3)
    print(json.dumps({
        "metric": "int8_train_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": 1, "steps": r["steps"] - 1, "warmup": 1,
        "ms_per_step": r["b200_ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8",
        "data": "synthetic", "via": "mnn",
        "config": {"workload": "MNIST NITIInt8 model (demo/mnistTrain.cpp:131-181), 1x28x28 synthetic, batch %d" % batch,
                   "per_gpu_batch": batch, "global_batch": batch, "parallelism": "dp1",
                   "launch": "reference Executor (express graph rebuilt per step, one Execution per op) on MNN_FORWARD_USER_0"},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": batch * 28 * 28 + batch * 10 * 4,
                "d2h_bytes_per_step": batch * 12 + 4, "note": "host tensors in, loss out, copies made by the plug-in's onCopyBuffer"},
        "gpu_launches": int(r["launches_per_step"] * (r["steps"] - 1)), "launches_per_step": r["launches_per_step"],
        "identical_to_cpu_backend": r["identical"],
        "cpu_baseline": {"value": batch / (r["cpu_ms_per_step"] * 1e-3), "unit": "samples/s", "cores": r["threads"], "kind": "reference",
                         "sample": "the same %d steps of the same model on the reference's CPU backend, same process" % (r["steps"] - 1),
                         "ms_per_step": r["cpu_ms_per_step"]},
        "roofline": None}))


def main():
    args = parse()
    if args.via == "mnn":
        return run_via_mnn(args)
    if args.impl == "reference":
        return run_reference_arm(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from mandheling_dsp_training_b200 import NitiB200
    from mandheling_dsp_training_b200.step import NitiNet, SPECS

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        # the dW all-reduce runs beside the contraction kernels: the library leaves NITI_B200_DP_RESERVE_SMS = 16 SMs free
        # for it (csrc/dp.cu) and NCCL is capped at the same 16 CTAs (never more CTAs than free SMs: 4xB200 with
        # reserve 8 / 16 CTAs collapsed to 51.7 ms per step).  Measured 2xB200, ms/step: (reserve, max CTAs) = (8,8) 0.810,
        # (8,16) 0.727, (16,16) 0.726, (8,32) 0.779, (24,32) 0.817, (4,4) 1.011; reserve 0 starves NCCL (41 ms)
        from mandheling_dsp_training_b200 import lib as _lib
        _lib.dp_env()                                        # NCCL_MAX_CTAS <= NITI_B200_DP_RESERVE_SMS, enforced by dp_connect
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    spec_key, dbatch, desc_text = WORKLOADS[args.workload]
    batch = args.batch or dbatch
    if args.scaling == "strong":                             # the workload's batch is the GLOBAL batch
        if batch % world:
            raise SystemExit(f"--scaling strong: global batch {batch} is not divisible by {world} GPUs")
        batch //= world
    spec_fn, chw = SPECS[spec_key]
    nb = NitiB200(local)
    net = NitiNet(nb, spec_fn(), batch, chw, device=local, dist=dist if world > 1 else None)
    net.init_weights(seed=1)

    # synthetic data: a few distinct pinned host batches (different shard per rank), rotated
    rng = np.random.default_rng(1000 + rank)
    nbuf = 4
    hx = [torch.from_numpy(rng.integers(-127, 128, size=(batch, chw[1], chw[2], chw[0]), dtype=np.int8)).pin_memory() for _ in range(nbuf)]
    hl = [torch.from_numpy(rng.integers(0, 10, size=batch).astype(np.int32)).pin_memory() for _ in range(nbuf)]
    dx = [t.cuda() for t in hx]; dl = [t.cuda() for t in hl]
    # the classifier's rows are padded to 64 channels on the device; only the real classes travel back (gathered on the device)
    dev_logits = torch.empty(net.logits.shape, dtype=torch.int8, device="cuda")
    host_logits = torch.empty(net.logits.shape, dtype=torch.int8).pin_memory()

    def read_back():
        dev_logits.copy_(net.logits)
        host_logits.copy_(dev_logits, non_blocking=True)

    flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda")      # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    use_graph = not args.no_graph
    step_fn = net.step
    net.set_input(dx[0], -7, dl[0])
    l0 = nb.launches
    net.step()                                               # also grows the library workspaces
    torch.cuda.synchronize()
    launches_per_step = nb.launches - l0
    if use_graph:
        step_fn = net.capture()
    net.init_weights(seed=1)

    sampler = ClockSampler(local); sampler.start()
    # ---- warm-up (through the e2e path, so pinned copies and the D2H are warm too)
    for i in range(max(3, args.warmup)):
        net.set_input(hx[i % nbuf], -7, hl[i % nbuf]); step_fn(); read_back()
    barrier()

    # ---- device-resident timed region: K steps, L2 flushed between steps, CUDA events per step
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches0 = nb.launches
    barrier()
    sampler.active = True
    for i in range(args.steps):
        net.set_input(dx[i % nbuf], -7, dl[i % nbuf])          # device-to-device: inputs already in HBM
        flush.zero_()
        ev[i][0].record(); step_fn(); ev[i][1].record()
    barrier()
    sampler.active = False
    step_ms = [a.elapsed_time(b) for a, b in ev]
    net.check()                                              # sticky device-side error flags (peer exchange / grid barrier)
    launches_timed = launches_per_step * args.steps         # graph replays launch the captured kernels
    del launches0
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    value = batch * world * args.steps / (total_ms * 1e-3)

    # ---- per-kernel timing of the contraction kernels, live, same inputs (eager launches + events)
    fam = {}
    pend = []

    def timer(kind, ops, fn):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        pend.append((kind + ":" + nb.last_path, ops, a, b))
    net.timer = timer
    for i in range(min(args.steps, 5)):
        net.set_input(dx[i % nbuf], -7, dl[i % nbuf]); flush.zero_()
        torch.cuda._sleep(6_000_000)        # ~3 ms spin: lets the host run ahead, so the event pairs bracket
        net.step()                          # back-to-back device execution and not host launch gaps
    torch.cuda.synchronize()
    net.timer = None
    for k, ops, a, b in pend:
        f = fam.setdefault(k, [0.0, 0.0, 0]); f[0] += ops; f[1] += a.elapsed_time(b) * 1e-3; f[2] += 1

    # ---- end to end: pinned host input -> H2D -> step -> D2H logits, every step inside the timed region
    barrier()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    sampler.active = True
    t0.record()
    for i in range(args.steps):
        net.set_input(hx[i % nbuf], -7, hl[i % nbuf])
        step_fn()
        read_back()
        torch.cuda.current_stream().synchronize()              # the caller consumes the result every step
    t1.record()
    barrier()
    sampler.active = False
    net.check()
    e2e_ms = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = batch * world * args.steps / (float(e2e_ms.item()) * 1e-3)

    # ---- data-parallel correctness witness: 3 steps from re-initialised weights on a fixed global batch; every rank's
    # weights must be identical, and equal to ONE GPU stepping the whole global batch (rank 0 runs that on a second,
    # unconnected context; layers whose tiles exceed tensor memory at that batch take the two-phase route by themselves)
    dp_check = None
    if world > 1 and not args.no_dp_check:
        crng = np.random.default_rng(4242)
        gx = torch.from_numpy(crng.integers(-127, 128, size=(3, batch * world, chw[1], chw[2], chw[0]), dtype=np.int8))
        gl = torch.from_numpy(crng.integers(0, 10, size=(3, batch * world)).astype(np.int32))
        net.init_weights(seed=1)
        for i in range(3):
            net.set_input(gx[i, rank * batch:(rank + 1) * batch].cuda(), -7, gl[i, rank * batch:(rank + 1) * batch].cuda())
            step_fn()                                        # the captured graph the timed region replays (eager under --no-graph)
        net.check()
        mine = torch.cat([L.w.reshape(-1) for L in net.convs] + [net.e_logits.reshape(-1)])
        allw = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allw, mine)
        same_ranks = all(torch.equal(allw[0], t) for t in allw[1:])
        equals_1gpu = None
        if rank == 0:
            nb1 = NitiB200(local)
            net1 = NitiNet(nb1, spec_fn(), batch * world, chw, device=local)
            net1.init_weights(seed=1)
            for i in range(3):
                net1.set_input(gx[i].cuda(), -7, gl[i].cuda())
                net1.step()
            net1.check()
            one = torch.cat([L.w.reshape(-1) for L in net1.convs] + [net1.e_logits.reshape(-1)])
            equals_1gpu = bool(torch.equal(one, mine)) and bool(torch.equal(net1.logits_buf[:batch], net.logits_buf))
            del net1, nb1
        dp_check = {"weights_equal_across_ranks": bool(same_ranks), "equals_1gpu_global_batch": equals_1gpu, "steps": 3,
                    "global_batch": batch * world, "launch": "cuda_graph replay on every rank" if use_graph else "eager"}
        dist.barrier()

    # ---- parity witness on the benchmarked configuration: the same net, stepped through a captured graph from seeded inputs,
    # against the logits / exponents / updated weights the UNMODIFIED reference produced for those inputs (committed
    # fixture; regenerate with tests/golden/make_golden_vgg.py where /root/reference exists)
    parity_check = None
    if world == 1 and args.workload == "vgg11" and batch in (4, 256) and not args.no_parity_check:
        try:
            from tests.gpu_util import vgg_reference_case
            nbp = NitiB200(local)                            # own context: its workspaces must not move under the timed graph
            parity_check = vgg_reference_case(nbp, batch, replay=use_graph)
            nbp.close()
        except Exception as ex:      # pragma: no cover
            parity_check = {"error": str(ex)}

    clocks = sampler.summary()         # sampled over warm-up + all timed regions
    if world > 1:
        dist.barrier(); torch.cuda.synchronize()
    if rank != 0:
        # captured NCCL work keeps the communicator busy: tearing it down can block, and nothing is left to flush
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)

    # ---- roofline of the dominant kernel family (tcgen05 contractions)
    peak_tops, peak_src = None, "unmeasured"
    try:
        peak_tops = nb.probe_i8_peak(4000); peak_src = "measured live: tcgen05.mma.kind::i8 128x256x32 issue-rate probe (niti_probe_i8_peak), all SMs"
    except Exception as ex:          # pragma: no cover
        peak_src = f"probe failed: {ex}"
    is_tc = lambda k: k.split(":")[1] in ("tc", "tma", "tma2", "fused", "stem", "chain")     # all of these are tcgen05 kernels
    tc_ops = sum(v[0] for k, v in fam.items() if is_tc(k))
    tc_s = sum(v[1] for k, v in fam.items() if is_tc(k))
    tc_n = sum(v[2] for k, v in fam.items() if is_tc(k))
    all_s = sum(v[1] for v in fam.values())
    achieved = tc_ops / tc_s / 1e12 if tc_s > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(args.workload)
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": "chain_kernel (layer-walking fwd chain and dX) / dw2_kernel / stem2_fwd_kernel (TMA im2col + tcgen05 kind::i8 implicit GEMM). fwd and dX launches are whole ops: contraction + max|acc| + grid barrier + rescale to int8 (+ ReLU, pool / pool-grad) in the timed launch",
                "achieved": achieved, "peak": peak_tops, "unit": "TOP/s",
                "frac": (achieved / peak_tops) if peak_tops else None, "traffic": traffic,
                "peak_source": peak_src, "launches_timed": tc_n,
                "share_of_contraction_time": (tc_s / all_s) if all_s else None,
                "families": {k: {"TOP/s": (v[0] / v[1] / 1e12) if v[1] else None, "ms": v[1] * 1e3 / max(1, min(args.steps, 5)),
                                 "launches": v[2] // max(1, min(args.steps, 5))} for k, v in sorted(fam.items())}}

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            cpu_baseline, _, _ = time_reference(args.workload, args.cpu_sample, 2, 1, budget_s=20.0, warm_batch=min(8, args.cpu_sample))
            cpu_baseline["sample"] += f" (bounded sample: {args.cpu_sample}-sample steps of the same net; the same-config number is the --impl reference arm)"
        except Exception as ex:      # pragma: no cover
            cpu_baseline = {"value": None, "unit": "samples/s", "cores": 0, "kind": "unavailable", "sample": str(ex)}

    ws_mb = (sum(L.y.numel() * 2 for L in net.ops) + net.acc.numel() * 4 + net.dw_acc.numel() * 4) / 2 ** 20
    line = {
        "metric": "int8_train_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "int8", "data": "synthetic",
        "config": {"workload": desc_text, "per_gpu_batch": batch, "global_batch": batch * world, "parallelism": f"dp{world}",
                   "notes": {"launch": "cuda_graph" if use_graph else "eager",
                             "dp_exchange": (("max|acc|: in-kernel NVLink peer-memory exchange; dW: bucketed " +
                                              ("two-shot sum over NVLink peer memory (niti_dp_allreduce_i32)" if getattr(net, "peer_sum", False)
                                               else "NCCL all-reduce(SUM)") + " overlapped with backward")
                                             if net.peer_max else ("NCCL" if world > 1 else "none")),
                             "l2": "flushed between timed steps (256 MiB write)", "step_working_set_mib": round(ws_mb, 1),
                             "algorithmic_gop_per_step_per_gpu": net.algorithmic_ops() / 1e9}},
        "e2e": {"value": e2e_value, "unit": "samples/s",
                "h2d_bytes_per_step": int(hx[0].numel() + hl[0].numel() * 4) * world,
                "d2h_bytes_per_step": int(host_logits.numel()) * world},
        "gpu_launches": int(launches_timed),
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "dp_check": dp_check,
        "parity_check": parity_check,
        "tensor_frac_of_step": (net.algorithmic_ops() / (total_ms / args.steps * 1e-3) / 1e12 / peak_tops) if peak_tops else None,
        # fused kernels store with the exponent predicted from the previous step and re-store on a miss (conv_chain.cu);
        # counted over the whole run (warm-up, timed, e2e and per-kernel passes): the stored values never depend on it
        "exponent_prediction": dict(zip(("hits", "misses", "not_speculated"), nb.speculate_stats())) if hasattr(nb, "speculate_stats") else None,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
