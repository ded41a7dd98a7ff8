"""BASELINE.json configs 1 and 2 in one process: the reference's OWN training loop (express graph, OpGrad rules, NITI_SGD —
the unmodified libMNN built into oracle/_ref) on the MNIST NITIInt8 model at batch 64, first on the reference's CPU backend,
then on the B200 plug-in (MNN_FORWARD_USER_0).  Prints one JSON object: ms/step of both, identical flag, CUDA launches.
Torch-free on purpose (the reference library must be loaded RTLD_GLOBAL before anything that brings its own libstdc++ users).
usage: python tools/mnn_lenet_bench.py [steps] [batch] [threads]"""
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.oracle import Reference            # noqa: E402

PLUGIN = os.path.join(ROOT, "mnn_backend", "libMNN_niti_b200.so")
LENET = [("conv", 20, 5, 1, 0), ("relu",), ("pool", 2, 2), ("conv", 52, 5, 1, 0), ("relu",), ("pool", 2, 2),
         ("flatten",), ("conv", 500, 1, 1, 0), ("relu",), ("conv", 12, 1, 1, 0)]      # demo/mnistTrain.cpp:131-181


def xavier_int8(rng, oc, c, k):
    w = rng.normal(0.0, np.sqrt(2.0 / ((c + oc) * k * k)), size=(oc, c, k, k)).astype(np.float32)
    r = float(np.abs(w).max())
    return np.round(w / r * 127.0).astype(np.int8), int(np.ceil(np.log2(r))) - 7


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    threads = int(sys.argv[3]) if len(sys.argv) > 3 else (os.cpu_count() or 4)
    rng = np.random.default_rng(1)
    ws = [xavier_int8(rng, oc, ic, k) for oc, ic, k in [(20, 1, 5), (52, 20, 5), (500, 832, 1), (12, 500, 1)]]
    rng = np.random.default_rng(0)
    x = rng.integers(-127, 128, size=(steps, batch, 1, 28, 28), dtype=np.int8)
    lab = rng.integers(0, 10, size=(steps, batch)).astype(np.int32)
    ref = Reference(threads=threads)
    cpu = ref.net_train(LENET, x, -7, lab, ws, tclasses=10)
    ref.use_backend(8, PLUGIN, threads=threads)                      # MNN_FORWARD_USER_0
    plug = ctypes.CDLL(PLUGIN)
    plug.MNNNitiB200LaunchCount.restype = ctypes.c_longlong
    before = plug.MNNNitiB200LaunchCount()
    gpu = ref.net_train(LENET, x, -7, lab, ws, tclasses=10)
    launches = int(plug.MNNNitiB200LaunchCount() - before)
    same = (np.array_equal(cpu["logits"], gpu["logits"]) and np.array_equal(cpu["e_logits"], gpu["e_logits"])
            and all(np.array_equal(a, b) for a, b in zip(cpu["weights"], gpu["weights"])))
    skip = 1 if steps > 2 else 0                                       # the first step builds the graph and the Executions
    print(json.dumps({"steps": steps, "batch": batch, "threads": threads,
                      "cpu_ms_per_step": float(np.mean(cpu["ms"][skip:])), "b200_ms_per_step": float(np.mean(gpu["ms"][skip:])),
                      "cpu_first_step_ms": float(cpu["ms"][0]), "b200_first_step_ms": float(gpu["ms"][0]),
                      "identical": bool(same), "launches_per_step": launches / steps}))


if __name__ == "__main__":
    main()
