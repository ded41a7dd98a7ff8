"""Generates tests/golden/vgg11_ref_steps.npz: whole training steps of VGG-11 (3x32x32, 12 logits) run by the UNMODIFIED
reference (oracle/_ref: its Module forward, OpGrad and NITI_SGD on its CPU backend, driven by oracle/ref_driver.cpp
ref_net_train) at batch 4 and at batch 256 — the configuration BASELINE.json's metric is quoted on.  Inputs and initial
weights are NOT stored: `case_inputs` below rebuilds them from seeds, and the GPU tests / bench.py import it.

Stored per batch size b: logits [steps, b, 12] int8, e_logits [steps] int8, loss [steps] float32 and, per conv layer, the
SHA-256 of the updated OIHW int8 weights after the last step (9.75 M weights would not fit a fixture).

Run in the build container (needs /root/reference compiled by oracle/Makefile.ref):  python tests/golden/make_golden_vgg.py
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

STEPS = 2
BATCHES = (4, 256)
CLASSES = 12
CHW = (3, 32, 32)
OUT = os.path.join(ROOT, "tests", "golden", "vgg11_ref_steps.npz")


def case_inputs(batch, steps=STEPS):
    """(spec, x [steps, batch, 3, 32, 32] int8 NCHW, labels [steps, batch] int32, weights [(OIHW int8, exponent)]).
    Weights = NitiNet.init_weights(seed=1) (the reference's int8 Xavier recipe, one generator walked layer by layer)."""
    from mandheling_dsp_training_b200.step import vgg11_spec, niti_xavier_int8
    spec = vgg11_spec(CLASSES)
    rng = np.random.default_rng(7000 + batch)
    x = rng.integers(-127, 128, size=(steps, batch) + CHW, dtype=np.int8)
    labels = rng.integers(0, 10, size=(steps, batch)).astype(np.int32)
    wrng = np.random.default_rng(1)
    ws, c, h = [], CHW[0], CHW[1]
    for it in spec:
        if it[0] == "conv":
            ws.append(niti_xavier_int8(wrng, it[1], c, it[2], it[2]))
            c = it[1]; h = (h + 2 * it[4] - it[2]) // it[3] + 1
        elif it[0] == "pool":
            h = (h - it[1]) // it[2] + 1
        elif it[0] == "flatten":
            c, h = c * h * h, 1
    return spec, x, labels, ws


def weight_digest(w_oihw):
    return hashlib.sha256(np.ascontiguousarray(w_oihw, dtype=np.int8).tobytes()).hexdigest()


def main():
    from oracle.oracle import Reference
    ref = Reference(threads=min(os.cpu_count() or 1, 32))
    out = {}
    for b in BATCHES:
        spec, x, labels, ws = case_inputs(b)
        r = ref.net_train(spec, x, -7, labels, ws, tclasses=CLASSES)
        out[f"b{b}.logits"] = r["logits"]
        out[f"b{b}.e_logits"] = r["e_logits"]
        out[f"b{b}.loss"] = r["loss"]
        out[f"b{b}.w_sha256"] = np.asarray([weight_digest(w) for w in r["weights"]])
        print(b, r["e_logits"], r["loss"], r["ms"], flush=True)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
