"""GPU: the reference's OWN graph (express API, OpGrad rules, Module, NITI_SGD — the unmodified libMNN built
into oracle/_ref) executed on the B200 plug-in (mnn_backend/libMNN_niti_b200.so, forward type
MNN_FORWARD_USER_0 = 8) must give byte-identical results to the same graph on the reference's CPU backend.
This is BASELINE.json config 2: "Same MNIST int8 CNN on 1 B200 via new CUDA backend, bit-exact vs CPU".

Order matters inside the one process: every case is first evaluated on the reference CPU backend, then the
global executor is switched (ref_use_backend) and the case is evaluated again through the plug-in."""
import os
import numpy as np
import pytest

from oracle.oracle import Reference, shape

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "mnn_backend", "libMNN_niti_b200.so")
MNN_FORWARD_USER_0 = 8          # include/MNN/MNNForwardType.h:38

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not Reference.available() or not os.path.exists(PLUGIN),
                                 reason="oracle/_ref or the MNN plug-in is not built")]

CONV_CASES = [(4, 4, 8, 8, 8, 3, 1, 1), (2, 8, 6, 6, 12, 3, 1, 0), (2, 20, 12, 12, 52, 5, 1, 0),
              (8, 16, 1, 1, 12, 1, 1, 0), (3, 12, 7, 7, 16, 3, 1, 1), (16, 64, 16, 16, 128, 3, 1, 1)]


def _conv_inputs(cfg, seed):
    n, c, h, w, oc, k, st, p = cfg
    s = shape(n, c, h, w, oc, k, stride=st, pad=p)
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, size=(n, c, h, w), dtype=np.int8)
    wt = rng.integers(-64, 65, size=(oc, c, k, k), dtype=np.int8)
    dy = rng.integers(-40, 41, size=(n, oc, s.oh, s.ow), dtype=np.int8)
    return s, x, wt, dy


def _lenet_case(steps, n, seed=0):
    from mandheling_dsp_training_b200.step import lenet_spec, niti_xavier_int8
    rng = np.random.default_rng(seed + 1)
    ws = [niti_xavier_int8(rng, oc, ic, k, k) for oc, ic, k in [(20, 1, 5), (52, 20, 5), (500, 832, 1), (12, 500, 1)]]
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, size=(steps, n, 1, 28, 28), dtype=np.int8)
    lab = rng.integers(0, 10, size=(steps, n)).astype(np.int32)
    return lenet_spec(), ws, x, lab


def _evaluate(ref):
    out = {}
    for i, cfg in enumerate(CONV_CASES):
        s, x, wt, dy = _conv_inputs(cfg, i)
        out[f"fwd{i}"] = ref.conv_fwd(s, x, wt, -5, -9)
        out[f"bwd{i}"] = ref.conv_bwd(s, x, wt, dy)
    rng = np.random.default_rng(7)
    a = rng.integers(-127, 128, size=(4, 8, 8, 8), dtype=np.int8)
    g = rng.integers(-127, 128, size=(4, 8, 8, 8), dtype=np.int8)
    out["relu"] = ref.relu(a)
    out["relu_grad"] = ref.relu_grad(a, g)
    out["maxpool"] = ref.maxpool(a, 2, 2)
    spec, ws, x, lab = _lenet_case(3, 64)
    out["lenet"] = ref.net_train(spec, x, -7, lab, ws, tclasses=10)
    return out


@pytest.fixture(scope="module")
def both():
    ref = Reference(threads=4)
    cpu = _evaluate(ref)
    ref.use_backend(MNN_FORWARD_USER_0, PLUGIN)
    import ctypes
    plug = ctypes.CDLL(PLUGIN)
    before = plug.MNNNitiB200LaunchCount() if hasattr(plug, "MNNNitiB200LaunchCount") else None
    gpu = _evaluate(ref)
    after = plug.MNNNitiB200LaunchCount() if before is not None else None
    return cpu, gpu, (before, after)


def test_plugin_really_ran(both):
    _, _, (before, after) = both
    assert before is not None, "plug-in does not export MNNNitiB200LaunchCount"
    assert after - before > 100, f"only {after - before} CUDA launches: the ops did not run on the B200 backend"


@pytest.mark.parametrize("i", range(len(CONV_CASES)))
def test_conv_ops_identical(both, i):
    cpu, gpu, _ = both
    y0, e0 = cpu[f"fwd{i}"]; y1, e1 = gpu[f"fwd{i}"]
    assert e0 == e1 and np.array_equal(y0, y1)
    dx0, dw0 = cpu[f"bwd{i}"]; dx1, dw1 = gpu[f"bwd{i}"]
    if dx0 is not None:
        assert np.array_equal(dx0, dx1)
    assert np.array_equal(dw0, dw1)


def test_elementwise_ops_identical(both):
    cpu, gpu, _ = both
    for k in ("relu", "relu_grad", "maxpool"):
        assert np.array_equal(cpu[k], gpu[k]), k


def test_lenet_training_steps_identical(both):
    """3 steps of the MNIST NITIInt8 model at the demo's batch size 64: logits, exponents, losses and all updated
    weights equal between the reference CPU backend and the B200 backend."""
    cpu, gpu, _ = both
    a, b = cpu["lenet"], gpu["lenet"]
    assert np.array_equal(a["logits"], b["logits"])
    assert np.array_equal(a["e_logits"], b["e_logits"])
    assert np.array_equal(a["loss"], b["loss"])
    for w0, w1 in zip(a["weights"], b["weights"]):
        assert np.array_equal(w0, w1)
    print("ms/step cpu", a["ms"], "b200 plug-in", b["ms"])
