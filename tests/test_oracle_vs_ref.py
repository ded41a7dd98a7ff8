"""CPU, build container only: the oracle against the unmodified reference run live
(oracle/_ref, built from /root/reference by oracle/Makefile.ref).  Skipped where the reference build is
absent (it travels to the GPU box, but these are `not gpu` tests run here)."""
import numpy as np
import pytest
from oracle.oracle import Reference, shape

pytestmark = pytest.mark.skipif(not Reference.available(), reason="oracle/_ref not built")


@pytest.fixture(scope="module")
def ref():
    return Reference(threads=4)


@pytest.mark.parametrize("cfg", [
    (4, 4, 8, 8, 8, 3, 1, 1), (2, 8, 6, 6, 12, 3, 1, 0), (2, 20, 12, 12, 52, 5, 1, 0),
    (8, 16, 1, 1, 12, 1, 1, 0), (4, 8, 10, 10, 8, 3, 2, 1), (3, 12, 7, 7, 16, 3, 1, 1)])
# (non-square images are outside the reference's contract: its grad rule swaps kernelX/kernelY,
#  grad/NITI_Conv_Int8_Grad.cpp:172-173, and the run crashes; the oracle and the CUDA path handle them.)
@pytest.mark.parametrize("seed", [0, 1])
def test_live_reference_conv(oracle, ref, cfg, seed):
    n, c, h, w, oc, k, st, p = cfg
    s = shape(n, c, h, w, oc, k, stride=st, pad=p)
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, size=(n, c, h, w), dtype=np.int8)
    wt = rng.integers(-64, 65, size=(oc, c, k, k), dtype=np.int8)
    dy = rng.integers(-40, 41, size=(n, oc, s.oh, s.ow), dtype=np.int8)
    y, e, _ = oracle.conv_fwd(s, x, wt, -5, -9)
    ry, re = ref.conv_fwd(s, x, wt, -5, -9)
    assert np.array_equal(y, ry) and e == re
    rdx, rdw = ref.conv_bwd(s, x, wt, dy)
    assert np.array_equal(oracle.conv_dx(s, dy, wt)[0], rdx)
    dw = oracle.conv_dw(s, x, dy)[0]
    if st == 1:
        assert np.array_equal(dw, rdw)
    else:
        # quirk Q9: for stride 2 the reference's grad rule applies a raw _Transpose to an NC4HW4
        # tensor (grad/NITI_Conv_Int8_Grad.cpp:164-166), which permutes the dW entries; the values
        # (accumulate + shared exponent + PSTO) are identical, so compare them as a multiset.
        assert np.array_equal(np.sort(dw.ravel()), np.sort(rdw.ravel()))


def _lenet_case(steps, n, seed=0):
    from mandheling_dsp_training_b200.step import lenet_spec, niti_xavier_int8
    rng = np.random.default_rng(seed + 1)
    ws = [niti_xavier_int8(rng, oc, ic, k, k) for oc, ic, k in [(20, 1, 5), (52, 20, 5), (500, 832, 1), (12, 500, 1)]]
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, size=(steps, n, 1, 28, 28), dtype=np.int8)
    lab = rng.integers(0, 10, size=(steps, n)).astype(np.int32)
    return lenet_spec(), ws, x, lab


def test_live_reference_training_steps(oracle, ref):
    """Whole training steps of the reference's NITIInt8 MNIST model (demo/mnistTrain.cpp:131-182) through its own
    Module / OpGrad / NITI_SGD code with injected weights, against oracle_step: logits, exponents and every
    updated weight tensor identical after 3 steps (the full-loop pin of SURVEY.md section 8c(2))."""
    from oracle.step_twin import oracle_step
    spec, ws, x, lab = _lenet_case(3, 8)
    out = ref.net_train(spec, x, -7, lab, ws, tclasses=10)
    w2 = [(w.copy(), e) for w, e in ws]
    for s in range(3):
        lg, e = oracle_step(oracle, spec, x[s], -7, lab[s], w2, shape)
        assert np.array_equal(lg.reshape(8, -1), out["logits"][s])
        assert e == out["e_logits"][s]
    for a, (b, _) in zip(out["weights"], w2):
        assert np.array_equal(a, b)
    assert np.all(np.isfinite(out["loss"]))


@pytest.mark.parametrize("dims", [(36, 64, 8), (52, 100, 16), (25 * 4, 36, 20)])
def test_live_reference_matmul(oracle, ref, dims):
    """OpType_NITI_MatMul_Int8 as NITI_GeometryConv2DBackPropFilter_Int8.cpp:78-126 emits it (transposeB), created with
    Expr::create in ref_driver.cpp; A / C are addressed as [channel/4][batch][4] blocks by the execution itself."""
    e, l, h = dims
    rng = np.random.default_rng(e)
    a_raw = rng.integers(-127, 128, size=e * l, dtype=np.int8)
    b = rng.integers(-90, 91, size=(h, l), dtype=np.int8)
    c, _ = oracle.matmul_nt(oracle.matmul_unpack_a(a_raw, e, l), b)
    assert np.array_equal(oracle.matmul_pack_c(c, e, h), ref.matmul_raw(a_raw, b, e, l, h))


@pytest.mark.parametrize("cfg", [(4, 8, 16, 48, 3, 1), (2, 4, 20, 48, 3, 1), (2, 8, 12, 24, 3, 1)])
def test_live_reference_gradient_split_conv(oracle, ref, cfg):
    n, c, hh, oc, k, pad = cfg
    s = shape(n, c, hh, hh, oc, k, stride=1, pad=pad)
    rng = np.random.default_rng(oc + hh)
    x = rng.integers(-127, 128, size=(n, c, hh, hh), dtype=np.int8)
    dy = rng.integers(-50, 51, size=(n, oc, s.oh, s.ow), dtype=np.int8)
    dy[:, oc // 2:] //= 4
    dw, slices = oracle.conv_dw_split(s, x, dy)
    assert np.array_equal(dw, ref.grad_conv(s, x, dy, split=True))
    assert np.array_equal(oracle.conv_dw(s, x, dy)[0], ref.grad_conv(s, x, dy, split=False))


@pytest.mark.parametrize("shp,k,s", [((4, 8, 8, 8), 2, 2), ((2, 16, 6, 6), 2, 2), ((3, 12, 9, 9), 3, 2), ((2, 20, 24, 24), 2, 2)])
def test_live_reference_step_glue_ops(oracle, ref, shp, k, s):
    """SURVEY.md section 8 f1/f2 against the unmodified reference: ReLU / ReLU-grad, max pool and its gradient (ties and
    overlapping windows: the first maximum in row-major order takes the gradient, contributions wrap in int8), the int8
    Binary SUB of the SGD update (clip to +-127) and the integer softmax-cross-entropy gradient."""
    rng = np.random.default_rng(shp[1] * 100 + k)
    x = rng.integers(-3, 4, size=shp, dtype=np.int8)
    oh = (shp[2] - k) // s + 1
    dy = rng.integers(-127, 128, size=(shp[0], shp[1], oh, oh), dtype=np.int8)
    g = rng.integers(-127, 128, size=shp, dtype=np.int8)
    assert np.array_equal(ref.relu(x), oracle.relu(x))
    assert np.array_equal(ref.relu_grad(x, g), oracle.relu_grad(x, g))
    y = oracle.maxpool(x, k, s)
    assert np.array_equal(ref.maxpool(x, k, s), y)
    assert np.array_equal(ref.maxpool_grad(x, dy, k, s), oracle.maxpool_grad(x, y, dy, k, s))
    w = rng.integers(-127, 128, size=4096, dtype=np.int8); dw = rng.integers(-127, 128, size=4096, dtype=np.int8)
    assert np.array_equal(ref.sgd_update(w, dw), oracle.sgd_update(w, dw))
    for asc in (-9, -3, 1):
        lg = rng.integers(-127, 128, size=(32, 12), dtype=np.int8)
        lab = rng.integers(0, 10, size=32)
        assert np.array_equal(ref.loss_grad(lg, asc, lab), oracle.loss_grad(lg, asc, lab))
        # the float loss scalar (NITI_CPULoss_Int8): same operations in the same order -> 1e-6 relative
        assert oracle.loss_value(lg, asc, lab) == pytest.approx(ref.loss_value(lg, asc, lab), rel=1e-6)
