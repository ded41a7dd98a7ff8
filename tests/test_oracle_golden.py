"""CPU: the oracle (oracle/niti_oracle.c) against the golden vectors produced by the unmodified
reference (tests/golden/make_golden.py).  This is the pin that makes the oracle trustworthy."""
import numpy as np
import pytest
from oracle.oracle import ConvShape


def conv_names(g):
    return sorted({k.split(".")[0] for k in g.files})


def load_case(g, name):
    s = ConvShape(*[int(v) for v in g[name + ".shape"]])
    return s, g[name + ".x"], g[name + ".w"], g[name + ".dy"]


def test_golden_has_cases(golden_conv):
    assert len(conv_names(golden_conv)) >= 10


@pytest.mark.parametrize("name", [
    "mnist_conv1", "mnist_conv2", "mnist_ip1", "mnist_ip2", "vgg_3x3_p1", "vgg_3x3_p1_b",
    "odd_hw_3x3_p1", "s2_3x3_p1", "s2_1x1_p0", "small_mag_cast", "small_mag_shift1", "zero_dy"])
def test_conv_ops_match_reference(oracle, golden_conv, name):
    g = golden_conv
    s, x, w, dy = load_case(g, name)
    y, e, acc = oracle.conv_fwd(s, x, w, -7, -7)
    assert np.array_equal(y, g[name + ".y"])
    assert e == int(g[name + ".exp"])
    dw, _ = oracle.conv_dw(s, x, dy)
    assert np.array_equal(dw, g[name + ".dw"])
    if name + ".dx" in g.files:
        dx, _ = oracle.conv_dx(s, dy, w)
        assert np.array_equal(dx, g[name + ".dx"])


def test_rescale_edge_paths_are_covered(oracle, golden_conv):
    """the golden set exercises all three branches of the forward shift rule and the b==0 dW rule"""
    seen = set()
    for name in conv_names(golden_conv):
        s, x, w, dy = load_case(golden_conv, name)
        acc, _ = oracle.conv_fwd_acc(s, x, w)
        sh = oracle.range_estimate(acc) - 7
        seen.add("gt1" if sh > 1 else ("eq1" if sh == 1 else "le0"))
    assert seen == {"gt1", "eq1", "le0"}
    s, x, w, dy = load_case(golden_conv, "zero_dy")
    assert not oracle.conv_dw(s, x, dy)[0].any()


def test_elementwise_match_reference(oracle, golden_ew):
    g = golden_ew
    x = g["relu.x"]
    assert np.array_equal(oracle.relu(x), g["relu.y"])
    assert np.array_equal(oracle.relu_grad(x, g["relu_grad.dy"]), g["relu_grad.dx"])
    assert np.array_equal(oracle.maxpool(x, 2, 2), g["maxpool.y"])
    assert np.array_equal(oracle.sgd_update(g["sgd.w"], g["sgd.dw"]), g["sgd.out"])
    for a in (-9, -7, -5, -3, 0, 1):
        got = oracle.loss_grad(g[f"loss_grad.{a}.logits"], a, g[f"loss_grad.{a}.labels"])
        assert np.array_equal(got, g[f"loss_grad.{a}.grad"]), a
        # float loss scalar of NITI_CPULoss_Int8: tolerance 1e-6 relative (same operations, same order)
        assert oracle.loss_value(g[f"loss_grad.{a}.logits"], a, g[f"loss_grad.{a}.labels"]) == pytest.approx(float(g[f"loss_value.{a}"]), rel=1e-6)


def test_psto_properties(oracle):
    rng = np.random.default_rng(3)
    acc = rng.integers(-(2 ** 20), 2 ** 20, size=10000).astype(np.int32)
    for shift in range(2, 16):
        out = oracle.psto(acc, shift)
        assert out.min() >= -127 and out.max() <= 127            # Q4 symmetric clip
        q = np.trunc(acc / float(1 << shift)).astype(np.int64)
        d = np.clip(out, -126, 126) - np.clip(q, -126, 126)
        assert np.all(np.abs(d) <= 1)                             # rounds by at most one step
        assert np.array_equal(oracle.psto(-acc, shift), -out)     # odd symmetry
    # Q2: power-of-two maximum -> bitwidth k, non power-of-two -> next k
    assert oracle.range_estimate(np.array([1], np.int32)) == 0
    assert oracle.range_estimate(np.array([-128], np.int32)) == 7
    assert oracle.range_estimate(np.array([129], np.int32)) == 8
    assert oracle.range_estimate(np.zeros(4, np.int32)) == 0


def test_ragged_and_empty(oracle):
    assert oracle.relu(np.zeros((0,), np.int8)).size == 0
    assert oracle.sgd_update(np.zeros((0,), np.int8), np.zeros((0,), np.int8)).size == 0
    out, inc = oracle.rescale_act(np.zeros((0,), np.int32))
    assert out.size == 0 and inc == 0


def _dsp_requant_numpy(acc, post_div):
    """independent restatement of the DSP requantisation (niti_op_supernode_new.c:4614-4634, quantize.c:337-400,
    NITI_DSPConv_Int8.cpp:447-452) in numpy, for checking oracle/niti_oracle.c niti_o_rescale_dsp"""
    a = acc.astype(np.int64)
    m = int(np.abs(a).max()) if a.size else 0
    cl = m - 1 if m < 4 else int(m - 1).bit_length()          # quantize.h:260-265 ceiling_log2
    shift = max(cl - 7, 0)
    q = np.floor_divide(a + (1 << (shift - 1)), 1 << shift) if shift > 0 else a
    q = np.clip(q, -128, 127)
    if post_div > 1:
        q = np.sign(q) * (np.abs(q) // post_div)               # C division truncates toward zero
    return q.astype(np.int8), shift


def test_dsp_requant_known_answers_and_restatement(oracle):
    """SURVEY.md section 8 a11 — the DSP path's rounding.  No cDSP here, so this rule is pinned only by the reference's
    source (parity unpinned): hand-worked answers + an independent numpy restatement."""
    acc = np.array([1000, -1000, 511, -513, 4, -4, 3, 0, 996, -996], dtype=np.int32)
    out, shift = oracle.rescale_dsp(acc, 1)
    assert shift == 3                                          # ceiling_log2(1000) = 10
    assert out.tolist() == [125, -125, 64, -64, 1, 0, 0, 0, 125, -124]      # round half up: 0.5 -> 1, -0.5 -> 0
    out16, _ = oracle.rescale_dsp(acc, 16)
    assert out16.tolist() == [7, -7, 4, -4, 0, 0, 0, 0, 7, -7]
    # powers of two sit on the ceil(log2) boundary: 1024 -> shift 3 -> 128 saturates to 127; -1024 -> -128 survives
    out, shift = oracle.rescale_dsp(np.array([1024, -1024], dtype=np.int32), 1)
    assert shift == 3 and out.tolist() == [127, -128]
    # small accumulators (the reference's `1 << shift` with a negative count is undefined): kept as they are
    out, shift = oracle.rescale_dsp(np.array([64, -37, 5], dtype=np.int32), 1)
    assert shift == 0 and out.tolist() == [64, -37, 5]
    for v, want in [(0, -1), (1, 0), (2, 1), (3, 2), (4, 2), (5, 3), (128, 7), (129, 8), (1 << 20, 20), ((1 << 20) + 1, 21)]:
        assert oracle.lib.niti_o_dsp_ceiling_log2(v) == want
    rng = np.random.default_rng(11)
    for mag in (3, 100, 129, 5000, 1 << 15, 1 << 22, (1 << 30) + 12345):
        for div in (1, 16, 32):
            a = rng.integers(-mag, mag + 1, size=1003).astype(np.int32)
            got, s1 = oracle.rescale_dsp(a, div)
            want, s2 = _dsp_requant_numpy(a, div)
            assert s1 == s2 and np.array_equal(got, want), (mag, div)
    out, shift = oracle.rescale_dsp(np.zeros(0, dtype=np.int32), 1)
    assert out.size == 0 and shift == 0


# ---- a4: OpType_NITI_MatMul_Int8, pinned by reference outputs (tests/golden/make_golden.py, ref_matmul) -------------
@pytest.mark.parametrize("i", range(6))
def test_matmul_matches_reference(oracle, golden_mm, i):
    g = golden_mm
    e, l, h = (int(v) for v in g[f"mm{i}.dims"])
    a = oracle.matmul_unpack_a(g[f"mm{i}.a_raw"], e, l)         # the execution's own [l/4][e][4] addressing of A
    c, acc = oracle.matmul_nt(a, g[f"mm{i}.b"])
    assert np.array_equal(oracle.matmul_pack_c(c, e, h), g[f"mm{i}.c_raw"])
    if i == 5:
        assert not c.any()                                       # b == 0 -> zeros (Q2)


# ---- a5: OpType_NITI_GradientSplitCONV_Int8 (per-slice exponents + alignment) ----------------------------------------
@pytest.mark.parametrize("i", range(4))
def test_gradient_split_conv_matches_reference(oracle, golden_split, i):
    g = golden_split
    s = ConvShape(*[int(v) for v in g[f"sp{i}.shape"]])
    slices = int(g[f"sp{i}.slices"])
    assert oracle.split_count(s.oc, s.oh, s.ow) == slices
    dw, k = oracle.conv_dw_split(s, g[f"sp{i}.x"], g[f"sp{i}.dy"])
    assert k == slices and np.array_equal(dw, g[f"sp{i}.dw"])
    assert np.array_equal(oracle.conv_dw(s, g[f"sp{i}.x"], g[f"sp{i}.dy"])[0], g[f"sp{i}.dw_unsplit"])


def test_split_cases_cover_the_rule(golden_split):
    counts = sorted(int(golden_split[f"sp{i}.slices"]) for i in range(4))
    assert counts == [1, 2, 2, 4]
    # the alignment really divides somewhere, and a zero slice exists
    assert any(not np.array_equal(golden_split[f"sp{i}.dw"], golden_split[f"sp{i}.dw_unsplit"]) for i in range(4))


def test_oracle_step_equals_reference_vgg11_steps(oracle):
    """oracle_step on the whole VGG-11 graph (C=3 stem, 3x3 pad 1, five pools, alias flatten, 12 classes) against what the
    unmodified reference's Module / OpGrad / NITI_SGD loop produced (tests/golden/vgg11_ref_steps.npz, batch 4)."""
    import os
    from oracle.oracle import shape
    from oracle.step_twin import oracle_step
    from tests.golden.make_golden_vgg import case_inputs, weight_digest, STEPS, OUT
    g = np.load(OUT)
    spec, x, labels, ws = case_inputs(4)
    w2 = [(w.copy(), e) for w, e in ws]
    for s in range(STEPS):
        lg, e = oracle_step(oracle, spec, x[s], -7, labels[s], w2, shape)
        assert np.array_equal(lg.reshape(4, -1), g["b4.logits"][s])
        assert e == g["b4.e_logits"][s]
    assert [weight_digest(w) for w, _ in w2] == [str(d) for d in g["b4.w_sha256"]]


RGB = ["rgb_stem_3x3_p1", "rgb_stem_3x3_p1_small", "rgb_stem_7x7_s2_p3"]


@pytest.mark.parametrize("name", RGB)
def test_rgb_stem_cases_match_reference(oracle, golden_rgb, name):
    """3-channel image layers (C % 4 != 0: the reference's NC4HW4 packing pads the channel block): forward + dW as the
    unmodified reference computed them.  dX of such a layer is outside the reference's contract (quirk Q6) and never needed
    (the image has no gradient)."""
    g = golden_rgb
    s, x, w, dy = load_case(g, name)
    assert s.c == 3
    y, e, _ = oracle.conv_fwd(s, x, w, -7, -7)
    assert np.array_equal(y, g[name + ".y"]) and e == int(g[name + ".exp"])
    assert np.array_equal(oracle.conv_dw(s, x, dy)[0], g[name + ".dw"])
