"""GPU parity tests proper: the CUDA path, called through the C ABI (ctypes), against the CPU oracle
and the committed golden vectors of the unmodified reference.  Bit-exact (integer work): every
comparison is np.array_equal."""
import numpy as np
import pytest

from oracle.oracle import ConvShape, shape

pytestmark = pytest.mark.gpu

GOLDEN = ["mnist_conv1", "mnist_conv2", "mnist_ip1", "mnist_ip2", "vgg_3x3_p1", "vgg_3x3_p1_b",
          "odd_hw_3x3_p1", "s2_3x3_p1", "s2_1x1_p0", "small_mag_cast", "small_mag_shift1", "zero_dy"]


@pytest.fixture(scope="module")
def nb():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from mandheling_dsp_training_b200 import NitiB200
    return NitiB200(0)


@pytest.fixture(scope="module")
def run(nb):
    from tests.gpu_util import Runner
    return Runner(nb)


def _case(g, name):
    s = ConvShape(*[int(v) for v in g[name + ".shape"]])
    return s, g[name + ".x"], g[name + ".w"], g[name + ".dy"]


def _rand(rng, shp, lo=-127, hi=127):
    return rng.integers(lo, hi + 1, size=shp, dtype=np.int8)


# ------------------------------------------------------------------ golden vectors of the reference
@pytest.mark.parametrize("name", GOLDEN)
def test_golden_reference_vectors(nb, run, golden_conv, name):
    from tests.gpu_util import PATH_AUTO
    nb.set_path(PATH_AUTO)
    g = golden_conv
    s, x, w, dy = _case(g, name)
    y, e = run.conv_fwd(s, x, w, -7, -7)
    assert np.array_equal(y, g[name + ".y"])
    assert e == int(g[name + ".exp"])
    assert np.array_equal(run.conv_dw(s, x, dy), g[name + ".dw"])
    if name + ".dx" in g.files:
        assert np.array_equal(run.conv_dx(s, dy, w), g[name + ".dx"])


# ------------------------------------------------------------------ both kernel families vs the oracle
TC_SHAPES = [
    # n, c, h, w, oc, k, stride, pad
    (2, 16, 8, 8, 32, 3, 1, 1),
    (4, 64, 8, 8, 128, 3, 1, 1),      # one full 128-byte K row per tap-half
    (2, 128, 6, 6, 256, 3, 1, 1),     # BN = 128/256, K = 1152
    (3, 32, 9, 7, 48, 3, 1, 1),       # ragged M (189 pixels), non-square
    (2, 16, 12, 12, 16, 5, 1, 0),     # 5x5, K = 400 (ragged K block)
    (64, 832, 1, 1, 500, 1, 1, 0),    # MNIST ip1 as 1x1 conv (fwd eligible only)
    (2, 32, 10, 10, 64, 3, 2, 1),     # stride 2
    (2, 64, 8, 8, 64, 1, 2, 0),       # 1x1 stride 2 (ResNet downsample)
    (1, 256, 4, 4, 512, 3, 1, 1),     # K = 2304, N = 512
    (2, 48, 5, 5, 80, 3, 1, 2),       # pad 2 > (k-1)/2
    (2, 64, 9, 9, 64, 3, 1, 1),       # TMA path, 64-byte swizzle, ragged M
    (3, 128, 7, 5, 192, 3, 1, 1),     # TMA path, non-square, N = 192 (ragged N tile)
    (2, 64, 12, 12, 128, 5, 1, 2),    # 5x5 pad 2
    (2, 128, 9, 9, 64, 3, 2, 1),      # stride 2 forward / dW via TMA, dX via cp.async
    (256, 512, 1, 1, 512, 1, 1, 0),   # FC as 1x1 conv, split-K
    (4, 16, 32, 32, 64, 3, 1, 1),     # VGG stem (3 -> 16 padded channels): dW with all 9 taps in one MMA
    (2, 16, 36, 36, 64, 7, 2, 3),     # ResNet stem 7x7 stride 2: 49 taps = 4 tap groups
    (3, 16, 12, 12, 32, 5, 1, 2),     # 5x5 stem, 25 taps = 16 + 9, ragged pixel count
    (2, 64, 10, 10, 64, 3, 1, 0),     # pad 0
    (2, 16, 9, 9, 32, 3, 1, 2),       # dilation handled below with dil=2
]


@pytest.mark.parametrize("cfg", TC_SHAPES)
@pytest.mark.parametrize("path", ["simt", "tc", "tma", "tma2"])
def test_accumulators_and_outputs_vs_oracle(nb, run, oracle, cfg, path):
    from tests.gpu_util import PATH_SIMT, PATH_TC, PATH_TMA, PATH_TMA2, to_desc
    n, c, h, w, oc, k, st, p = cfg
    dil = 2 if cfg == TC_SHAPES[-1] else 1
    s = shape(n, c, h, w, oc, k, stride=st, pad=p, dil=dil)
    rng = np.random.default_rng(hash(cfg) % (2 ** 31))
    x = _rand(rng, (n, c, h, w)); wt = _rand(rng, (oc, c, k, k), -64, 64)
    dy = _rand(rng, (n, oc, s.oh, s.ow), -50, 50)
    nb.set_path({"simt": PATH_SIMT, "tc": PATH_TC, "tma": PATH_TMA, "tma2": PATH_TMA2}[path])
    try:
        if path in ("tma", "tma2"):
            fwd_ok = c % 64 == 0 and oc % 4 == 0
            dx_ok = st == 1 and oc % 64 == 0 and c % 4 == 0 and (k - 1) * dil - p >= 0
            dw_ok = oc % 16 == 0 and (c % 64 == 0 or (c == 16 and path == "tma2"))   # "tma2": conv_tma2.cu / conv_dw2.cu
            if not (fwd_ok or dx_ok or dw_ok):
                pytest.skip("shape not eligible for the TMA path")
        else:
            fwd_ok = path == "simt" or (c % 16 == 0 and oc % 4 == 0)
            dx_ok = path == "simt" or (oc % 16 == 0 and c % 4 == 0)
            dw_ok = path == "simt" or (oc % 16 == 0 and c % 16 == 0)
        if fwd_ok:
            acc, mx = run.conv_fwd_acc(s, x, wt)
            oacc, _ = oracle.conv_fwd_acc(s, x, wt)
            assert np.array_equal(acc, oacc)
            assert mx == int(np.abs(oacc.astype(np.int64)).max())
            assert nb.last_path == path
            y, e = run.conv_fwd(s, x, wt, -3, -9)
            oy, oe, _ = oracle.conv_fwd(s, x, wt, -3, -9)
            assert np.array_equal(y, oy) and e == oe
        if dx_ok:
            acc, mx = run.conv_dx_acc(s, dy, wt)
            oacc, _ = oracle.conv_dx_acc(s, dy, wt)
            assert np.array_equal(acc, oacc)
            assert mx == int(np.abs(oacc.astype(np.int64)).max())
            assert np.array_equal(run.conv_dx(s, dy, wt), oracle.conv_dx(s, dy, wt)[0])
        if dw_ok:
            acc, mx = run.conv_dw_acc(s, x, dy)
            oacc, _ = oracle.conv_dw_acc(s, x, dy)
            assert np.array_equal(acc, oacc)
            assert mx == int(np.abs(oacc.astype(np.int64)).max())
            assert np.array_equal(run.conv_dw(s, x, dy), oracle.conv_dw(s, x, dy)[0])
    finally:
        nb.set_path(0)


STEM_SHAPES = [(4, 16, 32, 32, 64, 3, 1, 1), (2, 16, 36, 36, 64, 7, 2, 3), (3, 16, 12, 12, 32, 5, 1, 2),
               (2, 16, 9, 7, 48, 3, 1, 1), (1, 16, 28, 28, 16, 5, 1, 0), (70, 16, 32, 32, 64, 3, 1, 1), (2, 64, 8, 8, 64, 3, 1, 1)]


@pytest.mark.parametrize("cfg", [c for c in TC_SHAPES if c[4] % 16 == 0] + [(4, 16, 6, 6, 12, 1, 1, 0)])
def test_conv_dw_accumulate_adds_into_caller_zeroed_buffer(nb, oracle, cfg):
    """niti_conv_dw_accumulate (the barrier-free dW launch the step runs beside dX): acc == dW after one call into a
    zeroed buffer and 2*dW after a second call wherever the tcgen05 kernel ran (it adds); the SIMT kernel, taken for
    shapes the tcgen05 kernel does not cover, overwrites.  Neighbouring words of the bucket stay untouched."""
    import torch
    from tests.gpu_util import dev, nhwc, nchw, to_desc
    n, c, h, w, oc, k, st, p = cfg
    dil = 2 if cfg == TC_SHAPES[-1] else 1
    s = shape(n, c, h, w, oc, k, stride=st, pad=p, dil=dil)
    rng = np.random.default_rng(hash(cfg) % (2 ** 31))
    x = _rand(rng, (n, c, h, w)); dy = _rand(rng, (n, oc, s.oh, s.ow), -50, 50)
    oacc, _ = oracle.conv_dw_acc(s, x, dy)
    nw = oc * k * k * c
    bucket = torch.zeros(nw + 128, dtype=torch.int32, device="cuda:0")
    bucket[:64] = 7; bucket[64 + nw:] = -7
    acc = bucket[64: 64 + nw]
    dx_, dy_ = dev(nhwc(x)), dev(nhwc(dy))
    nb.conv_dw_accumulate(to_desc(s), dx_, dy_, acc)
    torch.cuda.synchronize()
    got = nchw(acc.view(oc, k, k, c).cpu().numpy())
    assert np.array_equal(got, oacc)
    adds = nb.last_path == "tma2"
    nb.conv_dw_accumulate(to_desc(s), dx_, dy_, acc)
    torch.cuda.synchronize()
    got = nchw(acc.view(oc, k, k, c).cpu().numpy())
    assert np.array_equal(got, 2 * oacc if adds else oacc)
    assert bool((bucket[:64] == 7).all()) and bool((bucket[64 + nw:] == -7).all())


@pytest.mark.parametrize("cfg", STEM_SHAPES)
@pytest.mark.parametrize("mag", [127, 3, 1])
def test_conv_fwd_fused_vs_oracle(nb, run, oracle, cfg, mag):
    """the one-launch forward op (stem kernel: recompute + grid barrier; other shapes: phase 1 + rescale) against the
    oracle's NITI_Conv_Int8 (+ ReLU), over magnitudes that hit all three branches of the shift rule; the (70, ...) case
    gives every CTA several tiles per pass, the last shape takes the generic two-phase route"""
    n, c, h, w, oc, k, st, p = cfg
    s = shape(n, c, h, w, oc, k, stride=st, pad=p)
    rng = np.random.default_rng(hash(cfg) % (2 ** 31) + mag)
    x = _rand(rng, (n, c, h, w), -mag, mag); wt = _rand(rng, (oc, c, k, k), -mag, mag)
    if mag == 1:
        x[:, 1:] = 0; wt[:, 1:] = 0; wt[:, :, 1:, :] = 0; wt[:, :, :, 1:] = 0      # max|acc| <= 1: the raw-cast branch
    for relu in (False, True):
        y, e, mx = run.conv_fwd_fused(s, x, wt, -3, -9, relu)
        oy, oe, oacc = oracle.conv_fwd(s, x, wt, -3, -9)
        if relu:
            oy = oracle.relu(oy)
        assert mx == int(np.abs(oacc.astype(np.int64)).max())
        assert e == oe
        assert np.array_equal(y, oy)
    if c == 16:
        assert nb.last_path == "stem"


FUSED_SHAPES = [(4, 64, 8, 8, 128, 3, 1, 1), (2, 128, 6, 6, 256, 3, 1, 1), (3, 128, 7, 5, 192, 3, 1, 1), (1, 256, 4, 4, 512, 3, 1, 1),
                (64, 64, 16, 16, 128, 3, 1, 1), (256, 512, 1, 1, 512, 1, 1, 0), (40, 64, 16, 16, 64, 3, 1, 1),
                (2, 64, 12, 12, 128, 5, 1, 2), (2, 128, 9, 9, 64, 3, 2, 1), (2, 64, 10, 10, 128, 1, 2, 0), (3, 64, 7, 7, 64, 3, 2, 1),
                (2, 64, 8, 8, 64, 3, 2, 0)]


@pytest.mark.parametrize("cfg", FUSED_SHAPES)
def test_fused_single_phase_ops_vs_oracle(nb, run, oracle, cfg):
    """accumulators resident in TMEM across the exponent barrier (igemm_fused_kernel): forward (+ReLU) and dX (+ReLU-grad
    gate) against the oracle; BN 64/128/256, several resident tiles per CTA pair, ragged M and N"""
    n, c, h, w, oc, k, st, p = cfg
    s = shape(n, c, h, w, oc, k, stride=st, pad=p)
    rng = np.random.default_rng(hash(cfg) % (2 ** 31))
    x = _rand(rng, (n, c, h, w)); wt = _rand(rng, (oc, c, k, k), -64, 64)
    dy = _rand(rng, (n, oc, s.oh, s.ow), -50, 50)
    for relu in (False, True):
        y, e, mx = run.conv_fwd_fused(s, x, wt, -3, -9, relu)
        oy, oe, oacc = oracle.conv_fwd(s, x, wt, -3, -9)
        assert mx == int(np.abs(oacc.astype(np.int64)).max()) and e == oe
        assert np.array_equal(y, oracle.relu(oy) if relu else oy)
    assert nb.last_path == "fused"
    # strided layers: dY with zeros inserted through the same fused kernel (the reference's LeftPoolGrad + DeConv lowering)
    odx, oacc = oracle.conv_dx(s, dy, wt)
    dx, mx = run.conv_dx_fused(s, dy, wt)
    assert mx == int(np.abs(oacc.astype(np.int64)).max())
    assert np.array_equal(dx, odx)
    assert nb.last_path == "fused"
    dxg, _ = run.conv_dx_fused(s, dy, wt, gate=x)
    assert np.array_equal(dxg, oracle.relu_grad(x, odx))


@pytest.mark.parametrize("mag", [3, 64, 129, 5000, 1 << 15, 1 << 22, (1 << 30) + 12345])
@pytest.mark.parametrize("div", [1, 16, 32])
def test_rescale_dsp_vs_oracle(nb, oracle, mag, div):
    """niti_rescale_dsp (the DSP-named op types' round-half-up requantisation) against the
    oracle's restatement of niti_op_supernode_new.c:4614-4634 + quantize.c:337-400, ragged length, exponent output"""
    import torch
    rng = np.random.default_rng(mag % 1000 + div)
    n = 16 * 517 + 11
    a = rng.integers(-mag, mag + 1, size=n).astype(np.int32)
    a[5] = mag; a[n - 1] = -mag
    want, shift = oracle.rescale_dsp(a, div)
    acc = torch.from_numpy(a).cuda()
    mx = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    out = torch.zeros(n, dtype=torch.int8, device="cuda:0")
    ex = torch.tensor([-5], dtype=torch.int8, device="cuda:0"); ew = torch.tensor([-9], dtype=torch.int8, device="cuda:0")
    ey = torch.zeros(1, dtype=torch.int8, device="cuda:0")
    nb.absmax(acc, n, mx)
    nb.rescale_dsp(acc, n, mx, out, div, ex, ew, ey)
    torch.cuda.synchronize()
    assert int(mx.item()) == mag
    assert np.array_equal(out.cpu().numpy(), want)
    assert int(ey.item()) == -14 + shift


@pytest.mark.parametrize("cfg", [(2, 16, 8, 8, 32, 3, 1, 1), (4, 64, 8, 8, 128, 3, 1, 1), (3, 8, 6, 6, 12, 3, 1, 0)])
def test_dsp_conv_ops_vs_oracle(nb, run, oracle, cfg):
    """the contractions behind OpType_NITI_DSP_CONV / _DECONV / _*GRADIENT*CONV with the DSP rounding: tcgen05 / SIMT
    accumulators + niti_rescale_dsp == oracle accumulators + restated DSP requantisation (/32 for TRANSPOSEGRADIENT)"""
    import torch
    from tests.gpu_util import dev, nhwc, to_desc
    n, c, h, w, oc, k, st, p = cfg
    s = shape(n, c, h, w, oc, k, stride=st, pad=p)
    rng = np.random.default_rng(hash(cfg) % (2 ** 31))
    x = _rand(rng, (n, c, h, w)); wt = _rand(rng, (oc, c, k, k), -64, 64)
    dy = _rand(rng, (n, oc, s.oh, s.ow), -50, 50)
    for kind, div in (("fwd", 1), ("dw", 32), ("dw", 16)):
        if kind == "fwd":
            oacc, _ = oracle.conv_fwd_acc(s, x, wt)
            acc = torch.empty((n, s.oh, s.ow, oc), dtype=torch.int32, device="cuda:0")
            mx = torch.zeros(1, dtype=torch.int32, device="cuda:0")
            nb.conv_fwd_acc(to_desc(s), dev(nhwc(x)), dev(nhwc(wt)), acc, mx)
            want, _ = oracle.rescale_dsp(nhwc(oacc), div)
        else:
            oacc, _ = oracle.conv_dw_acc(s, x, dy)
            acc = torch.empty((oc, k, k, c), dtype=torch.int32, device="cuda:0")
            mx = torch.zeros(1, dtype=torch.int32, device="cuda:0")
            nb.conv_dw_acc(to_desc(s), dev(nhwc(x)), dev(nhwc(dy)), acc, mx)
            want, _ = oracle.rescale_dsp(nhwc(oacc), div)
        out = torch.zeros(acc.numel(), dtype=torch.int8, device="cuda:0")
        nb.rescale_dsp(acc, acc.numel(), mx, out, div)
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy().reshape(want.shape), want), (kind, div)


def test_matmul_nt(nb, run, oracle):
    rng = np.random.default_rng(5)
    for (e, h, l) in [(64, 500, 832), (25, 52, 4096), (7, 12, 33)]:
        a = _rand(rng, (e, l), -30, 30); b = _rand(rng, (h, l), -30, 30)
        assert np.array_equal(run.matmul_nt(a, b), oracle.matmul_nt(a, b)[0])


@pytest.mark.parametrize("i", range(6))
def test_matmul_golden_reference_vectors(nb, run, oracle, golden_mm, i):
    """a4: niti_matmul_nt against outputs of the reference's NITI_Matmul_Int8 execution (tests/golden/make_golden.py);
    A / C cross the reference boundary in its [channel/4][batch][4] addressing, restated by matmul_unpack_a / pack_c"""
    g = golden_mm
    e, l, h = (int(v) for v in g[f"mm{i}.dims"])
    a = oracle.matmul_unpack_a(g[f"mm{i}.a_raw"], e, l)
    c = run.matmul_nt(a, g[f"mm{i}.b"])
    assert np.array_equal(oracle.matmul_pack_c(c, e, h), g[f"mm{i}.c_raw"])
    assert np.array_equal(c, oracle.matmul_nt(a, g[f"mm{i}.b"])[0])


@pytest.mark.parametrize("i", range(4))
def test_gradient_split_conv_golden_reference_vectors(nb, oracle, golden_split, i):
    """a5: niti_conv_dw_split (contraction + per-slice exponents + alignment, NITI_GradientSplitConv_Int8.cpp:318-362)
    against outputs of the reference execution"""
    import torch
    from tests.gpu_util import to_desc, dev, nhwc, nchw
    g = golden_split
    s = ConvShape(*[int(v) for v in g[f"sp{i}.shape"]])
    slices = int(g[f"sp{i}.slices"])
    assert nb.split_count(s.oc, s.oh, s.ow) == slices == oracle.split_count(s.oc, s.oh, s.ow)
    dw = torch.empty((s.oc, s.kh, s.kw, s.c), dtype=torch.int8, device="cuda:0")
    nb.conv_dw_split(to_desc(s), dev(nhwc(g[f"sp{i}.x"])), dev(nhwc(g[f"sp{i}.dy"])), slices, dw)
    torch.cuda.synchronize()
    assert np.array_equal(nchw(dw.cpu().numpy()), g[f"sp{i}.dw"])


@pytest.mark.parametrize("slices", [1, 2, 3, 8, 32])
@pytest.mark.parametrize("mags", [(0, 1, 5), (100, 1 << 12, 1 << 20), (3, 3, 3), ((1 << 30) + 7, 2, 0)])
def test_rescale_grad_split_vs_oracle(nb, oracle, slices, mags):
    """the split rule alone, over magnitudes that put slices up to 30 bits apart (1 << 31 divisor wraps like the x86 build)"""
    import torch
    rng = np.random.default_rng(slices * 7 + len(mags))
    per = 1000
    acc = np.zeros((slices, per), np.int32)
    for sl in range(slices):
        m = mags[sl % len(mags)]
        if m:
            acc[sl] = rng.integers(-m, m + 1, size=per)
            acc[sl, 0] = m
    want = oracle.rescale_grad_split(acc, slices)
    a = torch.from_numpy(acc.reshape(-1)).cuda()
    mx = torch.zeros(slices, dtype=torch.int32, device="cuda:0")
    for sl in range(slices):
        nb.absmax(a[sl * per:(sl + 1) * per], per, mx[sl:sl + 1])
    out = torch.empty(slices * per, dtype=torch.int8, device="cuda:0")
    nb.rescale_grad_split(a, per, slices, mx, 3, out)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().reshape(slices, per), want)


# ------------------------------------------------------------------ rescale / update kernels
@pytest.mark.parametrize("n", [0, 1, 15, 16, 17, 4096, 100003])
@pytest.mark.parametrize("mag", [0, 1, 3, 100, 128, 129, 255, 256, 257, 1 << 14, (1 << 23) + 7, (1 << 30) + 5])
def test_rescale_rules(run, oracle, n, mag):
    rng = np.random.default_rng(n * 131 + mag % 977)
    acc = rng.integers(-mag, mag + 1, size=n).astype(np.int32)
    if n:
        acc[rng.integers(0, n)] = -mag if (mag % 2) else mag
    out, eo = run.rescale_act(acc, -5, 2)
    oout, inc = oracle.rescale_act(acc)
    assert np.array_equal(out, oout)
    assert eo == int(np.int8(-5 + 2 + inc))
    if n:
        w = _rand(rng, (n,), -128, 127)
        for sub in (2, 3):
            o, wnew = run.rescale_grad(acc, sub, w)
            oo = oracle.rescale_grad(acc, sub)
            assert np.array_equal(o, oo)
            assert np.array_equal(wnew, oracle.sgd_update(w, oo))


def test_sgd_update(nb, oracle):
    import torch
    rng = np.random.default_rng(11)
    for n in (1, 31, 4096, 99999):
        w = _rand(rng, (n,), -128, 127); dw = _rand(rng, (n,), -128, 127)
        wd = torch.from_numpy(w.copy()).cuda(); dd = torch.from_numpy(dw).cuda()
        nb.sgd_update(wd, dd, n)
        assert np.array_equal(wd.cpu().numpy(), oracle.sgd_update(w, dw))


def test_update_batched_vs_oracle(nb, oracle):
    """one launch for every layer: max|acc| -> rule B (sub 2) -> int8 SGD -> transposed copy; full 32x32 tiles, ragged
    rows / columns, a zero segment (b == 0) and magnitudes on both sides of the shift rules, twice (barrier re-arm)"""
    import torch
    rng = np.random.default_rng(21)
    shapes = [(64, 9, 64, 1 << 20, True), (12, 1, 512, 300, True), (20, 25, 16, 5, False), (52, 25, 20, 129, True),
              (96, 9, 32, 0, True), (500, 1, 832, (1 << 23) + 3, True), (33, 3, 7, 2, True), (512, 9, 512, 70000, True)]
    for rep in range(2):
        layers, ref = [], []
        slots = torch.zeros(len(shapes), dtype=torch.int32, device="cuda")
        for i, (oc, taps, c, mag, with_wt) in enumerate(shapes):
            n = oc * taps * c
            acc = rng.integers(-mag, mag + 1, size=n).astype(np.int32)
            if mag:
                acc[rng.integers(0, n)] = mag
            w = _rand(rng, (oc, taps, c), -127, 127)
            n_pad = (n + 63) // 64 * 64
            accd = torch.zeros(n_pad, dtype=torch.int32, device="cuda"); accd[:n] = torch.from_numpy(acc).cuda()
            wd = torch.from_numpy(w.copy()).cuda()
            wtd = torch.full((c, taps, oc), 77, dtype=torch.int8, device="cuda") if with_wt else None
            layers.append((accd[:n], wd, wtd, slots[i:i + 1], oc, taps, c, 2))
            ref.append((acc, w, mag))
        nb.update_layers(layers)
        torch.cuda.synchronize()
        for (accd, wd, wtd, slot, oc, taps, c, _), (acc, w, mag) in zip(layers, ref):
            assert int(slot.item()) == mag
            want = oracle.sgd_update(w.reshape(-1), oracle.rescale_grad(acc, 2)).reshape(oc, taps, c)
            assert np.array_equal(wd.cpu().numpy(), want), (oc, taps, c, mag)
            if wtd is not None:
                assert np.array_equal(wtd.cpu().numpy(), np.transpose(want, (2, 1, 0))), (oc, taps, c, mag)


# ------------------------------------------------------------------ element-wise ops of the step
def test_elementwise_vs_oracle_and_golden(nb, oracle, golden_ew):
    import torch
    from tests.gpu_util import nhwc, nchw, dev
    g = golden_ew
    x = g["relu.x"]
    n, c, h, w = x.shape
    xd = dev(nhwc(x)); y = torch.empty_like(xd)
    nb.relu(xd, y, xd.numel())
    assert np.array_equal(nchw(y.cpu().numpy()), g["relu.y"])
    dyd = dev(nhwc(g["relu_grad.dy"])); dx = torch.empty_like(xd)
    nb.relu_grad(xd, dyd, dx, xd.numel())
    assert np.array_equal(nchw(dx.cpu().numpy()), g["relu_grad.dx"])
    yp = torch.empty((n, h // 2, w // 2, c), dtype=torch.int8, device="cuda")
    nb.maxpool(xd, n, h, w, c, 2, 2, yp)
    assert np.array_equal(nchw(yp.cpu().numpy()), g["maxpool.y"])
    # pool grad vs oracle, including ties (post-relu zeros) and an overlapping window (k=3, s=2)
    rng = np.random.default_rng(2)
    for (k, s, src) in [(2, 2, x), (2, 2, oracle.relu(x)), (3, 2, oracle.relu(x))]:
        oy = oracle.maxpool(src, k, s)
        dyp = _rand(rng, oy.shape)
        sd = dev(nhwc(src)); yd = dev(nhwc(oy)); gd = dev(nhwc(dyp)); out = torch.empty_like(sd)
        nb.maxpool_grad(sd, yd, gd, n, h, w, c, k, s, out)
        assert np.array_equal(nchw(out.cpu().numpy()), oracle.maxpool_grad(src, oy, dyp, k, s))
    for a in (-9, -7, -5, -3, 0, 1):
        lg, lab = g[f"loss_grad.{a}.logits"], g[f"loss_grad.{a}.labels"]
        out = torch.empty(lg.shape, dtype=torch.int8, device="cuda")
        nb.loss_grad(dev(lg), torch.tensor([a], dtype=torch.int8, device="cuda"), dev(lab.astype(np.int32)),
                     lg.shape[0], lg.shape[1], out)
        assert np.array_equal(out.cpu().numpy(), g[f"loss_grad.{a}.grad"]), a
        # the float loss scalar (NITI_CPULoss_Int8): floating point — tolerance 1e-5 relative against the reference's value
        # and the oracle; also through padded rows (stride 16)
        rows = torch.empty(lg.shape[0], dtype=torch.float32, device="cuda"); loss = torch.zeros(1, dtype=torch.float32, device="cuda")
        asc = torch.tensor([a], dtype=torch.int8, device="cuda")
        nb.loss_value(dev(lg), asc, dev(lab.astype(np.int32)), lg.shape[0], lg.shape[1], lg.shape[1], rows, loss)
        assert float(loss.item()) == pytest.approx(float(g[f"loss_value.{a}"]), rel=1e-5), a
        assert float(loss.item()) == pytest.approx(oracle.loss_value(lg, a, lab), rel=1e-5), a
        padded = torch.zeros((lg.shape[0], 16), dtype=torch.int8, device="cuda"); padded[:, :lg.shape[1]] = dev(lg)
        nb.loss_value(padded, asc, dev(lab.astype(np.int32)), lg.shape[0], lg.shape[1], 16, rows, loss)
        assert float(loss.item()) == pytest.approx(float(g[f"loss_value.{a}"]), rel=1e-5), a


def test_layout_round_trips(nb):
    import torch
    rng = np.random.default_rng(9)
    for (n, c, h, w) in [(3, 20, 5, 7), (2, 4, 3, 3), (1, 1, 28, 28), (4, 52, 8, 8)]:
        x = _rand(rng, (n, c, h, w))
        xd = torch.from_numpy(x).cuda()
        a = torch.empty((n, h, w, c), dtype=torch.int8, device="cuda")
        nb.nchw_to_nhwc(xd, n, c, h, w, a)
        assert np.array_equal(a.cpu().numpy(), np.transpose(x, (0, 2, 3, 1)))
        b = torch.empty_like(xd)
        nb.nhwc_to_nchw(a, n, c, h, w, b)
        assert np.array_equal(b.cpu().numpy(), x)
        c4 = (c + 3) // 4
        p = torch.empty((c4, n, h * w, 4), dtype=torch.int8, device="cuda")
        nb.nhwc_to_nc4hw4(a, n, c, h, w, p)
        ref = np.zeros((c4 * 4, n, h * w), np.int8)
        ref[:c] = np.transpose(x.reshape(n, c, h * w), (1, 0, 2))
        ref = np.transpose(ref.reshape(c4, 4, n, h * w), (0, 2, 3, 1))
        assert np.array_equal(p.cpu().numpy(), ref)
        back = torch.zeros_like(a)
        nb.nc4hw4_to_nhwc(p, n, c, h, w, back)
        assert np.array_equal(back.cpu().numpy(), a.cpu().numpy())


def test_host_buffer_call(nb, oracle):
    from tests.gpu_util import nhwc, nchw, to_desc
    rng = np.random.default_rng(4)
    s = shape(4, 32, 8, 8, 64, 3, pad=1)
    x = _rand(rng, (4, 32, 8, 8)); w = _rand(rng, (64, 32, 3, 3), -40, 40)
    y = np.empty((4, 8, 8, 64), np.int8)
    e = nb.conv_fwd_host(to_desc(s), nhwc(x), nhwc(w), -7, -7, y)
    oy, oe, _ = oracle.conv_fwd(s, x, w, -7, -7)
    assert np.array_equal(nchw(y), oy) and e == oe


# ------------------------------------------------------------------ full-size properties (BASELINE sizes)
def test_full_size_vgg_layer_properties(nb):
    """VGG-11 256->256 @ 8x8, batch 256 (M=16384, K=2304): too big for the CPU oracle in seconds, so
    check size-independent properties on the device: (1) the tensor-core and SIMT kernels agree
    element for element, (2) dW is additive over a batch split, (3) conv is odd in x."""
    import torch
    from tests.gpu_util import PATH_SIMT, PATH_TC, PATH_TMA
    from mandheling_dsp_training_b200.lib import desc
    torch.manual_seed(0)
    n, c, hw, oc = 256, 256, 8, 256
    d = desc(n, c, hw, hw, oc, 3, pad=1)
    x = torch.randint(-127, 128, (n, hw, hw, c), dtype=torch.int8, device="cuda")
    w = torch.randint(-20, 21, (oc, 3, 3, c), dtype=torch.int8, device="cuda")
    dy = torch.randint(-30, 31, (n, hw, hw, oc), dtype=torch.int8, device="cuda")
    outs = {}
    for path in (PATH_SIMT, PATH_TC, PATH_TMA):
        nb.set_path(path)
        acc = torch.empty((n, hw, hw, oc), dtype=torch.int32, device="cuda")
        dxa = torch.empty((n, hw, hw, c), dtype=torch.int32, device="cuda")
        dwa = torch.empty((oc, 3, 3, c), dtype=torch.int32, device="cuda")
        nb.conv_fwd_acc(d, x, w, acc); nb.conv_dx_acc(d, dy, w, dxa); nb.conv_dw_acc(d, x, dy, dwa)
        torch.cuda.synchronize()
        outs[path] = (acc, dxa, dwa)
    nb.set_path(0)
    for a, b, c_ in zip(outs[PATH_SIMT], outs[PATH_TC], outs[PATH_TMA]):
        assert torch.equal(a, b) and torch.equal(a, c_)
    acc, dxa, dwa = outs[PATH_TMA]
    # (2) additivity of dW over the batch
    nb.set_path(PATH_TMA)
    dh = desc(n // 2, c, hw, hw, oc, 3, pad=1)
    p0 = torch.empty_like(dwa); p1 = torch.empty_like(dwa)
    nb.conv_dw_acc(dh, x[: n // 2].contiguous(), dy[: n // 2].contiguous(), p0)
    nb.conv_dw_acc(dh, x[n // 2:].contiguous(), dy[n // 2:].contiguous(), p1)
    # (3) oddness: conv(-x) == -conv(x)   (x has no -128)
    neg = torch.empty_like(acc)
    nb.conv_fwd_acc(d, (-x).contiguous(), w, neg)
    torch.cuda.synchronize()
    nb.set_path(0)
    assert torch.equal(p0 + p1, dwa)
    assert torch.equal(neg, -acc)


@pytest.mark.parametrize("shape_", [(256, 256, 8, 256), (256, 512, 2, 512), (256, 64, 16, 128), (256, 512, 4, 512)])
def test_full_size_product_kernels_vs_pinned_simt_and_oracle_rescale(nb, oracle, shape_):
    """The three kernels bench.py times — igemm_fused_kernel (fwd + dX, incl. several resident tiles per CTA pair),
    the two-phase PATH_TMA2 route and dw2_kernel (niti_conv_dw_accumulate) — at the FULL VGG-11 batch-256 layer sizes.
    Oracle leg: the SIMT kernel's int32 accumulators (itself pinned against the CPU oracle element for element at small
    sizes, test_accumulators_and_outputs_vs_oracle) rescaled by the CPU oracle (niti_o_rescale_act)."""
    import torch
    from tests.gpu_util import PATH_SIMT, PATH_TMA2
    from mandheling_dsp_training_b200.lib import desc
    n, c, hw, oc = shape_
    torch.manual_seed(hw)
    d = desc(n, c, hw, hw, oc, 3, pad=1)
    i8 = dict(dtype=torch.int8, device="cuda"); i32 = dict(dtype=torch.int32, device="cuda")
    x = torch.randint(0, 128, (n, hw, hw, c), **i8)                      # post-ReLU activations
    w = torch.randint(-30, 31, (oc, 3, 3, c), **i8)
    dy = torch.randint(-40, 41, (n, hw, hw, oc), **i8)
    gate = torch.randint(-5, 6, (n, hw, hw, c), **i8)
    wt = w.permute(3, 1, 2, 0).contiguous()
    ref = {}
    for path in (PATH_SIMT, PATH_TMA2):
        nb.set_path(path)
        acc = torch.empty((n, hw, hw, oc), **i32); dxa = torch.empty((n, hw, hw, c), **i32); dwa = torch.empty((oc, 3, 3, c), **i32)
        nb.conv_fwd_acc(d, x, w, acc); nb.conv_dx_acc(d, dy, w, dxa); nb.conv_dw_acc(d, x, dy, dwa)
        torch.cuda.synchronize()
        ref[path] = (acc, dxa, dwa)
    nb.set_path(0)
    for a, b in zip(ref[PATH_SIMT], ref[PATH_TMA2]):
        assert torch.equal(a, b)
    acc, dxa, dwa = ref[PATH_SIMT]
    # fused forward (+ReLU) and fused dX (+ReLU-grad gate): single launch, accumulators never leave tensor memory
    y = torch.empty((n, hw, hw, oc), **i8); ey = torch.zeros(1, **i8); mx = torch.zeros(2, **i32)
    ex = torch.full((1,), -7, **i8); ew = torch.full((1,), -9, **i8)
    nb.conv_fwd_fused(d, x, w, ex, ew, y, ey, True, None, mx[0:1])
    assert nb.last_path == "fused"
    dx = torch.empty((n, hw, hw, c), **i8)
    nb.conv_dx_fused(d, dy, w, wt, dx, gate, None, mx[1:2])
    assert nb.last_path == "fused"
    dw2 = torch.zeros((oc, 3, 3, c), **i32)
    nb.conv_dw_accumulate(d, x, dy, dw2)
    torch.cuda.synchronize()
    want_y, inc = oracle.rescale_act(acc.cpu().numpy())
    assert np.array_equal(y.cpu().numpy(), np.maximum(want_y, 0))
    assert int(ey.item()) == -16 + inc and int(mx[0].item()) == int(acc.abs().max().item())
    want_dx, _ = oracle.rescale_act(dxa.cpu().numpy())
    want_dx = np.where(gate.cpu().numpy() > 0, want_dx, 0).astype(np.int8)
    assert np.array_equal(dx.cpu().numpy(), want_dx)
    assert torch.equal(dw2, dwa)


# ------------------------------------------------------------------ pool / pool-gradient folded into the fused ops, chains
POOL_SHAPES = [(8, 64, 16, 128), (16, 64, 8, 64), (4, 128, 4, 256), (32, 64, 2, 64), (2, 64, 12, 64)]   # (n, c, hw, oc); 12: unfused route


@pytest.mark.parametrize("cfg", POOL_SHAPES)
def test_conv_fwd_fused_pool_vs_oracle(nb, oracle, cfg):
    """NITI_Conv_Int8 + NITI_CPURelu_Int8 + NITI_Maxpool_Int8 (2x2/2) behind one call: y and y_pool against the oracle's
    op-by-op sequence"""
    import torch
    from tests.gpu_util import to_desc, dev, nhwc, nchw
    n, c, hw, oc = cfg
    s = shape(n, c, hw, hw, oc, 3, stride=1, pad=1)
    rng = np.random.default_rng(hw * oc)
    x = _rand(rng, (n, c, hw, hw)); w = _rand(rng, (oc, c, 3, 3), -40, 40)
    want_y, want_e, _ = oracle.conv_fwd(s, x, w, -7, -6)
    want_y = oracle.relu(want_y); want_p = oracle.maxpool(want_y, 2, 2)
    i8 = dict(dtype=torch.int8, device="cuda:0")
    y = torch.empty((n, hw, hw, oc), **i8); yp = torch.empty((n, hw // 2, hw // 2, oc), **i8)
    ey = torch.zeros(1, **i8); mx = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    acc = torch.empty(y.numel(), dtype=torch.int32, device="cuda:0")
    nb.conv_fwd_fused_pool(to_desc(s), dev(nhwc(x)), dev(nhwc(w)), torch.full((1,), -7, **i8), torch.full((1,), -6, **i8), y, ey, True, yp, acc, mx)
    torch.cuda.synchronize()
    assert np.array_equal(nchw(y.cpu().numpy()), want_y)
    assert np.array_equal(nchw(yp.cpu().numpy()), want_p)
    assert int(ey.item()) == want_e


# (n, c_live, h, oc, k, stride, pad, pool)
STEM2_SHAPES = [(8, 3, 32, 64, 3, 1, 1, True), (5, 3, 16, 32, 3, 1, 1, True), (3, 1, 28, 16, 5, 1, 2, False), (3, 1, 28, 16, 5, 1, 2, True),
                (4, 3, 32, 64, 7, 2, 3, True), (2, 4, 12, 128, 3, 1, 1, False), (7, 2, 20, 48, 3, 1, 0, False)]


@pytest.mark.parametrize("cfg", STEM2_SHAPES)
def test_conv_stem_fwd_vs_oracle(nb, oracle, cfg):
    """niti_conv_stem_fwd: image tensor stored with 16 channels of which c_live are real; conv + ReLU (+ 2x2/2 pool) against
    the oracle on the c_live-channel problem.  28x28 + pool: the pool does not fold (28 % 16 != 0), conv_stem2.cu then the pool kernel; the others conv_stem2.cu with the pool folded."""
    import torch
    from tests.gpu_util import to_desc, dev, nhwc, nchw
    n, cl, hw, oc, k, stride, pad, pool = cfg
    s = shape(n, cl, hw, hw, oc, k, stride=stride, pad=pad)
    rng = np.random.default_rng(hw * oc + cl)
    x = _rand(rng, (n, cl, hw, hw)); w = _rand(rng, (oc, cl, k, k), -60, 60)
    want_y, want_e, _ = oracle.conv_fwd(s, x, w, -7, -6)
    want_y = oracle.relu(want_y)
    oh = want_y.shape[2]
    i8 = dict(dtype=torch.int8, device="cuda:0")
    x16 = torch.zeros((n, hw, hw, 16), **i8); x16[..., :cl] = dev(nhwc(x))
    w16 = torch.zeros((oc, k, k, 16), **i8); w16[..., :cl] = dev(nhwc(w))
    w16[..., cl:] = 5                                   # dead channels of w meet zeros of x: must not matter
    s16 = shape(n, 16, hw, hw, oc, k, stride=stride, pad=pad)
    y = torch.full((n, oh, oh, oc), 99, **i8)
    yp = torch.full((n, oh // 2, oh // 2, oc), 99, **i8) if pool else None
    ey = torch.zeros(1, **i8); mx = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    acc = torch.empty(y.numel(), dtype=torch.int32, device="cuda:0")
    nb.conv_stem_fwd(to_desc(s16), cl, x16, w16, torch.full((1,), -7, **i8), torch.full((1,), -6, **i8), y, ey, True, yp, acc, mx)
    torch.cuda.synchronize()
    assert np.array_equal(nchw(y.cpu().numpy()), want_y)
    assert int(ey.item()) == want_e
    if pool:
        assert np.array_equal(nchw(yp.cpu().numpy()), oracle.maxpool(want_y, 2, 2))
    if cl <= 4 and oc <= 128:
        assert nb.last_path == "stem2"


@pytest.mark.parametrize("cfg", POOL_SHAPES)
@pytest.mark.parametrize("gate", [True, False])
def test_conv_dx_fused_poolgrad_vs_oracle(nb, oracle, cfg, gate):
    """NITI_DeConv_Int8 + NITI_CPUPoolGrad_Int8 (+ NITI_CPUReluGrad_Int8) behind one call.  The pre-pool tensor has many
    ties (values in [-3, 3]) so the 'first maximum in row-major order takes the gradient' rule is exercised."""
    import torch
    from tests.gpu_util import to_desc, dev, nhwc, nchw
    n, c, hw, oc = cfg                                  # the conv's input is the POOLED tensor [n, c, hw, hw]
    s = shape(n, c, hw, hw, oc, 3, stride=1, pad=1)
    rng = np.random.default_rng(hw * oc + 1)
    dy = _rand(rng, (n, oc, hw, hw), -50, 50); w = _rand(rng, (oc, c, 3, 3), -40, 40)
    xpre = _rand(rng, (n, c, 2 * hw, 2 * hw), -3, 3)
    dxp, _ = oracle.conv_dx(s, dy, w)
    want = oracle.maxpool_grad(xpre, oracle.maxpool(xpre, 2, 2), dxp, 2, 2)
    if gate:
        want = oracle.relu_grad(xpre, want)
    i8 = dict(dtype=torch.int8, device="cuda:0")
    wd = dev(nhwc(w)); wt = wd.permute(3, 1, 2, 0).contiguous()
    out = torch.full((n, 2 * hw, 2 * hw, c), 77, **i8)
    scratch = torch.empty((n, hw, hw, c), **i8)
    mx = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    acc = torch.empty(scratch.numel(), dtype=torch.int32, device="cuda:0")
    nb.conv_dx_fused_poolgrad(to_desc(s), dev(nhwc(dy)), wd, wt, out, dev(nhwc(xpre)), gate, scratch, acc, mx)
    torch.cuda.synchronize()
    assert np.array_equal(nchw(out.cpu().numpy()), want)


def test_chain_of_forward_ops_equals_the_ops_one_by_one(nb, oracle):
    """niti_chain_run: three forward convs (+ ReLU, + pools) in one persistent launch against the oracle's op sequence;
    layers differ in tile shape (BN 128 / 64), K-block (64 / 128 channels) and pipeline depth, one pair idles in layer 3."""
    import torch
    from tests.gpu_util import to_desc, dev, nhwc, nchw
    n = 8
    rng = np.random.default_rng(42)
    x = _rand(rng, (n, 64, 16, 16))
    ws = [_rand(rng, (128, 64, 3, 3), -40, 40), _rand(rng, (128, 128, 3, 3), -30, 30), _rand(rng, (64, 128, 1, 1), -50, 50)]
    shapes = [shape(n, 64, 16, 16, 128, 3, pad=1), shape(n, 128, 8, 8, 128, 3, pad=1), shape(n, 128, 8, 8, 64, 1, pad=0)]
    pools = [True, False, True]
    # oracle
    cur, e = x, -7
    want = []
    for s, w, p in zip(shapes, ws, pools):
        y, e, _ = oracle.conv_fwd(s, cur, w, e, -8)
        y = oracle.relu(y)
        yp = oracle.maxpool(y, 2, 2) if p else None
        want.append((y, yp, e))
        cur = yp if p else y
    # device
    i8 = dict(dtype=torch.int8, device="cuda:0")
    xin = dev(nhwc(x)); ex = torch.full((1,), -7, **i8); ew = torch.full((1,), -8, **i8)
    mx = torch.zeros(4, dtype=torch.int32, device="cuda:0")
    ops, outs, keep = [], [], []
    cur_t, cur_e = xin, ex
    for i, (s, w, p) in enumerate(zip(shapes, ws, pools)):
        d = to_desc(s)
        wd = dev(nhwc(w)); keep.append(wd)
        y = torch.empty((n, d.oh, d.ow, s.oc), **i8)
        yp = torch.empty((n, d.oh // 2, d.ow // 2, s.oc), **i8) if p else None
        ey = torch.zeros(1, **i8)
        op = nb.chain_op(d, 0, cur_t, wd, y, mx[i:i + 1], i, ex=cur_e, ew=ew, ey=ey, relu=True, out_pool=yp)
        assert nb.chain_op_eligible(op)
        ops.append(op); outs.append((y, yp, ey))
        cur_t, cur_e = (yp if p else y), ey
    ch = nb.chain_create(ops)
    for rep in range(2):                                 # run twice: barrier re-arm + phase bits across launches
        mx.zero_()
        nb.chain_run(ch)
    torch.cuda.synchronize()
    nb.chain_destroy(ch)
    for (y, yp, ey), (wy, wyp, we) in zip(outs, want):
        assert np.array_equal(nchw(y.cpu().numpy()), wy)
        if wyp is not None:
            assert np.array_equal(nchw(yp.cpu().numpy()), wyp)
        assert int(ey.item()) == we
    assert nb.sync_error() == 0


# ------------------------------------------------------------------ exponent prediction (niti_set_speculate)
@pytest.mark.parametrize("cfg", [(64, 64, 16, 16, 128, 3, 1, 1), (4, 128, 4, 4, 256, 3, 1, 1), (256, 512, 1, 1, 512, 1, 1, 0)])
def test_exponent_prediction_hit_miss_and_off_are_bit_identical(nb, run, oracle, cfg):
    """The fused kernels store with the shift predicted from the slot's previous max|acc| and re-store on a miss
    (conv_chain.cu).  Same op, same slot: prediction off (two sweeps) | on after a tensor of another magnitude (miss: re-store)
    | on again (hit: single sweep) | on after a SMALLER tensor (miss the other way) — y, exponent and max|acc| always the oracle's."""
    n, c, h, w, oc, k, st, p = cfg
    s = shape(n, c, h, w, oc, k, stride=st, pad=p)
    rng = np.random.default_rng(7)
    x = _rand(rng, (n, c, h, w)); wt = _rand(rng, (oc, c, k, k), -64, 64)
    x_small = _rand(rng, (n, c, h, w), -2, 2)
    dy = _rand(rng, (n, oc, s.oh, s.ow), -50, 50)
    oy, oe, oacc = oracle.conv_fwd(s, x, wt, -3, -9)
    oys, oes, _ = oracle.conv_fwd(s, x_small, wt, -3, -9)
    odx, _ = oracle.conv_dx(s, dy, wt)
    assert oe != oes                                        # the two inputs land in different bit widths

    def check_big():
        y, e, mx = run.conv_fwd_fused(s, x, wt, -3, -9, True)
        assert e == oe and mx == int(np.abs(oacc.astype(np.int64)).max())
        assert np.array_equal(y, oracle.relu(oy))
    h0, m0, k0 = nb.speculate_stats()
    try:
        nb.set_speculate(0); check_big()
        assert nb.speculate_stats() == (h0, m0, k0 + 1)
        nb.set_speculate(2)                                                    # always speculate: forces the miss paths
        y, e, _ = run.conv_fwd_fused(s, x_small, wt, -3, -9, False)          # leaves a small maximum in the slot
        assert e == oes and np.array_equal(y, oys)
        check_big()                                                            # miss (predicted shift too small)
        check_big()                                                            # hit
        y, e, _ = run.conv_fwd_fused(s, x_small, wt, -3, -9, False)          # miss (predicted shift too large)
        assert e == oes and np.array_equal(y, oys)
        for _ in range(2):                                                     # dX: miss, then hit
            dx, _ = run.conv_dx_fused(s, dy, wt, gate=x)
            assert np.array_equal(dx, oracle.relu_grad(x, odx))
        h1, m1, k1 = nb.speculate_stats()
        # every call above speculated and most missed (which calls hit depends on the predictors' state: the input-scale
        # predictor assumes the previous fused op produced this op's operand, which this sequence violates on purpose);
        # the repeated calls below settle into hits
        assert (h1 - h0) + (m1 - m0) == 6 and m1 - m0 >= 3 and k1 - k0 == 1
        nb.set_speculate(1)                                                    # adaptive: single-tile layers wait for two holds,
        for _ in range(7):                                                     # and the 2-bit predictor may need two steps to move
            check_big()
        h2, m2, k2 = nb.speculate_stats()
        assert h2 + m2 + k2 == h1 + m1 + k1 + 7 and m2 - m1 <= 3 and h2 - h1 >= 1
    finally:
        nb.set_speculate(1)
    assert nb.sync_error() == 0


# ------------------------------------------------------------------ streaming layers (outputs larger than tensor memory)
@pytest.mark.parametrize("shape_", [(64, 64, 56, 64, 1), (128, 128, 28, 128, 1), (64, 256, 28, 256, 1), (64, 64, 56, 128, 2)])
def test_streaming_fused_ops_vs_pinned_simt_and_oracle_rescale(nb, oracle, shape_):
    """Outputs of 6-13 M accumulators per launch exceed what the CTA pairs can keep in tensor memory (ResNet-18's 56^2 / 28^2
    stages): the fused kernel recycles its accumulator slots, stores with the predicted exponent and RECOMPUTES the layer on a
    miss (conv_chain.cu).  Forward (+ReLU) and dX (+gate; the stride-2 case through zero insertion) in all three schedules —
    prediction off (max pass, then a second contraction), miss (same), hit (one contraction) — against the pinned SIMT
    accumulators rescaled by the CPU oracle.  BN 64 / 128 / 256."""
    import torch
    from tests.gpu_util import PATH_SIMT
    from mandheling_dsp_training_b200.lib import desc
    n, c, hw, oc, st = shape_
    torch.manual_seed(hw + oc)
    d = desc(n, c, hw, hw, oc, 3, stride=st, pad=1)
    i8 = dict(dtype=torch.int8, device="cuda"); i32 = dict(dtype=torch.int32, device="cuda")
    x = torch.randint(0, 128, (n, hw, hw, c), **i8)
    w = torch.randint(-30, 31, (oc, 3, 3, c), **i8)
    dy = torch.randint(-40, 41, (n, d.oh, d.ow, oc), **i8)
    gate = torch.randint(-5, 6, (n, hw, hw, c), **i8)
    wt = w.permute(3, 1, 2, 0).contiguous()
    nb.set_path(PATH_SIMT)
    acc = torch.empty((n, d.oh, d.ow, oc), **i32); dxa = torch.empty((n, hw, hw, c), **i32)
    nb.conv_fwd_acc(d, x, w, acc); nb.conv_dx_acc(d, dy, w, dxa)
    torch.cuda.synchronize()
    nb.set_path(0)
    want_y, inc = oracle.rescale_act(acc.cpu().numpy()); want_y = np.maximum(want_y, 0)
    want_dx, _ = oracle.rescale_act(dxa.cpu().numpy())
    want_dx = np.where(gate.cpu().numpy() > 0, want_dx, 0).astype(np.int8)
    ex = torch.full((1,), -7, **i8); ew = torch.full((1,), -9, **i8)
    small = torch.randint(-1, 2, (n, hw, hw, c), **i8)
    try:
        for mode, what in ((0, "off"), (2, "miss"), (2, "hit"), (1, "adaptive")):
            nb.set_speculate(mode)
            if what == "miss":                       # leave another magnitude in both slots
                y = torch.empty((n, d.oh, d.ow, oc), **i8); ey = torch.zeros(1, **i8); mx = torch.zeros(2, **i32)
                nb.conv_fwd_fused(d, small, w, ex, ew, y, ey, True, None, mx[0:1], dp_slot=7)
                nb.conv_fwd_fused(d, small, w, ex, ew, y, ey, True, None, mx[0:1], dp_slot=8)
            y = torch.full((n, d.oh, d.ow, oc), 55, **i8); ey = torch.zeros(1, **i8); mx = torch.zeros(2, **i32)
            nb.conv_fwd_fused(d, x, w, ex, ew, y, ey, True, None, mx[0:1], dp_slot=7)
            assert nb.last_path == "fused"
            dx = torch.full((n, hw, hw, c), 55, **i8)
            nb.conv_dx_fused(d, dy, w, wt, dx, gate, None, mx[1:2], dp_slot=8)
            assert nb.last_path == "fused"
            torch.cuda.synchronize()
            assert np.array_equal(y.cpu().numpy(), want_y), what
            assert int(ey.item()) == -16 + inc and int(mx[0].item()) == int(acc.abs().max().item()), what
            assert np.array_equal(dx.cpu().numpy(), want_dx), what
            assert int(mx[1].item()) == int(dxa.abs().max().item()), what
    finally:
        nb.set_speculate(1)
    assert nb.sync_error() == 0


@pytest.mark.parametrize("name", ["rgb_stem_3x3_p1", "rgb_stem_3x3_p1_small", "rgb_stem_7x7_s2_p3"])
def test_rgb_stem_golden_reference_vectors(nb, run, golden_rgb, name):
    """a1/a3 on 3-channel image layers against outputs of the unmodified reference (tests/golden/conv_rgb_cases.npz): the
    generic C-ABI ops on the 3-channel tensors, then the product route — niti_conv_stem_fwd and the dW kernels on the image
    zero-padded to 16 channels, as the training step stores it."""
    import torch
    from tests.gpu_util import PATH_AUTO, to_desc, dev, nhwc, nchw
    nb.set_path(PATH_AUTO)
    g = golden_rgb
    s, x, w, dy = _case(g, name)
    y, e = run.conv_fwd(s, x, w, -7, -7)
    assert np.array_equal(y, g[name + ".y"]) and e == int(g[name + ".exp"])
    assert np.array_equal(run.conv_dw(s, x, dy), g[name + ".dw"])
    # product route: 16-channel storage, 3 live
    i8 = dict(dtype=torch.int8, device="cuda:0")
    x16 = torch.zeros((s.n, s.h, s.w, 16), **i8); x16[..., :3] = dev(nhwc(x))
    w16 = torch.zeros((s.oc, s.kh, s.kw, 16), **i8); w16[..., :3] = dev(nhwc(w))
    s16 = shape(s.n, 16, s.h, s.w, s.oc, s.kh, stride=s.stride, pad=s.pad)
    yd = torch.full((s.n, s.oh, s.ow, s.oc), 99, **i8)
    ey = torch.zeros(1, **i8); mx = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    acc = torch.empty(yd.numel(), dtype=torch.int32, device="cuda:0")
    nb.conv_stem_fwd(to_desc(s16), 3, x16, w16, torch.full((1,), -7, **i8), torch.full((1,), -7, **i8), yd, ey, False, None, acc, mx)
    torch.cuda.synchronize()
    assert np.array_equal(nchw(yd.cpu().numpy()), g[name + ".y"]) and int(ey.item()) == int(g[name + ".exp"])
    dw16 = run.conv_dw(s16, nchw(x16.cpu().numpy()), dy)
    assert np.array_equal(dw16[:, :3], g[name + ".dw"]) and not dw16[:, 3:].any()
