"""Generate tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref, built from
/root/reference by oracle/Makefile.ref) through oracle/ref_driver.cpp on seeded synthetic tensors.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
The committed .npz files carry inputs and reference outputs, so the oracle (and through it the CUDA
path) stays pinned to the reference on boxes where /root/reference does not exist.
"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle import Reference, Oracle, shape  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

# name: (n, c, h, w, oc, k, stride, pad, x_lo, x_hi, w_absmax, dy_absmax)
CONV_CASES = {
    # MNIST NITIInt8 layers (demo/mnistTrain.cpp:131-157), reduced batch
    "mnist_conv1": (4, 1, 28, 28, 20, 5, 1, 0, -127, 127, 127, 40),
    "mnist_conv2": (4, 20, 12, 12, 52, 5, 1, 0, 0, 127, 127, 40),
    "mnist_ip1": (64, 832, 1, 1, 500, 1, 1, 0, 0, 127, 127, 40),
    "mnist_ip2": (64, 500, 1, 1, 12, 1, 1, 0, 0, 127, 127, 40),
    # VGG-style 3x3 pad 1
    "vgg_3x3_p1": (4, 16, 8, 8, 32, 3, 1, 1, -127, 127, 100, 60),
    "vgg_3x3_p1_b": (2, 64, 4, 4, 64, 3, 1, 1, 0, 127, 127, 127),
    "odd_hw_3x3_p1": (3, 8, 7, 5 + 2, 12, 3, 1, 1, -127, 127, 64, 30),
    # ResNet-style strided (the reference lowers stride 2 through _NITI_LeftPoolGrad)
    "s2_3x3_p1": (4, 4, 8, 8, 8, 3, 2, 1, -127, 127, 127, 50),
    "s2_1x1_p0": (2, 8, 8, 8, 8, 1, 2, 0, -127, 127, 127, 50),
    # rescale edge cases: shift<=0 (raw cast), shift==1 (shifts by 2), tiny gradients
    "small_mag_cast": (2, 4, 4, 4, 4, 1, 1, 0, -3, 3, 3, 1),
    "small_mag_shift1": (2, 4, 4, 4, 4, 1, 1, 0, 5, 9, 7, 2),   # |w| used: 128 < max|acc| <= 256
    "zero_dy": (2, 4, 6, 6, 8, 3, 1, 1, -127, 127, 127, 0),
}

# 3-channel image stems (VGG-11 / ResNet-18 first layer; C % 4 != 0, so forward + dW only — dX is outside the reference's
# contract, quirk Q6).  Written to their own file, conv_rgb_cases.npz, so conv_cases.npz keeps its bytes.
RGB_CASES = {
    "rgb_stem_3x3_p1": (2, 3, 32, 32, 64, 3, 1, 1, -127, 127, 127, 40),
    "rgb_stem_3x3_p1_small": (4, 3, 16, 16, 32, 3, 1, 1, -127, 127, 90, 20),
    "rgb_stem_7x7_s2_p3": (2, 3, 32, 32, 64, 7, 2, 3, -127, 127, 127, 40),
}


def tensors(name, case):
    n, c, h, w, oc, k, st, p, xlo, xhi, wmax, dymax = case
    s = shape(n, c, h, w, oc, k, stride=st, pad=p)
    seed = int.from_bytes(name.encode(), "little") % (2 ** 31)
    rng = np.random.default_rng(seed)
    x = rng.integers(xlo, xhi + 1, size=(n, c, h, w), dtype=np.int8)
    wt = rng.integers(-wmax, wmax + 1, size=(oc, c, k, k), dtype=np.int8)
    dy = rng.integers(-dymax, dymax + 1, size=(n, oc, s.oh, s.ow), dtype=np.int8)
    if name == "small_mag_shift1":
        wt = np.maximum(np.abs(wt), 4).astype(np.int8)
    return s, x, wt, dy


def main_rgb(ref, orc):
    conv = {}
    for name, case in RGB_CASES.items():
        s, x, wt, dy = tensors(name, case)
        y, e = ref.conv_fwd(s, x, wt, -7, -7)
        _, dw = ref.conv_bwd(s, x, wt, dy)
        conv[name + ".shape"] = np.array([getattr(s, f) for f, _ in s._fields_], np.int32)
        conv[name + ".x"], conv[name + ".w"], conv[name + ".dy"] = x, wt, dy
        conv[name + ".y"], conv[name + ".exp"], conv[name + ".dw"] = y, np.int32(e), dw
        worst = max(orc.conv_fwd_acc(s, x, wt)[1], orc.conv_dw_acc(s, x, dy)[1])
        assert worst < 2 ** 24, (name, worst)
    np.savez_compressed(os.path.join(OUT, "conv_rgb_cases.npz"), **conv)


def main():
    ref = Reference(threads=4)
    orc = Oracle()
    conv = {}
    for name, case in CONV_CASES.items():
        s, x, wt, dy = tensors(name, case)
        y, e = ref.conv_fwd(s, x, wt, -7, -7)
        dx, dw = ref.conv_bwd(s, x, wt, dy)
        conv[name + ".shape"] = np.array([getattr(s, f) for f, _ in s._fields_], np.int32)
        conv[name + ".x"], conv[name + ".w"], conv[name + ".dy"] = x, wt, dy
        conv[name + ".y"], conv[name + ".exp"] = y, np.int32(e)
        conv[name + ".dw"] = dw
        # dX of a layer whose input-channel count is not a multiple of 4 is outside the reference's
        # contract (quirk Q6: acc buffer sized without the roundup-4 the GEMM writes).
        if case[1] % 4 == 0:
            conv[name + ".dx"] = dx
        # Q1: the reference accumulates in float; record that every accumulator stays exact.
        worst = max(orc.conv_fwd_acc(s, x, wt)[1], orc.conv_dw_acc(s, x, dy)[1])
        assert worst < 2 ** 24, (name, worst)
    np.savez_compressed(os.path.join(OUT, "conv_cases.npz"), **conv)
    main_rgb(ref, orc)

    rng = np.random.default_rng(7)
    ew = {}
    x = rng.integers(-127, 128, size=(4, 8, 12, 12), dtype=np.int8)
    dyf = rng.integers(-127, 128, size=x.shape, dtype=np.int8)
    ew["relu.x"], ew["relu.y"] = x, ref.relu(x)
    ew["relu_grad.dy"], ew["relu_grad.dx"] = dyf, ref.relu_grad(x, dyf)
    ew["maxpool.y"] = ref.maxpool(x, 2, 2)
    w = rng.integers(-128, 128, size=(4096,), dtype=np.int8)
    dw = rng.integers(-128, 128, size=(4096,), dtype=np.int8)
    ew["sgd.w"], ew["sgd.dw"], ew["sgd.out"] = w, dw, ref.sgd_update(w, dw)
    for a in (-9, -7, -5, -3, 0, 1):
        lg = rng.integers(-127, 128, size=(64, 12), dtype=np.int8)
        lab = rng.integers(0, 10, size=64).astype(np.int32)
        ew[f"loss_grad.{a}.logits"], ew[f"loss_grad.{a}.labels"] = lg, lab
        ew[f"loss_grad.{a}.grad"] = ref.loss_grad(lg, a, lab)
        ew[f"loss_value.{a}"] = np.float32(ref.loss_value(lg, a, lab))
    np.savez_compressed(os.path.join(OUT, "elementwise_cases.npz"), **ew)

    # ---- OpType_NITI_MatMul_Int8 (a4) and OpType_NITI_GradientSplitCONV_Int8 (a5): neither has a producer the demo
    # reaches, so the driver creates the ops directly (ref_matmul / ref_grad_conv in oracle/ref_driver.cpp)
    mm = {}
    rng = np.random.default_rng(11)
    for i, (e, l, h, amax, bmax) in enumerate([(36, 64, 8, 127, 60), (100, 128, 12, 127, 127), (108, 256, 20, 90, 40),
                                               (64, 832, 500, 127, 127), (16, 16, 4, 2, 1), (8, 4, 4, 1, 1)]):
        a_raw = rng.integers(-amax, amax + 1, size=e * l, dtype=np.int8)
        b = rng.integers(-bmax, bmax + 1, size=(h, l), dtype=np.int8)
        if i == 5:
            a_raw[:] = 0                                  # b == 0 -> zeros (Q2)
        mm[f"mm{i}.dims"] = np.array([e, l, h], np.int32)
        mm[f"mm{i}.a_raw"], mm[f"mm{i}.b"] = a_raw, b
        mm[f"mm{i}.c_raw"] = ref.matmul_raw(a_raw, b, e, l, h)
        assert orc.matmul_nt_acc(orc.matmul_unpack_a(a_raw, e, l), b)[1] < 2 ** 24
    np.savez_compressed(os.path.join(OUT, "matmul_cases.npz"), **mm)

    sp = {}
    rng = np.random.default_rng(13)
    # (n, c, h, oc, k, pad, dy_absmax): slice counts 2, 4, 1 (OC % 32 == 0 never splits), 2 with a zero slice
    for i, (n, c, hh, oc, k, pad, dymax) in enumerate([(4, 8, 16, 48, 3, 1, 50), (2, 4, 20, 48, 3, 1, 40),
                                                       (2, 4, 20, 96, 3, 1, 40), (2, 4, 18, 40, 3, 0, 20)]):
        s = shape(n, c, hh, hh, oc, k, stride=1, pad=pad)
        x = rng.integers(-127, 128, size=(n, c, hh, hh), dtype=np.int8)
        dy = rng.integers(-dymax, dymax + 1, size=(n, oc, s.oh, s.ow), dtype=np.int8)
        slices = orc.split_count(oc, s.oh, s.ow)
        per = oc // slices
        for sl in range(slices):
            dy[:, sl * per:(sl + 1) * per] //= (1 << (sl % 4))     # slices of different magnitude -> alignment divides
        if i == 3:
            dy[:, per:] = 0                                        # one slice with b_s == 0
        sp[f"sp{i}.shape"] = np.array([getattr(s, f) for f, _ in s._fields_], np.int32)
        sp[f"sp{i}.x"], sp[f"sp{i}.dy"] = x, dy
        sp[f"sp{i}.slices"] = np.int32(slices)
        sp[f"sp{i}.dw"] = ref.grad_conv(s, x, dy, split=True)
        sp[f"sp{i}.dw_unsplit"] = ref.grad_conv(s, x, dy, split=False)
        assert orc.conv_dw_acc(s, x, dy)[1] < 2 ** 24
    np.savez_compressed(os.path.join(OUT, "split_cases.npz"), **sp)
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    if "--rgb-only" in sys.argv:
        main_rgb(Reference(threads=4), Oracle())
    else:
        main()
