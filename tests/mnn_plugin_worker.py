"""Worker for tests/test_gpu_mnn_backend.py — runs in its OWN process, without torch.

The unmodified reference (oracle/_ref/libMNN_ref.so) has to be loaded RTLD_GLOBAL so the MNN plug-in's core
symbols resolve against it, exactly as in a real MNN host process; doing that inside the pytest process clashes
with libtorch's own copies of common third-party symbols, hence the subprocess.

Evaluates every case twice through the reference's express API / OpGrad / Module / NITI_SGD — first on the
reference CPU backend, then on the B200 plug-in (MNN_FORWARD_USER_0) — and writes both result sets to an .npz."""
import ctypes
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.oracle import Oracle, Reference, shape          # noqa: E402  (numpy + ctypes only)

PLUGIN = os.path.join(ROOT, "mnn_backend", "libMNN_niti_b200.so")
MNN_FORWARD_USER_0 = 8          # include/MNN/MNNForwardType.h:38
CONV_CASES = [(4, 4, 8, 8, 8, 3, 1, 1), (2, 8, 6, 6, 12, 3, 1, 0), (2, 20, 12, 12, 52, 5, 1, 0),
              (8, 16, 1, 1, 12, 1, 1, 0), (3, 12, 7, 7, 16, 3, 1, 1), (16, 64, 16, 16, 128, 3, 1, 1),
              (4, 8, 10, 10, 8, 3, 2, 1)]
LENET = [("conv", 20, 5, 1, 0), ("relu",), ("pool", 2, 2), ("conv", 52, 5, 1, 0), ("relu",), ("pool", 2, 2),
         ("flatten",), ("conv", 500, 1, 1, 0), ("relu",), ("conv", 12, 1, 1, 0)]      # demo/mnistTrain.cpp:131-181


MATMUL_CASES = [(36, 64, 8), (100, 128, 12), (64, 832, 500), (52, 100, 16)]           # (e, l, h)
SPLIT_CASES = [(4, 8, 16, 48, 3, 1), (2, 4, 20, 48, 3, 1), (2, 8, 12, 64, 3, 1)]      # (n, c, h, oc, k, pad)


def xavier_int8(rng, oc, c, k):
    w = rng.normal(0.0, np.sqrt(2.0 / ((c + oc) * k * k)), size=(oc, c, k, k)).astype(np.float32)
    r = float(np.abs(w).max())
    return np.round(w / r * 127.0).astype(np.int8), int(np.ceil(np.log2(r))) - 7


def evaluate(ref, tag, out, steps, batch):
    for i, cfg in enumerate(CONV_CASES):
        n, c, h, w, oc, k, st, p = cfg
        s = shape(n, c, h, w, oc, k, stride=st, pad=p)
        rng = np.random.default_rng(i)
        x = rng.integers(-127, 128, size=(n, c, h, w), dtype=np.int8)
        wt = rng.integers(-64, 65, size=(oc, c, k, k), dtype=np.int8)
        dy = rng.integers(-40, 41, size=(n, oc, s.oh, s.ow), dtype=np.int8)
        y, e = ref.conv_fwd(s, x, wt, -5, -9)
        dx, dw = ref.conv_bwd(s, x, wt, dy)
        out[f"{tag}_fwd{i}_y"], out[f"{tag}_fwd{i}_e"] = y, np.int32(e)
        out[f"{tag}_bwd{i}_dx"], out[f"{tag}_bwd{i}_dw"] = dx, dw
    # a4: OpType_NITI_MatMul_Int8 (raw buffers, the execution's own packed addressing) and
    # a5: OpType_NITI_GradientSplitCONV_Int8 (slice counts 2, 4, 1), created directly (ref_driver.cpp)
    for i, (e, l, h) in enumerate(MATMUL_CASES):
        rng = np.random.default_rng(200 + i)
        a_raw = rng.integers(-127, 128, size=e * l, dtype=np.int8)
        b = rng.integers(-90, 91, size=(h, l), dtype=np.int8)
        out[f"{tag}_matmul{i}"] = ref.matmul_raw(a_raw, b, e, l, h)
    for i, (n, c, hh, oc, k, p) in enumerate(SPLIT_CASES):
        s = shape(n, c, hh, hh, oc, k, stride=1, pad=p)
        rng = np.random.default_rng(300 + i)
        x = rng.integers(-127, 128, size=(n, c, hh, hh), dtype=np.int8)
        dy = rng.integers(-50, 51, size=(n, oc, s.oh, s.ow), dtype=np.int8)
        dy[:, oc // 2:] //= 4
        out[f"{tag}_split{i}"] = ref.grad_conv(s, x, dy, split=True)
        out[f"{tag}_nosplit{i}"] = ref.grad_conv(s, x, dy, split=False)
    rng = np.random.default_rng(7)
    a = rng.integers(-127, 128, size=(4, 8, 8, 8), dtype=np.int8)
    g = rng.integers(-127, 128, size=(4, 8, 8, 8), dtype=np.int8)
    out[f"{tag}_relu"] = ref.relu(a)
    out[f"{tag}_relu_grad"] = ref.relu_grad(a, g)
    out[f"{tag}_maxpool"] = ref.maxpool(a, 2, 2)
    # f1 / f2 rows: pool gradient (non-overlapping and overlapping windows, many ties), int8 Binary SUB (the SGD update), loss gradient
    gp = rng.integers(-127, 128, size=(4, 8, 4, 4), dtype=np.int8)
    out[f"{tag}_pool_grad"] = ref.maxpool_grad(a, gp, 2, 2)
    at = rng.integers(-2, 3, size=(3, 12, 9, 9), dtype=np.int8)
    out[f"{tag}_pool_grad_k3s2"] = ref.maxpool_grad(at, rng.integers(-127, 128, size=(3, 12, 4, 4), dtype=np.int8), 3, 2)
    out[f"{tag}_binary_sub"] = ref.sgd_update(rng.integers(-127, 128, size=5000, dtype=np.int8), rng.integers(-127, 128, size=5000, dtype=np.int8))
    for j, asc in enumerate((-3, -9, 1)):
        lg = rng.integers(-127, 128, size=(64, 12), dtype=np.int8)
        out[f"{tag}_loss_grad{j}"] = ref.loss_grad(lg, asc, rng.integers(0, 10, size=64))
    rng = np.random.default_rng(1)
    ws = [xavier_int8(rng, oc, ic, k) for oc, ic, k in [(20, 1, 5), (52, 20, 5), (500, 832, 1), (12, 500, 1)]]
    rng = np.random.default_rng(0)
    x = rng.integers(-127, 128, size=(steps, batch, 1, 28, 28), dtype=np.int8)
    lab = rng.integers(0, 10, size=(steps, batch)).astype(np.int32)
    r = ref.net_train(LENET, x, -7, lab, ws, tclasses=10)
    for k in ("logits", "e_logits", "loss", "ms"):
        out[f"{tag}_lenet_{k}"] = r[k]
    for i, w in enumerate(r["weights"]):
        out[f"{tag}_lenet_w{i}"] = w


# DSP-named op types: (n, c, h, w, oc, k, pad) stride-1 layers
DSP_CASES = [(2, 16, 8, 8, 32, 3, 1), (4, 64, 8, 8, 128, 3, 1), (3, 8, 6, 6, 12, 3, 0), (8, 32, 1, 1, 16, 1, 0)]


def evaluate_dsp(ref, out):
    """The DSP-named op types on the B200 plug-in against the oracle's restatement of the DSP arithmetic (the reference's
    own executions for them need the Hexagon FastRPC library, so there is no CPU leg).  Operands in the forms the
    reference's grad rule emits (grad/NITI_DSPConv_Int8_Grad.cpp:36-222): NHWC, weights [KH,KW,C,OC], x transposed
    {3,1,2,0} for the TRANSPOSEGRADIENT / SPLITBatch variants."""
    orc = Oracle()
    bad = []
    for i, (n, c, h, w, oc, k, p) in enumerate(DSP_CASES):
        s = shape(n, c, h, w, oc, k, stride=1, pad=p)
        rng = np.random.default_rng(100 + i)
        x = rng.integers(-127, 128, size=(n, c, h, w), dtype=np.int8)
        wt = rng.integers(-64, 65, size=(oc, c, k, k), dtype=np.int8)
        dy = rng.integers(-40, 41, size=(n, oc, s.oh, s.ow), dtype=np.int8)
        x_nhwc = np.ascontiguousarray(x.transpose(0, 2, 3, 1)); dy_nhwc = np.ascontiguousarray(dy.transpose(0, 2, 3, 1))
        w_hwio = np.ascontiguousarray(wt.transpose(2, 3, 1, 0))
        acc, _ = orc.conv_fwd_acc(s, x, wt)
        want, shift = orc.rescale_dsp(acc, 1)
        for kind in (0, 1):                                   # DSP_CONV, DSP_DECONV: forward-convolution form
            y, e = ref.dsp_op(kind, (k, k, 1, p, c, oc), x_nhwc, w_hwio, -5, -9, out_elems=want.size)
            ok = y.shape == (n, s.oh, s.ow, oc) and np.array_equal(y, want.transpose(0, 2, 3, 1)) and e == -14 + shift
            out[f"dsp_conv{kind}_{i}"] = np.int32(ok)
            if not ok:
                bad.append(("conv", kind, i, y.shape, e, shift))
        dacc, _ = orc.conv_dw_acc(s, x, dy)                   # [oc, c, kh, kw]
        xT = np.ascontiguousarray(x_nhwc.transpose(3, 1, 2, 0))
        # 6 = DSP_MATMUL_GRADIENT (the north star's op), 7 = DSP_MATMUL: operands as DSP_GRADIENTCONV (same size computer)
        for kind, div, xin in ((2, 32, xT), (3, 16, x_nhwc), (4, 16, xT), (5, 16, x_nhwc), (6, 16, x_nhwc), (7, 16, x_nhwc)):
            wantw, shiftw = orc.rescale_dsp(dacc, div)
            dw, e = ref.dsp_op(kind, (s.oh, s.ow, 1, p, n, oc), xin, dy_nhwc, 0, 0, out_elems=wantw.size)
            ok = dw.shape == (k, k, c, oc) and np.array_equal(dw, wantw.transpose(2, 3, 1, 0)) and e == shiftw
            out[f"dsp_dw{kind}_{i}"] = np.int32(ok)
            if not ok:
                bad.append(("dw", kind, i, dw.shape, e, shiftw))
    out["n_dsp"] = np.int32(len(DSP_CASES))
    print("dsp ops: mismatches", bad)


def main():
    dst = os.path.abspath(sys.argv[1])          # Reference() changes the cwd (DSP.txt / parallel.txt live there)
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    batch = int(sys.argv[3]) if len(sys.argv) > 3 else 64
    out = {"n_conv": np.int32(len(CONV_CASES)), "n_matmul": np.int32(len(MATMUL_CASES)), "n_split": np.int32(len(SPLIT_CASES))}
    ref = Reference(threads=4)
    if len(sys.argv) > 4 and sys.argv[4] == "dsp":          # DSP-named op types only (own process: no CPU leg exists)
        ref.use_backend(MNN_FORWARD_USER_0, PLUGIN)
        plug = ctypes.CDLL(PLUGIN)
        plug.MNNNitiB200LaunchCount.restype = ctypes.c_longlong
        before = plug.MNNNitiB200LaunchCount()
        evaluate_dsp(ref, out)
        out["launches"] = np.int64(plug.MNNNitiB200LaunchCount() - before)
        np.savez(dst, **out)
        print("worker ok: dsp launches", int(out["launches"]))
        return
    evaluate(ref, "cpu", out, steps, batch)
    ref.use_backend(MNN_FORWARD_USER_0, PLUGIN)
    plug = ctypes.CDLL(PLUGIN)
    plug.MNNNitiB200LaunchCount.restype = ctypes.c_longlong
    before = plug.MNNNitiB200LaunchCount()
    evaluate(ref, "b200", out, steps, batch)
    out["launches"] = np.int64(plug.MNNNitiB200LaunchCount() - before)
    plug.MNNNitiB200CreatedOps.restype = ctypes.c_longlong
    # Executions the plug-in created per op type (schema/current/MNN_generated.h OpType values)
    for name, code in (("binary", 7), ("pool_grad", 706), ("loss_grad", 711), ("loss", 710), ("pad", 714), ("left_pool_grad", 718)):
        out[f"created_{name}"] = np.int64(plug.MNNNitiB200CreatedOps(code))
    # op types that still ran on the executor's CPU backup backend (reported, not asserted: Raster / Convert / shape ops)
    plug.MNNNitiB200DeclinedOps.restype = ctypes.c_longlong
    buf = (ctypes.c_longlong * 128)()
    nd = int(plug.MNNNitiB200DeclinedOps(-1, buf, 64))
    print("declined op types {type: count}:", {int(buf[2 * i]): int(buf[2 * i + 1]) for i in range(min(nd, 64))})
    np.savez(dst, **{k: (v if v is not None else np.zeros(0, np.int8)) for k, v in out.items()})
    same = [k[4:] for k in out if k.startswith("cpu_") and not k.endswith("_ms") and out[k] is not None
            and np.array_equal(out[k], out["b200_" + k[4:]])]
    diff = [k[4:] for k in out if k.startswith("cpu_") and not k.endswith("_ms") and out[k] is not None
            and not np.array_equal(out[k], out["b200_" + k[4:]])]     # the float loss may differ in its last bits
    print("identical:", len(same), "different:", diff)
    print("worker ok: launches", int(out["launches"]), "cpu ms/step", out["cpu_lenet_ms"], "b200 ms/step", out["b200_lenet_ms"])


if __name__ == "__main__":
    main()
