"""TEST INFRASTRUCTURE ONLY — ctypes bindings for the CPU oracle (oracle/niti_oracle.c) and, where it
was built, for the unmodified reference driven through oracle/ref_driver.cpp (oracle/_ref/).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Tensors are numpy int8 in the mathematical layouts (NCHW activations, OIHW weights).
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libniti_oracle.so")
_REF_DIR = os.path.join(_HERE, "_ref")


def build_oracle(force=False):
    """gcc the C restatement into oracle/libniti_oracle.so (seconds)."""
    src = os.path.join(_HERE, "niti_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-std=c99", "-shared", "-fPIC", src, "-o", _LIB, "-lm"])
    return _LIB


def build_ref():
    """Build the unmodified reference + driver into oracle/_ref (needs /root/reference; minutes)."""
    if not os.path.isdir("/root/reference/execution-engine"):
        return False
    subprocess.check_call(["make", "-f", "Makefile.ref", "-j8", "all"], cwd=_HERE,
                          stdout=subprocess.DEVNULL)
    return True


class ConvShape(C.Structure):
    _fields_ = [(k, C.c_int) for k in ("n", "c", "h", "w", "oc", "kh", "kw", "stride", "pad", "dil")]

    @property
    def oh(self):
        return (self.h + 2 * self.pad - (self.dil * (self.kh - 1) + 1)) // self.stride + 1

    @property
    def ow(self):
        return (self.w + 2 * self.pad - (self.dil * (self.kw - 1) + 1)) // self.stride + 1


def shape(n, c, h, w, oc, kh, kw=None, stride=1, pad=0, dil=1):
    return ConvShape(n, c, h, w, oc, kh, kh if kw is None else kw, stride, pad, dil)


_i8p = np.ctypeslib.ndpointer(np.int8, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def _c(a, dt=np.int8):
    return np.ascontiguousarray(a, dtype=dt)


class Oracle:
    def __init__(self):
        self.lib = L = C.CDLL(build_oracle())
        sp = C.POINTER(ConvShape)
        L.niti_o_conv_fwd_acc.restype = C.c_int64
        L.niti_o_conv_fwd_acc.argtypes = [sp, _i8p, _i8p, _i32p]
        L.niti_o_conv_dx_acc.restype = C.c_int64
        L.niti_o_conv_dx_acc.argtypes = [sp, _i8p, _i8p, _i32p]
        L.niti_o_conv_dw_acc.restype = C.c_int64
        L.niti_o_conv_dw_acc.argtypes = [sp, _i8p, _i8p, _i32p]
        L.niti_o_matmul_nt_acc.restype = C.c_int64
        L.niti_o_matmul_nt_acc.argtypes = [C.c_int, C.c_int, C.c_int, _i8p, _i8p, _i32p]
        L.niti_o_rescale_act.restype = C.c_int
        L.niti_o_rescale_act.argtypes = [_i32p, C.c_long, _i8p]
        L.niti_o_rescale_dsp.restype = C.c_int
        L.niti_o_rescale_dsp.argtypes = [_i32p, C.c_long, C.c_int, _i8p]
        L.niti_o_dsp_ceiling_log2.restype = C.c_int
        L.niti_o_dsp_ceiling_log2.argtypes = [C.c_uint32]
        L.niti_o_rescale_grad.restype = None
        L.niti_o_rescale_grad.argtypes = [_i32p, C.c_long, C.c_int, _i8p]
        L.niti_o_split_count.restype = C.c_int
        L.niti_o_split_count.argtypes = [C.c_int, C.c_int, C.c_int]
        L.niti_o_rescale_grad_split.restype = None
        L.niti_o_rescale_grad_split.argtypes = [_i32p, C.c_long, C.c_int, _i8p]
        L.niti_o_range_estimate.restype = C.c_int
        L.niti_o_range_estimate.argtypes = [_i32p, C.c_long]
        L.niti_o_psto_shift.restype = None
        L.niti_o_psto_shift.argtypes = [_i32p, C.c_int, _i32p, C.c_long]
        L.niti_o_sgd_update.restype = None
        L.niti_o_sgd_update.argtypes = [_i8p, _i8p, C.c_long]
        L.niti_o_relu.argtypes = [_i8p, _i8p, C.c_long]
        L.niti_o_relu_grad.argtypes = [_i8p, _i8p, _i8p, C.c_long]
        L.niti_o_maxpool.argtypes = [_i8p] + [C.c_int] * 6 + [_i8p]
        L.niti_o_maxpool_grad.argtypes = [_i8p, _i8p, _i8p] + [C.c_int] * 6 + [_i8p]
        L.niti_o_loss_grad.argtypes = [_i8p, C.c_int, _i32p, C.c_int, C.c_int, _i8p]
        L.niti_o_loss_value.argtypes = [_i8p, C.c_int, _i32p, C.c_int, C.c_int]
        L.niti_o_loss_value.restype = C.c_float

    # ---- accumulators (int32) ; each returns (acc, worst_partial_sum)
    def conv_fwd_acc(self, s, x, w):
        acc = np.empty((s.n, s.oc, s.oh, s.ow), np.int32)
        worst = self.lib.niti_o_conv_fwd_acc(C.byref(s), _c(x), _c(w), acc)
        return acc, worst

    def conv_dx_acc(self, s, dy, w):
        acc = np.empty((s.n, s.c, s.h, s.w), np.int32)
        worst = self.lib.niti_o_conv_dx_acc(C.byref(s), _c(dy), _c(w), acc)
        return acc, worst

    def conv_dw_acc(self, s, x, dy):
        acc = np.empty((s.oc, s.c, s.kh, s.kw), np.int32)
        worst = self.lib.niti_o_conv_dw_acc(C.byref(s), _c(x), _c(dy), acc)
        return acc, worst

    def matmul_nt_acc(self, a, b):
        e, l = a.shape
        h = b.shape[0]
        acc = np.empty((e, h), np.int32)
        worst = self.lib.niti_o_matmul_nt_acc(e, h, l, _c(a), _c(b), acc)
        return acc, worst

    # ---- rescale
    def range_estimate(self, acc):
        a = _c(acc, np.int32)
        return self.lib.niti_o_range_estimate(a.reshape(-1), a.size)

    def psto(self, acc, shift):
        a = _c(acc, np.int32)
        out = np.empty_like(a)
        self.lib.niti_o_psto_shift(a.reshape(-1), shift, out.reshape(-1), a.size)
        return out

    def rescale_act(self, acc):
        a = _c(acc, np.int32)
        out = np.empty(a.shape, np.int8)
        inc = self.lib.niti_o_rescale_act(a.reshape(-1), a.size, out.reshape(-1))
        return out, inc

    def rescale_dsp(self, acc, post_div=1):
        """DSP-path requantisation (parity unpinned): returns (int8, shift)"""
        a = np.ascontiguousarray(acc, dtype=np.int32)
        out = np.empty(a.shape, dtype=np.int8)
        shift = self.lib.niti_o_rescale_dsp(a.reshape(-1), a.size, int(post_div), out.reshape(-1))
        return out, int(shift)

    def rescale_grad(self, acc, sub):
        a = _c(acc, np.int32)
        out = np.empty(a.shape, np.int8)
        self.lib.niti_o_rescale_grad(a.reshape(-1), a.size, sub, out.reshape(-1))
        return out

    def split_count(self, oc, oh, ow):
        """slice count of NITI_GradientSplitConv_Int8 (.cpp:209-222)"""
        return self.lib.niti_o_split_count(int(oc), int(oh), int(ow))

    def rescale_grad_split(self, acc, slices):
        """acc: int32, `slices` equal contiguous slices along axis 0 -> int8 (per-slice PSTO(b_s - 3) + alignment)"""
        a = _c(acc, np.int32)
        assert a.size % slices == 0
        out = np.empty(a.shape, np.int8)
        self.lib.niti_o_rescale_grad_split(a.reshape(-1), a.size // slices, int(slices), out.reshape(-1))
        return out

    def conv_dw_split(self, s, x, dy):
        """NITI_GradientSplitConv_Int8: dW [oc,c,kh,kw] with the output channels cut in split_count slices"""
        acc, _ = self.conv_dw_acc(s, x, dy)
        k = self.split_count(s.oc, s.oh, s.ow)
        return self.rescale_grad_split(acc, k), k

    @staticmethod
    def matmul_unpack_a(raw, e, l):
        """A as NITI_Matmul_Int8 addresses its input buffer: [l/4][e][4] blocks (NITI_Matmul_Int8.cpp:186-196) -> [e, l]"""
        return np.ascontiguousarray(np.asarray(raw, np.int8).reshape(l // 4, e, 4).transpose(1, 0, 2).reshape(e, l))

    @staticmethod
    def matmul_pack_c(c, e, h):
        """C [e, h] -> the [h/4][e][4] blocks NITI_Matmul_Int8 writes its output buffer in"""
        return np.ascontiguousarray(np.asarray(c, np.int8).reshape(e, h // 4, 4).transpose(1, 0, 2)).reshape(-1)

    # ---- whole ops
    def conv_fwd(self, s, x, w, ex, ew):
        acc, _ = self.conv_fwd_acc(s, x, w)
        y, inc = self.rescale_act(acc)
        return y, int(np.int8(ex + ew + inc)), acc

    def conv_dx(self, s, dy, w):
        acc, _ = self.conv_dx_acc(s, dy, w)
        return self.rescale_act(acc)[0], acc

    def conv_dw(self, s, x, dy):
        acc, _ = self.conv_dw_acc(s, x, dy)
        return self.rescale_grad(acc, 2), acc

    def matmul_nt(self, a, b):
        acc, _ = self.matmul_nt_acc(a, b)
        return self.rescale_grad(acc, 3), acc

    def sgd_update(self, w, dw):
        out = _c(w).copy()
        self.lib.niti_o_sgd_update(out.reshape(-1), _c(dw).reshape(-1), out.size)
        return out

    def relu(self, x):
        x = _c(x); y = np.empty_like(x)
        self.lib.niti_o_relu(x.reshape(-1), y.reshape(-1), x.size)
        return y

    def relu_grad(self, x, dy):
        x = _c(x); dx = np.empty_like(x)
        self.lib.niti_o_relu_grad(x.reshape(-1), _c(dy).reshape(-1), dx.reshape(-1), x.size)
        return dx

    def maxpool(self, x, k, s):
        n, c, h, w = x.shape
        y = np.empty((n, c, (h - k) // s + 1, (w - k) // s + 1), np.int8)
        self.lib.niti_o_maxpool(_c(x), n, c, h, w, k, s, y)
        return y

    def maxpool_grad(self, x, y, dy, k, s):
        n, c, h, w = x.shape
        dx = np.empty_like(_c(x))
        self.lib.niti_o_maxpool_grad(_c(x), _c(y), _c(dy), n, c, h, w, k, s, dx)
        return dx

    def loss_grad(self, logits, ascale, labels):
        b, k = logits.shape
        g = np.empty((b, k), np.int8)
        self.lib.niti_o_loss_grad(_c(logits), int(ascale), _c(labels, np.int32), b, k, g)
        return g

    def loss_value(self, logits, ascale, labels):
        b, k = logits.shape
        return float(self.lib.niti_o_loss_value(_c(logits), int(ascale), _c(labels, np.int32), b, k))


class Reference:
    """The unmodified reference, via oracle/_ref/libniti_ref_driver.so (None-able: see available())."""

    @staticmethod
    def available():
        return os.path.exists(os.path.join(_REF_DIR, "libniti_ref_driver.so"))

    def __init__(self, threads=4, workdir=None):
        # The reference build exports libstdc++ symbols of its own (default visibility, Makefile.ref) and is loaded
        # RTLD_GLOBAL so plug-ins can resolve MNN core symbols; a torch imported AFTER it would bind to those copies and
        # abort in std::locale.  Load torch first when it is there.
        try:
            import torch  # noqa: F401
        except ImportError:
            pass
        self.lib = L = C.CDLL(os.path.join(_REF_DIR, "libniti_ref_driver.so"), mode=C.RTLD_GLOBAL)  # plug-ins resolve MNN core symbols against it
        workdir = workdir or os.path.join(_REF_DIR, "_cwd")
        os.makedirs(workdir, exist_ok=True)
        rc = L.ref_init(workdir.encode(), threads)
        assert rc == 0, rc
        ip = C.POINTER(C.c_int)
        L.ref_conv_fwd.argtypes = [ip, _i8p, _i8p, C.c_int, C.c_int, _i8p]
        L.ref_conv_bwd.argtypes = [ip, _i8p, _i8p, _i8p, C.c_int, C.c_int, _i8p, _i8p]
        L.ref_sgd.argtypes = [_i8p, _i8p, C.c_int, _i8p]
        L.ref_relu.argtypes = [_i8p] + [C.c_int] * 4 + [_i8p]
        L.ref_relu_grad.argtypes = [_i8p, _i8p] + [C.c_int] * 4 + [_i8p]
        L.ref_maxpool.argtypes = [_i8p] + [C.c_int] * 6 + [_i8p]
        L.ref_maxpool_grad.argtypes = [_i8p, _i8p] + [C.c_int] * 6 + [_i8p]
        L.ref_loss_grad.argtypes = [_i8p, C.c_int, _i32p, C.c_int, C.c_int, _i8p]
        L.ref_loss_value.argtypes = [_i8p, C.c_int, _i32p, C.c_int, C.c_int, C.POINTER(C.c_float)]

    @staticmethod
    def _sp(s):
        return C.cast(C.byref(s), C.POINTER(C.c_int))

    def conv_fwd(self, s, x, w, ex, ew):
        y = np.empty((s.n, s.oc, s.oh, s.ow), np.int8)
        e = self.lib.ref_conv_fwd(self._sp(s), _c(x), _c(w), ex, ew, y)
        return y, e

    def conv_bwd(self, s, x, w, dy, ex=-7, ew=-7):
        """dX is only evaluated when C % 4 == 0: otherwise the reference overruns its accumulator
        buffer (quirk Q6, NITI_DeConv_Int8.cpp acc_int32 sizing) and corrupts the heap."""
        want_dx = s.c % 4 == 0
        dx = np.empty((s.n, s.c, s.h, s.w), np.int8) if want_dx else None
        dw = np.empty((s.oc, s.c, s.kh, s.kw), np.int8)
        L = self.lib
        L.ref_conv_bwd.argtypes = [C.POINTER(C.c_int), _i8p, _i8p, _i8p, C.c_int, C.c_int,
                                   _i8p if want_dx else C.c_void_p, _i8p]
        rc = L.ref_conv_bwd(self._sp(s), _c(x), _c(w), _c(dy), ex, ew, dx, dw)
        assert rc == 0, rc
        return dx, dw

    def matmul_raw(self, a_raw, b, e, l, h):
        """OpType_NITI_MatMul_Int8 (transposeB) on RAW buffers: a_raw e*l bytes, b [h, l]; returns the raw e*h output bytes"""
        f = self.lib.ref_matmul
        f.restype = C.c_int
        f.argtypes = [C.c_int, C.c_int, C.c_int, _i8p, _i8p, _i8p]
        out = np.empty(e * h, np.int8)
        rc = f(e, l, h, _c(a_raw).reshape(-1), _c(b).reshape(-1), out)
        assert rc == 0, rc
        return out

    def grad_conv(self, s, x, dy, split=False):
        """dW [oc,c,kh,kw] through OpType_NITI_GradientCONV_Int8 / OpType_NITI_GradientSplitCONV_Int8 (stride 1)"""
        f = self.lib.ref_grad_conv
        f.restype = C.c_int
        f.argtypes = [C.c_int, C.POINTER(C.c_int), _i8p, _i8p, _i8p]
        dw = np.empty((s.oc, s.c, s.kh, s.kw), np.int8)
        rc = f(int(bool(split)), self._sp(s), _c(x), _c(dy), dw)
        assert rc == 0, rc
        return dw

    def sgd_update(self, w, dw):
        out = np.empty_like(_c(w))
        self.lib.ref_sgd(_c(w).reshape(-1), _c(dw).reshape(-1), out.size, out.reshape(-1))
        return out

    def relu(self, x):
        y = np.empty_like(_c(x)); self.lib.ref_relu(_c(x), *x.shape, y); return y

    def relu_grad(self, x, dy):
        dx = np.empty_like(_c(x)); self.lib.ref_relu_grad(_c(x), _c(dy), *x.shape, dx); return dx

    def maxpool(self, x, k, s):
        n, c, h, w = x.shape
        y = np.empty((n, c, (h - k) // s + 1, (w - k) // s + 1), np.int8)
        self.lib.ref_maxpool(_c(x), n, c, h, w, k, s, y)
        return y

    def maxpool_grad(self, x, dy, k, s):
        dx = np.empty_like(_c(x))
        rc = self.lib.ref_maxpool_grad(_c(x), _c(dy), *x.shape, k, s, dx)
        assert rc == 0, rc
        return dx

    def loss_grad(self, logits, ascale, labels):
        b, k = logits.shape
        onehot = np.zeros((b, k), np.int32)
        onehot[np.arange(b), labels] = 1
        g = np.empty((b, k), np.int8)
        rc = self.lib.ref_loss_grad(_c(logits), int(ascale), onehot, b, k, g)
        assert rc == 0, rc
        return g

    def loss_value(self, logits, ascale, labels):
        b, k = logits.shape
        onehot = np.zeros((b, k), np.int32)
        onehot[np.arange(b), labels] = 1
        out = C.c_float(0)
        rc = self.lib.ref_loss_value(_c(logits), int(ascale), onehot, b, k, C.byref(out))
        assert rc == 0, rc
        return float(out.value)

    # ---- whole training steps through the reference's Module / OpGrad / NITI_SGD (ref_driver.cpp) ----
    def use_backend(self, forward_type, plugin=None, threads=4):
        """Switch the reference's global executor to another forward type (0 = CPU, 14 = MNN_FORWARD_USER_0,
        include/MNN/MNNForwardType.h) after dlopen-ing an MNN backend plug-in.  Raises when no runtime
        creator answers for that type — the reference itself would silently stay on the CPU."""
        self.lib.ref_use_backend.argtypes = [C.c_int, C.c_char_p, C.c_int]
        rc = self.lib.ref_use_backend(int(forward_type), plugin.encode() if plugin else None, threads)
        if rc != 0:
            raise RuntimeError(f"ref_use_backend({forward_type}, {plugin}) failed: {rc}")

    def dsp_op(self, kind, geo, a, b, e0=0, e1=0, out_elems=None):
        """One DSP-named op type through the reference's express API on the CURRENT backend (ref_driver.cpp ref_dsp_op).
        kind: 0 DSP_CONV, 1 DSP_DECONV, 2 TRANSPOSEGRADIENT_CONV, 3 PARALLEL_GRADIENTCONV, 4 GRADIENT_SPLITBatchCONV,
        5 DSP_GRADIENTCONV; geo = (kernelY, kernelX, stride, pad, inputCount, outputCount); a, b: 4-d int8 (NHWC order).
        Returns (out int8 in the shape the reference's shape rule gives, exponent)."""
        a = np.ascontiguousarray(a, dtype=np.int8); b = np.ascontiguousarray(b, dtype=np.int8)
        out = np.zeros(out_elems if out_elems else max(a.size, b.size) * 64, dtype=np.int8)
        od = (C.c_int * 4)()
        f = self.lib.ref_dsp_op
        f.restype = C.c_int
        f.argtypes = [C.c_int, C.POINTER(C.c_int), _i8p, C.POINTER(C.c_int), _i8p, C.POINTER(C.c_int), C.c_int, C.c_int,
                      _i8p, C.POINTER(C.c_int)]
        rc = f(int(kind), (C.c_int * 6)(*[int(v) for v in geo]), a.reshape(-1), (C.c_int * 4)(*a.shape), b.reshape(-1),
               (C.c_int * 4)(*b.shape), int(e0), int(e1), out, od)
        if rc <= -1000:
            raise RuntimeError(f"ref_dsp_op(kind={kind}) failed: {rc}")
        dims = tuple(od)
        return out[: int(np.prod(dims))].reshape(dims).copy(), rc

    def net_train(self, spec, x_steps, ascale, labels_steps, weights, tclasses=None):
        """spec: step.py layer list; x_steps [steps, n, c, h, w] int8; labels_steps [steps, n];
        weights: list of (OIHW int8, wscale).  Returns dict(logits [steps, n, classes], e_logits [steps],
        loss [steps], weights [list of OIHW], ms [steps])."""
        kinds = {"conv": 0, "relu": 1, "pool": 2, "flatten": 3}
        rows = []
        for it in spec:
            r = [kinds[it[0]]] + [int(v) for v in it[1:]]
            rows.append(r + [0] * (5 - len(r)))
        sp = np.asarray(rows, np.int32).reshape(-1)
        x_steps = _c(x_steps)
        steps, n, c, h, w = x_steps.shape
        classes = [it for it in spec if it[0] == "conv"][-1][1]
        tclasses = tclasses or classes
        onehot = np.zeros((steps, n, tclasses), np.int32)
        for s in range(steps):
            onehot[s, np.arange(n), labels_steps[s]] = 1
        wcat = np.concatenate([_c(W).reshape(-1) for W, _ in weights])
        wsc = np.asarray([e for _, e in weights], np.int8)
        logits = np.empty((steps, n, classes), np.int8)
        elog = np.empty(steps, np.int8)
        loss = np.empty(steps, np.float32)
        wout = np.empty_like(wcat)
        ms = np.empty(steps, np.float32)
        chw = np.asarray([c, h, w], np.int32)
        f = self.lib.ref_net_train
        f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
        f.argtypes = [_i32p, C.c_int, C.c_int, _i32p, C.c_int, C.c_int, _i8p, _i32p, C.c_int, _i8p, _i8p,
                      _i8p, _i8p, f32p, _i8p, f32p]
        rc = f(sp, len(rows), n, chw, int(ascale), steps, x_steps, onehot, tclasses, wcat, wsc,
               logits, elog, loss, wout, ms)
        if rc != classes:
            raise RuntimeError(f"ref_net_train failed: {rc}")
        outw, off = [], 0
        for W, _ in weights:
            outw.append(wout[off:off + W.size].reshape(W.shape).copy()); off += W.size
        return {"logits": logits, "e_logits": elog, "loss": loss, "weights": outw, "ms": ms}
