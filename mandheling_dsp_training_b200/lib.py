"""ctypes binding of libniti_b200.so (include/niti_b200.h).  No fallback: importing works anywhere,
but creating a context without the built library or without a B200 raises NitiError."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBPATH = os.environ.get("NITI_B200_LIB") or os.path.join(_HERE, "libniti_b200.so")     # override: A/B runs of two builds

PATH_AUTO, PATH_SIMT, PATH_TC, PATH_TMA, PATH_TMA2 = 0, 1, 2, 3, 4


class NitiError(RuntimeError):
    pass


class ConvDesc(C.Structure):
    """niti_conv_desc: input [n,h,w,c] NHWC, weight [oc,kh,kw,c] OHWI"""
    _fields_ = [(k, C.c_int32) for k in ("n", "c", "h", "w", "oc", "kh", "kw", "stride", "pad", "dil")]

    @property
    def oh(self):
        return (self.h + 2 * self.pad - (self.dil * (self.kh - 1) + 1)) // self.stride + 1

    @property
    def ow(self):
        return (self.w + 2 * self.pad - (self.dil * (self.kw - 1) + 1)) // self.stride + 1


def desc(n, c, h, w, oc, kh, kw=None, stride=1, pad=0, dil=1):
    return ConvDesc(n, c, h, w, oc, kh, kh if kw is None else kw, stride, pad, dil)


_P = C.c_void_p


class UpdateSeg(C.Structure):
    """niti_update_seg: one conv layer of niti_update_batched"""
    _fields_ = [("acc", _P), ("w", _P), ("wt", _P), ("maxabs", _P),
                ("oc", C.c_int32), ("taps", C.c_int32), ("c", C.c_int32), ("sub", C.c_int32)]


class ChainOp(C.Structure):
    """niti_chain_op: one forward / dX op of a chain (include/niti_b200.h)"""
    _fields_ = [("d", ConvDesc), ("op", C.c_int32), ("inp", _P), ("w", _P), ("ex", _P), ("ew", _P), ("ey", _P), ("out", _P),
                ("relu", C.c_int32), ("gate", _P), ("out_pool", _P), ("pool_x", _P), ("pool_gate", C.c_int32),
                ("maxabs", _P), ("dp_slot", C.c_int32)]


_SIGS = {
    "niti_ctx_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "niti_ctx_destroy": (C.c_int, [_P]),
    "niti_last_error": (C.c_char_p, [_P]),
    "niti_set_path": (C.c_int, [_P, C.c_int]),
    "niti_set_speculate": (C.c_int, [_P, C.c_int]),
    "niti_speculate_stats": (C.c_int, [_P, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]),
    "niti_version": (C.c_int, []),
    "niti_launch_count": (C.c_int64, [_P]),
    "niti_last_path": (C.c_char_p, [_P]),
    "niti_conv_fwd_acc": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P, _P]),
    "niti_conv_dx_acc": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P, _P]),
    "niti_conv_dw_acc": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P, _P]),
    "niti_conv_dw_accumulate": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P]),
    "niti_conv_dx_acc_wt": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P]),
    "niti_update_batched": (C.c_int, [_P, _P, C.c_int, _P]),
    "niti_update_batched_phases": (C.c_int, [_P, _P, C.c_int, _P]),
    "niti_matmul_nt_acc": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P, _P, _P, _P]),
    "niti_zero_u32": (C.c_int, [_P, _P, C.c_int, _P]),
    "niti_absmax": (C.c_int, [_P, _P, C.c_size_t, _P, _P]),
    "niti_rescale_act": (C.c_int, [_P, _P, C.c_size_t, _P, _P, _P, _P, _P, _P]),
    "niti_rescale_grad": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_int, _P, _P, _P]),
    "niti_rescale_dsp": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_int, _P, _P, _P, _P, _P]),
    "niti_rescale_act_fused": (C.c_int, [_P, _P, C.c_size_t, _P, _P, _P, _P, _P, C.c_int, _P, _P]),
    "niti_rescale_act_dp": (C.c_int, [_P, _P, C.c_size_t, _P, _P, _P, _P, _P, C.c_int, _P, C.c_int, _P]),
    "niti_dp_alloc": (C.c_int, [_P, _P]),
    "niti_dp_connect": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "niti_dp_close": (C.c_int, [_P]),
    "niti_dp_step_begin": (C.c_int, [_P, _P]),
    "niti_dp_error": (C.c_int, [_P, C.POINTER(C.c_int)]),
    "niti_dp_reserved_sms": (C.c_int, [_P]),
    "niti_dp_window_alloc": (C.c_int, [_P, C.c_size_t, C.POINTER(_P), _P]),
    "niti_dp_window_connect": (C.c_int, [_P, _P]),
    "niti_dp_window_free": (C.c_int, [_P]),
    "niti_dp_allreduce_i32": (C.c_int, [_P, C.c_size_t, C.c_size_t, C.c_int, _P]),
    "niti_sync_error": (C.c_int, [_P, C.POINTER(C.c_int)]),
    "niti_sgd_update": (C.c_int, [_P, _P, _P, C.c_size_t, _P]),
    "niti_conv_fwd": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, _P]),
    "niti_conv_fwd_fused": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, C.c_int, _P, _P, C.c_int, _P]),
    "niti_conv_dx_fused": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, _P, C.c_int, _P]),
    "niti_conv_fwd_fused_pool": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, C.c_int, _P, _P, _P, C.c_int, _P]),
    "niti_conv_dx_fused_poolgrad": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P, _P, C.c_int, _P, _P, _P, C.c_int, _P]),
    "niti_conv_stem_fwd": (C.c_int, [_P, C.POINTER(ConvDesc), C.c_int, _P, _P, _P, _P, _P, _P, C.c_int, _P, _P, _P, C.c_int, _P]),
    "niti_chain_op_eligible": (C.c_int, [_P, C.POINTER(ChainOp)]),
    "niti_chain_create": (C.c_int, [_P, _P, C.c_int, C.POINTER(_P)]),
    "niti_chain_run": (C.c_int, [_P, _P, _P]),
    "niti_chain_destroy": (C.c_int, [_P, _P]),
    "niti_loss_grad_strided": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, _P, _P]),
    "niti_conv_dx": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P]),
    "niti_conv_dw": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, _P, _P]),
    "niti_matmul_nt": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P, _P, _P]),
    "niti_matmul_nt_static": (C.c_int, [C.c_int, C.c_int, C.c_int, _P, _P, _P, _P, C.c_char_p, C.c_size_t]),
    "niti_rescale_grad_split": (C.c_int, [_P, _P, C.c_size_t, C.c_int, _P, C.c_int, _P, _P]),
    "niti_split_count": (C.c_int, [C.c_int, C.c_int, C.c_int]),
    "niti_conv_dw_split": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, C.c_int, _P, _P]),
    "niti_relu": (C.c_int, [_P, _P, _P, C.c_size_t, _P]),
    "niti_relu_grad": (C.c_int, [_P, _P, _P, _P, C.c_size_t, _P]),
    "niti_maxpool": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "niti_maxpool_grad": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "niti_maxpool_grad_gated": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, C.c_int, _P]),
    "niti_loss_grad": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, _P, _P]),
    "niti_loss_value": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, _P, _P, _P]),
    "niti_onehot_labels": (C.c_int, [_P, _P, C.c_int, C.c_int, _P, _P]),
    "niti_binary_i8": (C.c_int, [_P, C.c_int, _P, _P, C.c_size_t, C.c_int, _P, _P]),
    "niti_zero_insert_c4": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "niti_pad2d": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "niti_nchw_to_nhwc": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "niti_nhwc_to_nchw": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "niti_nc4hw4_to_nhwc": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "niti_nhwc_to_nc4hw4": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "niti_conv_fwd_host": (C.c_int, [_P, C.POINTER(ConvDesc), _P, _P, C.c_int8, C.c_int8, _P, _P]),
    "niti_debug_trace": (C.c_int, [_P, _P]),
    "niti_debug_timeline": (C.c_int, [_P, _P, C.c_int]),
    "niti_debug_timeline_count": (C.c_int, [_P]),
    "niti_debug_timeline_name": (C.c_char_p, [_P, C.c_int]),
    "niti_probe_i8_peak": (C.c_int, [_P, C.c_int, C.POINTER(C.c_double)]),
}
EXPORTED_SYMBOLS = tuple(_SIGS)


def load_library(path=_LIBPATH):
    if not os.path.exists(path):
        raise NitiError(f"{path} is not built: run `python -m mandheling_dsp_training_b200.build` "
                        "(there is no CPU fallback)")
    lib = C.CDLL(path)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


def _ptr(t):
    """device pointer of a torch tensor (or None / int passthrough)"""
    if t is None:
        return None
    if isinstance(t, int):
        return t
    return t.data_ptr()


def dp_env(reserve_sms=16):
    """Environment a data-parallel job needs BEFORE torch.distributed.init_process_group("nccl"): caps NCCL at the SMs the
    library leaves free for it (csrc/dp.cu).  Values already set are respected if they satisfy the rule."""
    os.environ.setdefault("NITI_B200_DP_RESERVE_SMS", str(int(reserve_sms)))
    res = int(os.environ["NITI_B200_DP_RESERVE_SMS"])
    cap = os.environ.get("NCCL_MAX_CTAS")
    if cap is None or not cap.isdigit() or int(cap) > max(res, 1):
        os.environ["NCCL_MAX_CTAS"] = str(max(res, 1))


class NitiB200:
    """One context per device.  Methods take torch CUDA tensors (int8 / int32 / uint32-as-int32) and
    enqueue on the given (default: torch current) stream; nothing synchronizes."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = _P()
        rc = self.lib.niti_ctx_create(int(device), C.byref(h))
        if rc != 0:
            raise NitiError(f"niti_ctx_create(device={device}) failed with code {rc}: no usable sm_100a "
                            "device (there is no CPU fallback)")
        self.h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "h", None):
            self.lib.niti_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self, stream):
        if stream is not None:
            return stream
        import torch
        return torch.cuda.current_stream(self.device).cuda_stream

    def _ck(self, rc):
        if rc != 0:
            raise NitiError(f"code {rc}: {self.lib.niti_last_error(self.h).decode()}")

    def set_path(self, path):
        self._ck(self.lib.niti_set_path(self.h, path))

    def set_speculate(self, mode):
        """exponent prediction of the fused kernels: 0 off, 1 adaptive (default), 2 always (include/niti_b200.h)"""
        self._ck(self.lib.niti_set_speculate(self.h, int(mode)))

    def speculate_stats(self):
        """(hits, misses, not speculated) of the fused kernels' exponent prediction since the context was created"""
        h, m, k = C.c_ulonglong(0), C.c_ulonglong(0), C.c_ulonglong(0)
        self._ck(self.lib.niti_speculate_stats(self.h, C.byref(h), C.byref(m), C.byref(k)))
        return int(h.value), int(m.value), int(k.value)

    @property
    def launches(self):
        return self.lib.niti_launch_count(self.h)

    @property
    def last_path(self):
        return self.lib.niti_last_path(self.h).decode()

    # ---- phase 1
    def conv_fwd_acc(self, d, x, w, acc, maxabs=None, stream=None):
        self._ck(self.lib.niti_conv_fwd_acc(self.h, C.byref(d), _ptr(x), _ptr(w), _ptr(acc), _ptr(maxabs), self._stream(stream)))

    def conv_dx_acc(self, d, dy, w, acc, maxabs=None, stream=None):
        self._ck(self.lib.niti_conv_dx_acc(self.h, C.byref(d), _ptr(dy), _ptr(w), _ptr(acc), _ptr(maxabs), self._stream(stream)))

    def conv_dx_acc_wt(self, d, dy, w, wt, acc, maxabs=None, stream=None):
        self._ck(self.lib.niti_conv_dx_acc_wt(self.h, C.byref(d), _ptr(dy), _ptr(w), _ptr(wt), _ptr(acc), _ptr(maxabs),
                                              self._stream(stream)))

    @staticmethod
    def make_update_segs(layers):
        """layers: list of (acc, w, wt|None, maxabs slot, oc, taps, c, sub) torch tensors / ints -> ctypes array of
        UpdateSeg for update_batched.  The caller keeps both the array and the tensors alive."""
        segs = (UpdateSeg * len(layers))()
        for s_, (acc, w, wt, slot, oc, taps, c, sub) in zip(segs, layers):
            s_.acc, s_.w, s_.wt, s_.maxabs = _ptr(acc), _ptr(w), _ptr(wt), _ptr(slot)
            s_.oc, s_.taps, s_.c, s_.sub = oc, taps, c, sub
        return segs

    def update_layers(self, layers, stream=None):
        self.update_batched(self.make_update_segs(layers), stream)

    def update_batched(self, segs, stream=None):
        """segs: ctypes array of UpdateSeg (host); read at call time"""
        self._ck(self.lib.niti_update_batched(self.h, C.cast(segs, _P), len(segs), self._stream(stream)))

    def update_batched_phases(self, segs, stream=None):
        """the same as two barrier-free launches (may run on a second stream beside the dX / dW kernels)"""
        self._ck(self.lib.niti_update_batched_phases(self.h, C.cast(segs, _P), len(segs), self._stream(stream)))

    def conv_dw_acc(self, d, x, dy, acc, maxabs=None, stream=None):
        self._ck(self.lib.niti_conv_dw_acc(self.h, C.byref(d), _ptr(x), _ptr(dy), _ptr(acc), _ptr(maxabs), self._stream(stream)))

    def rescale_dsp(self, acc, n, maxabs, out, post_div=1, ex=None, ew=None, ey=None, stream=None):
        """DSP-named op types' requantisation (round-half-up shift, saturate, optional /16 or /32)"""
        self._ck(self.lib.niti_rescale_dsp(self.h, _ptr(acc), n, _ptr(maxabs), int(post_div), _ptr(out), _ptr(ex), _ptr(ew),
                                           _ptr(ey), self._stream(stream)))

    def conv_dw_accumulate(self, d, x, dy, acc, stream=None):
        """acc += dW (acc zeroed by the caller); barrier-free launch, safe on a side stream"""
        self._ck(self.lib.niti_conv_dw_accumulate(self.h, C.byref(d), _ptr(x), _ptr(dy), _ptr(acc), self._stream(stream)))

    def matmul_nt_acc(self, e, h, l, a, b, acc, maxabs=None, stream=None):
        self._ck(self.lib.niti_matmul_nt_acc(self.h, e, h, l, _ptr(a), _ptr(b), _ptr(acc), _ptr(maxabs), self._stream(stream)))

    # ---- phase 2
    def zero_u32(self, p, count=1, stream=None):
        self._ck(self.lib.niti_zero_u32(self.h, _ptr(p), count, self._stream(stream)))

    def absmax(self, acc, n, maxabs, stream=None):
        self._ck(self.lib.niti_absmax(self.h, _ptr(acc), n, _ptr(maxabs), self._stream(stream)))

    def rescale_act(self, acc, n, maxabs, out, exp_a=None, exp_b=None, exp_out=None, stream=None):
        self._ck(self.lib.niti_rescale_act(self.h, _ptr(acc), n, _ptr(maxabs), _ptr(out), _ptr(exp_a), _ptr(exp_b),
                                           _ptr(exp_out), self._stream(stream)))

    def rescale_act_fused(self, acc, n, maxabs, out, exp_a=None, exp_b=None, exp_out=None, relu=False, gate=None, stream=None):
        self._ck(self.lib.niti_rescale_act_fused(self.h, _ptr(acc), n, _ptr(maxabs), _ptr(out), _ptr(exp_a), _ptr(exp_b),
                                                 _ptr(exp_out), int(bool(relu)), _ptr(gate), self._stream(stream)))

    def rescale_act_dp(self, acc, n, maxabs, out, exp_a=None, exp_b=None, exp_out=None, relu=False, gate=None, dp_slot=0, stream=None):
        self._ck(self.lib.niti_rescale_act_dp(self.h, _ptr(acc), n, _ptr(maxabs), _ptr(out), _ptr(exp_a), _ptr(exp_b),
                                              _ptr(exp_out), int(bool(relu)), _ptr(gate), int(dp_slot), self._stream(stream)))

    # ---- data-parallel scalar exchange over peer memory
    def dp_connect(self, dist):
        """gather every rank's IPC handle through torch.distributed and map the peers' exchange buffers.
        Enforces the SM-reserve rule of csrc/dp.cu: the collective library's CTA cap (NCCL_MAX_CTAS) must be set and must
        not exceed the SMs the library leaves free, otherwise NCCL CTAs wait for SMs held by kernels spinning on a peer."""
        import torch
        reserve = int(os.environ.get("NITI_B200_DP_RESERVE_SMS", "16"))
        nccl_sum = os.environ.get("NITI_B200_DP_ALLREDUCE", "peer") == "nccl"   # else NCCL never runs beside the contractions
        if dist.get_backend() == "nccl" and dist.get_world_size() > 1 and nccl_sum:
            cap = os.environ.get("NCCL_MAX_CTAS")
            if cap is None or not cap.isdigit() or int(cap) > reserve or int(cap) < 1:
                raise NitiError(f"data parallel needs NCCL_MAX_CTAS set (before the process group is created) to at most "
                                f"NITI_B200_DP_RESERVE_SMS = {reserve}; found {cap!r}.  Use lib.dp_env() before "
                                "init_process_group.")
        h = (C.c_ubyte * 64)()
        self._ck(self.lib.niti_dp_alloc(self.h, C.cast(h, _P)))
        world, rank = dist.get_world_size(), dist.get_rank()
        mine = torch.tensor(list(bytes(h)), dtype=torch.uint8, device=f"cuda:{self.device}")
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        blob = b"".join(bytes(t.cpu().tolist()) for t in allh)
        buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        self._ck(self.lib.niti_dp_connect(self.h, rank, world, C.cast(buf, _P)))
        dist.barrier()
        self.dp_connected = True

    def dp_window(self, dist, words):
        """Allocate this rank's gradient window (int32 words, zeroed), map every peer's, and return the local one as a
        torch tensor that aliases it (zero copy).  The dW accumulators of a data-parallel net live here so that
        dp_allreduce can add them over NVLink peer memory (include/niti_b200.h niti_dp_allreduce_i32)."""
        import torch
        h = (C.c_ubyte * 64)()
        ptr = _P()
        self._ck(self.lib.niti_dp_window_alloc(self.h, words, C.byref(ptr), C.cast(h, _P)))
        world = dist.get_world_size()
        mine = torch.tensor(list(bytes(h)), dtype=torch.uint8, device=f"cuda:{self.device}")
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        blob = b"".join(bytes(t.cpu().tolist()) for t in allh)
        buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        self._ck(self.lib.niti_dp_window_connect(self.h, C.cast(buf, _P)))
        dist.barrier()
        n = (words + 3) // 4 * 4

        class _Window:                                   # the cudaMalloc'ed window as a CUDA array for torch.as_tensor
            __cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr.value, False), "version": 3, "strides": None}
        t = torch.as_tensor(_Window(), device=f"cuda:{self.device}")
        assert t.data_ptr() == ptr.value and t.dtype == torch.int32
        # The memory belongs to the context, not to torch: keep the context alive as long as the tensor is referenced
        # from Python (views made by slicing share the base tensor), and never use it after dp_close / close.
        t._niti_owner = self
        self._dp_window_tensor = t
        return t

    def dp_allreduce(self, off, n, slot, stream=None):
        """in-place SUM over ranks of window words [off, off + n) (multiples of 4); one slot per call site and step"""
        self._ck(self.lib.niti_dp_allreduce_i32(self.h, off, n, slot, self._stream(stream)))

    def dp_reserved_sms(self):
        return self.lib.niti_dp_reserved_sms(self.h)

    def sync_error(self):
        v = C.c_int(0)
        self._ck(self.lib.niti_sync_error(self.h, C.byref(v)))
        return v.value

    def check_errors(self):
        """poll the sticky device-side error flags (synchronises): raises when a peer exchange or a grid barrier gave up"""
        if getattr(self, "dp_connected", False) and self.dp_error():
            raise NitiError("data-parallel max|acc| exchange timed out: a peer did not answer; replicas may have diverged")
        if self.sync_error():
            raise NitiError("a grid-wide barrier timed out (two barrier-holding calls of one context in flight?)")

    def dp_step_begin(self, stream=None):
        self._ck(self.lib.niti_dp_step_begin(self.h, self._stream(stream)))

    def dp_error(self):
        v = C.c_int(0)
        self._ck(self.lib.niti_dp_error(self.h, C.byref(v)))
        return v.value

    def rescale_grad(self, acc, n, maxabs, sub, dw_out=None, w_inout=None, stream=None):
        self._ck(self.lib.niti_rescale_grad(self.h, _ptr(acc), n, _ptr(maxabs), sub, _ptr(dw_out), _ptr(w_inout),
                                            self._stream(stream)))

    def rescale_grad_split(self, acc, n_per_slice, slices, maxabs, sub, out, stream=None):
        """maxabs: `slices` uint32 slots holding max|acc_s|; out = per-slice rule B + alignment (NITI_GradientSplitConv_Int8)"""
        self._ck(self.lib.niti_rescale_grad_split(self.h, _ptr(acc), n_per_slice, int(slices), _ptr(maxabs), int(sub), _ptr(out),
                                                  self._stream(stream)))

    def split_count(self, oc, oh, ow):
        return self.lib.niti_split_count(int(oc), int(oh), int(ow))

    def conv_dw_split(self, d, x, dy, slices, dw, stream=None):
        self._ck(self.lib.niti_conv_dw_split(self.h, C.byref(d), _ptr(x), _ptr(dy), int(slices), _ptr(dw), self._stream(stream)))

    def sgd_update(self, w, dw, n, stream=None):
        self._ck(self.lib.niti_sgd_update(self.h, _ptr(w), _ptr(dw), n, self._stream(stream)))

    # ---- whole ops
    def conv_fwd(self, d, x, w, ex, ew, y, ey, stream=None):
        self._ck(self.lib.niti_conv_fwd(self.h, C.byref(d), _ptr(x), _ptr(w), _ptr(ex), _ptr(ew), _ptr(y), _ptr(ey), self._stream(stream)))

    def conv_fwd_fused(self, d, x, w, ex, ew, y, ey, relu, acc, maxabs, dp_slot=0, stream=None):
        self._ck(self.lib.niti_conv_fwd_fused(self.h, C.byref(d), _ptr(x), _ptr(w), _ptr(ex), _ptr(ew), _ptr(y), _ptr(ey),
                                              int(bool(relu)), _ptr(acc), _ptr(maxabs), int(dp_slot), self._stream(stream)))

    def conv_dx_fused(self, d, dy, w, wt, dx, gate, acc, maxabs, dp_slot=0, stream=None):
        self._ck(self.lib.niti_conv_dx_fused(self.h, C.byref(d), _ptr(dy), _ptr(w), _ptr(wt), _ptr(dx), _ptr(gate), _ptr(acc),
                                             _ptr(maxabs), int(dp_slot), self._stream(stream)))

    def conv_fwd_fused_pool(self, d, x, w, ex, ew, y, ey, relu, y_pool, acc, maxabs, dp_slot=0, stream=None):
        """forward op + ReLU + the 2x2/2 max pool behind it: writes y and y_pool"""
        self._ck(self.lib.niti_conv_fwd_fused_pool(self.h, C.byref(d), _ptr(x), _ptr(w), _ptr(ex), _ptr(ew), _ptr(y), _ptr(ey),
                                                   int(bool(relu)), _ptr(y_pool), _ptr(acc), _ptr(maxabs), int(dp_slot),
                                                   self._stream(stream)))

    def conv_stem_fwd(self, d, c_live, x, w, ex, ew, y, ey, relu, y_pool, acc, maxabs, dp_slot=0, stream=None):
        """stem layer (image input, only the first c_live channels of x can be non-zero) + ReLU (+ 2x2/2 max pool -> y_pool)"""
        self._ck(self.lib.niti_conv_stem_fwd(self.h, C.byref(d), int(c_live), _ptr(x), _ptr(w), _ptr(ex), _ptr(ew), _ptr(y), _ptr(ey),
                                             int(bool(relu)), _ptr(y_pool) if y_pool is not None else None, _ptr(acc), _ptr(maxabs),
                                             int(dp_slot), self._stream(stream)))

    def conv_dx_fused_poolgrad(self, d, dy, w, wt, dx_prepool, pool_x, gate_relu, dx_pooled, acc, maxabs, dp_slot=0, stream=None):
        """dX op + max-pool gradient (+ ReLU gate) of the pool in front of this conv: writes the gradient w.r.t. pool_x"""
        self._ck(self.lib.niti_conv_dx_fused_poolgrad(self.h, C.byref(d), _ptr(dy), _ptr(w), _ptr(wt), _ptr(dx_prepool), _ptr(pool_x),
                                                      int(bool(gate_relu)), _ptr(dx_pooled), _ptr(acc), _ptr(maxabs), int(dp_slot),
                                                      self._stream(stream)))

    # ---- chains (several fused ops in one persistent launch)
    @staticmethod
    def chain_op(d, op, inp, w, out, maxabs, dp_slot, ex=None, ew=None, ey=None, relu=False, gate=None, out_pool=None,
                 pool_x=None, pool_gate=False):
        o = ChainOp()
        o.d = d; o.op = int(op); o.inp = _ptr(inp); o.w = _ptr(w); o.ex = _ptr(ex); o.ew = _ptr(ew); o.ey = _ptr(ey)
        o.out = _ptr(out); o.relu = int(bool(relu)); o.gate = _ptr(gate); o.out_pool = _ptr(out_pool)
        o.pool_x = _ptr(pool_x); o.pool_gate = int(bool(pool_gate)); o.maxabs = _ptr(maxabs); o.dp_slot = int(dp_slot)
        return o

    def chain_op_eligible(self, op):
        return bool(self.lib.niti_chain_op_eligible(self.h, C.byref(op)))

    def chain_create(self, ops):
        arr = (ChainOp * len(ops))(*ops)
        h = _P()
        self._ck(self.lib.niti_chain_create(self.h, C.cast(arr, _P), len(ops), C.byref(h)))
        return h

    def chain_run(self, chain, stream=None):
        self._ck(self.lib.niti_chain_run(self.h, chain, self._stream(stream)))

    def chain_destroy(self, chain):
        self.lib.niti_chain_destroy(self.h, chain)

    def conv_dx(self, d, dy, w, dx, stream=None):
        self._ck(self.lib.niti_conv_dx(self.h, C.byref(d), _ptr(dy), _ptr(w), _ptr(dx), self._stream(stream)))

    def conv_dw(self, d, x, dy, dw, stream=None):
        self._ck(self.lib.niti_conv_dw(self.h, C.byref(d), _ptr(x), _ptr(dy), _ptr(dw), self._stream(stream)))

    def matmul_nt(self, e, h, l, a, b, c, stream=None):
        self._ck(self.lib.niti_matmul_nt(self.h, e, h, l, _ptr(a), _ptr(b), _ptr(c), self._stream(stream)))

    # ---- element-wise
    def relu(self, x, y, n, stream=None):
        self._ck(self.lib.niti_relu(self.h, _ptr(x), _ptr(y), n, self._stream(stream)))

    def relu_grad(self, x, dy, dx, n, stream=None):
        self._ck(self.lib.niti_relu_grad(self.h, _ptr(x), _ptr(dy), _ptr(dx), n, self._stream(stream)))

    def maxpool(self, x, n, h, w, c, k, s, y, stream=None):
        self._ck(self.lib.niti_maxpool(self.h, _ptr(x), n, h, w, c, k, s, _ptr(y), self._stream(stream)))

    def maxpool_grad(self, x, y, dy, n, h, w, c, k, s, dx, stream=None):
        self._ck(self.lib.niti_maxpool_grad(self.h, _ptr(x), _ptr(y), _ptr(dy), n, h, w, c, k, s, _ptr(dx), self._stream(stream)))

    def maxpool_grad_gated(self, x, y, dy, n, h, w, c, k, s, dx, gate_relu, stream=None):
        self._ck(self.lib.niti_maxpool_grad_gated(self.h, _ptr(x), _ptr(y), _ptr(dy), n, h, w, c, k, s, _ptr(dx),
                                                  int(bool(gate_relu)), self._stream(stream)))

    def loss_grad_strided(self, logits, ascale, labels, batch, classes, stride, grad, stream=None):
        self._ck(self.lib.niti_loss_grad_strided(self.h, _ptr(logits), _ptr(ascale), _ptr(labels), batch, classes, stride, _ptr(grad),
                                                 self._stream(stream)))

    def loss_value(self, logits, ascale, labels, batch, classes, stride, row_scratch, loss_out, stream=None):
        """float loss scalar of NITI_CPULoss_Int8 (monitoring) into the 1-element device tensor loss_out"""
        self._ck(self.lib.niti_loss_value(self.h, _ptr(logits), _ptr(ascale), _ptr(labels), batch, classes, stride, _ptr(row_scratch),
                                          _ptr(loss_out), self._stream(stream)))

    def loss_grad(self, logits, ascale, labels, batch, classes, grad, stream=None):
        self._ck(self.lib.niti_loss_grad(self.h, _ptr(logits), _ptr(ascale), _ptr(labels), batch, classes, _ptr(grad), self._stream(stream)))

    # ---- layout
    def nchw_to_nhwc(self, src, n, c, h, w, dst, stream=None):
        self._ck(self.lib.niti_nchw_to_nhwc(self.h, _ptr(src), n, c, h, w, _ptr(dst), self._stream(stream)))

    def nhwc_to_nchw(self, src, n, c, h, w, dst, stream=None):
        self._ck(self.lib.niti_nhwc_to_nchw(self.h, _ptr(src), n, c, h, w, _ptr(dst), self._stream(stream)))

    def nc4hw4_to_nhwc(self, src, n, c, h, w, dst, stream=None):
        self._ck(self.lib.niti_nc4hw4_to_nhwc(self.h, _ptr(src), n, c, h, w, _ptr(dst), self._stream(stream)))

    def nhwc_to_nc4hw4(self, src, n, c, h, w, dst, stream=None):
        self._ck(self.lib.niti_nhwc_to_nc4hw4(self.h, _ptr(src), n, c, h, w, _ptr(dst), self._stream(stream)))

    # ---- host-buffer call (numpy int8 arrays, NHWC / OHWI)
    def conv_fwd_host(self, d, x_np, w_np, ex, ew, y_np):
        ey = C.c_int8(0)
        self._ck(self.lib.niti_conv_fwd_host(self.h, C.byref(d), x_np.ctypes.data, w_np.ctypes.data, ex, ew,
                                             y_np.ctypes.data, C.addressof(ey)))
        return ey.value

    def debug_trace(self, buf):
        self._ck(self.lib.niti_debug_trace(self.h, _ptr(buf)))

    def debug_timeline(self, buf, capacity=0):
        self._ck(self.lib.niti_debug_timeline(self.h, _ptr(buf), int(capacity)))

    def timeline_names(self):
        return [self.lib.niti_debug_timeline_name(self.h, i).decode() for i in range(self.lib.niti_debug_timeline_count(self.h))]

    def probe_i8_peak(self, iters=2000):
        t = C.c_double(0)
        self._ck(self.lib.niti_probe_i8_peak(self.h, iters, C.byref(t)))
        return t.value
