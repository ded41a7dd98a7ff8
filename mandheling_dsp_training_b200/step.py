"""The NITI int8 training step on the device, built only from C-ABI calls (lib.NitiB200).

Mirrors what the reference's express graph executes for one `sgd->step(loss)`
(tools/train/source/demo/MnistUtils.cpp:79-147, demo/mnistTrain.cpp:159-181,
grad/NITI_Conv_Int8_Grad.cpp:16-197, optimizer/NITI_SGD.hpp:20-54):

  forward   conv -> [relu] -> [maxpool] ...      NITI_Conv_Int8 / NITI_CPURelu_Int8 / NITI_Maxpool_Int8
  loss grad integer softmax-CE gradient          NITI_CPULossGrad_Int8
  backward  pool grad / relu grad, dX, dW        NITI_CPUPoolGrad / NITI_CPUReluGrad / NITI_DeConv / NITI_GradientConv
  update    W <- clip127(W - dW)                 NITI_SGD + int8 Binary SUB

Everything stays resident in HBM (NHWC activations, OHWI weights); all buffers are allocated once, so
the step is a fixed launch sequence (CUDA-graph capturable at world size 1).

Data parallel (world size g): the batch is sharded; to stay bit-identical to the single-device run the
two reductions that span the batch are all-reduced between phase 1 and phase 2 of each op:
  * max|acc| of every forward / dX tensor  -> all_reduce(MAX) on one uint32
  * the int32 dW accumulators              -> all_reduce(SUM), one bucket for the whole net
Integer sums and maxima are order-independent, so g devices reproduce the 1-device result exactly.
"""
import contextlib
import os

import numpy as np
import torch

from .lib import NitiB200, desc as make_desc

# ---- layer lists ------------------------------------------------------------------------------------
# ("conv", oc, k, stride, pad) | ("relu",) | ("pool", k, s) | ("flatten",)


def lenet_spec():
    """the reference's NITIInt8 model (demo/mnistTrain.cpp:131-181): 1x28x28 -> 12 logits"""
    return [("conv", 20, 5, 1, 0), ("relu",), ("pool", 2, 2),
            ("conv", 52, 5, 1, 0), ("relu",), ("pool", 2, 2), ("flatten",),
            ("conv", 500, 1, 1, 0), ("relu",), ("conv", 12, 1, 1, 0)]


def vgg11_spec(classes=12, first_c_pad=False):
    """VGG-11 for 32x32 inputs (CIFAR-shaped), NITI style: no bias, no batch-norm, int8 everywhere"""
    cfg = [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"]
    spec = []
    for v in cfg:
        if v == "M":
            spec.append(("pool", 2, 2))
        else:
            spec += [("conv", v, 3, 1, 1), ("relu",)]
    spec += [("flatten",), ("conv", 512, 1, 1, 0), ("relu",), ("conv", 512, 1, 1, 0), ("relu",),
             ("conv", classes, 1, 1, 0)]
    return spec


def resnet18_spec(classes=1000):
    """ResNet-18 trunk as a plain int8 chain (7x7 s2 stem, 3x3 stages with stride-2 transitions).
    The residual adds and fp32 batch-norm of the mixed-precision config are outside the int8 hot path
    (SURVEY.md section 8 f); this list carries every int8 contraction of the network."""
    spec = [("conv", 64, 7, 2, 3), ("relu",), ("pool", 2, 2)]
    for oc, first_stride in [(64, 1), (128, 2), (256, 2), (512, 2)]:
        for b in range(4):
            spec += [("conv", oc, 3, first_stride if b == 0 else 1, 1), ("relu",)]
    spec += [("pool", 7, 7), ("flatten",), ("conv", classes, 1, 1, 0)]
    return spec


SPECS = {"lenet": (lenet_spec, (1, 28, 28)), "vgg11": (vgg11_spec, (3, 32, 32)), "resnet18": (resnet18_spec, (3, 224, 224))}


def niti_xavier_int8(rng, oc, c, kh, kw):
    """the reference's weight recipe with an injectable seed (nn/Distributions.cpp:25-50,
    nn/Initializer.cpp:118-134): N(0, sqrt(2/(fan_in+fan_out))) -> /max|w| * 127 -> round;
    wscale = ceil(log2(max|w|)) - 7.  Returns (OIHW int8, wscale)."""
    fan_in, fan_out = c * kh * kw, oc * kh * kw
    w = rng.normal(0.0, np.sqrt(2.0 / (fan_in + fan_out)), size=(oc, c, kh, kw)).astype(np.float32)
    r = float(np.abs(w).max())
    wscale = int(np.ceil(np.log2(r))) - 7
    return np.round(w / r * 127.0).astype(np.int8), wscale


class _Op:
    pass


def fuse_spec(spec):
    """("conv", ...) directly followed by ("relu",) becomes one op: the ReLU is folded into the conv's
    rescale store (niti_rescale_act_fused), and its gradient gate into whatever produces the gradient."""
    out, i = [], 0
    while i < len(spec):
        it = spec[i]
        if it[0] == "conv":
            relu = i + 1 < len(spec) and spec[i + 1][0] == "relu"
            out.append(("conv",) + tuple(it[1:]) + (relu,))
            i += 2 if relu else 1
        else:
            out.append(it); i += 1
    return out


class NitiNet:
    """Device-resident int8 training step.  `fuse=False` keeps one kernel per reference op (the literal
    op sequence); `fuse=True` (default) folds ReLU / ReLU-grad / max-pool / max-pool-grad into neighbouring stores and runs
    consecutive forward convolutions as one persistent chain launch.  Both are bit-identical (tests/test_gpu_step.py)."""

    def __init__(self, nb: NitiB200, spec, batch, in_chw, device=0, dist=None, fuse=True, pad_c=16, chain=True):
        self.nb, self.batch = nb, batch
        self.dev = device if isinstance(device, torch.device) else torch.device("cuda", device)
        self.dist = dist if (dist is not None and dist.is_initialized() and dist.get_world_size() > 1) else None
        self.timer = None                      # optional callable(kind, ops, fn) used by bench.py
        c, h, w = in_chw
        self.in_c = c
        # Zero-pad channel counts to the tensor cores' granule (16): the input channels of the first conv (1 / 3 -> 16) and,
        # where the backend can address padded logit rows, every conv's output channels (12 classes -> 16).
        # Exact: padded activations are 0, padded weights start at 0 and their dW is 0 (a zero activation column on one
        # side, a zero dY column on the other), so they stay 0, and neither max|acc| nor any real element changes.
        c_real = c
        if pad_c and c % pad_c != 0:
            c = (c + pad_c - 1) // pad_c * pad_c
        pad_oc = pad_c if (fuse and pad_c and hasattr(nb, "loss_grad_strided")) else 0
        i8 = dict(dtype=torch.int8, device=self.dev)
        i32 = dict(dtype=torch.int32, device=self.dev)
        self.x_in = torch.zeros((batch, h, w, c), **i8)
        self.ex_in = torch.zeros(1, **i8)
        self.labels = torch.zeros(batch, **i32)
        self.ops, self.convs = [], []
        cur, cur_e = self.x_in, self.ex_in
        dw_total = 0
        fused = fuse_spec(spec) if fuse else [it + ((False,) if it[0] == "conv" else ()) for it in spec]
        for item in fused:
            kind = item[0]
            L = _Op(); L.kind = kind; L.relu = False; L.alias = False
            L.fwd_fused = L.bwd_fused = False              # pool ops: done inside the neighbouring conv launches
            L.pool_op = None                                # conv ops: the pool folded into the forward store
            if kind == "conv":
                _, oc, k, st, pad, relu = item
                oc_p = (oc + pad_oc - 1) // pad_oc * pad_oc if pad_oc else oc
                if pad_oc and os.environ.get("NITI_B200_PAD_OC64", "1") != "0":
                    # Output channels are the next layer's input channels and this layer's dX contraction length: whole
                    # 64-channel K blocks keep every layer on the TMA-fed fused kernels (LeNet's 20 / 52 / 12, a 12- or 1000-way
                    # classifier) instead of the cp.async kernel + transpose + rescale launches.  Zero rows / columns: exact.
                    oc_p = (oc_p + 63) // 64 * 64
                L.d = make_desc(batch, c, h, w, oc_p, k, stride=st, pad=pad)
                L.c_real, L.oc_real = c_real, oc
                L.relu = relu
                L.x, L.ex = cur, cur_e
                L.w = torch.zeros((oc_p, k, k, c), **i8)
                # transposed copy [c][kh][kw][oc]: the K-major B operand of the tensor-core dX kernels, kept in sync
                # by the batched update (no per-step transposes); the first conv has no dX
                L.wt = torch.zeros((c, k, k, oc_p), **i8) if (fuse and self.convs) else None
                L.ew = torch.zeros(1, **i8)
                L.y = torch.zeros((batch, L.d.oh, L.d.ow, oc_p), **i8)
                L.ey = torch.zeros(1, **i8)
                L.n_y, L.n_x, L.n_w = L.y.numel(), cur.numel(), L.w.numel()
                L.ops = 2.0 * batch * L.d.oh * L.d.ow * oc * L.c_real * k * k          # algorithmic: unpadded
                L.dw_off = dw_total
                dw_total += (L.n_w + 63) // 64 * 64
                L.index = len(self.convs)
                self.convs.append(L)
                cur, cur_e, c, c_real, h, w = L.y, L.ey, oc_p, oc, L.d.oh, L.d.ow
            elif kind == "relu":
                L.x = cur; L.y = torch.zeros_like(cur); cur = L.y
            elif kind == "pool":
                _, k, s = item
                L.k, L.s, L.geom = k, s, (batch, h, w, c)
                oh, ow = (h - k) // s + 1, (w - k) // s + 1
                L.x = cur; L.y = torch.zeros((batch, oh, ow, c), **i8)
                cur, h, w = L.y, oh, ow
            elif kind == "flatten":
                # the reference flattens in NCHW order (_Convert(NCHW) + _Reshape, mnistTrain.cpp:170-171); with h*w == 1
                # that is the same bytes as NHWC, so the op is an alias (no copy, and the gradient is written in place)
                L.geom = (batch, c, h, w)
                L.x = cur
                L.alias = bool(fuse) and h * w == 1
                L.y = cur.view(batch, 1, 1, c) if L.alias else torch.zeros((batch, 1, 1, c * h * w), **i8)
                cur, c, c_real, h, w = L.y, c * h * w, c_real * h * w, 1, 1
            self.ops.append(L)
        self.logits_buf, self.e_logits, self.classes = cur, cur_e, c_real
        self.logits = cur[..., :c_real]                      # real classes (the buffer's rows may be padded to 16)
        for i, L in enumerate(self.ops):                     # one gradient tensor per op output
            L.dy = self.ops[i - 1].dy.view_as(L.y) if L.alias else torch.zeros_like(L.y)
        self.acc = torch.zeros(max(max(L.n_y, L.n_x) for L in self.convs), **i32)   # fwd / dX accumulators
        self.dw_acc = torch.zeros(dw_total, **i32)                                   # one bucket: all dW accumulators
        self.maxabs = torch.zeros(4 * len(self.convs) + 4, **i32)                    # one slot per reduction
        self.batched_update = bool(fuse)
        self.fused_ops = False       # set below: whole-op calls (one launch where the kernels allow)
        nconv = len(self.convs)
        # data parallel: exchange the exponent scalars inside the rescale kernels over NVLink peer memory when the
        # backend offers it (lib.NitiB200.dp_connect); otherwise one all_reduce(MAX) per scalar
        self.peer_max = False
        if self.dist is not None and fuse and hasattr(nb, "dp_connect"):
            if not getattr(nb, "dp_connected", False):
                nb.dp_connect(self.dist)
            self.peer_max = True
        # dW all-reduce buckets, closed in backward order (last conv first) once they hold >= 8 MiB, so the exchange
        # of the deep layers overlaps the rest of backward: list of (lowest conv index, start word, end word)
        self._buckets, hi_end, size = {}, dw_total, 0
        # ... and the bucket that closes LAST holds only the first conv (an image layer: kilobytes), so no large exchange starts
        # behind the last dW (8xB200: the 41 us between the last dW and the update did not move — launch-chain latency, not bytes)
        for L in reversed(self.convs):
            size += L.n_w * 4
            if size >= (8 << 20) or L.index == 0 or (L.index == 1 and len(self.convs) >= 3):
                self._buckets[L.index] = (L.dw_off, hi_end)
                hi_end, size = L.dw_off, 0
        self._pending = []
        # Gradient sum without a collective library: the accumulators move into a peer-mapped window and every bucket is
        # added by one niti_dp_allreduce_i32 launch on a third stream (NCCL at 16 CTAs needed ~100 us per 9 MB bucket on
        # 8 GPUs and left a 190 us tail exposed: profiles/timeline_r2_vgg_8gpu_nccl.txt).  NITI_B200_DP_ALLREDUCE=nccl
        # keeps torch.distributed.all_reduce, and so does a job that reserves no SMs (the peer-sum CTAs need SMs of their own).
        self.peer_sum = (self.peer_max and hasattr(nb, "dp_window") and self.dist.get_world_size() > 1
                         and os.environ.get("NITI_B200_DP_ALLREDUCE", "peer") != "nccl"
                         and (not hasattr(nb, "dp_reserved_sms") or nb.dp_reserved_sms() > 0))
        if self.peer_sum:
            self.dw_acc = nb.dp_window(self.dist, dw_total)
            self.comm = torch.cuda.Stream(self.dev)
            self._bucket_slot = {idx: k for k, idx in enumerate(sorted(self._buckets, reverse=True))}
        self.fused_ops = bool(fuse) and hasattr(nb, "conv_fwd_fused") and (self.dist is None or self.peer_max)
        # (acc, w, wt, max-abs slot, oc, taps, c, sub) per layer: the segments of the one-launch weight update
        self._upd_layers = [(self.dw_acc[L.dw_off: L.dw_off + L.n_w], L.w, L.wt,
                             self.maxabs[2 * nconv + L.index: 2 * nconv + L.index + 1],
                             L.d.oc, L.d.kh * L.d.kw, L.d.c, 2) for L in self.convs]
        self._upd_segs = nb.make_update_segs(self._upd_layers) if self.batched_update else None
        # Early update: once backward has passed layer k, nothing reads the weights of the layers behind it any more and
        # their dW sums are complete, so their rule B + SGD runs on a third stream (two barrier-free launches) beside the
        # rest of backward; only the shallow layers' update stays in the step's tail (VGG-11: 26 -> 12 us; measured 463.0 ->
        # 460.8 us/step: backward is bound by SM time, not by the tail).  Single device only: data parallel, the gradient sums arrive bucket by bucket.
        self._early_from = None
        if (self.batched_update and self.dist is None and hasattr(nb, "update_batched_phases") and len(self.convs) >= 4
                and os.environ.get("NITI_B200_EARLY_UPDATE", "1") != "0"):
            k = len(self.convs) // 2
            # small nets (LeNet) are launch-bound: two more launches cost more than the shorter tail saves
            if sum(L.n_w for L in self.convs[k:]) >= int(os.environ.get("NITI_B200_EARLY_UPDATE_MIN", str(2 << 20))):
                self._early_from = k
                self.upd_stream = torch.cuda.Stream(self.dev)
                self._upd_segs_early = nb.make_update_segs(self._upd_layers[k:])
                self._upd_segs_late = nb.make_update_segs(self._upd_layers[:k])
        # dW beside dX: both only read dY, and the dX chain is the critical path of backward, so dW (and, data parallel,
        # the bucket all-reduces behind it) is enqueued on a second stream.  Needs the barrier-free accumulate entry
        # point (niti_conv_dw_accumulate) and the gradient bucket cleared once per step (done beside forward).
        self.overlap_dw = bool(fuse) and hasattr(nb, "conv_dw_accumulate") and self.dev.type == "cuda"
        self.side = torch.cuda.Stream(self.dev) if self.overlap_dw else None
        self._plan_fusions(chain)
        if self._early_from is not None and (not self.overlap_dw or self._bwd_chain is not None):
            self._early_from = None                # needs the dW stream; the opt-in backward chain reorders dX / dW

    def _plan_fusions(self, chain):
        """Decide which pools fold into the conv before them (forward store) and into the dX of the conv after them
        (max-pool gradient + ReLU gate), and group consecutive fused-eligible forward convs into chain launches."""
        nb, nconv = self.nb, len(self.convs)
        self._fwd_plan = [("op", i) for i in range(len(self.ops))]
        for i, L in enumerate(self.ops):
            # the op that really produced this op's input (alias flattens are views): its folded ReLU gates our dX
            j = i - 1
            while j >= 0 and self.ops[j].alias:
                j -= 1
            L.src = self.ops[j] if j >= 0 else None
            L.pg = None
        self._bwd_chain = None
        if not (self.fused_ops and hasattr(nb, "conv_fwd_fused_pool")):
            return

        def pool_2x2(P):
            if P.kind != "pool":
                return False
            n, h, w, c = P.geom
            return P.k == 2 and P.s == 2 and h % 2 == 0 and w % 2 == 0 and c % 16 == 0
        for i, L in enumerate(self.ops):
            nxt = self.ops[i + 1] if i + 1 < len(self.ops) else None
            if L.kind == "conv" and L.relu and nxt is not None and pool_2x2(nxt):
                L.pool_op = nxt; nxt.fwd_fused = True
            if L.kind == "conv" and L.index > 0:
                # a 2x2/2 pool in front of this conv folds its gradient into this conv's dX launch
                P = L.src
                if P is not None and pool_2x2(P) and P.src is not None and L.wt is not None:
                    src = P.src                                # the op whose output was pooled
                    L.pg = (P, src.dy, bool(src.relu))         # pool op, gradient target, ReLU gate of the pooled tensor
                    P.bwd_fused = True
        # forward chains
        if not (chain and hasattr(nb, "chain_create")):
            return
        plan, run = [], []

        def flush():
            if len(run) >= 2:
                plan.append(("chain", nb.chain_create([self._chain_op(self.ops[i]) for i in run]), list(run)))
            else:
                plan.extend(("op", i) for i in run)
            run.clear()
        for i, L in enumerate(self.ops):
            if L.kind == "conv" and nb.chain_op_eligible(self._chain_op(L)):
                run.append(i)
            elif (L.kind == "pool" and L.fwd_fused) or L.alias:
                if not run:
                    plan.append(("op", i))                    # nothing to launch, kept for symmetry
            else:
                flush(); plan.append(("op", i))
        flush()
        self._fwd_plan = plan
        self._plan_backward_chain()

    # Backward chain (opt-in, NITI_B200_BWD_CHAIN_PIXELS=<pixels per GPU>): the dX ops of the deepest layers (<= that many pixels)
    # as ONE persistent launch, their dW launches behind it.  Measured on VGG-11 b256 (profiles/ab/knob_bwd_chain.txt): off
    # 463.3 us/step, 256 px (3 FC layers) 469.2, 1024 px (+ conv8, conv7) 479.6, 4096 px 484.5 — the launches it saves cost less
    # than the dW kernels it pushes behind the chain (they no longer run beside these dX ops and pile up after the last dX).
    # Default: off.
    def _dx_chain_op(self, L):
        nconv = len(self.convs)
        prev = self.ops[self.ops.index(L) - 1]
        gate = L.src is not None and L.src.relu
        pg = getattr(L, "pg", None)
        slot = self.maxabs[nconv + L.index: nconv + L.index + 1]
        if pg is not None:
            return self.nb.chain_op(L.d, 1, L.dy, L.wt, pg[1], slot, nconv + L.index, pool_x=pg[0].x, pool_gate=pg[2])
        return self.nb.chain_op(L.d, 1, L.dy, L.wt, prev.dy, slot, nconv + L.index, gate=L.x if gate else None)

    def _plan_backward_chain(self):
        """The deepest layers (classifier, FC, 2x2 / 4x4 feature maps) in backward order, as long as each dX is fused-eligible,
        small and writes straight into the gradient tensor the next one reads."""
        self._bwd_chain = None
        nb = self.nb
        max_pixels = int(os.environ.get("NITI_B200_BWD_CHAIN_PIXELS", "0"))
        if not (self.fused_ops and hasattr(nb, "chain_create")) or max_pixels <= 0:
            return
        layers = []
        for L in reversed(self.convs):
            if L.index == 0 or L.wt is None:
                break
            i = self.ops.index(L)
            prev = self.ops[i - 1]
            pg = getattr(L, "pg", None)
            # the op in front must be the producer itself (ReLU folded), an alias flatten, or a pool folded into this dX
            simple = prev.kind == "conv" or (prev.kind == "flatten" and prev.alias) or (prev.kind == "pool" and pg is not None)
            if not simple or L.d.n * L.d.h * L.d.w > max_pixels or not nb.chain_op_eligible(self._dx_chain_op(L)):
                break
            layers.append(L)
        if len(layers) >= 2:
            self._bwd_chain = (nb.chain_create([self._dx_chain_op(L) for L in layers]), layers)

    def _chain_op(self, L):
        P = L.pool_op
        return self.nb.chain_op(L.d, 0, L.x, L.w, L.y, self.maxabs[L.index: L.index + 1], L.index, ex=L.ex, ew=L.ew, ey=L.ey,
                                relu=L.relu, out_pool=P.y if P is not None else None)

    # ---- parameters -------------------------------------------------------------------------------
    def init_weights(self, seed=1):
        rng = np.random.default_rng(seed)
        ws = [niti_xavier_int8(rng, L.oc_real, L.c_real, L.d.kh, L.d.kw) for L in self.convs]
        self.load_weights(ws)
        return ws

    def load_weights(self, ws):
        """ws: list of (OIHW int8 numpy, wscale)"""
        for L, (w, e) in zip(self.convs, ws):
            L.w.zero_()
            L.w[:L.oc_real, ..., :L.c_real].copy_(torch.from_numpy(np.ascontiguousarray(np.transpose(w, (0, 2, 3, 1)))))
            if L.wt is not None:
                L.wt.copy_(L.w.permute(3, 1, 2, 0))
            L.ew.fill_(int(e))

    def weights_oihw(self):
        return [np.ascontiguousarray(np.transpose(L.w[:L.oc_real, ..., :L.c_real].cpu().numpy(), (0, 3, 1, 2))) for L in self.convs]

    def set_input(self, x_nhwc, ascale, labels, non_blocking=True):
        """host (pinned) or device tensors -> the step's input buffers (the e2e H2D leg)"""
        if x_nhwc.shape[-1] == self.x_in.shape[-1]:
            self.x_in.copy_(x_nhwc, non_blocking=non_blocking)
        else:
            self.x_in[..., :x_nhwc.shape[-1]].copy_(x_nhwc.to(self.dev, non_blocking=non_blocking))
        self.labels.copy_(labels, non_blocking=non_blocking)
        if torch.is_tensor(ascale):
            self.ex_in.copy_(ascale, non_blocking=non_blocking)
        else:
            self.ex_in.fill_(int(ascale))

    # ---- helpers ----------------------------------------------------------------------------------
    def _timed(self, kind, ops, fn):
        if self.timer is not None:
            self.timer(kind, ops, fn)
        else:
            fn()

    def _beside(self):
        """context that enqueues on the side stream once everything already on the current stream is done; the
        per-kernel timing pass of bench.py (self.timer set) keeps everything on one stream"""
        if not self.overlap_dw or self.timer is not None:
            return contextlib.nullcontext()
        self.side.wait_stream(torch.cuda.current_stream(self.dev))
        return torch.cuda.stream(self.side)

    def _allreduce_max(self, slot):
        if self.dist is not None:
            self.dist.all_reduce(slot, op=self.dist.ReduceOp.MAX)

    def _rescale(self, n, slot_index, out, ex=None, ew=None, ey=None, relu=False, gate=None):
        """phase 2 of a forward / dX op: global max|acc| -> shared exponent -> PSTO -> int8 (+ fused ReLU / gate)"""
        nb = self.nb
        slot = self.maxabs[slot_index: slot_index + 1]
        if self.peer_max:
            nb.rescale_act_dp(self.acc, n, slot, out, ex, ew, ey, relu=relu, gate=gate, dp_slot=slot_index)
            return
        self._allreduce_max(slot)
        if relu or gate is not None:
            nb.rescale_act_fused(self.acc, n, slot, out, ex, ew, ey, relu=relu, gate=gate)
        else:
            nb.rescale_act(self.acc, n, slot, out, ex, ew, ey)

    # ---- the step ---------------------------------------------------------------------------------
    def forward(self):
        nb = self.nb
        nb.zero_u32(self.maxabs, self.maxabs.numel())
        if self.peer_max:
            nb.dp_step_begin()
        if self.overlap_dw:
            with self._beside():
                self.dw_acc.zero_()
        for unit in self._fwd_plan:
            if unit[0] == "chain":
                # consecutive fused forward convs (+ ReLU + pool) in ONE persistent launch (csrc/conv_chain.cu)
                layers = [self.ops[i] for i in unit[2]]
                self._timed("fwd", sum(L.ops for L in layers), lambda: nb.chain_run(unit[1]))
                continue
            L = self.ops[unit[1]]
            if L.kind == "conv":
                slot = self.maxabs[L.index: L.index + 1]
                if self.fused_ops:
                    # whole op (conv + exponent + rescale + ReLU [+ max pool]) behind one call: one launch for stem layers
                    # (conv_stem.cu) and whenever the output tiles fit in tensor memory (conv_chain.cu)
                    if L.index == 0 and L.c_real <= 4 and hasattr(nb, "conv_stem_fwd"):
                        # image layer: channels >= c_real of x_in are zero padding (set_input), conv_stem2.cu packs the live ones
                        self._timed("fwd", L.ops, lambda: nb.conv_stem_fwd(L.d, L.c_real, L.x, L.w, L.ex, L.ew, L.y, L.ey, L.relu,
                                                                          L.pool_op.y if L.pool_op is not None else None,
                                                                          self.acc, slot, dp_slot=L.index))
                    elif L.pool_op is not None:
                        self._timed("fwd", L.ops, lambda: nb.conv_fwd_fused_pool(L.d, L.x, L.w, L.ex, L.ew, L.y, L.ey, L.relu,
                                                                                L.pool_op.y, self.acc, slot, dp_slot=L.index))
                    else:
                        self._timed("fwd", L.ops, lambda: nb.conv_fwd_fused(L.d, L.x, L.w, L.ex, L.ew, L.y, L.ey, L.relu,
                                                                           self.acc, slot, dp_slot=L.index))
                    continue
                self._timed("fwd", L.ops, lambda: nb.conv_fwd_acc(L.d, L.x, L.w, self.acc, slot))
                self._rescale(L.n_y, L.index, L.y, L.ex, L.ew, L.ey, relu=L.relu)
            elif L.kind == "relu":
                nb.relu(L.x, L.y, L.x.numel())
            elif L.kind == "pool":
                if L.fwd_fused:
                    continue
                n, h, w, c = L.geom
                nb.maxpool(L.x, n, h, w, c, L.k, L.s, L.y)
            elif L.kind == "flatten":
                if L.alias:
                    continue
                n, c, h, w = L.geom
                if h * w == 1:
                    L.y.copy_(L.x.view_as(L.y))
                else:
                    nb.nhwc_to_nchw(L.x, n, c, h, w, L.y)

    def backward_and_update(self):
        nb = self.nb
        nconv = len(self.convs)
        last = self.ops[-1]
        stride = self.logits_buf.shape[-1]
        if stride != self.classes:
            nb.loss_grad_strided(self.logits_buf, self.e_logits, self.labels, self.batch, self.classes, stride, last.dy)
        else:
            nb.loss_grad(self.logits_buf, self.e_logits, self.labels, self.batch, self.classes, last.dy)
        if last.relu:
            nb.relu_grad(last.y, last.dy, last.dy, last.y.numel())
        for idx in range(len(self.ops) - 1, -1, -1):
            L = self.ops[idx]
            dy = L.dy
            prev = self.ops[idx - 1] if idx > 0 else None
            dx = prev.dy if prev is not None else None       # gradient w.r.t. this op's input
            gate = L.src is not None and L.src.relu          # input came through a folded ReLU
            if L.kind == "conv":
                chained = self._bwd_chain is not None and L in self._bwd_chain[1]
                if chained and L is not self._bwd_chain[1][0]:
                    continue                                   # its dX ran in the chain, its dW right behind the chain
                self._launch_dw(L)
                if self._early_from is not None and L.index == self._early_from - 1 and self.timer is None:
                    # behind dX of layer `early_from` (already enqueued on this stream) and behind every deeper dW (side stream);
                    # on its own stream, so that the shallow layers' dW launches do not queue up behind it
                    self.upd_stream.wait_stream(torch.cuda.current_stream(self.dev))
                    self.upd_stream.wait_stream(self.side)
                    with torch.cuda.stream(self.upd_stream):
                        nb.update_batched_phases(self._upd_segs_early)
                if chained:
                    # the latency-bound dX ops of the deepest layers in one persistent launch, then their dW launches
                    layers = self._bwd_chain[1]
                    self._timed("dx", sum(M.ops for M in layers), lambda: nb.chain_run(self._bwd_chain[0]))
                    for M in layers[1:]:
                        self._launch_dw(M)
                    continue
                if dx is not None:
                    slot = self.maxabs[nconv + L.index: nconv + L.index + 1]
                    if self.fused_ops:
                        pg = getattr(L, "pg", None)
                        if pg is not None:
                            # dX + the max-pool gradient (+ ReLU gate) of the pool in front of this conv, one launch
                            self._timed("dx", L.ops, lambda: nb.conv_dx_fused_poolgrad(L.d, dy, L.w, L.wt, pg[1], pg[0].x, pg[2], pg[0].dy,
                                                                                       self.acc, slot, dp_slot=nconv + L.index))
                        else:
                            self._timed("dx", L.ops, lambda: nb.conv_dx_fused(L.d, dy, L.w, L.wt, dx, L.x if gate else None,
                                                                              self.acc, slot, dp_slot=nconv + L.index))
                        continue
                    if L.wt is not None:
                        self._timed("dx", L.ops, lambda: nb.conv_dx_acc_wt(L.d, dy, L.w, L.wt, self.acc, slot))
                    else:
                        self._timed("dx", L.ops, lambda: nb.conv_dx_acc(L.d, dy, L.w, self.acc, slot))
                    self._rescale(L.n_x, nconv + L.index, dx, gate=L.x if gate else None)
            elif L.kind == "relu":
                nb.relu_grad(L.x, dy, dx, L.x.numel())
                if gate:
                    nb.relu_grad(L.x, dx, dx, L.x.numel())
            elif L.kind == "pool":
                if L.bwd_fused and self.fused_ops:
                    continue                                   # done by the dX launch of the conv behind this pool
                n, h, w, c = L.geom
                nb.maxpool_grad_gated(L.x, L.y, dy, n, h, w, c, L.k, L.s, dx, gate)
            elif L.kind == "flatten":
                if L.alias:
                    continue                                   # L.dy is a view of the producer's gradient tensor
                n, c, h, w = L.geom
                if h * w == 1:
                    dx.copy_(dy.view_as(dx))
                else:
                    nb.nchw_to_nhwc(dy, n, c, h, w, dx)
                if gate:
                    nb.relu_grad(L.x, dx, dx, L.x.numel())
        # gradient exchange (one bucket), then shared exponent + PSTO + int8 SGD, fused per layer
        if self.overlap_dw:
            torch.cuda.current_stream(self.dev).wait_stream(self.side)
        if self._early_from is not None and self.timer is None:
            torch.cuda.current_stream(self.dev).wait_stream(self.upd_stream)
        if self.peer_sum:
            torch.cuda.current_stream(self.dev).wait_stream(self.comm)
        for work in self._pending:
            work.wait()
        self._pending = []
        if self.batched_update:
            if self._early_from is not None and self.timer is None:
                nb.update_batched(self._upd_segs_late)        # the deep layers were updated beside backward
            else:
                nb.update_batched(self._upd_segs)             # one launch: max|acc|, rule B, SGD, transposed copies
            return
        for L in self.convs:
            acc_dw = self.dw_acc[L.dw_off: L.dw_off + L.n_w]
            slot = self.maxabs[2 * nconv + L.index: 2 * nconv + L.index + 1]
            nb.absmax(acc_dw, L.n_w, slot)
            nb.rescale_grad(acc_dw, L.n_w, slot, 2, None, L.w)

    def _launch_dw(self, L):
        """dW of one layer on the side stream (its dY is complete on the main stream) and, data parallel, the exchange of the
        gradient bucket this layer closes"""
        nb = self.nb
        acc_dw = self.dw_acc[L.dw_off: L.dw_off + L.n_w]
        with self._beside():
            if self.overlap_dw:
                self._timed("dw", L.ops, lambda: nb.conv_dw_accumulate(L.d, L.x, L.dy, acc_dw))
            else:
                self._timed("dw", L.ops, lambda: nb.conv_dw_acc(L.d, L.x, L.dy, acc_dw, None))
            if self.dist is not None and L.index in self._buckets:
                lo, hi = self._buckets[L.index]    # this bucket's dW sums are complete: start its exchange
                if self.peer_sum:
                    self.comm.wait_stream(torch.cuda.current_stream(self.dev))
                    with torch.cuda.stream(self.comm):
                        nb.dp_allreduce(lo, (hi - lo + 3) // 4 * 4, self._bucket_slot[L.index])
                else:
                    self._pending.append(self.dist.all_reduce(self.dw_acc[lo:hi], op=self.dist.ReduceOp.SUM, async_op=True))

    def step(self):
        self.forward()
        self.backward_and_update()

    def check(self):
        """Host sync point: poll the device-side error flags (a peer exchange or a grid barrier that gave up) and raise.
        Synchronises the device.  Data-parallel callers run it after warm-up and at the end of every timed / logged
        region: a timed-out exchange means the replicas no longer share exponents."""
        torch.cuda.synchronize(self.dev)
        self.nb.check_errors()

    def capture(self):
        """CUDA-graph the step.  Returns a callable that replays it.  Data-parallel: the gradient sum (by default the
        peer-sum launches on the third stream, NCCL all-reduces with NITI_B200_DP_ALLREDUCE=nccl) is captured into the graph
        with the kernels; every rank must capture and replay in lock-step.  A replay loop must still poll check() at its
        sync points: a peer that stops answering only raises the sticky device flag."""
        s = torch.cuda.Stream(self.dev)
        s.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(s):
            self.step()                                   # warm-up: workspaces grow here, not under capture
        torch.cuda.current_stream(self.dev).wait_stream(s)
        self.check()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.step()
        return g.replay

    def algorithmic_ops(self):
        """2*M*N*K (unpadded) of every contraction of one step: fwd + dW per conv, dX for all but the first"""
        return sum(2 * L.ops + (L.ops if L.index > 0 else 0.0) for L in self.convs)
