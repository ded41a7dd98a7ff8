/*
 * niti_b200.h — C ABI of libniti_b200.so: the B200 (sm_100a) implementation of Mandheling's NITI
 * int8 training hot path.  Plain C, raw device pointers + POD descriptors, int return codes
 * (0 = success), no torch/MNN types.  Modelled on the reference's own accelerator boundary, the
 * `struct HexagonNN` table of plain C calls loaded with dlopen/dlsym
 * (execution-engine/include/MNN/hexagon/hexagon_nn/hexagon_implementation.h:24-130,
 *  execution-engine/source/backend/cpu/hexagon_implementation.cpp:52-113), but op-level instead of
 * graph-level.  The MNN Backend/Execution creator set in mnn_backend/ and the ctypes mirror in
 * mandheling_dsp_training_b200/ are both thin callers of exactly these entry points.
 *
 * Data layout in HBM (all device memory, caller-owned):
 *   activations / gradients : int8  NHWC  [n][h][w][c]
 *   weights / weight grads  : int8  OHWI  [oc][kh][kw][c]
 *   accumulators            : int32, same index order as the int8 tensor they rescale into
 *   exponents               : int8 scalars in device memory (quirk Q7: carried beside each tensor)
 *   max-abs                 : uint32 scalar in device memory (input of the shared-exponent step)
 * Everything is asynchronous on the given stream and CUDA-graph capturable: no host sync and no
 * host-visible scalar on the path.  Allocation: the phase-1/phase-2 and fused entry points allocate nothing; the whole-op
 * conveniences that take no scratch pointer (niti_conv_fwd/dx/dw, niti_matmul_nt, niti_conv_dw_split, a NULL `acc`, a
 * NULL `wt`) and niti_conv_dx_fused on a STRIDED layer (its zero-inserted dY) grow a context workspace with cudaMalloc on first use — call them once eagerly before capturing (under
 * capture the growth fails with NITI_ERR_CUDA rather than corrupting the graph), and re-capture after a call that grew it.
 *
 * Concurrency rule: the persistent kernels behind niti_conv_fwd_fused, niti_conv_dx_fused, niti_conv_fwd/dx[_acc] (split-K),
 * niti_conv_dw[_acc] and niti_update_batched hold a grid-wide barrier whose counters belong to the context.  At most ONE
 * such call of a context may be in flight at a time: issue them on one stream, or order streams with events.  (A second
 * one running concurrently could share SMs with the first and neither barrier would ever fill; the spin is bounded and a
 * barrier that gives up sets a sticky flag, niti_sync_error.)  niti_conv_dw_accumulate holds no barrier and may run beside
 * them — that is how a training step overlaps dW with the dX chain.
 *
 * There is NO CPU fallback: every entry point fails with NITI_ERR_CUDA when no device is usable.
 */
#ifndef NITI_B200_H
#define NITI_B200_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct niti_ctx niti_ctx;
typedef void* niti_stream;           /* cudaStream_t */

enum {
    NITI_OK = 0,
    NITI_ERR_ARG = 1,                /* bad descriptor / null pointer          */
    NITI_ERR_CUDA = 2,               /* CUDA runtime error (see niti_last_error) */
    NITI_ERR_UNSUPPORTED = 3         /* shape outside every kernel's contract  */
};

/* Convolution geometry; mirrors MNN's Convolution2DCommon as the reference's ops read it
 * (schema/default/UserDefine.fbs:27-33; NITI_Conv_Int8.cpp:94-112): symmetric stride/pad/dilate. */
typedef struct {
    int32_t n, c, h, w;              /* input  [n,h,w,c]          */
    int32_t oc, kh, kw;              /* weight [oc,kh,kw,c]       */
    int32_t stride, pad, dil;
} niti_conv_desc;

/* Which kernel family to use.  AUTO picks, in order: TMA-fed tcgen05 (channel counts multiple of 64),
 * cp.async-fed tcgen05 (multiple of 16), SIMT.  The others force one family (tests, benchmarks). */
enum { NITI_PATH_AUTO = 0, NITI_PATH_SIMT = 1, NITI_PATH_TC = 2, NITI_PATH_TMA = 3, NITI_PATH_TMA2 = 4 };

/* ---- context ------------------------------------------------------------------------------- */
int         niti_ctx_create(int device, niti_ctx** out);
int         niti_ctx_destroy(niti_ctx* ctx);
const char* niti_last_error(niti_ctx* ctx);
int         niti_set_path(niti_ctx* ctx, int path);
/* Exponent prediction of the fused forward / dX kernels.  mode 0: off; 1: adaptive (default; NITI_B200_SPECULATE sets it at
 * creation); 2: always.  NITI's shared exponent needs max|acc| of the whole output tensor before anything can be stored.
 * With prediction, a fused kernel stores in the same sweep that reduces the maximum, using the shift the same tensor (same
 * dp_slot) had the last time it was computed, and re-stores (stem layer: recomputes) only when the true maximum lands in
 * another bit width.  Adaptive = speculate when a miss is cheap (several tiles per CTA pair, stem layer) or the slot's last
 * NITI_B200_SPEC_CONFIDENT predictions held (default 0: always).  The stored int8 values and the exponent are those of the true maximum in every mode (tests compare the
 * modes bit for bit). */
int         niti_set_speculate(niti_ctx* ctx, int mode);
/* predictions that held / missed / layers that did not speculate, since the context was created (synchronises the device;
 * skipped may be NULL) */
int         niti_speculate_stats(niti_ctx* ctx, unsigned long long* hits, unsigned long long* misses, unsigned long long* skipped);
int         niti_version(void);
/* kernels launched through this context since creation (bench.py's gpu_launches evidence) */
int64_t     niti_launch_count(niti_ctx* ctx);
/* name of the kernel family the last conv/matmul call used: "simt", "tc", "tma", "tma2", "fused", "chain", "stem" or "stem2" */
const char* niti_last_path(niti_ctx* ctx);

/* ---- phase 1: int32 accumulators + running max|acc| ---------------------------------------
 * Replaces the im2col + NITI_Int8GemmKernel loops of
 *   NITI_Conv_Int8::onExecute         (source/backend/cpu/NITI_Conv_Int8.cpp:212-249)
 *   NITI_DeConv_Int8::onExecute       (NITI_DeConv_Int8.cpp:208-290; rot180 folded into indexing)
 *   NITI_GradientConv_Int8::onExecute (NITI_GradientConv_Int8.cpp:218-268)
 *   NITI_Matmul_Int8::onExecute       (NITI_Matmul_Int8.cpp:186-208)
 * `maxabs` (optional) receives max(|acc|) via atomicMax and must be zeroed by the caller
 * (niti_zero_u32) or by a previous niti_rescale_* consuming it. */
int niti_conv_fwd_acc(niti_ctx*, const niti_conv_desc*, const int8_t* x, const int8_t* w,
                      int32_t* acc /*[n,oh,ow,oc]*/, uint32_t* maxabs, niti_stream);
int niti_conv_dx_acc(niti_ctx*, const niti_conv_desc*, const int8_t* dy, const int8_t* w,
                     int32_t* acc /*[n,h,w,c]*/, uint32_t* maxabs, niti_stream);
int niti_conv_dw_acc(niti_ctx*, const niti_conv_desc*, const int8_t* x, const int8_t* dy,
                     int32_t* acc /*[oc,kh,kw,c]*/, uint32_t* maxabs, niti_stream);
/* acc += dW into a buffer the caller has zeroed (a training step clears its whole gradient bucket once).  Same
 * contraction as niti_conv_dw_acc (NITI_GradientConv_Int8.cpp:218-268), but the launch holds no grid-wide barrier, so
 * it may run on a second stream beside the dX chain of the same layer (both only read dY). */
int niti_conv_dw_accumulate(niti_ctx*, const niti_conv_desc*, const int8_t* x, const int8_t* dy,
                            int32_t* acc /*[oc,kh,kw,c]*/, niti_stream);
/* niti_conv_dx_acc for callers that keep the transposed weights wt = [c][kh][kw][oc] resident (the K-major B operand
 * of the tensor-core dX kernels; the reference re-derives it per call: rot180 + transpose, NITI_DeConv_Int8.cpp:179-219).
 * niti_update_batched maintains wt next to w, so a training step launches no transposes.  wt may be NULL (then as
 * niti_conv_dx_acc). */
int niti_conv_dx_acc_wt(niti_ctx*, const niti_conv_desc*, const int8_t* dy, const int8_t* w, const int8_t* wt,
                        int32_t* acc /*[n,h,w,c]*/, uint32_t* maxabs, niti_stream);
/* C[e,h] = sum_l A[e,l] * B[h,l]   (A, B row-major int8) */
int niti_matmul_nt_acc(niti_ctx*, int e, int h, int l, const int8_t* a, const int8_t* b,
                       int32_t* acc, uint32_t* maxabs, niti_stream);

/* ---- phase 2: shared exponent + pseudo-stochastic shift-round ------------------------------
 * NITI_RangeEstimate + NITI_MNNPstoShiftInt32 (compute/CommonOptFunction.cpp:1565-1627) with the
 * callers' rules.  Between phase 1 and phase 2 a data-parallel job all-reduces `acc` (sum, dW) or
 * `maxabs` (max, activations). */
int niti_zero_u32(niti_ctx*, uint32_t* p, int count, niti_stream);
int niti_absmax(niti_ctx*, const int32_t* acc, size_t n, uint32_t* maxabs /*atomicMax*/, niti_stream);
/* rule A — forward / dX (NITI_Conv_Int8.cpp:255-307, NITI_DeConv_Int8.cpp:292-329):
 *   shift=b-7; >1: PSTO(shift); ==1: PSTO(2); else raw int8 cast.
 *   exp_out (optional) = (int8)(*exp_a + *exp_b + inc); exp_a/exp_b may be NULL (treated as 0). */
int niti_rescale_act(niti_ctx*, const int32_t* acc, size_t n, const uint32_t* maxabs,
                     int8_t* out, const int8_t* exp_a, const int8_t* exp_b, int8_t* exp_out, niti_stream);
/* rule A with the neighbouring element-wise op folded into the store (one pass less over HBM):
 *   relu != 0 : out = max(out, 0)                 (NITI_CPURelu_Int8 after a forward conv)
 *   gate != 0 : out = gate[i] > 0 ? out : 0       (NITI_CPUReluGrad_Int8 after a dX conv; gate = the
 *               forward tensor the gradient is taken against; relu(x) > 0 <=> x > 0)
 * max|acc| and the exponent are those of the un-gated tensor, exactly as in the unfused sequence. */
int niti_rescale_act_fused(niti_ctx*, const int32_t* acc, size_t n, const uint32_t* maxabs,
                           int8_t* out, const int8_t* exp_a, const int8_t* exp_b, int8_t* exp_out,
                           int relu, const int8_t* gate, niti_stream);
/* Data-parallel variant of niti_rescale_act_fused: the max|acc| that fixes the exponent is the max over all ranks,
 * exchanged INSIDE this kernel over NVLink peer memory (one 8-byte store per peer + a spin on local memory; see
 * csrc/dp.cuh) instead of a NCCL all-reduce per scalar.  dp_slot in [0,256) names the exchange word; a step must use
 * each slot at most once and at least two different slots.  Without niti_dp_connect (world 1) it equals the plain call. */
int niti_rescale_act_dp(niti_ctx*, const int32_t* acc, size_t n, const uint32_t* maxabs,
                        int8_t* out, const int8_t* exp_a, const int8_t* exp_b, int8_t* exp_out,
                        int relu, const int8_t* gate, int dp_slot, niti_stream);
/* peer-memory plumbing for the call above (one process per GPU, all GPUs of one NVLink box):
 *   niti_dp_alloc    allocates this rank's exchange buffer and returns its 64-byte CUDA IPC handle
 *   niti_dp_connect  maps every rank's buffer (handles = world x 64 bytes, rank-ordered, gathered by the host)
 *   niti_dp_step_begin bumps the device-side step epoch; call once at the start of every step (graph-capturable)
 *   niti_dp_error    *out != 0 when a peer did not answer within the spin budget (NITI_B200_DP_TIMEOUT_MS, default 60000):
 *                    results since the last clean poll are invalid.  Hosts poll it at their sync points. */
int niti_dp_alloc(niti_ctx*, unsigned char handle_out[64]);
int niti_dp_connect(niti_ctx*, int rank, int world, const unsigned char* handles);
int niti_dp_close(niti_ctx*);
int niti_dp_step_begin(niti_ctx*, niti_stream);
int niti_dp_error(niti_ctx*, int* out);
/* SMs the library leaves free for the collective kernels running beside the contractions once niti_dp_connect was called
 * with world > 1 (NITI_B200_DP_RESERVE_SMS, default 16).  The collective library must be capped at no more CTAs than this
 * (NCCL_MAX_CTAS): with more, its CTAs wait for SMs held by kernels that are themselves waiting on a peer. */
int niti_dp_reserved_sms(niti_ctx*);
/* Gradient sum over ranks without a collective library.  The reference's dW contracts over the whole batch
 * (NITI_GradientConv_Int8.cpp:218-268); with the batch sharded over g GPUs the int32 dW accumulators of all ranks are
 * added before the shared exponent is taken.  The accumulators live in a library-allocated WINDOW that
 * every peer maps over CUDA IPC; niti_dp_allreduce_i32 sums the words [off, off + n) of all ranks' windows in place in
 * one launch (two-shot over NVLink peer loads/stores, ~niti_dp_reserved_sms CTAs, csrc/dp.cu dp_allreduce_kernel).
 *   niti_dp_window_alloc    after niti_dp_connect: allocates `words` int32 (rounded up to 4), zeroed; returns the device
 *                           pointer and its 64-byte IPC handle
 *   niti_dp_window_connect  maps the peers' windows (handles = world x 64 bytes, rank-ordered)
 *   niti_dp_allreduce_i32   off, n multiples of 4 words; slot in [0,64): every call of one step uses its own slot, the
 *                           same slot on every rank; all ranks must issue the same calls in the same order.  The words
 *                           must not be written on any rank between this call's launch there and its completion there.
 *                           world 1: no-op.  A peer that never arrives sets the niti_dp_error flag. */
int niti_dp_window_alloc(niti_ctx*, size_t words, int32_t** ptr_out, unsigned char handle_out[64]);
int niti_dp_window_connect(niti_ctx*, const unsigned char* handles);
int niti_dp_window_free(niti_ctx*);
int niti_dp_allreduce_i32(niti_ctx*, size_t off, size_t n, int slot, niti_stream);
/* *out != 0 when a grid-wide barrier of one of this context's persistent kernels gave up (see the concurrency rule at the
 * top): every result since the last clean poll is invalid.  Synchronises with the device. */
int niti_sync_error(niti_ctx*, int* out);
/* rule B — dW / matmul (NITI_GradientConv_Int8.cpp:274-296 sub=2; NITI_Matmul_Int8.cpp:216-228 sub=3):
 *   b==0: zeros; else PSTO(b-sub).  dw_out optional; when w_inout is given the int8 SGD step
 *   w <- clip127(w - dw) (optimizer/NITI_SGD.hpp:49-52, BinaryUtils.hpp:276-299) is fused. */
int niti_rescale_grad(niti_ctx*, const int32_t* acc, size_t n, const uint32_t* maxabs, int sub,
                      int8_t* dw_out, int8_t* w_inout, niti_stream);
int niti_sgd_update(niti_ctx*, int8_t* w, const int8_t* dw, size_t n, niti_stream);
/* The weight-update tail of a whole training step in ONE launch: for every segment (= conv layer)
 *   *maxabs = max|acc| (atomicMax into a zeroed slot), b = ceil_log2, dW = b==0 ? 0 : PSTO(acc, b-sub),
 *   w <- clip127(w - dW), and, when wt != NULL, wt[c][tap][oc] <- the same updated weights (transposed copy).
 * Same arithmetic as niti_absmax + niti_rescale_grad(w_inout) per layer (NITI_GradientConv_Int8.cpp:274-296,
 * optimizer/NITI_SGD.hpp:20-54).  `segs` is a HOST array read at call time (captured by value: graph-safe). */
typedef struct {
    const int32_t* acc;              /* [oc][taps][c] int32 dW accumulator (already all-reduced when data-parallel) */
    int8_t* w;                       /* [oc][taps][c] in/out   */
    int8_t* wt;                      /* [c][taps][oc] out, optional */
    uint32_t* maxabs;                /* zeroed slot; holds max|acc| afterwards */
    int32_t oc, taps, c, sub;
} niti_update_seg;
int niti_update_batched(niti_ctx*, const niti_update_seg* segs, int nseg, niti_stream);
/* the same as two stream-ordered launches (max|acc| per segment, then rule B + SGD): holds no grid barrier, so it is exempt from
 * the one-barrier-call-at-a-time rule above and may run on a second stream beside the dX / dW kernels — the step updates the
 * deep layers this way while backward is still working on the shallow ones */
int niti_update_batched_phases(niti_ctx*, const niti_update_seg* segs, int nseg, niti_stream);
/* Requantisation of the DSP-named op types (OpType_NITI_DSP_*_Int8), which the reference runs as
 * hexagon_nn graphs: shift = ceiling_log2(max|acc|) - 7, out = sat[-128,127](round-half-up(acc / 2^shift)) / post_div,
 * exp_out = exp_a + exp_b + shift.  Replaces hexagon_nn/2.10.1/hexagon/ops/src/niti_op_supernode_new.c:4614-4634 +
 * src/quantize.c:337-400 and the u8 <-> int8 re-bias loops of NITI_DSPConv_Int8.cpp:375-452; post_div = 32 / 16 are the
 * gradient variants' divisions (NITI_DSPTransposeGradientConv_Int8.cpp:430-434, NITI_DSPMatmulGradientConv_Int8.cpp:548-550).
 * Parity of this rule is UNPINNED (the DSP path needs a Snapdragon cDSP); see oracle/niti_oracle.c niti_o_rescale_dsp. */
int niti_rescale_dsp(niti_ctx*, const int32_t* acc, size_t n, const uint32_t* maxabs, int post_div, int8_t* out,
                     const int8_t* exp_a, const int8_t* exp_b, int8_t* exp_out, niti_stream);
/* rule B per output-channel slice + post-alignment — NITI_GradientSplitConv_Int8.cpp:318-372 (sub = 3):
 *   acc holds `slices` equal contiguous slices of n_per_slice words; maxabs[s] = max|acc_s| (niti_absmax per slice);
 *   out_s = b_s==0 ? 0 : PSTO(acc_s, b_s-sub), then out_s /= 2^(max_s b - b_s) (C truncation). */
int niti_rescale_grad_split(niti_ctx*, const int32_t* acc, size_t n_per_slice, int slices,
                            const uint32_t* maxabs, int sub, int8_t* out, niti_stream);

/* ---- whole ops = phase 1 + phase 2 on the context's workspace -------------------------------
 * Drop-in for the four reference executions (same inputs/outputs, NHWC/OHWI instead of NC4HW4). */
int niti_conv_fwd(niti_ctx*, const niti_conv_desc*, const int8_t* x, const int8_t* w,
                  const int8_t* ex, const int8_t* ew, int8_t* y, int8_t* ey, niti_stream);
/* The forward op with everything of NITI_Conv_Int8::onExecute (+ the following NITI_CPURelu_Int8 when relu != 0) behind
 * one call, data-parallel aware: y = rule A(conv(x, w)) with the exponent taken from the max over all ranks.
 * Stem layers (c == 16: 1/3-channel inputs zero-padded to 16) run as ONE launch that recomputes the contraction instead
 * of round-tripping int32 accumulators through HBM (csrc/conv_stem.cu); other shapes run phase 1 + niti_rescale_act_dp on
 * `acc` (caller scratch of n*oh*ow*oc words, or NULL for the context workspace).  `maxabs`: zeroed slot, holds this
 * rank's max|acc| afterwards; dp_slot as in niti_rescale_act_dp. */
int niti_conv_fwd_fused(niti_ctx*, const niti_conv_desc*, const int8_t* x, const int8_t* w, const int8_t* ex,
                        const int8_t* ew, int8_t* y, int8_t* ey, int relu, int32_t* acc, uint32_t* maxabs, int dp_slot,
                        niti_stream);
/* The dX op of NITI_DeConv_Int8::onExecute (+ the NITI_CPUReluGrad_Int8 that follows when gate != NULL: dx = gate > 0 ? dx : 0)
 * behind one call, data-parallel aware.  When every output tile fits in tensor memory at once (csrc/conv_tma2.cu,
 * igemm_fused_kernel; also used by niti_conv_fwd_fused) the accumulators never leave the SM: contraction, max|acc|, grid
 * barrier (+ peer exchange), rescale and the int8 store are ONE launch.  Otherwise phase 1 + niti_rescale_act_dp on `acc`
 * (caller scratch of n*h*w*c words or NULL).  wt: transposed weights as in niti_conv_dx_acc_wt (NULL: per-call transpose). */
int niti_conv_dx_fused(niti_ctx*, const niti_conv_desc*, const int8_t* dy, const int8_t* w, const int8_t* wt, int8_t* dx,
                       const int8_t* gate, int32_t* acc, uint32_t* maxabs, int dp_slot, niti_stream);
/* niti_conv_fwd_fused with the 2x2 stride-2 max pool that follows the ReLU (NITI_Maxpool_Int8.cpp:24-70) folded into the store:
 * y [n,oh,ow,oc] is still written (backward needs it), y_pool [n,oh/2,ow/2,oc] is the pooled tensor.  One launch when the
 * fused kernel takes the shape and ow is 2, 4, 8 or 16 with oh even; otherwise niti_conv_fwd_fused + niti_maxpool. */
int niti_conv_fwd_fused_pool(niti_ctx*, const niti_conv_desc*, const int8_t* x, const int8_t* w, const int8_t* ex,
                             const int8_t* ew, int8_t* y, int8_t* ey, int relu, int8_t* y_pool, int32_t* acc, uint32_t* maxabs,
                             int dp_slot, niti_stream);
/* Stem layer: the input is an image stored with d->c channels of which only the first c_live can be non-zero (the caller's
 * guarantee — e.g. RGB padded to the 16-channel granule of the other kernels; w is [oc][kh][kw][d->c] as everywhere).  With
 * c_live <= 4 the operand is assembled from one 4-byte word per pixel and tap (K = taps x 4 instead of taps x 16) and, when
 * y_pool is given, the 2x2/2 max pool behind the conv is folded into the store (csrc/conv_stem2.cu).  Same results as
 * niti_conv_fwd_fused[_pool], which it calls for every other shape.  y_pool may be NULL. */
int niti_conv_stem_fwd(niti_ctx*, const niti_conv_desc*, int c_live, const int8_t* x, const int8_t* w, const int8_t* ex,
                       const int8_t* ew, int8_t* y, int8_t* ey, int relu, int8_t* y_pool, int32_t* acc, uint32_t* maxabs,
                       int dp_slot, niti_stream);
/* niti_conv_dx_fused whose result is the gradient w.r.t. a 2x2/2 max-POOLED tensor, with the pool's gradient
 * (NITI_CPUPoolGrad_Int8.cpp:36-72: the first window element equal to the maximum takes it) and, when gate_relu != 0, the
 * ReLU gradient of the layer before the pool folded into the store: pool_x [n,2h,2w,c] is the pre-pool forward tensor,
 * dx_prepool [n,2h,2w,c] receives the gradient w.r.t. it.  dx_pooled [n,h,w,c] is scratch for the unfused route (shapes the
 * fused kernel does not take) and may be NULL when the caller knows the fused kernel applies. */
int niti_conv_dx_fused_poolgrad(niti_ctx*, const niti_conv_desc*, const int8_t* dy, const int8_t* w, const int8_t* wt,
                                int8_t* dx_prepool, const int8_t* pool_x, int gate_relu, int8_t* dx_pooled, int32_t* acc,
                                uint32_t* maxabs, int dp_slot, niti_stream);

/* ---- chains: several fused ops behind ONE persistent launch ----------------------------------------------------------
 * The reference runs a training step as a serial loop over ops (express/Executor.cpp:333-359).  A chain is a list of
 * forward (or dX) ops of which each consumes what the previous one stored; niti_chain_run executes them in one launch that
 * keeps tensor memory, barriers and CTAs alive across layers (csrc/conv_chain.cu).  Results are identical to calling the
 * ops one by one.  Every op must be within the fused kernel's contract (niti_chain_create returns NITI_ERR_UNSUPPORTED
 * otherwise and the caller issues the ops individually).  Pointers are captured at creation. */
typedef struct {
    niti_conv_desc d;
    int32_t op;                      /* 0 forward (NITI_Conv_Int8), 1 dX (NITI_DeConv_Int8)                         */
    const int8_t* in;                /* forward: x [n,h,w,c]; dX: dy [n,oh,ow,oc]                                    */
    const int8_t* w;                 /* forward: w [oc,kh,kw,c]; dX: the transposed weights wt [c,kh,kw,oc]           */
    const int8_t* ex; const int8_t* ew; int8_t* ey;   /* forward: exponents in / out (dX: NULL)                      */
    int8_t* out;                     /* forward: y; dX: dx (with pool_x: the gradient w.r.t. the pre-pool tensor)     */
    int32_t relu;                    /* forward: fold NITI_CPURelu_Int8 into the store                               */
    const int8_t* gate;              /* dX: fold NITI_CPUReluGrad_Int8 (tensor of out's shape), or NULL               */
    int8_t* out_pool;                /* forward: also write the 2x2/2 max-pooled tensor, or NULL                      */
    const int8_t* pool_x; int32_t pool_gate;   /* dX: fold the max-pool gradient (+ ReLU gate) as in niti_conv_dx_fused_poolgrad */
    uint32_t* maxabs; int32_t dp_slot;         /* zeroed slot -> this rank's max|acc|; exchange slot (niti_rescale_act_dp) */
} niti_chain_op;
typedef struct niti_chain niti_chain;
int niti_chain_op_eligible(niti_ctx*, const niti_chain_op* op);      /* 1 when the fused kernel takes this op (shape + pooling geometry) */
int niti_chain_create(niti_ctx*, const niti_chain_op* ops, int n, niti_chain** out);
int niti_chain_run(niti_ctx*, niti_chain*, niti_stream);
int niti_chain_destroy(niti_ctx*, niti_chain*);
int niti_conv_dx(niti_ctx*, const niti_conv_desc*, const int8_t* dy, const int8_t* w, int8_t* dx, niti_stream);
int niti_conv_dw(niti_ctx*, const niti_conv_desc*, const int8_t* x, const int8_t* dy, int8_t* dw, niti_stream);
int niti_matmul_nt(niti_ctx*, int e, int h, int l, const int8_t* a, const int8_t* b, int8_t* c, niti_stream);
/* NITI_GradientSplitConv_Int8::onExecute (NITI_GradientSplitConv_Int8.cpp:195-386) as one call: dW contraction, then the
 * output channels cut into `slices` groups, each rescaled with its own exponent (PSTO(acc_s, b_s - 3)) and aligned to the
 * largest one (niti_rescale_grad_split).  niti_split_count restates the reference's slice rule (.cpp:209-222; host side,
 * oc = output channels, oh/ow = spatial size of dY); oc % slices == 0 and slices <= 32. */
int niti_split_count(int oc, int oh, int ow);
int niti_conv_dw_split(niti_ctx*, const niti_conv_desc*, const int8_t* x, const int8_t* dy, int slices, int8_t* dw, niti_stream);

/* ---- element-wise ops of the training step, NHWC ---------------------- */
int niti_relu(niti_ctx*, const int8_t* x, int8_t* y, size_t n, niti_stream);                 /* NITI_CPURelu_Int8.cpp:47-53 */
int niti_relu_grad(niti_ctx*, const int8_t* x, const int8_t* dy, int8_t* dx, size_t n, niti_stream); /* NITI_CPUReluGrad_Int8.cpp:51-57 */
int niti_maxpool(niti_ctx*, const int8_t* x, int n, int h, int w, int c, int k, int s, int8_t* y, niti_stream); /* NITI_Maxpool_Int8.cpp:24-70 */
int niti_maxpool_grad(niti_ctx*, const int8_t* x, const int8_t* y, const int8_t* dy, int n, int h, int w, int c,
                      int k, int s, int8_t* dx, niti_stream);                                 /* NITI_CPUPoolGrad_Int8.cpp:36-72 */
/* same as niti_maxpool_grad, with NITI_CPUReluGrad_Int8 folded in when gate_relu != 0: dx = x > 0 ? dx : 0 */
int niti_maxpool_grad_gated(niti_ctx*, const int8_t* x, const int8_t* y, const int8_t* dy, int n, int h, int w, int c,
                            int k, int s, int8_t* dx, int gate_relu, niti_stream);
int niti_loss_grad(niti_ctx*, const int8_t* logits, const int8_t* ascale, const int32_t* labels,
                   int batch, int classes, int8_t* grad, niti_stream);                        /* NITI_CPULossGrad_Int8.cpp:82-201 */
/* same, for logit / gradient rows of `stride` bytes of which the first `classes` are real (output channels zero-padded
 * to the tensor cores' granule); the padding columns of grad are written as 0 */
int niti_loss_grad_strided(niti_ctx*, const int8_t* logits, const int8_t* ascale, const int32_t* labels,
                           int batch, int classes, int stride, int8_t* grad, niti_stream);
/* the float loss scalar of NITI_CPULoss_Int8::onExecute (source/backend/cpu/NITI_CPULoss_Int8.cpp:68-131): -mean over the batch of
 * log softmax(2^ascale * logits)[label], with the reference's polynomial exp (compute/CommonOptFunction.cpp:1431-1456) and its
 * summation order.  Monitoring only (floating point: matches the reference to 1e-5 relative); rows `stride` bytes apart,
 * row_scratch: `batch` floats of device scratch, loss_out: one device float. */
int niti_loss_value(niti_ctx*, const int8_t* logits, const int8_t* ascale, const int32_t* labels, int batch, int classes,
                    int stride, float* row_scratch, float* loss_out, niti_stream);

/* one-hot int32 targets [batch][classes] (the form NITI_CPULossGrad_Int8 receives, :148,170-178) -> class indices */
int niti_onehot_labels(niti_ctx*, const int32_t* onehot, int batch, int classes, int32_t* labels, niti_stream);
/* int8 Binary op with the reference's clip to +-127 (CPUBinary.cpp:246-330, BinaryUtils.hpp:276-299): op 0 ADD, 1 SUB, 2 MUL;
 * broadcast -1: element-wise over n, 0: a is one element, 1: b is one element */
int niti_binary_i8(niti_ctx*, int op, const int8_t* a, const int8_t* b, size_t n, int broadcast, int8_t* out, niti_stream);
/* NITI_CPULeftPoolGrad_Int8.cpp:19-52 (zero insertion in front of a strided conv's dX): tensors as [slabs][h][w][4] blocks
 * (NC4HW4, slabs = batch * ceil(c/4)); out(y, x) = in(y / sy, x / sx) where sy | y and sx | x, else 0 */
int niti_zero_insert_c4(niti_ctx*, const int8_t* in, int slabs, int ih, int iw, int oh, int ow, int sy, int sx, int8_t* out, niti_stream);
/* NITI_Pad_Int8.cpp:26-66: planes = batch * channels of [h][w] (NCHW); each plane lands at (pad, pad) of a zeroed [oh][ow] */
int niti_pad2d(niti_ctx*, const int8_t* in, int planes, int ih, int iw, int pad, int oh, int ow, int8_t* out, niti_stream);

/* ---- layout conversion at the boundary ----------------------------------------------------- */
int niti_nchw_to_nhwc(niti_ctx*, const int8_t* src, int n, int c, int h, int w, int8_t* dst, niti_stream);
int niti_nhwc_to_nchw(niti_ctx*, const int8_t* src, int n, int c, int h, int w, int8_t* dst, niti_stream);
/* the reference CPU path's int8 "NC4HW4" is [c/4][n][h*w][4] (NITI_Conv_Int8.cpp:111,191,229-230) */
int niti_nc4hw4_to_nhwc(niti_ctx*, const int8_t* src, int n, int c, int h, int w, int8_t* dst, niti_stream);
int niti_nhwc_to_nc4hw4(niti_ctx*, const int8_t* src, int n, int c, int h, int w, int8_t* dst, niti_stream);

/* generic strided gather (layout conversion for the MNN plug-in: NC4HW4 / NCHW / transposed / rot180
 * operands as the reference's grad rule hands them over, tools/train/source/grad/NITI_Conv_Int8_Grad.cpp:84-191):
 *   dst is contiguous [dims[0]]..[dims[4]];  dst[i0..i4] = src[base + sum_k i_k * strides[k]]
 *   when split_hi >= 0 the pair (split_hi, split_lo) is one logical axis cut in blocks of dims[split_lo]
 *   (the 4-lane channel block of NC4HW4): positions with i_hi*dims[split_lo] + i_lo >= split_limit read as 0. */
int niti_gather_i8(niti_ctx*, const int8_t* src, int64_t base, int8_t* dst, const int32_t dims[5],
                   const int64_t strides[5], int split_hi, int split_lo, int split_limit, niti_stream);

/* ---- device memory for hosts that own no CUDA runtime (the MNN plug-in links no CUDA library) -- */
int niti_dev_alloc(niti_ctx*, size_t bytes, void** out);
int niti_dev_free(niti_ctx*, void* p);
int niti_dev_memset(niti_ctx*, void* p, int value, size_t bytes, niti_stream);
int niti_copy_h2d(niti_ctx*, void* dst_dev, const void* src_host, size_t bytes, niti_stream);
int niti_copy_d2h(niti_ctx*, void* dst_host, const void* src_dev, size_t bytes, niti_stream);
int niti_copy_d2d(niti_ctx*, void* dst_dev, const void* src_dev, size_t bytes, niti_stream);
int niti_stream_sync(niti_ctx*, niti_stream);

/* ---- for callers that may not allocate device memory ---------------------------------------
 * No counterpart in the reference (its NITI_Matmul_Int8 scratch comes from the MNN backend's buffer pool); this is an
 * addition of this library.  Same result as niti_matmul_nt — C[e,h] = rule B, sub 3, of sum_l A[e,l] * B[h,l], bit for bit —
 * but without a niti_ctx and without any cudaMalloc / cudaFree, now or on an earlier call.  It exists for hosts that get a
 * fixed device-memory budget at start-up and cannot ask the driver for more afterwards (a runtime embedded in someone
 * else's memory plan, an allocator that is sealed after set-up), and for call sites where an allocation is not allowed or
 * would stall the whole device (latency-critical sections).  The int32 accumulators and the max|acc| word, which the
 * other whole-op calls keep in a context workspace, live in an array compiled into the library's device image instead:
 * NITI_STATIC_SCRATCH_MAX_WORDS accumulators + a 256-byte header = 4 MiB + 256 B on EVERY GPU that loads the library,
 * held until the process exits whether or not this call is ever made.
 *   - e * h must not exceed NITI_STATIC_SCRATCH_MAX_WORDS: larger products return NITI_ERR_UNSUPPORTED with a message
 *     (split A by rows, or use niti_matmul_nt).  l is unlimited.  a, b, c: device memory of the CURRENT device, any
 *     alignment (4-byte aligned a, b with l % 4 == 0 take the faster load path).
 *   - The first call on each device looks up that device's address of the array (cudaGetSymbolAddress, which may also
 *     load the module).  Make that first call during set-up, with the device current and before any stream capture begins;
 *     later calls only launch kernels on `st`.
 *   - All calls on one device share the one array.  Each call resets the max|acc| word itself and overwrites the
 *     accumulators it reads, so back-to-back calls on ONE stream are safe; two calls that can overlap on the GPU (different
 *     streams without an event between them, or two host threads) race on it and give wrong results.
 *   - err (optional, err_len bytes) receives a NUL-terminated message when the return value is not NITI_OK. */
#define NITI_STATIC_SCRATCH_MAX_WORDS (1u << 20)
int niti_matmul_nt_static(int e, int h, int l, const int8_t* a, const int8_t* b, int8_t* c, niti_stream st,
                          char* err, size_t err_len);

/* ---- host-buffer convenience (the e2e leg): H2D, op, D2H inside the call ------------------- */
int niti_conv_fwd_host(niti_ctx*, const niti_conv_desc*, const int8_t* x_nhwc, const int8_t* w_ohwi,
                       int8_t ex, int8_t ew, int8_t* y_nhwc, int8_t* ey);

/* ---- probes ------------------------------------------------------------------------------- */
/* phase trace for kernel development: when device_buf != NULL the fused kernels write %globaltimer (ns) stamps to
 * device_buf[block * 8 + phase] (phases: 0 entry, 1 set-up done, 2 first tile accumulated, 3 max (+ predicted store) sweep done,
 * 4 grid barrier passed, 5 re-store done (prediction missed; else = 4), 6 next layer released, 7 exit); device_buf must hold 8 words per CTA of the largest grid.  NULL disables. */
int niti_debug_trace(niti_ctx*, unsigned long long* device_buf);
/* launch timeline for kernel development: while device_buf != NULL every instrumented launch issued through this context
 * takes the next record {first CTA entry, last CTA exit} (%globaltimer ns; the caller presets records to {~0, 0}) — records are
 * assigned at launch time, so a captured CUDA graph keeps writing the same records on every replay.  capacity in records. */
int niti_debug_timeline(niti_ctx*, unsigned long long* device_buf, int capacity);
int niti_debug_timeline_count(niti_ctx*);
const char* niti_debug_timeline_name(niti_ctx*, int i);
/* dense tcgen05 kind::i8 issue-rate probe (no global traffic): returns achieved TOPS */
int niti_probe_i8_peak(niti_ctx*, int iters, double* tops_out);

#ifdef __cplusplus
}
#endif
#endif
