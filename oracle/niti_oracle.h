/*
 * niti_oracle.h — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C) of the reference's NITI int8 training hot path, used as the parity
 * checker for the CUDA path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may use it; the product never links, imports or calls anything here.
 *
 * Parity status: PINNED.  The restatement is checked (tests/test_oracle_vs_ref.py, run where
 * /root/reference exists) against the unmodified reference compiled by oracle/Makefile.ref into
 * oracle/_ref/libMNN_ref.so and driven op-by-op by oracle/ref_driver.cpp; the vectors that check
 * produced are committed under tests/golden/ so the pin also holds on boxes without the reference.
 * (The reference ships no golden vectors or tests of its own for this path)
 *
 * Logical layouts here are the mathematical ones (NCHW activations, OIHW weights); every output of
 * the path is layout-independent up to a permutation because the rescale is element-wise and the
 * exponent is a max over the whole tensor.
 *
 * Reference files restated (under /root/reference/execution-engine/):
 *   source/backend/cpu/compute/CommonOptFunction.cpp:1548-1627  clip / sign / RangeEstimate / PSTO
 *   source/backend/cpu/NITI_Conv_Int8.cpp:162-310               forward conv + 3-way shift rule
 *   source/backend/cpu/NITI_DeConv_Int8.cpp:179-332             dX (rot180 + conv + same rule)
 *   source/backend/cpu/NITI_GradientConv_Int8.cpp:165-298       dW (b-2 rule)
 *   source/backend/cpu/NITI_Matmul_Int8.cpp:140-231             matmul (b-3 rule)
 *   source/backend/cpu/BinaryUtils.hpp:276-299 + optimizer/NITI_SGD.hpp:20-54   int8 SGD update
 *   source/backend/cpu/NITI_CPURelu_Int8.cpp / NITI_CPUReluGrad_Int8.cpp / NITI_Maxpool_Int8.cpp /
 *   NITI_CPUPoolGrad_Int8.cpp / NITI_CPULossGrad_Int8.cpp       the element-wise ops of the step
 */
#ifndef NITI_ORACLE_H
#define NITI_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int n, c, h, w;        /* input  [n,c,h,w]      */
    int oc, kh, kw;        /* weight [oc,c,kh,kw]   */
    int stride, pad, dil;  /* symmetric             */
} niti_conv_shape;

int  niti_o_out_h(const niti_conv_shape* s);
int  niti_o_out_w(const niti_conv_shape* s);

/* rescale primitives */
int32_t niti_o_clip(int32_t a);
int32_t niti_o_sign(int32_t a);
int     niti_o_range_estimate(const int32_t* acc, long n);   /* ceil(log2(max|acc|)), 0 if max==0 */
int     niti_o_bitwidth_of_max(int32_t maxabs);
void    niti_o_psto_shift(const int32_t* in, int shift, int32_t* out, long n);

/* rule A (forward / dX): shift=b-7; >1: PSTO(shift); ==1: PSTO(2); else raw int8 cast.
 * returns the exponent increment (shift, 2 or 0). */
int  niti_o_rescale_act(const int32_t* acc, long n, int8_t* out);
/* rule B (dW): b==0 -> zeros, else PSTO(b - sub) with sub = 2 (GradientConv) or 3 (Matmul). */
void niti_o_rescale_grad(const int32_t* acc, long n, int sub, int8_t* out);

/* NITI_GradientSplitConv_Int8.cpp:209-222 (slice count) and :318-362 (per-slice rule B with sub = 3 + post-alignment) */
int  niti_o_split_count(int oc, int oh, int ow);
void niti_o_rescale_grad_split(const int32_t* acc, long n_per_slice, int slices, int8_t* out);

/* DSP-named op types: round-half-up requantisation of the Hexagon path (niti_op_supernode_new.c:4614-4634,
 * quantize.c:337-400); returns the shift (exponent increment).  post_div: 1, 16 or 32.  PARITY UNPINNED (no cDSP). */
int  niti_o_dsp_ceiling_log2(uint32_t v);
int  niti_o_rescale_dsp(const int32_t* acc, long n, int post_div, int8_t* out);

/* int32 accumulators.  Returns the largest |partial sum| seen (Q1 check: must stay < 2^24 for the
 * reference's float accumulation to be exact). */
int64_t niti_o_conv_fwd_acc(const niti_conv_shape* s, const int8_t* x, const int8_t* w, int32_t* acc);
/* dX: dy [n,oc,oh,ow], w [oc,c,kh,kw] -> acc [n,c,h,w]  (stride-1 and strided via zero insertion) */
int64_t niti_o_conv_dx_acc(const niti_conv_shape* s, const int8_t* dy, const int8_t* w, int32_t* acc);
/* dW: x [n,c,h,w], dy [n,oc,oh,ow] -> acc [oc,c,kh,kw] */
int64_t niti_o_conv_dw_acc(const niti_conv_shape* s, const int8_t* x, const int8_t* dy, int32_t* acc);
/* C[e,h] = sum_l A[e,l] * B[h,l] */
int64_t niti_o_matmul_nt_acc(int e, int h, int l, const int8_t* a, const int8_t* b, int32_t* c);

/* whole ops */
int  niti_o_conv_fwd(const niti_conv_shape* s, const int8_t* x, const int8_t* w, int ex, int ew,
                     int8_t* y, int32_t* acc_opt);                 /* returns exp_out */
void niti_o_conv_dx(const niti_conv_shape* s, const int8_t* dy, const int8_t* w, int8_t* dx, int32_t* acc_opt);
void niti_o_conv_dw(const niti_conv_shape* s, const int8_t* x, const int8_t* dy, int8_t* dw, int32_t* acc_opt);
void niti_o_matmul_nt(int e, int h, int l, const int8_t* a, const int8_t* b, int8_t* c, int32_t* acc_opt);

void niti_o_sgd_update(int8_t* w, const int8_t* dw, long n);       /* w = clip127(w - dw) */

/* element-wise ops of the step (NCHW) */
void niti_o_relu(const int8_t* x, int8_t* y, long n);
void niti_o_relu_grad(const int8_t* x, const int8_t* dy, int8_t* dx, long n);
void niti_o_maxpool(const int8_t* x, int n, int c, int h, int w, int k, int s, int8_t* y);
void niti_o_maxpool_grad(const int8_t* x, const int8_t* y, const int8_t* dy, int n, int c, int h, int w,
                         int k, int s, int8_t* dx);
void niti_o_loss_grad(const int8_t* logits, int ascale, const int32_t* labels, int batch, int classes,
                      int8_t* grad);
/* the float loss scalar the demo prints (NITI_CPULoss_Int8.cpp:68-131): -mean_i log softmax(2^ascale * logits_i)[label_i] with
 * MNN's polynomial exp (compute/CommonOptFunction.cpp:1431-1456, the build without MNN_USE_SSE) */
float niti_o_loss_value(const int8_t* logits, int ascale, const int32_t* labels, int batch, int classes);

#ifdef __cplusplus
}
#endif
#endif
