/*
 * niti_oracle.c — TEST INFRASTRUCTURE ONLY (see niti_oracle.h for the header comment, scope rules
 * and the list of reference files restated).  Plain C, no dependencies beyond libc/libm.
 */
#include "niti_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

int niti_o_out_h(const niti_conv_shape* s) {
    return (s->h + 2 * s->pad - (s->dil * (s->kh - 1) + 1)) / s->stride + 1;
}
int niti_o_out_w(const niti_conv_shape* s) {
    return (s->w + 2 * s->pad - (s->dil * (s->kw - 1) + 1)) / s->stride + 1;
}

/* compute/CommonOptFunction.cpp:1548-1555 — symmetric clip, never -128 (quirk Q4) */
int32_t niti_o_clip(int32_t a) { return a > 127 ? 127 : (a < -127 ? -127 : a); }
/* compute/CommonOptFunction.cpp:1556-1563 */
int32_t niti_o_sign(int32_t a) { return a > 0 ? 1 : (a < 0 ? -1 : 0); }

/* ceil(log2(m)) for m>0 computed exactly in integers; the reference uses ceil(log2((double)m))
 * (CommonOptFunction.cpp:1575) which is exact for m < 2^31 because log2 of a power of two is exact
 * and any other m lies strictly between two powers. */
int niti_o_bitwidth_of_max(int32_t m) {
    if (m <= 0) return 0;
    int b = 0;
    while (((int64_t)1 << b) < (int64_t)m) ++b;
    return b;
}

/* compute/CommonOptFunction.cpp:1565-1576 */
int niti_o_range_estimate(const int32_t* acc, long n) {
    int32_t m = 0;
    for (long i = 0; i < n; ++i) {
        int32_t a = abs(acc[i]);
        if (m < a) m = a;
    }
    return niti_o_bitwidth_of_max(m);
}

/* `1 << s` as the reference's x86-64 build evaluates it: 32-bit shl masks the count to 5 bits, so a
 * negative count (dW/matmul with max|acc| <= 4) wraps instead of trapping.  This is undefined
 * behaviour in C; the oracle fixes the x86 meaning so the edge case has one defined answer. */
static int32_t shl1(int s) { return (int32_t)(1u << ((unsigned)s & 31u)); }

/* compute/CommonOptFunction.cpp:1595-1627 — pseudo-stochastic shift-round (deterministic, Q5) */
void niti_o_psto_shift(const int32_t* in, int shift, int32_t* out, long n) {
    const int32_t full = shl1(shift);
    const int32_t half = shl1(shift / 2);           /* C division: truncates toward zero */
    const int odd = (shift % 2 == 1);               /* -1 % 2 == -1 -> even branch        */
    for (long i = 0; i < n; ++i) {
        int32_t x = in[i];
        int32_t q = x / full;
        int32_t prob = abs(x - q * full);
        int32_t hi = prob / half;
        int32_t lo = prob - hi * half;
        if (odd) lo *= 2;
        int32_t r = (hi > lo);
        out[i] = niti_o_clip(q + r * niti_o_sign(x));
    }
}

/* NITI_Conv_Int8.cpp:260-305, NITI_DeConv_Int8.cpp:292-329 */
int niti_o_rescale_act(const int32_t* acc, long n, int8_t* out) {
    int b = niti_o_range_estimate(acc, n);
    int shift = b - 7;
    if (shift > 1 || shift == 1) {
        int s = shift > 1 ? shift : 2;              /* Q3: shift==1 shifts by 2 */
        int32_t* tmp = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
        niti_o_psto_shift(acc, s, tmp, n);
        for (long i = 0; i < n; ++i) out[i] = (int8_t)tmp[i];
        free(tmp);
        return s;
    }
    for (long i = 0; i < n; ++i) out[i] = (int8_t)acc[i];   /* Q3: raw cast, no clip */
    return 0;
}

/* NITI_GradientConv_Int8.cpp:274-296 (sub=2), NITI_Matmul_Int8.cpp:216-228 (sub=3) */
void niti_o_rescale_grad(const int32_t* acc, long n, int sub, int8_t* out) {
    int b = niti_o_range_estimate(acc, n);
    if (b == 0) { memset(out, 0, (size_t)n); return; }      /* Q2 */
    int32_t* tmp = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    niti_o_psto_shift(acc, b - sub, tmp, n);
    for (long i = 0; i < n; ++i) out[i] = (int8_t)tmp[i];
    free(tmp);
}

/* ---- NITI_GradientSplitConv_Int8 ----------------------------------------------------------
 * Slice count, NITI_GradientSplitConv_Int8.cpp:209-222: `oc` = weight->batch() (output channels of dyT [OC,N,OH,OW]),
 * oh/ow = weight->height()/width().  Note the reference only rounds OC up to a multiple of 32 when OC % 32 != 0; with
 * OC % 32 == 0 its `align32` is 0, the loop never runs and the op degenerates to one slice. */
int niti_o_split_count(int oc, int oh, int ow) {
    int align32 = oc % 32;
    if (align32 != 0) align32 = (oc / 32 + 1) * 32;
    int high = 8192 / ow / oh;
    int real = oc;
    while (align32 > high) { align32 /= 2; real /= 2; }
    return real > 0 ? oc / real : 0;       /* real == 0 divides by zero in the reference */
}
/* Per-slice rule + alignment, NITI_GradientSplitConv_Int8.cpp:318-362: acc holds `slices` equal contiguous slices (one per
 * group of output channels); scale[s] = RangeEstimate(acc_s); out_s = scale[s]==0 ? 0 : (int8)PSTO(acc_s, scale[s]-3);
 * then every slice with scale[s] < max scale is integer-divided (C truncation, int8 operand promoted to int) by
 * 1 << (max - scale[s]). */
void niti_o_rescale_grad_split(const int32_t* acc, long n_per_slice, int slices, int8_t* out) {
    int* scale = (int*)malloc(sizeof(int) * (size_t)(slices > 0 ? slices : 1));
    int32_t* tmp = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_per_slice > 0 ? n_per_slice : 1));
    for (int s = 0; s < slices; ++s) {
        const int32_t* a = acc + (long)s * n_per_slice;
        int8_t* o = out + (long)s * n_per_slice;
        scale[s] = niti_o_range_estimate(a, n_per_slice);
        if (scale[s] == 0) { memset(o, 0, (size_t)n_per_slice); continue; }
        niti_o_psto_shift(a, scale[s] - 3, tmp, n_per_slice);
        for (long i = 0; i < n_per_slice; ++i) o[i] = (int8_t)tmp[i];
    }
    int mx = scale[0];
    for (int s = 1; s < slices; ++s) if (mx < scale[s]) mx = scale[s];
    for (int s = 0; s < slices; ++s) {
        if (mx > scale[s]) {
            int8_t* o = out + (long)s * n_per_slice;
            const int32_t div = shl1(mx - scale[s]);
            for (long i = 0; i < n_per_slice; ++i) o[i] = (int8_t)((int32_t)o[i] / div);
        }
    }
    free(tmp); free(scale);
}

/* ---- DSP-named op types: requantisation of the Hexagon path.  PARITY UNPINNED: the code
 * below restates the reference's sources, but the reference's DSP path only runs on a Snapdragon cDSP, so no output of
 * it could be produced here to check the restatement against.
 *   hexagon_nn/2.10.1/hexagon/include/quantize.h:260-265          ceiling_log2
 *   hexagon_nn/2.10.1/hexagon/ops/src/niti_op_supernode_new.c:4614-4634
 *        m = max|acc|; shift = ceiling_log2(m) - 7; requantize(acc, in_level_size = 1 / (1 << shift), -128, 127)
 *   hexagon_nn/2.10.1/hexagon/src/quantize.c:337-400              gain < 1: sat_u8(asrrnd(x * 2^(31-shift), 31) + 128)
 *        = round-half-up of x / 2^shift; gain == 1: sat_u8((int)(x + 128.5))
 *   execution-engine/source/backend/cpu/NITI_DSPConv_Int8.cpp:447-452     int8 = u8 - 128, exponent += shift
 *   NITI_DSPTransposeGradientConv_Int8.cpp:430-434 (/32), NITI_DSPMatmulGradientConv_Int8.cpp:548-550 (/16):
 *        the gradient variants divide the int8 result (C truncation).
 * shift < 0 (max|acc| <= 64) makes `1 << shift` undefined in the reference; the restatement (and the kernel) keep the
 * accumulators as they are there (shift 0), which is what a zero shift count gives. */
int niti_o_dsp_ceiling_log2(uint32_t v) {
    if (v < 4) return (int)v - 1;
    return 32 - __builtin_clz(v - 1);
}

int niti_o_rescale_dsp(const int32_t* acc, long n, int post_div, int8_t* out) {
    uint32_t m = 0;
    for (long i = 0; i < n; ++i) {
        uint32_t a = acc[i] < 0 ? (uint32_t)(-(int64_t)acc[i]) : (uint32_t)acc[i];
        if (a > m) m = a;
    }
    int shift = niti_o_dsp_ceiling_log2(m) - 7;
    if (shift < 0) shift = 0;
    for (long i = 0; i < n; ++i) {
        int64_t q = shift > 0 ? (((int64_t)acc[i] + ((int64_t)1 << (shift - 1))) >> shift) : (int64_t)acc[i];
        if (q > 127) q = 127;
        if (q < -128) q = -128;
        int v = (int)q;
        if (post_div > 1) v = v / post_div;
        out[i] = (int8_t)v;
    }
    return shift;
}

#define TRACK(acc_) do { int64_t a_ = (acc_) < 0 ? -(acc_) : (acc_); if (a_ > worst) worst = a_; } while (0)

/* NITI_Conv_Int8.cpp:218-249 (im2col + 16x4 GEMM unit == this direct correlation) */
int64_t niti_o_conv_fwd_acc(const niti_conv_shape* s, const int8_t* x, const int8_t* w, int32_t* acc) {
    const int oh = niti_o_out_h(s), ow = niti_o_out_w(s);
    int64_t worst = 0;
    for (int n = 0; n < s->n; ++n)
    for (int o = 0; o < s->oc; ++o)
    for (int y = 0; y < oh; ++y)
    for (int xx = 0; xx < ow; ++xx) {
        int64_t a = 0;
        for (int c = 0; c < s->c; ++c)
        for (int i = 0; i < s->kh; ++i) {
            int iy = y * s->stride - s->pad + i * s->dil;
            if (iy < 0 || iy >= s->h) continue;
            for (int j = 0; j < s->kw; ++j) {
                int ix = xx * s->stride - s->pad + j * s->dil;
                if (ix < 0 || ix >= s->w) continue;
                a += (int32_t)x[((long)(n * s->c + c) * s->h + iy) * s->w + ix] *
                     (int32_t)w[((long)(o * s->c + c) * s->kh + i) * s->kw + j];
                TRACK(a);
            }
        }
        acc[((long)(n * s->oc + o) * oh + y) * ow + xx] = (int32_t)a;
    }
    return worst;
}

/* grad/NITI_Conv_Int8_Grad.cpp:27-132 builds dX = DeConv(pad(dilate(dY)), W^T) and
 * NITI_DeConv_Int8.cpp:179-290 rotates each plane 180 degrees and correlates: together that is the
 * transposed convolution below (scatter form; integer adds commute so the order is immaterial). */
int64_t niti_o_conv_dx_acc(const niti_conv_shape* s, const int8_t* dy, const int8_t* w, int32_t* acc) {
    const int oh = niti_o_out_h(s), ow = niti_o_out_w(s);
    const long nx = (long)s->n * s->c * s->h * s->w;
    int64_t* a64 = (int64_t*)calloc((size_t)(nx > 0 ? nx : 1), sizeof(int64_t));
    int64_t worst = 0;
    for (int n = 0; n < s->n; ++n)
    for (int o = 0; o < s->oc; ++o)
    for (int y = 0; y < oh; ++y)
    for (int xx = 0; xx < ow; ++xx) {
        int32_t g = dy[((long)(n * s->oc + o) * oh + y) * ow + xx];
        if (g == 0) continue;
        for (int c = 0; c < s->c; ++c)
        for (int i = 0; i < s->kh; ++i) {
            int iy = y * s->stride - s->pad + i * s->dil;
            if (iy < 0 || iy >= s->h) continue;
            for (int j = 0; j < s->kw; ++j) {
                int ix = xx * s->stride - s->pad + j * s->dil;
                if (ix < 0 || ix >= s->w) continue;
                a64[((long)(n * s->c + c) * s->h + iy) * s->w + ix] +=
                    g * (int32_t)w[((long)(o * s->c + c) * s->kh + i) * s->kw + j];
            }
        }
    }
    for (long i = 0; i < nx; ++i) { TRACK(a64[i]); acc[i] = (int32_t)a64[i]; }
    free(a64);
    return worst;   /* final magnitudes only: the summation order differs from the reference's */
}

/* grad/NITI_Conv_Int8_Grad.cpp:134-191 + NITI_GradientConv_Int8.cpp:218-268; the caller-side
 * transpose to [oc,c,kh,kw] (Grad.cpp:186-189) is folded in. */
int64_t niti_o_conv_dw_acc(const niti_conv_shape* s, const int8_t* x, const int8_t* dy, int32_t* acc) {
    const int oh = niti_o_out_h(s), ow = niti_o_out_w(s);
    int64_t worst = 0;
    for (int o = 0; o < s->oc; ++o)
    for (int c = 0; c < s->c; ++c)
    for (int i = 0; i < s->kh; ++i)
    for (int j = 0; j < s->kw; ++j) {
        int64_t a = 0;
        for (int n = 0; n < s->n; ++n)
        for (int y = 0; y < oh; ++y) {
            int iy = y * s->stride - s->pad + i * s->dil;
            if (iy < 0 || iy >= s->h) continue;
            for (int xx = 0; xx < ow; ++xx) {
                int ix = xx * s->stride - s->pad + j * s->dil;
                if (ix < 0 || ix >= s->w) continue;
                a += (int32_t)x[((long)(n * s->c + c) * s->h + iy) * s->w + ix] *
                     (int32_t)dy[((long)(n * s->oc + o) * oh + y) * ow + xx];
                TRACK(a);
            }
        }
        acc[((long)(o * s->c + c) * s->kh + i) * s->kw + j] = (int32_t)a;
    }
    return worst;
}

/* NITI_Matmul_Int8.cpp:186-208 */
int64_t niti_o_matmul_nt_acc(int e, int h, int l, const int8_t* a, const int8_t* b, int32_t* c) {
    int64_t worst = 0;
    for (int i = 0; i < e; ++i)
    for (int j = 0; j < h; ++j) {
        int64_t s = 0;
        for (int k = 0; k < l; ++k) { s += (int32_t)a[(long)i * l + k] * (int32_t)b[(long)j * l + k]; TRACK(s); }
        c[(long)i * h + j] = (int32_t)s;
    }
    return worst;
}

/* ---------------------------------------------------------------- whole ops */

static int32_t* acc_buf(int32_t* opt, long n) {
    return opt ? opt : (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
}

/* NITI_Conv_Int8.cpp:255-307: exp_out = ex + ew + {shift | 2 | 0}, stored as int8 (Q7) */
int niti_o_conv_fwd(const niti_conv_shape* s, const int8_t* x, const int8_t* w, int ex, int ew,
                    int8_t* y, int32_t* acc_opt) {
    long n = (long)s->n * s->oc * niti_o_out_h(s) * niti_o_out_w(s);
    int32_t* acc = acc_buf(acc_opt, n);
    niti_o_conv_fwd_acc(s, x, w, acc);
    int inc = niti_o_rescale_act(acc, n, y);
    if (!acc_opt) free(acc);
    return (int)(int8_t)(ex + ew + inc);
}

void niti_o_conv_dx(const niti_conv_shape* s, const int8_t* dy, const int8_t* w, int8_t* dx, int32_t* acc_opt) {
    long n = (long)s->n * s->c * s->h * s->w;
    int32_t* acc = acc_buf(acc_opt, n);
    niti_o_conv_dx_acc(s, dy, w, acc);
    niti_o_rescale_act(acc, n, dx);
    if (!acc_opt) free(acc);
}

void niti_o_conv_dw(const niti_conv_shape* s, const int8_t* x, const int8_t* dy, int8_t* dw, int32_t* acc_opt) {
    long n = (long)s->oc * s->c * s->kh * s->kw;
    int32_t* acc = acc_buf(acc_opt, n);
    niti_o_conv_dw_acc(s, x, dy, acc);
    niti_o_rescale_grad(acc, n, 2, dw);
    if (!acc_opt) free(acc);
}

void niti_o_matmul_nt(int e, int h, int l, const int8_t* a, const int8_t* b, int8_t* c, int32_t* acc_opt) {
    long n = (long)e * h;
    int32_t* acc = acc_buf(acc_opt, n);
    niti_o_matmul_nt_acc(e, h, l, a, b, acc);
    niti_o_rescale_grad(acc, n, 3, c);
    if (!acc_opt) free(acc);
}

/* optimizer/NITI_SGD.hpp:49-52 -> Binary SUB -> BinaryUtils.hpp:276-299 (clip to +-127) */
void niti_o_sgd_update(int8_t* w, const int8_t* dw, long n) {
    for (long i = 0; i < n; ++i) w[i] = (int8_t)niti_o_clip((int32_t)w[i] - (int32_t)dw[i]);
}

/* ---------------------------------------------------------------- element-wise ops of the step */

/* NITI_CPURelu_Int8.cpp:47-53 */
void niti_o_relu(const int8_t* x, int8_t* y, long n) {
    for (long i = 0; i < n; ++i) y[i] = x[i] >= 0 ? x[i] : 0;
}
/* NITI_CPUReluGrad_Int8.cpp:51-57 — gated on the forward INPUT being > 0 */
void niti_o_relu_grad(const int8_t* x, const int8_t* dy, int8_t* dx, long n) {
    for (long i = 0; i < n; ++i) dx[i] = x[i] > 0 ? dy[i] : 0;
}
/* NITI_Maxpool_Int8.cpp:24-70 (no padding in the demo; window clipped to the image) */
void niti_o_maxpool(const int8_t* x, int n, int c, int h, int w, int k, int s, int8_t* y) {
    int oh = (h - k) / s + 1, ow = (w - k) / s + 1;
    for (long z = 0; z < (long)n * c; ++z)
    for (int oy = 0; oy < oh; ++oy)
    for (int ox = 0; ox < ow; ++ox) {
        int8_t m = INT8_MIN;
        for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j) {
            int iy = oy * s + i, ix = ox * s + j;
            if (iy >= h || ix >= w) continue;
            int8_t v = x[(z * h + iy) * w + ix];
            if (v > m) m = v;
        }
        y[(z * oh + oy) * ow + ox] = m;
    }
}
/* NITI_CPUPoolGrad_Int8.cpp:36-72 — first window element (row-major) with value >= max takes the
 * gradient; overlapping windows accumulate with int8 wrap-around (dst[j] = dst[j] + diff). */
void niti_o_maxpool_grad(const int8_t* x, const int8_t* y, const int8_t* dy, int n, int c, int h, int w,
                         int k, int s, int8_t* dx) {
    int oh = (h - k) / s + 1, ow = (w - k) / s + 1;
    memset(dx, 0, (size_t)n * c * h * w);
    for (long z = 0; z < (long)n * c; ++z)
    for (int oy = 0; oy < oh; ++oy)
    for (int ox = 0; ox < ow; ++ox) {
        int8_t m = y[(z * oh + oy) * ow + ox];
        int8_t g = dy[(z * oh + oy) * ow + ox];
        int done = 0;
        for (int i = 0; i < k && !done; ++i) {
            int iy = oy * s + i;
            if (iy >= h) continue;
            for (int j = 0; j < k; ++j) {
                int ix = ox * s + j;
                if (ix >= w) continue;
                if (x[(z * h + iy) * w + ix] >= m) {
                    int8_t* d = &dx[(z * h + iy) * w + ix];
                    *d = (int8_t)(*d + g);
                    done = 1;
                    break;
                }
            }
        }
    }
}
/* NITI_CPULossGrad_Int8.cpp:82-201 — integer softmax cross-entropy gradient.
 * labels[i] = target class index (the reference scans a one-hot row for the first 1, :170-177). */
void niti_o_loss_grad(const int8_t* logits, int ascale, const int32_t* labels, int batch, int classes,
                      int8_t* grad) {
    long nel = (long)batch * classes;
    int64_t* e = (int64_t*)malloc(sizeof(int64_t) * (size_t)nel);
    int32_t* g = (int32_t*)malloc(sizeof(int32_t) * (size_t)nel);
    int8_t a = (int8_t)ascale;
    if (a > -7) {
        for (int i = 0; i < batch; ++i) {
            int64_t mx = 0;
            for (int j = 0; j < classes; ++j) {
                int64_t t = (int64_t)logits[i * classes + j] * 47274;
                t = t / (1 << 15);
                t = (a >= 0) ? t * (1 << a) : t / (1 << (-a));
                e[i * classes + j] = t;
                if (j == 0 || mx < t) mx = t;
            }
            mx -= 10;
            for (int j = 0; j < classes; ++j) {
                int64_t t = e[i * classes + j] - mx;
                t = t > 0 ? t : 0;
                e[i * classes + j] = (int64_t)((1 << t) - 1);
            }
        }
    } else {
        int64_t base = (int64_t)shl1(1 - 2 * (int)a);         /* int shift in the reference (:134) */
        int64_t shiftbase = (int64_t)shl1(1 - (int)a);
        for (long i = 0; i < nel; ++i) {
            int64_t t = logits[i];
            e[i] = base + t * shiftbase + t * t;
        }
    }
    for (int i = 0; i < batch; ++i) {
        int64_t sum = 0, gsum = 0;
        for (int j = 0; j < classes; ++j) sum += e[i * classes + j];
        for (int j = 0; j < classes; ++j) {
            int64_t t = e[i * classes + j] * (1 << 11);
            t = t / sum;
            e[i * classes + j] = t;
            gsum += t;
        }
        for (int j = 0; j < classes; ++j) g[i * classes + j] = (int32_t)e[i * classes + j];
        int t = labels[i];
        g[i * classes + t] = (int32_t)(e[i * classes + t] - gsum);
    }
    int32_t* r = (int32_t*)malloc(sizeof(int32_t) * (size_t)nel);
    niti_o_psto_shift(g, 4, r, nel);
    for (long i = 0; i < nel; ++i) grad[i] = (int8_t)r[i];
    free(r); free(g); free(e);
}

/* NITI_CPULoss_Int8.cpp:68-131 — the monitoring loss.  Float arithmetic in the reference's order: dequantise with powf(2, ascale),
 * MNNSoftmax per row (CommonOptFunction.cpp:1431-1456: max over the row, exp(x) = 2^div * degree-5 polynomial of the remainder,
 * normalise by the reciprocal of the sum), logf, pick the target class (one-hot row, :98-108), sum rows in order, -mean. */
#include <math.h>
float niti_o_loss_value(const int8_t* logits, int ascale, const int32_t* labels, int batch, int classes) {
    const float scale = powf(2.0f, (float)ascale);
    float total = 0.f;
    for (int i = 0; i < batch; ++i) {
        const int8_t* row = logits + (size_t)i * classes;
        float mx = scale * row[0];
        for (int j = 1; j < classes; ++j) { const float v = scale * row[j]; mx = v > mx ? v : mx; }
        const float xLimit = 87, param = 0.6931471805599453f;
        float sum = 0.f, picked = 0.f;
        for (int j = 0; j < classes; ++j) {
            float x = scale * row[j] - mx;
            x = x > -xLimit ? x : -xLimit;
            x = x < xLimit ? x : xLimit;
            const int div = (int)(x / param);
            const int div2 = (div + 127) << 23;
            const float rem = x - div * param;
            float basic; memcpy(&basic, &div2, 4);
            const float t = rem;
            const float e = basic * (((((1.0f / 120 * t + 1.0f / 24) * t + 1.0f / 6) * t + 0.5f) * t + 1.0f) * t + 1.0f);
            sum += e;
            if (j == labels[i]) picked = e;
        }
        total += logf(picked * (1.f / sum));
    }
    return total / batch * (-1);
}
