// elementwise.cu — the non-contraction int8 ops of the NITI training step and
// the layout conversions at the boundary.  All HBM-bound byte work on NHWC tensors.
//
// Reference behaviour restated (execution-engine/source/backend/cpu/):
//   NITI_CPURelu_Int8.cpp:47-53, NITI_CPUReluGrad_Int8.cpp:51-57, NITI_Maxpool_Int8.cpp:24-70,
//   NITI_CPUPoolGrad_Int8.cpp:36-72, NITI_CPULossGrad_Int8.cpp:82-201
#include "common.cuh"

namespace {
constexpr int kThreads = 256;

int grid_for(niti_ctx* ctx, size_t items) {
    size_t blocks = (items + kThreads - 1) / kThreads;
    size_t cap = (size_t)ctx->sm_count * 8;
    if (blocks > cap) blocks = cap;
    return (int)(blocks < 1 ? 1 : blocks);
}
bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// per-byte signed max with zero (relu) on packed int8x4
__device__ __forceinline__ uint32_t relu4(uint32_t v) { return __vmaxs4(v, 0u); }
// mask bytes of dy where x > 0
__device__ __forceinline__ uint32_t relu_grad4(uint32_t x, uint32_t dy) {
    uint32_t m = __vcmpgts4(x, 0u);     // 0xff where x > 0 (signed)
    return dy & m;
}

__global__ void __launch_bounds__(kThreads) relu_kernel(const int8_t* __restrict__ x, int8_t* __restrict__ y, size_t n, bool vec) {
    if (vec) {
        const size_t ng = n / 16;
        for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < ng; g += (size_t)gridDim.x * blockDim.x) {
            uint4 v = reinterpret_cast<const uint4*>(x)[g];
            reinterpret_cast<uint4*>(y)[g] = make_uint4(relu4(v.x), relu4(v.y), relu4(v.z), relu4(v.w));
        }
        if (blockIdx.x == 0)
            for (size_t i = ng * 16 + threadIdx.x; i < n; i += blockDim.x) y[i] = x[i] >= 0 ? x[i] : (int8_t)0;
    } else {
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
            y[i] = x[i] >= 0 ? x[i] : (int8_t)0;
    }
}

__global__ void __launch_bounds__(kThreads) relu_grad_kernel(const int8_t* __restrict__ x, const int8_t* __restrict__ dy,
                                                             int8_t* __restrict__ dx, size_t n, bool vec) {
    if (vec) {
        const size_t ng = n / 16;
        for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < ng; g += (size_t)gridDim.x * blockDim.x) {
            uint4 a = reinterpret_cast<const uint4*>(x)[g];
            uint4 d = reinterpret_cast<const uint4*>(dy)[g];
            reinterpret_cast<uint4*>(dx)[g] = make_uint4(relu_grad4(a.x, d.x), relu_grad4(a.y, d.y),
                                                         relu_grad4(a.z, d.z), relu_grad4(a.w, d.w));
        }
        if (blockIdx.x == 0)
            for (size_t i = ng * 16 + threadIdx.x; i < n; i += blockDim.x) dx[i] = x[i] > 0 ? dy[i] : (int8_t)0;
    } else {
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
            dx[i] = x[i] > 0 ? dy[i] : (int8_t)0;
    }
}

// NHWC max pooling, window clipped to the image; one thread per output element (c fastest => coalesced)
__global__ void __launch_bounds__(kThreads) maxpool_kernel(const int8_t* __restrict__ x, int n, int h, int w, int c,
                                                           int k, int s, int oh, int ow, int8_t* __restrict__ y) {
    const size_t total = (size_t)n * oh * ow * c;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int ch = (int)(i % c); size_t p = i / c;
        int ox = (int)(p % ow); p /= ow;
        int oy = (int)(p % oh); int b = (int)(p / oh);
        int m = -128;
        for (int ky = 0; ky < k; ++ky) {
            int iy = oy * s + ky; if (iy >= h) break;
            for (int kx = 0; kx < k; ++kx) {
                int ix = ox * s + kx; if (ix >= w) break;
                int v = x[(((size_t)b * h + iy) * w + ix) * c + ch];
                m = v > m ? v : m;
            }
        }
        y[i] = (int8_t)m;
    }
}

// First window element (row-major) with value >= max takes the gradient.  For non-overlapping
// windows (s >= k, the only case the reference's models use) each input element belongs to one
// window, so a gather formulation is exact: thread per input element.
// For overlapping windows the contributions accumulate with int8 wrap-around, also reproduced.
__global__ void __launch_bounds__(kThreads) maxpool_grad_kernel(const int8_t* __restrict__ x, const int8_t* __restrict__ y,
                                                                const int8_t* __restrict__ dy, int n, int h, int w, int c,
                                                                int k, int s, int oh, int ow, int8_t* __restrict__ dx, int gate_relu) {
    const size_t total = (size_t)n * h * w * c;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int ch = (int)(i % c); size_t p = i / c;
        int ix = (int)(p % w); p /= w;
        int iy = (int)(p % h); int b = (int)(p / h);
        int acc = 0;
        const int v = x[i];
        // windows (oy, ox) that contain (iy, ix): oy*s <= iy < oy*s + k
        int oy_hi = iy / s, ox_hi = ix / s;
        int oy_lo = (iy - k + s) / s; if (iy - k + 1 <= 0) oy_lo = 0;
        int ox_lo = (ix - k + s) / s; if (ix - k + 1 <= 0) ox_lo = 0;
        for (int oy = oy_lo; oy <= oy_hi && oy < oh; ++oy)
            for (int ox = ox_lo; ox <= ox_hi && ox < ow; ++ox) {
                size_t o = (((size_t)b * oh + oy) * ow + ox) * c + ch;
                int m = y[o];
                if (v < m) continue;
                // am I the first element >= m in row-major window order?
                bool first = true;
                for (int ky = 0; ky < k && first; ++ky) {
                    int yy = oy * s + ky; if (yy >= h) break;
                    for (int kx = 0; kx < k; ++kx) {
                        int xx = ox * s + kx; if (xx >= w) break;
                        if (yy == iy && xx == ix) { ky = k; break; }
                        if (x[(((size_t)b * h + yy) * w + xx) * c + ch] >= m) { first = false; break; }
                    }
                }
                if (first) acc += dy[o];
            }
        dx[i] = (gate_relu && v <= 0) ? (int8_t)0 : (int8_t)acc;
    }
}

// ---- vectorized variants: 16 channels (one 128-bit word) per thread, c % 16 == 0 -------------------
__device__ __forceinline__ uint4 vmax4x4(uint4 a, uint4 b) {
    return make_uint4(__vmaxs4(a.x, b.x), __vmaxs4(a.y, b.y), __vmaxs4(a.z, b.z), __vmaxs4(a.w, b.w));
}

__global__ void __launch_bounds__(kThreads) maxpool_vec_kernel(const uint4* __restrict__ x, int n, int h, int w, int c16,
                                                               int k, int s, int oh, int ow, uint4* __restrict__ y, unsigned long long* tl) {
    niti_pdl_trigger(); niti_pdl_wait();                     // programmatic dependent launch: see common.cuh
    niti_tl_enter(tl);
    const size_t total = (size_t)n * oh * ow * c16;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int cv = (int)(i % c16); size_t p = i / c16;
        int ox = (int)(p % ow); p /= ow;
        int oy = (int)(p % oh); int b = (int)(p / oh);
        uint4 m = make_uint4(0x80808080u, 0x80808080u, 0x80808080u, 0x80808080u);
        for (int ky = 0; ky < k; ++ky) {
            int iy = oy * s + ky; if (iy >= h) break;
            for (int kx = 0; kx < k; ++kx) {
                int ix = ox * s + kx; if (ix >= w) break;
                m = vmax4x4(m, x[(((size_t)b * h + iy) * w + ix) * c16 + cv]);
            }
        }
        y[i] = m;
    }
    niti_tl_exit(tl);
}

// non-overlapping windows (k == s): one thread owns a whole window x 16 channels, so the "first element
// >= max in row-major order wins" rule is a running mask and every dx word is written exactly once.
__global__ void __launch_bounds__(kThreads) maxpool_grad_vec_kernel(const uint4* __restrict__ x, const uint4* __restrict__ y,
                                                                    const uint4* __restrict__ dy, int n, int h, int w, int c16,
                                                                    int k, int oh, int ow, uint4* __restrict__ dx, int gate_relu,
                                                                    unsigned long long* tl) {
    niti_pdl_trigger(); niti_pdl_wait();                     // programmatic dependent launch: see common.cuh
    niti_tl_enter(tl);
    const size_t total = (size_t)n * oh * ow * c16;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int cv = (int)(i % c16); size_t p = i / c16;
        int ox = (int)(p % ow); p /= ow;
        int oy = (int)(p % oh); int b = (int)(p / oh);
        const uint4 g = dy[i];
        uint4 m;
        if (y != nullptr) m = y[i];
        else {                                               // window maximum taken from x itself
            m = make_uint4(0x80808080u, 0x80808080u, 0x80808080u, 0x80808080u);
            for (int ky = 0; ky < k; ++ky) {
                int iy = oy * k + ky; if (iy >= h) break;
                for (int kx = 0; kx < k; ++kx) {
                    int ix = ox * k + kx; if (ix >= w) break;
                    m = vmax4x4(m, x[(((size_t)b * h + iy) * w + ix) * c16 + cv]);
                }
            }
        }
        uint4 done = make_uint4(0u, 0u, 0u, 0u);
        for (int ky = 0; ky < k; ++ky) {
            int iy = oy * k + ky; if (iy >= h) break;
            for (int kx = 0; kx < k; ++kx) {
                int ix = ox * k + kx; if (ix >= w) break;
                const size_t o = (((size_t)b * h + iy) * w + ix) * c16 + cv;
                const uint4 v = x[o];
                uint4 ge = make_uint4(__vcmpges4(v.x, m.x), __vcmpges4(v.y, m.y), __vcmpges4(v.z, m.z), __vcmpges4(v.w, m.w));
                uint4 r = make_uint4(g.x & ge.x & ~done.x, g.y & ge.y & ~done.y, g.z & ge.z & ~done.z, g.w & ge.w & ~done.w);
                done = make_uint4(done.x | ge.x, done.y | ge.y, done.z | ge.z, done.w | ge.w);
                if (gate_relu) { r.x &= __vcmpgts4(v.x, 0u); r.y &= __vcmpgts4(v.y, 0u); r.z &= __vcmpgts4(v.z, 0u); r.w &= __vcmpgts4(v.w, 0u); }
                dx[o] = r;
            }
        }
    }
    niti_tl_exit(tl);
}

// integer softmax cross-entropy gradient; one thread per sample (batch is small, classes ~10).  Rows are `stride` bytes
// apart, the first `classes` entries are real, grad's padding columns are written as 0.  ev(j) and the 64-bit division are
// evaluated once per class (kept in a local array for classes <= 32, recomputed beyond that).
__global__ void loss_grad_kernel(const int8_t* __restrict__ logits, const int8_t* __restrict__ ascale_p,
                                 const int32_t* __restrict__ labels, int batch, int classes, int stride, int8_t* __restrict__ grad,
                                 unsigned long long* tl) {
    niti_pdl_trigger(); niti_pdl_wait();                     // programmatic dependent launch: see common.cuh
    niti_tl_enter(tl);
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= batch) return;
    const int a = (int)*ascale_p;
    const int8_t* row = logits + (size_t)i * stride;
    int8_t* out = grad + (size_t)i * stride;
    long long sum = 0;
    long long mx = 0;
    if (a > -7) {
        for (int j = 0; j < classes; ++j) {
            long long t = (long long)row[j] * 47274; t = t / (1 << 15);
            t = (a >= 0) ? t * (1 << a) : t / (1 << (-a));
            if (j == 0 || mx < t) mx = t;
        }
        mx -= 10;
    }
    const long long base = (long long)niti_shl1(1 - 2 * a), shiftbase = (long long)niti_shl1(1 - a);
    auto ev = [&](int j) -> long long {
        if (a > -7) {
            long long t = (long long)row[j] * 47274; t = t / (1 << 15);
            t = (a >= 0) ? t * (1 << a) : t / (1 << (-a));
            t -= mx; t = t > 0 ? t : 0;
            return (long long)((1 << (int)t) - 1);
        }
        long long t = row[j];
        return base + t * shiftbase + t * t;
    };
    constexpr int kKeep = 32;
    long long e[kKeep];
    for (int j = 0; j < classes; ++j) { const long long v = ev(j); if (j < kKeep) e[j] = v; sum += v; }
    long long gsum = 0;
    for (int j = 0; j < classes; ++j) {
        const long long g = ((j < kKeep ? e[j] : ev(j)) * (1 << 11)) / sum;
        if (j < kKeep) e[j] = g;
        gsum += g;
    }
    const int tgt = labels[i];
    for (int j = 0; j < classes; ++j) {
        long long g = j < kKeep ? e[j] : (ev(j) * (1 << 11)) / sum;
        if (j == tgt) g -= gsum;
        out[j] = (int8_t)niti_psto((int32_t)g, 4);
    }
    for (int j = classes; j < stride; ++j) out[j] = 0;
    niti_tl_exit(tl);
}

// The float loss scalar the demo prints (NITI_CPULoss_Int8.cpp:68-131): -mean_i log softmax(2^ascale * logits_i)[label_i] with MNN's
// polynomial exp (compute/CommonOptFunction.cpp:1431-1456).  Not on the training path (the gradient is loss_grad_kernel); here so
// that a step's loss is read back as 4 bytes instead of the logits.  One block; rows in parallel, then summed in row order like
// the reference (float addition is not associative).
__global__ void __launch_bounds__(256) loss_value_kernel(const int8_t* __restrict__ logits, const int8_t* __restrict__ ascale_p,
                                                         const int32_t* __restrict__ labels, int batch, int classes, int stride,
                                                         float* __restrict__ rows, float* __restrict__ out) {
    const float scale = exp2f((float)*ascale_p);             // exact for integer exponents, like powf(2, ascale)
    for (int i = threadIdx.x; i < batch; i += blockDim.x) {
        const int8_t* row = logits + (size_t)i * stride;
        float mx = scale * (float)row[0];
        for (int j = 1; j < classes; ++j) mx = fmaxf(mx, scale * (float)row[j]);
        const float xLimit = 87.f, param = 0.6931471805599453f;
        float sum = 0.f, picked = 0.f;
        const int lab = labels[i];
        for (int j = 0; j < classes; ++j) {
            float x = __fsub_rn(__fmul_rn(scale, (float)row[j]), mx);
            x = x > -xLimit ? x : -xLimit;
            x = x < xLimit ? x : xLimit;
            const int div = (int)__fdiv_rn(x, param);
            const float basic = __int_as_float((div + 127) << 23);
            const float t = __fsub_rn(x, __fmul_rn((float)div, param));
            // the reference's Horner chain with separate multiplies and adds (no contraction into FMAs)
            float pl = __fadd_rn(__fmul_rn(1.0f / 120, t), 1.0f / 24);
            pl = __fadd_rn(__fmul_rn(pl, t), 1.0f / 6);
            pl = __fadd_rn(__fmul_rn(pl, t), 0.5f);
            pl = __fadd_rn(__fmul_rn(pl, t), 1.0f);
            pl = __fadd_rn(__fmul_rn(pl, t), 1.0f);
            const float e = __fmul_rn(basic, pl);
            sum = __fadd_rn(sum, e);
            if (j == lab) picked = e;
        }
        rows[i] = logf(__fmul_rn(picked, __fdiv_rn(1.f, sum)));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float total = 0.f;
        for (int i = 0; i < batch; ++i) total = __fadd_rn(total, rows[i]);
        *out = __fmul_rn(__fdiv_rn(total, (float)batch), -1.f);
    }
}

// generic 4-d permutation of int8 tensors through index arithmetic (boundary only; small tensors)
// dst[a][b][c][d] (sizes A,B,C,D) = src[a*sa + b*sb + c*sc + d*sd]
__global__ void __launch_bounds__(kThreads) permute_kernel(const int8_t* __restrict__ src, int8_t* __restrict__ dst,
                                                           int A, int B, int C, int D, long sa, long sb, long sc, long sd) {
    const size_t total = (size_t)A * B * C * D;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int d = (int)(i % D); size_t p = i / D;
        int c = (int)(p % C); p /= C;
        int b = (int)(p % B); int a = (int)(p / B);
        dst[i] = src[a * sa + b * sb + c * sc + d * sd];
    }
}

// reference "NC4HW4" int8 (batch inside the channel block): [c/4][n][h*w][4]
__global__ void __launch_bounds__(kThreads) nc4_to_nhwc_kernel(const int8_t* __restrict__ src, int n, int c, int hw,
                                                               int8_t* __restrict__ dst, bool to_nc4) {
    const int c4 = (c + 3) / 4;
    const size_t total = (size_t)c4 * n * hw * 4;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int l = (int)(i & 3); size_t p = i >> 2;
        int px = (int)(p % hw); p /= hw;
        int b = (int)(p % n); int cb = (int)(p / n);
        int ch = cb * 4 + l;
        size_t nhwc = ((size_t)b * hw + px) * c + ch;
        if (to_nc4) dst[i] = ch < c ? src[nhwc] : (int8_t)0;
        else if (ch < c) dst[nhwc] = src[i];
    }
}

// generic 5-d strided gather with an optional split-axis bound (see niti_gather_i8 in niti_b200.h)
struct GatherArgs { int dims[5]; long long strides[5]; long long base; int split_hi, split_lo, split_limit; };
__global__ void __launch_bounds__(kThreads) gather_kernel(const int8_t* __restrict__ src, int8_t* __restrict__ dst, GatherArgs g) {
    const size_t total = (size_t)g.dims[0] * g.dims[1] * g.dims[2] * g.dims[3] * g.dims[4];
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int idx[5]; size_t p = i;
#pragma unroll
        for (int k = 4; k >= 0; --k) { idx[k] = (int)(p % g.dims[k]); p /= g.dims[k]; }
        long long off = g.base;
#pragma unroll
        for (int k = 0; k < 5; ++k) off += idx[k] * g.strides[k];
        bool ok = true;
        if (g.split_hi >= 0) ok = idx[g.split_hi] * g.dims[g.split_lo] + idx[g.split_lo] < g.split_limit;
        dst[i] = ok ? src[off] : (int8_t)0;
    }
}
} // namespace

// int8 Binary ADD / SUB / MUL with the result clipped to +-127 (BinaryUtils.hpp:276-299 niti_execute, NITI_int8_clip);
// bc: -1 element-wise, 0: a is one element, 1: b is one element
__global__ void __launch_bounds__(kThreads) binary_i8_kernel(int op, const int8_t* __restrict__ a, const int8_t* __restrict__ b,
                                                             size_t n, int bc, int8_t* __restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (size_t)gridDim.x * kThreads) {
        const int32_t x = (int32_t)a[bc == 0 ? 0 : i], y = (int32_t)b[bc == 1 ? 0 : i];
        const int32_t r = op == 0 ? x + y : (op == 1 ? x - y : x * y);
        out[i] = (int8_t)(r > 127 ? 127 : (r < -127 ? -127 : r));
    }
}

// NITI_CPULeftPoolGrad_Int8.cpp:19-52: zero insertion, [slab][h][w][4] blocks: out(y, x) = in(y/sy, x/sx) when both divide
__global__ void __launch_bounds__(kThreads) zero_insert_kernel(const uint32_t* __restrict__ in, int slabs, int ih, int iw, int oh, int ow,
                                                               int sy, int sx, uint32_t* __restrict__ out) {
    const size_t n = (size_t)slabs * oh * ow;
    for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (size_t)gridDim.x * kThreads) {
        const int x = (int)(i % ow); const size_t t = i / ow;
        const int y = (int)(t % oh); const size_t z = t / oh;
        out[i] = (y % sy == 0 && x % sx == 0) ? in[(z * ih + y / sy) * iw + x / sx] : 0u;
    }
}

// the same for NHWC tensors in 16-byte channel groups (strided dX: api.cu niti_conv_dx_fused)
__global__ void __launch_bounds__(kThreads) zero_insert_nhwc_kernel(const uint4* __restrict__ in, int n, int ih, int iw, int c16, int s, int oh, int ow,
                                                                    uint4* __restrict__ out) {
    const size_t total = (size_t)n * oh * ow * c16;
    for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < total; i += (size_t)gridDim.x * kThreads) {
        const int g = (int)(i % c16); size_t t = i / c16;
        const int x = (int)(t % ow); t /= ow;
        const int y = (int)(t % oh); const size_t b = t / oh;
        const int sy = y / s, sx = x / s;
        const bool live = sy * s == y && sx * s == x && sy < ih && sx < iw;
        out[i] = live ? in[((b * ih + sy) * iw + sx) * c16 + g] : make_uint4(0u, 0u, 0u, 0u);
    }
}

// NITI_Pad_Int8.cpp:26-66: every [h][w] plane copied into the middle of a zeroed [oh][ow] plane (rows pad..oh-pad-1, from column pad)
__global__ void __launch_bounds__(kThreads) pad2d_kernel(const int8_t* __restrict__ in, int planes, int ih, int iw, int pad, int oh, int ow,
                                                         int8_t* __restrict__ out) {
    const size_t n = (size_t)planes * oh * ow;
    for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (size_t)gridDim.x * kThreads) {
        const int x = (int)(i % ow); const size_t t = i / ow;
        const int y = (int)(t % oh); const size_t pl = t / oh;
        const bool in_rows = y >= pad && y < oh - pad, in_cols = x >= pad && x < pad + iw;
        out[i] = (in_rows && in_cols) ? in[(pl * ih + (y - pad)) * iw + (x - pad)] : (int8_t)0;
    }
}

// one-hot int32 targets [batch][classes] -> class index: the first j with target == 1 (NITI_CPULossGrad_Int8.cpp:170-178)
__global__ void onehot_labels_kernel(const int32_t* __restrict__ onehot, int batch, int classes, int32_t* __restrict__ labels) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= batch) return;
    int lab = 0;
    for (int j = 0; j < classes; ++j) if (onehot[(size_t)i * classes + j] == 1) { lab = j; break; }
    labels[i] = lab;
}

extern "C" int niti_relu(niti_ctx* ctx, const int8_t* x, int8_t* y, size_t n, niti_stream st) {
    if (!ctx || (n && (!x || !y))) return niti_fail(ctx, NITI_ERR_ARG, "niti_relu");
    if (!n) return NITI_OK;
    bool vec = aligned16(x) && aligned16(y);
    relu_kernel<<<grid_for(ctx, vec ? n / 16 + 1 : n), kThreads, 0, (cudaStream_t)st>>>(x, y, n, vec);
    return niti_check_launch(ctx, "relu");
}
extern "C" int niti_relu_grad(niti_ctx* ctx, const int8_t* x, const int8_t* dy, int8_t* dx, size_t n, niti_stream st) {
    if (!ctx || (n && (!x || !dy || !dx))) return niti_fail(ctx, NITI_ERR_ARG, "niti_relu_grad");
    if (!n) return NITI_OK;
    bool vec = aligned16(x) && aligned16(dy) && aligned16(dx);
    relu_grad_kernel<<<grid_for(ctx, vec ? n / 16 + 1 : n), kThreads, 0, (cudaStream_t)st>>>(x, dy, dx, n, vec);
    return niti_check_launch(ctx, "relu_grad");
}
extern "C" int niti_maxpool(niti_ctx* ctx, const int8_t* x, int n, int h, int w, int c, int k, int s, int8_t* y, niti_stream st) {
    if (!ctx || !x || !y || n <= 0 || h < k || w < k || c <= 0 || k <= 0 || s <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_maxpool");
    int oh = (h - k) / s + 1, ow = (w - k) / s + 1;
    if (c % 16 == 0 && aligned16(x) && aligned16(y)) {
        cudaError_t le = niti_launch(ctx, NITI_PDL_ELEMENTWISE, maxpool_vec_kernel, dim3(grid_for(ctx, (size_t)n * oh * ow * (c / 16))), dim3(kThreads), 0,
                                     (cudaStream_t)st, reinterpret_cast<const uint4*>(x), n, h, w, c / 16, k, s, oh, ow,
                                     reinterpret_cast<uint4*>(y), niti_tl_next(ctx, "maxpool"));
        if (le != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch maxpool_vec", le);
        return niti_check_launch(ctx, "maxpool_vec");
    }
    maxpool_kernel<<<grid_for(ctx, (size_t)n * oh * ow * c), kThreads, 0, (cudaStream_t)st>>>(x, n, h, w, c, k, s, oh, ow, y);
    return niti_check_launch(ctx, "maxpool");
}
extern "C" int niti_maxpool_grad_gated(niti_ctx* ctx, const int8_t* x, const int8_t* y, const int8_t* dy, int n, int h, int w, int c,
                                       int k, int s, int8_t* dx, int gate_relu, niti_stream st) {
    if (!ctx || !x || !y || !dy || !dx || n <= 0 || h < k || w < k || c <= 0 || k <= 0 || s <= 0)
        return niti_fail(ctx, NITI_ERR_ARG, "niti_maxpool_grad");
    int oh = (h - k) / s + 1, ow = (w - k) / s + 1;
    if (k == s && c % 16 == 0 && aligned16(x) && aligned16(y) && aligned16(dy) && aligned16(dx)) {
        if (oh * k != h || ow * k != w) {         // rows / columns no window covers stay zero
            cudaError_t e = cudaMemsetAsync(dx, 0, (size_t)n * h * w * c, (cudaStream_t)st);
            if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "maxpool_grad memset", e);
        }
        cudaError_t le = niti_launch(ctx, NITI_PDL_ELEMENTWISE, maxpool_grad_vec_kernel, dim3(grid_for(ctx, (size_t)n * oh * ow * (c / 16))), dim3(kThreads), 0,
                                     (cudaStream_t)st, reinterpret_cast<const uint4*>(x), reinterpret_cast<const uint4*>(y),
                                     reinterpret_cast<const uint4*>(dy), n, h, w, c / 16, k, oh, ow, reinterpret_cast<uint4*>(dx),
                                     gate_relu, niti_tl_next(ctx, "maxpool_grad"));
        if (le != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch maxpool_grad_vec", le);
        return niti_check_launch(ctx, "maxpool_grad_vec");
    }
    maxpool_grad_kernel<<<grid_for(ctx, (size_t)n * h * w * c), kThreads, 0, (cudaStream_t)st>>>(x, y, dy, n, h, w, c, k, s, oh, ow, dx, gate_relu);
    return niti_check_launch(ctx, "maxpool_grad");
}
extern "C" int niti_maxpool_grad(niti_ctx* ctx, const int8_t* x, const int8_t* y, const int8_t* dy, int n, int h, int w, int c,
                                 int k, int s, int8_t* dx, niti_stream st) {
    return niti_maxpool_grad_gated(ctx, x, y, dy, n, h, w, c, k, s, dx, 0, st);
}
extern "C" int niti_loss_grad_strided(niti_ctx* ctx, const int8_t* logits, const int8_t* ascale, const int32_t* labels,
                                      int batch, int classes, int stride, int8_t* grad, niti_stream st) {
    if (!ctx || !logits || !ascale || !labels || !grad || batch <= 0 || classes <= 0 || stride < classes)
        return niti_fail(ctx, NITI_ERR_ARG, "niti_loss_grad");
    cudaError_t le = niti_launch(ctx, NITI_PDL_ELEMENTWISE, loss_grad_kernel, dim3((batch + 31) / 32), dim3(32), 0, (cudaStream_t)st, logits, ascale, labels,
                                 batch, classes, stride, grad, niti_tl_next(ctx, "loss_grad"));
    if (le != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch loss_grad", le);
    return niti_check_launch(ctx, "loss_grad");
}
extern "C" int niti_loss_value(niti_ctx* ctx, const int8_t* logits, const int8_t* ascale, const int32_t* labels, int batch, int classes,
                               int stride, float* row_scratch, float* loss_out, niti_stream st) {
    if (!ctx || !logits || !ascale || !labels || !row_scratch || !loss_out || batch <= 0 || classes < 2 || stride < classes)
        return niti_fail(ctx, NITI_ERR_ARG, "niti_loss_value: bad argument (classes >= 2: the reference's softmax reads two entries)");
    loss_value_kernel<<<1, 256, 0, (cudaStream_t)st>>>(logits, ascale, labels, batch, classes, stride, row_scratch, loss_out);
    return niti_check_launch(ctx, "loss_value");
}
extern "C" int niti_loss_grad(niti_ctx* ctx, const int8_t* logits, const int8_t* ascale, const int32_t* labels,
                              int batch, int classes, int8_t* grad, niti_stream st) {
    return niti_loss_grad_strided(ctx, logits, ascale, labels, batch, classes, classes, grad, st);
}
int niti_maxpool_grad_from_x(niti_ctx* ctx, const int8_t* x, const int8_t* dy, int n, int h, int w, int c, int8_t* dx, int gate_relu,
                             niti_stream st) {
    if (!ctx || !x || !dy || !dx || n <= 0 || h < 2 || w < 2 || (h & 1) || (w & 1) || c % 16 != 0 || !aligned16(x) || !aligned16(dy) || !aligned16(dx))
        return niti_fail(ctx, NITI_ERR_UNSUPPORTED, "max-pool gradient from x: needs even h, w and c % 16 == 0");
    const int oh = h / 2, ow = w / 2;
    cudaError_t le = niti_launch(ctx, NITI_PDL_ELEMENTWISE, maxpool_grad_vec_kernel, dim3(grid_for(ctx, (size_t)n * oh * ow * (c / 16))), dim3(kThreads), 0,
                                 (cudaStream_t)st, reinterpret_cast<const uint4*>(x), (const uint4*)nullptr,
                                 reinterpret_cast<const uint4*>(dy), n, h, w, c / 16, 2, oh, ow, reinterpret_cast<uint4*>(dx), gate_relu,
                                 niti_tl_next(ctx, "maxpool_grad"));
    if (le != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch maxpool_grad_vec", le);
    return niti_check_launch(ctx, "maxpool_grad_vec");
}
extern "C" int niti_binary_i8(niti_ctx* ctx, int op, const int8_t* a, const int8_t* b, size_t n, int broadcast, int8_t* out, niti_stream st) {
    if (!ctx || !a || !b || !out || op < 0 || op > 2 || broadcast < -1 || broadcast > 1) return niti_fail(ctx, NITI_ERR_ARG, "niti_binary_i8: bad argument");
    if (n == 0) return NITI_OK;
    const size_t blocks = (n + kThreads - 1) / kThreads;
    binary_i8_kernel<<<(unsigned)(blocks < 4096 ? blocks : 4096), kThreads, 0, (cudaStream_t)st>>>(op, a, b, n, broadcast, out);
    return niti_check_launch(ctx, "binary_i8");
}

extern "C" int niti_zero_insert_c4(niti_ctx* ctx, const int8_t* in, int slabs, int ih, int iw, int oh, int ow, int sy, int sx,
                                   int8_t* out, niti_stream st) {
    if (!ctx || !in || !out || slabs <= 0 || ih <= 0 || iw <= 0 || oh <= 0 || ow <= 0 || sy <= 0 || sx <= 0)
        return niti_fail(ctx, NITI_ERR_ARG, "niti_zero_insert_c4: bad argument");
    if ((oh - 1) / sy >= ih || (ow - 1) / sx >= iw) return niti_fail(ctx, NITI_ERR_ARG, "niti_zero_insert_c4: input smaller than the strided output needs");
    if ((reinterpret_cast<uintptr_t>(in) & 3u) || (reinterpret_cast<uintptr_t>(out) & 3u)) return niti_fail(ctx, NITI_ERR_ARG, "niti_zero_insert_c4: 4-byte alignment");
    const size_t n = (size_t)slabs * oh * ow, blocks = (n + kThreads - 1) / kThreads;
    zero_insert_kernel<<<(unsigned)(blocks < 4096 ? blocks : 4096), kThreads, 0, (cudaStream_t)st>>>(
        reinterpret_cast<const uint32_t*>(in), slabs, ih, iw, oh, ow, sy, sx, reinterpret_cast<uint32_t*>(out));
    return niti_check_launch(ctx, "zero_insert");
}

int niti_zero_insert_nhwc(niti_ctx* ctx, const int8_t* in, int n, int ih, int iw, int c, int s, int oh, int ow, int8_t* out, niti_stream st) {
    if (!ctx || !in || !out || n <= 0 || ih <= 0 || iw <= 0 || c <= 0 || c % 16 != 0 || s <= 0 || oh <= 0 || ow <= 0 || !aligned16(in) || !aligned16(out))
        return niti_fail(ctx, NITI_ERR_ARG, "niti_zero_insert_nhwc: bad argument");
    zero_insert_nhwc_kernel<<<grid_for(ctx, (size_t)n * oh * ow * (c / 16)), kThreads, 0, (cudaStream_t)st>>>(
        reinterpret_cast<const uint4*>(in), n, ih, iw, c / 16, s, oh, ow, reinterpret_cast<uint4*>(out));
    return niti_check_launch(ctx, "zero_insert_nhwc");
}

extern "C" int niti_pad2d(niti_ctx* ctx, const int8_t* in, int planes, int ih, int iw, int pad, int oh, int ow, int8_t* out, niti_stream st) {
    if (!ctx || !in || !out || planes <= 0 || ih <= 0 || iw <= 0 || pad < 0 || oh - 2 * pad > ih || oh - 2 * pad < 0 || ow < iw + pad)
        return niti_fail(ctx, NITI_ERR_ARG, "niti_pad2d: bad argument");
    const size_t n = (size_t)planes * oh * ow, blocks = (n + kThreads - 1) / kThreads;
    pad2d_kernel<<<(unsigned)(blocks < 4096 ? blocks : 4096), kThreads, 0, (cudaStream_t)st>>>(in, planes, ih, iw, pad, oh, ow, out);
    return niti_check_launch(ctx, "pad2d");
}

extern "C" int niti_onehot_labels(niti_ctx* ctx, const int32_t* onehot, int batch, int classes, int32_t* labels, niti_stream st) {
    if (!ctx || !onehot || !labels || batch <= 0 || classes <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_onehot_labels: bad argument");
    onehot_labels_kernel<<<(batch + 127) / 128, 128, 0, (cudaStream_t)st>>>(onehot, batch, classes, labels);
    return niti_check_launch(ctx, "onehot_labels");
}

extern "C" int niti_nchw_to_nhwc(niti_ctx* ctx, const int8_t* src, int n, int c, int h, int w, int8_t* dst, niti_stream st) {
    if (!ctx || !src || !dst || n <= 0 || c <= 0 || h <= 0 || w <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_nchw_to_nhwc");
    permute_kernel<<<grid_for(ctx, (size_t)n * c * h * w), kThreads, 0, (cudaStream_t)st>>>(
        src, dst, n, h, w, c, (long)c * h * w, (long)w, 1L, (long)h * w);
    return niti_check_launch(ctx, "nchw_to_nhwc");
}
extern "C" int niti_nhwc_to_nchw(niti_ctx* ctx, const int8_t* src, int n, int c, int h, int w, int8_t* dst, niti_stream st) {
    if (!ctx || !src || !dst || n <= 0 || c <= 0 || h <= 0 || w <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_nhwc_to_nchw");
    permute_kernel<<<grid_for(ctx, (size_t)n * c * h * w), kThreads, 0, (cudaStream_t)st>>>(
        src, dst, n, c, h, w, (long)h * w * c, 1L, (long)w * c, (long)c);
    return niti_check_launch(ctx, "nhwc_to_nchw");
}
extern "C" int niti_nc4hw4_to_nhwc(niti_ctx* ctx, const int8_t* src, int n, int c, int h, int w, int8_t* dst, niti_stream st) {
    if (!ctx || !src || !dst || n <= 0 || c <= 0 || h <= 0 || w <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_nc4hw4_to_nhwc");
    nc4_to_nhwc_kernel<<<grid_for(ctx, (size_t)((c + 3) / 4) * n * h * w * 4), kThreads, 0, (cudaStream_t)st>>>(src, n, c, h * w, dst, false);
    return niti_check_launch(ctx, "nc4hw4_to_nhwc");
}
extern "C" int niti_nhwc_to_nc4hw4(niti_ctx* ctx, const int8_t* src, int n, int c, int h, int w, int8_t* dst, niti_stream st) {
    if (!ctx || !src || !dst || n <= 0 || c <= 0 || h <= 0 || w <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_nhwc_to_nc4hw4");
    nc4_to_nhwc_kernel<<<grid_for(ctx, (size_t)((c + 3) / 4) * n * h * w * 4), kThreads, 0, (cudaStream_t)st>>>(src, n, c, h * w, dst, true);
    return niti_check_launch(ctx, "nhwc_to_nc4hw4");
}
extern "C" int niti_gather_i8(niti_ctx* ctx, const int8_t* src, int64_t base, int8_t* dst, const int32_t dims[5],
                              const int64_t strides[5], int split_hi, int split_lo, int split_limit, niti_stream st) {
    if (!ctx || !src || !dst || !dims || !strides || split_hi > 4 || split_lo > 4 || (split_hi >= 0 && split_lo < 0))
        return niti_fail(ctx, NITI_ERR_ARG, "niti_gather_i8");
    GatherArgs g; size_t total = 1;
    for (int k = 0; k < 5; ++k) {
        if (dims[k] <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_gather_i8: dims");
        g.dims[k] = dims[k]; g.strides[k] = strides[k]; total *= (size_t)dims[k];
    }
    g.base = base; g.split_hi = split_hi; g.split_lo = split_lo; g.split_limit = split_limit;
    gather_kernel<<<grid_for(ctx, total), kThreads, 0, (cudaStream_t)st>>>(src, dst, g);
    return niti_check_launch(ctx, "gather_i8");
}
