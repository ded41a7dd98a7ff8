// conv_simt.cu — SIMT (CUDA-core, dp4a where aligned) int8 -> int32 contractions for the shapes the
// tcgen05 kernels do not take: channel counts that are not multiples of 16 (MNIST's 1/20/52/500/12,
// VGG's 3-channel stem, the 12-way classifier).  These layers are HBM/latency-bound (K = 25..500),
// so the rules that matter are coalescing (channel index fastest across a warp) and enough CTAs.
// Also the on-device cross-check for the tensor-core kernels in the tests.
//
// Replaces the im2col + NITI_MNNGemmInt8AddBiasScale_16x4_Unit loops of the reference
// (compute/Int8FunctionsOpt.cpp:201-232,270-407; NITI_Conv_Int8.cpp:218-249;
//  NITI_DeConv_Int8.cpp:208-290; NITI_GradientConv_Int8.cpp:218-268).  Integer accumulation is
// bit-identical to the reference's float accumulation while |partial sums| < 2^24 (quirk Q1).
#include "common.cuh"

namespace {

struct Geo {
    int n, c, h, w, oc, kh, kw, stride, pad, dil, oh, ow;
};

__device__ __forceinline__ void block_max_to_global(uint32_t m, uint32_t* maxabs) {
    if (maxabs == nullptr) return;
    m = warp_max_u32(m);
    if ((threadIdx.x & 31) == 0 && m) atomicMax(maxabs, m);
}

// ---- forward: one thread per (pixel, oc); threads of a warp share the pixel (x broadcast) -------
__global__ void __launch_bounds__(256) fwd_kernel(Geo g, const int8_t* __restrict__ x, const int8_t* __restrict__ w,
                                                  int32_t* __restrict__ acc, uint32_t* __restrict__ maxabs) {
    const size_t total = (size_t)g.n * g.oh * g.ow * g.oc;
    const bool vec4 = (g.c % 4) == 0;
    uint32_t m = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int o = (int)(i % g.oc); size_t p = i / g.oc;
        int ox = (int)(p % g.ow); p /= g.ow;
        int oy = (int)(p % g.oh); int b = (int)(p / g.oh);
        int32_t a = 0;
        const int8_t* wrow = w + (size_t)o * g.kh * g.kw * g.c;
        for (int ky = 0; ky < g.kh; ++ky) {
            int iy = oy * g.stride - g.pad + ky * g.dil;
            if (iy < 0 || iy >= g.h) continue;
            for (int kx = 0; kx < g.kw; ++kx) {
                int ix = ox * g.stride - g.pad + kx * g.dil;
                if (ix < 0 || ix >= g.w) continue;
                const int8_t* xp = x + (((size_t)b * g.h + iy) * g.w + ix) * g.c;
                const int8_t* wp = wrow + (size_t)(ky * g.kw + kx) * g.c;
                if (vec4) {
                    const int* x4 = reinterpret_cast<const int*>(xp);
                    const int* w4 = reinterpret_cast<const int*>(wp);
                    for (int cc = 0; cc < g.c / 4; ++cc) a = __dp4a(x4[cc], w4[cc], a);
                } else {
                    for (int cc = 0; cc < g.c; ++cc) a += (int)xp[cc] * (int)wp[cc];
                }
            }
        }
        acc[i] = a;
        m = max(m, niti_abs_u32(a));
    }
    block_max_to_global(m, maxabs);
}

// ---- dX: one thread per (n,h,w,c); c fastest => w reads coalesced, dy broadcast ----------------
// dx[n,iy,ix,c] = sum_{ky,kx,o} dy[n,oy,ox,o] * w[o,ky,kx,c],  oy*stride = iy + pad - ky*dil
__global__ void __launch_bounds__(256) dx_kernel(Geo g, const int8_t* __restrict__ dy, const int8_t* __restrict__ w,
                                                 int32_t* __restrict__ acc, uint32_t* __restrict__ maxabs) {
    const size_t total = (size_t)g.n * g.h * g.w * g.c;
    const size_t wstride_o = (size_t)g.kh * g.kw * g.c;
    uint32_t m = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int ch = (int)(i % g.c); size_t p = i / g.c;
        int ix = (int)(p % g.w); p /= g.w;
        int iy = (int)(p % g.h); int b = (int)(p / g.h);
        int32_t a = 0;
        for (int ky = 0; ky < g.kh; ++ky) {
            int ty = iy + g.pad - ky * g.dil;
            if (ty < 0 || (ty % g.stride) != 0) continue;
            int oy = ty / g.stride; if (oy >= g.oh) continue;
            for (int kx = 0; kx < g.kw; ++kx) {
                int tx = ix + g.pad - kx * g.dil;
                if (tx < 0 || (tx % g.stride) != 0) continue;
                int ox = tx / g.stride; if (ox >= g.ow) continue;
                const int8_t* dyp = dy + (((size_t)b * g.oh + oy) * g.ow + ox) * g.oc;
                const int8_t* wp = w + (size_t)(ky * g.kw + kx) * g.c + ch;
                for (int o = 0; o < g.oc; ++o) a += (int)dyp[o] * (int)wp[o * wstride_o];
            }
        }
        acc[i] = a;
        m = max(m, niti_abs_u32(a));
    }
    block_max_to_global(m, maxabs);
}

// ---- dW: tiled GEMM D[oc][(tap,c)] = sum_pix dy[pix][oc] * x[pix+tap][c], split over pixels ------
// block = 16x16 threads, tile 32(oc) x 32(c) for one tap, 32 pixels per smem stage;
// grid = (c tiles * taps, oc tiles, pixel splits); partial sums land with integer atomics
// (order-independent => still bit-exact).
constexpr int DW_T = 32;
__global__ void __launch_bounds__(256) dw_kernel(Geo g, const int8_t* __restrict__ x, const int8_t* __restrict__ dy,
                                                 int32_t* __restrict__ acc, int pix_per_split) {
    __shared__ int8_t s_dy[DW_T][DW_T + 4];   // [pix][oc]
    __shared__ int8_t s_x[DW_T][DW_T + 4];    // [pix][c]
    const int ctiles = (g.c + DW_T - 1) / DW_T;
    const int tap = blockIdx.x / ctiles, ct = blockIdx.x % ctiles;
    const int ky = tap / g.kw, kx = tap % g.kw;
    const int o0 = blockIdx.y * DW_T, c0 = ct * DW_T;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const long npix = (long)g.n * g.oh * g.ow;
    const long p_begin = (long)blockIdx.z * pix_per_split;
    const long p_end = min(npix, p_begin + pix_per_split);
    int32_t a00 = 0, a01 = 0, a10 = 0, a11 = 0;
    for (long p0 = p_begin; p0 < p_end; p0 += DW_T) {
        // cooperative load: 256 threads, 32x32 bytes each array => 4 bytes per thread per array
        for (int e = threadIdx.x; e < DW_T * DW_T; e += 256) {
            int pp = e / DW_T, col = e % DW_T;
            long p = p0 + pp;
            int8_t vdy = 0, vx = 0;
            if (p < p_end) {
                int ox = (int)(p % g.ow); long q = p / g.ow;
                int oy = (int)(q % g.oh); int b = (int)(q / g.oh);
                if (o0 + col < g.oc) vdy = dy[(size_t)p * g.oc + o0 + col];
                int iy = oy * g.stride - g.pad + ky * g.dil, ix = ox * g.stride - g.pad + kx * g.dil;
                if (c0 + col < g.c && iy >= 0 && iy < g.h && ix >= 0 && ix < g.w)
                    vx = x[(((size_t)b * g.h + iy) * g.w + ix) * g.c + c0 + col];
            }
            s_dy[pp][col] = vdy; s_x[pp][col] = vx;
        }
        __syncthreads();
#pragma unroll 8
        for (int pp = 0; pp < DW_T; ++pp) {
            int d0 = s_dy[pp][ty], d1 = s_dy[pp][ty + 16];
            int x0 = s_x[pp][tx], x1 = s_x[pp][tx + 16];
            a00 += d0 * x0; a01 += d0 * x1; a10 += d1 * x0; a11 += d1 * x1;
        }
        __syncthreads();
    }
    const size_t ostride = (size_t)g.kh * g.kw * g.c;
    auto put = [&](int o, int c, int32_t v) {
        if (o < g.oc && c < g.c && v != 0) atomicAdd(&acc[(size_t)o * ostride + (size_t)tap * g.c + c], v);
    };
    put(o0 + ty, c0 + tx, a00); put(o0 + ty, c0 + tx + 16, a01);
    put(o0 + ty + 16, c0 + tx, a10); put(o0 + ty + 16, c0 + tx + 16, a11);
}

int grid1d(niti_ctx* ctx, size_t items, int threads) {
    size_t blocks = (items + threads - 1) / threads;
    size_t cap = (size_t)ctx->sm_count * 8;
    if (blocks > cap) blocks = cap;
    return (int)(blocks < 1 ? 1 : blocks);
}

Geo make_geo(const niti_conv_desc* d) {
    Geo g{d->n, d->c, d->h, d->w, d->oc, d->kh, d->kw, d->stride, d->pad, d->dil, conv_out_h(d), conv_out_w(d)};
    return g;
}
} // namespace

int simt_conv_fwd_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, int32_t* acc,
                      uint32_t* maxabs, cudaStream_t st) {
    Geo g = make_geo(d);
    size_t total = (size_t)g.n * g.oh * g.ow * g.oc;
    fwd_kernel<<<grid1d(ctx, total, 256), 256, 0, st>>>(g, x, w, acc, maxabs);
    return niti_check_launch(ctx, "simt fwd");
}

int simt_conv_dx_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* w, int32_t* acc,
                     uint32_t* maxabs, cudaStream_t st) {
    Geo g = make_geo(d);
    size_t total = (size_t)g.n * g.h * g.w * g.c;
    dx_kernel<<<grid1d(ctx, total, 256), 256, 0, st>>>(g, dy, w, acc, maxabs);
    return niti_check_launch(ctx, "simt dx");
}

int simt_conv_dw_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* dy, int32_t* acc,
                     uint32_t* maxabs, cudaStream_t st) {
    Geo g = make_geo(d);
    size_t nout = (size_t)g.oc * g.kh * g.kw * g.c;
    cudaError_t e = cudaMemsetAsync(acc, 0, nout * sizeof(int32_t), st);
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "simt dw memset", e);
    const int ctiles = (g.c + DW_T - 1) / DW_T, otiles = (g.oc + DW_T - 1) / DW_T;
    const long npix = (long)g.n * g.oh * g.ow;
    long tiles = (long)ctiles * g.kh * g.kw * otiles;
    long want = ((long)ctx->sm_count * 4 + tiles - 1) / tiles;           // ~4 CTAs per SM
    long max_splits = (npix + 255) / 256;                                 // >= 256 pixels per split
    long splits = want < 1 ? 1 : (want > max_splits ? max_splits : want);
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
    int pps = (int)(((npix + splits - 1) / splits + DW_T - 1) / DW_T * DW_T);
    splits = (npix + pps - 1) / pps;
    dim3 grid((unsigned)(ctiles * g.kh * g.kw), (unsigned)otiles, (unsigned)splits);
    dw_kernel<<<grid, 256, 0, st>>>(g, x, dy, acc, pps);
    int rc = niti_check_launch(ctx, "simt dw");
    if (rc) return rc;
    if (maxabs) return niti_absmax(ctx, acc, nout, maxabs, st);
    return NITI_OK;
}
