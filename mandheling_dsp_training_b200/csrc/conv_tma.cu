// conv_tma.cu — TMA-fed tcgen05 int8 implicit GEMM (the main contraction kernels).
//
// Same math and tile shapes as conv_tc.cu, but operand tiles are staged by the Tensor Memory
// Accelerator instead of 128 gathering threads:
//   * activations: cp.async.bulk.tensor.4d ... im2col  — the hardware walks 128 output pixels of the
//     NHWC tensor for one filter tap (offsets {kx*dil, ky*dil}) and one channel chunk, zero-filling the
//     padding halo, and writes a swizzled [128 pixels][BKC bytes] tile: no im2col buffer, no index math;
//   * weights / dY matrices: cp.async.bulk.tensor.2d (tiled), zero-filling ragged rows.
// One elected thread issues two TMA instructions per stage; one elected thread issues the MMAs; four
// warps drain TMEM.  Split-K (grid.z) keeps all SMs busy on the deep, small-M layers; partial tiles are
// combined with int32 atomics (integer adds commute => bit-exact).
//
// MODE 0 (forward, dX):  D[pixel][n]   += A_im2col[pixel][k] * B[n][k]          (K-major operands)
// MODE 1 (dW):           D[oc][c|tap]  += dY[pixel][oc]^T    * X_im2col[pixel][c] (MN-major operands)
//
// Replaces NITI_Conv_Int8.cpp:212-249, NITI_DeConv_Int8.cpp:208-290, NITI_GradientConv_Int8.cpp:218-268
// (im2col + NITI_MNNGemmInt8AddBiasScale_16x4_Unit, compute/Int8FunctionsOpt.cpp:201-407).
#include "common.cuh"
#include "tc_ptx.cuh"
#include <cuda.h>
#include <mutex>

namespace {

constexpr int BM = 128;
constexpr int kThreads = 192;     // warp 0: TMA, warp 1: MMA + TMEM owner, warps 2-5: epilogue

struct TmaParams {
    int32_t* acc;
    uint32_t* maxabs;
    int M, Ncols;                 // D extents (MODE 0: pixels x out-channels; MODE 1: oc x c)
    int out_h, out_w;             // pixel space enumerated by the im2col rows
    int kh, kw, dil, stride;
    int lower_h, lower_w;         // im2col bounding-box lower corner (= coordinate of pixel 0, tap 0)
    int flip;                     // dX: tap offsets run backwards
    int cchunks;                  // MODE 0: channel chunks per tap
    int num_kb, kb_per_split, splits;
    int taps, c_tiles, oc, c;     // MODE 1
    int npix;                     // MODE 1: reduction length (pixels)
};

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint32_t dst, uint64_t* bar, int x, int y) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(tc::smem_u32(bar)), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void tma_load_im2col(const CUtensorMap* map, uint32_t dst, uint64_t* bar, int c, int w, int h, int n,
                                                uint16_t off_w, uint16_t off_h) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
                 " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(tc::smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n),
                   "h"(off_w), "h"(off_h) : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// smem matrix descriptor with selectable swizzle (layout_type: 2 = 128B, 4 = 64B)
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)layout << 61;
    return d;
}

template <int BN, int BKC, int MODE> struct Cfg {
    static constexpr int A_BYTES = (MODE == 0) ? BM * BKC : 128 * 128;          // MODE 1: [128 pixels][128 oc]
    static constexpr int B_BYTES = (MODE == 0) ? BN * BKC : 128 * BKC;          // MODE 1: [128 pixels][BKC c]
    static constexpr int STAGE = A_BYTES + B_BYTES;
    static constexpr int STAGES = (STAGE <= 16384) ? 6 : (STAGE <= 32768 ? 4 : 3);
    static constexpr int NCOLS = (MODE == 0) ? BN : BKC;                         // accumulator columns
    static constexpr int TMEM_COLS = NCOLS < 32 ? 32 : NCOLS;
    static constexpr int SMEM = STAGES * STAGE + 1024 + 256;
};

template <int BN, int BKC, int MODE>
__global__ void __launch_bounds__(kThreads)
igemm_tma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const TmaParams p) {
    using C = Cfg<BN, BKC, MODE>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE);
    uint64_t* empty_bar = full_bar + C::STAGES;
    uint64_t* accum_bar = empty_bar + C::STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kb_begin = blockIdx.z * p.kb_per_split;
    const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
    const int num_kb = kb_end - kb_begin;

    // tile coordinates
    int m0, n0, tap = 0;
    if (MODE == 0) { m0 = blockIdx.x * BM; n0 = blockIdx.y * BN; }
    else {
        const int ct = blockIdx.x % p.c_tiles;
        tap = (blockIdx.x / p.c_tiles) % p.taps;
        m0 = (blockIdx.x / (p.c_tiles * p.taps)) * BM;     // oc0
        n0 = ct * BKC;                                     // c0
    }

    if (tid == 0) {
        for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        tc::mbar_init(accum_bar, 1);
        tc::fence_barrier_init();
        prefetch_tmap(&mapA); prefetch_tmap(&mapB);
    }
    if (warp == 1) tc::tmem_alloc<C::TMEM_COLS>(tmem_slot);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ------------------------------ TMA producer ------------------------------
            int bw = 0, bh = 0, bn = 0;
            if (MODE == 0) {
                int q = m0 % p.out_w; int t = m0 / p.out_w;
                int pp = t % p.out_h; bn = t / p.out_h;
                bw = q * p.stride + p.lower_w; bh = pp * p.stride + p.lower_h;
            }
            for (int it = 0; it < num_kb; ++it) {
                const int kb = kb_begin + it;
                const int s = it % C::STAGES; const uint32_t ph = (uint32_t)(it / C::STAGES) & 1u;
                tc::mbar_wait(&empty_bar[s], ph ^ 1u);
                const uint32_t a_dst = tc::smem_u32(smem + s * C::STAGE);
                const uint32_t b_dst = a_dst + C::A_BYTES;
                mbar_arrive_expect_tx(&full_bar[s], (uint32_t)C::STAGE);
                if (MODE == 0) {
                    const int tp = kb / p.cchunks, cc = kb - tp * p.cchunks;
                    int ky = tp / p.kw, kx = tp - ky * p.kw;
                    if (p.flip) { ky = p.kh - 1 - ky; kx = p.kw - 1 - kx; }
                    tma_load_im2col(&mapA, a_dst, &full_bar[s], cc * BKC, bw, bh, bn, (uint16_t)(kx * p.dil), (uint16_t)(ky * p.dil));
                    tma_load_2d(&mapB, b_dst, &full_bar[s], kb * BKC, n0);
                } else {
                    const int pix0 = kb * 128;
                    int q = pix0 % p.out_w; int t = pix0 / p.out_w;
                    int pp = t % p.out_h; int nn = t / p.out_h;
                    const int ky = tap / p.kw, kx = tap - ky * p.kw;
                    tma_load_2d(&mapA, a_dst, &full_bar[s], m0, pix0);
                    tma_load_im2col(&mapB, b_dst, &full_bar[s], n0, q * p.stride + p.lower_w, pp * p.stride + p.lower_h, nn,
                                    (uint16_t)(kx * p.dil), (uint16_t)(ky * p.dil));
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ------------------------------ MMA issuer ------------------------------
            constexpr uint32_t idesc = tc::make_idesc_i8(BM, C::NCOLS, MODE, MODE);
            constexpr uint32_t lay = (BKC == 128) ? 2u : 4u;
            for (int it = 0; it < num_kb; ++it) {
                const int s = it % C::STAGES; const uint32_t ph = (uint32_t)(it / C::STAGES) & 1u;
                tc::mbar_wait(&full_bar[s], ph);
                tc::tc_fence_after();
                const uint32_t a_base = tc::smem_u32(smem + s * C::STAGE);
                const uint32_t b_base = a_base + C::A_BYTES;
                if (MODE == 0) {
#pragma unroll
                    for (int kk = 0; kk < BKC / 32; ++kk) {
                        uint64_t ad = make_desc(a_base + kk * 32, 16, 8 * BKC, lay);
                        uint64_t bd = make_desc(b_base + kk * 32, 16, 8 * BKC, lay);
                        tc::umma_i8(tmem_base, ad, bd, idesc, (uint32_t)((it | kk) != 0));
                    }
                } else {
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {       // 128 pixels = 4 x K32
                        uint64_t ad = make_desc(a_base + kk * 32 * 128, 128 * 128, 1024, 2u);
                        uint64_t bd = make_desc(b_base + kk * 32 * BKC, 128 * BKC, 8 * BKC, lay);
                        tc::umma_i8(tmem_base, ad, bd, idesc, (uint32_t)((it | kk) != 0));
                    }
                }
                tc::umma_commit(&empty_bar[s]);
            }
            tc::umma_commit(accum_bar);
        }
    } else {
        // ------------------------------ epilogue (warps 2..5) ------------------------------
        const int quarter = warp & 3;                       // TMEM lane quarter this warp may read
        tc::mbar_wait(accum_bar, 0);
        tc::tc_fence_after();
        const int row = m0 + quarter * 32 + lane;
        const bool row_ok = row < p.M;
        uint32_t mx = 0;
        for (int c0 = 0; c0 < C::NCOLS; c0 += 32) {
            if (n0 + c0 >= p.Ncols) break;                  // warp-uniform
            uint32_t v[32];
            tc::tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, v);
            tc::tmem_ld_wait();
            if (row_ok) {
                int32_t* dst = (MODE == 0) ? p.acc + (size_t)row * p.Ncols + n0 + c0
                                           : p.acc + ((size_t)row * p.taps + tap) * p.c + n0 + c0;
                if (p.splits > 1) {
#pragma unroll
                    for (int q = 0; q < 32; ++q)
                        if (n0 + c0 + q < p.Ncols && v[q] != 0u) atomicAdd(dst + q, (int)v[q]);
                } else {
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        if (n0 + c0 + q * 4 < p.Ncols) {
                            int4 o = make_int4((int)v[q * 4], (int)v[q * 4 + 1], (int)v[q * 4 + 2], (int)v[q * 4 + 3]);
                            *reinterpret_cast<int4*>(dst + q * 4) = o;
                            mx = max(mx, max(max(niti_abs_u32(o.x), niti_abs_u32(o.y)), max(niti_abs_u32(o.z), niti_abs_u32(o.w))));
                        }
                    }
                }
            }
        }
        if (p.maxabs != nullptr && p.splits == 1) {
            mx = warp_max_u32(mx);
            if (lane == 0 && mx) atomicMax(p.maxabs, mx);
        }
    }
    __syncwarp();
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// host side: tensor maps through the driver entry points (no link-time libcuda dependency)
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_tiled = nullptr;
EncodeIm2colFn g_im2col = nullptr;
int g_driver_version = 0;
std::once_flag g_once;

bool load_driver() {
    std::call_once(g_once, [] {
        void* f = nullptr; cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            g_tiled = (EncodeTiledFn)f;
        f = nullptr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            g_im2col = (EncodeIm2colFn)f;
        cudaDriverGetVersion(&g_driver_version);
    });
    return g_tiled && g_im2col;
}

CUtensorMapSwizzle swz_of(int bkc) { return bkc == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B; }

// row-major byte matrix [rows][cols]; box = [box_rows][box_cols]
bool make_tiled(CUtensorMap* m, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows, uint32_t box_cols, int swz_bytes) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t es[2] = {1, 1};
    return g_tiled(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz_of(swz_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// NHWC int8 tensor [n][h][w][c]; box = 128 pixels x bkc channels
bool make_im2col(CUtensorMap* m, const void* base, int n, int h, int w, int c, int lower_h, int lower_w, int upper_h, int upper_w,
                 int stride, int bkc) {
    cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
    cuuint64_t strides[3] = {(cuuint64_t)c, (cuuint64_t)w * c, (cuuint64_t)h * w * c};
    int lower[2] = {lower_w, lower_h}, upper[2] = {upper_w, upper_h};
    cuuint32_t es[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    CUresult r = g_im2col(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(base), dims, strides, lower, upper,
                          (cuuint32_t)bkc, 128u, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swz_of(bkc),
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    // driver <= 13.1: im2col descriptors of tensors under 128 KiB carry a stray bit (same fix-up CUTLASS applies,
    // cute/atom/copy_traits_sm90_im2col.hpp)
    if (g_driver_version <= 13010 && (size_t)n * h * w * c < 131072) reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
    return true;
}

template <int BN, int BKC, int MODE>
int launch(niti_ctx* ctx, const CUtensorMap& a, const CUtensorMap& b, const TmaParams& p, dim3 grid, cudaStream_t st) {
    using C = Cfg<BN, BKC, MODE>;
    if (int rc_cfg = niti_func_smem(ctx, (const void*)igemm_tma_kernel<BN, BKC, MODE>, C::SMEM, "cudaFuncSetAttribute(igemm_tma)")) return rc_cfg;
    igemm_tma_kernel<BN, BKC, MODE><<<grid, kThreads, C::SMEM, st>>>(a, b, p);
    return niti_check_launch(ctx, "igemm_tma");
}

void choose_splits(niti_ctx* ctx, long tiles, TmaParams& p) {
    int splits = 1;
    if (tiles < ctx->sm_count) splits = (int)((ctx->sm_count + tiles - 1) / tiles);
    if (splits > p.num_kb / 2) splits = p.num_kb / 2;            // at least two K-blocks per split
    if (splits < 1) splits = 1;
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    p.splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
}

template <int BKC>
int run_mode0(niti_ctx* ctx, const CUtensorMap& ma, const void* bmat, TmaParams& p, int ktotal, cudaStream_t st) {
    const long mt = (p.M + BM - 1) / BM;
    int bn = p.Ncols <= 64 ? 64 : ((p.Ncols > 128 && mt * ((p.Ncols + 255) / 256) >= ctx->sm_count) ? 256 : 128);
    CUtensorMap mb;
    if (!make_tiled(&mb, bmat, (uint64_t)p.Ncols, (uint64_t)ktotal, (uint32_t)bn, BKC, BKC))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(B) failed");
    const long nt = (p.Ncols + bn - 1) / bn;
    choose_splits(ctx, mt * nt, p);
    if (p.splits > 1) {
        cudaError_t e = cudaMemsetAsync(p.acc, 0, (size_t)p.M * p.Ncols * sizeof(int32_t), st);
        if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "split-K memset", e);
    }
    dim3 grid((unsigned)mt, (unsigned)nt, (unsigned)p.splits);
    int rc;
    if (bn == 64) rc = launch<64, BKC, 0>(ctx, ma, mb, p, grid, st);
    else if (bn == 128) rc = launch<128, BKC, 0>(ctx, ma, mb, p, grid, st);
    else rc = launch<256, BKC, 0>(ctx, ma, mb, p, grid, st);
    if (rc) return rc;
    if (p.splits > 1 && p.maxabs) return niti_absmax(ctx, p.acc, (size_t)p.M * p.Ncols, p.maxabs, st);
    return NITI_OK;
}
} // namespace

bool tma_fwd_eligible(const niti_conv_desc* d) {
    return d->c % 64 == 0 && d->oc % 4 == 0 && d->pad <= 127 && (d->kh - 1) * d->dil <= 255 && (d->kw - 1) * d->dil <= 255 && d->stride <= 8;
}
bool tma_dx_eligible(const niti_conv_desc* d) {
    const int padp_h = (d->kh - 1) * d->dil - d->pad, padp_w = (d->kw - 1) * d->dil - d->pad;
    return d->stride == 1 && d->oc % 64 == 0 && d->c % 4 == 0 && padp_h >= 0 && padp_w >= 0 && padp_h <= 127 && padp_w <= 127 &&
           (d->kh - 1) * d->dil <= 255 && (d->kw - 1) * d->dil <= 255;
}
bool tma_dw_eligible(const niti_conv_desc* d) {
    return d->oc % 16 == 0 && d->c % 64 == 0 && d->pad <= 127 && (d->kh - 1) * d->dil <= 255 && (d->kw - 1) * d->dil <= 255 && d->stride <= 8;
}

int tma_conv_fwd_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, int32_t* acc,
                     uint32_t* maxabs, cudaStream_t st) {
    if (!load_driver()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    const int bkc = d->c % 128 == 0 ? 128 : 64;
    TmaParams p{};
    p.acc = acc; p.maxabs = maxabs; p.M = d->n * oh * ow; p.Ncols = d->oc;
    p.out_h = oh; p.out_w = ow; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = d->stride;
    p.lower_h = -d->pad; p.lower_w = -d->pad; p.flip = 0;
    p.cchunks = d->c / bkc; p.num_kb = d->kh * d->kw * p.cchunks;
    CUtensorMap ma;
    if (!make_im2col(&ma, x, d->n, d->h, d->w, d->c, -d->pad, -d->pad, d->pad - (d->kh - 1) * d->dil, d->pad - (d->kw - 1) * d->dil, d->stride, bkc))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(x) failed");
    const int ktotal = d->kh * d->kw * d->c;
    return bkc == 128 ? run_mode0<128>(ctx, ma, w, p, ktotal, st) : run_mode0<64>(ctx, ma, w, p, ktotal, st);
}

// wt: [c][kh][kw][oc] (transposed weights, K-major over (tap, oc))
int tma_conv_dx_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* wt, int32_t* acc,
                    uint32_t* maxabs, cudaStream_t st) {
    if (!load_driver()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    const int bkc = d->oc % 128 == 0 ? 128 : 64;
    const int pph = (d->kh - 1) * d->dil - d->pad, ppw = (d->kw - 1) * d->dil - d->pad;   // padding of the equivalent forward conv
    TmaParams p{};
    p.acc = acc; p.maxabs = maxabs; p.M = d->n * d->h * d->w; p.Ncols = d->c;
    p.out_h = d->h; p.out_w = d->w; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = 1;
    p.lower_h = -pph; p.lower_w = -ppw; p.flip = 1;
    p.cchunks = d->oc / bkc; p.num_kb = d->kh * d->kw * p.cchunks;
    CUtensorMap ma;
    // upper corner chosen so the box enumerates exactly h x w base positions: upper = (H - OH) + lower
    if (!make_im2col(&ma, dy, d->n, oh, ow, d->oc, -pph, -ppw, d->h - oh - pph, d->w - ow - ppw, 1, bkc))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(dy) failed");
    const int ktotal = d->kh * d->kw * d->oc;
    return bkc == 128 ? run_mode0<128>(ctx, ma, wt, p, ktotal, st) : run_mode0<64>(ctx, ma, wt, p, ktotal, st);
}

int tma_conv_dw_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* dy, int32_t* acc,
                    uint32_t* maxabs, cudaStream_t st) {
    if (!load_driver()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    const int bkc = d->c % 128 == 0 ? 128 : 64;
    TmaParams p{};
    p.acc = acc; p.maxabs = maxabs; p.M = d->oc; p.Ncols = d->c; p.oc = d->oc; p.c = d->c;
    p.out_h = oh; p.out_w = ow; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = d->stride;
    p.lower_h = -d->pad; p.lower_w = -d->pad; p.flip = 0;
    p.taps = d->kh * d->kw; p.c_tiles = d->c / bkc; p.npix = d->n * oh * ow;
    p.num_kb = (p.npix + 127) / 128;
    CUtensorMap ma, mb;
    if (!make_tiled(&ma, dy, (uint64_t)p.npix, (uint64_t)d->oc, 128, 128, 128))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(dy) failed");
    if (!make_im2col(&mb, x, d->n, d->h, d->w, d->c, -d->pad, -d->pad, d->pad - (d->kh - 1) * d->dil, d->pad - (d->kw - 1) * d->dil, d->stride, bkc))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(x) failed");
    const long tiles = (long)((d->oc + BM - 1) / BM) * p.taps * p.c_tiles;
    int splits = (int)((2L * ctx->sm_count + tiles - 1) / tiles);
    if (splits > p.num_kb) splits = p.num_kb;
    if (splits < 1) splits = 1;
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    p.splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
    const size_t nout = (size_t)d->oc * p.taps * d->c;
    if (p.splits > 1) {
        cudaError_t e = cudaMemsetAsync(acc, 0, nout * sizeof(int32_t), st);
        if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "dw memset", e);
    }
    dim3 grid((unsigned)tiles, 1, (unsigned)p.splits);
    int rc = bkc == 128 ? launch<128, 128, 1>(ctx, ma, mb, p, grid, st) : launch<64, 64, 1>(ctx, ma, mb, p, grid, st);
    if (rc) return rc;
    if (maxabs) return niti_absmax(ctx, acc, nout, maxabs, st);
    return NITI_OK;
}
