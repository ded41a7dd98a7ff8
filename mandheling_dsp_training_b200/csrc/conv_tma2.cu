// conv_tma2.cu — persistent, CTA-pair (cta_group::2) tcgen05 int8 implicit GEMM for forward and dX.
//
// Why a second generation: conv_tma.cu runs one 128xBN tile per CTA; its operand traffic per MMA op
// (128 op/B at 128x128) saturates the L2->SM path (~12 TB/s chip-wide) at ~1/5 of the int8 tensor peak, and its
// epilogue (row-strided int32 stores) is serialized behind the main loop.  This kernel
//   * pairs two SMs on one 256 x BN accumulator tile (tcgen05.mma.cta_group::2, UMMA M = 256): each CTA stages its
//     own 128 im2col rows of A and HALF of the B tile, so bytes per op drop to 1/262 at BN = 256;
//   * is persistent: one CTA pair per TPC walks a static tile schedule, TMEM holds two accumulator stages
//     (2 x BN columns) so the epilogue of tile i overlaps the main loop of tile i+1;
//   * drains TMEM through swizzled shared memory and TMA stores (cp.async.bulk.tensor, full 128-byte lines), or TMA
//     reduce-add (cp.reduce.async.bulk ... .add) for split-K partial tiles — integer adds commute, so bit-exact;
//   * keeps the im2col-free operand path: A tiles come from cp.async.bulk.tensor.4d.im2col on the NHWC tensor.
//
// D[pixel][n] += A_im2col[pixel][k] * B[n][k], K-major operands, int8 x int8 -> int32 (+ max|acc| for NITI's exponent).
// Replaces NITI_Conv_Int8.cpp:212-249 and NITI_DeConv_Int8.cpp:208-290 (im2col + NITI_MNNGemmInt8AddBiasScale_16x4_Unit,
// compute/Int8FunctionsOpt.cpp:201-407).
#include "common.cuh"
#include "tc_ptx.cuh"
#include "pair_ptx.cuh"
#include <cuda.h>
#include <mutex>

namespace {

constexpr int kThreads2 = 192;      // warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-5: epilogue

struct Tma2Params {
    uint32_t* maxabs;
    int M, Ncols;                 // D extents: pixels x out-channels
    int out_h, out_w;             // pixel space enumerated by the im2col rows
    int kh, kw, dil, stride;
    int lower_h, lower_w;         // coordinate of (pixel 0, tap 0)
    int flip;                     // dX: tap offsets run backwards
    int cchunks;                  // channel chunks per tap
    int num_kb, kb_per_split, splits;
    int m_tiles, n_tiles, items;  // items = m_tiles * n_tiles * splits
    int32_t* acc;                 // split-K only: zeroed in-kernel before the first reduce-add, re-read for max|acc|
    uint32_t* sync;               // split-K only: {arrival counter, exit counter} of the grid barrier (self-resetting)
};

__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }      // the 4 epilogue warps
__device__ __forceinline__ void grid_barrier(uint32_t* counter, uint32_t target) { niti_grid_barrier(counter, target); }   // common.cuh

template <int BN, int BKC> struct Cfg2 {
    static constexpr int A_BYTES = 128 * BKC;               // this CTA's 128 im2col rows
    static constexpr int B_BYTES = (BN / 2) * BKC;          // this CTA's half of the B tile
    static constexpr int STAGE = A_BYTES + B_BYTES;
    static constexpr int EPI_BYTES = 4 * 2 * 4096;          // per epilogue warp: two [32 rows][128 B] staging tiles
    static constexpr int BUDGET = 232448 - 1024 - 512 - EPI_BYTES;
    static constexpr int STAGES = (BUDGET / STAGE) > 8 ? 8 : (BUDGET / STAGE);
    static constexpr int TMEM_COLS = 2 * BN;                // two accumulator stages
    static constexpr int SMEM = STAGES * STAGE + EPI_BYTES + 512 + 1024;
};

template <int BN, int BKC>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads2, 1)
igemm_tma2_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  const __grid_constant__ CUtensorMap mapD, const Tma2Params p) {
    using C = Cfg2<BN, BKC>;
    extern __shared__ uint8_t smem_raw2[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw2) + 1023) & ~(uintptr_t)1023);
    uint8_t* epi = smem + C::STAGES * C::STAGE;                                  // 1024-aligned (STAGE is a multiple of 1024)
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi + C::EPI_BYTES);        // used in the leader only
    uint64_t* empty_bar = full_bar + C::STAGES;                                  // per CTA
    uint64_t* tfull_bar = empty_bar + C::STAGES;                                 // per CTA: accumulator stage ready
    uint64_t* tempty_bar = tfull_bar + 2;                                        // leader only: accumulator stage drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

    if (tid == 0) {
        for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < 2; ++a) { tc::mbar_init(&tfull_bar[a], 1); tc::mbar_init(&tempty_bar[a], 8); }
        tc::fence_barrier_init();
        prefetch_tmap2(&mapA); prefetch_tmap2(&mapB); prefetch_tmap2(&mapD);
    }
    if (warp == 1) tmem_alloc2<C::TMEM_COLS>(tmem_slot);
    tc::tc_fence_before();
    cluster_sync_all();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // issue loops: whole warp on uniform values, single-thread instructions under elect.sync (see conv_chain.cu)
    if (warp == 0) {
        // ------------------------------ TMA producer (both CTAs) ------------------------------
        uint32_t it = 0;
        const int kx_first = p.flip ? (p.kw - 1) * p.dil : 0, kx_last = p.flip ? 0 : (p.kw - 1) * p.dil;
        const int kstep = p.flip ? -p.dil : p.dil;
        for (int item = pair; item < p.items; item += num_pairs) {
            const int n_t = item % p.n_tiles; const int rest = item / p.n_tiles;
            const int sp = rest % p.splits;   const int m_t = rest / p.splits;
            int m0 = m_t * 256 + (int)rank * 128;
            if (m0 >= p.M) m0 = 0;            // fully out-of-range half: load valid rows, results are clipped / masked
            const int nrow0 = n_t * BN + (int)rank * (BN / 2);
            const int q = m0 % p.out_w; const int t = m0 / p.out_w;
            const int pp = t % p.out_h; const int bn = t / p.out_h;
            const int bw = q * p.stride + p.lower_w, bh = pp * p.stride + p.lower_h;
            const int kb0 = sp * p.kb_per_split, kb1 = min(p.num_kb, kb0 + p.kb_per_split);
            // tap / channel-chunk walk without per-block divisions
            const int tp0 = kb0 / p.cchunks; int cc = kb0 - tp0 * p.cchunks;
            int ky = tp0 / p.kw, kx = tp0 - ky * p.kw;
            if (p.flip) { ky = p.kh - 1 - ky; kx = p.kw - 1 - kx; }
            ky *= p.dil; kx *= p.dil;
            for (int kb = kb0; kb < kb1; ++kb, ++it) {
                const int s = it % C::STAGES; const uint32_t ph = (it / C::STAGES) & 1u;
                tc::mbar_wait(&empty_bar[s], ph ^ 1u);
                const uint32_t a_dst = tc::smem_u32(smem + s * C::STAGE);
                if (tc::elect_one()) {
                    if (rank == 0) mbar_expect_tx(&full_bar[s], 2u * (uint32_t)C::STAGE);
                    tma2_load_im2col(&mapA, a_dst, &full_bar[s], cc * BKC, bw, bh, bn, (uint16_t)kx, (uint16_t)ky);
                    tma2_load_2d(&mapB, a_dst + C::A_BYTES, &full_bar[s], kb * BKC, nrow0);
                }
                if (++cc == p.cchunks) {
                    cc = 0;
                    if (kx == kx_last) { kx = kx_first; ky += kstep; } else kx += kstep;
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0) {
            // ------------------------------ MMA issuer (leader CTA only) ------------------------------
            constexpr uint32_t idesc = tc::make_idesc_i8(256, BN, 0, 0);
            constexpr uint32_t lay = (BKC == 128) ? 2u : 4u;
            const uint64_t dconst = make_desc2(0u, 16, 8 * BKC, lay);
            uint32_t it = 0, tile_iter = 0;
            for (int item = pair; item < p.items; item += num_pairs, ++tile_iter) {
                const int rest = item / p.n_tiles; const int sp = rest % p.splits;
                const int kb0 = sp * p.kb_per_split, kb1 = min(p.num_kb, kb0 + p.kb_per_split);
                const uint32_t a = tile_iter & 1u, aph = (tile_iter >> 1) & 1u;
                tc::mbar_wait(&tempty_bar[a], aph ^ 1u);
                tc::tc_fence_after();
                const uint32_t d_tmem = tmem_base + a * BN;
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int s = it % C::STAGES; const uint32_t ph = (it / C::STAGES) & 1u;
                    tc::mbar_wait(&full_bar[s], ph);
                    tc::tc_fence_after();
                    const uint32_t a_base = tc::smem_u32(smem + s * C::STAGE);
                    const uint64_t ad = dconst | (uint64_t)(a_base >> 4);
                    const uint64_t bd = dconst | (uint64_t)((a_base + C::A_BYTES) >> 4);
                    if (tc::elect_one()) {
#pragma unroll
                        for (int kk = 0; kk < BKC / 32; ++kk)
                            umma2_i8(d_tmem, ad + 2 * kk, bd + 2 * kk, idesc, (uint32_t)((kb != kb0) || (kk != 0)));
                        umma2_commit_mc(&empty_bar[s], 3);       // frees this stage in both CTAs
                        if (kb == kb1 - 1) umma2_commit_mc(&tfull_bar[a], 3);   // accumulator stage complete, both CTAs
                    }
                }
            }
        }
    } else {
        // ------------------------------ epilogue (warps 2..5, both CTAs) ------------------------------
        const int quarter = warp & 3;                        // TMEM lane quarter this warp may read
        uint8_t* my_epi = epi + (warp - 2) * 8192;
        uint32_t mx = 0, tile_iter = 0, nstore = 0;
        const int etid = tid - 64;                           // 0..127 within the epilogue warps
        if (p.splits > 1) {
            // split-K: partial tiles are combined with TMA reduce-add, so D starts at zero.  All CTAs clear it
            // cooperatively while the producers are already filling the pipeline; one grid barrier orders the
            // clears before the first reduce-add anywhere.
            const size_t n16 = ((size_t)p.M * p.Ncols) >> 2;                 // Ncols % 4 == 0
            int4* z = reinterpret_cast<int4*>(p.acc);
            for (size_t i = (size_t)blockIdx.x * 128 + etid; i < n16; i += (size_t)gridDim.x * 128) z[i] = make_int4(0, 0, 0, 0);
            epi_bar_sync();
            if (etid == 0) grid_barrier(p.sync, gridDim.x);
            epi_bar_sync();
            fence_proxy_async_all();
        }
        for (int item = pair; item < p.items; item += num_pairs, ++tile_iter) {
            const int n_t = item % p.n_tiles; const int rest = item / p.n_tiles; const int m_t = rest / p.splits;
            const int row0 = m_t * 256 + (int)rank * 128 + quarter * 32;
            const int n0 = n_t * BN;
            const uint32_t a = tile_iter & 1u, aph = (tile_iter >> 1) & 1u;
            tc::mbar_wait(&tfull_bar[a], aph);
            tc::tc_fence_after();
            const bool rows_live = row0 < p.M;               // warp-uniform
            const bool row_ok = row0 + lane < p.M;
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                if (n0 + c0 >= p.Ncols || !rows_live) break; // warp-uniform
                uint32_t v[32];
                tc::tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + a * BN + (uint32_t)c0, v);
                tc::tmem_ld_wait();
                if (p.splits == 1 && row_ok) {
#pragma unroll
                    for (int q = 0; q < 32; ++q) mx = max(mx, niti_abs_u32((int32_t)v[q]));
                }
                uint8_t* buf = my_epi + (nstore & 1u) * 4096;
                if (lane == 0) bulk_wait_read<1>();          // the store that last read this buffer has drained it
                __syncwarp();
                const uint32_t rowaddr = tc::smem_u32(buf) + (uint32_t)lane * 128u;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const uint32_t dst = rowaddr + (uint32_t)((q ^ (lane & 7)) << 4);       // 128-byte swizzle
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(v[4 * q]), "r"(v[4 * q + 1]),
                                 "r"(v[4 * q + 2]), "r"(v[4 * q + 3]) : "memory");
                }
                tc::fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if (p.splits == 1) tma_store_2d(&mapD, tc::smem_u32(buf), n0 + c0, row0);
                    else tma_reduce_add_2d(&mapD, tc::smem_u32(buf), n0 + c0, row0);
                    bulk_commit();
                }
                ++nstore;
            }
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_remote(&tempty_bar[a], 0);     // this warp is done with accumulator stage a
        }
        if (lane == 0) bulk_wait<0>();
        if (p.splits > 1) {
            // all reduce-adds of this CTA have completed; after the second grid barrier D is final everywhere
            fence_proxy_async_all();
            epi_bar_sync();
            if (etid == 0) grid_barrier(p.sync, 2u * gridDim.x);
            epi_bar_sync();
            if (p.maxabs != nullptr) {
                const size_t n16 = ((size_t)p.M * p.Ncols) >> 2;
                const int4* z = reinterpret_cast<const int4*>(p.acc);
                for (size_t i = (size_t)blockIdx.x * 128 + etid; i < n16; i += (size_t)gridDim.x * 128) {
                    const int4 o = __ldcg(z + i);
                    mx = max(mx, max(max(niti_abs_u32(o.x), niti_abs_u32(o.y)), max(niti_abs_u32(o.z), niti_abs_u32(o.w))));
                }
            }
            if (etid == 0) {                                 // last CTA out re-arms the barrier for the next launch
                if (atomicAdd(p.sync + 1, 1u) == gridDim.x - 1) { p.sync[0] = 0u; p.sync[1] = 0u; __threadfence(); }
            }
        }
        if (p.maxabs != nullptr) {
            mx = warp_max_u32(mx);
            if (lane == 0 && mx) atomicMax(p.maxabs, mx);
        }
    }
    __syncwarp();
    tc::tc_fence_before();
    cluster_sync_all();
    if (warp == 1) tmem_dealloc2<C::TMEM_COLS>(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn2 g2_tiled = nullptr;
EncodeIm2colFn2 g2_im2col = nullptr;
int g2_driver_version = 0;
std::once_flag g2_once;

bool load_driver2() {
    std::call_once(g2_once, [] {
        void* f = nullptr; cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            g2_tiled = (EncodeTiledFn2)f;
        f = nullptr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            g2_im2col = (EncodeIm2colFn2)f;
        cudaDriverGetVersion(&g2_driver_version);
    });
    return g2_tiled && g2_im2col;
}
CUtensorMapSwizzle swz2(int bkc) { return bkc == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B; }

// row-major byte matrix [rows][cols]; box = [box_rows][box_cols]
bool make_tiled_u8(CUtensorMap* m, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows, uint32_t box_cols, int swz_bytes) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t es[2] = {1, 1};
    return g2_tiled(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    swz2(swz_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// row-major int32 matrix [rows][cols]; box = 32 rows x 32 columns (128-byte rows, 128B swizzle) — the epilogue's store unit
bool make_tiled_i32(CUtensorMap* m, void* base, uint64_t rows, uint64_t cols) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 4};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t es[2] = {1, 1};
    return g2_tiled(m, CU_TENSOR_MAP_DATA_TYPE_INT32, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// NHWC int8 tensor [n][h][w][c]; box = 128 pixels x bkc channels
bool make_im2col2(CUtensorMap* m, const void* base, int n, int h, int w, int c, int lower_h, int lower_w, int upper_h, int upper_w,
                  int stride, int bkc) {
    cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
    cuuint64_t strides[3] = {(cuuint64_t)c, (cuuint64_t)w * c, (cuuint64_t)h * w * c};
    int lower[2] = {lower_w, lower_h}, upper[2] = {upper_w, upper_h};
    cuuint32_t es[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    CUresult r = g2_im2col(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(base), dims, strides, lower, upper,
                           (cuuint32_t)bkc, 128u, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swz2(bkc),
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    // driver <= 13.1: im2col descriptors of tensors under 128 KiB carry a stray bit (same fix-up CUTLASS applies,
    // cute/atom/copy_traits_sm90_im2col.hpp)
    if (g2_driver_version <= 13010 && (size_t)n * h * w * c < 131072) reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
    return true;
}

template <int BN, int BKC>
int launch2(niti_ctx* ctx, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& d, const Tma2Params& p, int pairs, cudaStream_t st) {
    using C = Cfg2<BN, BKC>;
    if (int rc = niti_func_smem(ctx, (const void*)igemm_tma2_kernel<BN, BKC>, C::SMEM, "cudaFuncSetAttribute(igemm_tma2)")) return rc;
    igemm_tma2_kernel<BN, BKC><<<dim3(2 * pairs), kThreads2, C::SMEM, st>>>(a, b, d, p);
    return niti_check_launch(ctx, "igemm_tma2");
}

template <int BKC>
int run2(niti_ctx* ctx, const CUtensorMap& ma, const void* bmat, int32_t* acc, Tma2Params& p, int ktotal, cudaStream_t st) {
    const int bn = p.Ncols <= 64 ? 64 : (p.Ncols <= 128 ? 128 : 256);
    CUtensorMap mb, md;
    if (!make_tiled_u8(&mb, bmat, (uint64_t)p.Ncols, (uint64_t)ktotal, (uint32_t)(bn / 2), BKC, BKC))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(B) failed");
    if (!make_tiled_i32(&md, acc, (uint64_t)p.M, (uint64_t)p.Ncols))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(D) failed");
    p.m_tiles = (p.M + 255) / 256; p.n_tiles = (p.Ncols + bn - 1) / bn;
    const int max_pairs = ctx->sm_count / 2;
    const long tiles = (long)p.m_tiles * p.n_tiles;
    int splits = 1;
    if (tiles * 2 <= max_pairs) splits = (int)(max_pairs / tiles);           // fill the machine on small-M layers
    if (splits > p.num_kb / 2) splits = p.num_kb / 2;                         // at least two K-blocks per split
    if (splits < 1) splits = 1;
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    p.splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
    p.items = (int)(tiles * p.splits);
    p.acc = acc; p.sync = ctx->ws_u32 + 32;                // in-kernel clear / grid barrier / max for split-K
    const int pairs = p.items < max_pairs ? p.items : max_pairs;
    int rc;
    if (bn == 64) rc = launch2<64, BKC>(ctx, ma, mb, md, p, pairs, st);
    else if (bn == 128) rc = launch2<128, BKC>(ctx, ma, mb, md, p, pairs, st);
    else rc = launch2<256, BKC>(ctx, ma, mb, md, p, pairs, st);
    return rc;
}
} // namespace

bool tma2_fwd_eligible(const niti_conv_desc* d) {
    return d->c % 64 == 0 && d->oc % 4 == 0 && d->pad <= 127 && (d->kh - 1) * d->dil <= 255 && (d->kw - 1) * d->dil <= 255 && d->stride <= 8;
}
bool tma2_dx_eligible(const niti_conv_desc* d) {
    const int padp_h = (d->kh - 1) * d->dil - d->pad, padp_w = (d->kw - 1) * d->dil - d->pad;
    return d->stride == 1 && d->oc % 64 == 0 && d->c % 4 == 0 && padp_h >= 0 && padp_w >= 0 && padp_h <= 127 && padp_w <= 127 &&
           (d->kh - 1) * d->dil <= 255 && (d->kw - 1) * d->dil <= 255;
}

int tma2_conv_fwd_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, int32_t* acc,
                      uint32_t* maxabs, cudaStream_t st) {
    if (!load_driver2()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    const int bkc = d->c % 128 == 0 ? 128 : 64;
    Tma2Params p{};
    p.maxabs = maxabs; p.M = d->n * oh * ow; p.Ncols = d->oc;
    p.out_h = oh; p.out_w = ow; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = d->stride;
    p.lower_h = -d->pad; p.lower_w = -d->pad; p.flip = 0;
    p.cchunks = d->c / bkc; p.num_kb = d->kh * d->kw * p.cchunks;
    CUtensorMap ma;
    if (!make_im2col2(&ma, x, d->n, d->h, d->w, d->c, -d->pad, -d->pad, d->pad - (d->kh - 1) * d->dil, d->pad - (d->kw - 1) * d->dil, d->stride, bkc))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(x) failed");
    const int ktotal = d->kh * d->kw * d->c;
    return bkc == 128 ? run2<128>(ctx, ma, w, acc, p, ktotal, st) : run2<64>(ctx, ma, w, acc, p, ktotal, st);
}

// wt: [c][kh][kw][oc] (transposed weights, K-major over (tap, oc))
int tma2_conv_dx_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* wt, int32_t* acc,
                     uint32_t* maxabs, cudaStream_t st) {
    if (!load_driver2()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    const int bkc = d->oc % 128 == 0 ? 128 : 64;
    const int pph = (d->kh - 1) * d->dil - d->pad, ppw = (d->kw - 1) * d->dil - d->pad;   // padding of the equivalent forward conv
    Tma2Params p{};
    p.maxabs = maxabs; p.M = d->n * d->h * d->w; p.Ncols = d->c;
    p.out_h = d->h; p.out_w = d->w; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = 1;
    p.lower_h = -pph; p.lower_w = -ppw; p.flip = 1;
    p.cchunks = d->oc / bkc; p.num_kb = d->kh * d->kw * p.cchunks;
    CUtensorMap ma;
    if (!make_im2col2(&ma, dy, d->n, oh, ow, d->oc, -pph, -ppw, d->h - oh - pph, d->w - ow - ppw, 1, bkc))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(dy) failed");
    const int ktotal = d->kh * d->kw * d->oc;
    return bkc == 128 ? run2<128>(ctx, ma, wt, acc, p, ktotal, st) : run2<64>(ctx, ma, wt, acc, p, ktotal, st);
}

// =====================================================================================================================
// Single-phase fused op: contraction -> max|acc| -> shared exponent -> PSTO -> (ReLU | ReLU-grad gate) -> int8, ONE launch.
//
// The two-phase path writes every int32 accumulator to HBM/L2 and reads it back in a second kernel only because the
// exponent needs max|acc| of the whole tensor first.  When all output tiles of the op fit in tensor memory at once
// (512 columns x 128 lanes x 148 SMs = 9.7 M accumulators: every VGG-11 b256 layer but the stem) they can simply STAY
// there: each CTA pair keeps its <= 512/BN tiles resident, the epilogue warps take the maximum straight from TMEM, one
// grid barrier (plus the NVLink peer exchange when data-parallel, dp.cuh) publishes it, and the same warps rescale
// TMEM -> int8 -> global.  No int32 tensor exists, no second launch, 1/4 of the store bytes.
// 16 epilogue warps: the rescale arithmetic needs the issue slots of the whole SM (see conv_stem.cu).
// =====================================================================================================================
namespace {

constexpr int kFusedThreads = 576;   // warp 0: TMA producer, warp 1: MMA issuer, warps 2-17: epilogue
constexpr int kFusedMaxTiles = 8;

struct FusedEpi {
    int8_t* y; const int8_t* gate;
    const int8_t* exp_a; const int8_t* exp_b; int8_t* exp_out;
    int relu;
    uint32_t* sync;                  // {arrival counter, exit counter}
    DpView dp; int dp_slot;
    unsigned long long* trace;
    int trigger_late;                // PDL: let the next launch start placing CTAs at the exponent barrier, not at entry
};

template <int BN, int BKC> struct CfgF {
    static constexpr int A_BYTES = 128 * BKC;
    static constexpr int B_BYTES = (BN / 2) * BKC;
    static constexpr int STAGE = A_BYTES + B_BYTES;
    static constexpr int BUDGET = 232448 - 1024 - 512;
    static constexpr int STAGES = (BUDGET / STAGE) > 8 ? 8 : (BUDGET / STAGE);
    static constexpr int SMEM = STAGES * STAGE + 512 + 1024;
};

__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}

template <int BN, int BKC>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kFusedThreads, 1)
igemm_fused_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const Tma2Params p,
                   const FusedEpi e) {
    using C = CfgF<BN, BKC>;
    extern __shared__ uint8_t smem_raw_f[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw_f) + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE);       // used in the leader only
    uint64_t* empty_bar = full_bar + C::STAGES;                                          // per CTA
    uint64_t* tfull_bar = empty_bar + C::STAGES;                                         // per CTA, one per resident tile
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull_bar + kFusedMaxTiles);
    uint32_t* bcast = tmem_slot + 1;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
    if (!e.trigger_late) niti_pdl_trigger();                 // the next launch of the stream may start placing its CTAs
    if (tid == 64) niti_trace(e.trace, 0);

    if (tid == 0) {
        for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < kFusedMaxTiles; ++a) tc::mbar_init(&tfull_bar[a], 1);
        tc::fence_barrier_init();
        prefetch_tmap2(&mapA); prefetch_tmap2(&mapB);
    }
    if (warp == 1) tmem_alloc2<512>(tmem_slot);
    tc::tc_fence_before();
    cluster_sync_all();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    niti_pdl_wait();                                         // prologue done; operands of the previous launch are now visible

    if (warp == 0) {
        if (lane == 0) {
            // ------------------------------ TMA producer (both CTAs) ------------------------------
            uint32_t it = 0;
            for (int item = pair; item < p.items; item += num_pairs) {
                const int n_t = item % p.n_tiles; const int m_t = item / p.n_tiles;
                int m0 = m_t * 256 + (int)rank * 128;
                if (m0 >= p.M) m0 = 0;
                const int nrow0 = n_t * BN + (int)rank * (BN / 2);
                const int q = m0 % p.out_w; const int t = m0 / p.out_w;
                const int pp = t % p.out_h; const int bn = t / p.out_h;
                const int bw = q * p.stride + p.lower_w, bh = pp * p.stride + p.lower_h;
                for (int kb = 0; kb < p.num_kb; ++kb, ++it) {
                    const int s = it % C::STAGES; const uint32_t ph = (it / C::STAGES) & 1u;
                    tc::mbar_wait(&empty_bar[s], ph ^ 1u);
                    const uint32_t a_dst = tc::smem_u32(smem + s * C::STAGE);
                    const uint32_t b_dst = a_dst + C::A_BYTES;
                    if (rank == 0) mbar_expect_tx(&full_bar[s], 2u * (uint32_t)C::STAGE);
                    const int tp = kb / p.cchunks, cc = kb - tp * p.cchunks;
                    int ky = tp / p.kw, kx = tp - ky * p.kw;
                    if (p.flip) { ky = p.kh - 1 - ky; kx = p.kw - 1 - kx; }
                    tma2_load_im2col(&mapA, a_dst, &full_bar[s], cc * BKC, bw, bh, bn, (uint16_t)(kx * p.dil), (uint16_t)(ky * p.dil));
                    tma2_load_2d(&mapB, b_dst, &full_bar[s], kb * BKC, nrow0);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            // ------------------------------ MMA issuer (leader CTA only) ------------------------------
            constexpr uint32_t idesc = tc::make_idesc_i8(256, BN, 0, 0);
            constexpr uint32_t lay = (BKC == 128) ? 2u : 4u;
            uint32_t it = 0, tile_iter = 0;
            for (int item = pair; item < p.items; item += num_pairs, ++tile_iter) {
                const uint32_t d_tmem = tmem_base + tile_iter * BN;          // every tile keeps its own columns
                for (int kb = 0; kb < p.num_kb; ++kb, ++it) {
                    const int s = it % C::STAGES; const uint32_t ph = (it / C::STAGES) & 1u;
                    tc::mbar_wait(&full_bar[s], ph);
                    tc::tc_fence_after();
                    const uint32_t a_base = tc::smem_u32(smem + s * C::STAGE);
                    const uint32_t b_base = a_base + C::A_BYTES;
#pragma unroll
                    for (int kk = 0; kk < BKC / 32; ++kk) {
                        const uint64_t ad = make_desc2(a_base + kk * 32, 16, 8 * BKC, lay);
                        const uint64_t bd = make_desc2(b_base + kk * 32, 16, 8 * BKC, lay);
                        umma2_i8(d_tmem, ad, bd, idesc, (uint32_t)((kb != 0) || (kk != 0)));
                    }
                    umma2_commit_mc(&empty_bar[s], 3);
                }
                umma2_commit_mc(&tfull_bar[tile_iter], 3);
            }
        }
    } else {
        // ------------------------------ epilogue (warps 2..17, both CTAs) ------------------------------
        const int quarter = warp & 3;                        // TMEM lane quarter this warp may read
        const int group = (warp - 2) >> 2;                   // four warps share a quarter: every 4th 16-column chunk each
        const int etid = tid - 64;
        int shift = 0, kind = 0;
        uint32_t mx = 0;
        if (etid == 0) niti_trace(e.trace, 1);
        for (int pass = 0; pass < 2; ++pass) {
            uint32_t tile_iter = 0;
            for (int item = pair; item < p.items; item += num_pairs, ++tile_iter) {
                const int n_t = item % p.n_tiles; const int m_t = item / p.n_tiles;
                const int row0 = m_t * 256 + (int)rank * 128 + quarter * 32;
                const int row = row0 + lane;
                const int n0 = n_t * BN;
                if (pass == 0) { tc::mbar_wait(&tfull_bar[tile_iter], 0); tc::tc_fence_after(); if (etid == 0 && tile_iter == 0) niti_trace(e.trace, 2); }
                if (row0 >= p.M) continue;                   // warp-uniform
                const bool row_ok = row < p.M;
#pragma unroll 1
                for (int c0 = group * 16; c0 < BN; c0 += 64) {
                    if (n0 + c0 >= p.Ncols) break;           // warp-uniform (Ncols % 16 == 0)
                    uint32_t v[16];
                    tmem_ld_x16(tmem_base + ((uint32_t)(quarter * 32) << 16) + tile_iter * BN + (uint32_t)c0, v);
                    tc::tmem_ld_wait();
                    if (pass == 0) {
                        if (row_ok) {
#pragma unroll
                            for (int q = 0; q < 16; ++q) mx = max(mx, niti_abs_u32((int32_t)v[q]));
                        }
                    } else if (row_ok) {
                        uint32_t o[4];
                        niti_rescale16(v, kind, shift, o);
                        const size_t off = (size_t)row * p.Ncols + n0 + c0;
                        if (e.relu) { o[0] = __vmaxs4(o[0], 0u); o[1] = __vmaxs4(o[1], 0u); o[2] = __vmaxs4(o[2], 0u); o[3] = __vmaxs4(o[3], 0u); }
                        if (e.gate != nullptr) {             // NITI_CPUReluGrad: pass where the forward value was > 0
                            const uint4 g = *reinterpret_cast<const uint4*>(e.gate + off);
                            o[0] &= __vcmpgts4(g.x, 0u); o[1] &= __vcmpgts4(g.y, 0u); o[2] &= __vcmpgts4(g.z, 0u); o[3] &= __vcmpgts4(g.w, 0u);
                        }
                        *reinterpret_cast<uint4*>(e.y + off) = make_uint4(o[0], o[1], o[2], o[3]);
                    }
                }
            }
            if (pass == 0) {
                // ---- global max|acc|: warp max -> atomicMax -> grid barrier -> (data parallel: peer exchange) ----
                mx = warp_max_u32(mx);
                if (lane == 0 && mx) atomicMax(p.maxabs, mx);
                asm volatile("bar.sync 1, 512;" ::: "memory");
                if (e.trigger_late) niti_pdl_trigger();
                if (etid == 0) {
                    niti_trace(e.trace, 3);
                    grid_barrier(e.sync, gridDim.x);
                    niti_trace(e.trace, 4);
                    const uint32_t m = dp_exchange_max(e.dp, e.dp_slot, __ldcg(p.maxabs), blockIdx.x == 0);
                    *bcast = m;
                }
                asm volatile("bar.sync 1, 512;" ::: "memory");
                const int b = niti_bitwidth(*bcast);
                const int s = b - 7;
                if (s > 1) { shift = s; kind = 0; } else if (s == 1) { shift = 2; kind = 0; } else { shift = 0; kind = 1; }
                if (blockIdx.x == 0 && etid == 0 && e.exp_out != nullptr)
                    *e.exp_out = (int8_t)((e.exp_a ? (int)*e.exp_a : 0) + (e.exp_b ? (int)*e.exp_b : 0) + (kind == 1 ? 0 : shift));
            }
        }
        asm volatile("bar.sync 1, 512;" ::: "memory");
        if (etid == 0) {                                     // last CTA out re-arms the grid barrier for the next launch
            niti_trace(e.trace, 5);
            __threadfence();
            if (atomicAdd(e.sync + 1, 1u) == gridDim.x - 1) { e.sync[0] = 0u; e.sync[1] = 0u; __threadfence(); }
        }
    }
    __syncwarp();
    tc::tc_fence_before();
    cluster_sync_all();
    if (warp == 1) tmem_dealloc2<512>(tmem_base);
    if (tid == 64) niti_trace(e.trace, 6);
}

// tile shape + residency plan: the largest BN that still gives >= ~2/3 of the CTA pairs a tile; every pair must be able
// to keep all its tiles in its 512 TMEM columns
bool fused_plan(const niti_ctx* ctx, int M, int Ncols, int* bn_out, int* pairs_out, int* items_out) {
    const int max_pairs = ctx->sm_count / 2;
    const long m_tiles = (M + 255) / 256;
    int bn = 64;
    for (int cand : {256, 128}) {
        if (Ncols > cand / 2 && m_tiles * ((Ncols + cand - 1) / cand) >= (2 * max_pairs) / 3) { bn = cand; break; }
    }
    if (Ncols <= 64) bn = 64; else if (Ncols <= 128 && bn > 128) bn = 128;
    const long items = m_tiles * ((Ncols + bn - 1) / bn);
    const int pairs = (int)(items < max_pairs ? items : max_pairs);
    const long per_pair = (items + pairs - 1) / pairs;
    if (per_pair * bn > 512 || per_pair > kFusedMaxTiles) return false;
    *bn_out = bn; *pairs_out = pairs; *items_out = (int)items;
    return true;
}

template <int BN, int BKC>
int launch_fused(niti_ctx* ctx, const CUtensorMap& a, const CUtensorMap& b, const Tma2Params& p, const FusedEpi& e, int pairs,
                 cudaStream_t st) {
    using C = CfgF<BN, BKC>;
    // One CTA per SM, enforced through the shared-memory request: every CTA allocates all 512 TMEM columns and then
    // waits at a grid barrier, so two CTAs of this launch sharing an SM (possible for the BKC = 64 configurations, < 114 KB,
    // once programmatic dependent launch places CTAs on partly occupied SMs) would deadlock — the second blocks in
    // tcgen05.alloc while the first waits for it at the barrier.
    constexpr size_t kSmemReq = C::SMEM > 120 * 1024 ? (size_t)C::SMEM : (size_t)120 * 1024;
    if (int rc = niti_func_smem(ctx, (const void*)igemm_fused_kernel<BN, BKC>, (int)kSmemReq, "cudaFuncSetAttribute(igemm_fused)")) return rc;
    cudaError_t lerr = niti_launch(ctx, NITI_PDL_FUSED, igemm_fused_kernel<BN, BKC>, dim3(2 * pairs), dim3(kFusedThreads), kSmemReq, st, a, b, p, e);
    if (lerr != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch igemm_fused", lerr);
    return niti_check_launch(ctx, "igemm_fused");
}

template <int BKC>
int run_fused(niti_ctx* ctx, const CUtensorMap& ma, const void* bmat, Tma2Params& p, FusedEpi& e, int ktotal, cudaStream_t st) {
    int bn, pairs, items;
    if (!fused_plan(ctx, p.M, p.Ncols, &bn, &pairs, &items)) return niti_fail(ctx, NITI_ERR_UNSUPPORTED, "fused op: tiles do not fit in TMEM");
    CUtensorMap mb;
    if (!make_tiled_u8(&mb, bmat, (uint64_t)p.Ncols, (uint64_t)ktotal, (uint32_t)(bn / 2), BKC, BKC))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(B) failed");
    p.m_tiles = (p.M + 255) / 256; p.n_tiles = (p.Ncols + bn - 1) / bn;
    p.splits = 1; p.kb_per_split = p.num_kb; p.items = items;
    e.sync = ctx->ws_u32 + 52;
    if (bn == 64) return launch_fused<64, BKC>(ctx, ma, mb, p, e, pairs, st);
    if (bn == 128) return launch_fused<128, BKC>(ctx, ma, mb, p, e, pairs, st);
    return launch_fused<256, BKC>(ctx, ma, mb, p, e, pairs, st);
}
} // namespace

bool fused_fwd_eligible(const niti_ctx* ctx, const niti_conv_desc* d) {
    int a, b, c;
    return tma2_fwd_eligible(d) && d->oc % 16 == 0 && fused_plan(ctx, d->n * conv_out_h(d) * conv_out_w(d), d->oc, &a, &b, &c);
}
bool fused_dx_eligible(const niti_ctx* ctx, const niti_conv_desc* d) {
    int a, b, c;
    return tma2_dx_eligible(d) && d->c % 16 == 0 && fused_plan(ctx, d->n * d->h * d->w, d->c, &a, &b, &c);
}

// y = ruleA(conv(x, w)) (+ ReLU), exponent out; maxabs: zeroed slot
int fused_conv_fwd(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, const int8_t* ex, const int8_t* ew,
                   int8_t* y, int8_t* ey, int relu, uint32_t* maxabs, int dp_slot, cudaStream_t st) {
    if (!load_driver2()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    const int bkc = d->c % 128 == 0 ? 128 : 64;
    Tma2Params p{};
    p.maxabs = maxabs; p.M = d->n * oh * ow; p.Ncols = d->oc;
    p.out_h = oh; p.out_w = ow; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = d->stride;
    p.lower_h = -d->pad; p.lower_w = -d->pad; p.flip = 0;
    p.cchunks = d->c / bkc; p.num_kb = d->kh * d->kw * p.cchunks;
    FusedEpi e{};
    e.y = y; e.gate = nullptr; e.exp_a = ex; e.exp_b = ew; e.exp_out = ey; e.relu = relu; e.dp = ctx->dp; e.dp_slot = dp_slot; e.trace = ctx->trace; e.trigger_late = (ctx->pdl & 512) ? 1 : 0;
    CUtensorMap ma;
    if (!make_im2col2(&ma, x, d->n, d->h, d->w, d->c, -d->pad, -d->pad, d->pad - (d->kh - 1) * d->dil, d->pad - (d->kw - 1) * d->dil, d->stride, bkc))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(x) failed");
    const int ktotal = d->kh * d->kw * d->c;
    return bkc == 128 ? run_fused<128>(ctx, ma, w, p, e, ktotal, st) : run_fused<64>(ctx, ma, w, p, e, ktotal, st);
}

// dx = ruleA(dX(dy, wt)) (+ ReLU-grad gate against `gate`, the forward tensor dx is the gradient of); wt = [c][kh][kw][oc]
int fused_conv_dx(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* wt, int8_t* dx, const int8_t* gate,
                  uint32_t* maxabs, int dp_slot, cudaStream_t st) {
    if (!load_driver2()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    const int bkc = d->oc % 128 == 0 ? 128 : 64;
    const int pph = (d->kh - 1) * d->dil - d->pad, ppw = (d->kw - 1) * d->dil - d->pad;
    Tma2Params p{};
    p.maxabs = maxabs; p.M = d->n * d->h * d->w; p.Ncols = d->c;
    p.out_h = d->h; p.out_w = d->w; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = 1;
    p.lower_h = -pph; p.lower_w = -ppw; p.flip = 1;
    p.cchunks = d->oc / bkc; p.num_kb = d->kh * d->kw * p.cchunks;
    FusedEpi e{};
    e.y = dx; e.gate = gate; e.relu = 0; e.dp = ctx->dp; e.dp_slot = dp_slot; e.trace = ctx->trace; e.trigger_late = (ctx->pdl & 512) ? 1 : 0;
    CUtensorMap ma;
    if (!make_im2col2(&ma, dy, d->n, oh, ow, d->oc, -pph, -ppw, d->h - oh - pph, d->w - ow - ppw, 1, bkc))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(dy) failed");
    const int ktotal = d->kh * d->kw * d->oc;
    return bkc == 128 ? run_fused<128>(ctx, ma, wt, p, e, ktotal, st) : run_fused<64>(ctx, ma, wt, p, e, ktotal, st);
}
