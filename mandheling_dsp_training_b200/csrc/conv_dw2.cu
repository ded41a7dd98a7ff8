// conv_dw2.cu — second-generation weight-gradient contraction (dW) on tcgen05.
//
//   dW[oc][tap][c] = sum over output pixels of dY[pixel][oc] * X[pixel shifted by tap][c]
// (NITI_GradientConv_Int8.cpp:218-268 runs this as a convolution whose "kernel" is dY; grad wiring
//  tools/train/source/grad/NITI_Conv_Int8_Grad.cpp:134-191).  The reduction runs over N*OH*OW pixels, the output is tiny,
// so the kernel is split-K by construction.  Compared with conv_tma.cu MODE 1 this generation
//   * groups the taps of one filter row: the dY tile (A operand, MN-major) is loaded once per K-block and multiplied
//     with G shifted im2col tiles of X into G accumulators in TMEM (G x BKC columns) -> 1/3 fewer operand bytes at 3x3;
//   * is persistent (grid <= #SMs, static item schedule), zeroes the output in-kernel behind one grid barrier, and
//     combines split-K partials with TMA reduce-add (cp.reduce.async.bulk.tensor .add, 128-byte rows) instead of one
//     red.global per element — integer adds commute, so the result is bit-exact in any order;
//   * drains TMEM through swizzled shared memory + TMA like conv_tma2.cu.
#include "tma_common.cuh"

namespace {

constexpr int kDwThreads = 192;      // warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-5: epilogue

struct Dw2Params {
    int oc, c, kh, kw, dil, stride;
    int out_h, out_w, lower_h, lower_w;
    int npix, num_kb, kb_per_split, splits;
    int oc_tiles, c_tiles, ngroups;   // tap groups (G consecutive taps in row-major tap order)
    int items;
    int rows_mode;                    // BKC == 16 only: X tiles loaded row-wise (tmax::make_rows16)
    int accumulate;                   // acc += dW into a caller-zeroed buffer: no in-kernel clear, no grid barrier
    int32_t* acc; size_t nout;
    uint32_t* sync;                   // {arrival counter, exit counter}
    unsigned long long* tl;
};

template <int BKC, int G> struct DwCfg {
    static constexpr int A_BYTES = 128 * 128;                 // [128 pixels][128 oc]
    static constexpr int B_BYTES = 128 * BKC;                 // [128 pixels][BKC c] per tap
    static constexpr int STAGE = A_BYTES + G * B_BYTES;
    static constexpr int EPI_BYTES = 4 * 2 * 4096;
    static constexpr int BUDGET = 232448 - 1024 - 512 - EPI_BYTES;
    static constexpr int STAGES = (BUDGET / STAGE) > 6 ? 6 : (BUDGET / STAGE);
    static constexpr int ACC_COLS = G * BKC;
    static constexpr int TMEM_COLS = ACC_COLS <= 32 ? 32 : (ACC_COLS <= 64 ? 64 : (ACC_COLS <= 128 ? 128 : (ACC_COLS <= 256 ? 256 : 512)));
    static constexpr int SMEM = STAGES * STAGE + EPI_BYTES + 512 + 1024;
    static_assert(ACC_COLS <= 512, "accumulators exceed TMEM");
    static_assert(STAGES >= 2, "need at least a double buffer");
};

struct DwItem { int ct, tap0, ng, ot, kb0, kb1; };

template <int G>
__device__ __forceinline__ DwItem decode_item(const Dw2Params& p, int item) {
    DwItem it;
    it.ct = item % p.c_tiles; int r = item / p.c_tiles;
    const int grp = r % p.ngroups; r /= p.ngroups;
    it.ot = r % p.oc_tiles; const int sp = r / p.oc_tiles;
    it.tap0 = grp * G;
    it.ng = min(G, p.kh * p.kw - it.tap0);
    it.kb0 = sp * p.kb_per_split; it.kb1 = min(p.num_kb, it.kb0 + p.kb_per_split);
    return it;
}

template <int BKC, int G>
__global__ void __launch_bounds__(kDwThreads, 1)
dw2_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
           const __grid_constant__ CUtensorMap mapD, const Dw2Params p) {
    using C = DwCfg<BKC, G>;
    extern __shared__ uint8_t smem_raw_dw2[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw_dw2) + 1023) & ~(uintptr_t)1023);
    uint8_t* epi = smem + C::STAGES * C::STAGE;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi + C::EPI_BYTES);
    uint64_t* empty_bar = full_bar + C::STAGES;
    uint64_t* tfull_bar = empty_bar + C::STAGES;
    uint64_t* tempty_bar = tfull_bar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    niti_tl_enter(p.tl);

    if (tid == 0) {
        for (int s = 0; s < C::STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        tc::mbar_init(tfull_bar, 1); tc::mbar_init(tempty_bar, 4);
        tc::fence_barrier_init();
        tmax::prefetch_map(&mapA); tmax::prefetch_map(&mapB); tmax::prefetch_map(&mapD);
    }
    if (warp == 1) tc::tmem_alloc<C::TMEM_COLS>(tmem_slot);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // Both issuing roles run with the whole warp on warp-uniform values; only the issue itself sits under elect.sync, so the
    // compiler feeds UTMALDG / UTCIMMA from uniform registers instead of wrapping each one in an ELECT / R2UR.BROADCAST loop
    // (12 MMAs + 4 loads per K block here; see conv_chain.cu).
    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        uint32_t n = 0;
        for (int item = blockIdx.x; item < p.items; item += gridDim.x) {
            const DwItem it = decode_item<G>(p, item);
            const int ky0 = it.tap0 / p.kw, kx0 = it.tap0 - ky0 * p.kw;
            for (int kb = it.kb0; kb < it.kb1; ++kb, ++n) {
                const int s = n % C::STAGES; const uint32_t ph = (n / C::STAGES) & 1u;
                tc::mbar_wait(&empty_bar[s], ph ^ 1u);
                const uint32_t a_dst = tc::smem_u32(smem + s * C::STAGE);
                const int pix0 = kb * 128;
                const int q = pix0 % p.out_w; const int t = pix0 / p.out_w;
                const int pp = t % p.out_h; const int nn = t / p.out_h;
                if (tc::elect_one()) {
                    tmax::mbar_expect_tx(&full_bar[s], (uint32_t)(C::A_BYTES + it.ng * C::B_BYTES));
                    tmax::load_2d(&mapA, a_dst, &full_bar[s], it.ot * 128, pix0);
                    int kx = kx0, ky = ky0;
                    for (int g = 0; g < it.ng; ++g) {
                        if (BKC == 16 && p.rows_mode)
                            tmax::load_3d(&mapB, a_dst + C::A_BYTES + g * C::B_BYTES, &full_bar[s], (kx + p.lower_w) * 2,
                                          pp + ky + p.lower_h, nn);
                        else
                            tmax::load_im2col(&mapB, a_dst + C::A_BYTES + g * C::B_BYTES, &full_bar[s], it.ct * BKC,
                                              q * p.stride + p.lower_w, pp * p.stride + p.lower_h, nn,
                                              (uint16_t)(kx * p.dil), (uint16_t)(ky * p.dil));
                        if (++kx == p.kw) { kx = 0; ++ky; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer ------------------------------
        constexpr uint32_t idesc = tc::make_idesc_i8(128, BKC, 1, 1);
        constexpr uint32_t lay = (BKC == 128) ? 2u : (BKC == 64 ? 4u : 0u);
        // shared-memory descriptors = constant | (address >> 4)
        const uint64_t adc = tmax::make_desc(0u, 128 * 128, 1024, 2u);
        const uint64_t bdc = BKC == 16 ? tmax::make_desc(0u, 128, 2048, 0u) : tmax::make_desc(0u, 128 * BKC, 8 * BKC, lay);
        uint32_t n = 0, tile_iter = 0;
        for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++tile_iter) {
            const DwItem it = decode_item<G>(p, item);
            tc::mbar_wait(tempty_bar, (tile_iter & 1u) ^ 1u);          // epilogue drained the previous item
            tc::tc_fence_after();
            const uint32_t idesc_n = tc::make_idesc_i8(128, it.ng * 16, 1, 1);
            for (int kb = it.kb0; kb < it.kb1; ++kb, ++n) {
                const int s = n % C::STAGES; const uint32_t ph = (n / C::STAGES) & 1u;
                tc::mbar_wait(&full_bar[s], ph);
                tc::tc_fence_after();
                const uint32_t a_base = tc::smem_u32(smem + s * C::STAGE);
                const uint64_t ad0 = adc | (uint64_t)(a_base >> 4);
                const uint64_t bd0 = bdc | (uint64_t)((a_base + C::A_BYTES) >> 4);
                const uint32_t acc_on = (uint32_t)(kb != it.kb0);
                if (tc::elect_one()) {
                    if (BKC == 16) {
                        // stem layers (<= 16 input channels): the ng tap tiles [128 pixels][16 B] sit back to back, i.e. one
                        // MN-major un-swizzled B operand of N = ng*16 columns: 16-byte column chunks SBO = 2048 B apart,
                        // 8-pixel K groups LBO = 128 B apart -> ONE MMA per K32 covers every tap of the group
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk)
                            tc::umma_i8(tmem_base, ad0 + (uint64_t)(kk * ((32 * 128) >> 4)), bd0 + (uint64_t)(kk * ((32 * 16) >> 4)), idesc_n,
                                        kk == 0 ? acc_on : 1u);
                    } else {
                        for (int g = 0; g < it.ng; ++g) {
                            const uint64_t bdg = bd0 + (uint64_t)(g * (C::B_BYTES >> 4));
#pragma unroll
                            for (int kk = 0; kk < 4; ++kk)                     // 128 pixels = 4 x K32
                                tc::umma_i8(tmem_base + g * BKC, ad0 + (uint64_t)(kk * ((32 * 128) >> 4)), bdg + (uint64_t)(kk * ((32 * BKC) >> 4)),
                                            idesc, kk == 0 ? acc_on : 1u);
                        }
                    }
                    tc::umma_commit(&empty_bar[s]);
                    if (kb == it.kb1 - 1) tc::umma_commit(tfull_bar);
                }
            }
        }
    } else {
        // ------------------------------ epilogue (warps 2..5) ------------------------------
        const int quarter = warp & 3;
        uint8_t* my_epi = epi + (warp - 2) * 8192;
        const int etid = tid - 64;
        uint32_t tile_iter = 0, nstore = 0;
        if (p.splits > 1 && !p.accumulate) {
            const size_t n16 = p.nout >> 2;
            int4* z = reinterpret_cast<int4*>(p.acc);
            for (size_t i = (size_t)blockIdx.x * 128 + etid; i < n16; i += (size_t)gridDim.x * 128) z[i] = make_int4(0, 0, 0, 0);
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (etid == 0) tmax::grid_barrier(p.sync, gridDim.x);
            asm volatile("bar.sync 1, 128;" ::: "memory");
            tmax::fence_proxy_async_all();
        }
        for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++tile_iter) {
            const DwItem it = decode_item<G>(p, item);
            const int row0 = it.ot * 128 + quarter * 32;
            tc::mbar_wait(tfull_bar, tile_iter & 1u);
            tc::tc_fence_after();
            if (row0 < p.oc) {                                   // warp-uniform
                // BKC == 16 (c == 16): the group's ng*16 accumulator columns are contiguous in dW as well, drained in 32-column
                // boxes; an odd ng only occurs in the last group, whose overhang lies outside the tensor map and is clipped
                const int ngrp = BKC == 16 ? 1 : it.ng;
                const int ncols = BKC == 16 ? it.ng * 16 : BKC;
                for (int g = 0; g < ngrp; ++g) {
                    const int colbase = ((it.tap0 + g) * p.c) + it.ct * BKC;
#pragma unroll 1
                    for (int c0 = 0; c0 < ncols; c0 += 32) {
                        uint32_t v[32];
                        tc::tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(g * BKC + c0), v);
                        tc::tmem_ld_wait();
                        uint8_t* buf = my_epi + (nstore & 1u) * 4096;
                        if (lane == 0) tmax::bulk_wait_read<1>();
                        __syncwarp();
                        const uint32_t rowaddr = tc::smem_u32(buf) + (uint32_t)lane * 128u;
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const uint32_t dst = rowaddr + (uint32_t)((q ^ (lane & 7)) << 4);
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(v[4 * q]), "r"(v[4 * q + 1]),
                                         "r"(v[4 * q + 2]), "r"(v[4 * q + 3]) : "memory");
                        }
                        tc::fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            if (p.splits == 1 && !p.accumulate) tmax::store_2d(&mapD, tc::smem_u32(buf), colbase + c0, row0);
                            else tmax::reduce_add_2d(&mapD, tc::smem_u32(buf), colbase + c0, row0);
                            tmax::bulk_commit();
                        }
                        ++nstore;
                    }
                }
            }
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(tempty_bar);
        }
        if (lane == 0) tmax::bulk_wait<0>();
        if (p.splits > 1 && !p.accumulate) {
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (etid == 0) {                                     // last CTA out re-arms the barrier for the next launch
                __threadfence();
                if (atomicAdd(p.sync + 1, 1u) == gridDim.x - 1) { p.sync[0] = 0u; p.sync[1] = 0u; __threadfence(); }
            }
        }
    }
    __syncwarp();
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc<C::TMEM_COLS>(tmem_base);
    niti_tl_exit(p.tl);
}

} // namespace

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
namespace {
template <int BKC, int G>
int launch_dw2(niti_ctx* ctx, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& d, const Dw2Params& p, int grid,
               cudaStream_t st) {
    using C = DwCfg<BKC, G>;
    if (int rc_cfg = niti_func_smem(ctx, (const void*)dw2_kernel<BKC, G>, C::SMEM, "cudaFuncSetAttribute(dw2)")) return rc_cfg;
    dw2_kernel<BKC, G><<<grid, kDwThreads, C::SMEM, st>>>(a, b, d, p);
    return niti_check_launch(ctx, "dw2");
}
} // namespace

bool dw2_eligible(const niti_conv_desc* d) {
    return d->oc % 16 == 0 && (d->c % 64 == 0 || d->c == 16) && d->pad <= 127 && (d->kh - 1) * d->dil <= 255 && (d->kw - 1) * d->dil <= 255 &&
           d->stride <= 8;
}

int dw2_conv_dw_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* dy, int32_t* acc,
                    uint32_t* maxabs, cudaStream_t st, int accumulate) {
    if (!tmax::driver_ok()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    const int bkc = d->c == 16 ? 16 : (d->c % 128 == 0 ? 128 : 64);
    const int spread = (ctx->dw_ctas > 0 && ctx->dw_ctas < ctx->sm_count) ? ctx->dw_ctas : ctx->sm_count;
    // Taps per item.  G = 3 (a filter row shares one dY tile: 1/3 fewer operand bytes per MAC) makes few, wide tiles, so the
    // split-K spread over the SMs is deep and every split adds a 128 x 3*BKC int32 tile to L2 (28 MB of reduce-add per VGG
    // launch, which set the 12-17 us of those launches); G = 1 makes 3x the tiles, 1/3 the reduce-add traffic and longer K
    // loops per CTA.  Pick by a two-term estimate: operand bytes one CTA pulls through its L2 port (~100 GB/s) + reduce-add
    // bytes of the launch (~3.5 TB/s aggregate).  NITI_B200_DW_G forces 1 or 3.
    int G = bkc == 16 ? 16 : (d->kw == 3 ? 3 : 1);
    if (G == 3) {
        const long num_kb = ((long)d->n * oh * ow + 127) / 128;
        double est[2];
        for (int i = 0; i < 2; ++i) {
            const int g = i == 0 ? 3 : 1;
            const long tiles = (long)((d->oc + 127) / 128) * ((d->kh * d->kw + g - 1) / g) * (d->c / bkc);
            long splits = spread / tiles; if (splits > num_kb) splits = num_kb; if (splits < 1) splits = 1;
            const long kb_per = (num_kb + splits - 1) / splits;
            const long waves = (tiles * splits + spread - 1) / spread;
            est[i] = (double)waves * kb_per * (128.0 * 128 + g * 128.0 * bkc) / 100e3 +            // us at 100 GB/s per CTA
                     (double)tiles * splits * (128.0 * g * bkc * 4) / 3.5e6;                      // us at 3.5 TB/s
        }
        if (est[1] < est[0]) G = 1;
        if (ctx->dw_g == 1 || ctx->dw_g == 3) G = ctx->dw_g;
    }
    Dw2Params p{};
    p.oc = d->oc; p.c = d->c; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = d->stride;
    p.out_h = oh; p.out_w = ow; p.lower_h = -d->pad; p.lower_w = -d->pad;
    p.npix = d->n * oh * ow; p.num_kb = (p.npix + 127) / 128;
    p.oc_tiles = (d->oc + 127) / 128; p.c_tiles = d->c / bkc; p.ngroups = (d->kh * d->kw + G - 1) / G;
    const long tiles = (long)p.oc_tiles * p.ngroups * p.c_tiles;
    int splits = (int)(spread / tiles);
    if (splits > p.num_kb) splits = p.num_kb;
    if (splits < 1) splits = 1;
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    p.splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
    p.items = (int)(tiles * p.splits);
    p.acc = acc; p.nout = (size_t)d->oc * d->kh * d->kw * d->c; p.sync = ctx->ws_u32 + 44;
    p.accumulate = accumulate; p.tl = niti_tl_next(ctx, "dw2");
    CUtensorMap ma, mb, md;
    if (!tmax::make_tiled_u8(&ma, dy, (uint64_t)p.npix, (uint64_t)d->oc, 128, 128, 128))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(dy) failed");
    p.rows_mode = (bkc == 16 && tmax::rows16_ok(d->c, d->stride, d->dil, oh, ow)) ? 1 : 0;
    if (p.rows_mode) {
        if (!tmax::make_rows16(&mb, x, d->n, d->h, d->w, ow))
            return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(x rows) failed");
    } else if (!tmax::make_im2col_u8(&mb, x, d->n, d->h, d->w, d->c, -d->pad, -d->pad, d->pad - (d->kh - 1) * d->dil,
                                     d->pad - (d->kw - 1) * d->dil, d->stride, bkc, 128, bkc == 16 ? 0 : bkc))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(x) failed");
    if (!tmax::make_tiled_i32(&md, acc, (uint64_t)d->oc, (uint64_t)d->kh * d->kw * d->c))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(dW) failed");
    const int grid = p.items < ctx->sm_count ? p.items : ctx->sm_count;
    int rc;
    if (bkc == 16) rc = launch_dw2<16, 16>(ctx, ma, mb, md, p, grid, st);
    else if (bkc == 128) rc = G == 3 ? launch_dw2<128, 3>(ctx, ma, mb, md, p, grid, st) : launch_dw2<128, 1>(ctx, ma, mb, md, p, grid, st);
    else rc = G == 3 ? launch_dw2<64, 3>(ctx, ma, mb, md, p, grid, st) : launch_dw2<64, 1>(ctx, ma, mb, md, p, grid, st);
    if (rc) return rc;
    if (maxabs) return niti_absmax(ctx, acc, p.nout, maxabs, st);
    return NITI_OK;
}
