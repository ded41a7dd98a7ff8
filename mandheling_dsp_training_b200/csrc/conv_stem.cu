// conv_stem.cu — the whole forward op of a stem layer (<= 16 input channels) in ONE launch:
//   conv -> max|acc| -> shared exponent -> PSTO -> (ReLU) -> int8 NHWC + exponent        (NITI_Conv_Int8.cpp:162-310)
//
// A stem layer's contraction is tiny (K = taps * 16) but its output is the largest tensor of the net: at VGG-11 b256 the
// int32 accumulators are 67 MB against 4 MB of input, so the two-phase path (write int32, read it back twice) is bound by
// HBM traffic that the algorithm does not need.  This kernel RECOMPUTES instead: pass 1 runs the contraction and keeps
// only max|acc| (TMEM -> registers -> warp max), a grid barrier (plus the NVLink peer exchange when data-parallel)
// publishes the global maximum, pass 2 runs the contraction again and stores the rescaled int8 tile directly.
// HBM traffic: 2 x |x| (second read from L2) + |y| int8.  Tensor work doubles, which is free here (K <= 800).
//
// Operands are un-swizzled K-major core-matrix layouts built by TMA itself:
//   A: one im2col box {16 channels, 128 pixels} per filter tap -> [tap][128 pixels][16 B]; a K32 MMA step spans two taps
//      (LBO = 2048 B between the taps, SBO = 128 B between 8-pixel groups)
//   B: weights [oc][tap][16] loaded once per CTA as [tap][oc][16 B] (LBO = oc*16 B, SBO = 128 B); an odd tap count is
//      padded with one zero tap in B, so whatever the matching A half holds contributes nothing.
#include "tma_common.cuh"

namespace {

constexpr int kStemThreads = 576;    // warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-17: epilogue
constexpr int kTapsPerBlock = 10;    // taps per pipeline stage (even): 20 KB of A
constexpr int kStemStages = 6;
constexpr int kStemMaxB = 65536;     // resident weight bytes

struct StemParams {
    int M, oc;                        // pixels x out-channels
    int out_h, out_w, kh, kw, dil, stride, lower_h, lower_w;
    int taps, taps_p, num_kb, m_tiles;
    int rows_mode;                    // A tiles loaded row-wise (tmax::make_rows16) instead of per pixel (im2col)
    int8_t* y; uint32_t* maxabs;
    const int8_t* exp_a; const int8_t* exp_b; int8_t* exp_out;
    int relu;
    uint32_t* sync;                   // {arrival counter, exit counter}
    DpView dp; int dp_slot;
    unsigned long long* tl;
};


__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}

__global__ void __launch_bounds__(kStemThreads, 1)
stem_fwd_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const StemParams p) {
    extern __shared__ uint8_t smem_raw_stem[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw_stem) + 1023) & ~(uintptr_t)1023);
    constexpr int STAGE = kTapsPerBlock * 2048;
    uint8_t* bsm = smem + kStemStages * STAGE;                                   // weights [taps_p][oc][16]
    const int b_bytes = p.taps_p * p.oc * 16;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(bsm + ((b_bytes + 127) & ~127));
    uint64_t* empty_bar = full_bar + kStemStages;
    uint64_t* tfull_bar = empty_bar + kStemStages;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint64_t* w_bar = tempty_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_bar + 1);
    uint32_t* bcast = tmem_slot + 1;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    niti_tl_enter(p.tl);
    const int accw = p.oc <= 32 ? 32 : (p.oc <= 64 ? 64 : (p.oc <= 128 ? 128 : 256));   // columns per accumulator stage

    if (tid == 0) {
        for (int s = 0; s < kStemStages; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < 2; ++a) { tc::mbar_init(&tfull_bar[a], 1); tc::mbar_init(&tempty_bar[a], 16); }
        tc::mbar_init(w_bar, 1);
        tc::fence_barrier_init();
        tmax::prefetch_map(&mapA); tmax::prefetch_map(&mapB);
    }
    if (warp == 1) tc::tmem_alloc<512>(tmem_slot);
    // zero the padding tap of B (generic proxy), made visible to the tensor-core (async) proxy below
    if (p.taps_p != p.taps)
        for (int i = tid; i < p.oc * 4; i += kStemThreads) reinterpret_cast<uint32_t*>(bsm + p.taps * p.oc * 16)[i] = 0u;
    tmax::fence_proxy_async_all();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    niti_pdl_trigger();
    niti_pdl_wait();                                         // programmatic dependent launch: see common.cuh
    const int my_tiles = (p.m_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    // issue loops: whole warp on uniform values, single-thread instructions under elect.sync (see conv_chain.cu)
    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        if (tc::elect_one()) {
            tmax::mbar_expect_tx(w_bar, (uint32_t)(p.taps * p.oc * 16));
            for (int t = 0; t < p.taps; ++t) tmax::load_2d(&mapB, tc::smem_u32(bsm + t * p.oc * 16), w_bar, t * 16, 0);
        }
        uint32_t n = 0;
        for (int pass = 0; pass < 2; ++pass)
            for (int i = 0; i < my_tiles; ++i) {
                const int m0 = ((int)blockIdx.x + i * (int)gridDim.x) * 128;
                const int q = m0 % p.out_w; const int t = m0 / p.out_w;
                const int pp = t % p.out_h; const int bn = t / p.out_h;
                const int bw = q * p.stride + p.lower_w, bh = pp * p.stride + p.lower_h;
                int kx = 0, ky = 0;
                for (int kb = 0; kb < p.num_kb; ++kb, ++n) {
                    const int s = n % kStemStages; const uint32_t ph = (n / kStemStages) & 1u;
                    tc::mbar_wait(&empty_bar[s], ph ^ 1u);
                    const int tap0 = kb * kTapsPerBlock, nt = min(kTapsPerBlock, p.taps - tap0);
                    const uint32_t a_dst = tc::smem_u32(smem + s * STAGE);
                    if (tc::elect_one()) {
                        tmax::mbar_expect_tx(&full_bar[s], (uint32_t)(nt * 2048));
                        int kxx = kx, kyy = ky;
                        for (int g = 0; g < nt; ++g) {
                            if (p.rows_mode) tmax::load_3d(&mapA, a_dst + g * 2048, &full_bar[s], (kxx + p.lower_w) * 2, pp + kyy + p.lower_h, bn);
                            else tmax::load_im2col(&mapA, a_dst + g * 2048, &full_bar[s], 0, bw, bh, bn,
                                                   (uint16_t)(kxx * p.dil), (uint16_t)(kyy * p.dil));
                            if (++kxx == p.kw) { kxx = 0; ++kyy; }
                        }
                    }
                    for (int g = 0; g < nt; ++g) if (++kx == p.kw) { kx = 0; ++ky; }
                }
            }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer ------------------------------
        const uint32_t idesc = tc::make_idesc_i8(128, p.oc, 0, 0);
        const uint32_t b_lbo = (uint32_t)p.oc * 16u;
        const uint64_t adc = tmax::make_desc(0u, 2048, 128, 0u);
        const uint64_t bdc = tmax::make_desc(0u, b_lbo, 128, 0u);
        tc::mbar_wait(w_bar, 0);
        tc::tc_fence_after();
        uint32_t n = 0, tile_iter = 0;
        for (int pass = 0; pass < 2; ++pass)
            for (int i = 0; i < my_tiles; ++i, ++tile_iter) {
                const uint32_t a = tile_iter & 1u, aph = (tile_iter >> 1) & 1u;
                tc::mbar_wait(&tempty_bar[a], aph ^ 1u);
                tc::tc_fence_after();
                const uint32_t d_tmem = tmem_base + a * accw;
                for (int kb = 0; kb < p.num_kb; ++kb, ++n) {
                    const int s = n % kStemStages; const uint32_t ph = (n / kStemStages) & 1u;
                    tc::mbar_wait(&full_bar[s], ph);
                    tc::tc_fence_after();
                    const uint32_t a_base = tc::smem_u32(smem + s * STAGE);
                    const int tap0 = kb * kTapsPerBlock, ntp = min(kTapsPerBlock, p.taps_p - tap0);     // even
                    const uint64_t ad0 = adc | (uint64_t)(a_base >> 4);
                    const uint64_t bd0 = bdc | (uint64_t)((tc::smem_u32(bsm) + (uint32_t)tap0 * b_lbo) >> 4);
                    if (tc::elect_one()) {
                        for (int j = 0; j < ntp / 2; ++j)
                            tc::umma_i8(d_tmem, ad0 + (uint64_t)(j * (4096 >> 4)), bd0 + (uint64_t)(j * ((2u * b_lbo) >> 4)), idesc,
                                        (uint32_t)((kb | j) != 0));
                        tc::umma_commit(&empty_bar[s]);
                        if (kb == p.num_kb - 1) tc::umma_commit(&tfull_bar[a]);
                    }
                }
            }
    } else {
        // ------------------------------ epilogue (warps 2..17) ------------------------------
        // 16 warps: the rescale arithmetic (~15 integer instructions per output) needs the issue slots of a full SM, one
        // warp per scheduler is 4x too few.  A warp may read TMEM lanes 32*(warp%4)..+31, so four warps share a lane
        // quarter and take every fourth 16-column chunk.
        const int quarter = warp & 3;
        const int group = (warp - 2) >> 2;
        const int etid = tid - 64;
        uint32_t tile_iter = 0, mx = 0;
        int shift = 0, kind = 0;
        for (int pass = 0; pass < 2; ++pass) {
            for (int i = 0; i < my_tiles; ++i, ++tile_iter) {
                const int m0 = ((int)blockIdx.x + i * (int)gridDim.x) * 128;
                const int row = m0 + quarter * 32 + lane;
                const bool row_ok = row < p.M;
                const uint32_t a = tile_iter & 1u, aph = (tile_iter >> 1) & 1u;
                tc::mbar_wait(&tfull_bar[a], aph);
                tc::tc_fence_after();
                if (m0 + quarter * 32 < p.M) {               // warp-uniform
#pragma unroll 1
                    for (int c0 = group * 16; c0 < p.oc; c0 += 64) {
                        uint32_t v[16];
                        tmem_ld_32x32b_x16(tmem_base + ((uint32_t)(quarter * 32) << 16) + a * accw + (uint32_t)c0, v);
                        tc::tmem_ld_wait();
                        if (pass == 0) {
                            if (row_ok) {
#pragma unroll
                                for (int q = 0; q < 16; ++q) mx = max(mx, niti_abs_u32((int32_t)v[q]));
                            }
                        } else if (row_ok) {
                            uint32_t o[4];
                            niti_rescale16(v, kind, shift, o, p.relu != 0);
                            *reinterpret_cast<uint4*>(p.y + (size_t)row * p.oc + c0) = make_uint4(o[0], o[1], o[2], o[3]);
                        }
                    }
                }
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&tempty_bar[a]);
            }
            if (pass == 0) {
                // ---- global max|acc|: warp max -> atomicMax -> grid barrier -> (data parallel: peer exchange) ----
                mx = warp_max_u32(mx);
                if (lane == 0 && mx) atomicMax(p.maxabs, mx);
                asm volatile("bar.sync 1, 512;" ::: "memory");
                if (etid == 0) {
                    tmax::grid_barrier(p.sync, gridDim.x);
                    const uint32_t m = dp_exchange_max(p.dp, p.dp_slot, __ldcg(p.maxabs), blockIdx.x == 0);
                    *bcast = m;
                }
                asm volatile("bar.sync 1, 512;" ::: "memory");
                const int b = niti_bitwidth(*bcast);
                const int s = b - 7;
                if (s > 1) { shift = s; kind = 0; } else if (s == 1) { shift = 2; kind = 0; } else { shift = 0; kind = 1; }
                if (blockIdx.x == 0 && etid == 0 && p.exp_out != nullptr)
                    *p.exp_out = (int8_t)((p.exp_a ? (int)*p.exp_a : 0) + (p.exp_b ? (int)*p.exp_b : 0) + (kind == 1 ? 0 : shift));
            }
        }
        asm volatile("bar.sync 1, 512;" ::: "memory");
        if (etid == 0) {                                     // last CTA out re-arms the grid barrier
            __threadfence();
            if (atomicAdd(p.sync + 1, 1u) == gridDim.x - 1) { p.sync[0] = 0u; p.sync[1] = 0u; __threadfence(); }
        }
    }
    __syncwarp();
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc<512>(tmem_base);
    niti_tl_exit(p.tl);
}

} // namespace

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
bool stem_fwd_eligible(const niti_conv_desc* d) {
    const int taps = d->kh * d->kw, taps_p = (taps + 1) & ~1;
    // im2col-mode TMA issues one 16-byte L2 request per pixel and tap (measured ~0.12 requests/clk/SM): only worth it for
    // small outputs; the row-wise mode (one request per image row) has no such limit
    const bool rows = tmax::rows16_ok(d->c, d->stride, d->dil, conv_out_h(d), conv_out_w(d));
    const long requests = (long)d->n * conv_out_h(d) * conv_out_w(d) * taps;
    return d->c == 16 && d->oc % 16 == 0 && d->oc <= 256 && taps_p * d->oc * 16 <= kStemMaxB && d->pad <= 127 &&
           (d->kh - 1) * d->dil <= 255 && (d->kw - 1) * d->dil <= 255 && d->stride <= 8 && (rows || requests <= 400000);
}

// y = rescale(conv(x, w)) in one launch; maxabs must be a zeroed slot, holds this rank's max|acc| afterwards
int stem_conv_fwd_fused(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, const int8_t* ex,
                        const int8_t* ew, int8_t* y, int8_t* ey, int relu, uint32_t* maxabs, int dp_slot, cudaStream_t st) {
    if (!tmax::driver_ok()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    StemParams p{};
    p.M = d->n * oh * ow; p.oc = d->oc;
    p.out_h = oh; p.out_w = ow; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = d->stride;
    p.lower_h = -d->pad; p.lower_w = -d->pad;
    p.taps = d->kh * d->kw; p.taps_p = (p.taps + 1) & ~1;
    p.num_kb = (p.taps_p + kTapsPerBlock - 1) / kTapsPerBlock;
    p.m_tiles = (p.M + 127) / 128;
    p.y = y; p.maxabs = maxabs; p.exp_a = ex; p.exp_b = ew; p.exp_out = ey; p.relu = relu;
    p.sync = ctx->ws_u32 + 48;
    p.dp = ctx->dp; p.dp_slot = dp_slot; p.tl = niti_tl_next(ctx, "stem_fwd");
    CUtensorMap ma, mb;
    p.rows_mode = tmax::rows16_ok(d->c, d->stride, d->dil, oh, ow) ? 1 : 0;
    if (p.rows_mode) {
        if (!tmax::make_rows16(&ma, x, d->n, d->h, d->w, ow))
            return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(x rows, stem) failed");
    } else if (!tmax::make_im2col_u8(&ma, x, d->n, d->h, d->w, d->c, -d->pad, -d->pad, d->pad - (d->kh - 1) * d->dil,
                                     d->pad - (d->kw - 1) * d->dil, d->stride, 16, 128, 0))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(x, stem) failed");
    if (!tmax::make_tiled_u8(&mb, w, (uint64_t)d->oc, (uint64_t)p.taps * 16, (uint32_t)d->oc, 16, 0))
        return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(w, stem) failed");
    const int smem = kStemStages * kTapsPerBlock * 2048 + ((p.taps_p * d->oc * 16 + 127) & ~127) + 512 + 1024;
    if (int rc_cfg = niti_func_smem(ctx, (const void*)stem_fwd_kernel, (int)(smem), "cudaFuncSetAttribute(stem_fwd)")) return rc_cfg;
    const int grid = p.m_tiles < ctx->sm_count ? p.m_tiles : ctx->sm_count;
    cudaError_t lerr = niti_launch(ctx, NITI_PDL_STEM, stem_fwd_kernel, dim3(grid), dim3(kStemThreads), (size_t)smem, st, ma, mb, p);
    if (lerr != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch stem_fwd", lerr);
    return niti_check_launch(ctx, "stem_fwd");
}
