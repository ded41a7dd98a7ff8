// conv_stem2.cu — second-generation stem layer (image input: <= 4 live channels) in ONE launch:
//   conv -> max|acc| -> shared exponent -> PSTO -> (ReLU) -> int8 NHWC (+ the 2x2/2 max pool behind it) + exponent
//   (NITI_Conv_Int8.cpp:162-310, NITI_CPURelu_Int8, NITI_Maxpool_Int8.cpp:24-70)
//
// conv_stem.cu lets TMA build the operand: one 16-byte (16-channel) slice per pixel and tap, of which an RGB image fills 3
// bytes — K = taps x 16 = 160 for 27 real products, five MMA steps per tile and a TMA request per pixel and tap.  Here four
// builder warps assemble the K-major operand themselves, one aligned 4-byte word (<= 4 live channels) per pixel and tap:
// K = taps x 4 = 36 -> two K32 steps, 8 KB per 128-pixel tile, read through L1 from the 4 MB image tensor.
// Because the builders choose which pixel a tile row holds, a warp's 32 accumulator rows can be made 2 image rows x 16
// columns: the 2x2 pooling partners of a pixel are then lanes l^1 and l^16 of the same warp and the pool that follows the
// stem conv is two shuffles in the store pass (the separate pool launch and its 17 MB re-read disappear).
// The exponent needs max|acc| of the whole tensor, 67 MB of int32 that is never parked in HBM: pass 0 reduces the maximum AND
// stores with the shift predicted from the previous step's maximum (see conv_chain.cu); only when the prediction misses is the
// contraction RECOMPUTED for a second store pass (conv_stem.cu always ran it twice).
//
// Operand layouts (un-swizzled K-major core matrices, 8 rows x 16 B):
//   A stage: [Kp/16 chunks][128 rows][16 B]   LBO (next 16 B of K) = 2048 B, SBO (next 8 rows) = 128 B
//   B      : [Kp/16 chunks][oc rows][16 B]    LBO = oc*16 B,              SBO = 128 B     (built once per CTA from w)
#include "tma_common.cuh"

namespace {

constexpr int kS2Threads = 800;      // warp 0: MMA issuer + TMEM owner, warps 1-8: operand builders, warps 9-24: epilogue
constexpr int kS2Stages = 4;         // A operand ring (when a CTA's tiles do not all fit in shared memory)
constexpr int kS2MaxStages = 32;     // resident mode: one slot per tile of the CTA, built once, used by both passes
constexpr int kS2AccStages = 4;      // accumulator ring in tensor memory
constexpr int kS2MaxTaps = 64;

// division by a run-time constant as multiply-high + shift (dividends < 2^31)
struct FastDiv {
    uint32_t mul, shr; int d;
    __host__ void init(int div) {
        d = div;
        if (div <= 1) { mul = 0; shr = 0; return; }
        int l = 0; while ((1u << l) < (uint32_t)div) ++l;
        const uint64_t pw = 1ull << (31 + l);
        mul = (uint32_t)((pw + (uint64_t)div - 1) / (uint64_t)div); shr = (uint32_t)(l - 1);
    }
    __device__ __forceinline__ int div(int n) const { return d <= 1 ? n : (int)(__umulhi((uint32_t)n, mul) >> shr); }
};

struct Stem2Params {
    const int8_t* x; const int8_t* w;            // x [n][h][w][C], w [oc][taps][C]; only the first 4 channels are read
    int n, h, wd, C;
    int oc, oh, ow, kh, kw, dil, stride, pad;
    int taps, Kp;                                // K = taps * 4 bytes, Kp = K rounded up to 32
    int M, m_tiles;
    int pool, units, units_per_img, xblocks;     // pool order: unit = (image, row pair, 16-column block) = 32 pixels
    FastDiv d_upi, d_xb, d_ow, d_oh;
    int stages;                                  // A slots in shared memory
    int resident;                                // every tile of a CTA keeps its A operand in shared memory for the second pass
    int8_t* y; int8_t* y_pool; uint32_t* maxabs;
    const int8_t* exp_a; const int8_t* exp_b; int8_t* exp_out;
    int relu;
    uint32_t* sync;                              // {arrival counter, exit counter}
    DpView dp; int dp_slot;
    uint32_t* pred; int speculate;               // exponent prediction: last max|acc| of this slot (conv_chain.cu)
    unsigned long long* tl;
};

__device__ __forceinline__ void s2_tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}

// which output pixel row r of tile `tile` holds; false: beyond the tensor
__device__ __forceinline__ bool s2_pixel(const Stem2Params& p, int tile, int r, int& n_, int& y_, int& x_) {
    if (p.pool) {
        const int u = tile * 4 + (r >> 5), l = r & 31;
        if (u >= p.units) return false;
        n_ = p.d_upi.div(u); const int rem = u - n_ * p.units_per_img;
        const int y2 = p.d_xb.div(rem), xb = rem - y2 * p.xblocks;
        y_ = 2 * y2 + (l >> 4); x_ = xb * 16 + (l & 15);
        return true;
    }
    const int idx = tile * 128 + r;
    if (idx >= p.M) return false;
    const int t = p.d_ow.div(idx); x_ = idx - t * p.ow;
    n_ = p.d_oh.div(t); y_ = t - n_ * p.oh;
    return true;
}

__global__ void __launch_bounds__(kS2Threads, 1)
stem2_fwd_kernel(const Stem2Params p) {
    extern __shared__ uint8_t smem_raw_s2[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw_s2) + 1023) & ~(uintptr_t)1023);
    const int stage_bytes = p.Kp * 128;
    uint8_t* bsm = smem + p.stages * stage_bytes;                               // weights [Kp/16][oc][16]
    const int b_bytes = p.Kp * p.oc;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(bsm + ((b_bytes + 127) & ~127));
    uint64_t* empty_bar = full_bar + kS2MaxStages;
    uint64_t* tfull_bar = empty_bar + kS2MaxStages;
    uint64_t* tempty_bar = tfull_bar + kS2AccStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + kS2AccStages);
    uint32_t* bcast = tmem_slot + 1;
    uint32_t* cta_max = tmem_slot + 2;
    uint32_t* redo_flag = tmem_slot + 3;                                         // pass 1 needed (prediction missed / speculation off)
    int* tapoff = reinterpret_cast<int*>(tmem_slot + 4);                         // [taps] (dy << 16) | (dx & 0xffff)
    uint64_t* decide_bar = reinterpret_cast<uint64_t*>(tapoff + kS2MaxTaps);     // redo_flag is valid

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    niti_tl_enter(p.tl);
    const int accw = p.oc <= 32 ? 32 : (p.oc <= 64 ? 64 : 128);                  // columns per accumulator stage (oc <= 128)

    if (tid == 0) {
        for (int s = 0; s < kS2MaxStages; ++s) { tc::mbar_init(&full_bar[s], 4); tc::mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < kS2AccStages; ++a) { tc::mbar_init(&tfull_bar[a], 1); tc::mbar_init(&tempty_bar[a], 16); }
        tc::mbar_init(decide_bar, 1);
        tc::fence_barrier_init();
        *cta_max = 0u;
    }
    if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
    if (tid < p.taps) {
        const int ky = tid / p.kw, kx = tid - ky * p.kw;
        tapoff[tid] = ((ky * p.dil - p.pad) << 16) | ((kx * p.dil - p.pad) & 0xffff);
    }
    // B operand: word (tap t, 4 channels) of output channel o -> chunk t/4, row o, word t%4; K padding is zero
    for (int i = tid; i < (p.Kp >> 2) * p.oc; i += kS2Threads) {
        const int t = i / p.oc, o = i - t * p.oc;
        const uint32_t v = t < p.taps ? *reinterpret_cast<const uint32_t*>(p.w + ((size_t)o * p.taps + t) * p.C) : 0u;
        *reinterpret_cast<uint32_t*>(bsm + ((size_t)(t >> 2) * p.oc + o) * 16 + (t & 3) * 4) = v;
    }
    tmax::fence_proxy_async_all();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    niti_pdl_trigger();
    niti_pdl_wait();                                         // programmatic dependent launch: see common.cuh
    const int my_tiles = (p.m_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    if (warp == 0) {
        // ------------------------------ MMA issuer ------------------------------
        const uint32_t idesc = tc::make_idesc_i8(128, p.oc, 0, 0);
        const uint32_t b_lbo = (uint32_t)p.oc * 16u;
        const uint64_t adc = tmax::make_desc(0u, 2048, 128, 0u);
        const uint64_t bdc = tmax::make_desc(0u, b_lbo, 128, 0u);
        const uint64_t bd0 = bdc | (uint64_t)(tc::smem_u32(bsm) >> 4);
        const int ksteps = p.Kp >> 5;
        uint32_t n = 0, tile_iter = 0;
        for (int pass = 0; pass < 2; ++pass) {
            if (pass == 1) {                                                      // second pass only when the epilogue asks for it
                tc::mbar_wait(decide_bar, 0u);
                if (*reinterpret_cast<volatile uint32_t*>(redo_flag) == 0u) break;
            }
            for (int i = 0; i < my_tiles; ++i, ++tile_iter, ++n) {
                const uint32_t a = tile_iter % kS2AccStages, aph = (tile_iter / kS2AccStages) & 1u;
                const uint32_t s = p.resident ? (uint32_t)i : n % kS2Stages, ph = p.resident ? 0u : (n / kS2Stages) & 1u;
                tc::mbar_wait(&tempty_bar[a], aph ^ 1u);
                tc::mbar_wait(&full_bar[s], ph);
                tc::tc_fence_after();
                const uint64_t ad0 = adc | (uint64_t)((tc::smem_u32(smem) + s * (uint32_t)stage_bytes) >> 4);
                if (tc::elect_one()) {
                    for (int j = 0; j < ksteps; ++j)
                        tc::umma_i8(tmem_base + a * accw, ad0 + (uint64_t)(j * (4096 >> 4)), bd0 + (uint64_t)(j * ((2u * b_lbo) >> 4)), idesc,
                                    (uint32_t)(j != 0));
                    tc::umma_commit(&empty_bar[s]);
                    tc::umma_commit(&tfull_bar[a]);
                }
            }
        }
    } else if (warp <= 8) {
        // ------------------------------ operand builders ------------------------------
        // two groups of 128 threads (one tile row each) take alternate tiles, so two tiles' loads are in flight; all of a
        // row's words are loaded before the first store (one L2 latency per tile, not one per 16-byte chunk)
        const int bt = tid - 32, grp = bt >> 7, r = bt & 127;
        const int chunks = p.Kp >> 4;
        const uint32_t total = (p.resident ? 1u : 2u) * (uint32_t)my_tiles;
        bool decided = false;
        for (uint32_t n = (uint32_t)grp; n < total; n += 2) {
            if (n >= (uint32_t)my_tiles && !decided) {                                // pass 2 (ring mode) only on request
                tc::mbar_wait(decide_bar, 0u);
                decided = true;
                if (*reinterpret_cast<volatile uint32_t*>(redo_flag) == 0u) break;
            }
            const int i = (int)(n >= (uint32_t)my_tiles ? n - my_tiles : n);          // pass 2 repeats the tiles of pass 1
            const int tile = (int)blockIdx.x + i * (int)gridDim.x;
            int n_ = 0, y_ = 0, x_ = 0;
            const bool valid = s2_pixel(p, tile, r, n_, y_, x_);
            const int iy0 = y_ * p.stride, ix0 = x_ * p.stride;
            const int8_t* img = p.x + (size_t)n_ * p.h * p.wd * p.C;
            const uint32_t s = p.resident ? n : n % kS2Stages, ph = (n / kS2Stages) & 1u;
            uint8_t* dst = smem + s * stage_bytes + r * 16;
            auto word = [&](int t) -> uint32_t {
                if (!valid || t >= p.taps) return 0u;
                const int to = tapoff[t];
                const int iy = iy0 + (to >> 16), ix = ix0 + (int)(int16_t)(to & 0xffff);
                if ((unsigned)iy >= (unsigned)p.h || (unsigned)ix >= (unsigned)p.wd) return 0u;
                return __ldg(reinterpret_cast<const uint32_t*>(img + ((size_t)iy * p.wd + ix) * p.C));
            };
            if (chunks <= 4) {                                                        // <= 16 taps (3x3, 4x4): registers
                uint32_t q[16];
#pragma unroll
                for (int t = 0; t < 16; ++t) q[t] = word(t);
                if (!p.resident) tc::mbar_wait(&empty_bar[s], ph ^ 1u);
#pragma unroll
                for (int g = 0; g < 4; ++g)
                    if (g < chunks) *reinterpret_cast<uint4*>(dst + g * 2048) = make_uint4(q[4 * g], q[4 * g + 1], q[4 * g + 2], q[4 * g + 3]);
            } else {
                // 16 words (four 16-byte chunks) in flight per thread: with one chunk at a time a 7x7 row (49 taps, 13 chunks) paid 13
                // L2 latencies per tile and the builders, not the tensor pipe or the stores, set the pace (ResNet stem: 0.66 ms)
                if (!p.resident) tc::mbar_wait(&empty_bar[s], ph ^ 1u);
                for (int g0 = 0; g0 < chunks; g0 += 4) {
                    uint32_t q[16];
#pragma unroll
                    for (int t = 0; t < 16; ++t) q[t] = word(4 * g0 + t);
#pragma unroll
                    for (int g = 0; g < 4; ++g)
                        if (g0 + g < chunks) *reinterpret_cast<uint4*>(dst + (g0 + g) * 2048) = make_uint4(q[4 * g], q[4 * g + 1], q[4 * g + 2], q[4 * g + 3]);
                }
            }
            tmax::fence_proxy_async_all();               // generic-proxy stores -> tensor-core (async proxy) reads
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&full_bar[s]);
        }
    } else {
        // ------------------------------ epilogue (warps 9..24) ------------------------------
        // 16 warps (the rescale needs the issue slots of the whole SM); a warp may read TMEM lanes 32*(warp%4)..+31, so four
        // warps share a lane quarter and take every fourth 16-column chunk.
        const int quarter = warp & 3;
        const int group = (warp - 9) >> 2;
        const int etid = tid - 288;
        uint32_t tile_iter = 0, mx = 0;
        int shift = 0, kind = 0;
        const bool spec = p.speculate != 0;                                      // a miss costs what the recompute schedule always paid: no gating
        // image layer: its input norm is the data's, so only the window predictor applies (in_now = 0 switches the ratio off)
        NitiPrediction pr = niti_predict(p.pred, p.dp_slot, 0u);
        pr.bw = pr.bw_window; pr.bw_ratio = pr.bw_window;
        niti_rule_a(pr.bw, kind, shift);                                         // predicted from the previous steps (common.cuh)
        const uint32_t eone = (uint32_t)(p.speculate >= 0);                      // 1, opaque to the compiler
        NitiRuleA rk = niti_rule_a_consts(shift, eone);
        int mxs = 0, mns = 0;
        for (int pass = 0; pass < 2; ++pass) {
            for (int i = 0; i < my_tiles; ++i, ++tile_iter) {
                const int tile = (int)blockIdx.x + i * (int)gridDim.x;
                int n_ = 0, y_ = 0, x_ = 0;
                const bool row_ok = s2_pixel(p, tile, quarter * 32 + lane, n_, y_, x_);
                const uint32_t a = tile_iter % kS2AccStages, aph = (tile_iter / kS2AccStages) & 1u;
                tc::mbar_wait_backoff(&tfull_bar[a], aph);
                tc::tc_fence_after();
#pragma unroll 1
                for (int c0 = group * 16; c0 < p.oc; c0 += 64) {
                    uint32_t v[16];
                    s2_tmem_ld16(tmem_base + ((uint32_t)(quarter * 32) << 16) + a * accw + (uint32_t)c0, v);
                    tc::tmem_ld_wait();
                    if (pass == 0) {
                        if (row_ok) {
#pragma unroll
                            for (int q = 0; q < 16; q += 2) {
                                mxs = __vimax3_s32(mxs, (int32_t)v[q], (int32_t)v[q + 1]);
                                mns = __vimin3_s32(mns, (int32_t)v[q], (int32_t)v[q + 1]);
                            }
                        }
                        if (!spec) continue;
                    }
                    uint32_t o[4];
                    niti_rescale16(v, kind, shift, rk, o, p.relu != 0);
                    if (row_ok)
                        *reinterpret_cast<uint4*>(p.y + (((size_t)n_ * p.oh + y_) * p.ow + x_) * p.oc + c0) = make_uint4(o[0], o[1], o[2], o[3]);
                    if (p.pool) {
                        // lanes l^1 (x parity) and l^16 (row parity) hold the other three pixels of the 2x2 window
                        uint32_t m[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            m[q] = __vmaxs4(o[q], __shfl_xor_sync(0xffffffffu, o[q], 1));
                            m[q] = __vmaxs4(m[q], __shfl_xor_sync(0xffffffffu, m[q], 16));
                        }
                        if (row_ok && !(lane & 17))
                            *reinterpret_cast<uint4*>(p.y_pool + (((size_t)n_ * (p.oh >> 1) + (y_ >> 1)) * (p.ow >> 1) + (x_ >> 1)) * p.oc + c0) =
                                make_uint4(m[0], m[1], m[2], m[3]);
                    }
                }
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&tempty_bar[a]);
            }
            if (pass == 1) break;
            // ---- global max|acc|: warp max -> CTA max -> one atomicMax -> grid barrier -> (data parallel: peer exchange) ----
            mx = warp_max_u32((uint32_t)max(mxs, -mns));
            if (lane == 0 && mx) atomicMax(cta_max, mx);
            asm volatile("bar.sync 1, 512;" ::: "memory");
            if (etid == 0) {
                const uint32_t cm = *cta_max;
                if (cm) atomicMax(p.maxabs, cm);
                tmax::grid_barrier(p.sync, gridDim.x);
                *bcast = dp_exchange_max(p.dp, p.dp_slot, __ldcg(p.maxabs), blockIdx.x == 0);
            }
            asm volatile("bar.sync 1, 512;" ::: "memory");
            const uint32_t gmax = *bcast;
            int kind2, shift2;
            niti_rule_a(niti_bitwidth(gmax), kind2, shift2);
            const bool hit = spec && kind2 == kind && shift2 == shift;          // grid-uniform
            if (etid == 0) {
                *reinterpret_cast<volatile uint32_t*>(redo_flag) = hit ? 0u : 1u;
                tc::mbar_arrive(decide_bar);                                     // releases the MMA issuer and the builders (release.cta)
                if (blockIdx.x == 0) {
                    if (p.exp_out != nullptr)
                        *p.exp_out = (int8_t)((p.exp_a ? (int)*p.exp_a : 0) + (p.exp_b ? (int)*p.exp_b : 0) + (kind2 == 1 ? 0 : shift2));
                    niti_predict_update(p.pred, p.dp_slot, pr, gmax, kind2 == kind && shift2 == shift, spec);   // every CTA read the old words before the barrier
                    p.pred[kPredLastNorm] = niti_out_norm(gmax, kind2, shift2);   // input norm of the chain's first layer
                }
            }
            if (hit) break;
            kind = kind2; shift = shift2;
            rk = niti_rule_a_consts(shift, eone);
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
    if (warp == 0) tc::tmem_dealloc<512>(tmem_base);
    niti_tl_exit(p.tl);
}

} // namespace

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int stem2_smem(const niti_conv_desc* d, int stages) {
    const int taps = d->kh * d->kw, kp = (taps * 4 + 31) & ~31;
    return stages * kp * 128 + ((kp * d->oc + 127) & ~127) + 2 * kS2MaxStages * 8 + 512 + kS2MaxTaps * 4 + 64 + 1024;
}

// c_live: how many leading channels of x may be non-zero (the caller's guarantee: an image tensor padded to d->c channels)
bool stem2_fwd_eligible(const niti_conv_desc* d, int c_live, bool pool) {
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    if (c_live < 1 || c_live > 4 || d->c < 4 || d->c % 4 != 0) return false;
    if (d->oc % 16 != 0 || d->oc < 16 || d->oc > 128 || d->kh * d->kw > kS2MaxTaps || d->pad > 127) return false;
    if (stem2_smem(d, kS2Stages) > 227 * 1024) return false;
    if (pool && (ow % 16 != 0 || oh % 2 != 0)) return false;
    return true;
}

// y = rescale(conv(x, w)) (+ y_pool = maxpool2x2(y)) in one launch; maxabs: zeroed slot, holds this rank's max|acc| afterwards
int stem2_conv_fwd_fused(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, const int8_t* ex,
                         const int8_t* ew, int8_t* y, int8_t* ey, int relu, int8_t* y_pool, uint32_t* maxabs, int dp_slot, cudaStream_t st) {
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    Stem2Params p{};
    p.x = x; p.w = w; p.n = d->n; p.h = d->h; p.wd = d->w; p.C = d->c;
    p.oc = d->oc; p.oh = oh; p.ow = ow; p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.stride = d->stride; p.pad = d->pad;
    p.taps = d->kh * d->kw; p.Kp = (p.taps * 4 + 31) & ~31;
    p.M = d->n * oh * ow;
    p.pool = y_pool != nullptr ? 1 : 0;
    if (p.pool) {
        p.xblocks = ow / 16; p.units_per_img = (oh / 2) * p.xblocks; p.units = d->n * p.units_per_img;
        p.m_tiles = (p.units + 3) / 4;
    } else {
        p.m_tiles = (p.M + 127) / 128;
    }
    p.d_upi.init(p.pool ? p.units_per_img : 1); p.d_xb.init(p.pool ? p.xblocks : 1); p.d_ow.init(ow); p.d_oh.init(oh);
    p.y = y; p.y_pool = y_pool; p.maxabs = maxabs; p.exp_a = ex; p.exp_b = ew; p.exp_out = ey; p.relu = relu;
    p.sync = ctx->ws_u32 + 48;
    p.dp = ctx->dp; p.dp_slot = dp_slot; p.pred = ctx->pred; p.speculate = ctx->speculate; p.tl = niti_tl_next(ctx, "stem2_fwd");
    // one CTA per SM, enforced through the shared-memory request: every CTA allocates all 512 TMEM columns and waits at a
    // grid barrier (see conv_tma2.cu launch_fused)
    const int grid = p.m_tiles < ctx->sm_count ? p.m_tiles : ctx->sm_count;
    const int per_cta = (p.m_tiles + grid - 1) / grid;
    p.resident = per_cta <= kS2MaxStages && stem2_smem(d, per_cta) <= 227 * 1024 ? 1 : 0;
    p.stages = p.resident ? per_cta : kS2Stages;
    const int need = stem2_smem(d, p.stages);
    const int smem = need > 120 * 1024 ? need : 120 * 1024;
    if (int rc_cfg = niti_func_smem(ctx, (const void*)stem2_fwd_kernel, smem, "cudaFuncSetAttribute(stem2_fwd)")) return rc_cfg;
    cudaError_t lerr = niti_launch(ctx, NITI_PDL_STEM, stem2_fwd_kernel, dim3(grid), dim3(kS2Threads), (size_t)smem, st, p);
    if (lerr != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch stem2_fwd", lerr);
    return niti_check_launch(ctx, "stem2_fwd");
}
