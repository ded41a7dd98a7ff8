// conv_tc.cu — tcgen05 (5th-gen tensor core) int8 implicit-GEMM kernels for the NITI contractions.
//
//   forward : D[pixel][oc]   = sum_{tap,c}  x[pixel+tap][c]   * w[oc][tap][c]     (A, B K-major)
//   dX      : D[pixel][c]    = sum_{tap,oc} dy[pixel-tap][oc] * wT[c][tap][oc]    (A, B K-major)
//   dW      : D[oc][tap,c]   = sum_{pixel}  dy[pixel][oc]     * x[pixel+tap][c]   (A, B MN-major)
//
// One CTA = one 128 x BN accumulator tile held in TMEM (int32).  Warps 0-3 are producers: they gather
// im2col-free operand tiles straight from the NHWC tensors with 16-byte cp.async (zero-fill for
// padding / ragged edges) into 128-byte-swizzled shared memory, 8 lanes per 128-byte row so every
// global request is a full line, and signal an mbarrier per stage.  Warp 4 lane 0 issues
// tcgen05.mma.kind::i8 (M=128, N=BN, K=32 per instruction, 4 per stage) and releases stages with
// tcgen05.commit.  When the K loop is done the producer warps become the epilogue: tcgen05.ld the
// accumulator, write int32 + fold max|acc| (warp reduce, one atomicMax per warp) for the
// shared-exponent step that follows (rescale.cu).  The same 128-byte-row shared-memory image serves
// the K-major (rows = pixels, bytes = channels = K) and the MN-major (rows = pixels = K, bytes =
// channels = M/N) interpretation; only the UMMA descriptors differ.
//
// Replaces: NITI_Conv_Int8.cpp:212-249, NITI_DeConv_Int8.cpp:208-290,
// NITI_GradientConv_Int8.cpp:218-268 and the GEMM unit compute/Int8FunctionsOpt.cpp:201-232 of the
// reference.  int32 accumulation == the reference's float accumulation while |partial| < 2^24 (Q1).
#include "common.cuh"
#include "tc_ptx.cuh"
#include <climits>

namespace {

constexpr int BM = 128;          // accumulator rows per CTA (TMEM lanes)
constexpr int BK = 128;          // bytes of K per stage row (K-major) / pixel rows per stage (MN-major)
constexpr int STAGES = 3;
constexpr int kProducerThreads = 128;
constexpr int kThreads = 160;    // 4 producer/epilogue warps + 1 MMA warp
constexpr int A_STAGE_BYTES = BM * 128;

struct TcParams {
    const int8_t* A;             // fwd: x      dx: dy     dw: dy
    const int8_t* B;             // fwd: w      dx: wT     dw: x
    int32_t* acc;
    uint32_t* maxabs;
    int M, Ncols, Ktotal;        // D is M x Ncols, reduction length Ktotal (bytes, K-major) / pixels (dw)
    int src_h, src_w, src_c;     // gathered tensor (fwd: x, dx: dy, dw: x)
    int out_h, out_w;            // pixel space that indexes the gather (fwd/dw: oh,ow; dx: h,w)
    int kh, kw, stride, pad, dil;
    int mode;                    // 0 forward-style gather, 1 dX-style gather
    int oc, taps, c_tiles, kb_per_split, splits;   // dw only
};

template <int BN> constexpr int smem_bytes() {
    return STAGES * (A_STAGE_BYTES + BN * 128) + 1024 /*align slack*/ + 256 /*barriers*/ + BM * 12 /*row info*/;
}

__device__ __forceinline__ uint32_t swz(int row, int chunk) { return (uint32_t)(row * 128 + ((chunk ^ (row & 7)) << 4)); }

// ------------------------------------------------------------------------------------------------
// K-major implicit GEMM (forward and dX)
// ------------------------------------------------------------------------------------------------
template <int BN>
__global__ void __launch_bounds__(kThreads) igemm_kmajor_kernel(const TcParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;
    uint8_t* sB = smem + STAGES * A_STAGE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + STAGES * BN * 128);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* accum_bar = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
    int* row_y = reinterpret_cast<int*>(tmem_slot + 2);
    int* row_x = row_y + BM;
    int* row_b = row_x + BM;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int num_kb = (p.Ktotal + BK - 1) / BK;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&full_bar[s], kProducerThreads); tc::mbar_init(&empty_bar[s], 1); }
        tc::mbar_init(accum_bar, 1);
        tc::fence_barrier_init();
    }
    if (warp == 4) tc::tmem_alloc<BN>(tmem_slot);
    if (tid < BM) {
        int m = m0 + tid;
        int y0 = INT_MIN / 4, x0 = INT_MIN / 4, base = 0;
        if (m < p.M) {
            int px = m % p.out_w; int t = m / p.out_w;
            int py = t % p.out_h; int b = t / p.out_h;
            if (p.mode == 0) { y0 = py * p.stride - p.pad; x0 = px * p.stride - p.pad; }
            else             { y0 = py + p.pad;            x0 = px + p.pad; }
            base = b * p.src_h * p.src_w;
        }
        row_y[tid] = y0; row_x[tid] = x0; row_b[tid] = base;
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 4) {
        // ------------------------------ producers ------------------------------
        const int j = tid & 7, r0 = tid >> 3;
        for (int kb = 0; kb < num_kb; ++kb) {
            const int s = kb % STAGES; const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
            tc::mbar_wait(&empty_bar[s], ph ^ 1u);
            const int k = kb * BK + j * 16;
            const bool kvalid = k < p.Ktotal;
            const int tap = k / p.src_c, ch = k - tap * p.src_c;
            const int ky = tap / p.kw, kx = tap - ky * p.kw;
            const uint32_t a_stage = tc::smem_u32(sA + s * A_STAGE_BYTES);
            const uint32_t b_stage = tc::smem_u32(sB + s * BN * 128);
#pragma unroll
            for (int i = 0; i < BM / 16; ++i) {
                const int r = r0 + 16 * i;
                int sy, sx; bool ok = kvalid;
                if (p.mode == 0) {
                    sy = row_y[r] + ky * p.dil; sx = row_x[r] + kx * p.dil;
                } else {
                    int ty = row_y[r] - ky * p.dil, tx = row_x[r] - kx * p.dil;
                    ok = ok && ty >= 0 && tx >= 0;
                    if (p.stride > 1) { ok = ok && (ty % p.stride == 0) && (tx % p.stride == 0); sy = ty / p.stride; sx = tx / p.stride; }
                    else { sy = ty; sx = tx; }
                }
                ok = ok && (unsigned)sy < (unsigned)p.src_h && (unsigned)sx < (unsigned)p.src_w;
                const int8_t* src = ok ? p.A + ((size_t)(row_b[r] + sy * p.src_w + sx) * p.src_c + ch) : p.A;
                tc::cp_async16(a_stage + swz(r, j), src, ok ? 16u : 0u);
            }
#pragma unroll
            for (int i = 0; i < BN / 16; ++i) {
                const int r = r0 + 16 * i;
                const int n = n0 + r;
                const bool ok = kvalid && n < p.Ncols;
                const int8_t* src = ok ? p.B + ((size_t)n * p.Ktotal + k) : p.B;
                tc::cp_async16(b_stage + swz(r, j), src, ok ? 16u : 0u);
            }
            tc::cp_async_mbar_arrive_noinc(&full_bar[s]);
        }
        // ------------------------------ epilogue ------------------------------
        tc::mbar_wait(accum_bar, 0);
        tc::tc_fence_after();
        const int m = m0 + warp * 32 + lane;
        uint32_t mx = 0;
        for (int c0 = 0; c0 < BN; c0 += 32) {
            if (n0 + c0 >= p.Ncols) break;                       // warp-uniform
            uint32_t v[32];
            tc::tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
            tc::tmem_ld_wait();
            if (m < p.M) {
                int32_t* dst = p.acc + (size_t)m * p.Ncols + n0 + c0;
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
        if (p.maxabs != nullptr) {
            mx = warp_max_u32(mx);
            if (lane == 0 && mx) atomicMax(p.maxabs, mx);
        }
    } else if (lane == 0) {
        // ------------------------------ MMA issuer ------------------------------
        constexpr uint32_t idesc = tc::make_idesc_i8(BM, BN, 0, 0);
        for (int kb = 0; kb < num_kb; ++kb) {
            const int s = kb % STAGES; const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
            tc::mbar_wait(&full_bar[s], ph);
            tc::tc_fence_after();
            tc::fence_proxy_async_smem();
            const uint32_t a_base = tc::smem_u32(sA + s * A_STAGE_BYTES);
            const uint32_t b_base = tc::smem_u32(sB + s * BN * 128);
#pragma unroll
            for (int kk = 0; kk < BK / 32; ++kk) {
                uint64_t ad = tc::make_smem_desc_sw128(a_base + kk * 32, 16, 1024);
                uint64_t bd = tc::make_smem_desc_sw128(b_base + kk * 32, 16, 1024);
                tc::umma_i8(tmem_base, ad, bd, idesc, (uint32_t)((kb | kk) != 0));
            }
            tc::umma_commit(&empty_bar[s]);
        }
        tc::umma_commit(accum_bar);
    }
    __syncwarp();
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 4) tc::tmem_dealloc<BN>(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// MN-major GEMM for dW: reduction over pixels, operands read as [pixel rows][channel bytes]
// grid.x = oc_tiles * taps * c_tiles, grid.y = pixel splits (split-K; integer atomics keep it exact)
// ------------------------------------------------------------------------------------------------
template <int BN>
__global__ void __launch_bounds__(kThreads) dw_mnmajor_kernel(const TcParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;
    uint8_t* sB = smem + STAGES * A_STAGE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + STAGES * BN * 128);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* accum_bar = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ct = blockIdx.x % p.c_tiles;
    const int tap = (blockIdx.x / p.c_tiles) % p.taps;
    const int ot = blockIdx.x / (p.c_tiles * p.taps);
    const int oc0 = ot * BM, c0 = ct * BN;
    const int ky = tap / p.kw, kx = tap - ky * p.kw;
    const int num_kb_total = (p.Ktotal + BK - 1) / BK;          // Ktotal = pixels
    const int kb_begin = blockIdx.y * p.kb_per_split;
    const int kb_end = min(num_kb_total, kb_begin + p.kb_per_split);
    const int num_kb = kb_end - kb_begin;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&full_bar[s], kProducerThreads); tc::mbar_init(&empty_bar[s], 1); }
        tc::mbar_init(accum_bar, 1);
        tc::fence_barrier_init();
    }
    if (warp == 4) tc::tmem_alloc<BN>(tmem_slot);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 4) {
        const int j = tid & 7, r0 = tid >> 3;
        const bool a_col_ok = oc0 + j * 16 < p.oc;
        for (int it = 0; it < num_kb; ++it) {
            const int s = it % STAGES; const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
            tc::mbar_wait(&empty_bar[s], ph ^ 1u);
            const uint32_t a_stage = tc::smem_u32(sA + s * A_STAGE_BYTES);
            const uint32_t b_stage = tc::smem_u32(sB + s * BN * 128);
            const int pix0 = (kb_begin + it) * BK;
#pragma unroll
            for (int i = 0; i < BK / 16; ++i) {
                const int r = r0 + 16 * i;
                const int pix = pix0 + r;
                const bool pv = pix < p.Ktotal;
                // A: dy[pix][oc0 + 16 j ..]
                {
                    const bool ok = pv && a_col_ok;
                    const int8_t* src = ok ? p.A + ((size_t)pix * p.oc + oc0 + j * 16) : p.A;
                    tc::cp_async16(a_stage + swz(r, j), src, ok ? 16u : 0u);
                }
                // B: x[pix shifted by tap][c0 + 16 (j + 8 h) ..], one 128-byte slab per 128 columns
                int ox = pix % p.out_w; int t = pix / p.out_w;
                int oy = t % p.out_h; int b = t / p.out_h;
                int sy = oy * p.stride - p.pad + ky * p.dil, sx = ox * p.stride - p.pad + kx * p.dil;
                const bool inb = pv && (unsigned)sy < (unsigned)p.src_h && (unsigned)sx < (unsigned)p.src_w;
                const size_t xoff = ((size_t)(b * p.src_h + sy) * p.src_w + sx) * p.src_c;
#pragma unroll
                for (int hcol = 0; hcol < BN / 128; ++hcol) {
                    const int c = c0 + hcol * 128 + j * 16;
                    const bool ok = inb && c < p.src_c;
                    const int8_t* src = ok ? p.B + xoff + c : p.B;
                    tc::cp_async16(b_stage + hcol * (BK * 128) + swz(r, j), src, ok ? 16u : 0u);
                }
            }
            tc::cp_async_mbar_arrive_noinc(&full_bar[s]);
        }
        // epilogue: rows = oc, columns = c of this tap
        tc::mbar_wait(accum_bar, 0);
        tc::tc_fence_after();
        const int o = oc0 + warp * 32 + lane;
        for (int cc = 0; cc < BN; cc += 32) {
            if (c0 + cc >= p.src_c) break;
            uint32_t v[32];
            tc::tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)cc, v);
            tc::tmem_ld_wait();
            if (o < p.oc) {
                int32_t* dst = p.acc + ((size_t)o * p.taps + tap) * p.src_c + c0 + cc;
                if (p.splits > 1) {
#pragma unroll
                    for (int q = 0; q < 32; ++q)
                        if (c0 + cc + q < p.src_c && v[q] != 0u) atomicAdd(dst + q, (int)v[q]);
                } else {
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        if (c0 + cc + q * 4 < p.src_c)
                            *reinterpret_cast<int4*>(dst + q * 4) = make_int4((int)v[q * 4], (int)v[q * 4 + 1], (int)v[q * 4 + 2], (int)v[q * 4 + 3]);
                }
            }
        }
    } else if (lane == 0) {
        constexpr uint32_t idesc = tc::make_idesc_i8(BM, BN, 1, 1);
        for (int it = 0; it < num_kb; ++it) {
            const int s = it % STAGES; const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
            tc::mbar_wait(&full_bar[s], ph);
            tc::tc_fence_after();
            tc::fence_proxy_async_smem();
            const uint32_t a_base = tc::smem_u32(sA + s * A_STAGE_BYTES);
            const uint32_t b_base = tc::smem_u32(sB + s * BN * 128);
#pragma unroll
            for (int kk = 0; kk < BK / 32; ++kk) {
                // MN-major, 128B swizzle: 8-row (K) groups every 1024 B (SBO); next 128 MN columns one slab on (LBO)
                uint64_t ad = tc::make_smem_desc_sw128(a_base + kk * 32 * 128, BK * 128, 1024);
                uint64_t bd = tc::make_smem_desc_sw128(b_base + kk * 32 * 128, BK * 128, 1024);
                tc::umma_i8(tmem_base, ad, bd, idesc, (uint32_t)((it | kk) != 0));
            }
            tc::umma_commit(&empty_bar[s]);
        }
        tc::umma_commit(accum_bar);
    }
    __syncwarp();
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 4) tc::tmem_dealloc<BN>(tmem_base);
}

// OHWI [oc][tap][c] -> [c][tap][oc] so dX's B operand is K-major over (tap, oc)
__global__ void __launch_bounds__(256) transpose_w_kernel(const int8_t* __restrict__ w, int8_t* __restrict__ wt, int oc, int taps, int c) {
    const size_t total = (size_t)oc * taps * c;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int o = (int)(i % oc); size_t t = i / oc;
        int tp = (int)(t % taps); int ch = (int)(t / taps);
        wt[i] = w[((size_t)o * taps + tp) * c + ch];
    }
}

// dense issue-rate probe: every CTA re-multiplies one resident smem tile; no global traffic
__global__ void __launch_bounds__(kThreads) probe_kernel(int iters) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + A_STAGE_BYTES + 256 * 128);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < (A_STAGE_BYTES + 256 * 128) / 4; i += kThreads) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u;
    if (tid == 0) { tc::mbar_init(bar, 1); tc::fence_barrier_init(); }
    if (warp == 4) tc::tmem_alloc<256>(tmem_slot);
    tc::fence_proxy_async_smem();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (warp == 4 && lane == 0) {
        constexpr uint32_t idesc = tc::make_idesc_i8(128, 256, 0, 0);
        const uint32_t a_base = tc::smem_u32(smem), b_base = tc::smem_u32(smem + A_STAGE_BYTES);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
                tc::umma_i8(tmem_base, tc::make_smem_desc_sw128(a_base + kk * 32, 16, 1024),
                            tc::make_smem_desc_sw128(b_base + kk * 32, 16, 1024), idesc, 1u);
        }
        tc::umma_commit(bar);
        tc::mbar_wait(bar, 0);
    }
    __syncwarp();
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 4) tc::tmem_dealloc<256>(tmem_base);
}

template <int BN>
int launch_kmajor(niti_ctx* ctx, const TcParams& p, cudaStream_t st) {
    if (int rc_cfg = niti_func_smem(ctx, (const void*)igemm_kmajor_kernel<BN>, (int)(smem_bytes<BN>()), "cudaFuncSetAttribute(igemm_kmajor)")) return rc_cfg;
    dim3 grid((unsigned)((p.M + BM - 1) / BM), (unsigned)((p.Ncols + BN - 1) / BN));
    igemm_kmajor_kernel<BN><<<grid, kThreads, smem_bytes<BN>(), st>>>(p);
    return niti_check_launch(ctx, "igemm_kmajor");
}

int dispatch_kmajor(niti_ctx* ctx, const TcParams& p, cudaStream_t st) {
    const long mt = (p.M + BM - 1) / BM;
    if (p.Ncols <= 64) return launch_kmajor<64>(ctx, p, st);
    if (p.Ncols <= 128) return launch_kmajor<128>(ctx, p, st);
    const long t256 = mt * ((p.Ncols + 255) / 256);
    if (t256 >= ctx->sm_count) return launch_kmajor<256>(ctx, p, st);
    return launch_kmajor<128>(ctx, p, st);
}
} // namespace

bool tc_fwd_eligible(const niti_conv_desc* d) { return d->c % 16 == 0 && d->oc % 4 == 0; }
bool tc_dx_eligible(const niti_conv_desc* d) { return d->oc % 16 == 0 && d->c % 4 == 0; }
bool tc_dw_eligible(const niti_conv_desc* d) { return d->oc % 16 == 0 && d->c % 16 == 0; }

int tc_conv_fwd_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, int32_t* acc,
                    uint32_t* maxabs, cudaStream_t st) {
    TcParams p{};
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    p.A = x; p.B = w; p.acc = acc; p.maxabs = maxabs;
    p.M = d->n * oh * ow; p.Ncols = d->oc; p.Ktotal = d->kh * d->kw * d->c;
    p.src_h = d->h; p.src_w = d->w; p.src_c = d->c; p.out_h = oh; p.out_w = ow;
    p.kh = d->kh; p.kw = d->kw; p.stride = d->stride; p.pad = d->pad; p.dil = d->dil; p.mode = 0;
    return dispatch_kmajor(ctx, p, st);
}

// OHWI -> [c][tap][oc] into the context's int8 workspace (the K-major B operand of dX)
int tc_transpose_w(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* w, int8_t** wt_out, cudaStream_t st) {
    const int taps = d->kh * d->kw;
    size_t total = (size_t)d->oc * taps * d->c;
    int8_t* wt; int rc = niti_ws_i8(ctx, total, &wt); if (rc) return rc;
    int blocks = (int)((total + 255) / 256); if (blocks > ctx->sm_count * 8) blocks = ctx->sm_count * 8;
    transpose_w_kernel<<<blocks, 256, 0, st>>>(w, wt, d->oc, taps, d->c);
    if ((rc = niti_check_launch(ctx, "transpose_w"))) return rc;
    *wt_out = wt;
    return NITI_OK;
}

int tc_conv_dx_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* w, int32_t* acc,
                   uint32_t* maxabs, cudaStream_t st) {
    const int oh = conv_out_h(d), ow = conv_out_w(d), taps = d->kh * d->kw;
    int8_t* wt; int rc = tc_transpose_w(ctx, d, w, &wt, st); if (rc) return rc;
    TcParams p{};
    p.A = dy; p.B = wt; p.acc = acc; p.maxabs = maxabs;
    p.M = d->n * d->h * d->w; p.Ncols = d->c; p.Ktotal = taps * d->oc;
    p.src_h = oh; p.src_w = ow; p.src_c = d->oc; p.out_h = d->h; p.out_w = d->w;
    p.kh = d->kh; p.kw = d->kw; p.stride = d->stride; p.pad = d->pad; p.dil = d->dil; p.mode = 1;
    return dispatch_kmajor(ctx, p, st);
}

int tc_conv_dw_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* dy, int32_t* acc,
                   uint32_t* maxabs, cudaStream_t st) {
    constexpr int BN = 128;
    if (int rc_cfg = niti_func_smem(ctx, (const void*)dw_mnmajor_kernel<BN>, (int)(smem_bytes<BN>()), "cudaFuncSetAttribute(dw_mnmajor)")) return rc_cfg;
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    TcParams p{};
    p.A = dy; p.B = x; p.acc = acc; p.maxabs = nullptr;
    p.oc = d->oc; p.taps = d->kh * d->kw; p.c_tiles = (d->c + BN - 1) / BN;
    p.M = d->oc; p.Ncols = d->c; p.Ktotal = d->n * oh * ow;
    p.src_h = d->h; p.src_w = d->w; p.src_c = d->c; p.out_h = oh; p.out_w = ow;
    p.kh = d->kh; p.kw = d->kw; p.stride = d->stride; p.pad = d->pad; p.dil = d->dil; p.mode = 0;
    const int tiles = ((d->oc + BM - 1) / BM) * p.taps * p.c_tiles;
    const int num_kb = (p.Ktotal + BK - 1) / BK;
    int splits = (2 * ctx->sm_count + tiles - 1) / tiles;        // aim at two resident CTAs per SM
    if (splits > num_kb) splits = num_kb;
    if (splits < 1) splits = 1;
    p.kb_per_split = (num_kb + splits - 1) / splits;
    splits = (num_kb + p.kb_per_split - 1) / p.kb_per_split;
    p.splits = splits;
    const size_t nout = (size_t)d->oc * p.taps * d->c;
    if (splits > 1) {
        cudaError_t e = cudaMemsetAsync(acc, 0, nout * sizeof(int32_t), st);
        if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "dw memset", e);
    }
    dim3 grid((unsigned)tiles, (unsigned)splits);
    dw_mnmajor_kernel<BN><<<grid, kThreads, smem_bytes<BN>(), st>>>(p);
    int rc = niti_check_launch(ctx, "dw_mnmajor");
    if (rc) return rc;
    if (maxabs) return niti_absmax(ctx, acc, nout, maxabs, st);
    return NITI_OK;
}

int tc_probe_i8_peak(niti_ctx* ctx, int iters, double* tops) {
    const int smem = A_STAGE_BYTES + 256 * 128 + 1024 + 64;
    cudaError_t e = cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "probe attr", e);
    cudaEvent_t t0, t1; cudaEventCreate(&t0); cudaEventCreate(&t1);
    // every SM of the device, whatever the data-parallel reserve: the roofline denominator must not depend on N
    probe_kernel<<<ctx->sm_total, kThreads, smem>>>(iters / 4 + 1);      // warm-up
    cudaEventRecord(t0);
    probe_kernel<<<ctx->sm_total, kThreads, smem>>>(iters);
    cudaEventRecord(t1);
    e = cudaEventSynchronize(t1);
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "probe run", e);
    float ms = 0; cudaEventElapsedTime(&ms, t0, t1);
    cudaEventDestroy(t0); cudaEventDestroy(t1);
    ctx->launches += 2;
    double ops = 2.0 * 128 * 256 * 128 * (double)iters * ctx->sm_total;
    *tops = ops / (ms * 1e-3) / 1e12;
    return NITI_OK;
}
