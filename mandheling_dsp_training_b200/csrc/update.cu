// update.cu — the weight-update tail of the int8 training step as ONE launch for the whole net.
//
// Per conv layer the reference runs (after the dW contraction): NITI_RangeEstimate over the dW accumulator,
// PSTO(acc, b-2) (NITI_GradientConv_Int8.cpp:274-296), then NITI_SGD's int8 Binary SUB
// W <- clip127(W - dW) (tools/train/source/optimizer/NITI_SGD.hpp:20-54, BinaryUtils.hpp:276-299).
// Done layer by layer that is 2 launches per layer of a few microseconds of HBM work each (VGG-11: 22 launches, 9.2 M
// weights); here every layer is a "segment" of one persistent kernel:
//   phase 1  segmented max|acc|  -> at most one atomicMax per block per segment
//   grid barrier (all CTAs resident: grid = SMs x occupancy)
//   phase 2  rule B + SGD on 16x128 tiles of the [oc][taps*c] weight matrix; the tile is also written transposed into
//            the [c][taps][oc] copy that the dX kernels use as their K-major B operand, so no per-step weight
//            transposes remain.
// HBM-bound: 4n (acc, phase 1; L2-resident for phase 2) + n (W read) + 2n (W, Wt write) bytes for n weights.
#include "common.cuh"

namespace {

constexpr int kUpdThreads = 256;
constexpr int kMaxSegs = 56;

struct UpdSegs {
    niti_update_seg s[kMaxSegs];
    uint32_t vec_begin[kMaxSegs + 1];        // prefix sums of phase-1 units (16-byte vectors of the accumulators)
    uint32_t tile_begin[kMaxSegs + 1];       // prefix sums of phase-2 units (tiles)
    int n;
};

__device__ __forceinline__ void grid_barrier_upd(uint32_t* counter, uint32_t target) { niti_grid_barrier(counter, target); }

__device__ __forceinline__ uint32_t pack4u(int32_t a, int32_t b, int32_t c, int32_t d) {
    return (uint32_t)(a & 0xff) | ((uint32_t)(b & 0xff) << 8) | ((uint32_t)(c & 0xff) << 16) | ((uint32_t)(d & 0xff) << 24);
}

__device__ __forceinline__ int find_seg(const uint32_t* begin, int n, uint32_t u) {
    int lo = 0, hi = n;                       // begin[lo] <= u < begin[hi]
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (begin[mid] <= u) lo = mid; else hi = mid; }
    return lo;
}

// 16-row x 128-column vector tiles need 16-byte accumulator rows, 4-byte weight rows and 16-byte runs of the transposed copy
__host__ __device__ __forceinline__ bool upd_vec_ok(const niti_update_seg& g) { return (g.c & 3) == 0 && (g.oc & 15) == 0; }

// rule B for one value: b == 0 -> 0, else PSTO(b - sub); then the SGD step against w.  kFast: b != 0 and shift >= 2 (decided
// once per tile — both are uniform over a segment), the branch-free path of common.cuh niti_psto_fast
template <bool kFast>
__device__ __forceinline__ int32_t upd_one(int32_t acc, int32_t w, int b, int shift) {
    const int32_t d = kFast ? niti_psto_fast(acc, shift) : (b == 0 ? 0 : (int32_t)(int8_t)niti_psto(acc, shift));
    const int32_t u = w - d;
    return u > 127 ? 127 : (u < -127 ? -127 : u);
}

// 16 rows x 4 columns of one lane (see phase 2)
template <bool kFast>
__device__ __forceinline__ void upd_vec_tile(const int32_t* ap, int8_t* wp, int cols, int b, int shift, uint32_t (&o)[4][4]) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        int4 av[8]; uint32_t wv[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {                    // eight rows in flight
            av[k] = __ldcg(reinterpret_cast<const int4*>(ap + (h * 8 + k) * cols));
            wv[k] = *reinterpret_cast<const uint32_t*>(wp + (h * 8 + k) * cols);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int32_t n0 = upd_one<kFast>(av[k].x, (int32_t)(int8_t)(wv[k] & 0xff), b, shift);
            const int32_t n1 = upd_one<kFast>(av[k].y, (int32_t)(int8_t)((wv[k] >> 8) & 0xff), b, shift);
            const int32_t n2 = upd_one<kFast>(av[k].z, (int32_t)(int8_t)((wv[k] >> 16) & 0xff), b, shift);
            const int32_t n3 = upd_one<kFast>(av[k].w, (int32_t)(int8_t)(wv[k] >> 24), b, shift);
            *reinterpret_cast<uint32_t*>(wp + (h * 8 + k) * cols) = pack4u(n0, n1, n2, n3);
            const int qd = h * 2 + (k >> 2), sh = 8 * (k & 3);
            if ((k & 3) == 0) { o[0][qd] = 0; o[1][qd] = 0; o[2][qd] = 0; o[3][qd] = 0; }
            o[0][qd] |= (uint32_t)(n0 & 0xff) << sh; o[1][qd] |= (uint32_t)(n1 & 0xff) << sh;
            o[2][qd] |= (uint32_t)(n2 & 0xff) << sh; o[3][qd] |= (uint32_t)(n3 & 0xff) << sh;
        }
    }
}

// mode 0: both phases around the grid barrier (persistent, every CTA resident); mode 1: phase 1 only; mode 2: phase 2 only —
// the two-launch form holds no barrier, so it may run beside other kernels of the context (niti_update_batched_phases)
__global__ void __launch_bounds__(kUpdThreads)
update_batched_kernel(const __grid_constant__ UpdSegs S, uint32_t* sync, unsigned long long* tl, int mode) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    niti_pdl_trigger();
    niti_pdl_wait();                                         // programmatic dependent launch: see common.cuh
    niti_tl_enter(tl);

    // ---------------- phase 1: max|acc| per segment ----------------
    if (mode != 2)
    // The accumulators of all segments form one concatenated run of 16-byte vectors; every WARP takes one contiguous slice of
    // it (1-2 segments), keeps eight loads per lane in flight and publishes with at most one atomic per segment it touched —
    // no block-level reduction, so nothing serialises the load latencies of different segments.  (Same-address atomics from
    // thousands of warps would serialise in L2: the atomic is skipped when the word already holds a value at least as large.)
    {
        const uint32_t nvec_all = S.vec_begin[S.n];
        const uint32_t nw = gridDim.x * (kUpdThreads / 32), wid = blockIdx.x * (kUpdThreads / 32) + warp;
        const uint32_t per = (nvec_all + nw - 1) / nw;
        uint32_t g = wid * per;
        const uint32_t g1 = g + per < nvec_all ? g + per : nvec_all;
        while (g < g1) {
            const int si = find_seg(S.vec_begin, S.n, g);
            const niti_update_seg& sg = S.s[si];
            const uint32_t e = S.vec_begin[si + 1] < g1 ? S.vec_begin[si + 1] : g1;
            const int4* v = reinterpret_cast<const int4*>(sg.acc) - S.vec_begin[si];
            uint32_t m = 0;
            for (uint32_t i = g + lane; i < e; i += 32 * 8) {
                int4 q[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) q[k] = i + 32 * k < e ? __ldcg(v + i + 32 * k) : make_int4(0, 0, 0, 0);
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    m = max(m, max(max(niti_abs_u32(q[k].x), niti_abs_u32(q[k].y)), max(niti_abs_u32(q[k].z), niti_abs_u32(q[k].w))));
            }
            m = warp_max_u32(m);
            if (lane == 0 && m > __ldcg(sg.maxabs)) atomicMax(sg.maxabs, m);
            g = e;
        }
        if (blockIdx.x == 0 && warp == 0) {                      // element counts that are not a multiple of 4: the last 1-3 words
            for (int si = 0; si < S.n; ++si) {
                const niti_update_seg& sg = S.s[si];
                const size_t n = (size_t)sg.oc * sg.taps * sg.c;
                uint32_t m = lane < (int)(n & 3) ? niti_abs_u32(sg.acc[(n & ~(size_t)3) + lane]) : 0u;
                m = warp_max_u32(m);
                if (lane == 0 && m) atomicMax(sg.maxabs, m);
            }
        }
    }

    // ---------------- grid barrier (self re-arming) ----------------
    if (mode == 1) { niti_tl_exit(tl); return; }
    if (mode == 0) {
        __syncthreads();
        if (tid == 0) grid_barrier_upd(sync, gridDim.x);
        __syncthreads();
    }

    // ---------------- phase 2: rule B + SGD (+ transposed copy) ----------------
    const uint32_t ntiles = S.tile_begin[S.n];
    const uint32_t nwarps = gridDim.x * (kUpdThreads / 32);
    for (uint32_t u = blockIdx.x * (kUpdThreads / 32) + warp; u < ntiles; u += nwarps) {
        const int si = find_seg(S.tile_begin, S.n, u);
        const niti_update_seg& sg = S.s[si];
        const int cols = sg.taps * sg.c, oc = sg.oc;
        const uint32_t t = u - S.tile_begin[si];
        const int b = niti_bitwidth(__ldcg(sg.maxabs));
        const int shift = b - sg.sub;
        if (upd_vec_ok(sg)) {
            // one warp = 16 rows x 128 columns of the [oc][taps*c] weight matrix; a lane owns 4 consecutive columns: one
            // 16-byte accumulator load and one 4-byte weight load / store per row, and after 16 rows it holds 16 consecutive
            // bytes of 4 rows of the transposed copy wt = [c][taps][oc] (one 16-byte store each)
            const int ctiles = (cols + 127) >> 7;
            const int r0 = (int)(t / ctiles) << 4, col = ((int)(t % ctiles) << 7) + lane * 4;
            if (col >= cols) continue;                            // cols % 4 == 0: a lane's four columns are in or out together
            const int32_t* ap = sg.acc + (size_t)r0 * cols + col;
            int8_t* wp = sg.w + (size_t)r0 * cols + col;
            uint32_t o[4][4];                                     // [column][row quad]
            if (b != 0 && shift >= 2) upd_vec_tile<true>(ap, wp, cols, b, shift, o);
            else upd_vec_tile<false>(ap, wp, cols, b, shift, o);
            if (sg.wt != nullptr) {
                const int tap = col / sg.c, ci = col - tap * sg.c;   // c % 4 == 0: the four columns share the tap
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    *reinterpret_cast<uint4*>(sg.wt + ((size_t)(ci + j) * sg.taps + tap) * oc + r0) = make_uint4(o[j][0], o[j][1], o[j][2], o[j][3]);
            }
            continue;
        }
        // general shapes: 32x32 tile, lanes along the columns, row by row
        const int ctiles = (cols + 31) >> 5;
        const int r0 = (int)(t / ctiles) << 5, col = ((int)(t % ctiles) << 5) + lane;
        if (col >= cols) continue;
        const int rows = min(32, oc - r0);
        const int tap = col / sg.c, ci = col - tap * sg.c;
        for (int r = 0; r < rows; ++r) {
            const size_t off = (size_t)(r0 + r) * cols + col;
            const int32_t nv = upd_one<false>(__ldcg(sg.acc + off), (int32_t)sg.w[off], b, shift);
            sg.w[off] = (int8_t)nv;
            if (sg.wt != nullptr) sg.wt[((size_t)ci * sg.taps + tap) * oc + r0 + r] = (int8_t)nv;
        }
    }

    // ---------------- last CTA out re-arms the barrier for the next launch ----------------
    __syncthreads();
    if (tid == 0 && mode == 0) {
        if (atomicAdd(sync + 1, 1u) == gridDim.x - 1) { sync[0] = 0u; sync[1] = 0u; __threadfence(); }
    }
    niti_tl_exit(tl);
}

} // namespace

static int update_launch(niti_ctx* ctx, const niti_update_seg* segs, int nseg, cudaStream_t st, bool phases) {
    if (!ctx || !segs || nseg <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_update_batched: bad argument");
    // occupancy is a property of the device: cached on the context (one context = one device), not process-wide
    if (ctx->upd_blocks_per_sm <= 0) {
        int occ = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, update_batched_kernel, kUpdThreads, 0);
        if (e != cudaSuccess || occ < 1) return niti_fail(ctx, NITI_ERR_CUDA, "niti_update_batched: occupancy query", e);
        ctx->upd_blocks_per_sm = occ > 4 ? 4 : occ;
    }
    const int blocks_per_sm = ctx->upd_blocks_per_sm;
    for (int first = 0; first < nseg; first += kMaxSegs) {
        UpdSegs S;
        S.n = nseg - first < kMaxSegs ? nseg - first : kMaxSegs;
        uint32_t tiles = 0, vecs = 0;
        for (int i = 0; i < S.n; ++i) {
            const niti_update_seg& g = segs[first + i];
            if (!g.acc || !g.w || !g.maxabs || g.oc <= 0 || g.taps <= 0 || g.c <= 0)
                return niti_fail(ctx, NITI_ERR_ARG, "niti_update_batched: bad segment");
            if ((reinterpret_cast<uintptr_t>(g.acc) & 15u) || (reinterpret_cast<uintptr_t>(g.w) & 15u) ||
                (g.wt && (reinterpret_cast<uintptr_t>(g.wt) & 15u)))
                return niti_fail(ctx, NITI_ERR_ARG, "niti_update_batched: buffers must be 16-byte aligned");
            S.s[i] = g;
            S.tile_begin[i] = tiles; S.vec_begin[i] = vecs;
            vecs += (uint32_t)(((size_t)g.oc * g.taps * g.c) >> 2);
            const size_t cols = (size_t)g.taps * g.c;
            tiles += upd_vec_ok(g) ? (uint32_t)((g.oc / 16) * ((cols + 127) / 128)) : (uint32_t)(((g.oc + 31) / 32) * ((cols + 31) / 32));
        }
        S.tile_begin[S.n] = tiles; S.vec_begin[S.n] = vecs;
        const int grid = ctx->sm_count * blocks_per_sm;            // mode 0: all CTAs co-resident, the grid barrier needs it
        if (!phases) {
            cudaError_t lerr = niti_launch(ctx, NITI_PDL_UPDATE, update_batched_kernel, dim3(grid), dim3(kUpdThreads), 0, st, S, ctx->ws_u32 + 40, niti_tl_next(ctx, "update_batched"), 0);
            if (lerr != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch update_batched", lerr);
            if (int rc = niti_check_launch(ctx, "update_batched")) return rc;
            continue;
        }
        for (int mode = 1; mode <= 2; ++mode) {                    // plain stream-ordered launches: no barrier, no residency rule
            update_batched_kernel<<<grid, kUpdThreads, 0, st>>>(S, ctx->ws_u32 + 40, niti_tl_next(ctx, mode == 1 ? "update_max" : "update_apply"), mode);
            if (int rc = niti_check_launch(ctx, "update_batched (phases)")) return rc;
        }
    }
    return NITI_OK;
}

extern "C" int niti_update_batched(niti_ctx* ctx, const niti_update_seg* segs, int nseg, niti_stream st) {
    return update_launch(ctx, segs, nseg, (cudaStream_t)st, false);
}
extern "C" int niti_update_batched_phases(niti_ctx* ctx, const niti_update_seg* segs, int nseg, niti_stream st) {
    return update_launch(ctx, segs, nseg, (cudaStream_t)st, true);
}
