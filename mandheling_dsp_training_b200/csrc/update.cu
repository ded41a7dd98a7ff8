// update.cu — the weight-update tail of the int8 training step as ONE launch for the whole net.
//
// Per conv layer the reference runs (after the dW contraction): NITI_RangeEstimate over the dW accumulator,
// PSTO(acc, b-2) (NITI_GradientConv_Int8.cpp:274-296), then NITI_SGD's int8 Binary SUB
// W <- clip127(W - dW) (tools/train/source/optimizer/NITI_SGD.hpp:20-54, BinaryUtils.hpp:276-299).
// Done layer by layer that is 2 launches per layer of a few microseconds of HBM work each (VGG-11: 22 launches, 9.2 M
// weights); here every layer is a "segment" of one persistent kernel:
//   phase 1  segmented max|acc|  -> one atomicMax per block per segment chunk
//   grid barrier (all CTAs resident: grid = SMs x occupancy)
//   phase 2  rule B + SGD on 32x32 tiles of the [oc][taps*c] weight matrix; the tile is also written transposed into
//            the [c][taps][oc] copy that the dX kernels use as their K-major B operand, so no per-step weight
//            transposes remain.
// HBM-bound: 4n (acc, phase 1; L2-resident for phase 2) + n (W read) + 2n (W, Wt write) bytes for n weights.
#include "common.cuh"

namespace {

constexpr int kUpdThreads = 256;
constexpr int kMaxSegs = 56;
constexpr int kChunk = 8192;                 // phase-1 unit: int32 words per block iteration

struct UpdSegs {
    niti_update_seg s[kMaxSegs];
    uint32_t chunk_begin[kMaxSegs + 1];      // prefix sums of phase-1 units
    uint32_t tile_begin[kMaxSegs + 1];       // prefix sums of phase-2 units (32x32 tiles)
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

// rule B for one value: b == 0 -> 0, else PSTO(b - sub); then the SGD step against w
__device__ __forceinline__ int32_t upd_one(int32_t acc, int32_t w, int b, int shift) {
    const int32_t d = b == 0 ? 0 : (int32_t)(int8_t)niti_psto(acc, shift);
    const int32_t u = w - d;
    return u > 127 ? 127 : (u < -127 ? -127 : u);
}

__global__ void __launch_bounds__(kUpdThreads)
update_batched_kernel(const __grid_constant__ UpdSegs S, uint32_t* sync, unsigned long long* tl) {
    __shared__ uint32_t red[kUpdThreads / 32];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    niti_pdl_trigger();
    niti_pdl_wait();                                         // programmatic dependent launch: see common.cuh
    niti_tl_enter(tl);

    // ---------------- phase 1: max|acc| per segment ----------------
    const uint32_t nchunks = S.chunk_begin[S.n];
    for (uint32_t u = blockIdx.x; u < nchunks; u += gridDim.x) {
        const int si = find_seg(S.chunk_begin, S.n, u);
        const niti_update_seg& sg = S.s[si];
        const size_t n = (size_t)sg.oc * sg.taps * sg.c;
        const size_t off = (size_t)(u - S.chunk_begin[si]) * kChunk;
        const size_t cnt = n - off < (size_t)kChunk ? n - off : (size_t)kChunk;
        const int32_t* a = sg.acc + off;
        uint32_t m = 0;
        const size_t nvec = cnt >> 2;                             // acc + off is 16-byte aligned (kChunk % 4 == 0)
        const int4* v = reinterpret_cast<const int4*>(a);
        for (size_t i = tid; i < nvec; i += kUpdThreads) {
            const int4 q = v[i];
            m = max(m, max(max(niti_abs_u32(q.x), niti_abs_u32(q.y)), max(niti_abs_u32(q.z), niti_abs_u32(q.w))));
        }
        for (size_t i = (nvec << 2) + tid; i < cnt; i += kUpdThreads) m = max(m, niti_abs_u32(a[i]));
        m = warp_max_u32(m);
        if (lane == 0) red[warp] = m;
        __syncthreads();
        if (warp == 0) {
            m = lane < kUpdThreads / 32 ? red[lane] : 0u;
            m = warp_max_u32(m);
            if (lane == 0 && m) atomicMax(sg.maxabs, m);
        }
        __syncthreads();
    }

    // ---------------- grid barrier (self re-arming) ----------------
    __syncthreads();
    if (tid == 0) grid_barrier_upd(sync, gridDim.x);
    __syncthreads();

    // ---------------- phase 2: rule B + SGD (+ transposed copy), one 32x32 tile per warp ----------------
    // Lanes run along the COLUMNS of the [oc][taps*c] weight matrix: every row of the tile is one coalesced 128-byte read
    // of the int32 accumulators and one 32-byte read / write of the weights.  After 32 rows lane j holds column c0+j of the
    // updated tile, i.e. 32 consecutive bytes of row (c0+j) of the transposed copy wt = [c][taps][oc] — written as two
    // 16-byte stores, no shared-memory transpose.
    const uint32_t ntiles = S.tile_begin[S.n];
    const uint32_t nwarps = gridDim.x * (kUpdThreads / 32);
    for (uint32_t u = blockIdx.x * (kUpdThreads / 32) + warp; u < ntiles; u += nwarps) {
        const int si = find_seg(S.tile_begin, S.n, u);
        const niti_update_seg& sg = S.s[si];
        const int cols = sg.taps * sg.c, oc = sg.oc;
        const int ctiles = (cols + 31) >> 5;
        const uint32_t t = u - S.tile_begin[si];
        const int r0 = (int)(t / ctiles) << 5, c0 = (int)(t % ctiles) << 5;
        const int b = niti_bitwidth(__ldcg(sg.maxabs));
        const int shift = b - sg.sub;
        const int col = c0 + lane;
        const bool col_ok = col < cols;
        const int rows = min(32, oc - r0);
        const int32_t* ap = sg.acc + (size_t)r0 * cols + col;
        int8_t* wp = sg.w + (size_t)r0 * cols + col;
        uint32_t q[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) q[j] = 0u;
        if (rows == 32) {
#pragma unroll
            for (int r = 0; r < 32; r += 8) {
                int32_t av[8], wv[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {                  // eight independent rows in flight
                    av[k] = col_ok ? __ldcg(ap + (size_t)(r + k) * cols) : 0;
                    wv[k] = col_ok ? (int32_t)wp[(size_t)(r + k) * cols] : 0;
                }
                uint32_t lo = 0, hi = 0;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int32_t nv = upd_one(av[k], wv[k], b, shift);
                    if (col_ok) wp[(size_t)(r + k) * cols] = (int8_t)nv;
                    if (k < 4) lo |= (uint32_t)(nv & 0xff) << (8 * k); else hi |= (uint32_t)(nv & 0xff) << (8 * (k - 4));
                }
                q[r >> 2] = lo; q[(r >> 2) + 1] = hi;
            }
        } else {
            for (int r = 0; r < rows; ++r) {
                if (col_ok) {
                    const int32_t nv = upd_one(__ldcg(ap + (size_t)r * cols), (int32_t)wp[(size_t)r * cols], b, shift);
                    wp[(size_t)r * cols] = (int8_t)nv;
                    q[r >> 2] |= (uint32_t)(nv & 0xff) << (8 * (r & 3));
                }
            }
        }
        if (sg.wt != nullptr && col_ok) {
            const int tap = col / sg.c, ci = col - tap * sg.c;
            int8_t* dst = sg.wt + ((size_t)ci * sg.taps + tap) * oc + r0;
            if (rows == 32 && (oc & 15) == 0) {
                uint4* d4 = reinterpret_cast<uint4*>(dst);
                d4[0] = make_uint4(q[0], q[1], q[2], q[3]);
                d4[1] = make_uint4(q[4], q[5], q[6], q[7]);
            } else {
                for (int k = 0; k < rows; ++k) dst[k] = (int8_t)((q[k >> 2] >> (8 * (k & 3))) & 0xff);
            }
        }
    }

    // ---------------- last CTA out re-arms the barrier for the next launch ----------------
    __syncthreads();
    if (tid == 0) {
        if (atomicAdd(sync + 1, 1u) == gridDim.x - 1) { sync[0] = 0u; sync[1] = 0u; __threadfence(); }
    }
    niti_tl_exit(tl);
}

} // namespace

extern "C" int niti_update_batched(niti_ctx* ctx, const niti_update_seg* segs, int nseg, niti_stream st) {
    if (!ctx || !segs || nseg <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_update_batched: bad argument");
    int blocks_per_sm = 0;                                    // occupancy is a per-device property: no process-wide cache
    {
        int occ = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, update_batched_kernel, kUpdThreads, 0);
        if (e != cudaSuccess || occ < 1) return niti_fail(ctx, NITI_ERR_CUDA, "niti_update_batched: occupancy query", e);
        blocks_per_sm = occ > 4 ? 4 : occ;
    }
    for (int first = 0; first < nseg; first += kMaxSegs) {
        UpdSegs S;
        S.n = nseg - first < kMaxSegs ? nseg - first : kMaxSegs;
        uint32_t chunks = 0, tiles = 0;
        for (int i = 0; i < S.n; ++i) {
            const niti_update_seg& g = segs[first + i];
            if (!g.acc || !g.w || !g.maxabs || g.oc <= 0 || g.taps <= 0 || g.c <= 0)
                return niti_fail(ctx, NITI_ERR_ARG, "niti_update_batched: bad segment");
            if ((reinterpret_cast<uintptr_t>(g.acc) & 15u) || (reinterpret_cast<uintptr_t>(g.w) & 15u) ||
                (g.wt && (reinterpret_cast<uintptr_t>(g.wt) & 15u)))
                return niti_fail(ctx, NITI_ERR_ARG, "niti_update_batched: buffers must be 16-byte aligned");
            S.s[i] = g;
            S.chunk_begin[i] = chunks; S.tile_begin[i] = tiles;
            const size_t n = (size_t)g.oc * g.taps * g.c;
            chunks += (uint32_t)((n + kChunk - 1) / kChunk);
            tiles += (uint32_t)(((g.oc + 31) / 32) * (((size_t)g.taps * g.c + 31) / 32));
        }
        S.chunk_begin[S.n] = chunks; S.tile_begin[S.n] = tiles;
        const int grid = ctx->sm_count * blocks_per_sm;            // all CTAs co-resident: the grid barrier needs it
        cudaError_t lerr = niti_launch(ctx, NITI_PDL_UPDATE, update_batched_kernel, dim3(grid), dim3(kUpdThreads), 0, (cudaStream_t)st, S, ctx->ws_u32 + 40, niti_tl_next(ctx, "update_batched"));
        if (lerr != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch update_batched", lerr);
        int rc = niti_check_launch(ctx, "update_batched");
        if (rc) return rc;
    }
    return NITI_OK;
}
