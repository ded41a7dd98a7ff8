// dp.cu — host side of the peer-memory scalar exchange (dp.cuh): buffer allocation, CUDA IPC export/open, epoch bump.
// One process per GPU; the handles travel between processes through whatever the host already has
// (torch.distributed all_gather in step.py, any byte transport elsewhere) — the C ABI only sees 64-byte blobs.
#include "common.cuh"
#include <cstring>
#include <cstdlib>

namespace {
__global__ void dp_step_begin_kernel(DpBuffer* b) { b->epoch = b->epoch + 1u; }

constexpr int kArMaxThreads = 1024;

__device__ __forceinline__ void ar_flag_store(unsigned long long* dst, uint32_t epoch) {
    const unsigned long long w = (unsigned long long)epoch;
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(dst), "l"(w) : "memory");
}
// spin until `src` (local memory, written by a peer) carries `epoch`; false when the budget ran out
__device__ __forceinline__ bool ar_flag_wait(const unsigned long long* src, uint32_t epoch, long long timeout_cycles) {
    const long long t0 = clock64();
    for (;;) {
        unsigned long long w;
        asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(w) : "l"(src) : "memory");
        if ((uint32_t)w == epoch) return true;
        if (clock64() - t0 > timeout_cycles) return false;
        __nanosleep(200);
    }
}
__device__ __forceinline__ int4 ld_peer(const int4* p) {      // never from a stale L1 line: the words change every step
    int4 v;
    asm volatile("ld.volatile.global.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

// In-place SUM over ranks of the window words [off, off + n) — the reference's gradient is the sum over the whole batch
// (NITI_GradientConv_Int8.cpp:218-268 contracts over every sample), so a batch sharded over g GPUs adds its int32 dW
// accumulators before the shared exponent is taken.  Two-shot over NVLink peer memory in ONE launch:
//   arrive   every rank tells every peer "my dW kernels for this bucket are done" (stream order on the sender)
//   reduce   rank r owns slice r: loads it from all g windows (g-1 NVLink reads in flight per thread), adds, and
//            stores the sum into all g windows (element i is read and written by its owner only: safe in place)
//   complete the last CTA of a rank to finish tells every peer; the launch ends when every peer has told this rank, so
//            whatever follows in the stream sees the complete sum.
// Integer adds commute: bit-exact in any order.  ~16 CTAs on the SMs the contractions leave free (niti_dp_reserved_sms).
template <int W>                                             // world rounded up to 2, 4 or 8: 8 / W vectors per thread and trip,
__global__ void __launch_bounds__(kArMaxThreads)              // so eight 16-byte loads are in flight per thread at any world size
dp_allreduce_kernel(const DpView v, const DpWindow w, size_t off, size_t n, int slot, unsigned long long* tl) {
    __shared__ int last;
    niti_tl_enter(tl);
    DpBuffer* mine = v.peer[v.rank];
    const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(&mine->epoch);
    const int tid = threadIdx.x;
    if (blockIdx.x == 0 && tid < v.world) ar_flag_store(&v.peer[tid]->ar_in[slot][v.rank], epoch);
    if (tid < v.world && tid != v.rank && !ar_flag_wait(&mine->ar_in[slot][tid], epoch, v.timeout_cycles)) mine->error = 1u;
    __syncthreads();

    constexpr int U = 8 / W;
    const size_t nvec = n >> 2;                               // off and n are multiples of 4 words (host)
    const size_t per = (nvec + v.world - 1) / v.world;
    const size_t v0 = (size_t)v.rank * per, v1 = v0 + per < nvec ? v0 + per : nvec;
    int4* win[W];
#pragma unroll
    for (int r = 0; r < W; ++r) win[r] = r < v.world ? reinterpret_cast<int4*>(w.peer[r] + off) : nullptr;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = v0 + (size_t)blockIdx.x * blockDim.x + tid; i < v1; i += stride * U) {
        int4 q[U][W];
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int r = 0; r < W; ++r)
                q[u][r] = (r < v.world && i + u * stride < v1) ? ld_peer(win[r] + i + u * stride) : make_int4(0, 0, 0, 0);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (i + u * stride >= v1) break;
            int4 s = q[u][0];
#pragma unroll
            for (int r = 1; r < W; ++r) { s.x += q[u][r].x; s.y += q[u][r].y; s.z += q[u][r].z; s.w += q[u][r].w; }
#pragma unroll
            for (int r = 0; r < W; ++r) if (r < v.world) win[r][i + u * stride] = s;
        }
    }
    __threadfence_system();                                   // this thread's peer stores are visible before the count below
    __syncthreads();
    if (tid == 0) last = atomicAdd(&mine->ar_count[slot], 1u) == gridDim.x - 1;
    __syncthreads();
    if (last) {
        __threadfence_system();
        if (tid == 0) mine->ar_count[slot] = 0u;
        if (tid < v.world) ar_flag_store(&v.peer[tid]->ar_out[slot][v.rank], epoch);
        if (tid < v.world && tid != v.rank && !ar_flag_wait(&mine->ar_out[slot][tid], epoch, v.timeout_cycles)) mine->error = 1u;
        __syncthreads();
        niti_tl_exit(tl);                                     // launch timeline: the last CTA's exit = the sum is complete everywhere
    }
}
}

extern "C" int niti_dp_alloc(niti_ctx* ctx, unsigned char handle_out[64]) {
    if (!ctx || !handle_out) return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_alloc");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (!ctx->dp_buf) {
        cudaError_t e = cudaMalloc(&ctx->dp_buf, sizeof(DpBuffer));
        if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "niti_dp_alloc: cudaMalloc", e);
        cudaMemset(ctx->dp_buf, 0, sizeof(DpBuffer));
        cudaDeviceSynchronize();
    }
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, ctx->dp_buf);
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "niti_dp_alloc: cudaIpcGetMemHandle", e);
    memcpy(handle_out, &h, 64);
    return NITI_OK;
}

extern "C" int niti_dp_connect(niti_ctx* ctx, int rank, int world, const unsigned char* handles /* world x 64 */) {
    if (!ctx || !handles || world < 1 || world > kDpMaxWorld || rank < 0 || rank >= world || !ctx->dp_buf)
        return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_connect: bad argument (call niti_dp_alloc first; world <= 8)");
    DpView v{};
    v.world = world; v.rank = rank;
    {   // spin budget of one exchange: generous, because ordinary first-step skew between ranks (lazy module load, NCCL
        // communicator set-up, workspace growth on one rank) is seconds; a dead peer still surfaces through niti_dp_error
        const char* t = getenv("NITI_B200_DP_TIMEOUT_MS");
        const double ms = t ? atof(t) : 60000.0;
        int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device);
        if (khz <= 0) khz = 2000000;
        v.timeout_cycles = (long long)(ms * (double)khz);
        if (v.timeout_cycles < 1000) v.timeout_cycles = 1000;
    }
    for (int r = 0; r < world; ++r) {
        if (r == rank) { v.peer[r] = ctx->dp_buf; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + 64 * r, 64);
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "niti_dp_connect: cudaIpcOpenMemHandle", e);
        v.peer[r] = reinterpret_cast<DpBuffer*>(p);
    }
    ctx->dp = v;
    // leave a few SMs to the NCCL kernels of the dW all-reduce that runs concurrently with backward: the persistent
    // contraction kernels otherwise occupy every SM (1 CTA/SM, ~200 KB smem) and the collective cannot start until one exits
    if (world > 1) {
        const char* env = getenv("NITI_B200_DP_RESERVE_SMS");
        // Keep reserve >= NCCL's CTA cap (bench.py sets NCCL_MAX_CTAS=16): with fewer free SMs than NCCL CTAs the
        // collective's CTAs wait for SMs held by contraction kernels that are themselves spinning on a peer (max exchange,
        // grid barrier) — measured 4xB200 with reserve 8 / 16 NCCL CTAs: 51.7 ms per step instead of 1.0.
        const int reserve = env ? atoi(env) : 16;
        ctx->sm_count = ctx->sm_total;                    // recomputed from the device's SM count: connecting twice reserves once
        if (reserve >= 0 && reserve < ctx->sm_total / 2) ctx->sm_count -= reserve;
        ctx->sm_count &= ~1;                              // CTA pairs
        ctx->dp_reserve = ctx->sm_total - ctx->sm_count;
    }
    return NITI_OK;
}

// ---- gradient window: the int32 dW accumulators of the whole net live in library-allocated memory every peer maps ----
extern "C" int niti_dp_window_alloc(niti_ctx* ctx, size_t words, int32_t** ptr_out, unsigned char handle_out[64]) {
    if (!ctx || !ptr_out || !handle_out || words == 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_window_alloc");
    if (ctx->dp.world < 1 || !ctx->dp_buf) return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_window_alloc: call niti_dp_connect first");
    int32_t*& mine = ctx->dp_win.peer[ctx->dp.rank];
    if (mine) return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_window_alloc: a window exists (niti_dp_window_free first)");
    words = (words + 3) & ~(size_t)3;
    cudaError_t e = cudaMalloc(&mine, words * sizeof(int32_t));
    if (e != cudaSuccess) { mine = nullptr; return niti_fail(ctx, NITI_ERR_CUDA, "niti_dp_window_alloc: cudaMalloc", e); }
    cudaMemset(mine, 0, words * sizeof(int32_t));
    cudaDeviceSynchronize();
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, mine);
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "niti_dp_window_alloc: cudaIpcGetMemHandle", e);
    memcpy(handle_out, &h, 64);
    ctx->dp_win.words = words;
    *ptr_out = mine;
    return NITI_OK;
}

extern "C" int niti_dp_window_connect(niti_ctx* ctx, const unsigned char* handles /* world x 64 */) {
    if (!ctx || !handles || !ctx->dp_win.peer[ctx->dp.rank]) return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_window_connect: no local window");
    for (int r = 0; r < ctx->dp.world; ++r) {
        if (r == ctx->dp.rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + 64 * r, 64);
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "niti_dp_window_connect: cudaIpcOpenMemHandle", e);
        ctx->dp_win.peer[r] = reinterpret_cast<int32_t*>(p);
    }
    return NITI_OK;
}

extern "C" int niti_dp_window_free(niti_ctx* ctx) {
    if (!ctx) return NITI_ERR_ARG;
    for (int r = 0; r < ctx->dp.world; ++r) {
        if (!ctx->dp_win.peer[r]) continue;
        if (r == ctx->dp.rank) cudaFree(ctx->dp_win.peer[r]); else cudaIpcCloseMemHandle(ctx->dp_win.peer[r]);
    }
    ctx->dp_win = DpWindow{};
    return NITI_OK;
}

extern "C" int niti_dp_allreduce_i32(niti_ctx* ctx, size_t off, size_t n, int slot, niti_stream st) {
    if (!ctx) return NITI_ERR_ARG;
    if (ctx->dp.world <= 1) return NITI_OK;
    if (slot < 0 || slot >= kDpArSlots || (off & 3) || (n & 3) || off + n > ctx->dp_win.words)
        return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_allreduce_i32: slot out of range, or range not 4-word aligned / outside the window");
    for (int r = 0; r < ctx->dp.world; ++r)
        if (!ctx->dp_win.peer[r]) return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_allreduce_i32: window not connected");
    if (n == 0) return NITI_OK;
    // shape of the launch: by default one 1024-thread CTA on each SM the contractions leave free; NITI_B200_DP_AR_CTAS /
    // NITI_B200_DP_AR_THREADS override (small CTAs fit beside a resident contraction CTA: 8 K registers, no shared memory)
    // Its CTAs spin on peers, so every one of them must be resident on an SM the one-CTA-per-SM contraction grids do not
    // own: with no reserve (NITI_B200_DP_RESERVE_SMS=0) a contraction kernel that cannot become fully resident would stall
    // the peers' max exchange until the spin budget runs out.  Refuse instead; callers fall back to a library all-reduce.
    if (ctx->dp_reserve <= 0)
        return niti_fail(ctx, NITI_ERR_UNSUPPORTED, "niti_dp_allreduce_i32: no SMs reserved for it (NITI_B200_DP_RESERVE_SMS=0); "
                                                    "sum the window with the collective library instead");
    int ctas = ctx->dp_reserve, threads = kArMaxThreads;
    if (const char* e = getenv("NITI_B200_DP_AR_CTAS")) { const int c = atoi(e); if (c >= 1 && c <= 1024) ctas = c; }
    if (const char* e = getenv("NITI_B200_DP_AR_THREADS")) { const int c = atoi(e); if (c >= 32 && c <= kArMaxThreads && c % 32 == 0) threads = c; }
    const int resident = ctx->dp_reserve * (2048 / threads);      // CTAs of this size the reserved SMs hold at once
    if (ctas > resident) ctas = resident;
    const size_t per = ((n >> 2) + ctx->dp.world - 1) / ctx->dp.world;
    const size_t need = (per + threads - 1) / threads;
    if ((size_t)ctas > need) ctas = (int)(need ? need : 1);
    const cudaStream_t cs = (cudaStream_t)st;
    unsigned long long* tl = niti_tl_next(ctx, "dp_allreduce");
    if (ctx->dp.world <= 2) dp_allreduce_kernel<2><<<ctas, threads, 0, cs>>>(ctx->dp, ctx->dp_win, off, n, slot, tl);
    else if (ctx->dp.world <= 4) dp_allreduce_kernel<4><<<ctas, threads, 0, cs>>>(ctx->dp, ctx->dp_win, off, n, slot, tl);
    else dp_allreduce_kernel<8><<<ctas, threads, 0, cs>>>(ctx->dp, ctx->dp_win, off, n, slot, tl);
    return niti_check_launch(ctx, "dp_allreduce");
}

extern "C" int niti_dp_close(niti_ctx* ctx) {
    if (!ctx) return NITI_ERR_ARG;
    niti_dp_window_free(ctx);
    for (int r = 0; r < ctx->dp.world; ++r)
        if (r != ctx->dp.rank && ctx->dp.peer[r]) cudaIpcCloseMemHandle(ctx->dp.peer[r]);
    ctx->dp = DpView{};
    ctx->sm_count = ctx->sm_total;
    ctx->dp_reserve = 0;
    return NITI_OK;
}

extern "C" int niti_dp_step_begin(niti_ctx* ctx, niti_stream st) {
    if (!ctx || !ctx->dp_buf) return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_step_begin: no exchange buffer");
    dp_step_begin_kernel<<<1, 1, 0, (cudaStream_t)st>>>(ctx->dp_buf);
    return niti_check_launch(ctx, "dp_step_begin");
}

// SMs left free for the collective kernels that run beside the contractions (0 when not connected)
extern "C" int niti_dp_reserved_sms(niti_ctx* ctx) { return ctx ? ctx->dp_reserve : 0; }

extern "C" int niti_dp_error(niti_ctx* ctx, int* out) {
    if (!ctx || !out || !ctx->dp_buf) return niti_fail(ctx, NITI_ERR_ARG, "niti_dp_error");
    uint32_t v = 0;
    cudaError_t e = cudaMemcpy(&v, &ctx->dp_buf->error, sizeof(v), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "niti_dp_error", e);
    *out = (int)v;
    return NITI_OK;
}
