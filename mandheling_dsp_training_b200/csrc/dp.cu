// dp.cu — host side of the peer-memory scalar exchange (dp.cuh): buffer allocation, CUDA IPC export/open, epoch bump.
// One process per GPU; the handles travel between processes through whatever the host already has
// (torch.distributed all_gather in step.py, any byte transport elsewhere) — the C ABI only sees 64-byte blobs.
#include "common.cuh"
#include <cstring>
#include <cstdlib>

namespace {
__global__ void dp_step_begin_kernel(DpBuffer* b) { b->epoch = b->epoch + 1u; }
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

extern "C" int niti_dp_close(niti_ctx* ctx) {
    if (!ctx) return NITI_ERR_ARG;
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
