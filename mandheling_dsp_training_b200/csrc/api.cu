// api.cu — context, workspace and the C-ABI dispatch of libniti_b200.so (see include/niti_b200.h).
#include <vector>
#include "common.cuh"
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <new>

// kernel families (conv_simt.cu / conv_tc.cu)
int simt_conv_fwd_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t);
int simt_conv_dx_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t);
int simt_conv_dw_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t);
bool tc_fwd_eligible(const niti_conv_desc*);
bool tc_dx_eligible(const niti_conv_desc*);
bool tc_dw_eligible(const niti_conv_desc*);
int tc_conv_fwd_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t);
int tc_conv_dx_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t);
int tc_conv_dw_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t);
int tc_probe_i8_peak(niti_ctx*, int iters, double* tops);
int tc_transpose_w(niti_ctx*, const niti_conv_desc*, const int8_t* w, int8_t** wt_out, cudaStream_t);
bool tma_fwd_eligible(const niti_conv_desc*);
bool tma_dx_eligible(const niti_conv_desc*);
bool tma_dw_eligible(const niti_conv_desc*);
int tma_conv_fwd_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t);
int tma_conv_dx_acc(niti_ctx*, const niti_conv_desc*, const int8_t* dy, const int8_t* wt, int32_t*, uint32_t*, cudaStream_t);
int tma_conv_dw_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t);
// conv_tma2.cu: persistent CTA-pair (cta_group::2) generation of the forward / dX kernels
bool tma2_fwd_eligible(const niti_conv_desc*);
bool tma2_dx_eligible(const niti_conv_desc*);
int tma2_conv_fwd_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t);
int tma2_conv_dx_acc(niti_ctx*, const niti_conv_desc*, const int8_t* dy, const int8_t* wt, int32_t*, uint32_t*, cudaStream_t);

// conv_dw2.cu: persistent, tap-grouped, TMA reduce-add generation of the dW kernel
bool dw2_eligible(const niti_conv_desc*);
int dw2_conv_dw_acc(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int32_t*, uint32_t*, cudaStream_t,
                    int accumulate = 0);

// conv_stem.cu: single-launch forward op (conv + exponent + rescale) of <= 16-channel stem layers
bool stem_fwd_eligible(const niti_conv_desc*);
bool stem2_fwd_eligible(const niti_conv_desc*, int c_live, bool pool);
int stem2_conv_fwd_fused(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, const int8_t*, const int8_t*, int8_t*,
                         int8_t*, int, int8_t*, uint32_t*, int, cudaStream_t);
int stem_conv_fwd_fused(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, const int8_t*, const int8_t*, int8_t*,
                        int8_t*, int, uint32_t*, int, cudaStream_t);

// conv_tma2.cu, single-phase fused ops (accumulators stay in TMEM across the exponent barrier)
bool fused_fwd_eligible(const niti_ctx*, const niti_conv_desc*);
bool fused_dx_eligible(const niti_ctx*, const niti_conv_desc*);
int fused_conv_fwd(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, const int8_t*, const int8_t*, int8_t*, int8_t*,
                   int, uint32_t*, int, cudaStream_t);
int fused_conv_dx(niti_ctx*, const niti_conv_desc*, const int8_t*, const int8_t*, int8_t*, const int8_t*, uint32_t*, int, cudaStream_t);

// conv_chain.cu: the layer-walking fused kernel; a chain of one op is the single-op fused entry point
int chain_run_single(niti_ctx*, const niti_chain_op*, cudaStream_t);
bool chain_op_eligible(niti_ctx*, const niti_chain_op*);

// kernel family selection: TMA-fed tcgen05 > cp.async-fed tcgen05 > SIMT, subject to shape eligibility
// and the context's forced path.  Returns 0 simt, 1 tc (cp.async), 2 tma, 3 tma2 (CTA pair), -1 unsupported under the
// forced path.
static int pick_path(niti_ctx* ctx, bool tma_ok, bool tc_ok, bool tma2_ok = false) {
    switch (ctx->path) {
        case NITI_PATH_SIMT: return 0;
        case NITI_PATH_TC: return tc_ok ? 1 : -1;
        case NITI_PATH_TMA: return tma_ok ? 2 : -1;
        case NITI_PATH_TMA2: return tma2_ok ? 3 : -1;
        default: return (tma2_ok && ctx->auto_tma2) ? 3 : (tma_ok ? 2 : (tc_ok ? 1 : 0));
    }
}
static const char* path_name(int p) { return p == 3 ? "tma2" : (p == 2 ? "tma" : (p == 1 ? "tc" : "simt")); }

int niti_fail(niti_ctx* ctx, int code, const char* what, cudaError_t e) {
    if (ctx) {
        if (e != cudaSuccess) snprintf(ctx->err, sizeof(ctx->err), "%s: %s", what, cudaGetErrorString(e));
        else snprintf(ctx->err, sizeof(ctx->err), "%s", what);
    }
    return code;
}

int niti_func_smem(niti_ctx* ctx, const void* fn, int bytes, const char* what) {
    for (int i = 0; i < ctx->n_smem_cfg; ++i)
        if (ctx->smem_cfg[i].fn == fn) {
            if (ctx->smem_cfg[i].bytes >= bytes) return NITI_OK;
            cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
            if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, what, e);
            ctx->smem_cfg[i].bytes = bytes;
            return NITI_OK;
        }
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, what, e);
    if (ctx->n_smem_cfg < 64) { ctx->smem_cfg[ctx->n_smem_cfg].fn = fn; ctx->smem_cfg[ctx->n_smem_cfg].bytes = bytes; ctx->n_smem_cfg++; }
    return NITI_OK;
}

int niti_check_launch(niti_ctx* ctx, const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, what, e);
    ctx->launches++;
    return NITI_OK;
}

// Workspaces only grow, and only outside stream capture (cudaMalloc is not capturable): callers
// that capture graphs warm up once eagerly, as the Python step and the MNN backend both do.
int niti_ws_acc(niti_ctx* ctx, size_t bytes, int32_t** out) {
    if (bytes > ctx->ws_acc_bytes) {
        if (ctx->ws_acc) cudaFree(ctx->ws_acc);
        ctx->ws_acc = nullptr; ctx->ws_acc_bytes = 0;
        size_t want = bytes + bytes / 4;
        cudaError_t e = cudaMalloc(&ctx->ws_acc, want);
        if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "workspace (acc) cudaMalloc", e);
        ctx->ws_acc_bytes = want;
    }
    *out = ctx->ws_acc;
    return NITI_OK;
}
int niti_ws_i8(niti_ctx* ctx, size_t bytes, int8_t** out) {
    if (bytes > ctx->ws_i8_bytes) {
        if (ctx->ws_i8) cudaFree(ctx->ws_i8);
        ctx->ws_i8 = nullptr; ctx->ws_i8_bytes = 0;
        size_t want = bytes + bytes / 4;
        cudaError_t e = cudaMalloc(&ctx->ws_i8, want);
        if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "workspace (i8) cudaMalloc", e);
        ctx->ws_i8_bytes = want;
    }
    *out = ctx->ws_i8;
    return NITI_OK;
}

// zero-inserted dY of strided dX ops (own buffer: the i8 workspace above may be in use for weights in the same call chain)
static int niti_ws_z(niti_ctx* ctx, size_t bytes, int8_t** out) {
    if (bytes > ctx->ws_z_bytes) {
        if (ctx->ws_z) cudaFree(ctx->ws_z);
        ctx->ws_z = nullptr; ctx->ws_z_bytes = 0;
        cudaError_t e = cudaMalloc(&ctx->ws_z, bytes);
        if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "workspace (zero-inserted dY) cudaMalloc", e);
        ctx->ws_z_bytes = bytes;
    }
    *out = ctx->ws_z;
    return NITI_OK;
}

static bool desc_ok(const niti_conv_desc* d) {
    if (!d) return false;
    if (d->n <= 0 || d->c <= 0 || d->h <= 0 || d->w <= 0 || d->oc <= 0 || d->kh <= 0 || d->kw <= 0) return false;
    if (d->stride <= 0 || d->dil <= 0 || d->pad < 0) return false;
    return conv_out_h(d) > 0 && conv_out_w(d) > 0;
}

extern "C" {

int niti_version(void) { return 100; }

int niti_ctx_create(int device, niti_ctx** out) {
    if (!out) return NITI_ERR_ARG;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || device < 0 || device >= count) return NITI_ERR_CUDA;   // no CPU fallback
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return NITI_ERR_CUDA;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return NITI_ERR_CUDA;
    niti_ctx* ctx = new (std::nothrow) niti_ctx();
    if (!ctx) return NITI_ERR_CUDA;
    memset(ctx, 0, sizeof(*ctx));
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->sm_total = prop.multiProcessorCount;
    ctx->path = NITI_PATH_AUTO;
    const char* env = getenv("NITI_B200_PATH");
    if (env && !strcmp(env, "simt")) ctx->path = NITI_PATH_SIMT;
    if (env && !strcmp(env, "tc")) ctx->path = NITI_PATH_TC;
    if (env && !strcmp(env, "tma")) ctx->path = NITI_PATH_TMA;
    if (env && !strcmp(env, "tma2")) ctx->path = NITI_PATH_TMA2;
    const char* a2 = getenv("NITI_B200_AUTO_TMA2");
    ctx->auto_tma2 = a2 ? atoi(a2) : 1;
    const char* uc = getenv("NITI_B200_CHAIN");           // 0: the round-1 per-shape templates (igemm_fused_kernel) for A/B runs
    ctx->use_chain = uc ? atoi(uc) : 1;
    const char* pdl = getenv("NITI_B200_PDL");
    ctx->pdl = pdl ? atoi(pdl) : (1 | 512);      // fused kernels, trigger at the exponent barrier
    ctx->last_path = "none";
    if (prop.major != 10) {
        snprintf(ctx->err, sizeof(ctx->err), "device is sm_%d%d; this library is built for sm_100a only", prop.major, prop.minor);
        delete ctx;
        return NITI_ERR_UNSUPPORTED;
    }
    e = cudaMalloc(&ctx->ws_u32, (64 + kPredWords) * sizeof(uint32_t));
    if (e != cudaSuccess) { delete ctx; return NITI_ERR_CUDA; }
    cudaMemset(ctx->ws_u32, 0, (64 + kPredWords) * sizeof(uint32_t));
    ctx->pred = ctx->ws_u32 + 64;
    {   // chooser bits start at 2: the input-scale predictor leads until the window predictor proves better (common.cuh)
        std::vector<uint32_t> init(kDpSlots, 2u << 16);
        cudaMemcpy(ctx->pred + kDpSlots, init.data(), kDpSlots * sizeof(uint32_t), cudaMemcpyHostToDevice);
    }
    const char* dwc = getenv("NITI_B200_DW_CTAS");        // CTAs a dW launch may spread its split-K over (default: every SM)
    ctx->dw_ctas = dwc ? atoi(dwc) : 0;
    const char* dwg = getenv("NITI_B200_DW_G");
    ctx->dw_g = dwg ? atoi(dwg) : 0;
    const char* sp = getenv("NITI_B200_SPECULATE");        // 0: two-sweep schedule (max pass | barrier | store pass) for A/B runs
    ctx->speculate = sp ? atoi(sp) : 1;
    if (ctx->speculate < 0 || ctx->speculate > 2) ctx->speculate = 1;
    const char* sc = getenv("NITI_B200_SPEC_CONFIDENT");
    ctx->spec_confident = sc ? atoi(sc) : 0;
    *out = ctx;
    return NITI_OK;
}

int niti_ctx_destroy(niti_ctx* ctx) {
    if (!ctx) return NITI_ERR_ARG;
    cudaSetDevice(ctx->device);
    if (ctx->ws_acc) cudaFree(ctx->ws_acc);
    if (ctx->ws_i8) cudaFree(ctx->ws_i8);
    if (ctx->ws_z) cudaFree(ctx->ws_z);
    if (ctx->ws_u32) cudaFree(ctx->ws_u32);
    niti_dp_close(ctx);
    if (ctx->dp_buf) cudaFree(ctx->dp_buf);
    delete ctx;
    return NITI_OK;
}

const char* niti_last_error(niti_ctx* ctx) { return ctx ? ctx->err : "null context"; }
int niti_set_path(niti_ctx* ctx, int path) {
    if (!ctx || path < NITI_PATH_AUTO || path > NITI_PATH_TMA2) return NITI_ERR_ARG;
    ctx->path = path; return NITI_OK;
}
int niti_set_speculate(niti_ctx* ctx, int on) {
    if (!ctx) return NITI_ERR_ARG;
    if (on < 0 || on > 2) return NITI_ERR_ARG;
    ctx->speculate = on; return NITI_OK;
}
int niti_speculate_stats(niti_ctx* ctx, unsigned long long* hits, unsigned long long* misses, unsigned long long* skipped) {
    if (!ctx || !hits || !misses) return NITI_ERR_ARG;
    uint32_t v[3] = {0, 0, 0};
    cudaError_t e = cudaMemcpy(v, ctx->pred + 2 * kDpSlots, sizeof(v), cudaMemcpyDeviceToHost);     // synchronises the device
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "niti_speculate_stats", e);
    *hits = v[0]; *misses = v[1];
    if (skipped) *skipped = v[2];
    return NITI_OK;
}
int64_t niti_launch_count(niti_ctx* ctx) { return ctx ? ctx->launches : -1; }
const char* niti_last_path(niti_ctx* ctx) { return ctx ? ctx->last_path : "none"; }

// ---- phase 1 ------------------------------------------------------------------------------------
int niti_conv_fwd_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, int32_t* acc,
                      uint32_t* maxabs, niti_stream st) {
    if (!ctx || !desc_ok(d) || !x || !w || !acc) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_fwd_acc: bad argument");
    const int path = pick_path(ctx, tma_fwd_eligible(d), tc_fwd_eligible(d), tma2_fwd_eligible(d));
    if (path < 0) return niti_fail(ctx, NITI_ERR_UNSUPPORTED, "conv_fwd: shape not eligible for the forced kernel path");
    ctx->last_path = path_name(path);
    if (path == 3) return tma2_conv_fwd_acc(ctx, d, x, w, acc, maxabs, (cudaStream_t)st);
    if (path == 2) return tma_conv_fwd_acc(ctx, d, x, w, acc, maxabs, (cudaStream_t)st);
    if (path == 1) return tc_conv_fwd_acc(ctx, d, x, w, acc, maxabs, (cudaStream_t)st);
    return simt_conv_fwd_acc(ctx, d, x, w, acc, maxabs, (cudaStream_t)st);
}

int niti_conv_dx_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* w, int32_t* acc,
                     uint32_t* maxabs, niti_stream st) {
    if (!ctx || !desc_ok(d) || !dy || !w || !acc) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dx_acc: bad argument");
    const int path = pick_path(ctx, tma_dx_eligible(d), tc_dx_eligible(d), tma2_dx_eligible(d));
    if (path < 0) return niti_fail(ctx, NITI_ERR_UNSUPPORTED, "conv_dx: shape not eligible for the forced kernel path");
    ctx->last_path = path_name(path);
    if (path == 3) {
        int8_t* wt; int rc = tc_transpose_w(ctx, d, w, &wt, (cudaStream_t)st); if (rc) return rc;
        return tma2_conv_dx_acc(ctx, d, dy, wt, acc, maxabs, (cudaStream_t)st);
    }
    if (path == 2) {
        int8_t* wt; int rc = tc_transpose_w(ctx, d, w, &wt, (cudaStream_t)st); if (rc) return rc;
        return tma_conv_dx_acc(ctx, d, dy, wt, acc, maxabs, (cudaStream_t)st);
    }
    if (path == 1) return tc_conv_dx_acc(ctx, d, dy, w, acc, maxabs, (cudaStream_t)st);
    return simt_conv_dx_acc(ctx, d, dy, w, acc, maxabs, (cudaStream_t)st);
}

int niti_conv_dx_acc_wt(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* w, const int8_t* wt,
                        int32_t* acc, uint32_t* maxabs, niti_stream st) {
    if (!wt) return niti_conv_dx_acc(ctx, d, dy, w, acc, maxabs, st);
    if (!ctx || !desc_ok(d) || !dy || !w || !acc) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dx_acc_wt: bad argument");
    const int path = pick_path(ctx, tma_dx_eligible(d), tc_dx_eligible(d), tma2_dx_eligible(d));
    if (path == 3) { ctx->last_path = path_name(path); return tma2_conv_dx_acc(ctx, d, dy, wt, acc, maxabs, (cudaStream_t)st); }
    if (path == 2) { ctx->last_path = path_name(path); return tma_conv_dx_acc(ctx, d, dy, wt, acc, maxabs, (cudaStream_t)st); }
    return niti_conv_dx_acc(ctx, d, dy, w, acc, maxabs, st);
}

int niti_conv_dw_acc(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* dy, int32_t* acc,
                     uint32_t* maxabs, niti_stream st) {
    if (!ctx || !desc_ok(d) || !x || !dy || !acc) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dw_acc: bad argument");
    const int path = pick_path(ctx, tma_dw_eligible(d), tc_dw_eligible(d), dw2_eligible(d));
    if (path < 0) return niti_fail(ctx, NITI_ERR_UNSUPPORTED, "conv_dw: shape not eligible for the forced kernel path");
    ctx->last_path = path_name(path);
    if (path == 3) return dw2_conv_dw_acc(ctx, d, x, dy, acc, maxabs, (cudaStream_t)st);
    if (path == 2) return tma_conv_dw_acc(ctx, d, x, dy, acc, maxabs, (cudaStream_t)st);
    if (path == 1) return tc_conv_dw_acc(ctx, d, x, dy, acc, maxabs, (cudaStream_t)st);
    return simt_conv_dw_acc(ctx, d, x, dy, acc, maxabs, (cudaStream_t)st);
}

int niti_conv_dw_accumulate(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* dy, int32_t* acc,
                            niti_stream st) {
    if (!ctx || !desc_ok(d) || !x || !dy || !acc) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dw_accumulate: bad argument");
    const int path = pick_path(ctx, tma_dw_eligible(d), tc_dw_eligible(d), dw2_eligible(d));
    if (path == 3) { ctx->last_path = path_name(path); return dw2_conv_dw_acc(ctx, d, x, dy, acc, nullptr, (cudaStream_t)st, 1); }
    return niti_conv_dw_acc(ctx, d, x, dy, acc, nullptr, st);     // these kernels overwrite acc and hold no grid barrier
}

// C[e,h] = sum_l A[e,l] B[h,l]  ==  1x1 convolution with e "pixels", l channels, h outputs
int niti_matmul_nt_acc(niti_ctx* ctx, int e, int h, int l, const int8_t* a, const int8_t* b, int32_t* acc,
                       uint32_t* maxabs, niti_stream st) {
    if (!ctx || e <= 0 || h <= 0 || l <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_matmul_nt_acc: bad argument");
    niti_conv_desc d{e, l, 1, 1, h, 1, 1, 1, 0, 1};
    return niti_conv_fwd_acc(ctx, &d, a, b, acc, maxabs, st);
}

// ---- whole ops ------------------------------------------------------------------------------------
int niti_conv_fwd(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, const int8_t* ex,
                  const int8_t* ew, int8_t* y, int8_t* ey, niti_stream st) {
    if (!ctx || !desc_ok(d) || !y) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_fwd: bad argument");
    size_t n = (size_t)d->n * conv_out_h(d) * conv_out_w(d) * d->oc;
    int32_t* acc; int rc = niti_ws_acc(ctx, n * 4, &acc); if (rc) return rc;
    if ((rc = niti_zero_u32(ctx, ctx->ws_u32, 1, st))) return rc;
    if ((rc = niti_conv_fwd_acc(ctx, d, x, w, acc, ctx->ws_u32, st))) return rc;
    return niti_rescale_act(ctx, acc, n, ctx->ws_u32, y, ex, ew, ey, st);
}

int niti_conv_fwd_fused(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, const int8_t* ex,
                        const int8_t* ew, int8_t* y, int8_t* ey, int relu, int32_t* acc, uint32_t* maxabs, int dp_slot,
                        niti_stream st) {
    if (!ctx || !desc_ok(d) || !x || !w || !y || !maxabs) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_fwd_fused: bad argument");
    if (dp_slot < 0 || dp_slot >= kDpSlots) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_fwd_fused: slot out of range");
    if ((ctx->path == NITI_PATH_AUTO || ctx->path == NITI_PATH_TMA2) && ctx->auto_tma2 && stem_fwd_eligible(d)) {
        ctx->last_path = "stem";
        return stem_conv_fwd_fused(ctx, d, x, w, ex, ew, y, ey, relu, maxabs, dp_slot, (cudaStream_t)st);
    }
    if ((ctx->path == NITI_PATH_AUTO || ctx->path == NITI_PATH_TMA2) && ctx->auto_tma2) {
        niti_chain_op op{};
        op.d = *d; op.op = 0; op.in = x; op.w = w; op.ex = ex; op.ew = ew; op.ey = ey; op.out = y; op.relu = relu;
        op.maxabs = maxabs; op.dp_slot = dp_slot;
        if (ctx->use_chain && chain_op_eligible(ctx, &op)) { ctx->last_path = "fused"; return chain_run_single(ctx, &op, (cudaStream_t)st); }
        if (fused_fwd_eligible(ctx, d)) {
            ctx->last_path = "fused";
            return fused_conv_fwd(ctx, d, x, w, ex, ew, y, ey, relu, maxabs, dp_slot, (cudaStream_t)st);
        }
    }
    const size_t n = (size_t)d->n * conv_out_h(d) * conv_out_w(d) * d->oc;
    int rc;
    if (!acc && (rc = niti_ws_acc(ctx, n * 4, &acc))) return rc;
    if ((rc = niti_conv_fwd_acc(ctx, d, x, w, acc, maxabs, st))) return rc;
    return niti_rescale_act_dp(ctx, acc, n, maxabs, y, ex, ew, ey, relu, nullptr, dp_slot, st);
}

int niti_conv_dx_fused(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* w, const int8_t* wt, int8_t* dx,
                       const int8_t* gate, int32_t* acc, uint32_t* maxabs, int dp_slot, niti_stream st) {
    if (!ctx || !desc_ok(d) || !dy || !w || !dx || !maxabs) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dx_fused: bad argument");
    if (dp_slot < 0 || dp_slot >= kDpSlots) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dx_fused: slot out of range");
    if (wt && (ctx->path == NITI_PATH_AUTO || ctx->path == NITI_PATH_TMA2) && ctx->auto_tma2) {
        niti_chain_op op{};
        op.d = *d; op.op = 1; op.in = dy; op.w = wt; op.out = dx; op.gate = gate; op.maxabs = maxabs; op.dp_slot = dp_slot;
        if (ctx->use_chain && chain_op_eligible(ctx, &op)) { ctx->last_path = "fused"; return chain_run_single(ctx, &op, (cudaStream_t)st); }
        if (ctx->use_chain && d->stride > 1 && d->oc % 16 == 0) {
            // Strided conv: dX is the stride-1 dX of dY with stride-1 zeros inserted between its pixels — the lowering the
            // reference's grad rule emits (NITI_CPULeftPoolGrad_Int8 in front of NITI_DeConv_Int8,
            // tools/train/source/grad/NITI_Conv_Int8_Grad.cpp:90-106).  The zeros cost stride^2 times the MMAs, on the fused
            // TMA/tcgen05 kernel instead of the cp.async kernel (55-80 TOP/s on the ResNet transitions) + transposes + rescale.
            niti_conv_desc d1 = *d; d1.stride = 1;
            niti_chain_op op1 = op; op1.d = d1;
            const int zh = conv_out_h(&d1), zw = conv_out_w(&d1);
            if (chain_op_eligible(ctx, &op1)) {
                int8_t* z; int rcz = niti_ws_z(ctx, (size_t)d->n * zh * zw * d->oc, &z); if (rcz) return rcz;
                if ((rcz = niti_zero_insert_nhwc(ctx, dy, d->n, conv_out_h(d), conv_out_w(d), d->oc, d->stride, zh, zw, z, st))) return rcz;
                op1.in = z;
                ctx->last_path = "fused";
                return chain_run_single(ctx, &op1, (cudaStream_t)st);
            }
        }
        if (fused_dx_eligible(ctx, d)) {
            ctx->last_path = "fused";
            return fused_conv_dx(ctx, d, dy, wt, dx, gate, maxabs, dp_slot, (cudaStream_t)st);
        }
    }
    const size_t n = (size_t)d->n * d->h * d->w * d->c;
    int rc;
    if (!acc && (rc = niti_ws_acc(ctx, n * 4, &acc))) return rc;
    if ((rc = niti_conv_dx_acc_wt(ctx, d, dy, w, wt, acc, maxabs, st))) return rc;
    return niti_rescale_act_dp(ctx, acc, n, maxabs, dx, nullptr, nullptr, nullptr, 0, gate, dp_slot, st);
}

int niti_conv_fwd_fused_pool(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, const int8_t* ex,
                             const int8_t* ew, int8_t* y, int8_t* ey, int relu, int8_t* y_pool, int32_t* acc, uint32_t* maxabs,
                             int dp_slot, niti_stream st) {
    if (!ctx || !desc_ok(d) || !x || !w || !y || !y_pool || !maxabs) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_fwd_fused_pool: bad argument");
    if (dp_slot < 0 || dp_slot >= kDpSlots) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_fwd_fused_pool: slot out of range");
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    if (oh < 2 || ow < 2) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_fwd_fused_pool: output smaller than the pool window");
    if ((ctx->path == NITI_PATH_AUTO || ctx->path == NITI_PATH_TMA2) && ctx->auto_tma2 && ctx->use_chain) {
        niti_chain_op op{};
        op.d = *d; op.op = 0; op.in = x; op.w = w; op.ex = ex; op.ew = ew; op.ey = ey; op.out = y; op.relu = relu; op.out_pool = y_pool;
        op.maxabs = maxabs; op.dp_slot = dp_slot;
        if (chain_op_eligible(ctx, &op)) { ctx->last_path = "fused"; return chain_run_single(ctx, &op, (cudaStream_t)st); }
    }
    int rc = niti_conv_fwd_fused(ctx, d, x, w, ex, ew, y, ey, relu, acc, maxabs, dp_slot, st);
    if (rc) return rc;
    return niti_maxpool(ctx, y, d->n, oh, ow, d->oc, 2, 2, y_pool, st);
}

int niti_conv_stem_fwd(niti_ctx* ctx, const niti_conv_desc* d, int c_live, const int8_t* x, const int8_t* w, const int8_t* ex,
                       const int8_t* ew, int8_t* y, int8_t* ey, int relu, int8_t* y_pool, int32_t* acc, uint32_t* maxabs,
                       int dp_slot, niti_stream st) {
    if (!ctx || !desc_ok(d) || !x || !w || !y || !maxabs) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_stem_fwd: bad argument");
    if (dp_slot < 0 || dp_slot >= kDpSlots) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_stem_fwd: slot out of range");
    if (c_live < 1 || c_live > d->c) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_stem_fwd: c_live outside [1, c]");
    if ((ctx->path == NITI_PATH_AUTO || ctx->path == NITI_PATH_TMA2) && ctx->auto_tma2 && stem2_fwd_eligible(d, c_live, y_pool != nullptr)) {
        ctx->last_path = "stem2";
        return stem2_conv_fwd_fused(ctx, d, x, w, ex, ew, y, ey, relu, y_pool, maxabs, dp_slot, (cudaStream_t)st);
    }
    if (y_pool && (ctx->path == NITI_PATH_AUTO || ctx->path == NITI_PATH_TMA2) && ctx->auto_tma2 && stem2_fwd_eligible(d, c_live, false)) {
        // the pool does not fold (output width not a multiple of 16: MNIST's 24): stem kernel, then the pool kernel
        ctx->last_path = "stem2";
        int rc = stem2_conv_fwd_fused(ctx, d, x, w, ex, ew, y, ey, relu, nullptr, maxabs, dp_slot, (cudaStream_t)st);
        if (rc) return rc;
        return niti_maxpool(ctx, y, d->n, conv_out_h(d), conv_out_w(d), d->oc, 2, 2, y_pool, st);
    }
    if (y_pool) return niti_conv_fwd_fused_pool(ctx, d, x, w, ex, ew, y, ey, relu, y_pool, acc, maxabs, dp_slot, st);
    return niti_conv_fwd_fused(ctx, d, x, w, ex, ew, y, ey, relu, acc, maxabs, dp_slot, st);
}

int niti_conv_dx_fused_poolgrad(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* w, const int8_t* wt,
                                int8_t* dx_prepool, const int8_t* pool_x, int gate_relu, int8_t* dx_pooled, int32_t* acc,
                                uint32_t* maxabs, int dp_slot, niti_stream st) {
    if (!ctx || !desc_ok(d) || !dy || !w || !dx_prepool || !pool_x || !maxabs) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dx_fused_poolgrad: bad argument");
    if (dp_slot < 0 || dp_slot >= kDpSlots) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dx_fused_poolgrad: slot out of range");
    if (wt && (ctx->path == NITI_PATH_AUTO || ctx->path == NITI_PATH_TMA2) && ctx->auto_tma2 && ctx->use_chain) {
        niti_chain_op op{};
        op.d = *d; op.op = 1; op.in = dy; op.w = wt; op.out = dx_prepool; op.pool_x = pool_x; op.pool_gate = gate_relu;
        op.maxabs = maxabs; op.dp_slot = dp_slot;
        if (chain_op_eligible(ctx, &op)) { ctx->last_path = "fused"; return chain_run_single(ctx, &op, (cudaStream_t)st); }
    }
    if (!dx_pooled) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dx_fused_poolgrad: shape needs the unfused route and dx_pooled is NULL");
    int rc = niti_conv_dx_fused(ctx, d, dy, w, wt, dx_pooled, nullptr, acc, maxabs, dp_slot, st);
    if (rc) return rc;
    // pool_y is recomputed by the gradient kernel from pool_x; it needs the pooled forward tensor: take the window maximum
    return niti_maxpool_grad_from_x(ctx, pool_x, dx_pooled, d->n, 2 * d->h, 2 * d->w, d->c, dx_prepool, gate_relu, st);
}

int niti_conv_dx(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* dy, const int8_t* w, int8_t* dx, niti_stream st) {
    if (!ctx || !desc_ok(d) || !dx) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dx: bad argument");
    size_t n = (size_t)d->n * d->h * d->w * d->c;
    int32_t* acc; int rc = niti_ws_acc(ctx, n * 4, &acc); if (rc) return rc;
    if ((rc = niti_zero_u32(ctx, ctx->ws_u32, 1, st))) return rc;
    if ((rc = niti_conv_dx_acc(ctx, d, dy, w, acc, ctx->ws_u32, st))) return rc;
    return niti_rescale_act(ctx, acc, n, ctx->ws_u32, dx, nullptr, nullptr, nullptr, st);
}

int niti_conv_dw(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* dy, int8_t* dw, niti_stream st) {
    if (!ctx || !desc_ok(d) || !dw) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dw: bad argument");
    size_t n = (size_t)d->oc * d->kh * d->kw * d->c;
    int32_t* acc; int rc = niti_ws_acc(ctx, n * 4, &acc); if (rc) return rc;
    if ((rc = niti_zero_u32(ctx, ctx->ws_u32, 1, st))) return rc;
    if ((rc = niti_conv_dw_acc(ctx, d, x, dy, acc, ctx->ws_u32, st))) return rc;
    return niti_rescale_grad(ctx, acc, n, ctx->ws_u32, 2, dw, nullptr, st);
}

// slice count of NITI_GradientSplitConv_Int8 (NITI_GradientSplitConv_Int8.cpp:209-222): OC is only rounded up to a multiple
// of 32 when OC % 32 != 0 (otherwise the reference's align32 is 0 and nothing is split)
int niti_split_count(int oc, int oh, int ow) {
    if (oc <= 0 || oh <= 0 || ow <= 0) return 0;
    int align32 = oc % 32;
    if (align32 != 0) align32 = (oc / 32 + 1) * 32;
    const int high = 8192 / ow / oh;
    int real = oc;
    while (align32 > high) { align32 /= 2; real /= 2; }
    return real > 0 ? oc / real : 0;
}

int niti_conv_dw_split(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* dy, int slices, int8_t* dw,
                       niti_stream st) {
    if (!ctx || !desc_ok(d) || !dw || slices <= 0 || slices > 32 || d->oc % slices != 0)
        return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_dw_split: bad argument (1 <= slices <= 32, oc % slices == 0)");
    const size_t n = (size_t)d->oc * d->kh * d->kw * d->c, per = n / slices;
    if (per % 4 != 0) return niti_fail(ctx, NITI_ERR_UNSUPPORTED, "niti_conv_dw_split: slice size must be a multiple of 4 words");
    int32_t* acc; int rc = niti_ws_acc(ctx, n * 4, &acc); if (rc) return rc;
    if ((rc = niti_zero_u32(ctx, ctx->ws_u32, 32, st))) return rc;
    if ((rc = niti_conv_dw_acc(ctx, d, x, dy, acc, nullptr, st))) return rc;
    for (int s = 0; s < slices; ++s)                           // an OC slice is contiguous in [oc][kh][kw][c]
        if ((rc = niti_absmax(ctx, acc + (size_t)s * per, per, ctx->ws_u32 + s, st))) return rc;
    return niti_rescale_grad_split(ctx, acc, per, slices, ctx->ws_u32, 3, dw, st);
}

int niti_matmul_nt(niti_ctx* ctx, int e, int h, int l, const int8_t* a, const int8_t* b, int8_t* c, niti_stream st) {
    if (!ctx || e <= 0 || h <= 0 || l <= 0 || !c) return niti_fail(ctx, NITI_ERR_ARG, "niti_matmul_nt: bad argument");
    size_t n = (size_t)e * h;
    int32_t* acc; int rc = niti_ws_acc(ctx, n * 4, &acc); if (rc) return rc;
    if ((rc = niti_zero_u32(ctx, ctx->ws_u32, 1, st))) return rc;
    if ((rc = niti_matmul_nt_acc(ctx, e, h, l, a, b, acc, ctx->ws_u32, st))) return rc;
    return niti_rescale_grad(ctx, acc, n, ctx->ws_u32, 3, c, nullptr, st);
}

// ---- host-buffer convenience: the reference-facing call shape (host tensors in, host tensors out) --
int niti_conv_fwd_host(niti_ctx* ctx, const niti_conv_desc* d, const int8_t* x, const int8_t* w, int8_t ex, int8_t ew,
                       int8_t* y, int8_t* ey) {
    if (!ctx || !desc_ok(d) || !x || !w || !y) return niti_fail(ctx, NITI_ERR_ARG, "niti_conv_fwd_host: bad argument");
    size_t nx = (size_t)d->n * d->h * d->w * d->c, nw = (size_t)d->oc * d->kh * d->kw * d->c;
    size_t ny = (size_t)d->n * conv_out_h(d) * conv_out_w(d) * d->oc;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    int8_t* buf; int rc = niti_ws_i8(ctx, up(nx) + up(nw) + up(ny) + 256, &buf); if (rc) return rc;
    int8_t *dx = buf, *dw = buf + up(nx), *dyv = dw + up(nw), *dexp = dyv + up(ny);
    cudaError_t e;
    if ((e = cudaMemcpyAsync(dx, x, nx, cudaMemcpyHostToDevice, 0)) != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "H2D x", e);
    if ((e = cudaMemcpyAsync(dw, w, nw, cudaMemcpyHostToDevice, 0)) != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "H2D w", e);
    int8_t hexp[2] = {ex, ew};
    if ((e = cudaMemcpyAsync(dexp, hexp, 2, cudaMemcpyHostToDevice, 0)) != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "H2D exp", e);
    if ((rc = niti_conv_fwd(ctx, d, dx, dw, dexp, dexp + 1, dyv, dexp + 2, 0))) return rc;
    if ((e = cudaMemcpyAsync(y, dyv, ny, cudaMemcpyDeviceToHost, 0)) != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "D2H y", e);
    int8_t out_e = 0;
    if ((e = cudaMemcpyAsync(&out_e, dexp + 2, 1, cudaMemcpyDeviceToHost, 0)) != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "D2H exp", e);
    if ((e = cudaStreamSynchronize(0)) != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "sync", e);
    if (ey) *ey = out_e;
    return NITI_OK;
}

// ---- device memory for CUDA-free hosts (mnn_backend/ links no CUDA library) ------------------------
int niti_dev_alloc(niti_ctx* ctx, size_t bytes, void** out) {
    if (!ctx || !out) return niti_fail(ctx, NITI_ERR_ARG, "niti_dev_alloc");
    *out = nullptr;
    cudaError_t e = cudaMalloc(out, bytes ? bytes : 1);
    return e == cudaSuccess ? NITI_OK : niti_fail(ctx, NITI_ERR_CUDA, "niti_dev_alloc", e);
}
int niti_dev_free(niti_ctx* ctx, void* p) {
    if (!ctx) return NITI_ERR_ARG;
    cudaError_t e = cudaFree(p);
    return e == cudaSuccess ? NITI_OK : niti_fail(ctx, NITI_ERR_CUDA, "niti_dev_free", e);
}
int niti_dev_memset(niti_ctx* ctx, void* p, int value, size_t bytes, niti_stream st) {
    if (!ctx || (!p && bytes)) return niti_fail(ctx, NITI_ERR_ARG, "niti_dev_memset");
    cudaError_t e = cudaMemsetAsync(p, value, bytes, (cudaStream_t)st);
    return e == cudaSuccess ? NITI_OK : niti_fail(ctx, NITI_ERR_CUDA, "niti_dev_memset", e);
}
static int copy_kind(niti_ctx* ctx, void* dst, const void* src, size_t bytes, cudaMemcpyKind k, niti_stream st, const char* what) {
    if (!ctx || (bytes && (!dst || !src))) return niti_fail(ctx, NITI_ERR_ARG, what);
    if (!bytes) return NITI_OK;
    cudaError_t e = cudaMemcpyAsync(dst, src, bytes, k, (cudaStream_t)st);
    return e == cudaSuccess ? NITI_OK : niti_fail(ctx, NITI_ERR_CUDA, what, e);
}
int niti_copy_h2d(niti_ctx* ctx, void* d, const void* s, size_t n, niti_stream st) { return copy_kind(ctx, d, s, n, cudaMemcpyHostToDevice, st, "niti_copy_h2d"); }
int niti_copy_d2h(niti_ctx* ctx, void* d, const void* s, size_t n, niti_stream st) { return copy_kind(ctx, d, s, n, cudaMemcpyDeviceToHost, st, "niti_copy_d2h"); }
int niti_copy_d2d(niti_ctx* ctx, void* d, const void* s, size_t n, niti_stream st) { return copy_kind(ctx, d, s, n, cudaMemcpyDeviceToDevice, st, "niti_copy_d2d"); }
int niti_stream_sync(niti_ctx* ctx, niti_stream st) {
    if (!ctx) return NITI_ERR_ARG;
    cudaError_t e = cudaStreamSynchronize((cudaStream_t)st);
    return e == cudaSuccess ? NITI_OK : niti_fail(ctx, NITI_ERR_CUDA, "niti_stream_sync", e);
}

int niti_sync_error(niti_ctx* ctx, int* out) {
    if (!ctx || !out) return niti_fail(ctx, NITI_ERR_ARG, "niti_sync_error");
    uint32_t v = 0;
    cudaError_t e = cudaMemcpy(&v, ctx->ws_u32 + 63, sizeof(v), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "niti_sync_error", e);
    *out = (int)v;
    return NITI_OK;
}

int niti_debug_trace(niti_ctx* ctx, unsigned long long* device_buf) {
    if (!ctx) return NITI_ERR_ARG;
    ctx->trace = device_buf;
    return NITI_OK;
}

int niti_debug_timeline(niti_ctx* ctx, unsigned long long* device_buf, int capacity) {
    if (!ctx) return NITI_ERR_ARG;
    ctx->tl = device_buf; ctx->tl_cap = device_buf ? capacity : 0; ctx->tl_n = 0;
    return NITI_OK;
}
int niti_debug_timeline_count(niti_ctx* ctx) { return ctx ? ctx->tl_n : 0; }
const char* niti_debug_timeline_name(niti_ctx* ctx, int i) { return (ctx && i >= 0 && i < ctx->tl_n) ? ctx->tl_name[i] : ""; }

int niti_probe_i8_peak(niti_ctx* ctx, int iters, double* tops_out) {
    if (!ctx || !tops_out || iters <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_probe_i8_peak");
    return tc_probe_i8_peak(ctx, iters, tops_out);
}

} // extern "C"
