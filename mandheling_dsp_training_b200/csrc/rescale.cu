// rescale.cu — phase 2 of the NITI path: per-tensor max-abs, shared exponent, pseudo-stochastic
// shift-round to int8, and the int8 SGD update.  HBM-bound streaming kernels: 16 elements per
// thread per iteration (4 x 128-bit loads, 1 x 128-bit store), grid = multiple of the SM count.
//
// Reference behaviour restated (execution-engine/):
//   source/backend/cpu/compute/CommonOptFunction.cpp:1565-1627   RangeEstimate / PstoShiftInt32
//   source/backend/cpu/NITI_Conv_Int8.cpp:255-307                rule A + exponent output
//   source/backend/cpu/NITI_GradientConv_Int8.cpp:274-296        rule B (sub = 2)
//   source/backend/cpu/NITI_Matmul_Int8.cpp:216-228              rule B (sub = 3)
//   tools/train/source/optimizer/NITI_SGD.hpp:49-52 + BinaryUtils.hpp:276-299   w = clip127(w - dw)
#include "common.cuh"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ int4 ldg_stream(const int4* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ uint32_t pack4(int32_t a, int32_t b, int32_t c, int32_t d) {
    return (uint32_t)(a & 0xff) | ((uint32_t)(b & 0xff) << 8) | ((uint32_t)(c & 0xff) << 16) |
           ((uint32_t)(d & 0xff) << 24);
}

__global__ void __launch_bounds__(kThreads) absmax_kernel(const int32_t* __restrict__ acc, size_t n,
                                                          uint32_t* __restrict__ maxabs) {
    uint32_t m = 0;
    const size_t nvec = n / 4;
    const int4* v = reinterpret_cast<const int4*>(acc);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
        int4 a = ldg_stream(v + i);
        m = max(m, max(max(niti_abs_u32(a.x), niti_abs_u32(a.y)), max(niti_abs_u32(a.z), niti_abs_u32(a.w))));
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) m = max(m, niti_abs_u32(acc[nvec * 4 + threadIdx.x]));
    m = warp_max_u32(m);
    __shared__ uint32_t sm[kThreads / 32];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = threadIdx.x < kThreads / 32 ? sm[threadIdx.x] : 0u;
        m = warp_max_u32(m);
        if (threadIdx.x == 0 && m) atomicMax(maxabs, m);
    }
}

// mode 0: rule A (activations).  mode 1: rule B (gradients; `sub`), optional fused SGD.
// mode 2: the DSP-named op types' requantisation (hexagon_nn/2.10.1/hexagon/ops/src/niti_op_supernode_new.c:4614-4634,
//         src/quantize.c:337-400): shift = ceiling_log2(max|acc|) - 7, round-half-up arithmetic shift, saturate to
//         [-128, 127], then the gradient variants' truncating int8 division (`sub` = 1, 16 or 32).
template <int MODE>
__global__ void __launch_bounds__(kThreads)
rescale_kernel(const int32_t* __restrict__ acc, size_t n, const uint32_t* __restrict__ maxabs, int sub,
               int8_t* __restrict__ out, int8_t* __restrict__ w_inout,
               const int8_t* __restrict__ exp_a, const int8_t* __restrict__ exp_b, int8_t* __restrict__ exp_out,
               int relu, const int8_t* __restrict__ gate, const DpView dp, int dp_slot) {
    __shared__ uint32_t dp_sh;
    uint32_t mglobal = *maxabs;
    if (MODE == 0) mglobal = dp_allreduce_max(dp, dp_slot, mglobal, &dp_sh);     // data parallel: max over ranks (dp.cuh)
    const int b = niti_bitwidth(mglobal);
    int shift;          // PSTO shift; kind: 0 = PSTO, 1 = raw cast, 2 = zeros
    int kind = 0;
    if (MODE == 2) {
        // quantize.h:260-265 ceiling_log2: v < 4 -> v - 1, else 32 - clz(v - 1); negative shifts are undefined in the
        // reference (1 << shift) and kept at 0 here (oracle/niti_oracle.c niti_o_rescale_dsp)
        const int cl = mglobal < 4u ? (int)mglobal - 1 : 32 - __clz(mglobal - 1u);
        shift = cl - 7 < 0 ? 0 : cl - 7;
        kind = 3;
        if (exp_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0)
            *exp_out = (int8_t)((exp_a ? (int)*exp_a : 0) + (exp_b ? (int)*exp_b : 0) + shift);
    } else if (MODE == 0) {
        const int s = b - 7;
        if (s > 1) shift = s; else if (s == 1) shift = 2; else { shift = 0; kind = 1; }
        if (exp_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
            int e = (exp_a ? (int)*exp_a : 0) + (exp_b ? (int)*exp_b : 0) + (kind == 1 ? 0 : shift);
            *exp_out = (int8_t)e;
        }
    } else {
        shift = b - sub;
        if (b == 0) kind = 2;
    }
    auto conv = [&](int32_t x) -> int32_t {
        if (kind == 0) return niti_psto(x, shift);
        if (kind == 1) return x;            // raw (int8_t) cast happens in pack / store (quirk Q3)
        if (kind == 3) {
            long long q = shift > 0 ? (((long long)x + (1ll << (shift - 1))) >> shift) : (long long)x;
            int32_t v = q > 127 ? 127 : (q < -128 ? -128 : (int32_t)q);
            return sub > 1 ? v / sub : v;
        }
        return 0;
    };
    const size_t ngrp = n / 16;
    const int4* v = reinterpret_cast<const int4*>(acc);
    for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < ngrp; g += (size_t)gridDim.x * blockDim.x) {
        int4 a0 = ldg_stream(v + g * 4 + 0), a1 = ldg_stream(v + g * 4 + 1);
        int4 a2 = ldg_stream(v + g * 4 + 2), a3 = ldg_stream(v + g * 4 + 3);
        int32_t r[16] = {conv(a0.x), conv(a0.y), conv(a0.z), conv(a0.w), conv(a1.x), conv(a1.y), conv(a1.z), conv(a1.w),
                         conv(a2.x), conv(a2.y), conv(a2.z), conv(a2.w), conv(a3.x), conv(a3.y), conv(a3.z), conv(a3.w)};
        if (out != nullptr) {
            uint4 o = make_uint4(pack4(r[0], r[1], r[2], r[3]), pack4(r[4], r[5], r[6], r[7]),
                                 pack4(r[8], r[9], r[10], r[11]), pack4(r[12], r[13], r[14], r[15]));
            if (MODE == 0 && relu) { o.x = __vmaxs4(o.x, 0u); o.y = __vmaxs4(o.y, 0u); o.z = __vmaxs4(o.z, 0u); o.w = __vmaxs4(o.w, 0u); }
            if (MODE == 0 && gate != nullptr) {       // fused NITI_CPUReluGrad: pass where the forward value was > 0
                uint4 gv = reinterpret_cast<const uint4*>(gate)[g];
                o.x &= __vcmpgts4(gv.x, 0u); o.y &= __vcmpgts4(gv.y, 0u); o.z &= __vcmpgts4(gv.z, 0u); o.w &= __vcmpgts4(gv.w, 0u);
            }
            reinterpret_cast<uint4*>(out)[g] = o;
        }
        if (MODE == 1 && w_inout != nullptr) {
            uint4 wv = reinterpret_cast<const uint4*>(w_inout)[g];
            uint32_t wi[4] = {wv.x, wv.y, wv.z, wv.w};
            uint32_t wo[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int32_t t[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    int32_t wk = (int32_t)(int8_t)((wi[j] >> (8 * k)) & 0xff);
                    int32_t d = (int32_t)(int8_t)(r[j * 4 + k] & 0xff);
                    int32_t u = wk - d;
                    t[k] = u > 127 ? 127 : (u < -127 ? -127 : u);
                }
                wo[j] = pack4(t[0], t[1], t[2], t[3]);
            }
            reinterpret_cast<uint4*>(w_inout)[g] = make_uint4(wo[0], wo[1], wo[2], wo[3]);
        }
    }
    // ragged tail (n % 16), handled by block 0
    if (blockIdx.x == 0) {
        for (size_t i = ngrp * 16 + threadIdx.x; i < n; i += blockDim.x) {
            int32_t r = conv(acc[i]);
            if (MODE == 0) {
                int8_t v = (int8_t)r;
                if (relu && v < 0) v = 0;
                if (gate != nullptr && gate[i] <= 0) v = 0;
                r = v;
            }
            if (out != nullptr) out[i] = (int8_t)r;
            if (MODE == 1 && w_inout != nullptr) {
                int32_t u = (int32_t)w_inout[i] - (int32_t)(int8_t)r;
                w_inout[i] = (int8_t)(u > 127 ? 127 : (u < -127 ? -127 : u));
            }
        }
    }
}

__global__ void __launch_bounds__(kThreads) sgd_kernel(int8_t* __restrict__ w, const int8_t* __restrict__ dw, size_t n) {
    const size_t ngrp = n / 16;
    for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < ngrp; g += (size_t)gridDim.x * blockDim.x) {
        uint4 wv = reinterpret_cast<const uint4*>(w)[g];
        uint4 dv = reinterpret_cast<const uint4*>(dw)[g];
        uint32_t wi[4] = {wv.x, wv.y, wv.z, wv.w}, di[4] = {dv.x, dv.y, dv.z, dv.w}, wo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int32_t t[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int32_t u = (int32_t)(int8_t)((wi[j] >> (8 * k)) & 0xff) - (int32_t)(int8_t)((di[j] >> (8 * k)) & 0xff);
                t[k] = u > 127 ? 127 : (u < -127 ? -127 : u);
            }
            wo[j] = pack4(t[0], t[1], t[2], t[3]);
        }
        reinterpret_cast<uint4*>(w)[g] = make_uint4(wo[0], wo[1], wo[2], wo[3]);
    }
    if (blockIdx.x == 0) {
        for (size_t i = ngrp * 16 + threadIdx.x; i < n; i += blockDim.x) {
            int32_t u = (int32_t)w[i] - (int32_t)dw[i];
            w[i] = (int8_t)(u > 127 ? 127 : (u < -127 ? -127 : u));
        }
    }
}


// rule B per OC slice with the post-alignment of NITI_GradientSplitConv_Int8 (NITI_GradientSplitConv_Int8.cpp:318-372):
// slice s: b_s = bitwidth(max|acc_s|); v = b_s == 0 ? 0 : PSTO(acc, b_s - sub); then slices whose b_s is below the
// largest one are integer-divided (C truncation) by 2^(max b - b_s).
__global__ void __launch_bounds__(kThreads)
rescale_split_kernel(const int32_t* __restrict__ acc, size_t per, int slices, const uint32_t* __restrict__ maxabs, int sub,
                     int8_t* __restrict__ out) {
    int bmax = 0;
    for (int s = 0; s < slices; ++s) bmax = max(bmax, niti_bitwidth(maxabs[s]));
    const size_t n = per * (size_t)slices;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int b = niti_bitwidth(maxabs[i / per]);
        int32_t v = b == 0 ? 0 : niti_psto(acc[i], b - sub);
        v = (int32_t)(int8_t)v;
        if (bmax > b) { const int32_t dv = niti_shl1(bmax - b); v = dv == INT32_MIN ? 0 : v / dv; }
        out[i] = (int8_t)v;
    }
}

__global__ void zero_u32_kernel(uint32_t* p, int count) {
    niti_pdl_trigger(); niti_pdl_wait();                     // programmatic dependent launch: see common.cuh
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) p[i] = 0u;
}

int grid_for(niti_ctx* ctx, size_t work_items) {
    size_t blocks = (work_items + kThreads - 1) / kThreads;
    size_t cap = (size_t)ctx->sm_count * 8;            // 8 resident 256-thread CTAs per SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

} // namespace

extern "C" int niti_zero_u32(niti_ctx* ctx, uint32_t* p, int count, niti_stream st) {
    if (!ctx || !p || count <= 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_zero_u32");
    cudaError_t le = niti_launch(ctx, NITI_PDL_ELEMENTWISE, zero_u32_kernel, dim3((count + 63) / 64), dim3(64), 0, (cudaStream_t)st, p, count);
    if (le != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch zero_u32", le);
    return niti_check_launch(ctx, "zero_u32");
}

extern "C" int niti_absmax(niti_ctx* ctx, const int32_t* acc, size_t n, uint32_t* maxabs, niti_stream st) {
    if (!ctx || !maxabs || (n && !acc)) return niti_fail(ctx, NITI_ERR_ARG, "niti_absmax");
    if (n == 0) return NITI_OK;
    if (!aligned16(acc)) return niti_fail(ctx, NITI_ERR_ARG, "niti_absmax: acc must be 16-byte aligned");
    absmax_kernel<<<grid_for(ctx, n / 4 + 1), kThreads, 0, (cudaStream_t)st>>>(acc, n, maxabs);
    return niti_check_launch(ctx, "absmax");
}

extern "C" int niti_rescale_act(niti_ctx* ctx, const int32_t* acc, size_t n, const uint32_t* maxabs, int8_t* out,
                                const int8_t* exp_a, const int8_t* exp_b, int8_t* exp_out, niti_stream st) {
    if (!ctx || !maxabs || (n && (!acc || !out))) return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_act");
    if (n && (!aligned16(acc) || !aligned16(out)))
        return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_act: buffers must be 16-byte aligned");
    rescale_kernel<0><<<grid_for(ctx, n / 16 + 1), kThreads, 0, (cudaStream_t)st>>>(
        acc, n, maxabs, 0, out, nullptr, exp_a, exp_b, exp_out, 0, nullptr, DpView{}, 0);
    return niti_check_launch(ctx, "rescale_act");
}

extern "C" int niti_rescale_act_fused(niti_ctx* ctx, const int32_t* acc, size_t n, const uint32_t* maxabs, int8_t* out,
                                      const int8_t* exp_a, const int8_t* exp_b, int8_t* exp_out, int relu,
                                      const int8_t* gate, niti_stream st) {
    if (!ctx || !maxabs || (n && (!acc || !out))) return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_act_fused");
    if (n && (!aligned16(acc) || !aligned16(out) || (gate && !aligned16(gate))))
        return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_act_fused: buffers must be 16-byte aligned");
    rescale_kernel<0><<<grid_for(ctx, n / 16 + 1), kThreads, 0, (cudaStream_t)st>>>(
        acc, n, maxabs, 0, out, nullptr, exp_a, exp_b, exp_out, relu, gate, DpView{}, 0);
    return niti_check_launch(ctx, "rescale_act_fused");
}

extern "C" int niti_rescale_act_dp(niti_ctx* ctx, const int32_t* acc, size_t n, const uint32_t* maxabs, int8_t* out,
                                   const int8_t* exp_a, const int8_t* exp_b, int8_t* exp_out, int relu,
                                   const int8_t* gate, int dp_slot, niti_stream st) {
    if (!ctx || !maxabs || (n && (!acc || !out))) return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_act_dp");
    if (dp_slot < 0 || dp_slot >= kDpSlots) return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_act_dp: slot out of range");
    if (n && (!aligned16(acc) || !aligned16(out) || (gate && !aligned16(gate))))
        return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_act_dp: buffers must be 16-byte aligned");
    rescale_kernel<0><<<grid_for(ctx, n / 16 + 1), kThreads, 0, (cudaStream_t)st>>>(
        acc, n, maxabs, 0, out, nullptr, exp_a, exp_b, exp_out, relu, gate, ctx->dp, dp_slot);
    return niti_check_launch(ctx, "rescale_act_dp");
}

extern "C" int niti_rescale_grad(niti_ctx* ctx, const int32_t* acc, size_t n, const uint32_t* maxabs, int sub,
                                 int8_t* dw_out, int8_t* w_inout, niti_stream st) {
    if (!ctx || !maxabs || (n && !acc) || (!dw_out && !w_inout)) return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_grad");
    if (n == 0) return NITI_OK;
    if (!aligned16(acc) || (dw_out && !aligned16(dw_out)) || (w_inout && !aligned16(w_inout)))
        return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_grad: buffers must be 16-byte aligned");
    rescale_kernel<1><<<grid_for(ctx, n / 16 + 1), kThreads, 0, (cudaStream_t)st>>>(
        acc, n, maxabs, sub, dw_out, w_inout, nullptr, nullptr, nullptr, 0, nullptr, DpView{}, 0);
    return niti_check_launch(ctx, "rescale_grad");
}

extern "C" int niti_rescale_dsp(niti_ctx* ctx, const int32_t* acc, size_t n, const uint32_t* maxabs, int post_div, int8_t* out,
                                const int8_t* exp_a, const int8_t* exp_b, int8_t* exp_out, niti_stream st) {
    if (!ctx || !maxabs || (n && (!acc || !out))) return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_dsp");
    if (post_div != 1 && post_div != 16 && post_div != 32) return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_dsp: post_div must be 1, 16 or 32");
    if (n && (!aligned16(acc) || !aligned16(out)))
        return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_dsp: buffers must be 16-byte aligned");
    rescale_kernel<2><<<grid_for(ctx, n / 16 + 1), kThreads, 0, (cudaStream_t)st>>>(
        acc, n, maxabs, post_div, out, nullptr, exp_a, exp_b, exp_out, 0, nullptr, DpView{}, 0);
    return niti_check_launch(ctx, "rescale_dsp");
}

extern "C" int niti_sgd_update(niti_ctx* ctx, int8_t* w, const int8_t* dw, size_t n, niti_stream st) {
    if (!ctx || (n && (!w || !dw))) return niti_fail(ctx, NITI_ERR_ARG, "niti_sgd_update");
    if (n == 0) return NITI_OK;
    if (!aligned16(w) || !aligned16(dw)) return niti_fail(ctx, NITI_ERR_ARG, "niti_sgd_update: 16-byte alignment");
    sgd_kernel<<<grid_for(ctx, n / 16 + 1), kThreads, 0, (cudaStream_t)st>>>(w, dw, n);
    return niti_check_launch(ctx, "sgd_update");
}

extern "C" int niti_rescale_grad_split(niti_ctx* ctx, const int32_t* acc, size_t n_per_slice, int slices,
                                       const uint32_t* maxabs, int sub, int8_t* out, niti_stream st) {
    if (!ctx || !acc || !maxabs || !out || slices <= 0 || n_per_slice == 0) return niti_fail(ctx, NITI_ERR_ARG, "niti_rescale_grad_split");
    rescale_split_kernel<<<grid_for(ctx, n_per_slice * slices), kThreads, 0, (cudaStream_t)st>>>(acc, n_per_slice, slices, maxabs, sub, out);
    return niti_check_launch(ctx, "rescale_grad_split");
}
