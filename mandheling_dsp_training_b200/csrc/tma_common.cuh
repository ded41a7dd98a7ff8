// tma_common.cuh — TMA (cp.async.bulk.tensor) PTX wrappers and tensor-map encoders shared by the newer kernel files
// (conv_dw2.cu, conv_stem.cu).  Tensor maps are encoded through the driver entry points fetched at run time, so the
// library has no link-time libcuda dependency.
#pragma once
#include "common.cuh"
#include "tc_ptx.cuh"
#include <cuda.h>
#include <mutex>

namespace tmax {

// ---- device: loads ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void load_2d(const CUtensorMap* map, uint32_t dst, uint64_t* bar, int x, int y) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(tc::smem_u32(bar)), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void load_3d(const CUtensorMap* map, uint32_t dst, uint64_t* bar, int x, int y, int z) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(tc::smem_u32(bar)), "r"(x), "r"(y), "r"(z) : "memory");
}
__device__ __forceinline__ void load_im2col(const CUtensorMap* map, uint32_t dst, uint64_t* bar, int c, int w, int h, int n,
                                            uint16_t off_w, uint16_t off_h) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
                 " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(tc::smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n),
                   "h"(off_w), "h"(off_h) : "memory");
}
__device__ __forceinline__ void prefetch_map(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
// ---- device: stores -----------------------------------------------------------------------------------------------
__device__ __forceinline__ void store_2d(const CUtensorMap* map, uint32_t src, int x, int y) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void reduce_add_2d(const CUtensorMap* map, uint32_t src, int x, int y) {
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// smem matrix descriptor; layout: 0 = no swizzle, 2 = 128B, 4 = 64B, 6 = 32B
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)layout << 61;
    return d;
}

// Grid-wide barrier for a persistent grid whose CTAs are all resident (grid <= SM count at 1 CTA/SM).  One thread per CTA.
// (common.cuh niti_grid_barrier: bounded spin, sets the context's sync-error flag instead of hanging.)
__device__ __forceinline__ void grid_barrier(uint32_t* counter, uint32_t target) { niti_grid_barrier(counter, target); }

// ---- host: tensor-map encoders --------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
struct Driver {
    EncodeTiledFn tiled = nullptr;
    EncodeIm2colFn im2col = nullptr;
    int version = 0;
};
inline Driver& driver() {
    static Driver d;
    static std::once_flag once;
    std::call_once(once, [] {
        void* f = nullptr; cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            d.tiled = (EncodeTiledFn)f;
        f = nullptr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            d.im2col = (EncodeIm2colFn)f;
        cudaDriverGetVersion(&d.version);
    });
    return d;
}
inline bool driver_ok() { Driver& d = driver(); return d.tiled && d.im2col; }

inline CUtensorMapSwizzle swizzle_of(int bytes) {
    return bytes >= 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B :
           (bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE));
}
// row-major byte matrix [rows][cols]; box = [box_rows][box_cols]; swz_bytes = swizzle span (0 = none)
inline bool make_tiled_u8(CUtensorMap* m, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows, uint32_t box_cols,
                          int swz_bytes) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t es[2] = {1, 1};
    return driver().tiled(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(swz_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// row-major int32 matrix [rows][cols]; box = 32 rows x 32 columns (128-byte rows, 128B swizzle): the epilogue store unit
inline bool make_tiled_i32(CUtensorMap* m, void* base, uint64_t rows, uint64_t cols) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 4};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t es[2] = {1, 1};
    return driver().tiled(m, CU_TENSOR_MAP_DATA_TYPE_INT32, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// 16-channel NHWC int8 tensor seen as rows of 8-byte words: dims {w*2, h, n}; box = {ow*2, rows, 1} with rows*ow == 128.
// One request per image ROW (ow*16 contiguous bytes) instead of one per pixel as in im2col mode, which is what bounds the
// stem kernels (16 B per pixel).  Lands in shared memory as [128 pixels][16 B], the same image an im2col box produces;
// the filter tap is a coordinate shift ({(kx - pad)*2, h0 + ky - pad, n}); out-of-range words read as zero (= padding).
// Valid for stride 1, dilation 1, ow | 128, (128/ow) | oh, ow <= 128.
inline bool rows16_ok(int c, int stride, int dil, int oh, int ow) {
    return c == 16 && stride == 1 && dil == 1 && ow <= 128 && 128 % ow == 0 && oh % (128 / ow) == 0;
}
inline bool make_rows16(CUtensorMap* m, const void* base, int n, int h, int w, int ow) {
    cuuint64_t dims[3] = {(cuuint64_t)w * 2, (cuuint64_t)h, (cuuint64_t)n};
    cuuint64_t strides[2] = {(cuuint64_t)w * 16, (cuuint64_t)h * w * 16};
    cuuint32_t box[3] = {(cuuint32_t)ow * 2, (cuuint32_t)(128 / ow), 1};
    cuuint32_t es[3] = {1, 1, 1};
    return driver().tiled(m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void*>(base), dims, strides, box, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// NHWC int8 tensor [n][h][w][c]; box = `pixels` pixels x bkc channels, swizzle span = swz_bytes
inline bool make_im2col_u8(CUtensorMap* m, const void* base, int n, int h, int w, int c, int lower_h, int lower_w, int upper_h,
                           int upper_w, int stride, int bkc, int pixels, int swz_bytes) {
    cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
    cuuint64_t strides[3] = {(cuuint64_t)c, (cuuint64_t)w * c, (cuuint64_t)h * w * c};
    int lower[2] = {lower_w, lower_h}, upper[2] = {upper_w, upper_h};
    cuuint32_t es[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    CUresult r = driver().im2col(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(base), dims, strides, lower, upper,
                                 (cuuint32_t)bkc, (cuuint32_t)pixels, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(swz_bytes),
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    // driver <= 13.1: im2col descriptors of tensors under 128 KiB carry a stray bit (the fix-up CUTLASS applies too,
    // cute/atom/copy_traits_sm90_im2col.hpp)
    if (driver().version <= 13010 && (size_t)n * h * w * c < 131072) reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
    return true;
}

} // namespace tmax
