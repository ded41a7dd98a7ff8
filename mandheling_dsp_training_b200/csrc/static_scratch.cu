// static_scratch.cu — niti_matmul_nt_static: the whole-op int8 matmul (NITI_Matmul_Int8: contraction, max|acc|,
// shared exponent, rule B with sub = 3) for callers that must not allocate device memory.
//
// Who that is: the other whole-op entry points keep their int32 accumulators in a workspace that a niti_ctx grows with
// cudaMalloc, and niti_ctx_create itself allocates.  That is a problem for a host that is handed a fixed device-memory
// budget up front and may not go back to the driver for more (an inference/training runtime embedded in another
// application's memory plan, or a process whose allocator is locked once set-up is over), and for code that runs while a
// stream is being captured or inside a latency-critical section, where cudaMalloc is either illegal or an implicit
// device-wide synchronisation.  Such a caller still owns its inputs and its output; what it cannot supply is the op's
// intermediate storage.  Here that storage is part of the library image instead: a fixed array in device global memory
// that every GPU which loads libniti_b200.so carries from module load until the process exits.
//
// This file is self-contained on purpose: it needs no niti_ctx, touches no workspace of api.cu and changes no other path.
#include "common.cuh"
#include <cstdio>
#include <cstdarg>
#include <mutex>

namespace {

// Capacity: NITI_STATIC_SCRATCH_MAX_WORDS int32 accumulators (e * h of one call) behind a 256-byte header whose first
// word is the running max|acc|.  4 MiB + 256 B per GPU, resident for the life of the process.
constexpr size_t kHeaderWords = 64;
constexpr size_t kMaxWords = NITI_STATIC_SCRATCH_MAX_WORDS;
static __device__ __align__(256) int32_t g_static_scratch[kHeaderWords + kMaxWords];

constexpr int kMaxDevices = 64;
struct DeviceSlot { int32_t* base; int sms; };
DeviceSlot g_slot[kMaxDevices];          // base == nullptr: this device has not been set up yet
std::mutex g_slot_mutex;

int fail(char* err, size_t err_len, int code, const char* fmt, ...) {
    if (err && err_len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

// In host code `g_static_scratch` is only a handle for the runtime's symbol table, not an address a kernel could use.  The
// real device address differs per GPU, so it is looked up once per device, with that device current, and remembered.
int resolve_scratch(int device, DeviceSlot* out, char* err, size_t err_len) {
    std::lock_guard<std::mutex> lock(g_slot_mutex);
    if (!g_slot[device].base) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p)
            return fail(err, err_len, NITI_ERR_CUDA, "niti_matmul_nt_static: cudaGetSymbolAddress on device %d: %s", device,
                        cudaGetErrorString(e));
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess || sms <= 0)
            return fail(err, err_len, NITI_ERR_CUDA, "niti_matmul_nt_static: SM count of device %d: %s", device, cudaGetErrorString(e));
        g_slot[device].sms = sms;
        g_slot[device].base = static_cast<int32_t*>(p);
    }
    *out = g_slot[device];
    return NITI_OK;
}

// The array keeps whatever the previous call left.  Only the max|acc| word has to start at zero (it is built with atomicMax);
// the accumulators are overwritten before they are read.
__global__ void static_clear_kernel(int32_t* scratch) { scratch[0] = 0; }

constexpr int kTile = 64;                 // output tile edge
constexpr int kChunk = 64;                // contraction bytes per shared-memory stage
constexpr int kRowWords = kChunk / 4 + 1; // +1 word: rows land in different banks

// one row-chunk word of A or B; out-of-range rows / contraction indices read as 0
__device__ __forceinline__ uint32_t load_word(const int8_t* __restrict__ m, int rows, int l, int row, int k, bool words_ok) {
    if (row >= rows || k >= l) return 0u;
    const int8_t* p = m + (size_t)row * l + k;
    if (words_ok) return __ldg(reinterpret_cast<const uint32_t*>(p));        // l % 4 == 0 and base 4-byte aligned: k + 3 < l
    uint32_t v = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if (k + j < l) v |= (uint32_t)(uint8_t)__ldg(p + j) << (8 * j);
    return v;
}

// acc[r][c] = sum_k A[r][k] * B[c][k] (both row-major, int8) into the scratch array, max|acc| into its header word.
// 256 threads own a 64 x 64 tile; each thread 4 x 4 outputs interleaved by 16 so shared-memory reads are conflict-free.
__global__ void __launch_bounds__(256)
static_matmul_acc_kernel(int e, int h, int l, const int8_t* __restrict__ a, const int8_t* __restrict__ b,
                         int32_t* __restrict__ scratch, int words_ok) {
    __shared__ uint32_t sa[kTile][kRowWords];
    __shared__ uint32_t sb[kTile][kRowWords];
    int32_t* acc_out = scratch + kHeaderWords;
    const int tiles_h = (h + kTile - 1) / kTile;
    const int tiles = ((e + kTile - 1) / kTile) * tiles_h;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    uint32_t m = 0;
    for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int r0 = (t / tiles_h) * kTile, c0 = (t % tiles_h) * kTile;
        int32_t acc[4][4] = {};
        for (int k0 = 0; k0 < l; k0 += kChunk) {
            __syncthreads();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int wi = threadIdx.x + 256 * i, row = wi >> 4, kw = wi & 15;
                sa[row][kw] = load_word(a, e, l, r0 + row, k0 + 4 * kw, words_ok != 0);
                sb[row][kw] = load_word(b, h, l, c0 + row, k0 + 4 * kw, words_ok != 0);
            }
            __syncthreads();
#pragma unroll
            for (int kw = 0; kw < kChunk / 4; ++kw) {
                uint32_t av[4], bv[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) { av[i] = sa[ty + 16 * i][kw]; bv[i] = sb[tx + 16 * i][kw]; }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = __dp4a((int)av[i], (int)bv[j], acc[i][j]);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int r = r0 + ty + 16 * i, c = c0 + tx + 16 * j;
                if (r < e && c < h) {
                    acc_out[(size_t)r * h + c] = acc[i][j];
                    m = max(m, niti_abs_u32(acc[i][j]));
                }
            }
    }
    m = warp_max_u32(m);
    if ((threadIdx.x & 31) == 0 && m) atomicMax(reinterpret_cast<uint32_t*>(scratch), m);
}

// rule B of NITI_Matmul_Int8.cpp:216-228: b = ceil(log2(max|acc|)); b == 0 -> zeros, else PSTO(acc, b - 3)
__global__ void __launch_bounds__(256)
static_rescale_kernel(const int32_t* __restrict__ scratch, size_t n, int8_t* __restrict__ out) {
    const int bw = niti_bitwidth(*reinterpret_cast<const uint32_t*>(scratch));
    const int32_t* acc = scratch + kHeaderWords;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = bw == 0 ? (int8_t)0 : (int8_t)niti_psto(acc[i], bw - 3);
}

}  // namespace

extern "C" int niti_matmul_nt_static(int e, int h, int l, const int8_t* a, const int8_t* b, int8_t* c, niti_stream st,
                                     char* err, size_t err_len) {
    if (err && err_len) err[0] = 0;
    if (e <= 0 || h <= 0 || l <= 0 || !a || !b || !c)
        return fail(err, err_len, NITI_ERR_ARG, "niti_matmul_nt_static: bad argument (e=%d h=%d l=%d a=%p b=%p c=%p)", e, h, l,
                    (const void*)a, (const void*)b, (void*)c);
    const size_t n = (size_t)e * (size_t)h;
    if (n > kMaxWords)
        return fail(err, err_len, NITI_ERR_UNSUPPORTED,
                    "niti_matmul_nt_static: e*h = %zu exceeds the built-in scratch of %zu accumulators (NITI_STATIC_SCRATCH_MAX_WORDS); "
                    "split the rows of A, or use niti_matmul_nt with a context", n, kMaxWords);
    int device = -1;
    cudaError_t ce = cudaGetDevice(&device);
    if (ce != cudaSuccess) return fail(err, err_len, NITI_ERR_CUDA, "niti_matmul_nt_static: cudaGetDevice: %s", cudaGetErrorString(ce));
    if (device < 0 || device >= kMaxDevices)
        return fail(err, err_len, NITI_ERR_UNSUPPORTED, "niti_matmul_nt_static: device ordinal %d outside [0, %d)", device, kMaxDevices);
    DeviceSlot slot;
    if (int rc = resolve_scratch(device, &slot, err, err_len)) return rc;

    cudaStream_t stream = (cudaStream_t)st;
    const int words_ok = (l % 4 == 0) && ((uintptr_t)a % 4 == 0) && ((uintptr_t)b % 4 == 0);
    const long tiles = (long)((e + kTile - 1) / kTile) * ((h + kTile - 1) / kTile);
    const long cap = (long)slot.sms * 4;
    static_clear_kernel<<<1, 1, 0, stream>>>(slot.base);
    static_matmul_acc_kernel<<<(unsigned)(tiles < cap ? tiles : cap), 256, 0, stream>>>(e, h, l, a, b, slot.base, words_ok);
    const long blocks = (long)((n + 255) / 256);
    static_rescale_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, stream>>>(slot.base, n, c);
    ce = cudaGetLastError();
    if (ce != cudaSuccess) return fail(err, err_len, NITI_ERR_CUDA, "niti_matmul_nt_static: launch: %s", cudaGetErrorString(ce));
    return NITI_OK;
}
