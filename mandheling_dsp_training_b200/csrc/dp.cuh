// dp.cuh — data-parallel exchange of the shared-exponent scalar over NVLink peer memory.
//
// NITI's exponent is b = ceil_log2(max|acc|) over the WHOLE batch tensor (NITI_Conv_Int8.cpp:260), so a batch sharded
// over g GPUs needs max over ranks of one uint32 between phase 1 and phase 2 of every forward / dX op.
// A NCCL all-reduce per scalar costs ~15 us of launch + protocol latency, 21 times per VGG-11 step.  Instead each rank
// owns a small exchange buffer that every peer maps (CUDA IPC); the rescale kernel itself
//   1. block 0, thread r < world: stores (epoch << 32 | local max) into rank r's buffer at [slot][my rank]  (one 8-byte
//      NVLink store per peer; tag and value travel in the same word, so no ordering between two stores is needed)
//   2. every block: spins on its OWN buffer (local L2) until all `world` words of the slot carry this step's epoch,
//      and takes their max.
// The epoch is a device-side counter bumped once per step (niti_dp_step_begin), so the sequence is CUDA-graph replayable.
// A slot is reused only one step later; a rank cannot run a full step ahead of a peer (every slot needs the peer's word),
// so a word is never overwritten before it was read as long as a step uses >= 2 slots.
#pragma once
#include <stdint.h>

constexpr int kDpMaxWorld = 8;
constexpr int kDpSlots = 256;
constexpr int kDpArSlots = 64;                        // gradient all-reduce calls per step (one per bucket)

// layout of one rank's exchange buffer (device memory, cudaMalloc'ed so it can be exported with cudaIpcGetMemHandle)
struct DpBuffer {
    unsigned long long word[kDpSlots][kDpMaxWorld];   // written by peers
    uint32_t epoch;                                   // local: current step number
    uint32_t error;                                   // local: set when a spin timed out (peer died)
    // gradient all-reduce (dp.cu dp_allreduce_kernel): epoch-tagged arrival / completion flags written by peers, and the
    // local count of CTAs that finished their share
    unsigned long long ar_in[kDpArSlots][kDpMaxWorld];
    unsigned long long ar_out[kDpArSlots][kDpMaxWorld];
    uint32_t ar_count[kDpArSlots];
};

// every rank's gradient window (int32 words, cudaMalloc'ed and exported like the exchange buffer)
struct DpWindow {
    int32_t* peer[kDpMaxWorld];
    size_t words;
};

struct DpView {                                       // passed by value to kernels
    DpBuffer* peer[kDpMaxWorld];                      // peer[r] = rank r's buffer as mapped in this process
    int world, rank;
    long long timeout_cycles;                         // spin budget per exchange (niti_dp_connect: NITI_B200_DP_TIMEOUT_MS, default 60 s)
};

// The exchange itself, run by ONE thread of the launch's publishing CTA (`publish`) and by one thread of every CTA that
// needs the result: store (epoch << 32 | local) into every rank's buffer, then spin on the local copy of each peer's word
// until it carries this step's epoch.  Returns the max over ranks.  A peer that never answers within timeout_cycles
// (its process died, or it is stuck behind something that waits for us) sets DpBuffer.error — sticky, read by
// niti_dp_error; hosts poll it at their sync points and must treat everything since the last clean poll as invalid.
__device__ __forceinline__ uint32_t dp_exchange_max(const DpView& v, int slot, uint32_t local, bool publish) {
    if (v.world <= 1) return local;
    DpBuffer* mine = v.peer[v.rank];
    const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(&mine->epoch);
    if (publish) {
        const unsigned long long w = ((unsigned long long)epoch << 32) | (unsigned long long)local;
        for (int r = 0; r < v.world; ++r) {
            unsigned long long* dst = &v.peer[r]->word[slot][v.rank];
            asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(dst), "l"(w) : "memory");
        }
    }
    uint32_t m = local;
    const long long t0 = clock64();
    for (int r = 0; r < v.world; ++r) {
        if (r == v.rank) continue;
        const unsigned long long* src = &mine->word[slot][r];
        unsigned long long w;
        for (;;) {
            asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(w) : "l"(src) : "memory");
            if ((uint32_t)(w >> 32) == epoch) break;
            if (clock64() - t0 > v.timeout_cycles) { mine->error = 1u; w = 0; break; }
            __nanosleep(100);
        }
        m = max(m, (uint32_t)w);
    }
    return m;
}

// returns max over ranks of `local`; must be called by ALL threads of the block (uses one __shared__ word).
// Block 0 publishes.  world <= 1: returns local.
__device__ __forceinline__ uint32_t dp_allreduce_max(const DpView& v, int slot, uint32_t local, uint32_t* sh) {
    if (v.world <= 1) return local;
    if (threadIdx.x == 0) *sh = dp_exchange_max(v, slot, local, blockIdx.x == 0);
    __syncthreads();
    return *sh;
}
