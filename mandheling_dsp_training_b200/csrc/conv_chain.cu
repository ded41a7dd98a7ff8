// conv_chain.cu — the layer-walking fused NITI op: one persistent launch runs a CHAIN of forward (or dX) convolutions.
//
// Per layer (= one reference op, NITI_Conv_Int8.cpp:162-310 / NITI_DeConv_Int8.cpp:187-332, plus the element-wise ops the
// reference runs around it):
//     contraction (TMA im2col -> tcgen05.mma.cta_group::2 kind::i8, int32 in tensor memory)
//  -> max|acc| straight from tensor memory -> grid barrier (+ NVLink peer exchange when data-parallel, dp.cuh)
//  -> shared exponent -> PSTO -> int8, with the neighbouring element-wise ops folded into the store:
//        forward : ReLU (NITI_CPURelu_Int8) and the 2x2/2 max pool that follows it (NITI_Maxpool_Int8.cpp:24-70)
//        dX      : ReLU-grad gate (NITI_CPUReluGrad_Int8) or max-pool gradient + gate (NITI_CPUPoolGrad_Int8.cpp:36-72)
//  -> grid barrier -> next layer (its operand is what this layer just stored).
// The reference executes these as separate ops in a serial per-op loop (express/Executor.cpp:333-359); round 1 ran one
// launch per op (igemm_fused_kernel<BN,BKC>, conv_tma2.cu) and measured 7-14 us of launch + prologue + drain per op on top
// of 9-24 us in the kernel (profiles/phase_r2_vgg_before.txt).  Here tile shape and pipeline depth are run-time per layer,
// tensor memory is allocated once, the mbarrier ring is re-used across layers with per-stage phase bits, and a chain of one
// layer is the single-op entry point (niti_conv_fwd_fused / niti_conv_dx_fused).
#include "pair_ptx.cuh"
#include "tma_common.cuh"
#include <new>
#include <vector>

bool tma2_fwd_eligible(const niti_conv_desc*);
bool tma2_dx_eligible(const niti_conv_desc*);

namespace {

constexpr int kChainThreads = 576;   // warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-17: epilogue
constexpr int kChainMaxTiles = 8;    // resident accumulator tiles per CTA pair (512 TMEM columns / 64)
constexpr int kChainMaxStages = 8;
constexpr int kChainSmem = 232448;   // all of it: one CTA per SM is a requirement (512 TMEM columns + grid barriers)
constexpr int kChainStageArea = ((kChainSmem - 1024 - 512) / 1024) * 1024;

struct alignas(128) ChainLayer {
    CUtensorMap mapA;                 // activations (forward: x, dX: dy), im2col mode, box {bkc, 128 pixels}
    CUtensorMap mapB;                 // weights [Ncols][K] K-major, box {bkc, bn/2}
    int M, Ncols;                     // output extents: pixels x channels
    int out_h, out_w;                 // pixel space enumerated by the im2col rows
    int kh, kw, dil, stride, lower_h, lower_w, flip;
    int cchunks, num_kb;              // channel chunks per tap; K blocks per tile
    int n_tiles, items, pairs;        // items = m_tiles * n_tiles, distributed over `pairs` CTA pairs
    int bn, bkc, stages;
    int nslots, stream;               // accumulator slots of bn columns in tensor memory; stream: tiles per pair > nslots (slots are recycled)
    int8_t* y; const int8_t* gate;
    const int8_t* exp_a; const int8_t* exp_b; int8_t* exp_out;
    uint32_t* maxabs;
    int relu, dp_slot;
    int8_t* y_pool; int pool;         // forward: 2x2/2 max pool of y fused into the store (out_w in {2,4,8,16}, out_h even)
    const int8_t* pg_x; int pg_gate;  // dX: the result is the gradient w.r.t. a pooled tensor; pg_x = the pre-pool forward
                                      // tensor [n][2*out_h][2*out_w][Ncols]; y receives the gradient w.r.t. it
};

struct ChainArgs {
    const ChainLayer* table;          // device array, or nullptr: use the by-value layer
    int nlayers;
    uint32_t* sync;                   // {arrival counter, exit counter}
    DpView dp;
    unsigned long long* trace;        // [layer][block][8] globaltimer stamps, or nullptr
    unsigned long long* tl;           // launch timeline record, or nullptr
    uint32_t* pred;                   // [kDpSlots] predictor state per slot (common.cuh niti_pred_*), [kDpSlots] confidence, {hits, misses, not speculated}
    int spec_confident;               // consecutive predictions that must have held before a single-tile layer speculates
    int speculate;                    // 0: pass 0 only reduces the maximum, pass 1 always stores (two-sweep schedule); 1: adaptive; 2: always
    int trigger_late;
};

__device__ __forceinline__ void chain_trace(unsigned long long* trace, int layer, int slot) {
    if (trace != nullptr) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        trace[((size_t)layer * gridDim.x + blockIdx.x) * 8 + slot] = t;
    }
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ uint4 ldg_u4(const int8_t* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ uint4 vmax4(uint4 a, uint4 b) {
    return make_uint4(__vmaxs4(a.x, b.x), __vmaxs4(a.y, b.y), __vmaxs4(a.z, b.z), __vmaxs4(a.w, b.w));
}
__device__ __forceinline__ uint4 vand4(uint4 a, uint4 b) { return make_uint4(a.x & b.x, a.y & b.y, a.z & b.z, a.w & b.w); }
__device__ __forceinline__ uint4 vandn4(uint4 a, uint4 b) { return make_uint4(a.x & ~b.x, a.y & ~b.y, a.z & ~b.z, a.w & ~b.w); }
__device__ __forceinline__ uint4 vor4(uint4 a, uint4 b) { return make_uint4(a.x | b.x, a.y | b.y, a.z | b.z, a.w | b.w); }
__device__ __forceinline__ uint4 vge4(uint4 a, uint4 b) {
    return make_uint4(__vcmpges4(a.x, b.x), __vcmpges4(a.y, b.y), __vcmpges4(a.z, b.z), __vcmpges4(a.w, b.w));
}
__device__ __forceinline__ uint4 vgt0(uint4 a) {
    return make_uint4(__vcmpgts4(a.x, 0u), __vcmpgts4(a.y, 0u), __vcmpgts4(a.z, 0u), __vcmpgts4(a.w, 0u));
}

// kTable: walk a.table[0..nlayers) in global memory; otherwise the one layer passed by value (its tensor maps are then read
// straight from the kernel-parameter bank instead of through a generic pointer that may point either way)
//
// Tensor memory is a ring of L.nslots accumulator slots of L.bn columns; tile t of a CTA pair accumulates in slot
// t % nslots.  RESIDENT layers (tiles per pair <= nslots) keep every tile until the layer's exponent is settled, so a
// missed prediction (or no speculation) re-stores straight from tensor memory.  STREAMING layers (outputs larger than the
// 9.7 M accumulators of tensor memory: ResNet's 56^2 / 28^2 stages, GEMMs with M >= 64 k) hand a slot back to the MMA
// issuer as soon as its tile is swept — the sweep of tile i runs under the K loop of tile i+1 — and a miss RECOMPUTES the
// layer (the stem kernels' schedule); with speculation a streaming layer costs one contraction and no int32 ever reaches HBM.
// Per layer the epilogue tells the two issuing warps what comes next through `flag` + layer_bar: 0 = layer complete
// everywhere, go on; 1 = run this layer's loads / MMAs once more (store pass of a streaming layer).
// kStream: some layer of the chain is a streaming layer.  Chains of resident layers only (all of VGG-11 at batch 256) compile
// without the slot hand-back and the per-layer verdict wait of the issuing warps (measured: 13 us per 10-layer chain).
template <bool kTable, bool kStream>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kChainThreads, 1)
chain_kernel(const __grid_constant__ ChainLayer single, const ChainArgs a) {
    extern __shared__ uint8_t smem_raw_chain[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw_chain) + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kChainStageArea);            // used in the leader only
    uint64_t* empty_bar = full_bar + kChainMaxStages;                                    // per CTA
    uint64_t* tfull_bar = empty_bar + kChainMaxStages;                                   // per CTA, one per accumulator slot
    uint64_t* tempty_bar = tfull_bar + kChainMaxTiles;                                   // leader only: slot swept by all 32 epilogue warps of the pair
    uint64_t* layer_bar = tempty_bar + kChainMaxTiles;                                   // per CTA: `flag` is valid
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(layer_bar + 1);
    uint32_t* bcast = tmem_slot + 1;
    uint32_t* cta_max = tmem_slot + 2;                                                   // this CTA's max|acc| of the current layer
    volatile uint32_t* flag = tmem_slot + 3;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1;
    if (tid == 0) *cta_max = 0u;
    if (!a.trigger_late) niti_pdl_trigger();                 // the next launch of the stream may start placing its CTAs
    if (tid == 64) chain_trace(a.trace, 0, 0);
    niti_tl_enter(a.tl);

    if (tid == 0) {
        for (int s = 0; s < kChainMaxStages; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        for (int t = 0; t < kChainMaxTiles; ++t) { tc::mbar_init(&tfull_bar[t], 1); tc::mbar_init(&tempty_bar[t], 32); }
        tc::mbar_init(layer_bar, 1);
        tc::fence_barrier_init();
    }
    // tensor maps of every layer are fetched now, behind the set-up, not in front of each layer's first load (a cold
    // descriptor fetch cost 1.5-2.3 us per layer on the critical path: profiles/phase_r2_vgg_chain_v1.txt)
    if (tid >= 32 && tid < 32 + 2 * a.nlayers && tid < 64) {
        const ChainLayer& Lp = kTable ? a.table[(tid - 32) >> 1] : single;
        prefetch_tmap2(((tid - 32) & 1) ? &Lp.mapB : &Lp.mapA);
    }
    // ... and the layer records themselves: every role reads record l+1 on the critical path of the layer transition, a DRAM
    // miss (~1 us) per layer whenever the table has left L2 (the bench flushes L2 between steps); pull every line into L2 now
    if (kTable && tid >= 64) {
        const char* tb = reinterpret_cast<const char*>(a.table);
        for (int off = (tid - 64) * 128; off < a.nlayers * (int)sizeof(ChainLayer); off += (kChainThreads - 64) * 128) prefetch_l2(tb + off);
    }
    if (warp == 1) tmem_alloc2<512>(tmem_slot);
    tc::tc_fence_before();
    cluster_sync_all();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    niti_pdl_wait();                                         // prologue done; everything the previous launch wrote is visible

    // Both issuing roles run their loops with the WHOLE warp (all values warp-uniform) and only the instruction that must
    // come from one thread sits under elect.sync.  Written as `if (lane == 0) { loop }` the compiler cannot prove that the
    // operands of UTMALDG / UTCIMMA are uniform and wraps every one of them in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop:
    // ~120 dependent single-thread instructions per K block, which — not TMA or the tensor pipe — set the 330 ns per K block
    // measured in round 1 (profiles/phase_r2_vgg_before.txt: same time per block for 20 KB and 24 KB stages, 64 or 128 CTAs).
    if (warp == 0) {
        // ------------------------------ TMA producer (both CTAs) ------------------------------
        uint32_t pphase = 0;                             // per-stage phase bits of empty_bar
        uint32_t lphase = 0;                             // phase of layer_bar
        for (int l = 0; l < a.nlayers; ++l) {
            const ChainLayer& L = kTable ? a.table[l] : single;
            // the layer record is read ONCE into registers: every asm below clobbers memory, so fields left in the
            // (global or param) record would be re-loaded on each use — measured +25 % on the K loop
            const int Lpairs = L.pairs, Litems = L.items, Lntiles = L.n_tiles, LM = L.M, Lbn = L.bn, Lbkc = L.bkc;
            const int Low = L.out_w, Loh = L.out_h, Lstride = L.stride, Llw = L.lower_w, Llh = L.lower_h, Lnumkb = L.num_kb;
            const int Lcch = L.cchunks, Lkw = L.kw, Lkh = L.kh, Lflip = L.flip, Ldil = L.dil, Lstages = L.stages;
            const uint32_t a_bytes = 128u * (uint32_t)Lbkc, stage = a_bytes + (uint32_t)(Lbn / 2) * (uint32_t)Lbkc;
            // tap walk without divisions: forward runs (ky, kx) up from (0, 0), dX (flipped kernel) down from (kh-1, kw-1)
            const int kx_first = Lflip ? (Lkw - 1) * Ldil : 0, kx_last = Lflip ? 0 : (Lkw - 1) * Ldil;
            const int kstep = Lflip ? -Ldil : Ldil;
            const int ky_first = Lflip ? (Lkh - 1) * Ldil : 0;
            int s = 0;
            for (;;) {
                if (pair < Lpairs)
                for (int item = pair; item < Litems; item += Lpairs) {
                    const int n_t = item % Lntiles; const int m_t = item / Lntiles;
                    int m0 = m_t * 256 + (int)rank * 128;
                    if (m0 >= LM) m0 = 0;                    // fully out-of-range half: load valid rows, results are masked
                    const int nrow0 = n_t * Lbn + (int)rank * (Lbn / 2);
                    const int q = m0 % Low; const int t = m0 / Low;
                    const int pp = t % Loh; const int bn_ = t / Loh;
                    const int bw = q * Lstride + Llw, bh = pp * Lstride + Llh;
                    int cc = 0, kx = kx_first, ky = ky_first, kcol = 0;      // channel chunk, tap offsets, K coordinate of block kb
                    for (int kb = 0; kb < Lnumkb; ++kb) {
                        tc::mbar_wait(&empty_bar[s], ((pphase >> s) & 1u) ^ 1u);
                        pphase ^= 1u << s;
                        const uint32_t a_dst = tc::smem_u32(smem) + (uint32_t)s * stage;
                        if (tc::elect_one()) {
                            if (rank == 0) mbar_expect_tx(&full_bar[s], 2u * stage);
                            tma2_load_im2col(&L.mapA, a_dst, &full_bar[s], cc, bw, bh, bn_, (uint16_t)kx, (uint16_t)ky);
                            tma2_load_2d(&L.mapB, a_dst + a_bytes, &full_bar[s], kcol, nrow0);
                        }
                        if (++s == Lstages) s = 0;
                        kcol += Lbkc; cc += Lbkc;
                        if (cc == Lcch * Lbkc) {
                            cc = 0;
                            if (kx == kx_last) { kx = kx_first; ky += kstep; } else kx += kstep;
                        }
                    }
                }
                if (!kStream && l + 1 == a.nlayers) break;                   // nothing follows and nothing is ever re-run
                tc::mbar_wait(layer_bar, lphase); lphase ^= 1u;              // the epilogue's verdict on this layer
                if (!kStream || *flag == 0u) break;                          // complete everywhere (grid barrier passed)
            }
            tmax::fence_proxy_async_all();                                   // generic-proxy stores -> async-proxy (TMA) reads
        }
    } else if (warp == 1) {
        if (rank == 0) {
            // ------------------------------ MMA issuer (leader CTA only) ------------------------------
            uint32_t cphase = 0;                             // per-stage phase bits of full_bar
            uint32_t euse = 0;                               // per-slot use parity (tempty_bar)
            uint32_t lphase = 0;
            for (int l = 0; l < a.nlayers; ++l) {
                const ChainLayer& L = kTable ? a.table[l] : single;
                const int Lpairs = L.pairs, Litems = L.items, Lbn = L.bn, Lbkc = L.bkc, Lnumkb = L.num_kb, Lstages = L.stages;
                const uint32_t Lslots = (uint32_t)L.nslots;
                const uint32_t idesc = tc::make_idesc_i8(256, Lbn, 0, 0);
                const uint32_t a_bytes = 128u * (uint32_t)Lbkc, stage = a_bytes + (uint32_t)(Lbn / 2) * (uint32_t)Lbkc;
                // shared-memory descriptor = layer constant | (address >> 4): K-major, LBO 16 B, SBO 8 rows, swizzle of the row
                const uint64_t dconst = make_desc2(0u, 16, 8u * (uint32_t)Lbkc, (Lbkc == 128) ? 2u : 4u);
                int s = 0;
                for (;;) {
                    uint32_t slot = 0;
                    if (pair < Lpairs)
                    for (int item = pair; item < Litems; item += Lpairs) {
                        const uint32_t d_tmem = tmem_base + slot * (uint32_t)Lbn;
                        // the slot's previous tile has been swept by all 32 epilogue warps of the pair (first use: passes)
                        if (kStream) {
                            tc::mbar_wait(&tempty_bar[slot], ((euse >> slot) & 1u) ^ 1u);
                            euse ^= 1u << slot;
                            tc::tc_fence_after();
                        }
                        for (int kb = 0; kb < Lnumkb; ++kb) {
                            tc::mbar_wait(&full_bar[s], (cphase >> s) & 1u);
                            cphase ^= 1u << s;
                            tc::tc_fence_after();
                            const uint32_t a_base = tc::smem_u32(smem) + (uint32_t)s * stage;
                            const uint64_t ad = dconst | (uint64_t)(a_base >> 4);
                            const uint64_t bd = dconst | (uint64_t)((a_base + a_bytes) >> 4);
                            if (tc::elect_one()) {
                                if (Lbkc == 128) {
                                    umma2_i8(d_tmem, ad, bd, idesc, (uint32_t)(kb != 0));
                                    umma2_i8(d_tmem, ad + 2, bd + 2, idesc, 1u);
                                    umma2_i8(d_tmem, ad + 4, bd + 4, idesc, 1u);
                                    umma2_i8(d_tmem, ad + 6, bd + 6, idesc, 1u);
                                } else {
                                    umma2_i8(d_tmem, ad, bd, idesc, (uint32_t)(kb != 0));
                                    umma2_i8(d_tmem, ad + 2, bd + 2, idesc, 1u);
                                }
                                umma2_commit_mc(&empty_bar[s], 3);   // frees this stage in both CTAs
                                if (kb == Lnumkb - 1) umma2_commit_mc(&tfull_bar[slot], 3);
                            }
                            if (++s == Lstages) s = 0;
                        }
                        if (++slot == Lslots) slot = 0;
                    }
                    if (!kStream) break;                     // resident chains: the next layer's loads are gated by the producer
                    tc::mbar_wait(layer_bar, lphase); lphase ^= 1u;
                    if (*flag == 0u) break;
                }
            }
        }
    } else {
        // ------------------------------ epilogue (warps 2..17, both CTAs) ------------------------------
        const int quarter = warp & 3;                        // TMEM lane quarter this warp may read
        const int group = (warp - 2) >> 2;                   // four warps share a quarter: every 4th 16-column chunk each
        const int etid = tid - 64;
        uint32_t tuse = 0;                                   // per-slot use parity (tfull_bar)
        uint32_t nbar = 0;                                   // grid barriers passed so far
        const uint32_t eone = (uint32_t)(a.speculate >= 0);  // 1, opaque to the compiler (common.cuh niti_rule_a_consts)
        uint32_t prev_norm = 0u;                             // norm of the previous layer's output (exponent prediction)
        if (etid == 0) chain_trace(a.trace, 0, 1);
        for (int l = 0; l < a.nlayers; ++l) {
            const ChainLayer& L = kTable ? a.table[l] : single;
            // layer record -> registers (see the producer)
            const int eM = L.M, eN = L.Ncols, eoh = L.out_h, eow = L.out_w, ent = L.n_tiles, eitems = L.items, epairs = L.pairs;
            const int ebn = L.bn, erelu = L.relu, epool = L.pool, epgg = L.pg_gate, edps = L.dp_slot;
            const uint32_t eslots = (uint32_t)L.nslots; const bool estream = kStream && L.stream != 0;
            int8_t* const ey = L.y; const int8_t* const egate = L.gate; int8_t* const eyp = L.y_pool; const int8_t* const epgx = L.pg_x;
            uint32_t* const emax = L.maxabs; const int8_t* const eea = L.exp_a; const int8_t* const eeb = L.exp_b; int8_t* const eeo = L.exp_out;
            const bool active = pair < epairs;
            const bool last_layer = l + 1 == a.nlayers;
            int shift = 0, kind = 0;
            uint32_t mx = 0;
            const int hw = eoh * eow;
            // dX: the store pass reads the forward tensor (ReLU gate, or the 2x2 windows of the pool gradient) row by row, a
            // dependent global load per 16-column chunk.  Its addresses do not depend on the accumulators: prefetch those
            // lines into L2 now, under the K loop (measured: the pool-gradient store pass was bound by that latency).
            if (active && (epgx != nullptr || egate != nullptr)) {
                int ahead = 0;
                for (int item = pair; item < eitems && ahead < kChainMaxTiles; item += epairs, ++ahead) {
                    const int n_t = item % ent; const int m_t = item / ent;
                    const int row = m_t * 256 + (int)rank * 128 + quarter * 32 + lane;
                    if (row >= eM) continue;
                    const int n0 = n_t * ebn;
                    const int nlines = (min(ebn, eN - n0) + 127) >> 7;
                    if (epgx != nullptr) {
                        const int img = row / hw, r = row - img * hw, py = r / eow, px = r - py * eow, w2 = eow * 2;
                        const int8_t* b00 = epgx + (((size_t)img * (eoh * 2) + 2 * py) * w2 + 2 * px) * eN + n0;
                        for (int j = group; j < 4 * nlines; j += 4)
                            prefetch_l2(b00 + (size_t)((j & 1) + ((j >> 1) & 1) * w2) * eN + (j >> 2) * 128);
                    } else {
                        for (int j = group; j < nlines; j += 4) prefetch_l2(egate + (size_t)row * eN + n0 + j * 128);
                    }
                }
            }
            // ---- speculative single pass ------------------------------------------------------------------------------
            // NITI's exponent needs max|acc| over the WHOLE tensor, which used to cost two sweeps over tensor memory with a
            // grid barrier between them (max pass | barrier | store pass | layer barrier).  The bit width of that maximum
            // seldom changes from one training step to the next, so pass 0 reduces the maximum AND stores with the
            // shift predicted from the previous step's maximum (a.pred[slot]), tile by tile as the MMAs complete — the
            // epilogue of tile i overlaps the K loop of tile i+1 — and ONE grid barrier both publishes the maximum and
            // completes the layer.  Every CTA then checks the prediction against the true global maximum (a grid-uniform
            // decision); only on a miss does pass 1 store again (resident layers: from tensor memory, still intact because
            // the next layer's MMAs are held back by layer_bar; streaming layers: from a second run of the contraction)
            // behind a second barrier.  What is stored is always rounded with the true exponent.
            // A miss costs a second sweep and a second barrier, so a layer speculates when that is cheap or unlikely: when its
            // CTA pairs hold several tiles (the predicted-store work hides under the K loops of the later tiles, a miss then
            // costs what the two-sweep schedule always pays) or when the slot's prediction held the last `spec_confident`
            // times (a.pred[kDpSlots + slot], updated whether or not the layer speculated).  Measured on VGG-11 with a new
            // batch every step: the window predictor alone missed 23-36 % (maxima hover around powers of two), with the
            // input-scale predictor 21 % miss — at that rate speculating always (spec_confident = 0, the default) beats waiting
            // for two holds: 444.9 vs 448.9 us/step.
            // Two predictors (common.cuh): the window over the slot's recent bit widths, and last max|acc| scaled by how much of
            // the int8 range this layer's INPUT fills now against last time — the norm the producing op left behind (previous
            // layer of this chain: in a register; first layer: the context word the previous fused launch wrote).
            const uint32_t in_now = l == 0 ? __ldcg(a.pred + kPredLastNorm) : prev_norm;
            bool espec;
            {   // (the prediction record is not kept live over the layer: CTA 0 re-reads the same words when it folds the truth in)
                const NitiPrediction pr = niti_predict(a.pred, edps, in_now);
                espec = a.speculate == 2 || (a.speculate == 1 && ((pr.confw & 0xffffu) >= (uint32_t)a.spec_confident || eitems > epairs));
                niti_rule_a(pr.bw, kind, shift);
            }
            NitiRuleA rk = niti_rule_a_consts(shift, eone);
            int mxs = 0, mns = 0;                            // signed running max / min: one VIMNMX3 each per two accumulators
            for (int pass = 0; pass < 2; ++pass) {
                uint32_t slot = 0;
                if (active)
                for (int item = pair; item < eitems; item += epairs) {
                    const int n_t = item % ent; const int m_t = item / ent;
                    const int row0 = m_t * 256 + (int)rank * 128 + quarter * 32;
                    const int row = row0 + lane;
                    const int n0 = n_t * ebn;
                    const uint32_t myslot = slot;
                    if (++slot == eslots) slot = 0;
                    if (pass == 0 || estream) {              // a fresh set of accumulators: wait for its last MMA
                        tc::mbar_wait_backoff(&tfull_bar[myslot], (tuse >> myslot) & 1u);
                        tuse ^= 1u << myslot;
                        tc::tc_fence_after();
                        if (etid == 0 && item == pair && pass == 0) chain_trace(a.trace, l, 2);
                    }
                    if (row0 < eM) {                         // warp-uniform
                    const bool row_ok = row < eM;
                    // pixel coordinates of this row (pool / pool-gradient addressing)
                    int img = 0, py = 0, px = 0;
                    if (epool || epgx != nullptr) { img = row / hw; const int r = row - img * hw; py = r / eow; px = r - py * eow; }
#pragma unroll 1
                    for (int c0 = group * 16; c0 < ebn; c0 += 64) {
                        if (n0 + c0 >= eN) break;       // warp-uniform (Ncols % 16 == 0)
                        // the global loads the store needs are issued before the tensor-memory read so the latencies overlap
                        const int col = n0 + c0;
                        uint4 x00, x01, x10, x11, gv;
                        size_t b00 = 0, b10 = 0;
                        if (row_ok && (pass == 1 || espec)) {
                            if (epgx != nullptr) {
                                const int w2 = eow * 2;
                                b00 = (((size_t)img * (eoh * 2) + 2 * py) * w2 + 2 * px) * eN + col;
                                b10 = b00 + (size_t)w2 * eN;
                                x00 = ldg_u4(epgx + b00); x01 = ldg_u4(epgx + b00 + eN);
                                x10 = ldg_u4(epgx + b10); x11 = ldg_u4(epgx + b10 + eN);
                            } else if (egate != nullptr) {
                                gv = ldg_u4(egate + (size_t)row * eN + col);
                            }
                        }
                        uint32_t v[16];
                        tmem_ld16(tmem_base + ((uint32_t)(quarter * 32) << 16) + myslot * (uint32_t)ebn + (uint32_t)c0, v);
                        tc::tmem_ld_wait();
                        if (pass == 0) {
                            if (row_ok) {
#pragma unroll
                                for (int q = 0; q < 16; q += 2) {
                                    mxs = __vimax3_s32(mxs, (int32_t)v[q], (int32_t)v[q + 1]);
                                    mns = __vimin3_s32(mns, (int32_t)v[q], (int32_t)v[q + 1]);
                                }
                            }
                            if (!espec) continue;
                        }
                        uint32_t o[4];
                        niti_rescale16(v, kind, shift, rk, o, erelu != 0);
                        if (epgx == nullptr) {
                            const size_t off = (size_t)row * eN + col;
                            if (row_ok) {
                                if (egate != nullptr) {     // NITI_CPUReluGrad: pass where the forward value was > 0
                                    const uint4 g = vgt0(gv);
                                    o[0] &= g.x; o[1] &= g.y; o[2] &= g.z; o[3] &= g.w;
                                }
                                *reinterpret_cast<uint4*>(ey + off) = make_uint4(o[0], o[1], o[2], o[3]);
                            }
                            if (epool) {
                                // 2x2/2 max pool across lanes: lane bit 0 = x parity, lane bit log2(out_w) = y parity
                                uint32_t m[4];
#pragma unroll
                                for (int q = 0; q < 4; ++q) {
                                    m[q] = __vmaxs4(o[q], __shfl_xor_sync(0xffffffffu, o[q], 1));
                                    m[q] = __vmaxs4(m[q], __shfl_xor_sync(0xffffffffu, m[q], eow));
                                }
                                if (row_ok && !(px & 1) && !(py & 1)) {
                                    const size_t po = (((size_t)img * (eoh >> 1) + (py >> 1)) * (eow >> 1) + (px >> 1)) * eN + col;
                                    *reinterpret_cast<uint4*>(eyp + po) = make_uint4(m[0], m[1], m[2], m[3]);
                                }
                            }
                        } else if (row_ok) {
                            // max-pool gradient (+ ReLU-grad gate): this row is one pooled pixel; route its gradient to the
                            // first window element (row-major) that equals the window maximum, zero elsewhere
                            const uint4 mxv = vmax4(vmax4(x00, x01), vmax4(x10, x11));
                            const uint4 g = make_uint4(o[0], o[1], o[2], o[3]);
                            uint4 e0 = vge4(x00, mxv), done = e0;
                            uint4 e1 = vandn4(vge4(x01, mxv), done); done = vor4(done, e1);
                            uint4 e2 = vandn4(vge4(x10, mxv), done); done = vor4(done, e2);
                            uint4 e3 = vandn4(vge4(x11, mxv), done);
                            if (epgg) { e0 = vand4(e0, vgt0(x00)); e1 = vand4(e1, vgt0(x01)); e2 = vand4(e2, vgt0(x10)); e3 = vand4(e3, vgt0(x11)); }
                            *reinterpret_cast<uint4*>(ey + b00) = vand4(g, e0);
                            *reinterpret_cast<uint4*>(ey + b00 + eN) = vand4(g, e1);
                            *reinterpret_cast<uint4*>(ey + b10) = vand4(g, e2);
                            *reinterpret_cast<uint4*>(ey + b10 + eN) = vand4(g, e3);
                        }
                    }
                    }
                    if (estream) {                           // this warp is done with the slot: hand it back to the MMA issuer
                        tc::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_remote_relaxed(&tempty_bar[myslot], 0u);
                    }
                }
                if (pass == 0) {
                    // ---- global max|acc|: warp max -> one atomicMax per CTA -> grid barrier -> (data parallel: peer exchange) ----
                    // (16 x 148 atomics on one word serialise in L2 for microseconds, hence the shared-memory stage)
                    mx = warp_max_u32((uint32_t)max(mxs, -mns));
                    if (lane == 0 && mx) atomicMax(cta_max, mx);
                    tc::tc_fence_before();
                    asm volatile("bar.sync 1, 512;" ::: "memory");       // this CTA's maximum is final and its stores are issued
                    if (a.trigger_late && last_layer) niti_pdl_trigger();
                    if (etid == 0) {
                        const uint32_t cm = *cta_max; *cta_max = 0u;
                        if (cm) atomicMax(emax, cm);
                        chain_trace(a.trace, l, 3);
                        ++nbar; niti_grid_barrier(a.sync, nbar * gridDim.x);
                        chain_trace(a.trace, l, 4);
                        *bcast = dp_exchange_max(a.dp, edps, __ldcg(emax), blockIdx.x == 0);
                    }
                    asm volatile("bar.sync 1, 512;" ::: "memory");
                    const uint32_t gmax = *bcast;
                    int kind2, shift2;
                    niti_rule_a(niti_bitwidth(gmax), kind2, shift2);
                    const bool held = kind2 == kind && shift2 == shift;
                    prev_norm = niti_out_norm(gmax, kind2, shift2);      // the next layer's input norm (grid-uniform)
                    if (blockIdx.x == 0 && etid == 0) {
                        if (eeo != nullptr) *eeo = (int8_t)((eea ? (int)*eea : 0) + (eeb ? (int)*eeb : 0) + (kind2 == 1 ? 0 : shift2));
                        niti_predict_update(a.pred, edps, niti_predict(a.pred, edps, in_now), gmax, held, espec);   // every CTA read the old words before it arrived above
                        if (last_layer) a.pred[kPredLastNorm] = prev_norm;          // for the next fused launch of the stream
                    }
                    if (espec && held) {                                 // prediction held (grid-uniform): the layer is complete
                        if (etid == 0) chain_trace(a.trace, l, 5);
                        break;
                    }
                    kind = kind2; shift = shift2;
                    rk = niti_rule_a_consts(shift, eone);
                    // run the contraction once more.  Only pairs that hold tiles are told: their issuing warps cannot fall a
                    // phase behind (the epilogue needs their work before its next arrive); an idle pair sees one arrive per layer
                    if (estream && active && etid == 0) { *flag = 1u; tc::mbar_arrive(layer_bar); }
                } else {
                    // second pass done in this CTA; the next layer needs it complete everywhere
                    tc::tc_fence_before();
                    asm volatile("bar.sync 1, 512;" ::: "memory");
                    if (etid == 0) {
                        chain_trace(a.trace, l, 5);
                        if (!last_layer) { ++nbar; niti_grid_barrier(a.sync, nbar * gridDim.x); }
                    }
                }
            }
            // resident layers: every slot this pair used goes back to the MMA issuer now (all reads of it are done)
            if (kStream && active && !estream) {
                tc::tc_fence_before();
                __syncwarp();
                const int mine = (eitems - pair + epairs - 1) / epairs;
                if (lane == 0)
                    for (int t = 0; t < mine; ++t) mbar_arrive_remote_relaxed(&tempty_bar[t], 0u);
            }
            if (etid == 0 && (kStream || !last_layer)) {
                *flag = 0u;
                tc::mbar_arrive(layer_bar);                                        // release this CTA's producer (and MMA issuer)
                chain_trace(a.trace, l, 6);
            }
        }
        if (etid == 0) {                                     // last CTA out re-arms the grid barrier for the next launch
            __threadfence();
            if (atomicAdd(a.sync + 1, 1u) == gridDim.x - 1) { a.sync[0] = 0u; a.sync[1] = 0u; __threadfence(); }
        }
    }
    __syncwarp();
    tc::tc_fence_before();
    cluster_sync_all();
    if (warp == 1) tmem_dealloc2<512>(tmem_base);
    if (tid == 64) chain_trace(a.trace, a.nlayers - 1, 7);
    niti_tl_exit(a.tl);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
// tile shape + schedule: resident when every pair can keep all its tiles in its 512 TMEM columns — then the largest BN that
// still gives >= ~2/3 of the CTA pairs a tile — otherwise streaming (slots recycled, conv_chain.cu kernel comment) with the
// widest tile the output allows (operand bytes per MMA fall with BN)
bool chain_plan(int max_pairs, int M, int Ncols, int* bn_out, int* pairs_out, int* items_out, int* stream_out = nullptr) {
    const long m_tiles = (M + 255) / 256;
    int bn = 64;
    for (int cand : {256, 128}) {
        if (Ncols > cand / 2 && m_tiles * ((Ncols + cand - 1) / cand) >= (2 * max_pairs) / 3) { bn = cand; break; }
    }
    if (Ncols <= 64) bn = 64; else if (Ncols <= 128 && bn > 128) bn = 128;
    long items = m_tiles * ((Ncols + bn - 1) / bn);
    int pairs = (int)(items < max_pairs ? items : max_pairs);
    long per_pair = (items + pairs - 1) / pairs;
    int stream = 0;
    if (per_pair * bn > 512 || per_pair > kChainMaxTiles) {
        stream = 1;
        bn = Ncols > 128 ? 256 : (Ncols > 64 ? 128 : 64);
        items = m_tiles * ((Ncols + bn - 1) / bn);
        if (items > 0x3fffffff) return false;
        pairs = (int)(items < max_pairs ? items : max_pairs);
    }
    *bn_out = bn; *pairs_out = pairs; *items_out = (int)items;
    if (stream_out) *stream_out = stream;
    return true;
}

bool pool_geometry_ok(int out_h, int out_w) {
    return (out_w == 2 || out_w == 4 || out_w == 8 || out_w == 16) && out_h % 2 == 0 && (out_h * out_w) % 2 == 0;
}

}  // namespace

// One op of a chain as the C ABI describes it (niti_b200.h niti_chain_op), lowered to the kernel's layer record.
// Returns NITI_OK, or NITI_ERR_UNSUPPORTED when the op cannot run in the fused kernel (caller falls back).
int chain_lower(niti_ctx* ctx, const niti_chain_op* op, ChainLayer* L) {
    if (!tmax::driver_ok()) return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncode* entry points unavailable");
    const niti_conv_desc* d = &op->d;
    const int oh = conv_out_h(d), ow = conv_out_w(d);
    const int max_pairs = ctx->sm_count / 2;
    *L = ChainLayer{};
    int bn, pairs, items, stream = 0;
    if (op->op == 0) {
        if (!tma2_fwd_eligible(d) || d->oc % 16 != 0) return NITI_ERR_UNSUPPORTED;
        L->M = d->n * oh * ow; L->Ncols = d->oc;
        if (!chain_plan(max_pairs, L->M, L->Ncols, &bn, &pairs, &items, &stream)) return NITI_ERR_UNSUPPORTED;
        L->bkc = d->c % 128 == 0 ? 128 : 64;
        L->out_h = oh; L->out_w = ow; L->kh = d->kh; L->kw = d->kw; L->dil = d->dil; L->stride = d->stride;
        L->lower_h = -d->pad; L->lower_w = -d->pad; L->flip = 0;
        L->cchunks = d->c / L->bkc; L->num_kb = d->kh * d->kw * L->cchunks;
        if (!tmax::make_im2col_u8(&L->mapA, op->in, d->n, d->h, d->w, d->c, -d->pad, -d->pad, d->pad - (d->kh - 1) * d->dil,
                                  d->pad - (d->kw - 1) * d->dil, d->stride, L->bkc, 128, L->bkc))
            return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(x) failed");
        if (!tmax::make_tiled_u8(&L->mapB, op->w, (uint64_t)d->oc, (uint64_t)d->kh * d->kw * d->c, (uint32_t)(bn / 2), (uint32_t)L->bkc, L->bkc))
            return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(w) failed");
        if (op->out_pool != nullptr) {
            if (!pool_geometry_ok(oh, ow)) return NITI_ERR_UNSUPPORTED;
            L->pool = 1; L->y_pool = op->out_pool;
        }
        L->relu = op->relu; L->exp_a = op->ex; L->exp_b = op->ew; L->exp_out = op->ey;
    } else {
        if (!tma2_dx_eligible(d) || d->c % 16 != 0) return NITI_ERR_UNSUPPORTED;
        const int pph = (d->kh - 1) * d->dil - d->pad, ppw = (d->kw - 1) * d->dil - d->pad;   // padding of the equivalent forward conv
        L->M = d->n * d->h * d->w; L->Ncols = d->c;
        if (!chain_plan(max_pairs, L->M, L->Ncols, &bn, &pairs, &items, &stream)) return NITI_ERR_UNSUPPORTED;
        L->bkc = d->oc % 128 == 0 ? 128 : 64;
        L->out_h = d->h; L->out_w = d->w; L->kh = d->kh; L->kw = d->kw; L->dil = d->dil; L->stride = 1;
        L->lower_h = -pph; L->lower_w = -ppw; L->flip = 1;
        L->cchunks = d->oc / L->bkc; L->num_kb = d->kh * d->kw * L->cchunks;
        if (!tmax::make_im2col_u8(&L->mapA, op->in, d->n, oh, ow, d->oc, -pph, -ppw, d->h - oh - pph, d->w - ow - ppw, 1, L->bkc, 128, L->bkc))
            return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeIm2col(dy) failed");
        if (!tmax::make_tiled_u8(&L->mapB, op->w, (uint64_t)d->c, (uint64_t)d->kh * d->kw * d->oc, (uint32_t)(bn / 2), (uint32_t)L->bkc, L->bkc))
            return niti_fail(ctx, NITI_ERR_CUDA, "cuTensorMapEncodeTiled(wt) failed");
        L->gate = op->gate;
        if (op->pool_x != nullptr) { L->pg_x = op->pool_x; L->pg_gate = op->pool_gate; L->gate = nullptr; }
    }
    L->n_tiles = (L->Ncols + bn - 1) / bn; L->items = items; L->pairs = pairs; L->bn = bn;
    L->nslots = 512 / bn > kChainMaxTiles ? kChainMaxTiles : 512 / bn; L->stream = stream;
    const int stage = 128 * L->bkc + (bn / 2) * L->bkc;
    L->stages = kChainStageArea / stage > kChainMaxStages ? kChainMaxStages : kChainStageArea / stage;
    L->y = op->out; L->maxabs = op->maxabs; L->dp_slot = op->dp_slot;
    return NITI_OK;
}

struct niti_chain {
    ChainLayer* table;                // device
    int nlayers, grid;
    bool streams;                     // some layer recycles its accumulator slots
    std::vector<ChainLayer> host;
};

template <bool kTable, bool kStream>
static int chain_launch_t(niti_ctx* ctx, const ChainLayer& byvalue, const ChainArgs& a, int grid, cudaStream_t st) {
    if (int rc = niti_func_smem(ctx, (const void*)chain_kernel<kTable, kStream>, kChainSmem, "cudaFuncSetAttribute(chain_kernel)")) return rc;
    cudaError_t e = niti_launch(ctx, NITI_PDL_FUSED, chain_kernel<kTable, kStream>, dim3(grid), dim3(kChainThreads), (size_t)kChainSmem, st, byvalue, a);
    if (e != cudaSuccess) return niti_fail(ctx, NITI_ERR_CUDA, "launch chain_kernel", e);
    return niti_check_launch(ctx, "chain_kernel");
}

static int chain_launch(niti_ctx* ctx, const ChainLayer* single, const ChainLayer* table, int nlayers, int grid, bool streams, cudaStream_t st) {
    ChainArgs a{};
    a.table = table; a.nlayers = nlayers; a.sync = ctx->ws_u32 + 52; a.dp = ctx->dp; a.trace = ctx->trace; a.pred = ctx->pred; a.speculate = ctx->speculate; a.spec_confident = ctx->spec_confident;
    a.tl = niti_tl_next(ctx, nlayers > 1 ? "chain" : (single && single->flip ? "dx_fused" : "fwd_fused"));
    a.trigger_late = (ctx->pdl & NITI_PDL_TRIGGER_LATE) ? 1 : 0;
    static const ChainLayer zero{};
    if (table) return streams ? chain_launch_t<true, true>(ctx, zero, a, grid, st) : chain_launch_t<true, false>(ctx, zero, a, grid, st);
    return streams ? chain_launch_t<false, true>(ctx, *single, a, grid, st) : chain_launch_t<false, false>(ctx, *single, a, grid, st);
}

// single op through the chain kernel (the fused forward / dX entry points of api.cu)
int chain_run_single(niti_ctx* ctx, const niti_chain_op* op, cudaStream_t st) {
    ChainLayer L;
    int rc = chain_lower(ctx, op, &L);
    if (rc != NITI_OK) return rc;
    return chain_launch(ctx, &L, nullptr, 1, 2 * L.pairs, L.stream != 0, st);
}

bool chain_op_eligible(niti_ctx* ctx, const niti_chain_op* op) {
    const niti_conv_desc* d = &op->d;
    const int max_pairs = ctx->sm_count / 2;
    int bn, pairs, items;
    if (op->op == 0) {
        if (!tma2_fwd_eligible(d) || d->oc % 16 != 0) return false;
        if (!chain_plan(max_pairs, d->n * conv_out_h(d) * conv_out_w(d), d->oc, &bn, &pairs, &items)) return false;
        return op->out_pool == nullptr || pool_geometry_ok(conv_out_h(d), conv_out_w(d));
    }
    if (!tma2_dx_eligible(d) || d->c % 16 != 0) return false;
    return chain_plan(max_pairs, d->n * d->h * d->w, d->c, &bn, &pairs, &items);
}

extern "C" int niti_chain_op_eligible(niti_ctx* ctx, const niti_chain_op* op) {
    if (!ctx || !op) return 0;
    if (!(ctx->path == NITI_PATH_AUTO || ctx->path == NITI_PATH_TMA2) || !ctx->auto_tma2 || !ctx->use_chain) return 0;
    return chain_op_eligible(ctx, op) ? 1 : 0;
}

extern "C" int niti_chain_create(niti_ctx* ctx, const niti_chain_op* ops, int n, niti_chain** out) {
    if (!ctx || !ops || n <= 0 || !out) return niti_fail(ctx, NITI_ERR_ARG, "niti_chain_create: bad argument");
    *out = nullptr;
    niti_chain* c = new (std::nothrow) niti_chain();
    if (!c) return niti_fail(ctx, NITI_ERR_CUDA, "niti_chain_create: out of host memory");
    c->host.resize(n);
    int max_pairs = 0;
    for (int i = 0; i < n; ++i) {
        if (!ops[i].in || !ops[i].w || !ops[i].out || !ops[i].maxabs || ops[i].dp_slot < 0 || ops[i].dp_slot >= kDpSlots) {
            delete c; return niti_fail(ctx, NITI_ERR_ARG, "niti_chain_create: op with a null operand or a slot out of range");
        }
        int rc = chain_lower(ctx, &ops[i], &c->host[i]);
        if (rc != NITI_OK) {
            delete c;
            return rc == NITI_ERR_UNSUPPORTED ? niti_fail(ctx, rc, "niti_chain_create: an op is outside the fused kernel's contract") : rc;
        }
        if (c->host[i].pairs > max_pairs) max_pairs = c->host[i].pairs;
        if (c->host[i].stream) c->streams = true;
    }
    c->nlayers = n; c->grid = 2 * max_pairs;
    cudaError_t e = cudaMalloc(&c->table, sizeof(ChainLayer) * (size_t)n);
    if (e != cudaSuccess) { delete c; return niti_fail(ctx, NITI_ERR_CUDA, "niti_chain_create: cudaMalloc", e); }
    e = cudaMemcpy(c->table, c->host.data(), sizeof(ChainLayer) * (size_t)n, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { cudaFree(c->table); delete c; return niti_fail(ctx, NITI_ERR_CUDA, "niti_chain_create: upload", e); }
    *out = c;
    return NITI_OK;
}

extern "C" int niti_chain_run(niti_ctx* ctx, niti_chain* c, niti_stream st) {
    if (!ctx || !c || !c->table) return niti_fail(ctx, NITI_ERR_ARG, "niti_chain_run: bad argument");
    ctx->last_path = "chain";
    return chain_launch(ctx, nullptr, c->table, c->nlayers, c->grid, c->streams, (cudaStream_t)st);
}

extern "C" int niti_chain_destroy(niti_ctx* ctx, niti_chain* c) {
    if (!c) return NITI_ERR_ARG;
    if (c->table) cudaFree(c->table);
    delete c;
    return NITI_OK;
}
