// spl_batch.cu -- batched multi-query search as a dense Q x E^T contraction on the 5th-gen tensor
// cores (SPL_MODE_BATCH_BF16, config C3).
//
// The reference scores one query at a time (cmd_search is one process per query,
// /root/reference/splinter_cli_cmd_search.c:373-409); for nq queries the same arithmetic is
// S[q][e] = sum_d Q[q][d] * E[e][d], a GEMM with M = nq, N = slots, K = D.  Scores are never
// written out: the epilogue turns every accumulator element into an ordering value and keeps
// only the few that can still reach the top-k.
//
//   gemm_topk_kernel  persistent, one CTA per SM, warp-specialised:
//       warp 0   TMA producer: Q tile (128 x 64 bf16) + E tile (256 x 64 bf16) per k-block into a
//                4-stage SWIZZLE_128B ring (cp.async.bulk.tensor.2d, mbarrier complete_tx)
//       warp 1   one elected thread issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=256, K=16),
//                fp32 accumulators in TMEM (2 x 256 columns, double-buffered against the epilogue);
//                tcgen05.commit releases smem stages / publishes accumulators
//       warps 2-3 per-tile column coefficients (label predicate, row norms) -> smem, one tile ahead
//       warps 4-7 epilogue: tcgen05.ld 32 columns at a time; thread = query (TMEM lane);
//                t = fma(s, cA[col], cB[col])  (L2: |e|^2 - 2s  [+|q|^2 is constant per query];
//                cosine: -s/|e|); pass 1 keeps min t per (query, tile), pass 2 appends the rare
//                t < thr[q] to a global candidate list
//   threshold_kernel  k'-th smallest per-tile minimum of a strided sample of tiles = a valid upper
//                bound of the k'-th smallest t over the whole shard (each such tile holds one
//                element <= it), tight because the best rows rarely share a tile
//   select_rerank_kernel  per query: sort the candidates by t, re-score the best k' = 2k from the
//                fp32 mirror with the same formulas as the single-query scan, sort by the
//                reference's comparator and emit k records.  Final scores are fp32-exact; bf16 only
//                decides which k' rows are re-scored.
#include <cuda.h>
#include <cuda_bf16.h>
#include <float.h>
#include <stdlib.h>

#include "spl_batch.cuh"
#include "spl_scan_epilogue.cuh"

namespace spl {

namespace {

constexpr int BM = 128, BN = 256, BK = 64, UK = 16;
constexpr int kStages = 4, kAcc = 2;
constexpr int kThreadsGemm = 256;
constexpr uint32_t kStageA = BM * BK * 2, kStageB = BN * BK * 2, kStageBytes = kStageA + kStageB;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kMaxSampledTiles = 16384;  // pass-1 sample size cap (a power of two): 192 KB of sort space per query

// ---- smem carve-up (dynamic, 1024-byte aligned for SWIZZLE_128B) ----
struct GemmSmem {
    static constexpr uint32_t stages_off = 0;
    static constexpr uint32_t coef_off = kStages * kStageBytes;              // cA[2][BN], cB[2][BN]
    static constexpr uint32_t bar_off = coef_off + 2 * 2 * BN * 4;
    static constexpr uint32_t total = bar_off + 256;
};

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint32_t b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(c)); }
__device__ __forceinline__ void bar_arrive(uint32_t b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(b) : "memory"); }
__device__ __forceinline__ void bar_expect_tx(uint32_t b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bar_wait(uint32_t b, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(b), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // clears the CTA-rank bit of a shared::cluster address: "CTA 0 of the pair"
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
    // issued by both CTAs of a pair; the bytes are accounted on the LEADER's mbarrier
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bar_arrive_leader(uint32_t b) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(b & kPeerBitMask) : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {  // arrives on the same barrier in BOTH CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_f16_pair(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor layout):
// start address >> 4 in [0,14), LBO (ignored for swizzled K-major) = 1 in [16,30), SBO = 1024 B >> 4
// in [32,46), descriptor version 1 in [46,48), layout type SWIZZLE_128B = 2 in [61,64).
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16 instruction descriptor (cute::UMMA::InstrDescriptor): fp32 accumulate (bit 4), A = B = bf16
// (bits 7, 10), both K-major, N >> 3 at [17,23), M >> 4 at [24,29).
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
constexpr uint32_t kIdescPair = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
// Pair variant.  Per k-block each CTA needs its own 128 queries (16 KB, always an L2 hit: the whole query operand is
// 512 KB) and HALF of the 256-row operand (16 KB, from HBM).  The two operands get rings of their own, made of 16 KB
// units.  What was measured while looking for the main loop's limiter (C3: 256 x 2^20 x 1024, pass 2 = 420 us):
//   * re-fetching the query tile for every row tile doubles the L2 -> SM bytes (4.3 GB against 2.16 GB of rows), but
//     it is NOT what paces the loop: a probe that skipped the query loads after the first tile (wrong results, same
//     MMAs) ran only 2.5 % faster (0.480 -> 0.468 ms per batch);
//   * keeping half of the query tile resident in shared memory (L2 -> SM 3.2 GB, but only 5 units left for the ring)
//     was SLOWER (494 us): ring depth matters more than those bytes;
//   * ring shapes beyond 4 query units + the rest for rows change nothing (query units 2 / 3 / 4 / 5 / 6: 0.551 /
//     0.499 / 0.490 / 0.490 / 0.490 ms per batch).
// The launch runs at 1.30-1.35 PFLOP/s at a power-capped 1.42 GHz -- 93-96 % of what cuBLAS sustains under the same cap.
constexpr uint32_t kUnitBytes = kStageA;                    // 16 KB: 128 rows x 64 bf16, one SWIZZLE_128B operand block
constexpr uint32_t kStageBPair = (BN / 2) * BK * 2;
static_assert(kStageBPair == kUnitBytes, "half of the row operand is one unit");
constexpr uint32_t kPairQUnits = 4, kPairEUnitsMax = 11;
constexpr uint32_t kPairBarBytes = 512;                     // up to 56 barriers
constexpr uint32_t kPairSmemBudget = 227u * 1024u - 2048u;  // static shared memory (TMEM base word, ...) takes the rest
struct PairLayout { uint32_t q_units, e_units, e_off, coef_off, bar_off, total; };
__host__ __device__ inline PairLayout pair_layout(uint32_t q_units_req) {
    PairLayout L;
    L.q_units = q_units_req >= 2 && q_units_req <= 6 ? q_units_req : kPairQUnits;
    L.e_off = L.q_units * kUnitBytes;
    const uint32_t fixed = 2 * 2 * BN * 4 + kPairBarBytes;  // coefficients + barriers
    const uint32_t u = (kPairSmemBudget - L.e_off - fixed) / kUnitBytes;
    L.e_units = u > kPairEUnitsMax ? kPairEUnitsMax : u;
    L.coef_off = L.e_off + L.e_units * kUnitBytes;
    L.bar_off = L.coef_off + 2 * 2 * BN * 4;
    L.total = L.bar_off + kPairBarBytes;
    return L;
}

}  // namespace

// -------------------------------------------------------------------------------------------------
// PAIR = false: one CTA = 128 queries x 256 rows per tile (cta_group::1).
// PAIR = true : a 2-CTA cluster = 256 queries x 256 rows per tile (cta_group::2, M = 256): each CTA stages its own
//               128 queries and half of the row operand, so the bytes every SM pulls from L2 per MMA drop by a third
//               and six stages fit; CTA 0 issues the MMAs for both, each CTA drains its own 128 TMEM lanes.
template <bool PAIR>
__global__ void __launch_bounds__(kThreadsGemm, 1)
gemm_topk_kernel(const __grid_constant__ CUtensorMap mapQ, const __grid_constant__ CUtensorMap mapE, const BatchArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t s_tmem_base;
    const PairLayout PL = pair_layout(a.pair_q_units);      // (PAIR only)
    constexpr uint32_t NS = kStages, STAGE = kStageBytes;   // (1-CTA kernel: 48 KB stages holding both operands)
    const uint32_t NQU = PL.q_units, NEU = PL.e_units;      // (PAIR: separate rings of 16 KB units)
    const uint32_t COEF_OFF = PAIR ? PL.coef_off : GemmSmem::coef_off;
    const uint32_t BAR_OFF = PAIR ? PL.bar_off : GemmSmem::bar_off;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = PAIR ? cluster_rank() : 0u;
    const bool leader = crank == 0;
    const uint32_t bars = s32(smem + BAR_OFF);
    // barrier map: full[s] 0..5, empty[s] 6..11, tfull[a] 12..13, tempty[a] 14..15, cfull[c] 16..17, cempty[c] 18..19;
    // PAIR: query ring full 20..25 / empty 26..31, row ring full 32..43 / empty 44..55
    auto FULL = [&](uint32_t s) { return bars + 8 * s; };
    auto EMPTY = [&](uint32_t s) { return bars + 8 * (6 + s); };
    auto TFULL = [&](uint32_t x) { return bars + 8 * (12 + x); };
    auto TEMPTY = [&](uint32_t x) { return bars + 8 * (14 + x); };
    auto CFULL = [&](uint32_t x) { return bars + 8 * (16 + x); };
    auto CEMPTY = [&](uint32_t x) { return bars + 8 * (18 + x); };
    auto QFULLB = [&](uint32_t u) { return bars + 8 * (20 + u); };
    auto QEMPTYB = [&](uint32_t u) { return bars + 8 * (26 + u); };
    auto EFULLB = [&](uint32_t u) { return bars + 8 * (32 + u); };
    auto EEMPTYB = [&](uint32_t u) { return bars + 8 * (44 + u); };
    float *cA = reinterpret_cast<float *>(smem + COEF_OFF);  // [2][BN]
    float *cB = cA + 2 * BN;                                 // [2][BN]

    if (threadIdx.x == 0) {
        // PAIR: the leader's FULL barrier collects its own arrive.expect_tx + the peer's remote arrive, and its
        // TEMPTY barrier the 4 epilogue warps of BOTH CTAs
        if (PAIR) {
            for (uint32_t u = 0; u < NQU; u++) { bar_init(QFULLB(u), 2); bar_init(QEMPTYB(u), 1); }
            for (uint32_t u = 0; u < NEU; u++) { bar_init(EFULLB(u), 2); bar_init(EEMPTYB(u), 1); }
        } else {
            for (uint32_t s = 0; s < NS; s++) { bar_init(FULL(s), 1); bar_init(EMPTY(s), 1); }
        }
        for (uint32_t x = 0; x < kAcc; x++) {
            bar_init(TFULL(x), 1); bar_init(TEMPTY(x), PAIR ? 8 : 4);
            bar_init(CFULL(x), 64); bar_init(CEMPTY(x), 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {  // TMEM allocation: one warp (the same warp index in both CTAs of a pair)
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&s_tmem_base)), "n"(kTmemCols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&s_tmem_base)), "n"(kTmemCols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();  // the peer's barriers must be initialised before anything signals them
    tc_fence_after();
    const uint32_t tmem_base = s_tmem_base;

    // ---- static tile schedule: unit = CTA (or CTA pair) = (query tile [pair], group); groups stride over the tiles ----
    const uint32_t unit = PAIR ? blockIdx.x / 2 : blockIdx.x, nunits = PAIR ? gridDim.x / 2 : gridDim.x;
    const uint32_t qunits = PAIR ? a.n_qtiles / 2 : a.n_qtiles;   // query tiles (pairs of tiles) in flight
    const uint32_t qtile = PAIR ? (unit % qunits) * 2 + crank : unit % qunits;
    const uint32_t group = unit / qunits, ngroups = nunits / qunits;
    const uint32_t kblocks = a.dim / BK;
    const bool active = group < ngroups;  // units beyond a whole number of groups idle

    if (warp == 0) {
        if (lane == 0 && active) {
            uint32_t st = 0, ph = 0, qu = 0, qph = 0, eu = 0, eph = 0;
            for (uint32_t j = group; j < a.n_tiles; j += ngroups) {
                const int row0 = (int)((a.tile_first + j * a.tile_stride) * BN);
                for (uint32_t kb = 0; kb < kblocks; kb++) {
                    // both operands are stored tiled ([tile][k-block][rows][64]): each load is one contiguous burst
                    if (PAIR) {
                        // Each ring unit: wait until the MMAs that read its previous contents have retired, arm its barrier
                        // (leader: expect both CTAs' 16 KB; peer: arrive remotely), load.  The row operand first: it is
                        // the one that comes from HBM.
                        bar_wait(EEMPTYB(eu), eph ^ 1u);
                        if (leader) bar_expect_tx(EFULLB(eu), 2 * kUnitBytes); else bar_arrive_leader(EFULLB(eu));
                        tma_load_2d_pair(s32(smem + PL.e_off + eu * kUnitBytes), &mapE, EFULLB(eu), 0,
                                         (int)((((uint32_t)row0 / BN) * kblocks + kb) * BN + crank * (BN / 2)));
                        if (++eu == NEU) { eu = 0; eph ^= 1u; }
                        bar_wait(QEMPTYB(qu), qph ^ 1u);
                        if (leader) bar_expect_tx(QFULLB(qu), 2 * kUnitBytes); else bar_arrive_leader(QFULLB(qu));
                        tma_load_2d_pair(s32(smem + qu * kUnitBytes), &mapQ, QFULLB(qu), 0, (int)((qtile * kblocks + kb) * BM));
                        if (++qu == NQU) { qu = 0; qph ^= 1u; }
                    } else {
                        bar_wait(EMPTY(st), ph ^ 1u);
                        const uint32_t sa = s32(smem + st * STAGE), sb = sa + kStageA;
                        bar_expect_tx(FULL(st), STAGE);
                        tma_load_2d(sa, &mapQ, FULL(st), 0, (int)((qtile * kblocks + kb) * BM));
                        tma_load_2d(sb, &mapE, FULL(st), 0, (int)((((uint32_t)row0 / BN) * kblocks + kb) * BN));
                        if (++st == NS) { st = 0; ph ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && active && leader) {
            uint32_t st = 0, ph = 0, acc = 0, aph = 0, qu = 0, qph = 0, eu = 0, eph = 0;
            for (uint32_t j = group; j < a.n_tiles; j += ngroups) {
                bar_wait(TEMPTY(acc), aph ^ 1u);
                tc_fence_after();
                for (uint32_t kb = 0; kb < kblocks; kb++) {
                    if (PAIR) {
                        bar_wait(QFULLB(qu), qph);
                        bar_wait(EFULLB(eu), eph);
                        tc_fence_after();
                        const uint64_t da = make_desc(s32(smem + qu * kUnitBytes)), db = make_desc(s32(smem + PL.e_off + eu * kUnitBytes));
#pragma unroll
                        for (uint32_t k = 0; k < BK / UK; k++)
                            tc_mma_f16_pair(tmem_base + acc * BN, da + 2 * k, db + 2 * k, kIdescPair, (kb | k) != 0 ? 1u : 0u);
                        // the units are free (in both CTAs of the pair) once these MMAs have read them
                        tc_commit_pair(QEMPTYB(qu));
                        tc_commit_pair(EEMPTYB(eu));
                        if (++qu == NQU) { qu = 0; qph ^= 1u; }
                        if (++eu == NEU) { eu = 0; eph ^= 1u; }
                    } else {
                        bar_wait(FULL(st), ph);
                        tc_fence_after();
                        const uint32_t sa = s32(smem + st * STAGE), sb = sa + kStageA;
                        const uint64_t da = make_desc(sa), db = make_desc(sb);
#pragma unroll
                        for (uint32_t k = 0; k < BK / UK; k++)
                            tc_mma_f16(tmem_base + acc * BN, da + 2 * k, db + 2 * k, kIdesc, (kb | k) != 0 ? 1u : 0u);
                        tc_commit(EMPTY(st));
                        if (++st == NS) { st = 0; ph ^= 1u; }
                    }
                }
                if (PAIR) tc_commit_pair(TFULL(acc)); else tc_commit(TFULL(acc));  // accumulator complete
                if (++acc == kAcc) { acc = 0; aph ^= 1u; }
            }
        }
    } else if (warp < 4) {
        // ---- per-tile column coefficients, one tile ahead of the epilogue ----
        if (active) {
            const uint32_t t = threadIdx.x - 64;  // 0..63, four columns each
            const uint32_t need = a.mask ? SPL_ROW_ENUM : SPL_ROW_LIST;
            // rows that score exactly (zero vector / no embedding) and have k' rows of their own kind before them can
            // never be results: exact_bound_kernel found the row index of the k'-th of each kind
            const uint32_t zero_bound = a.pass == 2 ? a.exact_bound[0] : kIdxSentinel;
            const uint32_t noemb_bound = a.pass == 2 ? a.exact_bound[1] : kIdxSentinel;
            uint32_t cs = 0, cph = 0;
            for (uint32_t j = group; j < a.n_tiles; j += ngroups) {
                const uint32_t row0 = (a.tile_first + j * a.tile_stride) * BN;
                float va[4], vb[4];
                uint32_t nzero = 0, nnoemb = 0;   // pass 1: candidates of either exactly-scoring kind in this tile
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const uint32_t r = row0 + t + 64 * i;
                    float ca = 0.f, cb = INFINITY;
                    if (r < a.n) {
                        const uint32_t f = a.flags[r];
                        const bool pass = (f & need) && ((a.bloom[r] & a.mask) == a.mask) && r != a.exclude;
                        if (pass && (f & SPL_ROW_EMB)) {
                            const float n2 = a.norm2[r];
                            const bool zero_row = n2 == 0.f && a.err2[r] == 0.f;
                            nzero += zero_row;
                            if (zero_row && r > zero_bound) {
                                // a zero vector with k' zero vectors before it: same score, later index -- out
                            } else if (a.order == SPL_ORDER_L2) { ca = -2.f; cb = n2; }
                            else { ca = n2 > 0.f ? -rsqrtf(n2) : 0.f; cb = 0.f; }
                        } else if (pass && r > noemb_bound) {
                            nnoemb++;   // likewise for rows without an embedding copy
                        } else if (pass) {
                            nnoemb++;
                            // no consistent embedding copy: scores like _search.c:388-391 (sim 0, dist FLT_MAX)
                            ca = 0.f; cb = a.order == SPL_ORDER_L2 ? 3.0e38f : 0.f;
                        }
                    }
                    va[i] = ca; vb[i] = cb;
                }
                if (a.pass == 1 && qtile == 0) {   // every query tile sees the same rows: one of them counts
                    nzero = __reduce_add_sync(0xffffffffu, nzero);
                    nnoemb = __reduce_add_sync(0xffffffffu, nnoemb);
                    if (lane == 0 && (nzero | nnoemb)) {
                        if (nzero) atomicAdd(&a.tile_kinds[2 * j], nzero);
                        if (nnoemb) atomicAdd(&a.tile_kinds[2 * j + 1], nnoemb);
                    }
                }
                bar_wait(CEMPTY(cs), cph ^ 1u);
#pragma unroll
                for (int i = 0; i < 4; i++) { cA[cs * BN + t + 64 * i] = va[i]; cB[cs * BN + t + 64 * i] = vb[i]; }
                bar_arrive(CFULL(cs));
                if (++cs == kAcc) { cs = 0; cph ^= 1u; }
            }
        }
    } else {
        // ---- epilogue: thread = query (TMEM lane) ----
        if (active) {
            const uint32_t quad = warp - 4;                       // TMEM lane quadrant this warp may read
            const uint32_t ql = quad * 32 + lane;                 // query within the tile
            const uint32_t q = qtile * BM + ql;
            const bool qvalid = q < a.nq;
            // composite bound (t, row): ties on t -- e.g. thousands of never-embedded zero rows -- are cut by row index
            const float thr = (a.pass == 2 && qvalid) ? a.thr[q] : -INFINITY;
            const uint32_t thr_row = (a.pass == 2 && qvalid) ? a.thr_row[q] : 0u;
            uint32_t acc = 0, aph = 0;
            // selection proof (select_rerank_kernel): the best ordering value among the rows this thread leaves OUT of
            // its query's candidate list, and whether rows tying with the bound were cut off by their row index
            float out_min = INFINITY;
            uint32_t out_tie = 0;
            for (uint32_t j = group; j < a.n_tiles; j += ngroups) {
                const uint32_t row0 = (a.tile_first + j * a.tile_stride) * BN;
                bar_wait(CFULL(acc), aph);
                bar_wait(TFULL(acc), aph);
                tc_fence_after();
                const float *pa = cA + acc * BN, *pb = cB + acc * BN;
                float tmin = INFINITY;
                uint32_t imin = kIdxSentinel;
#pragma unroll 1
                for (uint32_t c0 = 0; c0 < (uint32_t)BN; c0 += 32) {
                    uint32_t v[32];
                    tc_ld32(tmem_base + ((quad * 32u) << 16) + acc * BN + c0, v);
                    const float4 *pa4 = reinterpret_cast<const float4 *>(pa + c0), *pb4 = reinterpret_cast<const float4 *>(pb + c0);
                    // branch-free common path: ordering values + their minimum over the 32 columns
                    float gmin[8];   // minimum of each group of four columns: the slow path below looks only into groups that hit
#pragma unroll
                    for (int i4 = 0; i4 < 8; i4++) {
                        const float4 ca = pa4[i4], cb = pb4[i4];
                        float t0 = fmaf(__uint_as_float(v[4 * i4 + 0]), ca.x, cb.x), t1 = fmaf(__uint_as_float(v[4 * i4 + 1]), ca.y, cb.y);
                        float t2 = fmaf(__uint_as_float(v[4 * i4 + 2]), ca.z, cb.z), t3 = fmaf(__uint_as_float(v[4 * i4 + 3]), ca.w, cb.w);
                        v[4 * i4 + 0] = __float_as_uint(t0); v[4 * i4 + 1] = __float_as_uint(t1);
                        v[4 * i4 + 2] = __float_as_uint(t2); v[4 * i4 + 3] = __float_as_uint(t3);
                        gmin[i4] = fminf(fminf(t0, t1), fminf(t2, t3));
                    }
                    const float cmin = fminf(fminf(fminf(gmin[0], gmin[1]), fminf(gmin[2], gmin[3])),
                                             fminf(fminf(gmin[4], gmin[5]), fminf(gmin[6], gmin[7])));
                    if (a.pass == 1) {
                        if (cmin < tmin) {   // rare after the first few tiles: find the first column holding the new minimum
                            tmin = cmin;
#pragma unroll
                            for (int i = 31; i >= 0; i--) if (__uint_as_float(v[i]) == cmin) imin = row0 + c0 + i;
                        }
                    } else if (cmin <= thr) {  // some column of this chunk may beat the bound (static register indexing)
                        // A thread takes this path for a few percent of its chunks, but a WARP (32 queries) for most of them:
                        // what the whole warp pays must stay small, so only the groups of four that hit are opened.
#pragma unroll
                        for (int j = 0; j < 8; j++) {
                          if (gmin[j] > thr) { out_min = fminf(out_min, gmin[j]); continue; }
#pragma unroll
                          for (int e4 = 0; e4 < 4; e4++) {
                            const int i = 4 * j + e4;
                            const float t = __uint_as_float(v[i]);
                            // t = +inf marks a row that is not a candidate (or lies past the shard's end); it must stay out
                            // even when the bound itself is +inf (fewer sampled tiles than k')
                            if (t < INFINITY && t <= thr && (t < thr || row0 + c0 + i <= thr_row)) {
                                const uint32_t pos = atomicAdd(&a.cand_cnt[q], 1u);
                                if (pos < a.cand_cap) {
                                    a.cand_t[(size_t)q * a.cand_cap + pos] = t;
                                    a.cand_idx[(size_t)q * a.cand_cap + pos] = row0 + c0 + i;
                                }
                            } else if (t < INFINITY && t == thr) {
                                // ties with the bound that the row index cut off: rows whose coefficient is 0 score exactly
                                // (zero vector / no embedding), the others are ordinary rows that happen to tie in bf16
                                // (coefficient 0: no embedding, or a zero vector in cosine order; L2 keeps -2 for every
                                // embedded row, there a zero vector is the one whose norm term is 0)
                                const bool exact_row = pa[c0 + i] == 0.f || (a.order == SPL_ORDER_L2 && pb[c0 + i] == 0.f);
                                out_tie |= exact_row ? 1u : 2u;
                            } else {
                                out_min = fminf(out_min, t);
                            }
                          }
                        }
                    } else {
                        out_min = fminf(out_min, cmin);   // the whole chunk stays out: remember the best value left out
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (PAIR) bar_arrive_leader(TEMPTY(acc)); else bar_arrive(TEMPTY(acc));  // the leader's MMA waits for both CTAs
                    bar_arrive(CEMPTY(acc));
                }
                if (a.pass == 1 && qvalid) {
                    a.tile_min[(size_t)j * a.nq_pad + q] = tmin;
                    a.tile_min_row[(size_t)j * a.nq_pad + q] = imin;
                }
                if (++acc == kAcc) { acc = 0; aph ^= 1u; }
            }
            if (a.pass == 2 && qvalid) {
                if (out_min < INFINITY) atomicMax(&a.out_min[q], ~f2ord(out_min));   // stored inverted: 0 = nothing kept out
                if (out_tie) atomicOr(&a.out_tie[q], out_tie);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();  // neither CTA may free TMEM / exit while the other still uses the pair
    if (warp == 2) {
        tc_fence_after();
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols));
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols));
    }
}

// -------------------------------------------------------------------------------------------------
// Rows whose score does not depend on the row's content -- a zero vector scores (0, |q|), a row without a consistent
// embedding copy (0, FLT_MAX), _search.c:388-391 -- tie with every other row of their kind and are ordered by slot index
// alone.  Only the first k' of each kind can ever be results, whatever the query.  The synthetic corpus of the benchmark
// has 2 % never-embedded zero rows: without this cut every other 32-column chunk of the epilogue holds one, ties with the
// bound and takes the slow per-element path.  Pass 1 (the sampled PREFIX of tiles) counts, per tile, the rows of either
// kind that pass the candidate predicate; exact_bound_block (an extra block of threshold_kernel) turns the counts into
// bound[0] / bound[1] = row index of the k'-th zero / no-embedding candidate, or kIdxSentinel when the prefix holds fewer
// than k' of them (then all are kept: correct, merely not pruned).  Pass 2 drops rows of a kind beyond its bound.
__device__ __forceinline__ uint32_t exact_kind(const BatchArgs &a, uint32_t r, uint32_t need) {   // 0 none, 1 zero, 2 no embedding
    const uint32_t f = a.flags[r];
    if (!(f & need) || (a.bloom[r] & a.mask) != a.mask || r == a.exclude) return 0;
    if (!(f & SPL_ROW_EMB)) return 2;
    return (a.norm2[r] == 0.f && a.err2[r] == 0.f) ? 1u : 0u;
}

// -------------------------------------------------------------------------------------------------
// fp32 queries -> bf16 (zero-padded to a multiple of 128 rows) + |bf16(q)|^2
__global__ void __launch_bounds__(256) qprep_kernel(const float *__restrict__ q, uint32_t nq, uint32_t nq_pad, uint32_t dim,
                                                    __nv_bfloat16 *__restrict__ q16) {
    const uint32_t row = blockIdx.x;
    // tiled like the corpus shadow: [query tile of 128][k-block][128][64]
    const uint32_t qt = row / BM, rr = row % BM, kblocks = dim / BK;
    for (uint32_t d = threadIdx.x; d < dim; d += blockDim.x)
        q16[(((size_t)qt * kblocks + d / BK) * BM + rr) * BK + d % BK] = __float2bfloat16_rn(row < nq ? q[(size_t)row * dim + d] : 0.f);
    (void)nq_pad;
}

// thr[q] = k'-th smallest tile minimum over the sampled tiles (+inf if fewer than k' finite ones)
// One block: from the per-tile counts pass 1 left in tile_kinds, the row index of the k'-th candidate of either
// exactly-scoring kind (tiles of the prefix are consecutive: tile j = rows [256 j, 256 j + 256)).
__device__ void exact_bound_block(const BatchArgs &a, uint32_t n_tiles, uint32_t kprime) {
    __shared__ uint32_t s_tile[2], s_res[2], s_w[8], s_run;
    const uint32_t need = a.mask ? SPL_ROW_ENUM : SPL_ROW_LIST;
    // thread i sums the counts of tiles [i*per, (i+1)*per); a block-wide scan of the 256 partial sums finds the thread
    // whose range holds the k'-th row of a kind, and that thread walks its own range
    __shared__ uint32_t s_part[2][256];
    const uint32_t per = (n_tiles + blockDim.x - 1) / blockDim.x;
    const uint32_t j0 = threadIdx.x * per, j1 = min(n_tiles, j0 + per);
    uint32_t sum[2] = { 0, 0 };
    for (uint32_t j = j0; j < j1; j++) { sum[0] += a.tile_kinds[2 * j]; sum[1] += a.tile_kinds[2 * j + 1]; }
    s_part[0][threadIdx.x] = sum[0]; s_part[1][threadIdx.x] = sum[1];
    if (threadIdx.x < 2) { s_tile[threadIdx.x] = kIdxSentinel; s_res[threadIdx.x] = 0; }
    __syncthreads();
    for (uint32_t kind = 0; kind < 2; kind++) {
        uint32_t before = 0;
        for (uint32_t i = 0; i < threadIdx.x; i++) before += s_part[kind][i];   // 256 shared-memory reads at most
        if (before < kprime && kprime <= before + sum[kind]) {
            uint32_t cum = before;
            for (uint32_t j = j0; j < j1; j++) {
                const uint32_t c = a.tile_kinds[2 * j + kind];
                if (cum + c >= kprime) { s_tile[kind] = j; s_res[kind] = kprime - cum; break; }
                cum += c;
            }
        }
    }
    __syncthreads();
    for (uint32_t kind = 1; kind <= 2; kind++) {
        const uint32_t tile = s_tile[kind - 1], res = s_res[kind - 1];
        if (tile == kIdxSentinel) { if (threadIdx.x == 0) a.exact_bound_out[kind - 1] = kIdxSentinel; continue; }
        const uint32_t r = tile * BN + threadIdx.x;           // blockDim.x == BN == 256: one row per thread
        const bool is = r < a.n && exact_kind(a, r, need) == kind;
        const unsigned bal = __ballot_sync(0xffffffffu, is);
        if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = __popc(bal);
        __syncthreads();
        uint32_t before = 0;
        for (uint32_t w = 0; w < (threadIdx.x >> 5); w++) before += s_w[w];
        before += __popc(bal & ((1u << (threadIdx.x & 31)) - 1u));
        if (is && before + 1 == res) a.exact_bound_out[kind - 1] = r;
        __syncthreads();
    }
    (void)s_run;
}

__global__ void __launch_bounds__(256) threshold_kernel(const float *__restrict__ tile_min, const uint32_t *__restrict__ tile_min_row,
                                                        uint32_t n_tiles, uint32_t nq, uint32_t nq_pad, uint32_t kprime,
                                                        float *__restrict__ thr, uint32_t *__restrict__ thr_row,
                                                        const BatchArgs ba, uint32_t kprime_exact) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t *keys = reinterpret_cast<uint64_t *>(smem_raw);
    const uint32_t q = blockIdx.x;
    if (q == nq) { exact_bound_block(ba, n_tiles, kprime_exact); return; }
    const uint32_t P = next_pow2(n_tiles < 2 ? 2 : n_tiles);
    uint32_t *idx = reinterpret_cast<uint32_t *>(keys + P);
    for (uint32_t i = threadIdx.x; i < P; i += blockDim.x) {
        const float v = i < n_tiles ? tile_min[(size_t)i * nq_pad + q] : INFINITY;
        keys[i] = f2ord(v); idx[i] = i < n_tiles ? tile_min_row[(size_t)i * nq_pad + q] : kIdxSentinel;
    }
    __syncthreads();
    bitonic_sort_cta(keys, idx, P);
    if (threadIdx.x == 0) {
        // the k'-th smallest (t, row) pair bounds the k' best pairs of the whole shard from above
        float t = INFINITY; uint32_t r = kIdxSentinel;
        if (kprime <= n_tiles) { t = ord2f((uint32_t)keys[kprime - 1]); r = idx[kprime - 1]; }
        if (!(t < INFINITY)) r = kIdxSentinel;
        thr[q] = t; thr_row[q] = r;
    }
    (void)nq;
}

// ---- selection proof: how far a row's true score can sit from what the bf16 tensor-core pass saw ----------------
// Operands are rounded to bf16 with round-to-nearest: every element moves by at most 2^-8 of itself, so the rounded
// vector e16 = e + de has |de| <= 2^-8 |e|.  The mirror keeps |de|^2 per row (MirrorPtrs::err2, usually ~0.4 of the
// worst case) and this kernel computes |dq| of its query, which gives two rigorous bounds that need no assumption
// on how the rounding errors line up:
//   L2      | |q - e| - |q16 - e16| |  <=  |dq| + |de|                          (triangle inequality)
//   cosine  angle(q, e) >= angle(q16, e16) - asin(|dq|/|q|) - asin(|de|/|e16|),  and cos is 1-Lipschitz in the angle
// On top of that the tensor core accumulates dim products in fp32 (relative error <= dim * 2^-24 of |q16||e16|, taken
// as 2^-12) and the fp32 scores of the re-score themselves carry the scan's 1e-5 tolerance.
constexpr uint32_t kRescoreCap = 1024;             // most rows one query may have re-scored
constexpr float kBf16Unit = 0.00390625f;           // 2^-8
constexpr float kTcAccRel = 2.44140625e-4f;        // 2^-12
constexpr float kScanTol = 1e-5f;
__device__ __forceinline__ float asin_upper(float x) {   // >= asin(x) on [0, 0.5]; "no bound" (pi) beyond
    return x <= 0.5f ? x * (1.0f + 0.2f * x * x) * 1.00001f : 3.2f;
}

// per query: sort candidates by t, exact fp32 re-score of the best k', reference order, PROVE that no row left
// out can belong to the top-k (re-scoring the few more that could), emit k
__global__ void __launch_bounds__(256) select_rerank_kernel(const BatchArgs a, const float *__restrict__ q32, uint32_t k,
                                                           uint32_t kprime, spl_gpu_result *__restrict__ out,
                                                           uint32_t out_stride, uint32_t *__restrict__ out_count,
                                                           uint32_t *__restrict__ overflow) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t q = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t total = a.cand_cnt[q];
    if (total > a.cand_cap) {  // threshold too loose for this query: the fp32 scan answers it
        if (threadIdx.x == 0) { overflow[q] = 1; out_count[q] = SPL_GPU_COUNT_UNANSWERED; }
        return;
    }
    // The list is NOT sorted as a whole (a bitonic sort of ~1000 entries per query was the bulk of this kernel): only
    // its m best entries by (t, row) are needed in order, the rest are merely visited by the proof below.  A radix
    // select (8 passes of 8 bits over the 64-bit composite f2ord(t) << 32 | row, which is unique per row) finds the
    // m-th smallest composite; the entries up to it are compacted into the re-score area.
    uint64_t *ckey = reinterpret_cast<uint64_t *>(smem_raw);          // [total]
    for (uint32_t i = threadIdx.x; i < total; i += blockDim.x)
        ckey[i] = ((uint64_t)f2ord(a.cand_t[(size_t)q * a.cand_cap + i]) << 32) | a.cand_idx[(size_t)q * a.cand_cap + i];
    const uint32_t m = total < kprime ? total : kprime;   // candidates re-scored unconditionally
    __shared__ uint64_t rk[kRescoreCap];
    __shared__ uint32_t ri[kRescoreCap];
    __shared__ uint32_t s_extra, s_fail, s_nout, s_hist[256], s_pick, s_fill;
    if (threadIdx.x == 0) { s_extra = 0; s_fail = 0; s_fill = 0; }
    uint64_t pivot = kKeySentinel;                          // m == total: everything is taken
    if (m < total) {
        uint64_t prefix = 0, pmask = 0;
        uint32_t want = m;                                  // the want-th smallest among the keys matching the prefix
        for (int shift = 56; shift >= 0; shift -= 8) {
            s_hist[threadIdx.x] = 0;
            __syncthreads();
            for (uint32_t i = threadIdx.x; i < total; i += blockDim.x)
                if ((ckey[i] & pmask) == prefix) atomicAdd(&s_hist[(uint32_t)(ckey[i] >> shift) & 255u], 1u);
            __syncthreads();
            if (warp == 0) {   // the bin in which the cumulative count reaches `want`: 8 bins per lane
                uint32_t loc[8], sum = 0;
#pragma unroll
                for (int b = 0; b < 8; b++) { loc[b] = s_hist[lane * 8 + b]; sum += loc[b]; }
                uint32_t incl = sum;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= (uint32_t)off) incl += o; }
                const uint32_t excl = incl - sum;
                if (excl < want && want <= incl) {
                    uint32_t cum = excl;
#pragma unroll
                    for (int b = 0; b < 8; b++) {
                        if (cum < want && want <= cum + loc[b]) { s_pick = (lane * 8 + b) | ((want - cum) << 8); }
                        cum += loc[b];
                    }
                }
            }
            __syncthreads();
            prefix |= (uint64_t)(s_pick & 255u) << shift;
            pmask |= 0xFFull << shift;
            want = s_pick >> 8;
            __syncthreads();
        }
        pivot = prefix;
    }
    const uint32_t R = next_pow2(m < 2 ? 2 : m);
    for (uint32_t i = threadIdx.x; i < R; i += blockDim.x) { rk[i] = kKeySentinel; ri[i] = kIdxSentinel; }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < total; i += blockDim.x)
        if (ckey[i] <= pivot) ri[atomicAdd(&s_fill, 1u)] = (uint32_t)ckey[i];   // exactly m of them, in any order
    __syncthreads();
    // Exact fp32 re-score.  Each row is 4*dim bytes somewhere in HBM: the loop is written so that a warp has two
    // whole rows (2 x dim/128 independent 16-byte loads per lane) in flight at a time -- with scalar loads in a
    // rolled loop this step was latency-bound and took longer than everything else in the kernel.
    // Same per-lane arithmetic order for every row, so equal rows get equal scores.
    const float4 *qv4 = reinterpret_cast<const float4 *>(q32 + (size_t)q * a.dim);
    const uint32_t nv = a.dim >> 2;  // float4s per row; dim is a multiple of 64 on this path
    float ma = 0.f, mq16 = 0.f, dq2 = 0.f;   // |q|^2, |bf16(q)|^2, |q - bf16(q)|^2
    for (uint32_t d = lane; d < nv; d += 32) {
        const float4 y = qv4[d];
        ma = fmaf(y.x, y.x, ma); ma = fmaf(y.y, y.y, ma); ma = fmaf(y.z, y.z, ma); ma = fmaf(y.w, y.w, ma);
        const float yy[4] = { y.x, y.y, y.z, y.w };
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float h = __bfloat162float(__float2bfloat16_rn(yy[j])), e = yy[j] - h;
            mq16 = fmaf(h, h, mq16); dq2 = fmaf(e, e, dq2);
        }
    }
    ma = warp_sum(ma); mq16 = warp_sum(mq16); dq2 = warp_sum(dq2);
    auto score_row = [&](const float4 *e4, float &dot, float &mb, float &ss) {
        dot = 0.f; mb = 0.f; ss = 0.f;
#pragma unroll 8
        for (uint32_t d = lane; d < nv; d += 32) {
            const float4 x = __ldg(e4 + d), y = qv4[d];
            float t;
            dot = fmaf(y.x, x.x, dot); mb = fmaf(x.x, x.x, mb); t = y.x - x.x; ss = fmaf(t, t, ss);
            dot = fmaf(y.y, x.y, dot); mb = fmaf(x.y, x.y, mb); t = y.y - x.y; ss = fmaf(t, t, ss);
            dot = fmaf(y.z, x.z, dot); mb = fmaf(x.z, x.z, mb); t = y.z - x.z; ss = fmaf(t, t, ss);
            dot = fmaf(y.w, x.w, dot); mb = fmaf(x.w, x.w, mb); t = y.w - x.w; ss = fmaf(t, t, ss);
        }
    };
    auto emit = [&](uint32_t c, uint32_t row, float dot, float mb, float ss) {
        dot = warp_sum(dot); mb = warp_sum(mb); ss = warp_sum(ss);
        float sim = cosine_from_sums(dot, ma, mb), dist = __fsqrt_rn(ss);
        if (!(a.flags[row] & SPL_ROW_EMB)) { sim = 0.f; dist = FLT_MAX; }
        if (lane == 0) rk[c] = make_key(a.order, sim, dist);
    };
    const uint32_t nwarps = blockDim.x / 32;
    auto rescore = [&](uint32_t lo, uint32_t hi) {   // rows ri[lo..hi) -> rk[lo..hi)
        for (uint32_t c = lo + warp; c < hi; c += 2 * nwarps) {
            const uint32_t c2 = c + nwarps;
            const bool two = c2 < hi;
            const uint32_t row0 = ri[c], row1 = two ? ri[c2] : row0;
            float d0, b0, s0, d1, b1, s1;
            score_row(reinterpret_cast<const float4 *>(a.E + (size_t)row0 * a.dim), d0, b0, s0);
            score_row(reinterpret_cast<const float4 *>(a.E + (size_t)row1 * a.dim), d1, b1, s1);
            emit(c, row0, d0, b0, s0);
            if (two) emit(c2, row1, d1, b1, s1);
        }
    };
    rescore(0, m);
    __syncthreads();
    bitonic_sort_cta(rk, ri, R);

    // ---- selection proof ----------------------------------------------------------------------------------------
    // X = the k-th best exact score so far (or the threshold, whichever is harder to beat).  A row that was not
    // re-scored can only matter if its true score could reach X; its bf16 ordering value bounds its true score
    // (bounds above).  Every such row still in the candidate list is re-scored now -- X can only get better by
    // that, so one round is enough -- and rows that never made the list are covered through the list's bound thr.
    const bool cosine = a.order == SPL_ORDER_COSINE;
    const bool have_k = m >= k;
    float simk = 0.f, distk = 0.f;
    if (have_k) split_key(a.order, rk[k - 1], simk, distk);
    const float qn = sqrtf(ma), qn16 = sqrtf(mq16), dqn = sqrtf(dq2) * 1.000001f;
    float S = have_k ? simk : -INFINITY;          // cosine: a row matters only if its similarity can reach S
    if (a.min_sim > 0.f) S = fmaxf(S, a.min_sim);
    float Dk = have_k ? distk : INFINITY;         // L2: ... only if its distance can get down to Dk
    if (a.max_dist > 0.f) Dk = fminf(Dk, a.max_dist);
    const float DkUp = Dk * (1.0f + kScanTol);
    const float thq = qn > 0.f ? asin_upper(dqn / qn) : 3.2f;
    // Rows whose score is known without reading them: a zero vector scores (0, |q|) -- with exactly the bits the
    // re-score produces, ma above is the same sum in the same order -- and a row without a consistent embedding copy
    // (0, FLT_MAX), _search.c:388-391.  They compare EXACTLY, ties by row index like everywhere else; that matters:
    // the synthetic corpus holds thousands of never-embedded zero rows that tie at distance |q|.
    const uint64_t key_zero = make_key(a.order, 0.f, __fsqrt_rn(ma)), key_noemb = make_key(a.order, 0.f, FLT_MAX);
    const uint64_t xkey = have_k ? rk[k - 1] : kKeySentinel;
    const uint32_t xidx = have_k ? ri[k - 1] : kIdxSentinel;
    const bool thresholds = a.min_sim > 0.f || a.max_dist > 0.f;
    auto exact_excluded = [&](uint64_t key, uint32_t row, bool emb) {
        if (thresholds) {   // rows failing the threshold are never results
            float sim, dist;
            split_key(a.order, key, sim, dist);
            if (!emb || (a.min_sim > 0.f && sim < a.min_sim) || (a.max_dist > 0.f && dist > a.max_dist)) return true;
        }
        return have_k && !better(key, row, xkey, xidx);
    };
    for (uint32_t c = threadIdx.x; c < total; c += blockDim.x) {
        if (ckey[c] <= pivot) continue;                     // re-scored already
        const uint32_t row = (uint32_t)ckey[c];
        const float t = ord2f((uint32_t)(ckey[c] >> 32));
        bool excluded;
        const bool emb = (a.flags[row] & SPL_ROW_EMB) != 0;
        const float n2 = emb ? a.norm2[row] : 0.f, er2 = emb ? a.err2[row] : 0.f;
        if (!emb) {
            excluded = exact_excluded(key_noemb, row, false);
        } else if (n2 == 0.f && er2 == 0.f) {
            excluded = exact_excluded(key_zero, row, true);
        } else {
            const float en16 = sqrtf(n2), de = sqrtf(er2) * 1.000001f;
            if (cosine) {
                float sim_up;
                if (!(qn16 > 0.f) || en16 <= 2.f * de) sim_up = INFINITY;  // no usable bound
                else sim_up = -t / qn16 + thq + asin_upper(de / en16) + kTcAccRel + kScanTol;
                excluded = sim_up < S;
            } else {
                const float C = t + mq16;                                      // |q16 - e16|^2 as the pass computed it
                const float delta = 2.f * kTcAccRel * qn16 * en16 + 1e-5f * (n2 + mq16);
                const float lo = sqrtf(fmaxf(0.f, C - delta)) * (1.0f - 1e-6f) - (dqn + de) * (1.0f + kScanTol);
                excluded = lo > DkUp;
            }
        }
        if (!excluded) {
            const uint32_t pos = atomicAdd(&s_extra, 1u);
            if (m + pos < kRescoreCap) ri[m + pos] = row;   // (slots >= R are written here for the first time)
        }
    }
    if (threadIdx.x == 0) {
        // Rows that never entered the list.  Pass 2 left behind, per query, the best ordering value among the rows it
        // kept out because they were WORSE than the bound (out_min), and whether rows TYING with the bound were cut off
        // by their row index (out_tie: bit 0 = rows that score exactly, bit 1 = ordinary rows).
        auto bound_clears = [&](float t16) {   // can a row whose bf16 ordering value is >= t16 still reach X?
            if (cosine) {
                const float sim_up = qn16 > 0.f ? -t16 / qn16 + thq + asin_upper(kBf16Unit) + kTcAccRel + kScanTol : INFINITY;
                return sim_up < S;
            }
            // |e16| of an unknown row: at most the shard's maximum, and a row with |e| > Dk + |q| is farther than Dk anyway
            float emax = sqrtf(__uint_as_float(*a.norm2_max));
            if (Dk < INFINITY) emax = fminf(emax, 1.004f * (Dk + qn));
            const float de = kBf16Unit * 1.004f * emax;
            const float delta = 2.f * kTcAccRel * qn16 * emax + 1e-5f * (emax * emax + mq16);
            const float lo = sqrtf(fmaxf(0.f, t16 + mq16 - delta)) * (1.0f - 1e-6f) - (dqn + de) * (1.0f + kScanTol);
            return lo > DkUp;
        };
        bool ok = true;
        const uint32_t omin_inv = a.out_min[q], tie = a.out_tie[q];
        if (omin_inv != 0u) ok = bound_clears(ord2f(~omin_inv));
        if (ok && (tie & 2u)) ok = bound_clears(a.thr[q]);
        if (ok && (tie & 1u)) {
            // exactly-scoring rows tying with the bound, all with a row index beyond thr_row: the best they can be is a
            // zero row (or, at the far end of the L2 order, a row without embedding) at index thr_row + 1
            const bool noemb = !cosine && a.thr[q] > 1e38f;
            ok = exact_excluded(noemb ? key_noemb : key_zero, a.thr_row[q] + 1u, !noemb);
        }
        if (!ok) s_fail = 1;
    }
    __syncthreads();
    const uint32_t extra = s_extra;
    if (s_fail || m + extra > kRescoreCap) {
        // too many near-ties inside the bf16 error bound (or a bound that does not clear): the fp32 scan answers it
        if (threadIdx.x == 0) { overflow[q] = 1; out_count[q] = SPL_GPU_COUNT_UNANSWERED; }
        return;
    }
    const uint32_t m2 = m + extra;
    if (extra) {
        const uint32_t R2 = next_pow2(m2);
        for (uint32_t i = m2 + threadIdx.x; i < R2; i += blockDim.x) { rk[i] = kKeySentinel; ri[i] = kIdxSentinel; }
        for (uint32_t i = R + threadIdx.x; i < m2; i += blockDim.x) rk[i] = kKeySentinel;   // filled by the re-score below
        __syncthreads();
        rescore(m, m2);
        __syncthreads();
        bitonic_sort_cta(rk, ri, R2);
    }
    // threshold on the ordering score (_search.c:398-399): the rows that fail are the tail of the sorted list
    if (threadIdx.x == 0) { s_nout = m2 < k ? m2 : k; overflow[q] = extra << 1; }
    __syncthreads();
    if (a.min_sim > 0.f || a.max_dist > 0.f) {
        const uint32_t lim = s_nout;
        __syncthreads();
        for (uint32_t t = threadIdx.x; t < lim; t += blockDim.x) {
            float sim, dist;
            split_key(a.order, rk[t], sim, dist);
            const bool emb = (a.flags[ri[t]] & SPL_ROW_EMB) != 0;   // no embedding + any threshold: dropped (_search.c:392-393)
            if (!emb || (a.min_sim > 0.f && sim < a.min_sim) || (a.max_dist > 0.f && dist > a.max_dist)) atomicMin(&s_nout, t);
        }
        __syncthreads();
    }
    const uint32_t nout = s_nout;
    for (uint32_t t = threadIdx.x; t < nout; t += blockDim.x) {
        spl_gpu_result r;
        split_key(a.order, rk[t], r.similarity, r.distance);
        const uint32_t row = ri[t];
        r.slot_idx = (uint32_t)(a.row0 + row);
        r.has_embedding = (a.flags[row] & SPL_ROW_EMB) ? 1u : 0u;
        r.epoch = a.epoch[row];
        r.bloom = a.bloom[row];
        out[(size_t)q * out_stride + t] = r;
    }
    if (threadIdx.x == 0) out_count[q] = nout;
}

// -------------------------------------------------------------------------------------------------
// host side
// -------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// tiled bf16 operand seen as a 2-D [rows * k-blocks][64] tensor (row pitch 128 B, fully contiguous),
// box = {64 elements (128 B), box_rows}, 128-byte swizzle
static bool make_map(CUtensorMap *map, const void *base, uint64_t rows, uint32_t dim, uint32_t box_rows) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return false;
    cuuint64_t gdim[2] = { (cuuint64_t)BK, rows * (dim / BK) };
    cuuint64_t gstride[1] = { (cuuint64_t)BK * 2 };
    cuuint32_t box[2] = { (cuuint32_t)BK, box_rows };
    cuuint32_t estr[2] = { 1, 1 };
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), gdim, gstride, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

cudaError_t batch_prepare() {
    cudaError_t e = cudaFuncSetAttribute(gemm_topk_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GemmSmem::total);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(gemm_topk_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmemBudget);
    if (e != cudaSuccess) return e;
    // function attributes are per device: set here (once per context, its device current), not lazily behind a flag
    e = cudaFuncSetAttribute(select_rerank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kBatchCandCap * 12));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(threshold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kMaxSampledTiles * 12));
}

static cudaError_t launch_gemm(const CUtensorMap &mapQ, const CUtensorMap &mapE, const BatchArgs &a, bool pair, int grid,
                               cudaStream_t s) {
    if (!pair) {
        gemm_topk_kernel<false><<<grid, kThreadsGemm, GemmSmem::total, s>>>(mapQ, mapE, a);
        return cudaGetLastError();
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreadsGemm);
    cfg.dynamicSmemBytes = pair_layout(a.pair_q_units).total;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, gemm_topk_kernel<true>, mapQ, mapE, a);
}

bool batch_supported(uint32_t dim, uint32_t k) { return dim % BK == 0 && dim >= BK && k >= 1 && k <= 128; }

size_t batch_scratch_bytes(uint32_t n_rows, uint32_t nq_max, uint32_t dim, BatchScratchLayout *L) {
    const uint32_t nq_pad = (nq_max + BM - 1) / BM * BM;
    const uint32_t tiles = (n_rows + BN - 1) / BN;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    L->q16_off = take((size_t)nq_pad * dim * 2);
    L->thr_off = take((size_t)nq_pad * 4);
    L->thr_row_off = take((size_t)nq_pad * 4);
    L->cnt_off = take((size_t)nq_pad * 4);
    L->ovf_off = take((size_t)nq_pad * 4);
    L->out_min_off = take((size_t)nq_pad * 4);
    L->out_tie_off = take((size_t)nq_pad * 4);
    L->exact_off = take((size_t)(4 + 2 * kMaxSampledTiles) * 4);   // bound[2], pad[2], per-tile counts [sampled tiles][2]
    L->tile_min_off = take((size_t)tiles * nq_pad * 4);
    L->tile_min_row_off = take((size_t)tiles * nq_pad * 4);
    L->cand_t_off = take((size_t)nq_pad * kBatchCandCap * 4);
    L->cand_idx_off = take((size_t)nq_pad * kBatchCandCap * 4);
    L->nq_pad = nq_pad;
    return off;
}

cudaError_t batch_search(const BatchArgs &base, const BatchScratchLayout &L, unsigned char *scratch, const float *d_q32,
                         uint32_t k, int sm_count, spl_gpu_result *d_out, uint32_t out_stride, uint32_t *d_counts,
                         uint32_t **d_overflow_out, uint32_t *launches, cudaStream_t s) {
    BatchArgs a = base;
    a.nq_pad = (a.nq + BM - 1) / BM * BM;
    a.n_qtiles = a.nq_pad / BM;
    __nv_bfloat16 *q16 = reinterpret_cast<__nv_bfloat16 *>(scratch + L.q16_off);
    a.thr = reinterpret_cast<float *>(scratch + L.thr_off);
    a.thr_row = reinterpret_cast<uint32_t *>(scratch + L.thr_row_off);
    a.tile_min_row = reinterpret_cast<uint32_t *>(scratch + L.tile_min_row_off);
    a.cand_cnt = reinterpret_cast<uint32_t *>(scratch + L.cnt_off);
    uint32_t *overflow = reinterpret_cast<uint32_t *>(scratch + L.ovf_off);
    a.out_min = reinterpret_cast<uint32_t *>(scratch + L.out_min_off);
    a.out_tie = reinterpret_cast<uint32_t *>(scratch + L.out_tie_off);
    a.tile_min = reinterpret_cast<float *>(scratch + L.tile_min_off);
    a.cand_t = reinterpret_cast<float *>(scratch + L.cand_t_off);
    a.cand_idx = reinterpret_cast<uint32_t *>(scratch + L.cand_idx_off);
    a.cand_cap = kBatchCandCap;
    *d_overflow_out = overflow;
    const uint32_t kprime = 2 * k;

    CUtensorMap mapQ, mapE;
    const uint64_t n_pad = ((uint64_t)a.n + BN - 1) / BN * BN;
    // 256 queries and up run on CTA pairs (cta_group::2); the row operand is then fetched in two 128-row halves
    const bool pair = a.n_qtiles % 2 == 0 && getenv("SPL_GPU_BATCH_1CTA") == nullptr;
    {   // developer knob: depth of the query ring of the pair kernel (the row ring takes the rest)
        const char *qu = getenv("SPL_GPU_PAIR_QUNITS");
        a.pair_q_units = qu ? (uint32_t)atoi(qu) : 0u;
    }
    if (!make_map(&mapQ, q16, a.nq_pad, a.dim, BM) || !make_map(&mapE, a.E16, n_pad, a.dim, pair ? BN / 2 : BN))
        return cudaErrorNotSupported;

    cudaError_t e;
    qprep_kernel<<<a.nq_pad, 256, 0, s>>>(d_q32, a.nq, a.nq_pad, a.dim, q16);
    if ((e = cudaGetLastError())) return e;

    {   // per-kind bounds for the rows that score exactly (exact_bound_block)
        uint32_t *ex = reinterpret_cast<uint32_t *>(scratch + L.exact_off);
        a.exact_bound = ex; a.exact_bound_out = ex; a.tile_kinds = ex + 4;
    }
    const uint32_t tiles = (a.n + BN - 1) / BN;
    int grid = sm_count / (int)a.n_qtiles * (int)a.n_qtiles;
    if (grid < (int)a.n_qtiles) grid = (int)a.n_qtiles;
    if (pair) grid &= ~1;
    // pass 1: a PREFIX of the tiles (1/64 of them, at least 64, all when the shard is small) -> per-tile minima ->
    // thresholds.  A prefix (not a strided sample) keeps the estimate below true for ties on t as well, because the
    // bound is the pair (t, row).  The sample only has to place the bound roughly: whatever it yields, the selection
    // proof in select_rerank_kernel decides whether the list it produced can be trusted.
    uint32_t sampled = tiles / 64;
    if (sampled < 64) sampled = 64;
    if (sampled > tiles) sampled = tiles;
    if (sampled > kMaxSampledTiles) sampled = kMaxSampledTiles;   // threshold_kernel sorts the sample in shared memory
    a.pass = 1; a.tile_first = 0; a.tile_stride = 1; a.n_tiles = sampled;
    // one fill for everything that must start at zero: candidate counts, overflow flags, out_min (stored inverted), tie
    // flags, the exact-kind bounds and per-tile counts -- laid out back to back by batch_scratch_bytes
    if ((e = cudaMemsetAsync(scratch + L.cnt_off, 0, L.exact_off + (size_t)(4 + 2 * sampled) * 4 - L.cnt_off, s))) return e;
    if ((e = launch_gemm(mapQ, mapE, a, pair, grid, s))) return e;
    const uint32_t P = next_pow2(sampled < 2 ? 2 : sampled);
    // Which order statistic of the sampled tile minima becomes the bound: the list should hold about 16 k' entries, so
    // that a crowd of near-ties around the k-th score is IN the list (the selection proof can then re-score it) and
    // the rows left out are far enough from the k-th for the proof's bound to clear them.  With a 1/f sample the
    // k'-th smallest minimum already sits near rank f k' of the shard; a small shard that is sampled whole needs a
    // later one (but no later than half of its tiles: one candidate per tile is all the minima can vouch for).  The order
    // statistic may be smaller than k: then the minima alone do not guarantee k rows under the bound, but in expectation
    // ~16 k' rows are, and a list that did come out short fails the proof and goes to the fp32 scan.
    uint32_t thr_rank = (uint32_t)(((uint64_t)16 * kprime * sampled + tiles - 1) / tiles);
    const uint32_t rank_hi = tiles <= 64 ? sampled : sampled / 2;
    if (thr_rank > rank_hi) thr_rank = rank_hi;
    if (thr_rank < 8) thr_rank = 8;
    if (a.n <= kBatchCandCap) thr_rank = sampled + 1;   // the whole shard fits the list: no bound (+inf), every candidate is listed
    threshold_kernel<<<a.nq + 1, 256, (size_t)P * 12, s>>>(a.tile_min, a.tile_min_row, sampled, a.nq, a.nq_pad, thr_rank, a.thr,
                                                           a.thr_row, a, kprime);
    if ((e = cudaGetLastError())) return e;
    // pass 2: every tile, append what beats the threshold
    a.pass = 2; a.tile_first = 0; a.tile_stride = 1; a.n_tiles = tiles;
    if ((e = launch_gemm(mapQ, mapE, a, pair, grid, s))) return e;
    const size_t sel_smem = (size_t)kBatchCandCap * 12;   // opted in by batch_prepare()
    select_rerank_kernel<<<a.nq, 256, sel_smem, s>>>(a, d_q32, k, kprime, d_out, out_stride, d_counts, overflow);
    if ((e = cudaGetLastError())) return e;
    *launches = 5;
    return cudaSuccess;
}

}  // namespace spl
