// spl_scan_tma.cu -- fused scan, TMA-staged (the default SPL_MODE_FAST kernel): one query, or 2 / 4 / 8 queries per pass.
//
// Same contract as spl_scan.cu (predicate -> cosine + L2 -> thresholds -> top-k; reference:
// /root/reference/splinter_cli_cmd_search.c:336-416) but the embedding rows travel HBM -> shared
// memory through the bulk-copy engine (cp.async.bulk, SASS UBLKCP) instead of through registers:
//
//   4 producer warps  evaluate the label / validity predicate for a chunk of up to 256 slots from 9 bytes
//                     per slot (bloom u64 + flags u8), compact the slots that pass and issue bulk copies of
//                     their rows into a shared-memory ring of PAIR slots (2 rows each; two passing rows that
//                     are neighbours in HBM travel as one 8*dim-byte copy).  Rows that fail the predicate
//                     never leave HBM.
//   16 consumer warps each take every 16th pair slot: wait on the slot's mbarrier, pull both rows into
//                     registers with conflict-free 128-bit LDS, hand the slot back, then score the two rows
//                     (dot, |e|^2, sum (q-e)^2 with packed FFMA2, warp-shuffle tree) and feed the per-warp
//                     top-k list.  NQ = 1: the query sits in registers (in shared memory for 1024-d rows).
//                     NQ = 2 / 4 / 8: the queries sit in shared memory, every row is scored against each of them,
//                     the partial sums are reduced transposed and each (query, row) is finished by its own lane.
//
// One persistent CTA per SM keeps up to ~190 KB of row data in flight per SM no matter what the
// compute warps are doing -- what an HBM-bound gather of 3 KB rows needs.
#include <float.h>
#include <string.h>

#include <type_traits>

#include "spl_scan.cuh"
#include "spl_scan_epilogue.cuh"

namespace spl {

constexpr int kConsumers = 16;
constexpr int kProducers = 4;
constexpr int kTmaThreads = (kConsumers + kProducers) * 32;  // consumers = warps 0..15, producers = warps 16..19
constexpr uint32_t kSmemBudget = 227u * 1024u;
constexpr uint32_t kMaxPairs = 32;
constexpr uint32_t kMinPairs = 17;  // a 32-slot group may straddle 17 pair slots; all must be distinct

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
// Slot hand-back uses an exact generation count instead of an mbarrier parity: with several
// producer warps one of them may run a whole ring ahead of another, which a 1-bit phase cannot
// tell apart from "the previous round".
__device__ __forceinline__ void wait_gen(uint32_t addr, uint32_t gen) {
    uint32_t v;
    do {
        asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    } while (v != gen);
}
// producer flavour: when the consumers are the slower side (small-batch scans) 128 producer lanes spinning on
// shared memory would steal load/store bandwidth and issue slots from them, so back off between polls
__device__ __forceinline__ void wait_gen_backoff(uint32_t addr, uint32_t gen) {
    uint32_t v;
    int spins = 0;
    for (;;) {
        asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
        if (v == gen) return;
        // The ring holds ~4x the bytes the HBM latency-bandwidth product needs, so a refill that comes a
        // fraction of a microsecond late costs nothing, while a tight poll costs the consumers issue slots.
        if (++spins > 2) __nanosleep(spins > 8 ? 500 : 200);
    }
}
__device__ __forceinline__ void publish_gen(uint32_t addr, uint32_t gen) {
    asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" ::"r"(addr), "r"(gen) : "memory");
}
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// L2 residency: a query streams hundreds of MB of rows through a 126 MB L2 once and never looks at them again, while the
// 9 bytes per slot of predicate data (bloom + flags, 9.4 MB at 2^20 slots) are read again by every query.  The rows are
// therefore fetched evict-first and the predicate bytes evict-last: the predicate columns stay in L2 from one query to
// the next instead of being pushed out by rows on their way to nowhere.
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t ldg_u64_hint(const uint64_t *ptr, uint64_t policy) {
    uint64_t v;
    asm volatile("ld.global.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(ptr), "l"(policy));
    return v;
}
__device__ __forceinline__ uint32_t ldg_u8_hint(const uint8_t *ptr, uint64_t policy) {
    uint32_t v;
    asm volatile("ld.global.L2::cache_hint.u8 %0, [%1], %2;" : "=r"(v) : "l"(ptr), "l"(policy));
    return v;
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy) : "memory");
}

// Packed fp32 pairs (FFMA2, sm_100): one instruction does the (x, y) or (z, w) half of a float4.  Every sum below
// is kept as two partial sums (x/z components, y/w components) that are added once at the end -- in the single-
// and the multi-query loop alike, so both produce the same bits.
__device__ __forceinline__ float2 lo2(const float4 &v) { return make_float2(v.x, v.y); }
__device__ __forceinline__ float2 hi2(const float4 &v) { return make_float2(v.z, v.w); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 sub2(float2 a, float2 b) { return __ffma2_rn(b, make_float2(-1.0f, -1.0f), a); }  // a - b, one rounding
// one float4 of query q against one float4 of row e: dot += q*e, ss += (q-e)^2
__device__ __forceinline__ void score4(const float4 &q, const float4 &e, float2 &dot, float2 &ss) {
    dot = fma2(lo2(q), lo2(e), dot);
    float2 t = sub2(lo2(q), lo2(e));
    ss = fma2(t, t, ss);
    dot = fma2(hi2(q), hi2(e), dot);
    t = sub2(hi2(q), hi2(e));
    ss = fma2(t, t, ss);
}
__device__ __forceinline__ void norm4(const float4 &e, float2 &n) {
    n = fma2(lo2(e), lo2(e), n);
    n = fma2(hi2(e), hi2(e), n);
}

struct TmaLayout {
    uint32_t pairs, row_bytes, rows_off, full_off, empty_off, meta_off, clist_off, state_off, q_off, list_entries, total;
};

// Which consumer loop a launch uses.  Small batches keep the queries in shared memory by construction; so does a
// single 1024-d query: with the query in registers as well (32 + 64 row registers of the 96 available) that
// instantiation spilled and ran a dense scan at 79 % of the bandwidth the shared-memory variant reaches.
__host__ __device__ constexpr bool tma_query_in_smem(uint32_t dim, uint32_t nq) { return nq > 1 || dim >= 1024; }

// nq > 1 (small-batch scan): nq * kConsumers lists of k entries, their bookkeeping words and the nq
// queries themselves live in shared memory as well; the ring gets what is left.
__host__ __device__ inline TmaLayout tma_layout(uint32_t dim, uint32_t k, uint32_t nq = 1, bool stage_q = false) {
    TmaLayout L;
    L.row_bytes = dim * 4;
    L.list_entries = nq * scan_topk_entries(k, kConsumers);   // per query: kConsumers lists of k entries, padded to a power of two
    const uint32_t topk = L.list_entries * 12u;                    // u64 keys[] then u32 idx[]
    L.full_off = (topk + 15u) & ~15u;                             // top-k lists sit at offset 0
    L.empty_off = L.full_off + kMaxPairs * 8;
    L.meta_off = L.empty_off + kMaxPairs * 8;                     // uint4 {row0, flags0, row1, flags1}
    L.clist_off = L.meta_off + kMaxPairs * 16;                    // per producer warp: 256 x {row, flags}
    L.state_off = L.clist_off + kProducers * 256 * 8;             // nq > 1: uint4 per (query, consumer warp)
    const bool qs = tma_query_in_smem(dim, nq) || stage_q;   // stage_q: a query that arrived in the launch's parameter space
    L.q_off = L.state_off + (qs ? nq * kConsumers * 20u + ((nq * 4u + 15u) & ~15u) : 0u);  // + u32 survivors per list, |q|^2 per query
                                                                  // then the queries, fp32 [nq][dim]
    L.rows_off = (L.q_off + (qs ? nq * L.row_bytes : 0u) + 127u) & ~127u;
    const uint32_t avail = kSmemBudget - 8192u - L.rows_off;      // 8 KB headroom for static smem
    const uint32_t p = avail / (2 * L.row_bytes);
    L.pairs = p > kMaxPairs ? kMaxPairs : p;
    const uint32_t ring = L.pairs * 2 * L.row_bytes, scratch = epilogue_scratch_bytes<kConsumers>();
    L.total = L.rows_off + (ring > scratch ? ring : scratch);     // the idle ring doubles as the epilogue's scratch
    return L;
}

// The host-buffer call's single query can travel INSIDE the launch (kernel parameter space, 4*dim bytes) instead of
// through a host-to-device copy of its own: one operation less on the stream between the caller and its answer.
template <int V>
struct QueryArg { float4 v[V * 32]; };
struct NoQueryArg { float4 v[1]; };

template <int V, int NQ, bool QP>
__device__ __forceinline__ void scan_tma_body(const ScanArgs &a, const float4 *qparam);

template <int V, int NQ>
__global__ void __launch_bounds__(kTmaThreads, 1) scan_tma_kernel(const ScanArgs a) {
    scan_tma_body<V, NQ, false>(a, nullptr);
}

// (only instantiated for NQ = 1 and dimensions whose query fits the classic 4 KB parameter space next to ScanArgs)
template <int V>
__global__ void __launch_bounds__(kTmaThreads, 1) scan_tma_qparam_kernel(const ScanArgs a, const __grid_constant__ QueryArg<V> q) {
    scan_tma_body<V, 1, true>(a, q.v);
}

template <int V, int NQ, bool QP>
__device__ __forceinline__ void scan_tma_body(const ScanArgs &a, const float4 *qparam) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint32_t s_flag;
    __shared__ uint32_t s_seq;  // rows issued so far by all producer warps
    // consumers with the query in shared memory: the k smallest primary key words (the key's high half: similarity in cosine order, distance
    // in L2 order) any warp of this CTA has offered to its list so far, per query, and the largest of them (see cta_note)
    constexpr uint32_t kCtaHi = !tma_query_in_smem(V * 128, NQ) ? 1u : (NQ == 8 ? 8u * 32u : (uint32_t)NQ * 256u);   // NQ = 8: k <= 32
    __shared__ uint32_t s_cta_hi[kCtaHi];
    __shared__ uint32_t s_cta_kth[8];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const TmaLayout L = tma_layout(a.dim, a.k, NQ, QP);
    const uint32_t list_entries = L.list_entries;
    const uint32_t qstride = list_entries / NQ;   // list entries per query
    uint64_t *skeys = reinterpret_cast<uint64_t *>(smem);
    uint32_t *sidx = reinterpret_cast<uint32_t *>(skeys + list_entries);
    const uint32_t full0 = smem_u32(smem + L.full_off), empty0 = smem_u32(smem + L.empty_off);
    uint4 *meta = reinterpret_cast<uint4 *>(smem + L.meta_off);
    const uint32_t NP = L.pairs, RB = L.row_bytes;

    if (threadIdx.x == 0) {
        if (a.dbg) a.dbg[blockIdx.x * 8 + 0] = gtime();
        s_seq = 0;
        for (uint32_t s = 0; s < NP; s++) { mbar_init(full0 + 8 * s, 2); publish_gen(empty0 + 8 * s, 0); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    topk_lists_init(skeys, sidx, list_entries);
    constexpr bool QS = tma_query_in_smem(V * 128, NQ);
    if constexpr (QS) {
        uint4 *wstate = reinterpret_cast<uint4 *>(smem + L.state_off);
        for (uint32_t i = threadIdx.x; i < NQ * kConsumers; i += blockDim.x) {
            wstate[i] = make_uint4((uint32_t)kKeySentinel, (uint32_t)(kKeySentinel >> 32), kIdxSentinel, 0u);
            reinterpret_cast<uint32_t *>(wstate + NQ * kConsumers)[i] = 0u;
        }
        for (uint32_t i = threadIdx.x; i < kCtaHi; i += blockDim.x) s_cta_hi[i] = 0xffffffffu;
        if (threadIdx.x < 8) s_cta_kth[threadIdx.x] = 0xffffffffu;
        float4 *sq4 = reinterpret_cast<float4 *>(smem + L.q_off);
        const float4 *gq4 = reinterpret_cast<const float4 *>(a.q);
        for (uint32_t i = threadIdx.x; i < NQ * (a.dim >> 2); i += blockDim.x) sq4[i] = gq4[i];
    }
    if constexpr (QP) {
        // Parameter space is read through the constant cache, which serves ONE address per request: a lane-indexed read
        // (32 addresses per instruction, 16 warps x 6 instructions) cost ~100 us per launch.  Read it warp-uniformly
        // instead -- every warp takes a slice, one address per instruction -- and park the query in shared memory.
        float4 *sq4 = reinterpret_cast<float4 *>(smem + L.q_off);
        for (uint32_t i = warp; i < (uint32_t)V * 32u; i += kConsumers + kProducers) {
            const float4 v = qparam[i];
            if (lane == 0) sq4[i] = v;
        }
    }
    __syncthreads();

    WarpTopK W;
    W.init(skeys + (warp < kConsumers ? warp : 0) * a.k, sidx + (warp < kConsumers ? warp : 0) * a.k, a.k);

    if (warp >= kConsumers) {
        // ================================ producers ================================
        // kProducers warps run the same loop on different chunks; ring positions are handed out
        // by one shared-memory atomic per 32-slot group, so the rows of all producers interleave
        // into one sequence that the consumers drain in order.
        const uint32_t need = a.mask ? SPL_ROW_ENUM : SPL_ROW_LIST;
        const uint32_t ngroups = (a.n + 31u) >> 5;
        const unsigned char *Eb = reinterpret_cast<const unsigned char *>(a.E);
        const uint32_t rows_s = smem_u32(smem + L.rows_off);
        const uint64_t pol_keep = l2_policy_evict_last(), pol_stream = l2_policy_evict_first();
        // Work is handed out in chunks of kChunk groups (kChunk*32 slots) by one atomic per chunk
        // (a single counter cannot sustain one atomic per 32 slots at HBM speed).  The NEXT chunk's
        // base and its predicate bytes are requested before the current chunk's rows are issued, so
        // neither the atomic's round trip nor the 9-byte-per-slot reads sit on the critical path.
        // The loop body is written out twice (register sets A / B) so the prefetched registers are
        // never copied -- a copy would wait for the loads it is meant to hide.
        constexpr int kChunk = 8;
        const uint32_t cg = a.chunk_groups;  // largest chunk, in 32-slot groups (small stores use small chunks)
        // A chunk is (first group, length 1..8), packed as first << 4 | length.  Every producer warp's FIRST chunk is
        // fixed (CTA * kProducers + producer, cg groups), so the scan starts without waiting for an atomic's round
        // trip; the shared counter -- it counts GROUPS -- hands out the rest, and each warp asks for a length that
        // suits the density it is seeing: where most slots pass (every chunk is hundreds of KB of rows) short chunks
        // even out the CTAs' finishing times, where few pass long chunks keep the atomics off the critical path.
        // Measured at 2^20 x 768 with fixed lengths 8 / 4 / 2: all-match 476 / 463 / 457 us, mask 0x3 82.4 / 80.6 /
        // 83.1 us, mask 0x7 (1.6 %) 28.9 / 35.7 / 49.1 us.
        const uint32_t first_dynamic = gridDim.x * kProducers * cg;
        uint32_t want = cg;   // length of the next chunk to ask for (warp-uniform)
        auto fetch_chunk = [&]() -> uint32_t {
            uint32_t c = 0;
            if (lane == 0) c = ((atomicAdd(&a.counters[0], want) + first_dynamic) << 4) | want;
            return c;  // lane 0 only; broadcast lazily so the atomic's latency is not exposed here
        };
        auto chunk_live = [&](uint32_t c) -> bool { return (c >> 4) < ngroups; };
        auto load_chunk_pred = [&](uint32_t c, uint32_t (&f)[kChunk], uint64_t (&b)[kChunk]) {
            const uint32_t g0 = c >> 4, len = c & 15u;
#pragma unroll
            for (int j = 0; j < kChunk; j++) {
                const uint32_t row = ((g0 + j) << 5) + lane;   // g0 < 2^27, so this cannot wrap
                f[j] = 0; b[j] = 0;
                if (j < len && g0 + j < ngroups && row < a.n) { f[j] = ldg_u8_hint(a.flags + row, pol_keep); b[j] = ldg_u64_hint(a.bloom + row, pol_keep); }
            }
        };
        // Rows that pass are first compacted into a per-warp list in shared memory, then issued 32 at a
        // time by ONE convergent instruction sequence (wait for the slot, write its metadata, arm the
        // mbarrier, bulk copy).  With a selective mask this turns ~8 short divergent sequences per chunk
        // into one full-width pass; the long-latency steps (shared atomics, mbarrier ops, the per-lane
        // serialised UBLKCP issue) are paid per pass, not per 32-slot group.
        uint2 *clist = reinterpret_cast<uint2 *>(smem + L.clist_off) + (warp - kConsumers) * (kChunk * 32);
        uint32_t *meta32 = reinterpret_cast<uint32_t *>(meta);
        auto issue_chunk = [&](uint32_t c, const uint32_t (&f)[kChunk], const uint64_t (&b)[kChunk]) {
            uint32_t T = 0;
            const uint32_t g0 = c >> 4;
#pragma unroll
            for (int j = 0; j < kChunk; j++) {
                const uint32_t row = ((g0 + j) << 5) + lane;
                const bool pass = (f[j] & need) && ((b[j] & a.mask) == a.mask) && row != a.exclude;
                const unsigned m = __ballot_sync(0xffffffffu, pass);
                if (pass) clist[T + __popc(m & ((1u << lane) - 1u))] = make_uint2(row, f[j]);
                T += __popc(m);
            }
            {   // passing rows per group seen in this chunk -> length to ask for next (>= 40 %: 2, >= 5 %: 4, else 8)
                const uint32_t len = c & 15u;
                const uint32_t pick = T * 10u >= 260u * len ? 1u : (T * 10u >= 130u * len ? 2u : (T * 10u >= 16u * len ? 4u : 8u));
                want = pick < cg ? pick : cg;
            }
            if (T == 0) return;
            __syncwarp();  // clist writes -> reads by other lanes
            uint32_t seq0 = 0;
            if (lane == 0) seq0 = atomicAdd(&s_seq, T);
            seq0 = __shfl_sync(0xffffffffu, seq0, 0);
            for (uint32_t t0 = 0; t0 < T; t0 += 32) {
                const uint32_t i = t0 + lane;
                if (i < T) {
                    const uint2 me = clist[i];
                    const uint32_t my = seq0 + i, sub = my & 1u, pair = my >> 1;
                    const uint32_t gen = pair / NP, slot = pair - gen * NP;
                    const uint32_t fullb = full0 + 8 * slot;
                    // a pair whose two rows are neighbours in HBM and sit in the same pass travels as one copy
                    const bool next_adj = sub == 0 && lane < 31 && i + 1 < T && clist[i + 1].x == me.x + 1;
                    const bool prev_adj = sub == 1 && lane > 0 && clist[i - 1].x + 1 == me.x;
                    const uint32_t bytes = prev_adj ? 0u : (next_adj ? 2 * RB : RB);
                    // every lane waits for the slot's previous round to be handed back before it touches
                    // the slot's metadata or rows
                    wait_gen_backoff(empty0 + 8 * slot, gen);
                    meta32[slot * 4 + sub * 2] = me.x;
                    meta32[slot * 4 + sub * 2 + 1] = me.y;
                    if (bytes) {
                        mbar_arrive_expect_tx(fullb, bytes);
                        bulk_g2s(rows_s + (slot * 2 + sub) * RB, Eb + (size_t)me.x * RB, bytes, fullb, pol_stream);
                    } else {
                        mbar_arrive(fullb);
                    }
                }
                __syncwarp();
            }
        };
        uint32_t cA = (((blockIdx.x * kProducers + (warp - kConsumers)) * cg) << 4) | cg;
        uint32_t fA[kChunk], fB[kChunk]; uint64_t bA[kChunk], bB[kChunk];
        load_chunk_pred(cA, fA, bA);
        uint32_t next_raw = fetch_chunk();
        while (chunk_live(cA)) {
            const uint32_t cB = __shfl_sync(0xffffffffu, next_raw, 0);
            load_chunk_pred(cB, fB, bB);
            next_raw = fetch_chunk();
            issue_chunk(cA, fA, bA);
            if (!chunk_live(cB)) break;
            cA = __shfl_sync(0xffffffffu, next_raw, 0);
            load_chunk_pred(cA, fA, bA);
            next_raw = fetch_chunk();
            issue_chunk(cB, fB, bB);
        }
        // ---- end of stream: wait for the other producers, then one warp closes the sequence ----
        asm volatile("bar.sync 1, %0;" ::"n"(kProducers * 32) : "memory");
        if (a.dbg && threadIdx.x == kConsumers * 32) a.dbg[blockIdx.x * 8 + 1] = gtime();
        if (warp == kConsumers) {
            const uint32_t seq = *reinterpret_cast<volatile uint32_t *>(&s_seq);
            uint32_t pair = seq >> 1;
            if (seq & 1u) {  // a half-filled pair: close it with an empty second row
                if (lane == 0) {
                    const uint32_t slot = pair % NP;
                    meta[slot].z = kIdxSentinel; meta[slot].w = 0;
                    mbar_arrive(full0 + 8 * slot);
                }
                pair++;
            }
            if (lane < kConsumers) {  // one sentinel pair per consumer
                pair += lane;
                const uint32_t gen = pair / NP, slot = pair - gen * NP;
                wait_gen(empty0 + 8 * slot, gen);
                meta[slot] = make_uint4(kIdxSentinel, 0, kIdxSentinel, 0);
                mbar_arrive(full0 + 8 * slot);
                mbar_arrive(full0 + 8 * slot);
            }
        }
    } else if constexpr (!QS) {
        // ================================ consumers ================================
        const float4 *q4 = QP ? reinterpret_cast<const float4 *>(smem + L.q_off) : reinterpret_cast<const float4 *>(a.q);
        float4 qv[V];
        float ma = 0.f;
#pragma unroll
        for (int j = 0; j < V; j++) {
            qv[j] = q4[j * 32 + lane];
            ma = fmaf(qv[j].x, qv[j].x, ma); ma = fmaf(qv[j].y, qv[j].y, ma);
            ma = fmaf(qv[j].z, qv[j].z, ma); ma = fmaf(qv[j].w, qv[j].w, ma);
        }
        ma = warp_sum(ma);
        const bool thresholds = a.min_sim > 0.f || a.max_dist > 0.f;
        uint32_t nsurv = 0;  // rows that passed predicate + thresholds (the reference's "active_keys")
        auto finish = [&](uint32_t row, uint32_t f, float dot, float mb, float ss) {
            float sim, dist;
            if (f & SPL_ROW_EMB) {
                sim = cosine_from_sums(dot, ma, mb);
                dist = __fsqrt_rn(ss);
                if (a.min_sim > 0.f && sim < a.min_sim) return;
                if (a.max_dist > 0.f && dist > a.max_dist) return;
            } else {  // _search.c:388-393
                if (thresholds) return;
                sim = 0.f; dist = FLT_MAX;
            }
            nsurv++;
            const uint64_t key = make_key(a.order, sim, dist);
            if (a.keys_out) {   // unlimited / large-k search: one key per surviving row, sorted afterwards (spl_sortall.cu)
                if (lane == 0) a.keys_out[row] = key;
            } else if (W.admits(key, row)) {
                W.insert(key, row, lane);
            }
        };
        uint32_t slot = warp, gen = 0;  // this warp's pair index = gen * NP + slot
        if (slot >= NP) { slot -= NP; gen++; }
        for (;;) {
            // When the ring size is not a multiple of the consumer count a slot's successive rounds belong to
            // DIFFERENT warps, and one of them can get here before the previous round has even been filled; a
            // 1-bit phase would then read as "complete".  The generation word says the previous round is done.
            wait_gen(empty0 + 8 * slot, gen);
            mbar_wait(full0 + 8 * slot, gen & 1u);
            const uint4 mt = meta[slot];
            if (mt.x == kIdxSentinel) break;
            const float4 *r0 = reinterpret_cast<const float4 *>(smem + L.rows_off + (size_t)slot * 2 * RB) + lane;
            const float4 *r1 = reinterpret_cast<const float4 *>(reinterpret_cast<const unsigned char *>(r0) + RB);
            const bool two = mt.z != kIdxSentinel;
            float4 x[V], y[V];
#pragma unroll
            for (int j = 0; j < V; j++) x[j] = r0[32 * j];
#pragma unroll
            for (int j = 0; j < V; j++) y[j] = r1[32 * j];   // a lone row's partner slot holds stale bytes: scored, never used
            __syncwarp();
            if (lane == 0) publish_gen(empty0 + 8 * slot, gen + 1);  // both rows now live in registers
            slot += kConsumers;
            if (slot >= NP) { slot -= NP; gen++; }
            float2 pd0 = make_float2(0.f, 0.f), pe0 = pd0, ps0 = pd0, pd1 = pd0, pe1 = pd0, ps1 = pd0;
#pragma unroll
            for (int j = 0; j < V; j++) {
                score4(qv[j], x[j], pd0, ps0); norm4(x[j], pe0);
                score4(qv[j], y[j], pd1, ps1); norm4(y[j], pe1);
            }
            float d0 = pd0.x + pd0.y, e0 = pe0.x + pe0.y, s0 = ps0.x + ps0.y;
            float d1 = pd1.x + pd1.y, e1 = pe1.x + pe1.y, s1 = ps1.x + ps1.y;
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                d0 += __shfl_xor_sync(0xffffffffu, d0, off);
                e0 += __shfl_xor_sync(0xffffffffu, e0, off);
                s0 += __shfl_xor_sync(0xffffffffu, s0, off);
                d1 += __shfl_xor_sync(0xffffffffu, d1, off);
                e1 += __shfl_xor_sync(0xffffffffu, e1, off);
                s1 += __shfl_xor_sync(0xffffffffu, s1, off);
            }
            finish(mt.x, mt.y, d0, e0, s0);
            if (two) finish(mt.z, mt.w, d1, e1, s1);
        }
        if (a.total && lane == 0 && nsurv) atomicAdd(&a.counters[2], nsurv);
    } else {
        // ===================== consumers, NQ queries per pass over the store =====================
        // The pair of rows sits in registers as before; the queries are streamed from shared memory, so every
        // row byte fetched from HBM is scored NQ times.  The 4*NQ partial sums of a pair (dot and sum of squared
        // differences, per row, per query) are reduced across the warp TRANSPOSED: each halving step swaps half
        // of the values, so after five steps lane L holds the complete sum of value (L / TB) in 4*NQ + ~8
        // shuffles instead of 5 per value -- in the same pairwise order as the plain butterfly, so the scores are
        // bit-identical to the single-query kernel's.  Each (query, row) result is then scored, thresholded and
        // tested against its list by its own lane, all 2*NQ of them at once; only an admitted candidate (rare
        // once the lists have warmed up) takes a warp-wide detour.
        constexpr int NV = 4 * NQ;        // value index = (query * 2 + row) * 2 + {0: dot, 1: sum (q-e)^2}
        constexpr uint32_t TB = 32 / NV;  // lane L ends up holding value L / TB
        const float4 *sq4 = reinterpret_cast<const float4 *>(smem + L.q_off) + lane;
        uint4 *wstate = reinterpret_cast<uint4 *>(smem + L.state_off);
        float *smaq = reinterpret_cast<float *>(reinterpret_cast<uint32_t *>(wstate + NQ * kConsumers) + NQ * kConsumers);
        const uint32_t qstride4 = a.dim >> 2;
#pragma unroll 1
        for (int qi = 0; qi < NQ; qi++) {
            float m = 0.f;
#pragma unroll
            for (int j = 0; j < V; j++) {
                const float4 v = sq4[qi * qstride4 + 32 * j];
                m = fmaf(v.x, v.x, m); m = fmaf(v.y, v.y, m); m = fmaf(v.z, v.z, m); m = fmaf(v.w, v.w, m);
            }
            m = warp_sum(m);
            if (lane == 0) smaq[qi] = m;   // every consumer warp stores the same value
        }
        __syncwarp();
        const uint32_t my_q = lane / (4 * TB), my_r = (lane / (2 * TB)) & 1u, my_t = (lane / TB) & 1u;
        const bool active = (lane % (2 * TB)) == 0;        // one lane per (query, row)
        const uint32_t group_mask = (uint32_t)((1ull << (4 * TB)) - 1ull) << (my_q * 4 * TB);
        const float maq = smaq[my_q];
        const bool thresholds = a.min_sim > 0.f || a.max_dist > 0.f;
        uint4 *my_state = wstate + my_q * kConsumers + warp;
        uint32_t nsurv = 0;  // survivors of this lane's query
        const bool lazy_ok = a.order == SPL_ORDER_COSINE && !(a.max_dist > 0.f) && !a.keys_out;
        auto tsum = [&](auto &v, auto n_tag) -> float {      // transposed warp sum of v[N]
            constexpr int N = decltype(n_tag)::value;
            uint32_t width = 16;
#pragma unroll
            for (int n = N; n > 1; n >>= 1, width >>= 1) {
                const bool hi = (lane & width) != 0;
#pragma unroll
                for (int i = 0; i < n / 2; i++) {
                    const float send = hi ? v[i] : v[i + n / 2];
                    const float keep = hi ? v[i + n / 2] : v[i];
                    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, width);
                }
            }
            float r = v[0];
            for (; width > 0; width >>= 1) r += __shfl_xor_sync(0xffffffffu, r, width);
            return r;
        };
        // ---- a CTA-wide admission bound on the key's primary word ---------------------------------------------------
        // A warp sees a sixteenth of the CTA's rows; its own list's k-th key is a loose bound (with a selective mask a
        // warp meets a few dozen rows per query and admits every other one).  So the CTA also keeps, per query, the k
        // smallest PRIMARY words (32 bits: one CAS) of the rows its warps have admitted, lock-free: a row whose primary
        // word is larger than the largest of those k has k rows of this CTA strictly ahead of it and cannot be in the
        // answer, whatever its warp's list looks like.  Words only ever shrink, so a stale read is a looser bound, never a
        // wrong one; ties on the word are let through to the exact 96-bit comparison of the warp's list.
        auto cta_note = [&](uint32_t cq, uint32_t hi) __attribute__((always_inline)) {   // warp-uniform arguments
            volatile uint32_t *sl = s_cta_hi + cq * a.k;
            for (;;) {
                uint32_t mx = 0, mp = 0;
                for (uint32_t p = lane; p < a.k; p += kWarp) { const uint32_t v = sl[p]; if (v >= mx) { mx = v; mp = p; } }
                const uint32_t wmx = __reduce_max_sync(0xffffffffu, mx);
                if (hi >= wmx) {   // nothing (more) to replace: wmx bounds the k-th smallest word from above
                    if (lane == 0) atomicMin(&s_cta_kth[cq], wmx);
                    break;
                }
                const unsigned who = __ballot_sync(0xffffffffu, lane < a.k && mx == wmx);
                const uint32_t pos = __shfl_sync(0xffffffffu, mp, __ffs(who) - 1);
                uint32_t ok = 0;
                if (lane == 0) ok = atomicCAS(&s_cta_hi[cq * a.k + pos], wmx, hi) == wmx;
                if (__shfl_sync(0xffffffffu, ok, 0)) hi = 0xffffffffu;   // placed: one more pass publishes the new maximum
            }
        };
        uint32_t slot = warp, gen = 0;
        if (slot >= NP) { slot -= NP; gen++; }
        for (;;) {
            wait_gen(empty0 + 8 * slot, gen);   // see the single-query loop for why both waits are needed
            mbar_wait(full0 + 8 * slot, gen & 1u);
            const uint4 mt = meta[slot];
            if (mt.x == kIdxSentinel) break;
            const float4 *r0 = reinterpret_cast<const float4 *>(smem + L.rows_off + (size_t)slot * 2 * RB) + lane;
            const float4 *r1 = reinterpret_cast<const float4 *>(reinterpret_cast<const unsigned char *>(r0) + RB);
            const bool two = mt.z != kIdxSentinel;
            float4 x[V], y[V];
#pragma unroll
            for (int j = 0; j < V; j++) x[j] = r0[32 * j];
#pragma unroll
            for (int j = 0; j < V; j++) y[j] = r1[32 * j];   // a lone row's partner slot holds stale bytes: scored, never used
            __syncwarp();
            if (lane == 0) publish_gen(empty0 + 8 * slot, gen + 1);
            slot += kConsumers;
            if (slot >= NP) { slot -= NP; gen++; }
            float2 pe0 = make_float2(0.f, 0.f), pe1 = pe0;
#pragma unroll
            for (int j = 0; j < V; j++) { norm4(x[j], pe0); norm4(y[j], pe1); }
            float ev[2] = { pe0.x + pe0.y, pe1.x + pe1.y };
            if (lazy_ok) {
                // ---- lazy body: with the CTA-wide bound few candidates are admitted.  In the reference's order (similarity
                // first) the distance then only matters for the few that are: score the dot products alone (a third of
                // the FMAs, half of the reduction), test the similarity against the bounds, and compute sum (q-e)^2 --
                // same per-lane order, same butterfly, hence the same bits as the full body -- only for a candidate
                // whose similarity does not already rule it out.
                constexpr int NVL = 2 * NQ;          // value index = query * 2 + row
                constexpr uint32_t TBL = 32 / NVL;   // lane L ends up holding value L / TBL
                float accl[NVL];
#pragma unroll
                for (int qi = 0; qi < NQ; qi++) {
                    const float4 *qrow = sq4 + qi * qstride4;
                    float2 pd0 = make_float2(0.f, 0.f), pd1 = pd0;
#pragma unroll
                    for (int j = 0; j < V; j++) {
                        const float4 qv = qrow[32 * j];
                        pd0 = fma2(lo2(qv), lo2(x[j]), pd0); pd0 = fma2(hi2(qv), hi2(x[j]), pd0);
                        pd1 = fma2(lo2(qv), lo2(y[j]), pd1); pd1 = fma2(hi2(qv), hi2(y[j]), pd1);
                    }
                    accl[qi * 2 + 0] = pd0.x + pd0.y; accl[qi * 2 + 1] = pd1.x + pd1.y;
                }
                const float etot = tsum(ev, std::integral_constant<int, 2>{});
                const float dotl = tsum(accl, std::integral_constant<int, NVL>{});
                const uint32_t lq = lane / (2 * TBL), lr = (lane / TBL) & 1u;
                const bool lactive = (lane % TBL) == 0;          // one lane per (query, row)
                const float mbl = __shfl_sync(0xffffffffu, etot, lr << 4);
                const uint32_t rowl = lr ? mt.z : mt.x, fl = lr ? mt.w : mt.y;
                bool passl = lactive && (lr == 0 || two);
                float siml = 0.f;
                if (fl & SPL_ROW_EMB) {
                    siml = cosine_from_sums(dotl, smaq[lq], mbl);
                    if (a.min_sim > 0.f && siml < a.min_sim) passl = false;
                } else if (a.min_sim > 0.f) {   // _search.c:388-393
                    passl = false;
                }
                static_assert(2 * TBL == 4 * TB, "both bodies give a query the same group of lanes");
                nsurv += __popc(__ballot_sync(0xffffffffu, passl) & group_mask);
                const uint4 stl = *(wstate + lq * kConsumers + warp);
                // similarity alone decides unless it ties with the bound's: ~ord(sim) is the key's high word
                const uint32_t hil = ~f2ord(siml);
                const bool admitl = passl && hil <= *reinterpret_cast<volatile uint32_t *>(&s_cta_kth[lq]) &&
                                    ((stl.w & 0xffffu) < a.k || hil <= stl.y);
                unsigned ml = __ballot_sync(0xffffffffu, admitl);
                while (ml) {   // warp-uniform
                    const int l = __ffs(ml) - 1;
                    ml &= ml - 1;
                    const uint32_t cq = (uint32_t)l / (2 * TBL), cr = ((uint32_t)l / TBL) & 1u;
                    const uint32_t crow = __shfl_sync(0xffffffffu, rowl, l);
                    const uint32_t cf = __shfl_sync(0xffffffffu, fl, l);
                    const float csim = __shfl_sync(0xffffffffu, siml, l);
                    cta_note(cq, ~f2ord(csim));
                    float cdist = FLT_MAX;
                    if (cf & SPL_ROW_EMB) {
                        const float4 *qrow = sq4 + cq * qstride4;
                        float2 ps = make_float2(0.f, 0.f);
                        if (cr) {
#pragma unroll
                            for (int j = 0; j < V; j++) {
                                const float4 qv = qrow[32 * j];
                                float2 t = sub2(lo2(qv), lo2(y[j])); ps = fma2(t, t, ps);
                                t = sub2(hi2(qv), hi2(y[j])); ps = fma2(t, t, ps);
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < V; j++) {
                                const float4 qv = qrow[32 * j];
                                float2 t = sub2(lo2(qv), lo2(x[j])); ps = fma2(t, t, ps);
                                t = sub2(hi2(qv), hi2(x[j])); ps = fma2(t, t, ps);
                            }
                        }
                        float ssum = ps.x + ps.y;
#pragma unroll
                        for (int off = 16; off > 0; off >>= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, off);
                        cdist = __fsqrt_rn(ssum);
                    }
                    const uint64_t ckey = make_key(a.order, csim, cdist);
                    uint4 *stp = wstate + cq * kConsumers + warp;
                    const uint4 cs = *stp;
                    WarpTopK T;
                    T.keys = skeys + cq * qstride + warp * a.k; T.idxs = sidx + cq * qstride + warp * a.k;
                    T.k = a.k; T.cnt = cs.w & 0xffffu; T.thr_pos = cs.w >> 16;
                    T.thr_key = ((uint64_t)cs.y << 32) | cs.x; T.thr_idx = cs.z;
                    if (T.admits(ckey, crow)) {
                        T.insert(ckey, crow, lane);
                        if (lane == 0) *stp = make_uint4((uint32_t)T.thr_key, (uint32_t)(T.thr_key >> 32), T.thr_idx, (T.thr_pos << 16) | T.cnt);
                    }
                    __syncwarp();
                }
                continue;
            }
            float acc[NV];
#pragma unroll
            for (int qi = 0; qi < NQ; qi++) {
                const float4 *qrow = sq4 + qi * qstride4;
                float2 pd0 = make_float2(0.f, 0.f), ps0 = pd0, pd1 = pd0, ps1 = pd0;
#pragma unroll
                for (int j = 0; j < V; j++) {
                    const float4 qv = qrow[32 * j];
                    score4(qv, x[j], pd0, ps0);
                    score4(qv, y[j], pd1, ps1);
                }
                acc[qi * 4 + 0] = pd0.x + pd0.y; acc[qi * 4 + 1] = ps0.x + ps0.y;
                acc[qi * 4 + 2] = pd1.x + pd1.y; acc[qi * 4 + 3] = ps1.x + ps1.y;
            }
            const float etot = tsum(ev, std::integral_constant<int, 2>{});   // lanes 0-15: |row0|^2, 16-31: |row1|^2
            const float tot = tsum(acc, std::integral_constant<int, NV>{});
            const float other = __shfl_xor_sync(0xffffffffu, tot, TB);
            const float dot = my_t ? other : tot, ss = my_t ? tot : other;
            const float mb = __shfl_sync(0xffffffffu, etot, my_r << 4);
            const uint32_t row = my_r ? mt.z : mt.x, f = my_r ? mt.w : mt.y;
            bool pass = active && (my_r == 0 || two);
            float sim, dist;
            if (f & SPL_ROW_EMB) {
                sim = cosine_from_sums(dot, maq, mb);
                dist = __fsqrt_rn(ss);
                if (a.min_sim > 0.f && sim < a.min_sim) pass = false;
                if (a.max_dist > 0.f && dist > a.max_dist) pass = false;
            } else {  // _search.c:388-393
                if (thresholds) pass = false;
                sim = 0.f; dist = FLT_MAX;
            }
            const uint64_t key = make_key(a.order, sim, dist);
            nsurv += __popc(__ballot_sync(0xffffffffu, pass) & group_mask);
            if (a.keys_out) {   // key-per-row mode (single query with the query in shared memory): see the loop above
                if (pass) a.keys_out[row] = key;
                continue;
            }
            const uint4 st = *my_state;
            const bool admit = pass && (uint32_t)(key >> 32) <= *reinterpret_cast<volatile uint32_t *>(&s_cta_kth[my_q]) &&
                               ((st.w & 0xffffu) < a.k || better(key, row, ((uint64_t)st.y << 32) | st.x, st.z));
            unsigned m = __ballot_sync(0xffffffffu, admit);
            while (m) {   // warp-uniform
                const int l = __ffs(m) - 1;
                m &= m - 1;
                const uint64_t ckey = __shfl_sync(0xffffffffu, key, l);
                const uint32_t crow = __shfl_sync(0xffffffffu, row, l);
                const uint32_t cq = (uint32_t)l / (4 * TB);
                cta_note(cq, (uint32_t)(ckey >> 32));
                uint4 *stp = wstate + cq * kConsumers + warp;
                const uint4 cs = *stp;   // re-read: an earlier candidate of this pair may have moved the bound
                WarpTopK T;
                T.keys = skeys + cq * qstride + warp * a.k; T.idxs = sidx + cq * qstride + warp * a.k;
                T.k = a.k; T.cnt = cs.w & 0xffffu; T.thr_pos = cs.w >> 16;
                T.thr_key = ((uint64_t)cs.y << 32) | cs.x; T.thr_idx = cs.z;
                if (T.admits(ckey, crow)) {
                    T.insert(ckey, crow, lane);
                    if (lane == 0) *stp = make_uint4((uint32_t)T.thr_key, (uint32_t)(T.thr_key >> 32), T.thr_idx, (T.thr_pos << 16) | T.cnt);
                }
                __syncwarp();
            }
        }
        if (a.total && (lane % (4 * TB)) == 0 && nsurv) atomicAdd(&a.counters[2 + my_q], nsurv);
    }
    if (a.dbg && threadIdx.x == 0) a.dbg[blockIdx.x * 8 + 2] = gtime();
    // ---- back-to-back scans (programmatic dependent launch) ---------------------------------------------------------
    // What is left of this launch after the scan loop is a few microseconds of merging, most of it in ONE CTA while
    // the other SMs idle.  If the next launch on the stream is another scan (the host says so with the launch
    // attribute), let it begin now: its CTAs take the SMs as this launch's CTAs leave.  Two rules keep that safe.
    //  * Consecutive launches use alternating scratch sets (work counter, tickets, per-CTA lists), so launch i+1
    //    shares nothing with launch i but the read-only mirror.  Launch i+2 reuses launch i's set: by then launch i
    //    must be over, and it is, because every CTA WAITS for the launch before its own to have completed (and
    //    flushed) before it gives the go-ahead -- launch i+2 cannot start unless all of launch i+1 have done so.
    //  * The same wait orders the result writes of consecutive launches (a caller may hand every query the same
    //    output buffer).
    // Without the launch attribute both instructions do nothing.
    if (warp == kConsumers && lane == 0) {   // a producer lane: done as soon as the sequence is closed
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    }
    if (a.keys_out) {   // nothing to merge: the last CTA re-arms the counters for the next launch
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(&a.counters[1], 1u) == gridDim.x - 1) {
                if (a.total) *a.total = *reinterpret_cast<volatile uint32_t *>(&a.counters[2]);
                a.counters[0] = 0; a.counters[1] = 0; a.counters[2] = 0;
            }
        }
        return;
    }
    // the row ring is idle now: it becomes the workspace of the final cross-CTA merge
    uint32_t ws_entries = topk_pow2((NP * 2 * RB) / 12 + 1) >> 1;  // largest power of two that fits
    uint64_t *ws_keys = reinterpret_cast<uint64_t *>(smem + L.rows_off);
    uint32_t *ws_idx = reinterpret_cast<uint32_t *>(ws_keys + ws_entries);
    if (a.k <= 32) {
        scan_epilogue_compact<kConsumers>(a, NQ, qstride, skeys, sidx, smem + L.rows_off, &s_flag, &s_seq);
    } else {
#pragma unroll 1
        for (uint32_t qi = 0; qi < NQ; qi++)
            scan_epilogue<kConsumers, false>(a, skeys + qi * qstride, sidx + qi * qstride, ws_keys, ws_idx, ws_entries, &s_flag, qi);
    }
    if (a.dbg && threadIdx.x == 0) a.dbg[blockIdx.x * 8 + 3] = gtime();
    if (a.done_seq_ptr && s_flag) {   // the last CTA, after its last result write: tell a polling host the answer is complete
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();
            *reinterpret_cast<volatile uint32_t *>(a.done_seq_ptr) = a.done_seq;
        }
    }
}

bool scan_tma_supported(uint32_t dim, uint32_t k) {
    if (dim != 768 && dim != 1024 && dim != 512 && dim != 256 && dim != 128) return false;
    return tma_layout(dim, k).pairs >= kMinPairs;
}

// small-batch variant: nq in {2, 4, 8} queries per pass
bool scan_tma_multi_supported(uint32_t dim, uint32_t k, uint32_t nq) {
    if (nq != 2 && nq != 4 && nq != 8) return false;
    if (nq == 8 && (k > 32 || dim > 768)) return false;   // eight queries: the register/shuffle epilogue, rows of up to 768 floats
    if (k == 0 || k > 256 || !scan_tma_supported(dim, k)) return false;
    const TmaLayout L = tma_layout(dim, k, nq);
    return L.pairs >= kMinPairs;
}

template <int V, int NQ>
static cudaError_t launch_v(const ScanArgs &a, int grid, cudaStream_t stream) {
    const TmaLayout L = tma_layout(a.dim, a.k, NQ);
    if (!a.overlap_prev) {
        scan_tma_kernel<V, NQ><<<grid, kTmaThreads, L.total, stream>>>(a);
        return cudaGetLastError();
    }
    // may begin while the previous kernel on the stream (another scan: the caller of this function vouches for it) is
    // still in its tail; see the protocol in the kernel
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kTmaThreads);
    cfg.dynamicSmemBytes = L.total;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, scan_tma_kernel<V, NQ>, a);
}

template <int V>
static cudaError_t launch_nq(const ScanArgs &a, int grid, cudaStream_t stream) {
    switch (a.nq) {
        case 0: case 1: return launch_v<V, 1>(a, grid, stream);
        case 2: return launch_v<V, 2>(a, grid, stream);
        case 4: return launch_v<V, 4>(a, grid, stream);
        case 8: return launch_v<V, 8>(a, grid, stream);
        default: return cudaErrorInvalidValue;
    }
}

template <int V>
static cudaError_t prepare_v(int bytes) {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(scan_tma_kernel<V, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    if ((e = cudaFuncSetAttribute(scan_tma_kernel<V, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    if ((e = cudaFuncSetAttribute(scan_tma_kernel<V, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    return cudaFuncSetAttribute(scan_tma_kernel<V, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

// opt in to > 48 KB dynamic shared memory; call once per device with that device current
cudaError_t scan_tma_prepare() {
    cudaError_t e;
    const int bytes = (int)kSmemBudget - 8192;  // static smem (exchange buffers, flags) takes the rest
    if ((e = prepare_v<6>(bytes))) return e;
    if ((e = prepare_v<8>(bytes))) return e;
    if ((e = prepare_v<4>(bytes))) return e;
    if ((e = prepare_v<2>(bytes))) return e;
    if ((e = prepare_v<1>(bytes))) return e;
    if ((e = cudaFuncSetAttribute(scan_tma_qparam_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    if ((e = cudaFuncSetAttribute(scan_tma_qparam_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    if ((e = cudaFuncSetAttribute(scan_tma_qparam_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes))) return e;
    return cudaFuncSetAttribute(scan_tma_qparam_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

template <int V>
static cudaError_t launch_qparam_v(const ScanArgs &a, const float *q_host, int grid, cudaStream_t stream) {
    const TmaLayout L = tma_layout(a.dim, a.k, 1, true);
    QueryArg<V> q;
    memcpy(q.v, q_host, sizeof q.v);
    scan_tma_qparam_kernel<V><<<grid, kTmaThreads, L.total, stream>>>(a, q);
    return cudaGetLastError();
}

bool scan_tma_qparam_supported(uint32_t dim, uint32_t k) {
    return (dim == 768 || dim == 512 || dim == 256 || dim == 128) && !tma_query_in_smem(dim, 1) &&
           tma_layout(dim, k, 1, true).pairs >= kMinPairs;
}

// single query handed over in the launch's parameter space (host pointer q_host, 4*dim bytes, read before returning)
cudaError_t launch_scan_tma_qparam(const ScanArgs &a_in, const float *q_host, int grid, cudaStream_t stream) {
    ScanArgs a = a_in;
    const uint32_t ngroups = (a.n + 31u) >> 5;
    uint32_t cg = 8;
    while (cg > 1 && ngroups / cg < (uint32_t)grid * kProducers * 4u) cg >>= 1;
    a.chunk_groups = cg;
    a.nq = 1;
    switch (a.dim) {
        case 768: return launch_qparam_v<6>(a, q_host, grid, stream);
        case 512: return launch_qparam_v<4>(a, q_host, grid, stream);
        case 256: return launch_qparam_v<2>(a, q_host, grid, stream);
        case 128: return launch_qparam_v<1>(a, q_host, grid, stream);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_scan_tma(const ScanArgs &a_in, int grid, cudaStream_t stream) {
    ScanArgs a = a_in;
    // aim for >= 4 chunks per producer warp so the dynamic hand-out can balance; cap at 8 groups
    const uint32_t ngroups = (a.n + 31u) >> 5;
    uint32_t cg = 8;
    while (cg > 1 && ngroups / cg < (uint32_t)grid * kProducers * 4u) cg >>= 1;
    a.chunk_groups = cg;
    switch (a.dim) {
        case 768: return launch_nq<6>(a, grid, stream);
        case 1024: return launch_nq<8>(a, grid, stream);
        case 512: return launch_nq<4>(a, grid, stream);
        case 256: return launch_nq<2>(a, grid, stream);
        case 128: return launch_nq<1>(a, grid, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace spl
