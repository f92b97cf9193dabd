// spl_scan_epilogue.cuh -- the top-k epilogue every single-query scan kernel ends with.
//
// The reference sorts ALL survivors with qsort(compare_results) and prints the first `limit`
// (/root/reference/splinter_cli_cmd_search.c:412-416).  Here each warp has kept a bounded
// candidate list while scanning; the epilogue turns those lists into the same k rows:
//   1. the CTA's NW lists (contiguous in shared memory, unused entries = sentinel) are sorted
//      together by one block-wide bitonic network; the first k entries are the CTA's top-k;
//   2. every CTA publishes its k entries; the LAST CTA to finish (ticket counter) gathers all
//      gridDim.x * k published entries into a larger shared-memory workspace, sorts again and
//      writes the k result records, already in the reference's order.
#pragma once
#include "spl_scan.cuh"

namespace spl {

__device__ __forceinline__ unsigned long long gtime_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}


__host__ __device__ inline uint32_t next_pow2(uint32_t x) {
    uint32_t p = 1;
    while (p < x) p <<= 1;
    return p;
}

// Ascending bitonic sort of n = 2^m (key, idx) pairs in shared memory by the whole CTA.
// Order is `better`: key ascending, then idx ascending.  `aux` (nullable) is carried along.
__device__ __forceinline__ void bitonic_sort_cta(uint64_t *keys, uint32_t *idx, uint32_t n, uint32_t *aux = nullptr) {
    for (uint32_t size = 2; size <= n; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            for (uint32_t t = threadIdx.x; t < (n >> 1); t += blockDim.x) {
                const uint32_t i = ((t & ~(stride - 1)) << 1) | (t & (stride - 1));
                const uint32_t j = i + stride;
                const bool up = (i & size) == 0;
                const uint64_t ki = keys[i], kj = keys[j];
                const uint32_t ii = idx[i], ij = idx[j];
                const bool j_first = better(kj, ij, ki, ii);
                if (j_first == up) {
                    keys[i] = kj; idx[i] = ij; keys[j] = ki; idx[j] = ii;
                    if (aux) { const uint32_t x = aux[i]; aux[i] = aux[j]; aux[j] = x; }
                }
            }
            __syncthreads();
        }
    }
}

// ---- k <= 32: one (key, idx) per lane, everything in registers + shuffles -------------------
// ascending bitonic sort of the 32 lane values
__device__ __forceinline__ void warp_sort32(uint64_t &key, uint32_t &idx, uint32_t lane) {
#pragma unroll
    for (uint32_t size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            const uint64_t ok = __shfl_xor_sync(0xffffffffu, key, stride);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, idx, stride);
            const bool up = (lane & size) == 0, lower = (lane & stride) == 0;
            const bool other_first = better(ok, oi, key, idx);
            if ((lower == up) ? other_first : !other_first) { key = ok; idx = oi; }
        }
    }
}
// a, b: two ascending 32-lane lists -> a := the 32 smallest of their union, ascending
__device__ __forceinline__ void warp_merge32(uint64_t &key, uint32_t &idx, uint64_t bkey, uint32_t bidx, uint32_t lane) {
    const uint64_t rk = __shfl_sync(0xffffffffu, bkey, 31 - lane);
    const uint32_t ri = __shfl_sync(0xffffffffu, bidx, 31 - lane);
    if (better(rk, ri, key, idx)) { key = rk; idx = ri; }  // bitonic sequence holding the 32 smallest
#pragma unroll
    for (uint32_t stride = 16; stride > 0; stride >>= 1) {
        const uint64_t ok = __shfl_xor_sync(0xffffffffu, key, stride);
        const uint32_t oi = __shfl_xor_sync(0xffffffffu, idx, stride);
        const bool lower = (lane & stride) == 0;
        const bool other_first = better(ok, oi, key, idx);
        if (lower ? other_first : !other_first) { key = ok; idx = oi; }
    }
}
// tree-merge the per-warp sorted lists of warps 0..NW-1 through `xk/xi` (NW*32 entries of smem);
// afterwards warp 0 holds the CTA-wide top-32 (ascending).  Called by ALL threads of the CTA.
template <int NW>
__device__ __forceinline__ void cta_tree_merge32(uint64_t &key, uint32_t &idx, uint64_t *xk, uint32_t *xi) {
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int step = 1; step < NW; step <<= 1) {
        if (warp < NW && (warp & (2 * step - 1)) == step) { xk[warp * 32 + lane] = key; xi[warp * 32 + lane] = idx; }
        __syncthreads();
        if (warp < NW && (warp & (2 * step - 1)) == 0 && warp + step < NW)
            warp_merge32(key, idx, xk[(warp + step) * 32 + lane], xi[(warp + step) * 32 + lane], lane);
        __syncthreads();
    }
}

template <int NW>
__device__ __forceinline__ void scan_epilogue_small(const ScanArgs &a, uint64_t *skeys, uint32_t *sidx, uint32_t *s_flag) {
    const uint32_t k = a.k, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    // 1. every list-holding warp sorts its own (unsorted, sentinel-padded) list
    uint64_t key = kKeySentinel; uint32_t idx = kIdxSentinel;
    if (warp < NW && lane < k) { key = skeys[warp * k + lane]; idx = sidx[warp * k + lane]; }
    if (warp < NW) warp_sort32(key, idx, lane);
    // 2. CTA top-32 by tree merge; the exchange buffer is a dedicated smem area passed by the caller
    __shared__ uint64_t xk[NW * 32];
    __shared__ uint32_t xi[NW * 32];
    cta_tree_merge32<NW>(key, idx, xk, xi);
    if (a.dbg && threadIdx.x == 0) a.dbg[blockIdx.x * 8 + 4] = gtime_ns();
    if (warp == 0 && lane < k) {
        Cand c; c.key = key; c.idx = idx; c.pad = 0;
        a.cta_cands[(size_t)blockIdx.x * k + lane] = c;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t ticket = atomicAdd(&a.counters[1], 1u);
        *s_flag = (ticket == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (a.dbg && threadIdx.x == 0) a.dbg[blockIdx.x * 8 + 5] = gtime_ns();
    if (*s_flag == 0) return;
    // 3. last CTA: each warp folds a strided subset of the gridDim.x published (sorted) lists
    __threadfence();
    key = kKeySentinel; idx = kIdxSentinel;
    if (warp < NW) {
        constexpr int kBatch = 10;  // lists loaded per round trip (148 CTAs / 16 warps: one round)
        for (uint32_t c0 = warp; c0 < gridDim.x; c0 += NW * kBatch) {
            uint64_t bk[kBatch]; uint32_t bi[kBatch];
#pragma unroll
            for (int j = 0; j < kBatch; j++) {
                const uint32_t c = c0 + j * NW;
                bk[j] = kKeySentinel; bi[j] = kIdxSentinel;
                if (c < gridDim.x && lane < k) {
                    const int4 raw = __ldcg(reinterpret_cast<const int4 *>(a.cta_cands + (size_t)c * k + lane));
                    bk[j] = ((uint64_t)(uint32_t)raw.y << 32) | (uint32_t)raw.x;
                    bi[j] = (uint32_t)raw.z;
                }
            }
#pragma unroll
            for (int j = 0; j < kBatch; j++)
                if (c0 + j * NW < gridDim.x) warp_merge32(key, idx, bk[j], bi[j], lane);
        }
    }
    if (a.dbg && threadIdx.x == 0) a.dbg[blockIdx.x * 8 + 6] = gtime_ns();
    cta_tree_merge32<NW>(key, idx, xk, xi);
    if (a.dbg && threadIdx.x == 0) a.dbg[blockIdx.x * 8 + 7] = gtime_ns();
    if (warp == 0) {
        if (lane == 0) {
            if (a.total) *a.total = *reinterpret_cast<volatile uint32_t *>(&a.counters[2]);
            a.counters[0] = 0; a.counters[1] = 0; a.counters[2] = 0;
        }
        const unsigned valid = __ballot_sync(0xffffffffu, lane < k && idx != kIdxSentinel);
        if (lane == 0) *a.out_count = __popc(valid);
        if (lane < k && idx != kIdxSentinel) {
            spl_gpu_result r;
            split_key(a.order, key, r.similarity, r.distance);
            r.slot_idx = (uint32_t)(a.row0 + idx);
            r.has_embedding = (a.flags[idx] & SPL_ROW_EMB) ? 1u : 0u;
            r.epoch = a.epoch[idx];
            r.bloom = a.bloom[idx];
            a.out[lane] = r;
        }
    }
}

// ---- compact epilogue for k <= 32 (TMA scan kernel; nq = 1, 2, 4 or 8 queries) ------------------------------
// Same three steps as scan_epilogue_small, written as ROLLED loops around one instance of each building block.
// The epilogue runs once per CTA, so its instructions are always cold: measured, the time of the unrolled
// version (~30 KB of straight-line code per query) was mostly instruction fetch, about 0.45 us per KB.
// Lists of (query qi, warp w) sit at skeys/sidx[qi*qstride + w*k].  `scratch` (the idle row ring, at least
// kEpilogueScratchBytes) holds the tree-merge exchange buffer and, in the last CTA, a staging area so that the
// loads of all published lists of a round are in flight together while the merges run as a rolled loop.
// Query qi writes a.out + qi*out_stride, a.out_count[qi*count_stride], a.total[qi] (from counters[2 + qi]).
// `width` (a power of two): only lanes [0, width) hold entries, the rest hold sentinels and need no sorting
__device__ __forceinline__ void warp_sort32_rolled(uint64_t &key, uint32_t &idx, uint32_t lane, uint32_t width = 32) {
#pragma unroll 1
    for (uint32_t size = 2; size <= width; size <<= 1) {
#pragma unroll 1
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            const uint64_t ok = __shfl_xor_sync(0xffffffffu, key, stride);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, idx, stride);
            const bool up = (lane & size) == 0, lower = (lane & stride) == 0;
            const bool other_first = better(ok, oi, key, idx);
            if ((lower == up) ? other_first : !other_first) { key = ok; idx = oi; }
        }
    }
}
// The merge and the tree merge exist ONCE per kernel (noinline) and are shared by the per-CTA phase and the last
// CTA's fold.  Measured on the last CTA (2^20 x 768, k = 10): fence 0.5 us, loading + staging 148 lists 0.9 us,
// 160 merges on 16 warps 4.5 us, final tree 2 us, records 1 us.  The merges are bound by the four schedulers'
// issue rate (about 20 dependent instructions per stage and warp, four warps per scheduler), not by shuffle
// latency: interleaving four independent merge chains per warp did not shorten them.
struct KeyIdx { uint64_t key; uint32_t idx; };
static __device__ __noinline__ KeyIdx warp_merge32_shared(KeyIdx a, KeyIdx b) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t rk = __shfl_sync(0xffffffffu, b.key, 31 - lane);
    const uint32_t ri = __shfl_sync(0xffffffffu, b.idx, 31 - lane);
    if (better(rk, ri, a.key, a.idx)) { a.key = rk; a.idx = ri; }
#pragma unroll 1
    for (uint32_t stride = 16; stride > 0; stride >>= 1) {
        const uint64_t ok = __shfl_xor_sync(0xffffffffu, a.key, stride);
        const uint32_t oi = __shfl_xor_sync(0xffffffffu, a.idx, stride);
        const bool lower = (lane & stride) == 0;
        const bool other_first = better(ok, oi, a.key, a.idx);
        if (lower ? other_first : !other_first) { a.key = ok; a.idx = oi; }
    }
    return a;
}
__device__ __forceinline__ void warp_merge32_rolled(uint64_t &key, uint32_t &idx, uint64_t bkey, uint32_t bidx, uint32_t) {
    const KeyIdx r = warp_merge32_shared(KeyIdx{ key, idx }, KeyIdx{ bkey, bidx });
    key = r.key; idx = r.idx;
}
// tree merge of the sorted lists of warps 0..nw-1 through xk/xi; afterwards warp 0 holds the CTA's best 32.
// Called by ALL threads of the CTA.
// `group` (a power of two dividing nw): the warps merge in independent groups of that many -- afterwards the first warp
// of every group holds its group's best 32 (group = nw: one tree over all warps).
static __device__ __noinline__ KeyIdx cta_tree_merge32_shared(KeyIdx v, uint64_t *xk, uint32_t *xi, uint32_t nw, uint32_t group) {
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wg = warp & (group - 1);
#pragma unroll 1
    for (uint32_t step = 1; step < group; step <<= 1) {
        if (warp < nw && (wg & (2 * step - 1)) == step) { xk[warp * 32 + lane] = v.key; xi[warp * 32 + lane] = v.idx; }
        __syncthreads();
        if (warp < nw && (wg & (2 * step - 1)) == 0 && warp + step < nw)
            v = warp_merge32_shared(v, KeyIdx{ xk[(warp + step) * 32 + lane], xi[(warp + step) * 32 + lane] });
        // no barrier here: a region written at one level is not written again at a later one
    }
    __syncthreads();  // the buffer is free again for the next call
    return v;
}
template <int NW>
__device__ __forceinline__ void cta_tree_merge32_rolled(uint64_t &key, uint32_t &idx, uint64_t *xk, uint32_t *xi, uint32_t group = NW) {
    const KeyIdx r = cta_tree_merge32_shared(KeyIdx{ key, idx }, xk, xi, NW, group);
    key = r.key; idx = r.idx;
}

constexpr int kFoldBatch = 10;  // published lists a warp loads per round (148 CTAs / 16 warps: one round)
template <int NW>
constexpr uint32_t epilogue_scratch_bytes() { return NW * 32 * 12 + NW * kFoldBatch * 32 * 12; }

template <int NW>
__device__ __forceinline__ void scan_epilogue_compact(const ScanArgs &a, uint32_t nq, uint32_t qstride, const uint64_t *skeys,
                                                      const uint32_t *sidx, unsigned char *scratch, uint32_t *s_flag,
                                                      const uint32_t *rows_issued = nullptr) {
    const uint32_t k = a.k, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t *xk = reinterpret_cast<uint64_t *>(scratch);
    uint32_t *xi = reinterpret_cast<uint32_t *>(xk + NW * 32);
    uint64_t *fk = reinterpret_cast<uint64_t *>(scratch + NW * 32 * 12);
    uint32_t *fi = reinterpret_cast<uint32_t *>(fk + NW * kFoldBatch * 32);
    __syncthreads();  // lists final, ring idle
    // a CTA whose producers issued no row at all (a very selective label mask) has nothing to sort or merge
    const bool cta_empty = rows_issued && *rows_issued == 0;
    // 1 + 2, small batch: the queries' merges are independent, so they run side by side -- NW / nq warps per query, each
    // sorting and merging nq of the query's NW lists (any warp can read any list: they sit in shared memory), then one
    // tree per group.  Four queries took four times the single query's 16-level tree before.
    // 1 + 2 by counting, for short lists: a query's NW*k list entries are n <= a few hundred (key, slot) pairs in shared
    // memory, all distinct.  One thread per entry counts the entries ahead of it -- the key's high word decides except
    // between entries that share it, and two keys come with every 128-bit load -- and an entry with fewer than k ahead
    // of it IS the CTA's entry number `rank`: it goes straight to its place in the published list.  No sort, no
    // merge tree, one barrier; about n*n/2 compare pairs per query instead of sixteen sorts and four tree levels
    // (measured on one query, k = 10: 6 us -> see DESIGN.md).  The quadratic cost limits it to small k.
    const uint32_t n = NW * k;
    const bool counted = !cta_empty && (n & 1u) == 0 && nq * n * n <= 210000u;
    if (counted) {
        __shared__ uint32_t s_valid[8];
        if (threadIdx.x < 8) s_valid[threadIdx.x] = 0;
        __syncthreads();
        uint32_t nvalid[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
#pragma unroll 1
        for (uint32_t c = threadIdx.x; c < nq * n; c += blockDim.x) {
            const uint32_t qi = c / n, i = c - qi * n;
            const uint64_t *lk = skeys + qi * qstride;
            const uint32_t *li = sidx + qi * qstride;
            const uint64_t mk = lk[i]; const uint32_t mi = li[i];
            if (mi == kIdxSentinel) continue;
#pragma unroll
            for (uint32_t v = 0; v < 8; v++) nvalid[v] += (qi == v) ? 1u : 0u;   // (constant indices: the counts stay in registers)
            const uint32_t mh = (uint32_t)(mk >> 32);
            const uint4 *k4 = reinterpret_cast<const uint4 *>(lk);
            uint32_t lt = 0, eq = 0;
#pragma unroll 4
            for (uint32_t j = 0; j < n / 2; j++) {
                const uint4 v = k4[j];   // {lo, hi} of two keys; the same address in every lane: a broadcast
                lt += (v.y < mh) + (v.w < mh);
                eq += (v.y == mh) + (v.w == mh);
            }
            uint32_t rank = lt;
            if (eq > 1) {   // others share the high word (rows without an embedding all score 0): the full comparison
                rank = 0;
                for (uint32_t j = 0; j < n; j++) rank += better(lk[j], li[j], mk, mi) ? 1u : 0u;
            }
            if (rank < k) {
                Cand cd; cd.key = mk; cd.idx = mi; cd.pad = 0;
                a.cta_cands[((size_t)qi * gridDim.x + blockIdx.x) * k + rank] = cd;
            }
        }
#pragma unroll
        for (uint32_t qi = 0; qi < 8; qi++) {
            if (qi < nq) {
                const uint32_t t = __reduce_add_sync(0xffffffffu, nvalid[qi]);
                if (lane == 0 && t) atomicAdd(&s_valid[qi], t);
            }
        }
        __syncthreads();
        if (threadIdx.x < nq * k) {   // fewer than k entries in all: the rest of the published list is padding
            const uint32_t qi = threadIdx.x / k, r = threadIdx.x - qi * k;
            if (r >= s_valid[qi]) {
                Cand cd; cd.key = kKeySentinel; cd.idx = kIdxSentinel; cd.pad = 0;
                a.cta_cands[((size_t)qi * gridDim.x + blockIdx.x) * k + r] = cd;
            }
        }
        __threadfence();   // every thread's part of the published lists, before the CTA's ticket below
        __syncthreads();
    }
    const bool grouped = !counted && nq > 1 && !cta_empty && (NW % nq) == 0 && ((NW / nq) & (NW / nq - 1)) == 0;
    if (grouped) {
        const uint32_t g = NW / nq, qi = warp / g, wg = warp & (g - 1);
        uint64_t key = kKeySentinel; uint32_t idx = kIdxSentinel;
        if (warp < NW) {
#pragma unroll 1
            for (uint32_t j = 0; j < nq; j++) {
                const uint32_t src = qi * qstride + (wg * nq + j) * k + lane;
                uint64_t lk = kKeySentinel; uint32_t li = kIdxSentinel;
                if (lane < k) { lk = skeys[src]; li = sidx[src]; }
                warp_sort32_rolled(lk, li, lane, next_pow2(k < 2 ? 2 : k));
                if (j == 0) { key = lk; idx = li; } else warp_merge32_rolled(key, idx, lk, li, lane);
            }
        }
        cta_tree_merge32_rolled<NW>(key, idx, xk, xi, g);
        if (warp < NW && wg == 0) {
            if (lane < k) {
                Cand c; c.key = key; c.idx = idx; c.pad = 0;
                a.cta_cands[((size_t)qi * gridDim.x + blockIdx.x) * k + lane] = c;
            }
            __threadfence();   // this warp's published list, before the CTA's ticket below
        }
        __syncthreads();
    }
#pragma unroll 1
    for (uint32_t qi = 0; qi < (grouped || counted ? 0u : nq); qi++) {
        uint64_t key = kKeySentinel; uint32_t idx = kIdxSentinel;
        if (cta_empty) {
            if (warp == 0 && lane < k) {
                Cand c; c.key = key; c.idx = idx; c.pad = 0;
                a.cta_cands[((size_t)qi * gridDim.x + blockIdx.x) * k + lane] = c;
            }
            continue;
        }
        if (warp < NW && lane < k) { key = skeys[qi * qstride + warp * k + lane]; idx = sidx[qi * qstride + warp * k + lane]; }
        if (warp < NW) warp_sort32_rolled(key, idx, lane, next_pow2(k < 2 ? 2 : k));
        cta_tree_merge32_rolled<NW>(key, idx, xk, xi);
        if (warp == 0 && lane < k) {
            Cand c; c.key = key; c.idx = idx; c.pad = 0;
            a.cta_cands[((size_t)qi * gridDim.x + blockIdx.x) * k + lane] = c;
        }
    }
    if (a.dbg && threadIdx.x == 0) a.dbg[blockIdx.x * 8 + 4] = gtime_ns();
    if (warp == 0) {   // the only warp that published: its fence, then its ticket -- no block-wide barrier needed before
        __threadfence();
        __syncwarp();
        if (lane == 0) *s_flag = atomicAdd(&a.counters[1], 1u);
    }
    __syncthreads();
    if (a.dbg && threadIdx.x == 0) a.dbg[blockIdx.x * 8 + 5] = gtime_ns();
    // The final merges of a small batch are independent, one per query: instead of the last CTA doing them one after the
    // other while 147 SMs idle, the last F = min(nq, grid) CTAs to publish take a query each.  The very last one starts
    // at once; the others first wait until every list is published.  The wait ends: a launch never has more than F - 1
    // <= 7 CTAs waiting here, every other CTA runs to its exit without waiting for anything, so the CTAs still to come
    // always find an SM (it would take some twenty small-batch launches in flight on one GPU at once -- each context
    // runs its searches one after another -- for the waiters to hold every SM).  Whichever of the F finishes last
    // re-arms the counters.
    const uint32_t F = nq < gridDim.x ? nq : gridDim.x;
    const uint32_t ticket = *s_flag;
    __syncthreads();   // everyone has read the ticket: s_flag becomes "this CTA wrote the launch's last result"
    if (threadIdx.x == 0) *s_flag = 0;
    __syncthreads();
    if (ticket + F < gridDim.x) return;
    const uint32_t fin = gridDim.x - 1 - ticket;   // 0 .. F-1: the queries fin, fin + F, ... are this CTA's
    if (fin != 0) {
        if (threadIdx.x == 0) {
            while (*reinterpret_cast<volatile uint32_t *>(&a.counters[1]) < gridDim.x) __nanosleep(100);
        }
        __syncthreads();
    }
    // 3. last CTA, per query: the k best of the gridDim.x published (sorted) lists.
    //
    // Fast path.  Let T be the k-th best of the lists' HEADS (their best entries).  At least k candidates -- those
    // k heads -- are at least as good as T, so nothing worse than T can be among the k best overall: only the
    // candidates <= T survive.  The k best of a scan are normally spread over about k different CTAs, so T is
    // close to the k-th best overall and a dozen or two candidates survive; one warp sorts them.  This replaces
    // ~160 list merges and a tree merge (6 + 2 us of the 14 us tail on a 2^20-row scan) by a rank count over the
    // heads and one filter pass.  If more than 64 survive (all the best rows sit in one CTA's range, or k is
    // close to 32) the general fold below runs instead.
    __threadfence();
    __shared__ uint32_t s_cnt, s_tidx;
    __shared__ uint64_t s_tkey;
    const uint32_t G = gridDim.x;
    const bool one_round = G <= 160u && G * k <= 8u * blockDim.x;   // the fast path's register budget (below)
    uint64_t *hk = xk, *sk = xk + 256;        // heads / survivors live in the (idle) tree-merge exchange buffer
    uint32_t *hi = xi, *si = xi + 256, *hh = xi + 320;   // hh: the heads' high words, 160 entries (16-byte aligned)
#pragma unroll 1
    for (uint32_t qi = fin; qi < nq; qi += F) {
        uint64_t key = kKeySentinel; uint32_t idx = kIdxSentinel;
        const Cand *lists = a.cta_cands + (size_t)qi * G * k;
        bool done = false;
        if (one_round) {
            // every candidate is loaded once, straight into registers (all loads of a thread in flight together);
            // a list's first entry doubles as its head
            constexpr int kPerThread = 8;
            uint64_t ck[kPerThread]; uint32_t ci[kPerThread];
            const uint32_t total = G * k;   // <= 160 * 32 = 5120 <= kPerThread * 640
            if (threadIdx.x == 0) { s_cnt = 0; s_tkey = kKeySentinel; s_tidx = kIdxSentinel; }
#pragma unroll
            for (int j = 0; j < kPerThread; j++) {
                const uint32_t i = threadIdx.x + j * blockDim.x;
                ck[j] = kKeySentinel; ci[j] = kIdxSentinel;
                if (i < total) {
                    const int4 raw = __ldcg(reinterpret_cast<const int4 *>(lists + i));
                    ck[j] = ((uint64_t)(uint32_t)raw.y << 32) | (uint32_t)raw.x;
                    ci[j] = (uint32_t)raw.z;
                }
            }
#pragma unroll
            for (int j = 0; j < kPerThread; j++) {
                const uint32_t i = threadIdx.x + j * blockDim.x;
                if (i < total) {
                    const uint32_t c = i / k;
                    if (i == c * k) { hk[c] = ck[j]; hi[c] = ci[j]; hh[c] = (uint32_t)(ck[j] >> 32); }
                }
            }
            if (threadIdx.x >= G && threadIdx.x < 160) hh[threadIdx.x] = 0xFFFFFFFFu;   // padding: never below a head
            __syncthreads();
            {   // Rank of every head among all heads (the order is total, so ranks are distinct): four threads per
                // head, a quarter of the comparisons each.  The high word of a key (the ordering score) almost always
                // decides, so it is compared first, four heads per 128-bit load; only a head whose high word occurs
                // more than once goes through the full (key, slot) comparison.
                const uint32_t h = threadIdx.x >> 2, part = threadIdx.x & 3u;
                uint32_t lt = 0, eq = 0;
                uint64_t mk = kKeySentinel; uint32_t mi = kIdxSentinel;
                if (h < G) { mk = hk[h]; mi = hi[h]; }
                const uint32_t mh = (uint32_t)(mk >> 32);
                if (h < G && mi != kIdxSentinel) {
                    const uint4 *h4 = reinterpret_cast<const uint4 *>(hh) + part * 10;   // 160 padded heads / 4 parts
#pragma unroll
                    for (int j = 0; j < 10; j++) {
                        const uint4 v = h4[j];
                        lt += (v.x < mh) + (v.y < mh) + (v.z < mh) + (v.w < mh);
                        eq += (v.x == mh) + (v.y == mh) + (v.z == mh) + (v.w == mh);
                    }
                }
                lt += __shfl_xor_sync(0xffffffffu, lt, 1); eq += __shfl_xor_sync(0xffffffffu, eq, 1);
                lt += __shfl_xor_sync(0xffffffffu, lt, 2); eq += __shfl_xor_sync(0xffffffffu, eq, 2);
                uint32_t rank = lt;
                if (eq > 1) {   // rare: several heads share the high word
                    rank = 0;
                    const uint32_t q4 = (G + 3) >> 2, j0 = part * q4, j1 = j0 + q4 < G ? j0 + q4 : G;
                    for (uint32_t j = j0; j < j1; j++) rank += better(hk[j], hi[j], mk, mi) ? 1u : 0u;
                    rank += __shfl_xor_sync(0xfu << (threadIdx.x & 28u), rank, 1);
                    rank += __shfl_xor_sync(0xfu << (threadIdx.x & 28u), rank, 2);
                }
                if (h < G && part == 0 && mi != kIdxSentinel && rank == k - 1) { s_tkey = mk; s_tidx = mi; }
                // fewer than k heads: T stays +inf and every candidate survives
            }
            __syncthreads();
            const uint64_t tk = s_tkey; const uint32_t ti = s_tidx;
#pragma unroll
            for (int j = 0; j < kPerThread; j++) {
                if (ci[j] != kIdxSentinel && !better(tk, ti, ck[j], ci[j])) {   // e <= T
                    const uint32_t pos = atomicAdd(&s_cnt, 1u);
                    if (pos < 64) { sk[pos] = ck[j]; si[pos] = ci[j]; }
                    // the record of a winner needs three more words of its row; the scan has pushed them out of
                    // L2 long ago, so start fetching while the sort runs
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(a.flags + ci[j]));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(a.epoch + ci[j]));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(a.bloom + ci[j]));
                }
            }
            __syncthreads();
            const uint32_t cnt = s_cnt;
            if (cnt <= 64) {
                done = true;
                if (warp == 0) {
                    if (lane < cnt) { key = sk[lane]; idx = si[lane]; }
                    warp_sort32_rolled(key, idx, lane, cnt >= 32 ? 32u : next_pow2(cnt < 2 ? 2 : cnt));
                    if (cnt > 32) {
                        uint64_t k2 = kKeySentinel; uint32_t i2 = kIdxSentinel;
                        if (32 + lane < cnt) { k2 = sk[32 + lane]; i2 = si[32 + lane]; }
                        warp_sort32_rolled(k2, i2, lane);
                        warp_merge32_rolled(key, idx, k2, i2, lane);
                    }
                }
            }
            __syncthreads();   // s_cnt / the survivor area may be rewritten by the next query
        }
        if (a.dbg && threadIdx.x == 0 && qi == 0) a.dbg[blockIdx.x * 8 + 6] = gtime_ns();
        if (!done) {
            // general fold: each warp merges a strided subset of the lists, then a tree merge over the warps
            key = kKeySentinel; idx = kIdxSentinel;
            if (warp < NW) {
                uint64_t *wk = fk + warp * kFoldBatch * 32;
                uint32_t *wi = fi + warp * kFoldBatch * 32;
#pragma unroll 1
                for (uint32_t c0 = warp; c0 < G; c0 += NW * kFoldBatch) {
                    uint64_t bk[kFoldBatch]; uint32_t bi[kFoldBatch];
#pragma unroll
                    for (int j = 0; j < kFoldBatch; j++) {  // all loads of the round in flight together
                        const uint32_t c = c0 + j * NW;
                        bk[j] = kKeySentinel; bi[j] = kIdxSentinel;
                        if (c < G && lane < k) {
                            const int4 raw = __ldcg(reinterpret_cast<const int4 *>(lists + (size_t)c * k + lane));
                            bk[j] = ((uint64_t)(uint32_t)raw.y << 32) | (uint32_t)raw.x;
                            bi[j] = (uint32_t)raw.z;
                        }
                    }
#pragma unroll
                    for (int j = 0; j < kFoldBatch; j++) { wk[j * 32 + lane] = bk[j]; wi[j * 32 + lane] = bi[j]; }
                    __syncwarp();
                    const uint32_t left = (G - c0 + NW - 1) / NW;
                    const uint32_t nl = left < (uint32_t)kFoldBatch ? left : (uint32_t)kFoldBatch;
#pragma unroll 1
                    for (uint32_t j = 0; j < nl; j++) warp_merge32_rolled(key, idx, wk[j * 32 + lane], wi[j * 32 + lane], lane);
                    __syncwarp();
                }
            }
            cta_tree_merge32_rolled<NW>(key, idx, xk, xi);
        }
        if (a.dbg && threadIdx.x == 0 && qi == 0) a.dbg[blockIdx.x * 8 + 7] = gtime_ns();
        if (warp == 0) {
            const unsigned valid = __ballot_sync(0xffffffffu, lane < k && idx != kIdxSentinel);
            if (lane == 0) {
                a.out_count[(size_t)qi * a.count_stride] = __popc(valid);
                if (a.total) a.total[qi] = *reinterpret_cast<volatile uint32_t *>(&a.counters[2 + qi]);
            }
            if (lane < k && idx != kIdxSentinel) {
                spl_gpu_result r;
                split_key(a.order, key, r.similarity, r.distance);
                r.slot_idx = (uint32_t)(a.row0 + idx);
                r.has_embedding = (a.flags[idx] & SPL_ROW_EMB) ? 1u : 0u;
                r.epoch = a.epoch[idx];
                r.bloom = a.bloom[idx];
                a.out[(size_t)qi * a.out_stride + lane] = r;
            }
        }
    }
    if (warp == 0) {
        __syncwarp();
        uint32_t last = 1;
        if (F > 1) {   // result writes of this CTA, then its place among the finishers
            __threadfence();
            if (lane == 0) last = atomicAdd(&a.counters[12], 1u) == F - 1 ? 1u : 0u;
            last = __shfl_sync(0xffffffffu, last, 0);
        }
        if (last) {
            if (F > 1) __threadfence();
            if (lane < 2 + nq || lane == 12) a.counters[lane] = 0;
            if (lane == 0) *s_flag = 1;
        }
    }
    __syncthreads();
}

// workspace the final merge may use: `ws_entries` (power of two, >= 2*k) entries at ws_keys/ws_idx
template <int NW, bool SMALL = true>
__device__ __forceinline__ void scan_epilogue(const ScanArgs &a, uint64_t *skeys, uint32_t *sidx, uint64_t *ws_keys,
                                              uint32_t *ws_idx, uint32_t ws_entries, uint32_t *s_flag, uint32_t qi = 0) {
    // qi: which query of a small batch this call finishes (skeys/sidx already point at that query's lists).  Each
    // query has its own published-list area, ticket counter (1, then 9, 10, 11), survivor counter (2 + qi) and outputs.
    const uint32_t k = a.k;
    const uint32_t tix = qi == 0 ? 1u : 8u + qi;
    Cand *cands_q = a.cta_cands + (size_t)qi * gridDim.x * k;
    if constexpr (SMALL) {
        if (k <= 32) { scan_epilogue_small<NW>(a, skeys, sidx, s_flag); return; }  // registers + shuffles
    }
    const uint32_t P = next_pow2(NW * k);  // the lists region is padded to P entries (sentinel-filled)
    __syncthreads();
    bitonic_sort_cta(skeys, sidx, P);
    {
        Cand *dst = cands_q + (size_t)blockIdx.x * k;
        for (uint32_t p = threadIdx.x; p < k; p += blockDim.x) {
            Cand c;
            c.key = skeys[p]; c.idx = sidx[p]; c.pad = 0;
            dst[p] = c;
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t ticket = atomicAdd(&a.counters[tix], 1u);
        *s_flag = (ticket == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (*s_flag == 0) return;

    // ---- last CTA ----
    __threadfence();
    const uint32_t G = gridDim.x, total = G * k;
    // Bounded selection first (cf. scan_epilogue_compact).  Every published list is sorted; take its entry at
    // position p as an ANCHOR and let T be the s-th best anchor, s = ceil(k / (p+1)): s lists hold p+1 entries at
    // least as good as T each, i.e. at least k candidates are <= T, so nothing worse than T can be among the k
    // best.  With the best rows spread over the CTAs only a few hundred of the G*k candidates survive, and ONE
    // sort of them replaces sorting all G*k (a 2^20-row top-100 spent 240 us there, three times the scan itself).
    bool done = false;
    {
        __shared__ uint32_t s_cnt2, s_ti2;
        __shared__ uint64_t s_tk2;
        const uint32_t p = k / 100, s = (k + p) / (p + 1);   // p = 0 up to k = 99, 1 up to 199, 2 up to 256
        if (G <= ws_entries && s <= G) {   // (uniform)
            if (threadIdx.x == 0) { s_cnt2 = 0; s_tk2 = kKeySentinel; s_ti2 = kIdxSentinel; }
            for (uint32_t c = threadIdx.x; c < G; c += blockDim.x) {
                const int4 raw = __ldcg(reinterpret_cast<const int4 *>(cands_q + (size_t)c * k + p));
                ws_keys[c] = ((uint64_t)(uint32_t)raw.y << 32) | (uint32_t)raw.x;
                ws_idx[c] = (uint32_t)raw.z;
            }
            __syncthreads();
            for (uint32_t c = threadIdx.x; c < G; c += blockDim.x) {   // rank of my anchor; ranks are distinct
                const uint64_t mk = ws_keys[c]; const uint32_t mi = ws_idx[c];
                if (mi == kIdxSentinel) continue;
                uint32_t rank = 0;
#pragma unroll 4
                for (uint32_t j = 0; j < G; j++) rank += better(ws_keys[j], ws_idx[j], mk, mi) ? 1u : 0u;
                if (rank == s - 1) { s_tk2 = mk; s_ti2 = mi; }   // fewer than s anchors: T stays +inf
            }
            __syncthreads();   // anchors are dead from here on: the workspace now collects the survivors
            const uint64_t tk = s_tk2; const uint32_t ti = s_ti2;
#pragma unroll 4
            for (uint32_t i = threadIdx.x; i < total; i += blockDim.x) {
                const int4 raw = __ldcg(reinterpret_cast<const int4 *>(cands_q + i));
                const uint64_t ek = ((uint64_t)(uint32_t)raw.y << 32) | (uint32_t)raw.x;
                const uint32_t ei = (uint32_t)raw.z;
                if (ei != kIdxSentinel && !better(tk, ti, ek, ei)) {   // e <= T
                    const uint32_t pos = atomicAdd(&s_cnt2, 1u);
                    if (pos < ws_entries) { ws_keys[pos] = ek; ws_idx[pos] = ei; }
                }
            }
            __syncthreads();
            const uint32_t cnt = s_cnt2;
            if (cnt <= ws_entries) {
                uint32_t lo = cnt > k ? cnt : k;
                const uint32_t W2 = next_pow2(lo < 2 ? 2 : lo);
                for (uint32_t q = cnt + threadIdx.x; q < W2; q += blockDim.x) { ws_keys[q] = kKeySentinel; ws_idx[q] = kIdxSentinel; }
                __syncthreads();
                bitonic_sort_cta(ws_keys, ws_idx, W2);
                done = true;
            }
            __syncthreads();
        }
    }
    // general path: running top-k over the G * k published candidates, a workspace-full at a time
    const uint32_t need = next_pow2(total + k);
    const uint32_t W = need < ws_entries ? need : ws_entries;  // workspace actually sorted
    if (!done)
        for (uint32_t p = threadIdx.x; p < k; p += blockDim.x) { ws_keys[p] = kKeySentinel; ws_idx[p] = kIdxSentinel; }
    for (uint32_t base = done ? total : 0u; base < total; base += W - k) {
        for (uint32_t p = k + threadIdx.x; p < W; p += blockDim.x) {
            const uint32_t src = base + (p - k);
            uint64_t key = kKeySentinel; uint32_t id = kIdxSentinel;
            if (src < total) {
                const int4 raw = __ldcg(reinterpret_cast<const int4 *>(cands_q + src));
                key = ((uint64_t)(uint32_t)raw.y << 32) | (uint32_t)raw.x;
                id = (uint32_t)raw.z;
            }
            ws_keys[p] = key; ws_idx[p] = id;
        }
        __syncthreads();
        bitonic_sort_cta(ws_keys, ws_idx, W);  // running top-k stays in [0, k)
    }
    if (threadIdx.x == 0) {
        if (a.total) a.total[qi] = *reinterpret_cast<volatile uint32_t *>(&a.counters[2 + qi]);
        if (qi == 0) a.counters[0] = 0;   // every CTA is past its scan once any ticket has been taken
        a.counters[tix] = 0; a.counters[2 + qi] = 0;
    }
    uint32_t cnt = 0;
    for (uint32_t t = threadIdx.x; t < k; t += blockDim.x) {
        const uint32_t idx = ws_idx[t];
        if (idx == kIdxSentinel) continue;
        spl_gpu_result r;
        split_key(a.order, ws_keys[t], r.similarity, r.distance);
        r.slot_idx = (uint32_t)(a.row0 + idx);
        r.has_embedding = (a.flags[idx] & SPL_ROW_EMB) ? 1u : 0u;
        r.epoch = a.epoch[idx];
        r.bloom = a.bloom[idx];
        a.out[(size_t)qi * a.out_stride + t] = r;
        // survivors are a prefix of the sorted array: the last valid one knows the count
        if (t + 1 == k || ws_idx[t + 1] == kIdxSentinel) cnt = t + 1;
    }
    if (cnt) a.out_count[(size_t)qi * a.count_stride] = cnt;
    else if (threadIdx.x == 0 && ws_idx[0] == kIdxSentinel) a.out_count[(size_t)qi * a.count_stride] = 0;
    __syncthreads();   // the workspace and *s_flag are reused by the next query of a batch
}

// sentinel-fill the padded lists region; every scan kernel calls this before its main loop
__device__ __forceinline__ void topk_lists_init(uint64_t *skeys, uint32_t *sidx, uint32_t entries) {
    for (uint32_t p = threadIdx.x; p < entries; p += blockDim.x) { skeys[p] = kKeySentinel; sidx[p] = kIdxSentinel; }
}

}  // namespace spl
