// spl_scan.cu -- single-query fused scan (SPL_MODE_FAST) for sm_100a.
//
// One launch does, per query, everything the reference's cmd_search does after it has its
// query vector (/root/reference/splinter_cli_cmd_search.c:336-416):
//   candidate predicate   splinter_enumerate_matches (splinter.c:977-989) / splinter_list
//                         (splinter.c:466-477) + the epoch!=0 / key / scratch-key skips
//                         (_search.c:365-375), evaluated from 9 bytes per slot (bloom u64 +
//                         flags u8) BEFORE any embedding byte of the slot is requested;
//   scoring               cosine_similarity + euclidean_distance (_search.c:43-62) fused in one
//                         pass over the row: dot, |e|^2 and sum (q-e)^2 from 128-bit coalesced
//                         streaming loads, per-lane partial sums, warp-shuffle tree;
//   thresholds            --similarity / --distance (_search.c:398-399);
//   sort + limit          compare_results + qsort + limit (_search.c:64-72,412-416) as a
//                         per-warp bounded list -> per-CTA list -> last-CTA merge + sort.
//
// HBM-bound by design: algorithmic bytes per query = 9*n + 4*dim*M (M = rows passing the
// predicate) + 4*dim.  Rows are handed out dynamically in groups of 32 slots so a selective
// mask does not unbalance the warps.
#include "spl_scan.cuh"
#include "spl_scan_epilogue.cuh"

#include <float.h>

namespace spl {

// ---------------------------------------------------------------------------------------------
// SPL_MODE_FAST: one warp per row, V float4 per lane (dim = 128*V), two rows in flight.
// ---------------------------------------------------------------------------------------------
template <int V, bool KEYS>
__global__ void __launch_bounds__(kScanThreads, 3) scan_fast_kernel(const ScanArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_flag;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // lists region doubles as the final-merge workspace: ws entries, keys then idx
    const uint32_t ws = KEYS ? 0u : scan_fast_ws_entries(a.k);
    uint64_t *skeys = reinterpret_cast<uint64_t *>(smem_raw);
    uint32_t *sidx = reinterpret_cast<uint32_t *>(skeys + ws);
    WarpTopK W;
    W.init(skeys + warp * a.k, sidx + warp * a.k, a.k);
    if (!KEYS) { topk_lists_init(skeys, sidx, ws); __syncthreads(); }

    // query -> registers; |q|^2 once per warp
    const float4 *q4 = reinterpret_cast<const float4 *>(a.q);
    float4 qv[V];
    float ma = 0.f;
#pragma unroll
    for (int j = 0; j < V; j++) {
        qv[j] = q4[j * 32 + lane];
        ma = fmaf(qv[j].x, qv[j].x, ma); ma = fmaf(qv[j].y, qv[j].y, ma);
        ma = fmaf(qv[j].z, qv[j].z, ma); ma = fmaf(qv[j].w, qv[j].w, ma);
    }
    ma = warp_sum(ma);

    const uint32_t need = a.mask ? SPL_ROW_ENUM : SPL_ROW_LIST;
    const uint32_t ngroups = (a.n + 31u) >> 5;
    const float4 *E4 = reinterpret_cast<const float4 *>(a.E);
    const bool thresholds = a.min_sim > 0.f || a.max_dist > 0.f;

    auto fetch_group = [&]() -> uint32_t {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(&a.counters[0], 1u);
        return __shfl_sync(0xffffffffu, g, 0);
    };
    auto load_pred = [&](uint32_t g, uint32_t &f, uint64_t &b) {
        const uint32_t row = (g << 5) + lane;
        f = 0; b = 0;
        if (g < ngroups && row < a.n) { f = a.flags[row]; b = a.bloom[row]; }
    };
    uint32_t nsurv = 0;
    auto finish_row = [&](uint32_t row, uint32_t f, float dot, float mb, float ss) {
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
        if (KEYS) {  // unlimited / large-k path: publish the key, a device-wide sort follows
            if (lane == 0) a.keys_out[row] = key;
        } else if (W.admits(key, row)) {
            W.insert(key, row, lane);
        }
    };

    uint32_t g = fetch_group(), f; uint64_t b;
    load_pred(g, f, b);
    while (g < ngroups) {
        const uint32_t gn = fetch_group();
        const uint32_t row_l = (g << 5) + lane;
        const bool pass = (f & need) && ((b & a.mask) == a.mask) && row_l != a.exclude;
        unsigned m = __ballot_sync(0xffffffffu, pass);
        uint32_t fn; uint64_t bn;
        bool prefetched = false;
        if (m == 0) { load_pred(gn, fn, bn); prefetched = true; }
        while (m) {
            const int l0 = __ffs(m) - 1; m &= m - 1;
            const bool two = m != 0;
            const int l1 = two ? __ffs(m) - 1 : l0;
            if (two) m &= m - 1;
            const uint32_t r0 = (g << 5) + l0, r1 = (g << 5) + l1;
            const float4 *p0 = E4 + (size_t)r0 * (V * 32) + lane;
            const float4 *p1 = E4 + (size_t)r1 * (V * 32) + lane;
            float4 x[V], y[V];
#pragma unroll
            for (int j = 0; j < V; j++) x[j] = ldg_stream(p0 + 32 * j);
            if (two) {
#pragma unroll
                for (int j = 0; j < V; j++) y[j] = ldg_stream(p1 + 32 * j);
            }
            // the next group's predicate bytes ride behind the first row loads of this group
            if (!prefetched) { load_pred(gn, fn, bn); prefetched = true; }
            float d0 = 0.f, b0 = 0.f, s0 = 0.f, d1 = 0.f, b1 = 0.f, s1 = 0.f;
#pragma unroll
            for (int j = 0; j < V; j++) {
                float t;
                d0 = fmaf(qv[j].x, x[j].x, d0); b0 = fmaf(x[j].x, x[j].x, b0); t = qv[j].x - x[j].x; s0 = fmaf(t, t, s0);
                d0 = fmaf(qv[j].y, x[j].y, d0); b0 = fmaf(x[j].y, x[j].y, b0); t = qv[j].y - x[j].y; s0 = fmaf(t, t, s0);
                d0 = fmaf(qv[j].z, x[j].z, d0); b0 = fmaf(x[j].z, x[j].z, b0); t = qv[j].z - x[j].z; s0 = fmaf(t, t, s0);
                d0 = fmaf(qv[j].w, x[j].w, d0); b0 = fmaf(x[j].w, x[j].w, b0); t = qv[j].w - x[j].w; s0 = fmaf(t, t, s0);
            }
            if (two) {
#pragma unroll
                for (int j = 0; j < V; j++) {
                    float t;
                    d1 = fmaf(qv[j].x, y[j].x, d1); b1 = fmaf(y[j].x, y[j].x, b1); t = qv[j].x - y[j].x; s1 = fmaf(t, t, s1);
                    d1 = fmaf(qv[j].y, y[j].y, d1); b1 = fmaf(y[j].y, y[j].y, b1); t = qv[j].y - y[j].y; s1 = fmaf(t, t, s1);
                    d1 = fmaf(qv[j].z, y[j].z, d1); b1 = fmaf(y[j].z, y[j].z, b1); t = qv[j].z - y[j].z; s1 = fmaf(t, t, s1);
                    d1 = fmaf(qv[j].w, y[j].w, d1); b1 = fmaf(y[j].w, y[j].w, b1); t = qv[j].w - y[j].w; s1 = fmaf(t, t, s1);
                }
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                d0 += __shfl_xor_sync(0xffffffffu, d0, off);
                b0 += __shfl_xor_sync(0xffffffffu, b0, off);
                s0 += __shfl_xor_sync(0xffffffffu, s0, off);
                d1 += __shfl_xor_sync(0xffffffffu, d1, off);
                b1 += __shfl_xor_sync(0xffffffffu, b1, off);
                s1 += __shfl_xor_sync(0xffffffffu, s1, off);
            }
            const uint32_t f0 = __shfl_sync(0xffffffffu, f, l0), f1 = __shfl_sync(0xffffffffu, f, l1);
            finish_row(r0, f0, d0, b0, s0);
            if (two) finish_row(r1, f1, d1, b1, s1);
        }
        g = gn; f = fn; b = bn;
    }
    if (a.total && lane == 0 && nsurv) atomicAdd(&a.counters[2], nsurv);
    if (KEYS) {
        // re-arm the work counter once every CTA has drained it
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(&a.counters[1], 1u) == gridDim.x - 1) { a.counters[0] = 0; a.counters[1] = 0; a.counters[2] = 0; }
        }
        return;
    }
    scan_epilogue<kScanWarps>(a, skeys, sidx, skeys, sidx, ws, &s_flag);
}

// generic dimension (dim % 4 == 0): query in shared memory, strided float4 loop, one row at a time
template <bool KEYS>
__global__ void __launch_bounds__(kScanThreads, 3) scan_fast_generic_kernel(const ScanArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_flag;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t ws = KEYS ? 0u : scan_fast_ws_entries(a.k);
    uint64_t *skeys = reinterpret_cast<uint64_t *>(smem_raw);
    uint32_t *sidx = reinterpret_cast<uint32_t *>(skeys + ws);
    float4 *sq = reinterpret_cast<float4 *>(smem_raw + (((size_t)ws * 12 + 15) & ~(size_t)15));
    WarpTopK W;
    W.init(skeys + warp * a.k, sidx + warp * a.k, a.k);
    if (!KEYS) topk_lists_init(skeys, sidx, ws);
    const uint32_t nv = a.dim >> 2;
    for (uint32_t c = threadIdx.x; c < nv; c += blockDim.x) sq[c] = reinterpret_cast<const float4 *>(a.q)[c];
    __syncthreads();
    float ma = 0.f;
    for (uint32_t c = lane; c < nv; c += 32) {
        float4 v = sq[c];
        ma = fmaf(v.x, v.x, ma); ma = fmaf(v.y, v.y, ma); ma = fmaf(v.z, v.z, ma); ma = fmaf(v.w, v.w, ma);
    }
    ma = warp_sum(ma);
    const uint32_t need = a.mask ? SPL_ROW_ENUM : SPL_ROW_LIST;
    const uint32_t ngroups = (a.n + 31u) >> 5;
    const float4 *E4 = reinterpret_cast<const float4 *>(a.E);
    const bool thresholds = a.min_sim > 0.f || a.max_dist > 0.f;
    uint32_t nsurv = 0;
    for (;;) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(&a.counters[0], 1u);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= ngroups) break;
        const uint32_t row_l = (g << 5) + lane;
        uint32_t f = 0; uint64_t b = 0;
        if (row_l < a.n) { f = a.flags[row_l]; b = a.bloom[row_l]; }
        const bool pass = (f & need) && ((b & a.mask) == a.mask) && row_l != a.exclude;
        unsigned m = __ballot_sync(0xffffffffu, pass);
        while (m) {
            const int l0 = __ffs(m) - 1; m &= m - 1;
            const uint32_t r0 = (g << 5) + l0;
            const float4 *p0 = E4 + (size_t)r0 * nv;
            float d0 = 0.f, b0 = 0.f, s0 = 0.f;
            for (uint32_t c = lane; c < nv; c += 32) {
                const float4 x = ldg_stream(p0 + c), qq = sq[c];
                float t;
                d0 = fmaf(qq.x, x.x, d0); b0 = fmaf(x.x, x.x, b0); t = qq.x - x.x; s0 = fmaf(t, t, s0);
                d0 = fmaf(qq.y, x.y, d0); b0 = fmaf(x.y, x.y, b0); t = qq.y - x.y; s0 = fmaf(t, t, s0);
                d0 = fmaf(qq.z, x.z, d0); b0 = fmaf(x.z, x.z, b0); t = qq.z - x.z; s0 = fmaf(t, t, s0);
                d0 = fmaf(qq.w, x.w, d0); b0 = fmaf(x.w, x.w, b0); t = qq.w - x.w; s0 = fmaf(t, t, s0);
            }
            d0 = warp_sum(d0); b0 = warp_sum(b0); s0 = warp_sum(s0);
            const uint32_t f0 = __shfl_sync(0xffffffffu, f, l0);
            float sim, dist;
            bool keep = true;
            if (f0 & SPL_ROW_EMB) {
                sim = cosine_from_sums(d0, ma, b0);
                dist = __fsqrt_rn(s0);
                if (a.min_sim > 0.f && sim < a.min_sim) keep = false;
                if (a.max_dist > 0.f && dist > a.max_dist) keep = false;
            } else {
                keep = !thresholds; sim = 0.f; dist = FLT_MAX;
            }
            if (keep) {
                nsurv++;
                const uint64_t key = make_key(a.order, sim, dist);
                if (KEYS) {
                    if (lane == 0) a.keys_out[r0] = key;
                } else if (W.admits(key, r0)) {
                    W.insert(key, r0, lane);
                }
            }
        }
    }
    if (a.total && lane == 0 && nsurv) atomicAdd(&a.counters[2], nsurv);
    if (KEYS) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(&a.counters[1], 1u) == gridDim.x - 1) { a.counters[0] = 0; a.counters[1] = 0; a.counters[2] = 0; }
        }
        return;
    }
    scan_epilogue<kScanWarps>(a, skeys, sidx, skeys, sidx, ws, &s_flag);
}

int scan_fast_ctas_per_sm(uint32_t, uint32_t) { return 3; }

template <bool KEYS>
static cudaError_t launch_scan_fast_t(const ScanArgs &a, int grid, cudaStream_t stream) {
    const size_t smem = KEYS ? 0 : (size_t)scan_fast_ws_entries(a.k) * 12;
    if (a.dim == 768) {
        scan_fast_kernel<6, KEYS><<<grid, kScanThreads, smem, stream>>>(a);
    } else if (a.dim == 1024) {
        scan_fast_kernel<8, KEYS><<<grid, kScanThreads, smem, stream>>>(a);
    } else if (a.dim == 512) {
        scan_fast_kernel<4, KEYS><<<grid, kScanThreads, smem, stream>>>(a);
    } else if (a.dim == 256) {
        scan_fast_kernel<2, KEYS><<<grid, kScanThreads, smem, stream>>>(a);
    } else if (a.dim == 128) {
        scan_fast_kernel<1, KEYS><<<grid, kScanThreads, smem, stream>>>(a);
    } else {
        const size_t smem_g = ((smem + 15) & ~(size_t)15) + (size_t)a.dim * 4;
        scan_fast_generic_kernel<KEYS><<<grid, kScanThreads, smem_g, stream>>>(a);
    }
    return cudaGetLastError();
}

cudaError_t launch_scan_fast(const ScanArgs &a, int grid, cudaStream_t stream) {
    return a.keys_out ? launch_scan_fast_t<true>(a, grid, stream) : launch_scan_fast_t<false>(a, grid, stream);
}

// ---------------------------------------------------------------------------------------------
// merge of result-record lists (after the per-GPU all-gather): one CTA, one bitonic sort
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kMergeMaxEntries = 8192;  // nlists * stride candidates per call (96 KB of smem)

// One CTA per query.  Gathered layout: shard s holds its [query][stride] records at lists + s * shard_pitch (records) and
// its [query] counts at counts + s * count_pitch (uint32s): shard_pitch = nq * stride, count_pitch = nq for two separate
// contiguous arrays; the packed form (records then counts in ONE block per shard, so one transfer per shard suffices)
// uses the block size for both.  Output: out[query][out_stride], out_count[query].
__global__ void __launch_bounds__(1024, 1)
merge_results_kernel(const spl_gpu_result *lists, const uint32_t *counts, uint32_t nlists, uint32_t nq, uint32_t stride,
                     uint32_t shard_pitch, uint32_t count_pitch, uint32_t order, uint32_t k, uint32_t entries,
                     spl_gpu_result *out, uint32_t out_stride, uint32_t *out_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t *keys = reinterpret_cast<uint64_t *>(smem_raw);
    uint32_t *slot = reinterpret_cast<uint32_t *>(keys + entries);
    uint32_t *pos = slot + entries;
    const uint32_t q = blockIdx.x;
    out += (size_t)q * out_stride;
    out_count += q;
    // a shard that could not answer this query (SPL_GPU_COUNT_UNANSWERED) makes the merged answer unknowable:
    // hand the marker on instead of merging the shards that did answer into a wrong top-k
    __shared__ uint32_t s_unanswered;
    if (threadIdx.x == 0) s_unanswered = 0;
    __syncthreads();
    for (uint32_t l = threadIdx.x; l < nlists; l += blockDim.x)
        if (counts[(size_t)l * count_pitch + q] == SPL_GPU_COUNT_UNANSWERED) s_unanswered = 1;
    __syncthreads();
    if (s_unanswered) {
        if (threadIdx.x == 0) *out_count = SPL_GPU_COUNT_UNANSWERED;
        return;
    }
    for (uint32_t p = threadIdx.x; p < entries; p += blockDim.x) {
        const uint32_t l = p / stride, i = p - l * stride;
        uint64_t key = kKeySentinel; uint32_t id = kIdxSentinel;
        uint32_t src = 0;
        if (l < nlists && i < min(counts[(size_t)l * count_pitch + q], stride)) {
            src = (uint32_t)((size_t)l * shard_pitch + (size_t)q * stride + i);
            const spl_gpu_result r = lists[src];
            key = make_key(order, r.similarity, r.distance);
            id = r.slot_idx;  // ties between shards break on the GLOBAL slot index
        }
        keys[p] = key; slot[p] = id; pos[p] = src;
    }
    __syncthreads();
    bitonic_sort_cta(keys, slot, entries, pos);
    uint32_t cnt = 0;
    for (uint32_t t = threadIdx.x; t < k; t += blockDim.x) {
        if (slot[t] == kIdxSentinel) continue;
        out[t] = lists[pos[t]];
        if (t + 1 == k || slot[t + 1] == kIdxSentinel) cnt = t + 1;
    }
    if (cnt) *out_count = cnt;
    else if (threadIdx.x == 0 && slot[0] == kIdxSentinel) *out_count = 0;
}

// opt in to > 48 KB dynamic shared memory; function attributes are per device: call once per context
cudaError_t merge_prepare() {
    return cudaFuncSetAttribute(merge_results_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kMergeMaxEntries * 16));
}

cudaError_t launch_merge_results(const spl_gpu_result *lists, const uint32_t *counts, uint32_t nlists, uint32_t nq,
                                 uint32_t stride, uint32_t order, uint32_t k, spl_gpu_result *out, uint32_t out_stride,
                                 uint32_t *out_count, cudaStream_t stream, uint32_t shard_pitch, uint32_t count_pitch) {
    const uint32_t entries = topk_pow2(nlists * stride > k ? nlists * stride : k + 1);
    if (entries > kMergeMaxEntries || nq == 0) return cudaErrorInvalidValue;
    if (shard_pitch == 0) { shard_pitch = nq * stride; count_pitch = nq; }
    const size_t smem = (size_t)entries * 16;
    const int threads = entries >= 2048 ? 1024 : (entries >= 512 ? 256 : 128);
    merge_results_kernel<<<nq, threads, smem, stream>>>(lists, counts, nlists, nq, stride, shard_pitch, count_pitch, order, k,
                                                       entries, out, out_stride, out_count);
    return cudaGetLastError();
}

}  // namespace spl
