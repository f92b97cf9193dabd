// spl_exact.cu -- SPL_MODE_EXACT: scores bit-identical to the reference's CPU search.
//
// The reference sums 768 products one after the other in fp32, without FMA
// (/root/reference/splinter_cli_cmd_search.c:43-62; built -O3, no -ffast-math, no -march:
// CMakeLists.txt:78).  Any tree reduction differs from that in the last bits, which is enough to
// swap near-ties.  EXACT reproduces the reference's rounding step by step:
//
//   fused (k <= 128)  the fast fused scan (spl_scan_tma.cu) selects k' > k candidates; then
//       exact_rescore_kernel recomputes both scores of every candidate in the reference's order
//       (__fmul_rn / __fadd_rn, one sequential chain per accumulator, IEEE sqrt and divide), sorts by
//       the reference comparator and emits k records.  It also PROVES the selection: every row the
//       fast scan left out has a fast primary score no better than the k'-th candidate's, and fast
//       and sequential scores differ by at most a stated bound, so if the k-th exact score clears
//       the k'-th fast score by that bound no outsider can belong to the exact top-k.  If the proof
//       fails (a crowd of near-ties) the host retries with k' = 256 and finally falls back to
//   full   exact_all_kernel: one thread per row, the same sequential arithmetic straight from HBM,
//       one key per surviving row, device-wide sort (spl_sortall.cu).  Also serves k = 0 / k > 128.
#include <float.h>

#include "spl_exact.cuh"
#include "spl_scan_epilogue.cuh"

namespace spl {

// the reference's two loops, rounding for rounding
struct ExactSums { float dot, ma, mb, ss; };

__device__ __forceinline__ float ref_cosine(float dot, float ma, float mb) {
    const float den = __fmul_rn(__fsqrt_rn(ma), __fsqrt_rn(mb));  // sqrtf(mag_a) * sqrtf(mag_b)
    return den < 1e-9f ? 0.0f : __fdiv_rn(dot, den);
}

// ---- fused path: one CTA (one warp) per candidate ------------------------------------------------
// lanes 0..3 each own one of the four accumulators the reference keeps (dot, mag_a, mag_b, sum) and
// walk the row in index order from shared memory.
__global__ void __launch_bounds__(32) exact_rescore_kernel(const ExactArgs a) {
    extern __shared__ __align__(16) float srow[];  // [dim] row, [dim] query
    __shared__ uint32_t s_last;
    const uint32_t lane = threadIdx.x, c = blockIdx.x;
    const uint32_t ncand = *a.cand_count < a.kprime ? *a.cand_count : a.kprime;
    float *sq = srow + a.dim;
    if (c < ncand) {
        const spl_gpu_result cand = a.cand[c];
        const uint32_t row = (uint32_t)(cand.slot_idx - a.row0);
        const float4 *src = reinterpret_cast<const float4 *>(a.E + (size_t)row * a.dim);
        const float4 *q4 = reinterpret_cast<const float4 *>(a.q);
        for (uint32_t i = lane; i < a.dim / 4; i += 32) {
            reinterpret_cast<float4 *>(srow)[i] = src[i];
            reinterpret_cast<float4 *>(sq)[i] = q4[i];
        }
        __syncwarp();
        // lanes 0..3: dot, mag_a, mag_b, sum -- same instruction stream, operands picked by lane (no divergence)
        float acc = 0.0f;
        const float4 *sq4 = reinterpret_cast<const float4 *>(sq), *sr4 = reinterpret_cast<const float4 *>(srow);
        const bool l1 = lane == 1, l2 = lane == 2, l3 = lane == 3;
#pragma unroll 4
        for (uint32_t i = 0; i < a.dim / 4; i++) {
            const float4 xq = sq4[i], yr = sr4[i];   // a = query, b = slot vector (_search.c:395-396)
            const float xs[4] = { xq.x, xq.y, xq.z, xq.w }, ys[4] = { yr.x, yr.y, yr.z, yr.w };
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float x = xs[j], y = ys[j];
                const float diff = __fsub_rn(x, y);
                const float u = l3 ? diff : (l2 ? y : x);
                const float v = l3 ? diff : (l1 ? x : y);
                acc = __fadd_rn(acc, __fmul_rn(u, v));   // strictly in index order, no FMA
            }
        }
        const float dot = __shfl_sync(0xffffffffu, acc, 0), ma = __shfl_sync(0xffffffffu, acc, 1);
        const float mb = __shfl_sync(0xffffffffu, acc, 2), ss = __shfl_sync(0xffffffffu, acc, 3);
        if (lane == 0) {
            float sim, dist;
            if (cand.has_embedding) { sim = ref_cosine(dot, ma, mb); dist = __fsqrt_rn(ss); }
            else { sim = 0.0f; dist = FLT_MAX; }   // _search.c:388-391
            a.exact_key[c] = make_key(a.order, sim, dist);
        }
    }
    // ---- last CTA: sort, prove the selection, emit ----
    __threadfence();
    __syncwarp();
    if (lane == 0) s_last = atomicAdd(a.ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncwarp();
    if (!s_last) return;
    __threadfence();
    if (lane == 0) *a.ticket = 0;
    // stage the candidates' slot indices and has_embedding flags in shared memory for the sort below
    __shared__ uint64_t s_key[kExactMaxCand];
    __shared__ uint32_t s_slot[kExactMaxCand], s_pos[kExactMaxCand];
    // thresholds apply to the exact scores (a candidate can sit on either side of one by an ulp)
    uint32_t kept = 0;
    for (uint32_t base = 0; base < ncand; base += 32) {
        const uint32_t i = base + lane;
        bool keep = false;
        uint64_t key = 0;
        if (i < ncand) {
            key = __ldcg(&a.exact_key[i]);
            float sim, dist;
            split_key(a.order, key, sim, dist);
            keep = true;
            if (a.cand[i].has_embedding) {
                if (a.min_sim > 0.f && sim < a.min_sim) keep = false;
                if (a.max_dist > 0.f && dist > a.max_dist) keep = false;
            } else if (a.min_sim > 0.f || a.max_dist > 0.f) {
                keep = false;
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const uint32_t pos = kept + __popc(m & ((1u << lane) - 1u));
            s_key[pos] = key;
            s_pos[pos] = i;
            s_slot[pos] = a.cand[i].slot_idx;
        }
        kept += __popc(m);
    }
    __syncwarp();
    // rank sort of <= 256 survivors by (exact key, slot index)
    const uint32_t nout = kept < a.k ? kept : a.k;
    uint64_t kth_key = 0;
    for (uint32_t base = 0; base < kept; base += 32) {
        const uint32_t i = base + lane;
        if (i < kept) {
            const uint64_t key = s_key[i];
            const uint32_t pos = s_pos[i], slot = s_slot[i];
            uint32_t rank = 0;
            for (uint32_t j = 0; j < kept; j++) rank += better(s_key[j], s_slot[j], key, slot) ? 1u : 0u;
            if (rank < nout) {
                spl_gpu_result r = a.cand[pos];
                split_key(a.order, key, r.similarity, r.distance);
                a.out[rank] = r;
            }
            if (rank + 1 == nout) a.kth_key_out[0] = key;
        }
    }
    __syncwarp();
    __threadfence();
    if (lane == 0) {
        *a.out_count = nout;
        // ---- proof that no row outside the candidate set belongs to the exact top-k ----
        uint32_t valid = 1;
        if (*a.cand_count >= a.kprime && nout > 0) {         // the fast scan may have left rows out
            kth_key = a.kth_key_out[0];
            float ks, kd, ws, wd;                               // exact k-th, fast k'-th (worst candidate)
            split_key(a.order, kth_key, ks, kd);
            const spl_gpu_result worst = a.cand[a.kprime - 1];  // fast list is sorted: last = worst
            ws = worst.similarity; wd = worst.distance;
            if (nout < a.k) valid = 0;                          // thresholds emptied part of the list: cannot prove
            else if (a.order == SPL_ORDER_COSINE) valid = (ks > ws + kExactSimBound) ? 1u : 0u;
            else valid = (kd < wd - kExactDistRel * wd) ? 1u : 0u;
        }
        *a.valid = valid;
        // device-buffer callers have only the count to go by: an unproven selection must not read as an answer
        if (!valid) *a.out_count = SPL_GPU_COUNT_UNANSWERED;
    }
}

// ---- full path: one thread per row, sequential arithmetic from HBM, one key per survivor -----------
__global__ void __launch_bounds__(128) exact_all_kernel(const ScanArgs a) {
    extern __shared__ __align__(16) float sq[];
    for (uint32_t i = threadIdx.x; i < a.dim; i += blockDim.x) sq[i] = a.q[i];
    __syncthreads();
    const uint32_t need = a.mask ? SPL_ROW_ENUM : SPL_ROW_LIST;
    const bool thresholds = a.min_sim > 0.f || a.max_dist > 0.f;
    for (uint32_t row = blockIdx.x * blockDim.x + threadIdx.x; row < a.n; row += gridDim.x * blockDim.x) {
        const uint32_t f = a.flags[row];
        if (!(f & need) || (a.bloom[row] & a.mask) != a.mask || row == a.exclude) continue;
        float sim, dist;
        if (f & SPL_ROW_EMB) {
            const float4 *e4 = reinterpret_cast<const float4 *>(a.E + (size_t)row * a.dim);
            float dot = 0.f, ma = 0.f, mb = 0.f, ss = 0.f;
            for (uint32_t i = 0; i < a.dim / 4; i++) {
                const float4 y4 = e4[i];
                const float yy[4] = { y4.x, y4.y, y4.z, y4.w };
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const float x = sq[4 * i + j], y = yy[j];
                    dot = __fadd_rn(dot, __fmul_rn(x, y));
                    ma = __fadd_rn(ma, __fmul_rn(x, x));
                    mb = __fadd_rn(mb, __fmul_rn(y, y));
                    const float diff = __fsub_rn(x, y);
                    ss = __fadd_rn(ss, __fmul_rn(diff, diff));
                }
            }
            sim = ref_cosine(dot, ma, mb);
            dist = __fsqrt_rn(ss);
            if (a.min_sim > 0.f && sim < a.min_sim) continue;
            if (a.max_dist > 0.f && dist > a.max_dist) continue;
        } else {
            if (thresholds) continue;
            sim = 0.f; dist = FLT_MAX;
        }
        a.keys_out[row] = make_key(a.order, sim, dist);
    }
}

cudaError_t launch_exact_rescore(const ExactArgs &a, cudaStream_t s) {
    exact_rescore_kernel<<<a.kprime, 32, (size_t)a.dim * 8, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_exact_all(const ScanArgs &a, int sm_count, cudaStream_t s) {
    exact_all_kernel<<<sm_count * 8, 128, (size_t)a.dim * 4, s>>>(a);
    return cudaGetLastError();
}

}  // namespace spl
