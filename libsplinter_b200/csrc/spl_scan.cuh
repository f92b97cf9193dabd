// spl_scan.cuh -- argument block shared by the single-query scan kernels and their launchers.
#pragma once
#include "spl_device.cuh"

namespace spl {

struct ScanArgs {
    // mirror (device)
    const float *E;          // [n][dim] fp32, row pitch = dim*4 bytes
    const uint64_t *bloom;   // [n]
    const uint8_t *flags;    // [n] SPL_ROW_*
    const uint64_t *epoch;   // [n]
    // query
    const float *q;          // [dim]
    uint32_t n, dim, k, order;
    uint64_t mask;
    float min_sim, max_dist;
    uint32_t exclude;        // local row index to skip, kIdxSentinel = none
    uint64_t row0;           // global slot index of local row 0
    // scratch (device): one Cand[k] list per CTA + {work counter, done counter}
    Cand *cta_cands;
    uint32_t *counters;
    // output (device)
    spl_gpu_result *out;
    uint32_t *out_count;
    // unlimited / large-k path: when non-null the kernel writes one key per surviving row here
    // (pre-filled with kKeySentinel by the caller) instead of keeping a top-k
    uint64_t *keys_out;
    uint32_t *total;         // nullable: where the LAST CTA stores the number of rows that passed predicate and
                             // thresholds (accumulated in counters[2]; may point to mapped host memory)
    uint32_t nq;             // TMA kernel: queries scored per pass (0/1, 2 or 4); q holds nq vectors back to back,
    uint32_t out_stride;     //   query i writes out + i*out_stride, out_count[i*count_stride], total[i] (counters[2+i])
    uint32_t count_stride;
    uint32_t chunk_groups;   // TMA kernel: 32-slot groups handed out per work-counter increment (1..8)
    uint32_t *done_seq_ptr;  // TMA kernel, nullable: (mapped host) word the LAST CTA sets to done_seq after every result of the
    uint32_t done_seq;       //   launch is written -- lets the host-buffer call poll for its answer instead of synchronising the stream
    uint32_t overlap_prev;   // TMA kernel, host side only: launch so that it may start during the previous scan's tail
    // developer timeline (nullable): per CTA {start, producers done, consumers done, end} in ns
    unsigned long long *dbg;
};

constexpr int kScanThreads = 256;
constexpr int kScanWarps = kScanThreads / 32;

__host__ __device__ inline uint32_t topk_pow2(uint32_t x) {
    uint32_t p = 1;
    while (p < x) p <<= 1;
    return p;
}
// entries / bytes of the shared-memory region holding the per-warp top-k lists (padded to a power of
// two so the epilogue can sort it in place): u64 keys[entries] followed by u32 idx[entries]
__host__ __device__ inline uint32_t scan_topk_entries(uint32_t k, int nwarps = kScanWarps) { return topk_pow2((uint32_t)nwarps * k); }
__host__ __device__ inline size_t scan_topk_smem(uint32_t k, int nwarps = kScanWarps) {
    return (size_t)scan_topk_entries(k, nwarps) * (sizeof(uint64_t) + sizeof(uint32_t));
}
// the register-path kernel reuses its lists region as the final-merge workspace: at least 2048 entries
__host__ __device__ inline uint32_t scan_fast_ws_entries(uint32_t k) {
    const uint32_t p = scan_topk_entries(k);
    return p > 2048u ? p : 2048u;
}

cudaError_t launch_scan_fast(const ScanArgs &a, int grid, cudaStream_t stream);
int scan_fast_ctas_per_sm(uint32_t dim, uint32_t k);

// TMA-staged variant (spl_scan_tma.cu): one persistent CTA per SM
bool scan_tma_supported(uint32_t dim, uint32_t k);
bool scan_tma_multi_supported(uint32_t dim, uint32_t k, uint32_t nq);
cudaError_t scan_tma_prepare();
cudaError_t launch_scan_tma(const ScanArgs &a, int grid, cudaStream_t stream);
// one query handed over inside the launch (kernel parameter space) instead of through a device buffer
bool scan_tma_qparam_supported(uint32_t dim, uint32_t k);
cudaError_t launch_scan_tma_qparam(const ScanArgs &a, const float *q_host, int grid, cudaStream_t stream);

cudaError_t merge_prepare();
// merge nlists lists of Cand-like results (spl_gpu_result records) into the top-k; shard_pitch (records) / count_pitch
// (uint32s) = distance between consecutive shards' blocks, 0 = two separate contiguous arrays
cudaError_t launch_merge_results(const spl_gpu_result *lists, const uint32_t *counts, uint32_t nlists, uint32_t nq,
                                 uint32_t stride, uint32_t order, uint32_t k, spl_gpu_result *out, uint32_t out_stride,
                                 uint32_t *out_count, cudaStream_t stream, uint32_t shard_pitch = 0, uint32_t count_pitch = 0);

// unlimited / k > SPL_GPU_KMAX_FUSED: compact the rows holding a key, then sort only those (key, row) pairs and emit
// records.  keys[n] holds the scan's output; the workspace is caller-allocated (see sortall_tmp_bytes).
size_t sortall_tmp_bytes(uint32_t n);
cudaError_t launch_sortall_select(const ScanArgs &a, const uint64_t *keys, void *tmp, size_t tmp_bytes, uint32_t *d_total,
                                  cudaStream_t stream);
cudaError_t launch_sortall_sort(const ScanArgs &a, const uint64_t *keys, void *tmp, size_t tmp_bytes, uint32_t total,
                                uint32_t out_cap, cudaStream_t stream);

}  // namespace spl
