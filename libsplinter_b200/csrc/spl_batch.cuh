// spl_batch.cuh -- argument block + host entry points of the batched tensor-core search (spl_batch.cu)
#pragma once
#include <cuda_bf16.h>

#include "spl_device.cuh"

namespace spl {

constexpr uint32_t kBatchCandCap = 4096;  // per-query candidate list capacity (entries)
constexpr uint32_t kBatchMaxQ = 1024;     // queries per batched launch (larger requests are split)

struct BatchArgs {
    // mirror
    const float *E;               // fp32 rows (exact re-score)
    const __nv_bfloat16 *E16;     // bf16 shadow rows (tensor-core operand)
    const float *norm2;           // |bf16(e)|^2
    const float *err2;            // |e - bf16(e)|^2 (selection proof)
    const uint32_t *norm2_max;    // float bits of an upper bound of max |bf16(e)|^2 over the shard
    const uint64_t *bloom;
    const uint8_t *flags;
    const uint64_t *epoch;
    uint32_t n, dim;
    // request
    uint32_t nq, nq_pad, n_qtiles, order;
    uint64_t mask;
    uint32_t exclude;             // local row or kIdxSentinel
    uint64_t row0;
    float min_sim, max_dist;      // threshold on the ORDERING score only (cosine: min_sim, L2: max_dist); 0 = none.
                                  // Rows failing it are the tail of the sorted result, so it is applied after the re-score.
    // schedule
    uint32_t pass;                // 1: per-tile minima, 2: append below threshold
    uint32_t pair_q_units;        // pair kernel: depth of the query ring in 16 KB units (0 = default)
    uint32_t tile_first, tile_stride, n_tiles;
    // scratch
    float *thr;                   // [nq_pad]   bound on t ...
    uint32_t *thr_row;            // [nq_pad]   ... and, for ties on t, on the row index
    float *tile_min;              // [n_tiles][nq_pad]
    uint32_t *tile_min_row;       // [n_tiles][nq_pad]
    uint32_t *out_min;            // [nq_pad]   f2ord of the best ordering value pass 2 kept OUT of the list as worse than thr
    uint32_t *out_tie;            // [nq_pad]   bit 0 / 1: exactly-scoring / ordinary rows tying with thr were cut off by row index
    const uint32_t *exact_bound;  // [2]        row index of the k'-th zero-vector / no-embedding candidate (kIdxSentinel: fewer)
    uint32_t *exact_bound_out;    //            (the same words, written by exact_bound_block between the passes)
    uint32_t *tile_kinds;         // [sampled tiles][2]  pass 1: candidates of either exactly-scoring kind per tile
    uint32_t *cand_cnt;           // [nq_pad]
    float *cand_t;                // [nq_pad][cand_cap]
    uint32_t *cand_idx;           // [nq_pad][cand_cap]
    uint32_t cand_cap;
};

struct BatchScratchLayout {
    size_t q16_off, thr_off, thr_row_off, cnt_off, ovf_off, out_min_off, out_tie_off, exact_off, tile_min_off, tile_min_row_off, cand_t_off, cand_idx_off;
    uint32_t nq_pad;
};

cudaError_t batch_prepare();  // once per device
bool batch_supported(uint32_t dim, uint32_t k);
size_t batch_scratch_bytes(uint32_t n_rows, uint32_t nq_max, uint32_t dim, BatchScratchLayout *L);
// enqueues: query prep, sampling pass, thresholds, full pass, select + exact re-rank with its selection proof.
// d_counts[q] = SPL_GPU_COUNT_UNANSWERED and bit 0 of the flag array returned in *d_overflow_out mark the queries
// the bf16 pass could not answer (list overflow / unprovable selection); the other bits of a flag word hold the
// number of rows re-scored beyond k' (<< 1).
cudaError_t batch_search(const BatchArgs &base, const BatchScratchLayout &L, unsigned char *scratch, const float *d_q32,
                         uint32_t k, int sm_count, spl_gpu_result *d_out, uint32_t out_stride, uint32_t *d_counts,
                         uint32_t **d_overflow_out, uint32_t *launches, cudaStream_t s);

}  // namespace spl
