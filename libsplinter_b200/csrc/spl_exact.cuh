// spl_exact.cuh -- SPL_MODE_EXACT plumbing (spl_exact.cu)
#pragma once
#include "spl_scan.cuh"

namespace spl {

// How far a fast (FMA, tree-reduced) score can sit from the reference's sequential fp32 score.
// Worst-case forward error of a length-n sequential fp32 sum is n*2^-24 relative to sum|terms|
// (768 * 6e-8 = 4.6e-5; dim 4096: 2.4e-4), the tree sum's is far smaller; the cosine adds the two norms'
// errors.  1e-3 covers every supported dimension with margin and is still far below the spacing of
// the top-k scores of a real corpus, so the proof in exact_rescore_kernel almost always succeeds.
constexpr float kExactSimBound = 1e-3f;   // absolute, cosine similarity
constexpr float kExactDistRel = 1e-3f;    // relative, euclidean distance
constexpr uint32_t kExactMaxCand = 256;   // k' ceiling (= SPL_GPU_KMAX_FUSED)
constexpr uint32_t kExactMaxK = 128;      // fused exact path serves k <= 128

struct ExactArgs {
    const float *E;
    const float *q;
    uint32_t dim, order, k, kprime;
    uint64_t row0;
    float min_sim, max_dist;
    // candidates written by the fast scan (sorted by fast key) + how many
    const spl_gpu_result *cand;
    const uint32_t *cand_count;
    // scratch
    uint64_t *exact_key;   // [kprime]
    uint64_t *sort_key;    // [kprime]
    uint32_t *sort_pos;    // [kprime]
    uint64_t *kth_key_out; // [1]
    uint32_t *ticket;      // [1], zero before launch, re-armed by the kernel
    // results
    spl_gpu_result *out;
    uint32_t *out_count;
    uint32_t *valid;       // 1 = selection proven, 0 = retry with more candidates / full pass
};

cudaError_t launch_exact_rescore(const ExactArgs &a, cudaStream_t s);
cudaError_t launch_exact_all(const ScanArgs &a, int sm_count, cudaStream_t s);

}  // namespace spl
