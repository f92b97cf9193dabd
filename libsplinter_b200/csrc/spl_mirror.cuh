// spl_mirror.cuh -- HBM mirror pointers and the staging record used by the sync path.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/splinter_gpu.h"

namespace spl {

struct MirrorPtrs {
    float *E;
    uint64_t *bloom;
    uint8_t *flags;
    uint64_t *epoch;
    __nv_bfloat16 *E16;  // nullable
    float *norm2;        // nullable
    uint32_t dim;
};

// one staged slot: local row index + the metadata snapshot taken inside the seqlock window
struct alignas(8) StagedMeta {
    uint64_t bloom;
    uint64_t epoch;
    uint32_t idx;
    uint32_t flags;
};

cudaError_t launch_scatter_rows(const MirrorPtrs &m, const float *rows, const StagedMeta *meta, uint32_t nrows,
                                int sm_count, cudaStream_t s);
cudaError_t launch_scatter_meta(const MirrorPtrs &m, const StagedMeta *meta, uint32_t n, cudaStream_t s);
cudaError_t launch_synth(const MirrorPtrs &m, uint64_t n, uint64_t row0, uint64_t seed_c, uint64_t seed_b,
                         float *scale_scratch, int sm_count, cudaStream_t s);

}  // namespace spl
