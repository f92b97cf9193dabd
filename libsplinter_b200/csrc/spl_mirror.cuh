// spl_mirror.cuh -- HBM mirror pointers and the staging record used by the sync path.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/splinter_gpu.h"

namespace spl {

// The bf16 shadow is stored TILED for the tensor-core path: [row tile of 256][k-block of 64][256 rows][64
// elements], so the operand tile one MMA stage needs (256 rows x 64 elements = 32 KB) is ONE contiguous
// burst in HBM instead of 256 separate 128-byte pieces 2*dim bytes apart.
constexpr uint32_t kE16TileRows = 256, kE16TileK = 64;
__host__ __device__ inline size_t e16_offset(uint64_t row, uint32_t k, uint32_t dim) {
    const uint64_t tile = row / kE16TileRows, rr = row % kE16TileRows;
    const uint32_t kb = k / kE16TileK, kk = k % kE16TileK;
    return (((tile * (dim / kE16TileK) + kb) * kE16TileRows + rr) * kE16TileK) + kk;
}
__host__ __device__ inline uint64_t e16_padded_rows(uint64_t n) { return (n + kE16TileRows - 1) / kE16TileRows * kE16TileRows; }

struct MirrorPtrs {
    float *E;
    uint64_t *bloom;
    uint8_t *flags;
    uint64_t *epoch;
    __nv_bfloat16 *E16;  // nullable; tiled layout, see e16_offset()
    float *norm2;        // nullable: |bf16(e)|^2
    float *err2;         // nullable: |e - bf16(e)|^2, what rounding the row to bf16 moved it by (selection proof, spl_batch.cu)
    uint32_t *norm2_max; // nullable: float bits of max norm2 ever written (only grows: an upper bound)
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
