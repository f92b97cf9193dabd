// spl_mirror.cu -- device side of the HBM mirror: scatter of staged dirty rows, metadata-only
// updates, bf16 shadow rows + squared norms, and the on-device synthetic corpus generator.
//
// Layout in HBM (one shard = slots [slot_begin, slot_begin+n)):
//   E      fp32 [n][dim]   dense copy of splinter_slot::embedding (splinter.h:252-254); the
//                          reference's AoS pitch of 128+4*dim bytes is dropped so a row is one
//                          contiguous 4*dim-byte burst
//   bloom  u64  [n]        splinter_slot::bloom  (splinter.h:248)
//   flags  u8   [n]        SPL_ROW_* validity bits derived from hash / val_len / epoch / key[0]
//   epoch  u64  [n]        splinter_slot::epoch at the time the row was copied
//   E16    bf16 [n][dim]   optional shadow for the tcgen05 batched path
//   norm2  fp32 [n]        |bf16(e)|^2 for that path
//   err2   fp32 [n]        |e - bf16(e)|^2: how far rounding moved the row (selection proof of that path)
#include "spl_mirror.cuh"

#include <cuda_bf16.h>

namespace spl {

// one warp per staged row: 128-bit copies staging -> E[idx]; lane 0 writes the metadata
__global__ void __launch_bounds__(256) scatter_rows_kernel(MirrorPtrs m, const float *__restrict__ rows,
                                                           const StagedMeta *__restrict__ meta, uint32_t nrows) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nv = m.dim >> 2;
    for (uint32_t r = w; r < nrows; r += nw) {
        const StagedMeta mt = meta[r];
        const float4 *src = reinterpret_cast<const float4 *>(rows + (size_t)r * m.dim);
        float4 *dst = reinterpret_cast<float4 *>(m.E + (size_t)mt.idx * m.dim);
        float n2 = 0.f, er2 = 0.f;
        for (uint32_t c = lane; c < nv; c += 32) {
            const float4 v = src[c];
            dst[c] = v;
            if (m.E16) {
                __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
                uint2 packed;
                packed.x = *reinterpret_cast<uint32_t *>(&lo);
                packed.y = *reinterpret_cast<uint32_t *>(&hi);
                *reinterpret_cast<uint2 *>(m.E16 + e16_offset(mt.idx, c * 4, m.dim)) = packed;
                const float2 a = __bfloat1622float2(lo), b = __bfloat1622float2(hi);
                n2 += a.x * a.x + a.y * a.y + b.x * b.x + b.y * b.y;
                const float ex = v.x - a.x, ey = v.y - a.y, ez = v.z - b.x, ew = v.w - b.y;
                er2 += ex * ex + ey * ey + ez * ez + ew * ew;
            }
        }
        if (m.E16) {
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                n2 += __shfl_xor_sync(0xffffffffu, n2, off);
                er2 += __shfl_xor_sync(0xffffffffu, er2, off);
            }
        }
        if (lane == 0) {
            m.bloom[mt.idx] = mt.bloom;
            m.epoch[mt.idx] = mt.epoch;
            m.flags[mt.idx] = (uint8_t)mt.flags;
            if (m.E16) {
                m.norm2[mt.idx] = n2;
                m.err2[mt.idx] = er2 * 1.0001f;   // rounded up a little: it is used as a bound
                if (__float_as_uint(n2) > *m.norm2_max) atomicMax(m.norm2_max, __float_as_uint(n2));  // n2 >= 0: bits order like values
            }
        }
    }
}

__global__ void __launch_bounds__(256) scatter_meta_kernel(MirrorPtrs m, const StagedMeta *__restrict__ meta,
                                                           uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const StagedMeta mt = meta[i];
    m.bloom[mt.idx] = mt.bloom;
    m.epoch[mt.idx] = mt.epoch;
    m.flags[mt.idx] = (uint8_t)mt.flags;
}

cudaError_t launch_scatter_rows(const MirrorPtrs &m, const float *rows, const StagedMeta *meta, uint32_t nrows,
                                int sm_count, cudaStream_t s) {
    if (!nrows) return cudaSuccess;
    uint32_t blocks = (nrows + 7) / 8;
    const uint32_t cap = (uint32_t)sm_count * 8;
    if (blocks > cap) blocks = cap;
    scatter_rows_kernel<<<blocks, 256, 0, s>>>(m, rows, meta, nrows);
    return cudaGetLastError();
}

cudaError_t launch_scatter_meta(const MirrorPtrs &m, const StagedMeta *meta, uint32_t n, cudaStream_t s) {
    if (!n) return cudaSuccess;
    scatter_meta_kernel<<<(n + 255) / 256, 256, 0, s>>>(m, meta, n);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// synthetic corpus; restates oracle/spl_oracle.c splo_synth_* with IEEE-rounded
// intrinsics in the same (sequential) order, so host and device agree bit-for-bit.
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__device__ __forceinline__ float synth_raw(uint64_t hr, uint32_t d) {
    const uint64_t h = mix64(hr + d);
    const int32_t mnt = (int32_t)(h >> 40) - (1 << 23);
    return __fmul_rn((float)mnt, 1.0f / 8388608.0f);
}

// one thread per (row, 4-float chunk) would break the sequential norm; instead: kernel 1 computes
// the per-row scale with one thread per row, kernel 2 writes the scaled values fully coalesced.
__global__ void __launch_bounds__(256) synth_scale_kernel(MirrorPtrs m, uint64_t n, uint64_t row0, uint64_t seed_c,
                                                          uint64_t seed_b, float *__restrict__ scale) {
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const uint64_t g = row0 + r;
    const uint64_t c = mix64(mix64(seed_c ^ 0xC1A55C1A55ull) ^ (g * 0xD1342543DE82EF95ull)) % 1000u;
    float sc = 0.f;
    uint64_t bloom = 0;
    uint8_t flags = 0;
    if (c >= 1) {
        flags = SPL_ROW_LIST | SPL_ROW_ENUM | SPL_ROW_EMB;
        const uint64_t hb = mix64(seed_b ^ (g * 0xA24BAED4963EE407ull));
        bloom = 1ull << 63;
        for (unsigned b = 0; b < 63; b++) {
            const uint64_t h = mix64(hb + b);
            const unsigned nb = 1 + (b & 3);
            if ((h & ((1u << nb) - 1)) == 0) bloom |= 1ull << b;
        }
        if (c >= 21) {
            const uint64_t hr = mix64(seed_c ^ (g * 0x9E3779B97F4A7C15ull));
            float ss = 0.f;
            for (uint32_t d = 0; d < m.dim; d++) {
                const float v = synth_raw(hr, d);
                ss = __fadd_rn(ss, __fmul_rn(v, v));
            }
            sc = __fdiv_rn(1.0f, __fsqrt_rn(ss));
        }
    }
    scale[r] = sc;
    m.bloom[r] = bloom;
    m.flags[r] = flags;
    m.epoch[r] = flags ? (sc != 0.f ? 4 : 2) : 0;  // set -> 2, set_embedding -> 4
}

__global__ void __launch_bounds__(256) synth_fill_kernel(MirrorPtrs m, uint64_t n, uint64_t row0, uint64_t seed_c,
                                                         const float *__restrict__ scale) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t w = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t nw = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t r = w; r < n; r += nw) {
        const float sc = scale[r];
        const uint64_t hr = mix64(seed_c ^ ((row0 + r) * 0x9E3779B97F4A7C15ull));
        float n2 = 0.f, er2 = 0.f;
        for (uint32_t d = lane; d < m.dim; d += 32) {
            const float v = sc != 0.f ? __fmul_rn(synth_raw(hr, d), sc) : 0.f;
            m.E[(size_t)r * m.dim + d] = v;
            if (m.E16) {
                const __nv_bfloat16 hb = __float2bfloat16_rn(v);
                m.E16[e16_offset(r, d, m.dim)] = hb;
                const float fb = __bfloat162float(hb);
                n2 += fb * fb;
                er2 += (v - fb) * (v - fb);
            }
        }
        if (m.E16) {
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                n2 += __shfl_xor_sync(0xffffffffu, n2, off);
                er2 += __shfl_xor_sync(0xffffffffu, er2, off);
            }
            if (lane == 0) {
                m.norm2[r] = n2;
                m.err2[r] = er2 * 1.0001f;
                if (__float_as_uint(n2) > *m.norm2_max) atomicMax(m.norm2_max, __float_as_uint(n2));
            }
        }
    }
}

cudaError_t launch_synth(const MirrorPtrs &m, uint64_t n, uint64_t row0, uint64_t seed_c, uint64_t seed_b,
                         float *scale_scratch, int sm_count, cudaStream_t s) {
    if (!n) return cudaSuccess;
    synth_scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(m, n, row0, seed_c, seed_b, scale_scratch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    synth_fill_kernel<<<sm_count * 8, 256, 0, s>>>(m, n, row0, seed_c, scale_scratch);
    return cudaGetLastError();
}

}  // namespace spl
