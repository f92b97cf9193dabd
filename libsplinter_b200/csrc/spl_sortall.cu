// spl_sortall.cu -- the reference's DEFAULT behaviour: `--limit` absent means every survivor is
// printed, sorted (/root/reference/splinter_cli_cmd_search.c:412-416, limit <= 0).  For k == 0 or
// k > SPL_GPU_KMAX_FUSED the scan kernel publishes one 64-bit ordering key per surviving row and
// this file sorts (key, row) pairs for the whole shard.  The sort itself is a plain library
// radix sort (CUB, part of the CUDA toolkit); it is stable, so equal keys keep ascending row
// order -- the same tie rule as the fused top-k.  This is the cold path; the headline path is
// the fused top-k in spl_scan.cu.
#include <cub/device/device_radix_sort.cuh>

#include "spl_scan.cuh"

namespace spl {

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static size_t cub_temp_bytes(uint32_t n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t *)nullptr, (uint64_t *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n);
    return bytes;
}

size_t sortall_tmp_bytes(uint32_t n) {
    return align_up((size_t)n * 8, 256) + 2 * align_up((size_t)n * 4, 256) + align_up(cub_temp_bytes(n), 256) + 256;
}

__global__ void iota_kernel(uint32_t *p, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = i;
}

__global__ void emit_sorted_kernel(ScanArgs a, const uint64_t *__restrict__ keys, const uint32_t *__restrict__ rows,
                                   uint32_t out_cap) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.n) return;
    const uint64_t key = keys[t];
    if (key == kKeySentinel) {
        if (t == 0) *a.out_count = 0;
        return;
    }
    if (t == a.n - 1 || keys[t + 1] == kKeySentinel) *a.out_count = t + 1;  // total survivors
    if (t >= out_cap) return;
    const uint32_t idx = rows[t];
    spl_gpu_result r;
    split_key(a.order, key, r.similarity, r.distance);
    r.slot_idx = (uint32_t)(a.row0 + idx);
    r.has_embedding = (a.flags[idx] & SPL_ROW_EMB) ? 1u : 0u;
    r.epoch = a.epoch[idx];
    r.bloom = a.bloom[idx];
    a.out[t] = r;
}

cudaError_t launch_sortall(const ScanArgs &a, uint64_t *keys, void *tmp, size_t tmp_bytes, uint32_t out_cap,
                           cudaStream_t stream) {
    const uint32_t n = a.n;
    unsigned char *p = static_cast<unsigned char *>(tmp);
    uint64_t *keys_alt = reinterpret_cast<uint64_t *>(p);      p += align_up((size_t)n * 8, 256);
    uint32_t *rows_in = reinterpret_cast<uint32_t *>(p);       p += align_up((size_t)n * 4, 256);
    uint32_t *rows_out = reinterpret_cast<uint32_t *>(p);      p += align_up((size_t)n * 4, 256);
    size_t cub_bytes = tmp_bytes - (size_t)(p - static_cast<unsigned char *>(tmp));
    iota_kernel<<<(n + 255) / 256, 256, 0, stream>>>(rows_in, n);
    cudaError_t e = cub::DeviceRadixSort::SortPairs(p, cub_bytes, keys, keys_alt, rows_in, rows_out, (int)n, 0, 64, stream);
    if (e != cudaSuccess) return e;
    emit_sorted_kernel<<<(n + 255) / 256, 256, 0, stream>>>(a, keys_alt, rows_out, out_cap);
    return cudaGetLastError();
}

}  // namespace spl
