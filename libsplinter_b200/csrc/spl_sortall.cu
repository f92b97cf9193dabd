// spl_sortall.cu -- the reference's DEFAULT behaviour: `--limit` absent means every survivor is
// printed, sorted (/root/reference/splinter_cli_cmd_search.c:412-416, limit <= 0).  For k == 0 or
// k > SPL_GPU_KMAX_FUSED the scan kernel publishes one 64-bit ordering key per surviving row and
// this file compacts the rows that hold a key (in row order) and sorts ONLY those (key, row) pairs
// -- with a 12 %-selective label mask that is an eighth of the shard, not all of it.  Select and
// sort are plain library primitives (CUB, part of the CUDA toolkit); the sort is stable, so equal
// keys keep ascending row order -- the same tie rule as the fused top-k.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>

#include "spl_scan.cuh"

namespace spl {

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct NotSentinel {
    const uint64_t *keys;
    __host__ __device__ bool operator()(uint32_t row) const { return keys[row] != kKeySentinel; }
};

static size_t cub_temp_bytes(uint32_t n) {
    size_t sort_bytes = 0, sel_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const uint64_t *)nullptr, (uint64_t *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n);
    cub::DeviceSelect::If(nullptr, sel_bytes, cub::CountingInputIterator<uint32_t>(0), (uint32_t *)nullptr, (uint32_t *)nullptr,
                          (int)n, NotSentinel{ nullptr });
    return sort_bytes > sel_bytes ? sort_bytes : sel_bytes;
}

// workspace: keys_c[n] u64, keys_s[n] u64, rows_c[n] u32, rows_s[n] u32, library temp
size_t sortall_tmp_bytes(uint32_t n) {
    return 2 * align_up((size_t)n * 8, 256) + 2 * align_up((size_t)n * 4, 256) + align_up(cub_temp_bytes(n), 256) + 256;
}

namespace {
struct SortWs {
    uint64_t *keys_c, *keys_s;
    uint32_t *rows_c, *rows_s;
    unsigned char *lib;
    size_t lib_bytes;
};
SortWs carve(uint32_t n, void *tmp, size_t tmp_bytes) {
    unsigned char *p = static_cast<unsigned char *>(tmp);
    SortWs w;
    w.keys_c = reinterpret_cast<uint64_t *>(p);  p += align_up((size_t)n * 8, 256);
    w.keys_s = reinterpret_cast<uint64_t *>(p);  p += align_up((size_t)n * 8, 256);
    w.rows_c = reinterpret_cast<uint32_t *>(p);  p += align_up((size_t)n * 4, 256);
    w.rows_s = reinterpret_cast<uint32_t *>(p);  p += align_up((size_t)n * 4, 256);
    w.lib = p;
    w.lib_bytes = tmp_bytes - (size_t)(p - static_cast<unsigned char *>(tmp));
    return w;
}
}  // namespace

__global__ void gather_keys_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ rows, uint32_t n,
                                   uint64_t *__restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = keys[rows[i]];
}

__global__ void emit_sorted_kernel(ScanArgs a, const uint64_t *__restrict__ keys, const uint32_t *__restrict__ rows,
                                   uint32_t n_emit) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_emit) return;
    const uint32_t idx = rows[t];
    spl_gpu_result r;
    split_key(a.order, keys[t], r.similarity, r.distance);
    r.slot_idx = (uint32_t)(a.row0 + idx);
    r.has_embedding = (a.flags[idx] & SPL_ROW_EMB) ? 1u : 0u;
    r.epoch = a.epoch[idx];
    r.bloom = a.bloom[idx];
    a.out[t] = r;
}

// Step 1 (no host round trip yet): the rows that hold a key, in ascending row order, and how many there are.
// keys[n] is the scan's output (kKeySentinel where a row did not survive).
cudaError_t launch_sortall_select(const ScanArgs &a, const uint64_t *keys, void *tmp, size_t tmp_bytes, uint32_t *d_total,
                                  cudaStream_t stream) {
    SortWs w = carve(a.n, tmp, tmp_bytes);
    return cub::DeviceSelect::If(w.lib, w.lib_bytes, cub::CountingInputIterator<uint32_t>(0), w.rows_c, d_total, (int)a.n,
                                 NotSentinel{ keys }, stream);
}

// Step 2, once the host knows `total`: sort only the survivors.  The select kept them in ascending row order and the
// radix sort is stable, so equal keys come out by ascending slot index -- the tie rule of the fused top-k.
cudaError_t launch_sortall_sort(const ScanArgs &a, const uint64_t *keys, void *tmp, size_t tmp_bytes, uint32_t total,
                                uint32_t out_cap, cudaStream_t stream) {
    if (total == 0) return cudaSuccess;
    SortWs w = carve(a.n, tmp, tmp_bytes);
    gather_keys_kernel<<<(total + 255) / 256, 256, 0, stream>>>(keys, w.rows_c, total, w.keys_c);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    e = cub::DeviceRadixSort::SortPairs(w.lib, w.lib_bytes, w.keys_c, w.keys_s, w.rows_c, w.rows_s, (int)total, 0, 64, stream);
    if (e != cudaSuccess) return e;
    const uint32_t n_emit = total < out_cap ? total : out_cap;
    emit_sorted_kernel<<<(n_emit + 255) / 256, 256, 0, stream>>>(a, w.keys_s, w.rows_s, n_emit);
    return cudaGetLastError();
}

}  // namespace spl
