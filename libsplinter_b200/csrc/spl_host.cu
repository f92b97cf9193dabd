// spl_host.cu -- host side of libsplinter_gpu.so: the C ABI declared in include/splinter_gpu.h.
//
//  * attach: map the reference's store (same shm name / file as splinter_open,
//    /root/reference/splinter.c:235-248), gate on magic/version (:246) and slot geometry;
//  * sync:   epoch-consistent snapshot of the slot array into HBM.  A slot is copied only while
//    its seqlock epoch is even and unchanged over the copy (reader rule, splinter.c:598-605);
//    only slots whose cached {hash, epoch, bloom, validity} moved are re-sent (writers:
//    splinter.c:326-429, 567-588, 799-867);
//  * search: launches the sm_100a kernels (spl_scan.cu, spl_exact.cu, spl_batch.cu).
//
// No CPU fallback exists: without a CUDA device every entry point fails with errno = EIO.
#include <errno.h>
#include <fcntl.h>
#include <float.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <poll.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <sys/uio.h>
#include <time.h>
#include <unistd.h>

#include <array>
#include <atomic>
#include <chrono>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "spl_batch.cuh"
#include "spl_exact.cuh"
#include "spl_mirror.cuh"
#include "spl_scan.cuh"

using namespace spl;

// ---- store layout (splinter.h:176-255; offsets measured with offsetof on the reference) --------
namespace layout {
constexpr uint32_t kMagic = 0x534C4E54u, kVersion = 4u;
constexpr size_t kHeaderSize = 5440, kSlotMeta = 128;
constexpr size_t H_MAGIC = 0, H_VERSION = 4, H_SLOTS = 8, H_MAXVAL = 12, H_EPOCH = 16;
constexpr size_t H_BUS_OWNER_FD = 4352, H_BUS_OWNER_PID = 4356;  // H->event_bus.owner_fd / owner_pid (splinter.h:144-148)
constexpr size_t S_HASH = 0, S_EPOCH = 8, S_VAL_LEN = 20, S_BLOOM = 56, S_KEY = 64, S_EMBED = 128;
}  // namespace layout

static thread_local char g_err[512] = "";
static inline uint64_t ld_acq64(const uint8_t *p);

static int fail(int rc, int err, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    errno = err;
    return rc;
}
namespace spl {
// lets other translation units report through spl_gpu_last_error()
void set_last_error(const char *msg) { snprintf(g_err, sizeof g_err, "%s", msg); }
}  // namespace spl

#define CUDA_TRY(expr)                                                                                   \
    do {                                                                                                 \
        cudaError_t e__ = (expr);                                                                        \
        if (e__ != cudaSuccess) return fail(-1, EIO, "%s failed: %s", #expr, cudaGetErrorString(e__));   \
    } while (0)

struct StageBuf {
    float *h_rows = nullptr, *d_rows = nullptr;
    StagedMeta *h_meta = nullptr, *d_meta = nullptr;
    uint32_t cap = 0, n = 0;
    cudaEvent_t ev = nullptr;
    bool in_flight = false;
};

struct spl_gpu_ctx {
    int device = 0, sm_count = 0;
    cudaStream_t stream = nullptr;
    // store mapping
    const uint8_t *base = nullptr;
    size_t map_size = 0;
    bool own_map = false;
    uint64_t total_slots = 0;
    uint32_t dim = 0;
    size_t slot_size = 0;
    uint64_t slot_begin = 0, n = 0;  // shard: local row r <-> global slot slot_begin + r
    bool dense = false;
    uint32_t sync_threads = 0;
    // device mirror
    MirrorPtrs m{};
    uint64_t hbm_bytes = 0;
    // host caches for dirty detection (store-backed contexts)
    std::vector<uint64_t> c_hash, c_epoch, c_bloom;
    std::vector<uint8_t> c_flags;   // SPL_ROW_* incl. EMB
    std::vector<uint8_t> c_state;   // bit0: mirrored at least once, bit1: mirrored row is non-zero
    std::vector<uint64_t> c_fp;     // fingerprint of the mirrored row's first 64 bytes (epoch-ABA detector)
    std::vector<uint64_t> c_full;   // hash of the whole mirrored row (periodic re-validation of clean rows)
    uint64_t sweep_no = 0;
    uint64_t c_header_epoch = 0;
    bool first_sync = true;
    int bus_fd = -1;               // this process's duplicate of the owner's eventfd (spl_gpu_wait)
    int32_t bus_owner_pid = 0, bus_owner_fd = -1;
    std::atomic<uint64_t> generation{0};
    // per-sync-thread staging (2 row buffers + 2 meta buffers each)
    std::vector<StageBuf> stage;
    uint32_t stage_rows = 0, stage_meta = 0;
    // search scratch
    int scan_grid = 0;
    // Two scratch sets (per-CTA candidate lists + counters) used by alternate scan launches, so that a scan may start
    // while its predecessor is still merging (see scan_tma_kernel); d_cands / d_counters = set 0.
    Cand *d_cands = nullptr, *d_cands_set[2] = { nullptr, nullptr };
    uint32_t *d_counters = nullptr, *d_counters_set[2] = { nullptr, nullptr };
    uint32_t scan_set = 0;                 // set the NEXT scan launch takes
    cudaStream_t last_tma_stream = nullptr;  // stream of the most recent launch if that was a TMA scan, else null
    bool last_was_tma = false;
    float *d_q = nullptr, *h_q = nullptr;
    spl_gpu_result *d_out = nullptr, *h_out = nullptr;
    uint32_t *d_count = nullptr, *h_count = nullptr;
    spl_gpu_result h_bounce[SPL_GPU_KMAX_FUSED];  // small host bounce buffers for the retry paths
    uint32_t h_bounce_count = 0;
    uint32_t *h_total = nullptr, *m_total = nullptr;  // mapped: rows that survived predicate + thresholds, per query
    uint32_t *h_seq = nullptr, *m_seq = nullptr;      // mapped: completion word of the host-buffer fused search (see wait_answer)
    uint32_t seq = 0;
    uint32_t last_nq = 0;
    spl_gpu_result *m_out = nullptr;  // device views of the mapped host buffers h_out / h_count
    uint32_t *m_count = nullptr;
    uint32_t q_cap = 0, out_cap = 0;
    // unlimited-k scratch (lazy)
    uint64_t *d_keys = nullptr;
    void *d_sort_tmp = nullptr;
    size_t sort_tmp_bytes = 0;
    spl_gpu_result *d_big_out = nullptr;
    uint32_t big_out_cap = 0;
    float *d_synth_scale = nullptr;
    // SPL_MODE_EXACT scratch: per query of a chunk, kExactMaxCand candidates + sort space (lazy)
    unsigned char *d_exact = nullptr;
    uint32_t *h_xvalid = nullptr, *m_xvalid = nullptr;  // mapped: proof flags per query of a chunk
    // batched tensor-core path scratch (lazy)
    unsigned char *d_batch = nullptr;
    BatchScratchLayout batch_layout{};
    float *d_qbatch = nullptr;
    spl_gpu_result *d_bout = nullptr;
    uint32_t *d_bcount = nullptr;
    spl_gpu_result *h_bout = nullptr;
    uint32_t *h_bcount = nullptr;
    uint32_t bout_cap = 0;
    // bookkeeping
    std::mutex mu;  // serialises enqueueing on `stream` and the scratch above
    std::mutex sweep_mu;  // one sweep at a time per context keeps the host caches single-writer
    // GPU-side ordering between mirror updates (always on `stream`) and searches (on `stream` or a caller's stream):
    // the seqlock's guarantee carried onto the device.  ev_flush = the last scatter enqueued; ev_search = the last
    // search enqueued on search_stream.  Searches also share the scratch above, so consecutive searches on different
    // streams are chained through ev_search as well.
    cudaEvent_t ev_flush = nullptr, ev_search = nullptr;
    bool flush_recorded = false, search_recorded = false;
    cudaStream_t search_stream = nullptr;
    bool counters_dirty = false;   // a failed call may have left d_counters non-zero: re-zero before the next scan
    bool totals_valid = false;     // spl_gpu_last_totals: false after a search path that does not count survivors
    std::atomic<uint64_t> n_batch_queries{0}, n_batch_unproven{0}, n_exact_unproven{0}, n_stale_rows{0}, n_batch_extra{0};
    std::atomic<uint64_t> launches{0};
    bool force_ldg = false;
    bool no_multi = false;  // developer knob SPL_GPU_SCAN_MULTI=0: one query per launch even for small batches
    bool no_nq8 = false;    // developer knob SPL_GPU_SCAN_NQ8=0: at most four queries per pass
    bool no_overlap = false;  // developer knob SPL_GPU_SCAN_OVERLAP=0: never let a scan start during its predecessor's tail
    unsigned long long *d_dbg = nullptr;  // SPL_GPU_DEBUG_TIMELINE=1: per-CTA timestamps of the last scan
    bool timing = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    double last_scan_ms = -1.0;
};

static constexpr uint32_t kMaxBatchQ = 4096;

// ------------------------------------------------------------------------------------------------
static int device_setup(spl_gpu_ctx *c, const spl_gpu_opts *o) {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(-1, EIO, "no CUDA device (%s); libsplinter_gpu has no CPU fallback", cudaGetErrorString(e));
    int dev = (o && o->device >= 0) ? o->device : -1;
    if (dev < 0) CUDA_TRY(cudaGetDevice(&dev));
    if (dev >= ndev) return fail(-2, EINVAL, "device %d out of range (%d devices)", dev, ndev);
    CUDA_TRY(cudaSetDevice(dev));
    cudaDeviceProp p;
    CUDA_TRY(cudaGetDeviceProperties(&p, dev));
    if (p.major < 10)
        return fail(-1, EIO, "device %d is sm_%d%d; this library carries sm_100a code only", dev, p.major, p.minor);
    c->device = dev;
    c->sm_count = p.multiProcessorCount;
    CUDA_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreate(&c->ev0));
    CUDA_TRY(cudaEventCreate(&c->ev1));
    CUDA_TRY(cudaEventCreateWithFlags(&c->ev_flush, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&c->ev_search, cudaEventDisableTiming));
    CUDA_TRY(scan_tma_prepare());
    CUDA_TRY(batch_prepare());
    CUDA_TRY(merge_prepare());
    const char *impl = getenv("SPL_GPU_SCAN_IMPL");  // developer knob: "ldg" forces the register-path kernel
    c->force_ldg = impl && strcmp(impl, "ldg") == 0;
    const char *multi = getenv("SPL_GPU_SCAN_MULTI");
    c->no_multi = multi && strcmp(multi, "0") == 0;
    { const char *e8 = getenv("SPL_GPU_SCAN_NQ8"); c->no_nq8 = e8 && strcmp(e8, "0") == 0; }
    const char *ovl = getenv("SPL_GPU_SCAN_OVERLAP");
    c->no_overlap = ovl && strcmp(ovl, "0") == 0;
    if (getenv("SPL_GPU_DEBUG_TIMELINE")) {
        CUDA_TRY(cudaMalloc((void **)&c->d_dbg, 1024 * 8 * sizeof(unsigned long long)));
        CUDA_TRY(cudaMemset(c->d_dbg, 0, 1024 * 8 * sizeof(unsigned long long)));
    }
    return 0;
}

template <typename T>
static int dev_alloc(spl_gpu_ctx *c, T **p, size_t count, bool zero) {
    size_t bytes = count * sizeof(T);
    if (bytes == 0) bytes = sizeof(T);
    cudaError_t e = cudaMalloc((void **)p, bytes);
    if (e != cudaSuccess) return fail(-1, ENOMEM, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
    c->hbm_bytes += bytes;
    if (zero) CUDA_TRY(cudaMemsetAsync(*p, 0, bytes, c->stream));
    return 0;
}

static int alloc_mirror(spl_gpu_ctx *c, bool bf16) {
    int rc;
    if ((rc = dev_alloc(c, &c->m.E, c->n * c->dim, true))) return rc;
    if ((rc = dev_alloc(c, &c->m.bloom, c->n, true))) return rc;
    if ((rc = dev_alloc(c, &c->m.flags, c->n, true))) return rc;
    if ((rc = dev_alloc(c, &c->m.epoch, c->n, true))) return rc;
    if (bf16 && c->dim % kE16TileK == 0) {  // the tiled bf16 shadow needs whole 64-element k-blocks
        if ((rc = dev_alloc(c, &c->m.E16, e16_padded_rows(c->n) * c->dim, true))) return rc;
        if ((rc = dev_alloc(c, &c->m.norm2, c->n, true))) return rc;
        if ((rc = dev_alloc(c, &c->m.err2, c->n, true))) return rc;
        if ((rc = dev_alloc(c, &c->m.norm2_max, 1, true))) return rc;
    }
    c->m.dim = c->dim;
    // search scratch
    c->scan_grid = c->sm_count * scan_fast_ctas_per_sm(c->dim, 0);
    // one k-entry list per CTA and query: up to 4 queries x 256 entries (small batches) per launch
    if ((rc = dev_alloc(c, &c->d_cands, (size_t)c->scan_grid * SPL_GPU_KMAX_FUSED * 4 * 2, false))) return rc;
    if ((rc = dev_alloc(c, &c->d_counters, 32, true))) return rc;   // per set: chunk, ticket, 4 survivor counts, 3 more tickets
    c->d_cands_set[0] = c->d_cands; c->d_cands_set[1] = c->d_cands + (size_t)c->scan_grid * SPL_GPU_KMAX_FUSED * 4;
    c->d_counters_set[0] = c->d_counters; c->d_counters_set[1] = c->d_counters + 16;
    c->q_cap = 64;
    c->out_cap = 64 * SPL_GPU_KMAX_FUSED;
    if ((rc = dev_alloc(c, &c->d_q, (size_t)c->q_cap * c->dim, false))) return rc;
    if ((rc = dev_alloc(c, &c->d_out, c->out_cap, false))) return rc;
    if ((rc = dev_alloc(c, &c->d_count, c->q_cap, true))) return rc;
    CUDA_TRY(cudaMallocHost((void **)&c->h_q, (size_t)c->q_cap * c->dim * sizeof(float)));
    // results land straight in mapped pinned host memory: the last CTA's few hundred bytes of stores
    // cross PCIe inside the kernel, which saves two cudaMemcpyAsync round trips per search
    CUDA_TRY(cudaHostAlloc((void **)&c->h_out, (size_t)c->out_cap * sizeof(spl_gpu_result), cudaHostAllocMapped));
    CUDA_TRY(cudaHostAlloc((void **)&c->h_count, (size_t)c->q_cap * sizeof(uint32_t), cudaHostAllocMapped));
    CUDA_TRY(cudaHostGetDevicePointer((void **)&c->m_out, c->h_out, 0));
    CUDA_TRY(cudaHostGetDevicePointer((void **)&c->m_count, c->h_count, 0));
    CUDA_TRY(cudaHostAlloc((void **)&c->h_total, (size_t)kMaxBatchQ * sizeof(uint32_t), cudaHostAllocMapped));
    CUDA_TRY(cudaHostGetDevicePointer((void **)&c->m_total, c->h_total, 0));
    memset(c->h_total, 0, (size_t)kMaxBatchQ * sizeof(uint32_t));
    CUDA_TRY(cudaHostAlloc((void **)&c->h_seq, 64, cudaHostAllocMapped));
    CUDA_TRY(cudaHostGetDevicePointer((void **)&c->m_seq, c->h_seq, 0));
    memset(c->h_seq, 0, 64);
    // staging for sync / dense upload.  Sized by the shard: a 4096-slot store must not pay for (and wait on) half a
    // gigabyte of pinned host memory -- pinning is the slowest part of attaching a small store.
    if (c->sync_threads == 0) {
        unsigned hc = std::thread::hardware_concurrency();
        c->sync_threads = hc ? (hc > 16 ? 16 : hc) : 4;
        const uint64_t by_size = c->n / 16384 + 1;          // one sweep thread per 16 K slots is plenty
        if (c->sync_threads > by_size) c->sync_threads = (uint32_t)by_size;
    }
    const size_t row_bytes = (size_t)c->dim * 4;
    c->stage_rows = (uint32_t)((8u << 20) / row_bytes);
    {
        const uint64_t share = c->n / c->sync_threads + 1;  // rows one thread stages if every row is dirty
        if (c->stage_rows > share) c->stage_rows = (uint32_t)share;
    }
    if (c->stage_rows < 64) c->stage_rows = 64;
    c->stage_meta = 32768;
    if (c->stage_meta > c->n + 64) c->stage_meta = (uint32_t)c->n + 64;
    c->stage.resize((size_t)c->sync_threads * 4);
    for (size_t i = 0; i < c->stage.size(); i++) {
        StageBuf &b = c->stage[i];
        const bool rows = (i & 2) == 0;  // per thread: [0,1] row buffers, [2,3] meta-only buffers
        b.cap = rows ? c->stage_rows : c->stage_meta;
        if (rows) {
            CUDA_TRY(cudaMallocHost((void **)&b.h_rows, (size_t)b.cap * row_bytes));
            if ((rc = dev_alloc(c, &b.d_rows, (size_t)b.cap * c->dim, false))) return rc;
        }
        CUDA_TRY(cudaMallocHost((void **)&b.h_meta, (size_t)b.cap * sizeof(StagedMeta)));
        if ((rc = dev_alloc(c, &b.d_meta, b.cap, false))) return rc;
        CUDA_TRY(cudaEventCreateWithFlags(&b.ev, cudaEventDisableTiming));
    }
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

static void destroy(spl_gpu_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->search_recorded) cudaEventSynchronize(c->ev_search);   // a caller-stream search may still read the mirror
    cudaFree(c->m.E); cudaFree(c->m.bloom); cudaFree(c->m.flags); cudaFree(c->m.epoch);
    cudaFree(c->m.E16); cudaFree(c->m.norm2); cudaFree(c->m.err2); cudaFree(c->m.norm2_max);
    cudaFree(c->d_cands); cudaFree(c->d_counters); cudaFree(c->d_q); cudaFree(c->d_out); cudaFree(c->d_count);
    cudaFree(c->d_exact); cudaFreeHost(c->h_xvalid);
    cudaFree(c->d_batch); cudaFree(c->d_qbatch); cudaFree(c->d_bout); cudaFree(c->d_bcount);
    cudaFreeHost(c->h_bout); cudaFreeHost(c->h_bcount);
    cudaFree(c->d_keys); cudaFree(c->d_sort_tmp); cudaFree(c->d_big_out); cudaFree(c->d_synth_scale);
    cudaFreeHost(c->h_q); cudaFreeHost(c->h_out); cudaFreeHost(c->h_count); cudaFreeHost(c->h_total); cudaFreeHost(c->h_seq);
    for (StageBuf &b : c->stage) {
        cudaFreeHost(b.h_rows); cudaFreeHost(b.h_meta); cudaFree(b.d_rows); cudaFree(b.d_meta);
        if (b.ev) cudaEventDestroy(b.ev);
    }
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->ev_flush) cudaEventDestroy(c->ev_flush);
    if (c->ev_search) cudaEventDestroy(c->ev_search);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->bus_fd >= 0) close(c->bus_fd);
    if (c->own_map && c->base) munmap((void *)c->base, c->map_size);
    delete c;
}

// ---- geometry gate: slot size must be 128 + 4*D ------------------------------------
static int parse_geometry(spl_gpu_ctx *c, const uint8_t *base, size_t size) {
    using namespace layout;
    if (size < kHeaderSize) return fail(-1, EINVAL, "mapping smaller than a splinter header");
    uint32_t magic, version, slots, maxval;
    memcpy(&magic, base + H_MAGIC, 4); memcpy(&version, base + H_VERSION, 4);
    memcpy(&slots, base + H_SLOTS, 4); memcpy(&maxval, base + H_MAXVAL, 4);
    if (magic != kMagic || version != kVersion)
        return fail(-1, EINVAL, "not a splinter v%u store (magic %08x version %u)", kVersion, magic, version);
    if (slots == 0) return fail(-1, EINVAL, "store has no slots");
    uint64_t rest = size - kHeaderSize, arena = (uint64_t)slots * maxval;
    if (rest < arena || (rest - arena) % slots) return fail(-1, EINVAL, "store size does not match its header");
    uint64_t ss = (rest - arena) / slots;
    if (ss <= kSlotMeta || (ss - kSlotMeta) % 16)
        return fail(-1, EINVAL, "slot size %llu: store was built without embeddings (need 128 + 4*D)",
                    (unsigned long long)ss);
    c->total_slots = slots;
    c->slot_size = (size_t)ss;
    c->dim = (uint32_t)((ss - kSlotMeta) / 4);
    c->base = base;
    c->map_size = size;
    return 0;
}

static double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

static spl_gpu_ctx *attach_common(spl_gpu_ctx *c, const spl_gpu_opts *o) {
    const bool trace = getenv("SPL_GPU_DEBUG_ATTACH") != nullptr;   // developer knob: where does attach time go
    const double t0 = now_ms();
    c->slot_begin = o ? o->slot_begin : 0;
    uint64_t cnt = o ? o->slot_count : 0;
    if (c->slot_begin >= c->total_slots) { fail(-2, EINVAL, "slot_begin beyond the store"); destroy(c); return nullptr; }
    if (cnt == 0 || c->slot_begin + cnt > c->total_slots) cnt = c->total_slots - c->slot_begin;
    c->n = cnt;
    c->sync_threads = o ? o->sync_threads : 0;
    if (device_setup(c, o)) { destroy(c); return nullptr; }
    const double t1 = now_ms();
    if (alloc_mirror(c, o && (o->flags & SPL_OPT_BF16_MIRROR))) { destroy(c); return nullptr; }
    const double t2 = now_ms();
    c->c_hash.assign(c->n, 0); c->c_epoch.assign(c->n, 0); c->c_bloom.assign(c->n, 0);
    c->c_flags.assign(c->n, 0); c->c_state.assign(c->n, 0); c->c_fp.assign(c->n, 0); c->c_full.assign(c->n, 0);
    if (spl_gpu_sync(c, nullptr) != 0) { destroy(c); return nullptr; }
    if (trace)
        fprintf(stderr, "spl_gpu_attach: device + kernels %.1f ms, mirror + staging allocation %.1f ms, first sync %.1f ms\n",
                t1 - t0, t2 - t1, now_ms() - t2);
    return c;
}

extern "C" spl_gpu_ctx *spl_gpu_attach(const char *name_or_path, const spl_gpu_opts *o) {
    if (!name_or_path) { fail(-2, EINVAL, "null store name"); return nullptr; }
    int fd = (o && (o->flags & SPL_OPT_PERSISTENT)) ? open(name_or_path, O_RDONLY) : shm_open(name_or_path, O_RDONLY, 0);
    if (fd < 0) { fail(-1, errno, "cannot open store '%s': %s", name_or_path, strerror(errno)); return nullptr; }
    struct stat st;
    if (fstat(fd, &st) != 0) { int e = errno; close(fd); fail(-1, e, "fstat: %s", strerror(e)); return nullptr; }
    void *base = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_SHARED, fd, 0);
    close(fd);
    if (base == MAP_FAILED) { fail(-1, errno, "mmap: %s", strerror(errno)); return nullptr; }
    spl_gpu_ctx *c = new spl_gpu_ctx();
    c->own_map = true;
    c->base = (const uint8_t *)base;
    c->map_size = (size_t)st.st_size;
    if (parse_geometry(c, (const uint8_t *)base, (size_t)st.st_size)) { destroy(c); return nullptr; }
    return attach_common(c, o);
}

extern "C" spl_gpu_ctx *spl_gpu_attach_image(const void *base, size_t size, const spl_gpu_opts *o) {
    if (!base) { fail(-2, EINVAL, "null image"); return nullptr; }
    spl_gpu_ctx *c = new spl_gpu_ctx();
    if (parse_geometry(c, (const uint8_t *)base, size)) { destroy(c); return nullptr; }
    return attach_common(c, o);
}

extern "C" spl_gpu_ctx *spl_gpu_open_dense(uint64_t n_rows, uint32_t dim, uint64_t global_row0, const spl_gpu_opts *o) {
    if (n_rows == 0 || n_rows > 0xFFFFFFE0ull || dim == 0 || dim % 4) {   // 32-row groups are counted in 32 bits
        fail(-2, EINVAL, "dense shard needs 0 < rows <= 2^32-32 and dim %% 4 == 0");
        return nullptr;
    }
    spl_gpu_ctx *c = new spl_gpu_ctx();
    c->dense = true;
    c->dim = dim;
    c->n = n_rows;
    c->total_slots = n_rows;
    c->slot_begin = global_row0;
    c->sync_threads = o ? o->sync_threads : 0;
    if (c->sync_threads == 0) c->sync_threads = 2;
    if (device_setup(c, o) || alloc_mirror(c, o && (o->flags & SPL_OPT_BF16_MIRROR))) { destroy(c); return nullptr; }
    return c;
}

extern "C" void spl_gpu_detach(spl_gpu_ctx *c) { destroy(c); }

extern "C" const char *spl_gpu_last_error(void) { return g_err; }

extern "C" int spl_gpu_get_info(const spl_gpu_ctx *c, spl_gpu_info *out) {
    if (!c || !out) return fail(-2, EINVAL, "null argument");
    memset(out, 0, sizeof *out);
    out->abi_version = SPL_GPU_ABI_VERSION;
    out->dim = c->dim;
    out->total_slots = c->total_slots;
    out->slot_begin = c->slot_begin;
    out->slot_count = c->n;
    out->generation = c->generation.load();
    out->device = c->device;
    out->sm_count = c->sm_count;
    out->has_bf16_mirror = c->m.E16 != nullptr;
    out->is_dense = c->dense;
    out->hbm_bytes = c->hbm_bytes;
    return 0;
}

// developer aid, not part of the public header: dump the per-CTA timeline of the last TMA scan
extern "C" int spl_gpu_debug_timeline(spl_gpu_ctx *c, unsigned long long *out, int max_ctas) {
    if (!c || !c->d_dbg) return -1;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    int n = max_ctas < c->sm_count ? max_ctas : c->sm_count;
    cudaMemcpy(out, c->d_dbg, (size_t)n * 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    return n;
}

extern "C" int spl_gpu_slot_meta(const spl_gpu_ctx *c, uint32_t slot_idx, char key_out[64], uint32_t *val_len,
                                 uint8_t *type_flag) {
    if (!c || !key_out) return fail(-2, EINVAL, "null argument");
    if (c->dense || !c->base) return fail(-2, EINVAL, "dense contexts have no keys");
    if (slot_idx >= c->total_slots) return fail(-2, EINVAL, "slot index beyond the store");
    const uint8_t *sl = c->base + layout::kHeaderSize + (size_t)slot_idx * c->slot_size;
    memcpy(key_out, sl + layout::S_KEY, 64);
    key_out[63] = '\0';
    if (val_len) *val_len = __atomic_load_n(reinterpret_cast<const uint32_t *>(sl + layout::S_VAL_LEN), __ATOMIC_ACQUIRE);
    if (type_flag) *type_flag = sl[24];
    return 0;
}

extern "C" int spl_gpu_find_key(const spl_gpu_ctx *c, const char *key, uint32_t *slot_idx) {
    if (!c || !key || !slot_idx) return fail(-2, EINVAL, "null argument");
    if (c->dense || !c->base) return fail(-2, EINVAL, "dense contexts have no keys");
    uint64_t h = 14695981039346656037ull;                                   // FNV-1a, splinter.c:67-71
    for (const unsigned char *s = (const unsigned char *)key; *s; ++s) h = (h ^ *s) * 1099511628211ull;
    const uint64_t slots = c->total_slots, first = h % slots;
    for (uint64_t i = 0; i < slots; i++) {                                   // linear probe, splinter.c:535-539
        const uint64_t idx = (first + i) % slots;
        const uint8_t *sl = c->base + layout::kHeaderSize + idx * c->slot_size;
        if (ld_acq64(sl + layout::S_HASH) == h && strncmp((const char *)sl + layout::S_KEY, key, 64) == 0) {
            *slot_idx = (uint32_t)idx;
            return 0;
        }
    }
    return fail(-1, ENOENT, "no slot holds key '%.64s'", key);
}

extern "C" int spl_gpu_last_totals(const spl_gpu_ctx *c, uint32_t *totals, uint32_t nq) {
    if (!c || !totals) return fail(-2, EINVAL, "null argument");
    if (!c->totals_valid)
        return fail(-1, ENODATA, "the last search on this context took the batched tensor-core path, which does not count survivors");
    if (nq > kMaxBatchQ) nq = kMaxBatchQ;
    memcpy(totals, c->h_total, (size_t)nq * sizeof(uint32_t));
    return 0;
}

extern "C" int spl_gpu_get_counters(const spl_gpu_ctx *c, spl_gpu_counters *out) {
    if (!c || !out) return fail(-2, EINVAL, "null argument");
    if (out->struct_size != sizeof(spl_gpu_counters)) return fail(-2, EINVAL, "spl_gpu_counters size mismatch");
    out->batch_queries = c->n_batch_queries.load();
    out->batch_unproven = c->n_batch_unproven.load();
    out->exact_unproven = c->n_exact_unproven.load();
    out->stale_rows_found = c->n_stale_rows.load();
    out->batch_rescored_extra = c->n_batch_extra.load();
    return 0;
}

extern "C" uint64_t spl_gpu_launch_count(const spl_gpu_ctx *c) { return c ? c->launches.load() : 0; }
extern "C" int spl_gpu_set_timing(spl_gpu_ctx *c, int on) {
    if (!c) return fail(-2, EINVAL, "null context");
    c->timing = on != 0;
    return 0;
}
extern "C" double spl_gpu_last_scan_ms(const spl_gpu_ctx *c) { return c ? c->last_scan_ms : -1.0; }

// ------------------------------------------------------------------------------------------------
// sync
// ------------------------------------------------------------------------------------------------
struct SyncCounters {
    std::atomic<uint64_t> rows{0}, meta{0}, busy{0}, aba{0}, bytes{0}, reval{0}, stale{0};
    std::atomic<int> err{0};
};

// ---- GPU-side ordering of mirror updates against searches ---------------------------------------------------------
// The reference's reader never returns a half-written slot (seqlock, splinter.c:598-605).  On the device the same
// promise needs stream order: a scatter kernel must not rewrite a row (its fp32 copy, bf16 shadow, norm, flags,
// bloom) while a scan on ANOTHER stream is reading it, and the next scan must see the scatter complete.  All three
// helpers are called with c->mu held.  A stream that is being captured into a CUDA graph is left alone: events
// recorded there would not order anything outside the graph, so such a caller orders graph launches against
// spl_gpu_sync itself.
static bool stream_capturing(cudaStream_t s) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &st) != cudaSuccess) { cudaGetLastError(); return false; }
    return st != cudaStreamCaptureStatusNone;
}

// before a mirror update is enqueued on c->stream: wait for the last search if it ran elsewhere
static int order_before_update(spl_gpu_ctx *c) {
    if (c->search_recorded && c->search_stream != c->stream) CUDA_TRY(cudaStreamWaitEvent(c->stream, c->ev_search, 0));
    return 0;
}
static int order_after_update(spl_gpu_ctx *c) {
    CUDA_TRY(cudaEventRecord(c->ev_flush, c->stream));
    c->flush_recorded = true;
    return 0;
}
// before a search is enqueued on s: wait for the last mirror update (if s is not the update stream) and for the
// last search on a different stream (they share the context's scratch: candidate lists, counters, batch lists)
static int order_before_search(spl_gpu_ctx *c, cudaStream_t s) {
    if (stream_capturing(s)) return 0;
    if (s != c->stream && c->flush_recorded) CUDA_TRY(cudaStreamWaitEvent(s, c->ev_flush, 0));
    if (c->search_recorded && c->search_stream != s) CUDA_TRY(cudaStreamWaitEvent(s, c->ev_search, 0));
    if (c->counters_dirty) {
        CUDA_TRY(cudaMemsetAsync(c->d_counters, 0, 32 * sizeof(uint32_t), s));
        c->counters_dirty = false;
        c->last_was_tma = false;
    }
    return 0;
}
static int order_after_search(spl_gpu_ctx *c, cudaStream_t s) {
    if (stream_capturing(s)) return 0;
    CUDA_TRY(cudaEventRecord(c->ev_search, s));
    c->search_recorded = true;
    c->search_stream = s;
    return 0;
}

static int flush_buf(spl_gpu_ctx *c, StageBuf &b, bool rows, SyncCounters &sc) {
    if (b.n == 0) return 0;
    std::lock_guard<std::mutex> lk(c->mu);
    const size_t row_bytes = (size_t)c->dim * 4;
    c->last_was_tma = false;
    if (order_before_update(c)) return -1;
    if (rows) {
        CUDA_TRY(cudaMemcpyAsync(b.d_rows, b.h_rows, (size_t)b.n * row_bytes, cudaMemcpyHostToDevice, c->stream));
        CUDA_TRY(cudaMemcpyAsync(b.d_meta, b.h_meta, (size_t)b.n * sizeof(StagedMeta), cudaMemcpyHostToDevice, c->stream));
        CUDA_TRY(launch_scatter_rows(c->m, b.d_rows, b.d_meta, b.n, c->sm_count, c->stream));
        sc.bytes += (uint64_t)b.n * (row_bytes + sizeof(StagedMeta));
    } else {
        CUDA_TRY(cudaMemcpyAsync(b.d_meta, b.h_meta, (size_t)b.n * sizeof(StagedMeta), cudaMemcpyHostToDevice, c->stream));
        CUDA_TRY(launch_scatter_meta(c->m, b.d_meta, b.n, c->stream));
        sc.bytes += (uint64_t)b.n * sizeof(StagedMeta);
    }
    c->launches++;
    CUDA_TRY(cudaEventRecord(b.ev, c->stream));
    if (order_after_update(c)) return -1;
    b.in_flight = true;
    b.n = 0;
    return 0;
}

static int acquire_buf(StageBuf &b) {
    if (b.in_flight) {
        CUDA_TRY(cudaEventSynchronize(b.ev));
        b.in_flight = false;
    }
    b.n = 0;
    return 0;
}

static inline uint64_t ld_acq64(const uint8_t *p) { return __atomic_load_n(reinterpret_cast<const uint64_t *>(p), __ATOMIC_ACQUIRE); }
static inline uint32_t ld_acq32(const uint8_t *p) { return __atomic_load_n(reinterpret_cast<const uint32_t *>(p), __ATOMIC_ACQUIRE); }

static inline uint8_t validity_bits(uint64_t hash, uint64_t epoch, uint32_t val_len, uint8_t key0) {
    uint8_t f = 0;
    if (hash && epoch && key0) {       // enumerate_matches (splinter.c:982) + _search.c:366,375
        f |= SPL_ROW_ENUM;
        if (val_len > 0) f |= SPL_ROW_LIST;  // splinter_list (splinter.c:470-471)
    }
    return f;
}

// 64-bit fingerprint of the first cache line of an embedding (16 floats)
static inline uint64_t row_fingerprint(const uint8_t *p) {
    const uint64_t *w = reinterpret_cast<const uint64_t *>(p);
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < 8; i++) { h ^= w[i]; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 29; }
    return h;
}

// one pass over a row: 64-bit hash of all its bytes + "is it all zero"
static inline uint64_t row_hash(const uint8_t *p, size_t bytes, bool *is_zero) {
    const uint64_t *w = reinterpret_cast<const uint64_t *>(p);
    uint64_t h0 = 0x9E3779B97F4A7C15ull, h1 = 0xD1342543DE82EF95ull, acc = 0;
    size_t i = 0;
    for (; i + 1 < bytes / 8; i += 2) {   // two independent chains keep the multiplier pipelined
        acc |= w[i] | w[i + 1];
        h0 = (h0 ^ w[i]) * 0xBF58476D1CE4E5B9ull; h0 ^= h0 >> 32;
        h1 = (h1 ^ w[i + 1]) * 0x94D049BB133111EBull; h1 ^= h1 >> 32;
    }
    for (; i < bytes / 8; i++) { acc |= w[i]; h0 = (h0 ^ w[i]) * 0xBF58476D1CE4E5B9ull; h0 ^= h0 >> 32; }
    if (is_zero) *is_zero = acc == 0;
    return h0 ^ (h1 * 0x9E3779B97F4A7C15ull);
}

// Every sweep re-validates 1/kRevalEvery of the rows it believes clean by hashing the whole row in the store and
// comparing with the hash taken when the row was mirrored.  That bounds the life of a stale row that slipped past
// both {hash, epoch} and the first-line fingerprint (a rewrite that restores the epoch and keeps the first 16 floats)
// to kRevalEvery sweeps, at a cost of n/kRevalEvery row reads per sweep.
constexpr uint64_t kRevalEvery = 1024;

// Sweep threads pull fixed-size slot ranges from a shared counter: on a host whose cores are also running
// the store's writers a static split would wait for whichever thread shares a core with one of them.
static inline uint64_t sweep_chunk(const spl_gpu_ctx *c) {
    const uint64_t per = c->n / (8ull * (c->sync_threads ? c->sync_threads : 1));
    return per < 256 ? 256 : (per > 4096 ? 4096 : per);
}

static void sync_worker(spl_gpu_ctx *c, unsigned tid, std::atomic<uint64_t> *next_chunk, bool header_moved,
                        uint64_t reval_phase, SyncCounters *scp) {
    using namespace layout;
    SyncCounters &sc = *scp;
    cudaSetDevice(c->device);
    StageBuf *rb = &c->stage[(size_t)tid * 4], *mb = &c->stage[(size_t)tid * 4 + 2];
    int rcur = 0, mcur = 0;
    if (acquire_buf(rb[0]) || acquire_buf(mb[0])) { sc.err = 1; return; }
    const size_t row_bytes = (size_t)c->dim * 4;
    const uint64_t chunk = sweep_chunk(c);
    for (;;) {
    const uint64_t lo = next_chunk->fetch_add(1, std::memory_order_relaxed) * chunk;
    if (lo >= c->n) break;
    const uint64_t hi = lo + chunk < c->n ? lo + chunk : c->n;
    for (uint64_t r = lo; r < hi; r++) {
        const uint8_t *sl = c->base + kHeaderSize + (c->slot_begin + r) * c->slot_size;
        // slots are 128 + 4*dim bytes apart, so every slot's metadata is its own cache line (and, on a big store, its
        // own TLB entry): the hardware prefetcher does not follow the stride, ask for the lines a few slots ahead
        if (r + 12 < hi) {
            __builtin_prefetch(sl + 12 * c->slot_size, 0, 0);
            __builtin_prefetch(sl + 12 * c->slot_size + S_KEY, 0, 0);
            if (header_moved) __builtin_prefetch(sl + 12 * c->slot_size + S_EMBED, 0, 0);   // the fingerprint's line
        }
        const uint64_t h = ld_acq64(sl + S_HASH);
        const uint64_t e1 = ld_acq64(sl + S_EPOCH);
        const uint8_t st = c->c_state[r];
        bool row_dirty = !(st & 1) || h != c->c_hash[r] || e1 != c->c_epoch[r];
        if (!(st & 1) && h == 0 && e1 == 0) {
            // never-used slot on a never-mirrored row: the zero-initialised mirror is already right
            c->c_state[r] = 1;
            continue;
        }
        if (!row_dirty && header_moved && h != 0) {
            // Epoch ABA: splinter_retrain_slot forces the epoch to 4 from anywhere (splinter.c:807-825) and
            // later writes walk it up again, so an unchanged {hash, epoch} does NOT prove an unchanged row
            // once H->epoch has moved.  Compare a fingerprint of the row's first cache line (it sits right
            // behind the metadata line this sweep reads anyway); a scrub or a re-embedding changes it.
            if (row_fingerprint(sl + S_EMBED) != c->c_fp[r]) { row_dirty = true; sc.aba++; }
        }
        if (!row_dirty && h != 0 && !(e1 & 1) && ((r ^ reval_phase) & (kRevalEvery - 1)) == 0) {
            sc.reval++;
            const uint64_t fh = row_hash(sl + S_EMBED, row_bytes, nullptr);
            std::atomic_thread_fence(std::memory_order_acquire);
            // (an epoch that moved while hashing is an ordinary dirty row for the next sweep)
            if (ld_acq64(sl + S_EPOCH) == e1 && fh != c->c_full[r]) { row_dirty = true; sc.stale++; }
        }
        if (!row_dirty) {
            // label changes move no slot epoch (splinter.c:843): compare bloom + validity bits
            const uint64_t b = ld_acq64(sl + S_BLOOM);
            const uint8_t f = validity_bits(h, e1, ld_acq32(sl + S_VAL_LEN), sl[S_KEY]) | (c->c_flags[r] & SPL_ROW_EMB);
            if (b != c->c_bloom[r] || f != c->c_flags[r]) {
                StageBuf &m = mb[mcur];
                m.h_meta[m.n++] = StagedMeta{ b, e1, (uint32_t)r, f };
                c->c_bloom[r] = b; c->c_flags[r] = f;
                sc.meta++;
                if (m.n == m.cap) {
                    if (flush_buf(c, m, false, sc)) { sc.err = 1; return; }
                    mcur ^= 1;
                    if (acquire_buf(mb[mcur])) { sc.err = 1; return; }
                }
            }
            continue;
        }
        // ---- seqlock-validated row copy ----
        StageBuf &b = rb[rcur];
        float *dst = b.h_rows + (size_t)b.n * c->dim;
        bool ok = false;
        uint64_t es = e1, hh = h, bloom = 0;
        uint32_t vl = 0; uint8_t key0 = 0;
        for (int attempt = 0; attempt < 4 && !ok; attempt++) {
            if (attempt) { es = ld_acq64(sl + S_EPOCH); }
            if (es & 1) continue;                       // writer active: splinter.c:599 returns EAGAIN
            memcpy(dst, sl + S_EMBED, row_bytes);
            hh = ld_acq64(sl + S_HASH);
            bloom = ld_acq64(sl + S_BLOOM);
            vl = ld_acq32(sl + S_VAL_LEN);
            key0 = sl[S_KEY];
            std::atomic_thread_fence(std::memory_order_acquire);
            ok = ld_acq64(sl + S_EPOCH) == es;          // splinter.c:602-603
            // Hardening beyond the reference's own reader.  The epoch is not monotonic: retrain forces it to 4
            // from anywhere (splinter.c:807-825) and unset + set restart it, so a reader that is descheduled in
            // the middle of its copy can come back to the SAME even epoch it started with after any number of
            // writes (seen under the mixed-workload stress test).  Re-compare the staged bytes with the slot
            // and re-check the epoch: a write that overlapped the copy changes one or the other.
            if (ok) ok = memcmp(dst, sl + S_EMBED, row_bytes) == 0 && ld_acq64(sl + S_EPOCH) == es;
        }
        if (!ok) {
            sc.busy++;
            if (!(st & 1)) {
                // never mirrored and busy right now: publish validity without an embedding so the
                // slot scores like a failed splinter_get_embedding (_search.c:388-391)
                const uint8_t f = validity_bits(h, e1 | 1, ld_acq32(sl + S_VAL_LEN), sl[S_KEY]);
                StageBuf &m = mb[mcur];
                m.h_meta[m.n++] = StagedMeta{ ld_acq64(sl + S_BLOOM), es, (uint32_t)r, f };
                if (m.n == m.cap) {
                    if (flush_buf(c, m, false, sc)) { sc.err = 1; return; }
                    mcur ^= 1;
                    if (acquire_buf(mb[mcur])) { sc.err = 1; return; }
                }
            }
            continue;  // keep the previous mirror row and the previous cache: retried next sync
        }
        const uint8_t f = validity_bits(hh, es, vl, key0) | SPL_ROW_EMB;
        b.h_meta[b.n] = StagedMeta{ bloom, es, (uint32_t)r, f };
        b.n++;
        c->c_hash[r] = hh; c->c_epoch[r] = es; c->c_bloom[r] = bloom; c->c_flags[r] = f;
        bool zero_row = false;
        c->c_full[r] = row_hash(reinterpret_cast<const uint8_t *>(dst), row_bytes, &zero_row);
        c->c_state[r] = 1 | (zero_row ? 0 : 2);
        c->c_fp[r] = row_fingerprint(reinterpret_cast<const uint8_t *>(dst));
        sc.rows++;
        if (b.n == b.cap) {
            if (flush_buf(c, b, true, sc)) { sc.err = 1; return; }
            rcur ^= 1;
            if (acquire_buf(rb[rcur])) { sc.err = 1; return; }
        }
    }
    }
    if (flush_buf(c, rb[rcur], true, sc) || flush_buf(c, mb[mcur], false, sc)) sc.err = 1;
}

extern "C" int spl_gpu_sync(spl_gpu_ctx *c, spl_gpu_sync_stats *stats) {
    if (!c) return fail(-2, EINVAL, "null context");
    if (c->dense) return fail(-2, EINVAL, "dense contexts have no backing store to sync");
    std::lock_guard<std::mutex> sweep_lock(c->sweep_mu);  // per context: several GPUs' sweeps run side by side
    auto t0 = std::chrono::steady_clock::now();
    const uint64_t hdr_epoch = ld_acq64(c->base + layout::H_EPOCH);
    const bool header_moved = c->first_sync || hdr_epoch != c->c_header_epoch;
    SyncCounters sc;
    const unsigned T = c->sync_threads;
    std::vector<std::thread> th;
    std::atomic<uint64_t> next_chunk{0};
    const uint64_t reval_phase = c->sweep_no++;
    for (unsigned t = 0; t < T; t++) th.emplace_back(sync_worker, c, t, &next_chunk, header_moved, reval_phase, &sc);
    for (auto &x : th) x.join();
    auto t1 = std::chrono::steady_clock::now();
    if (sc.err.load()) return fail(-1, EIO, "sync worker failed while staging rows to the device");
    cudaSetDevice(c->device);
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    for (StageBuf &b : c->stage) b.in_flight = false;
    auto t2 = std::chrono::steady_clock::now();
    c->c_header_epoch = hdr_epoch;
    c->first_sync = false;
    const uint64_t gen = ++c->generation;
    if (stats) {
        memset(stats, 0, sizeof *stats);
        stats->generation = gen;
        stats->slots_swept = c->n;
        stats->rows_uploaded = sc.rows;
        stats->meta_updated = sc.meta;
        stats->skipped_busy = sc.busy;
        stats->aba_rechecked = sc.aba;
        stats->bytes_h2d = sc.bytes;
        stats->sweep_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        stats->upload_ms = std::chrono::duration<double, std::milli>(t2 - t1).count();
        stats->header_epoch = hdr_epoch;
        stats->rows_revalidated = sc.reval;
        stats->stale_rows_found = sc.stale;
    }
    c->n_stale_rows += sc.stale.load();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// wake-up: the reference's event bus as a hint, H->epoch as the gate
// ------------------------------------------------------------------------------------------------
static void bus_attach(spl_gpu_ctx *c) {
    using namespace layout;
    const int32_t fd = (int32_t)ld_acq32(c->base + H_BUS_OWNER_FD), pid = (int32_t)ld_acq32(c->base + H_BUS_OWNER_PID);
    if (fd == c->bus_owner_fd && pid == c->bus_owner_pid && c->bus_fd >= 0) return;  // same bus as before
    if (c->bus_fd >= 0) { close(c->bus_fd); c->bus_fd = -1; }
    c->bus_owner_fd = fd; c->bus_owner_pid = pid;
    if (fd < 0 || pid <= 0) return;  // splinter_event_bus_init never ran (splinter.c:991-999): timed checks only
    if ((pid_t)pid == getpid()) { c->bus_fd = dup(fd); return; }  // splinter.c:1009-1010
#if defined(SYS_pidfd_open) && defined(SYS_pidfd_getfd)
    const int pidfd = (int)syscall(SYS_pidfd_open, (pid_t)pid, 0);  // splinter.c:1013-1018
    if (pidfd < 0) return;
    c->bus_fd = (int)syscall(SYS_pidfd_getfd, pidfd, fd, 0);
    close(pidfd);
    if (c->bus_fd >= 0) {  // the owner may have died and its pid/fd numbers been reused: accept eventfds only
        char link[64], target[64];
        snprintf(link, sizeof link, "/proc/self/fd/%d", c->bus_fd);
        const ssize_t n = readlink(link, target, sizeof target - 1);
        if (n <= 0 || strncmp(target, "anon_inode:[eventfd", 19) != 0) { close(c->bus_fd); c->bus_fd = -1; }
    }
#endif
}

extern "C" int spl_gpu_wait(spl_gpu_ctx *c, uint32_t timeout_ms) {
    using namespace layout;
    if (!c) return fail(-2, EINVAL, "null context");
    if (c->dense) return fail(-2, EINVAL, "dense contexts have no backing store to wait on");
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
    constexpr int kSliceMs = 2;  // how stale a write by a process that does not own the bus can get
    for (;;) {
        if (ld_acq64(c->base + H_EPOCH) != c->c_header_epoch || c->first_sync) return 1;
        const auto now = std::chrono::steady_clock::now();
        if (now >= deadline) return 0;
        long left = (long)std::chrono::duration_cast<std::chrono::milliseconds>(deadline - now).count();
        const int slice = left < kSliceMs ? (int)(left > 0 ? left : 1) : kSliceMs;
        bus_attach(c);
        if (c->bus_fd >= 0) {
            struct pollfd pfd = { c->bus_fd, POLLIN, 0 };
            const int pr = poll(&pfd, 1, slice);
            if (pr > 0 && (pfd.revents & POLLIN)) {
                // consume the count like splinter_event_bus_wait does (splinter.c:1032), but without blocking if
                // another watcher drained it first and without touching the shared O_NONBLOCK file flag
                uint64_t v;
                struct iovec iov = { &v, sizeof v };
                if (preadv2(c->bus_fd, &iov, 1, -1, RWF_NOWAIT) < 0 && errno != EAGAIN) { close(c->bus_fd); c->bus_fd = -1; }
            } else if (pr > 0) {  // POLLERR / POLLHUP: the bus is gone
                close(c->bus_fd); c->bus_fd = -1;
            }
        } else {
            struct timespec ts = { 0, slice * 1000000L };
            nanosleep(&ts, nullptr);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// dense shards
// ------------------------------------------------------------------------------------------------
extern "C" int spl_gpu_dense_upload(spl_gpu_ctx *c, uint64_t first_row, uint64_t n_rows, const float *E,
                                    const uint64_t *bloom, const uint8_t *flags, const uint64_t *epoch) {
    if (!c || !E) return fail(-2, EINVAL, "null argument");
    if (!c->dense) return fail(-2, EINVAL, "store-backed contexts are filled by spl_gpu_sync");
    if (first_row + n_rows > c->n) return fail(-2, EINVAL, "row range beyond the shard");
    cudaSetDevice(c->device);
    SyncCounters sc;
    StageBuf *rb = &c->stage[0];
    int cur = 0;
    if (acquire_buf(rb[0])) return -1;
    const size_t row_bytes = (size_t)c->dim * 4;
    for (uint64_t r = 0; r < n_rows; r++) {
        StageBuf &b = rb[cur];
        memcpy(b.h_rows + (size_t)b.n * c->dim, E + r * c->dim, row_bytes);
        b.h_meta[b.n] = StagedMeta{ bloom ? bloom[r] : 0, epoch ? epoch[r] : 2, (uint32_t)(first_row + r),
                                    flags ? flags[r] : (uint32_t)(SPL_ROW_LIST | SPL_ROW_ENUM | SPL_ROW_EMB) };
        if (++b.n == b.cap) {
            if (flush_buf(c, b, true, sc)) return -1;
            cur ^= 1;
            if (acquire_buf(rb[cur])) return -1;
        }
    }
    if (flush_buf(c, rb[cur], true, sc)) return -1;
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    rb[0].in_flight = rb[1].in_flight = false;
    ++c->generation;
    return 0;
}

extern "C" int spl_gpu_dense_fill_synthetic(spl_gpu_ctx *c, uint64_t seed_corpus, uint64_t seed_bloom) {
    if (!c) return fail(-2, EINVAL, "null context");
    if (!c->dense) return fail(-2, EINVAL, "only dense contexts can be filled synthetically");
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->d_synth_scale) {
        int rc = dev_alloc(c, &c->d_synth_scale, c->n, false);
        if (rc) return rc;
    }
    if (order_before_update(c)) return -1;
    CUDA_TRY(launch_synth(c->m, c->n, c->slot_begin, seed_corpus, seed_bloom, c->d_synth_scale, c->sm_count, c->stream));
    c->launches += 2;
    if (order_after_update(c)) return -1;
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    ++c->generation;
    return 0;
}

extern "C" int spl_gpu_read_mirror(spl_gpu_ctx *c, uint64_t first_slot, uint64_t n, float *E, uint64_t *bloom,
                                   uint8_t *flags, uint64_t *epoch) {
    if (!c) return fail(-2, EINVAL, "null context");
    if (first_slot < c->slot_begin || first_slot + n > c->slot_begin + c->n) return fail(-2, EINVAL, "range outside the shard");
    const uint64_t r0 = first_slot - c->slot_begin;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    if (E) CUDA_TRY(cudaMemcpy(E, c->m.E + r0 * c->dim, n * c->dim * 4, cudaMemcpyDeviceToHost));
    if (bloom) CUDA_TRY(cudaMemcpy(bloom, c->m.bloom + r0, n * 8, cudaMemcpyDeviceToHost));
    if (flags) CUDA_TRY(cudaMemcpy(flags, c->m.flags + r0, n, cudaMemcpyDeviceToHost));
    if (epoch) CUDA_TRY(cudaMemcpy(epoch, c->m.epoch + r0, n * 8, cudaMemcpyDeviceToHost));
    return 0;
}

// ------------------------------------------------------------------------------------------------
// search
// ------------------------------------------------------------------------------------------------
static int check_params(const spl_gpu_ctx *c, const spl_gpu_search_params *p) {
    if (!c || !p) return fail(-2, EINVAL, "null argument");
    if (p->struct_size != sizeof(spl_gpu_search_params)) return fail(-2, EINVAL, "spl_gpu_search_params size mismatch");
    if (p->dim != c->dim) return fail(-2, EINVAL, "query dim %u != mirror dim %u", p->dim, c->dim);
    if (p->nq == 0 || p->nq > kMaxBatchQ) return fail(-2, EINVAL, "nq must be in 1..%u", kMaxBatchQ);
    if (p->order > SPL_ORDER_L2) return fail(-2, EINVAL, "unknown order %u", p->order);
    if (p->mode > SPL_MODE_BATCH_BF16) return fail(-2, EINVAL, "unknown mode %u", p->mode);
    return 0;
}

static void fill_args(const spl_gpu_ctx *c, const spl_gpu_search_params *p, ScanArgs &a) {
    memset(&a, 0, sizeof a);
    a.E = c->m.E; a.bloom = c->m.bloom; a.flags = c->m.flags; a.epoch = c->m.epoch;
    a.n = (uint32_t)c->n; a.dim = c->dim; a.k = p->k; a.order = p->order;
    a.mask = p->bloom_mask; a.min_sim = p->min_similarity; a.max_dist = p->max_distance;
    a.exclude = kIdxSentinel;
    if (p->exclude_slot != SPL_NO_SLOT && p->exclude_slot >= c->slot_begin && p->exclude_slot < c->slot_begin + c->n)
        a.exclude = (uint32_t)(p->exclude_slot - c->slot_begin);
    a.row0 = c->slot_begin;
    a.cta_cands = c->d_cands_set[c->scan_set]; a.counters = c->d_counters_set[c->scan_set];
    a.dbg = c->d_dbg;
}

// enqueue one fused top-k scan per query (k in 1..KMAX_FUSED); caller holds c->mu
static int enqueue_fused(spl_gpu_ctx *c, const spl_gpu_search_params *p, const float *d_q, spl_gpu_result *d_out,
                         uint32_t out_stride, uint32_t *d_counts, cudaStream_t s, uint32_t *d_totals = nullptr,
                         uint32_t count_stride = 1) {
    uint32_t mode = p->mode == SPL_MODE_AUTO ? SPL_MODE_FAST : p->mode;
    if (mode != SPL_MODE_FAST) return fail(-1, ENOSYS, "mode %u is not available for this query shape", mode);
    ScanArgs a;
    fill_args(c, p, a);
    a.k = p->k;
    const bool tma = !c->force_ldg && scan_tma_supported(a.dim, a.k);
    a.out_stride = out_stride;
    a.count_stride = count_stride;
    for (uint32_t qi = 0; qi < p->nq;) {
        // small batches: 4 (then 2) queries share one pass over the store, so each row read from HBM is scored
        // for all of them; the rest go one per launch
        uint32_t g = 1;
        if (tma && !c->no_multi) {
            // eight per pass where only similarities have to be scored for every row (the kernel's lazy loop body)
            if (p->nq - qi >= 8 && !c->no_nq8 && a.order == SPL_ORDER_COSINE && !(a.max_dist > 0.f) &&
                scan_tma_multi_supported(a.dim, a.k, 8)) g = 8;
            else if (p->nq - qi >= 4 && scan_tma_multi_supported(a.dim, a.k, 4)) g = 4;
            else if (p->nq - qi >= 2 && scan_tma_multi_supported(a.dim, a.k, 2)) g = 2;
        }
        a.nq = g;
        a.q = d_q + (size_t)qi * c->dim;
        a.out = d_out + (size_t)qi * out_stride;
        a.out_count = d_counts + (size_t)qi * count_stride;
        a.total = d_totals ? d_totals + qi : nullptr;
        // alternate scratch sets from launch to launch; a TMA scan that directly follows another TMA scan of this
        // context on the same stream may start during its predecessor's tail: always inside one call, across calls
        // only when the caller says nothing else sits between them (SPL_SEARCH_PIPELINED)
        a.cta_cands = c->d_cands_set[c->scan_set]; a.counters = c->d_counters_set[c->scan_set];
        c->scan_set ^= 1u;
        if (tma) {
            a.overlap_prev = c->last_was_tma && c->last_tma_stream == s && !c->no_overlap &&
                             (qi > 0 || (p->flags & SPL_SEARCH_PIPELINED)) ? 1u : 0u;
            CUDA_TRY(launch_scan_tma(a, c->sm_count, s));
            c->last_was_tma = true; c->last_tma_stream = s;
        } else {
            CUDA_TRY(launch_scan_fast(a, c->scan_grid, s));
            c->last_was_tma = false;
        }
        c->launches++;
        qi += g;
    }
    return 0;
}

// ---- SPL_MODE_EXACT ---------------------------------------------------------------------------
struct ExactSlices {
    spl_gpu_result *cand; uint32_t *cand_count; uint64_t *exact_key, *sort_key; uint32_t *sort_pos; uint64_t *kth;
    uint32_t *ticket;
};
static constexpr size_t kExactPerQuery = kExactMaxCand * (sizeof(spl_gpu_result) + 8 + 8 + 4) + 64;

static int ensure_exact_scratch(spl_gpu_ctx *c) {
    if (c->d_exact) return 0;
    int rc;
    if ((rc = dev_alloc(c, &c->d_exact, kExactPerQuery * c->q_cap, true))) return rc;
    CUDA_TRY(cudaHostAlloc((void **)&c->h_xvalid, c->q_cap * sizeof(uint32_t), cudaHostAllocMapped));
    CUDA_TRY(cudaHostGetDevicePointer((void **)&c->m_xvalid, c->h_xvalid, 0));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

static ExactSlices exact_slices(spl_gpu_ctx *c, uint32_t qi) {
    unsigned char *p = c->d_exact + kExactPerQuery * qi;
    ExactSlices s;
    s.cand = reinterpret_cast<spl_gpu_result *>(p);            p += kExactMaxCand * sizeof(spl_gpu_result);
    s.exact_key = reinterpret_cast<uint64_t *>(p);             p += kExactMaxCand * 8;
    s.sort_key = reinterpret_cast<uint64_t *>(p);              p += kExactMaxCand * 8;
    s.sort_pos = reinterpret_cast<uint32_t *>(p);              p += kExactMaxCand * 4;
    s.kth = reinterpret_cast<uint64_t *>(p);                   p += 8;
    s.cand_count = reinterpret_cast<uint32_t *>(p);            p += 4;
    s.ticket = reinterpret_cast<uint32_t *>(p);
    return s;
}

// fast scan with k' candidates (thresholds relaxed by the fast-vs-sequential bound), then the exact
// re-score + proof.  nq <= q_cap queries; caller holds c->mu.
static int enqueue_exact(spl_gpu_ctx *c, const spl_gpu_search_params *p, const float *d_q, uint32_t nq, uint32_t kprime,
                         spl_gpu_result *d_out, uint32_t out_stride, uint32_t *d_counts, uint32_t *d_valid, cudaStream_t s,
                         uint32_t *d_totals = nullptr) {
    spl_gpu_search_params fastp = *p;
    fastp.mode = SPL_MODE_FAST;
    fastp.k = kprime;
    fastp.nq = nq;
    if (p->min_similarity > 0.f) fastp.min_similarity = p->min_similarity - kExactSimBound;
    if (p->max_distance > 0.f) fastp.max_distance = p->max_distance * (1.f + kExactDistRel);
    {   // candidate scans for all queries (small batches share passes), each into its own scratch slice
        static_assert(kExactPerQuery % sizeof(spl_gpu_result) == 0, "slices must be a whole number of records apart");
        const ExactSlices x0 = exact_slices(c, 0);
        int rc = enqueue_fused(c, &fastp, d_q, x0.cand, (uint32_t)(kExactPerQuery / sizeof(spl_gpu_result)), x0.cand_count, s,
                               d_totals, (uint32_t)(kExactPerQuery / sizeof(uint32_t)));
        if (rc) return rc;
    }
    for (uint32_t qi = 0; qi < nq; qi++) {
        const ExactSlices x = exact_slices(c, qi);
        ExactArgs a;
        memset(&a, 0, sizeof a);
        a.E = c->m.E; a.q = d_q + (size_t)qi * c->dim; a.dim = c->dim; a.order = p->order; a.k = p->k; a.kprime = kprime;
        a.row0 = c->slot_begin; a.min_sim = p->min_similarity; a.max_dist = p->max_distance;
        a.cand = x.cand; a.cand_count = x.cand_count; a.exact_key = x.exact_key; a.sort_key = x.sort_key;
        a.sort_pos = x.sort_pos; a.kth_key_out = x.kth; a.ticket = x.ticket;
        a.out = d_out + (size_t)qi * out_stride; a.out_count = d_counts + qi; a.valid = d_valid + qi;
        CUDA_TRY(launch_exact_rescore(a, s));
        c->launches++;
    }
    c->last_was_tma = false;
    return 0;
}

static uint32_t exact_kprime(uint32_t k) {
    if (k <= 16) return 32;  // stays on the scan's register/shuffle epilogue (k' <= 32)
    uint32_t kp = 2 * k > k + 32 ? 2 * k : k + 32;
    return kp > kExactMaxCand ? kExactMaxCand : kp;
}

static int ensure_batch_scratch(spl_gpu_ctx *c, uint32_t k) {
    int rc;
    if (!c->d_batch) {
        const size_t bytes = batch_scratch_bytes((uint32_t)c->n, kBatchMaxQ, c->dim, &c->batch_layout);
        if ((rc = dev_alloc(c, &c->d_batch, bytes, false))) return rc;
        if ((rc = dev_alloc(c, &c->d_qbatch, (size_t)kBatchMaxQ * c->dim, false))) return rc;
    }
    const uint32_t need = kBatchMaxQ * k;
    if (c->bout_cap < need) {
        cudaFree(c->d_bout); cudaFree(c->d_bcount); cudaFreeHost(c->h_bout); cudaFreeHost(c->h_bcount);
        c->d_bout = nullptr; c->d_bcount = nullptr; c->h_bout = nullptr; c->h_bcount = nullptr;   // a failed re-allocation
        c->bout_cap = 0;                                                                        // must not leave them dangling
        if ((rc = dev_alloc(c, &c->d_bout, need, false))) return rc;
        if ((rc = dev_alloc(c, &c->d_bcount, 2 * kBatchMaxQ, false))) return rc;
        CUDA_TRY(cudaMallocHost((void **)&c->h_bout, (size_t)need * sizeof(spl_gpu_result)));
        CUDA_TRY(cudaMallocHost((void **)&c->h_bcount, (size_t)2 * kBatchMaxQ * sizeof(uint32_t)));
        c->bout_cap = need;
    }
    return 0;
}

// The batched path re-scores a fixed number of candidates chosen by the ordering score, so it can honour a
// threshold on THAT score (the failing rows are the tail of the result) but not one on the other score.
static bool batch_thresholds_ok(const spl_gpu_search_params *p) {
    return p->order == SPL_ORDER_COSINE ? !(p->max_distance > 0.f) : !(p->min_similarity > 0.f);
}

static bool wants_batch(const spl_gpu_ctx *c, const spl_gpu_search_params *p) {
    if (p->mode == SPL_MODE_BATCH_BF16) return true;
    // The batched pass streams the whole bf16 shadow whatever the mask (~0.35 ms per 2^20 x 1024 rows, up to 128
    // queries); the fp32 scan reads only the rows that pass.  Unfiltered, four queries already cost the scan twice
    // that; with a label mask of unknown selectivity the scan keeps the small batches.
    const uint32_t from = p->bloom_mask == 0 ? 4u : 16u;
    return p->mode == SPL_MODE_AUTO && p->nq >= from && c->m.E16 && batch_supported(c->dim, p->k) && batch_thresholds_ok(p);
}

static int check_batch(const spl_gpu_ctx *c, const spl_gpu_search_params *p) {
    if (!c->m.E16) return fail(-1, ENOSYS, "SPL_MODE_BATCH_BF16 is not available: attach with SPL_OPT_BF16_MIRROR");
    if (!batch_supported(c->dim, p->k)) return fail(-1, ENOSYS, "SPL_MODE_BATCH_BF16 is not available for dim %u / k %u (dim %% 64 == 0, 1 <= k <= 128)", c->dim, p->k);
    if (!batch_thresholds_ok(p))
        return fail(-1, ENOSYS, "SPL_MODE_BATCH_BF16 is not available with a threshold on the score that does not order the result "
                                "(cosine order takes --similarity, L2 order takes --distance)");
    return 0;
}

// enqueue one batched search of nq <= kBatchMaxQ queries (fp32, on the device); caller holds c->mu
static int enqueue_batch(spl_gpu_ctx *c, const spl_gpu_search_params *p, const float *d_q, uint32_t nq, spl_gpu_result *d_out,
                         uint32_t out_stride, uint32_t *d_counts, uint32_t **d_overflow, cudaStream_t s) {
    BatchArgs b;
    memset(&b, 0, sizeof b);
    b.E = c->m.E; b.E16 = c->m.E16; b.norm2 = c->m.norm2; b.err2 = c->m.err2; b.norm2_max = c->m.norm2_max; b.bloom = c->m.bloom; b.flags = c->m.flags; b.epoch = c->m.epoch;
    b.n = (uint32_t)c->n; b.dim = c->dim; b.nq = nq; b.order = p->order; b.mask = p->bloom_mask;
    b.min_sim = p->min_similarity; b.max_dist = p->max_distance;
    b.exclude = kIdxSentinel;
    if (p->exclude_slot != SPL_NO_SLOT && p->exclude_slot >= c->slot_begin && p->exclude_slot < c->slot_begin + c->n)
        b.exclude = (uint32_t)(p->exclude_slot - c->slot_begin);
    b.row0 = c->slot_begin;
    uint32_t launches = 0;
    c->last_was_tma = false;
    CUDA_TRY(batch_search(b, c->batch_layout, c->d_batch, d_q, p->k, c->sm_count, d_out, out_stride, d_counts, d_overflow,
                          &launches, s));
    c->launches += launches;
    return 0;
}

// body of spl_gpu_search_device; c->mu held, stream order already established by the caller
static int search_device_locked(spl_gpu_ctx *c, const spl_gpu_search_params *p, const float *d_q, spl_gpu_result *d_out,
                                uint32_t out_stride, uint32_t *d_counts, cudaStream_t s) {
    int rc;
    if (p->mode == SPL_MODE_EXACT) {
        // no host round trip is possible here, so take the widest candidate margin at once; a query whose selection
        // still cannot be proven reports SPL_GPU_COUNT_UNANSWERED (the host-buffer call falls back to a full exact pass)
        if (p->k > kExactMaxK) return fail(-2, EINVAL, "device-buffer exact search needs k <= %u", kExactMaxK);
        if (p->nq > c->q_cap) return fail(-2, EINVAL, "device-buffer exact search takes at most %u queries per call", c->q_cap);
        if ((rc = ensure_exact_scratch(c))) return rc;
        return enqueue_exact(c, p, d_q, p->nq, kExactMaxCand, d_out, out_stride, d_counts, c->m_xvalid, s);
    }
    if (wants_batch(c, p)) {
        if ((rc = check_batch(c, p))) return rc;
        if (p->nq > kBatchMaxQ) return fail(-2, EINVAL, "device-buffer batched search takes at most %u queries per call", kBatchMaxQ);
        if ((rc = ensure_batch_scratch(c, p->k))) return rc;
        uint32_t *ovf = nullptr;
        // a query whose candidate list overflowed, or whose selection the bf16 scores cannot prove, reports
        // SPL_GPU_COUNT_UNANSWERED here; the host-buffer call re-runs it on the fp32 scan
        c->n_batch_queries += p->nq;
        return enqueue_batch(c, p, d_q, p->nq, d_out, out_stride, d_counts, &ovf, s);
    }
    return enqueue_fused(c, p, d_q, d_out, out_stride, d_counts, s);
}

extern "C" int spl_gpu_search_device(spl_gpu_ctx *c, const spl_gpu_search_params *p, const float *d_q,
                                     spl_gpu_result *d_out, uint32_t out_stride, uint32_t *d_counts, void *stream) {
    int rc = check_params(c, p);
    if (rc) return rc;
    if (!d_q || !d_out || !d_counts) return fail(-2, EINVAL, "null device buffer");
    if (p->k == 0 || p->k > SPL_GPU_KMAX_FUSED) return fail(-2, EINVAL, "device search needs k in 1..%u", SPL_GPU_KMAX_FUSED);
    if (out_stride < p->k) return fail(-2, EINVAL, "out_stride < k");
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    if ((rc = order_before_search(c, s))) return rc;
    rc = search_device_locked(c, p, d_q, d_out, out_stride, d_counts, s);
    if (rc) c->counters_dirty = true;   // a partly enqueued search may leave the work / ticket counters non-zero
    const int rc2 = order_after_search(c, s);   // also after a failure: whatever was enqueued still reads the mirror
    return rc ? rc : rc2;
}

static int ensure_sort_scratch(spl_gpu_ctx *c, uint32_t out_cap) {
    int rc;
    if (!c->d_keys) {
        if ((rc = dev_alloc(c, &c->d_keys, c->n, false))) return rc;
        c->sort_tmp_bytes = sortall_tmp_bytes((uint32_t)c->n);
        unsigned char *tmp = nullptr;
        if ((rc = dev_alloc(c, &tmp, c->sort_tmp_bytes, false))) return rc;
        c->d_sort_tmp = tmp;
    }
    if (c->big_out_cap < out_cap) {
        if (c->d_big_out) { cudaFree(c->d_big_out); c->d_big_out = nullptr; }
        if ((rc = dev_alloc(c, &c->d_big_out, out_cap, false))) return rc;
        c->big_out_cap = out_cap;
    }
    return 0;
}

// one query, every survivor scored and sorted on the device (the reference's default when --limit is absent);
// `exact` selects the sequential-order scorer.  Caller holds c->mu.
static int search_one_full(spl_gpu_ctx *c, const spl_gpu_search_params *p, const float *q_host, bool exact,
                           spl_gpu_result *out, uint32_t out_stride, uint32_t *count, uint32_t qi) {
    const uint32_t want = p->k == 0 ? out_stride : p->k;
    const uint32_t cap = want < c->n ? want : (uint32_t)c->n;
    int rc;
    if ((rc = ensure_sort_scratch(c, cap))) return rc;
    ScanArgs a;
    fill_args(c, p, a);
    memcpy(c->h_q, q_host, (size_t)c->dim * sizeof(float));
    CUDA_TRY(cudaMemcpyAsync(c->d_q, c->h_q, (size_t)c->dim * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(cudaMemsetAsync(c->d_keys, 0xFF, c->n * sizeof(uint64_t), c->stream));
    a.q = c->d_q;
    a.keys_out = c->d_keys;
    a.out = c->d_big_out;
    a.out_count = c->d_count;
    a.k = 0;
    c->last_was_tma = false;
    if (c->timing) CUDA_TRY(cudaEventRecord(c->ev0, c->stream));
    if (exact) {
        CUDA_TRY(launch_exact_all(a, c->sm_count, c->stream));
    } else if (!c->force_ldg && scan_tma_supported(a.dim, 1)) {
        // the TMA-staged scan in its key-per-row mode: same row traffic as the fused top-k, one 8-byte store per survivor
        a.k = 1; a.nq = 1;
        a.cta_cands = c->d_cands_set[c->scan_set]; a.counters = c->d_counters_set[c->scan_set];
        c->scan_set ^= 1u;
        CUDA_TRY(launch_scan_tma(a, c->sm_count, c->stream));
        a.k = 0;
    } else {
        CUDA_TRY(launch_scan_fast(a, c->scan_grid, c->stream));
    }
    if (c->timing) CUDA_TRY(cudaEventRecord(c->ev1, c->stream));
    CUDA_TRY(launch_sortall_select(a, c->d_keys, c->d_sort_tmp, c->sort_tmp_bytes, c->d_count, c->stream));
    CUDA_TRY(cudaMemcpyAsync(&c->h_bounce_count, c->d_count, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    if (c->timing) {
        float ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        c->last_scan_ms = ms;   // the scan alone; select + sort + copy-out come on top
    }
    const uint32_t total = c->h_bounce_count;
    const uint32_t ncopy = total < cap ? total : cap;
    CUDA_TRY(launch_sortall_sort(a, c->d_keys, c->d_sort_tmp, c->sort_tmp_bytes, total, cap, c->stream));
    c->launches += 5;
    CUDA_TRY(cudaMemcpyAsync(out, c->d_big_out, (size_t)ncopy * sizeof(spl_gpu_result), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    *count = p->k == 0 ? total : ncopy;
    c->h_total[qi] = total;  // the sorted pass counts every survivor
    return 0;
}

static int search_host_locked(spl_gpu_ctx *c, const spl_gpu_search_params *p, const float *q, spl_gpu_result *out,
                              uint32_t out_stride, uint32_t *counts);

// The fused scan writes its records straight into mapped host memory; its last CTA then sets a completion word.  Polling
// that word returns the answer a few microseconds before cudaStreamSynchronize would (which waits for the driver to
// notice the kernel's retirement).  A launch that faults never sets the word: after a bounded spin the stream is
// synchronised the ordinary way, which also surfaces the error.
static int wait_answer(spl_gpu_ctx *c, uint32_t expect) {
    const auto t0 = std::chrono::steady_clock::now();
    for (uint32_t spins = 0;; spins++) {
        if (__atomic_load_n(c->h_seq, __ATOMIC_ACQUIRE) == expect) return 0;
        if ((spins & 1023u) == 1023u && std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(20)) break;
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
    }
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int spl_gpu_search(spl_gpu_ctx *c, const spl_gpu_search_params *p, const float *q, spl_gpu_result *out,
                              uint32_t out_stride, uint32_t *counts) {
    int rc = check_params(c, p);
    if (rc) return rc;
    if (!q || !out || !counts) return fail(-2, EINVAL, "null buffer");
    if (p->k > 0 && out_stride < p->k) return fail(-2, EINVAL, "out_stride < k");
    if (p->k == 0 && out_stride == 0) return fail(-2, EINVAL, "out_stride must be > 0");
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    if ((rc = order_before_search(c, c->stream))) return rc;
    rc = search_host_locked(c, p, q, out, out_stride, counts);
    if (rc) {
        c->counters_dirty = true;
        order_after_search(c, c->stream);
        return rc;
    }
    // every path above ends with a cudaStreamSynchronize(c->stream): nothing of this search (nor of a caller-stream
    // search it waited for) is still in flight, so later updates and searches need not wait on it
    c->search_recorded = false;
    return 0;
}

static int search_host_locked(spl_gpu_ctx *c, const spl_gpu_search_params *p, const float *q, spl_gpu_result *out,
                              uint32_t out_stride, uint32_t *counts) {
    int rc;
    const size_t qbytes = (size_t)c->dim * sizeof(float);
    const bool fused = p->k >= 1 && p->k <= SPL_GPU_KMAX_FUSED;
    c->totals_valid = !wants_batch(c, p);
    if (wants_batch(c, p)) {
        if ((rc = check_batch(c, p))) return rc;
        if ((rc = ensure_batch_scratch(c, p->k))) return rc;
        c->n_batch_queries += p->nq;
        for (uint32_t q0 = 0; q0 < p->nq; q0 += kBatchMaxQ) {
            const uint32_t nq = (p->nq - q0 < kBatchMaxQ) ? p->nq - q0 : kBatchMaxQ;
            CUDA_TRY(cudaMemcpyAsync(c->d_qbatch, q + (size_t)q0 * c->dim, nq * qbytes, cudaMemcpyHostToDevice, c->stream));
            uint32_t *d_ovf = nullptr;
            if (c->timing) CUDA_TRY(cudaEventRecord(c->ev0, c->stream));
            if ((rc = enqueue_batch(c, p, c->d_qbatch, nq, c->d_bout, p->k, c->d_bcount, &d_ovf, c->stream))) return rc;
            if (c->timing) CUDA_TRY(cudaEventRecord(c->ev1, c->stream));
            CUDA_TRY(cudaMemcpyAsync(c->h_bout, c->d_bout, (size_t)nq * p->k * sizeof(spl_gpu_result), cudaMemcpyDeviceToHost, c->stream));
            CUDA_TRY(cudaMemcpyAsync(c->h_bcount, c->d_bcount, nq * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
            CUDA_TRY(cudaMemcpyAsync(c->h_bcount + kBatchMaxQ, d_ovf, nq * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
            CUDA_TRY(cudaStreamSynchronize(c->stream));
            if (c->timing) {
                float ms = 0.f;
                CUDA_TRY(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
                c->last_scan_ms = ms;
            }
            for (uint32_t i = 0; i < nq; i++) {
                c->n_batch_extra += c->h_bcount[kBatchMaxQ + i] >> 1;
                if (c->h_bcount[kBatchMaxQ + i] & 1u) {
                    // candidate list overflowed (threshold from the sample was too loose for this query) or the bf16
                    // scores could not prove the selection (a crowd of near-ties inside the bf16 error bound):
                    // answer it with the fp32 single-query scan instead -- still on the GPU, never on the CPU
                    c->n_batch_unproven++;
                    spl_gpu_search_params one = *p;
                    one.nq = 1; one.mode = SPL_MODE_FAST;
                    memcpy(c->h_q, q + (size_t)(q0 + i) * c->dim, qbytes);
                    CUDA_TRY(cudaMemcpyAsync(c->d_q, c->h_q, qbytes, cudaMemcpyHostToDevice, c->stream));
                    if ((rc = enqueue_fused(c, &one, c->d_q, c->m_out, p->k, c->m_count, c->stream))) return rc;
                    CUDA_TRY(cudaStreamSynchronize(c->stream));
                    counts[q0 + i] = c->h_count[0];
                    memcpy(out + (size_t)(q0 + i) * out_stride, c->h_out, (size_t)c->h_count[0] * sizeof(spl_gpu_result));
                } else {
                    counts[q0 + i] = c->h_bcount[i];
                    memcpy(out + (size_t)(q0 + i) * out_stride, c->h_bout + (size_t)i * p->k, (size_t)c->h_bcount[i] * sizeof(spl_gpu_result));
                }
            }
        }
        return 0;
    }
    const bool exact = p->mode == SPL_MODE_EXACT;
    if (exact && p->k >= 1 && p->k <= kExactMaxK) {
        if ((rc = ensure_exact_scratch(c))) return rc;
        const uint32_t kp0 = exact_kprime(p->k);
        for (uint32_t q0 = 0; q0 < p->nq; q0 += c->q_cap) {
            const uint32_t nq = (p->nq - q0 < c->q_cap) ? p->nq - q0 : c->q_cap;
            memcpy(c->h_q, q + (size_t)q0 * c->dim, nq * qbytes);
            CUDA_TRY(cudaMemcpyAsync(c->d_q, c->h_q, nq * qbytes, cudaMemcpyHostToDevice, c->stream));
            if (c->timing) CUDA_TRY(cudaEventRecord(c->ev0, c->stream));
            memset(c->h_total + q0, 0, nq * sizeof(uint32_t));
            if ((rc = enqueue_exact(c, p, c->d_q, nq, kp0, c->m_out, p->k, c->m_count, c->m_xvalid, c->stream, c->m_total + q0))) return rc;
            if (c->timing) CUDA_TRY(cudaEventRecord(c->ev1, c->stream));
            CUDA_TRY(cudaStreamSynchronize(c->stream));
            if (c->timing) {
                float ms = 0.f;
                CUDA_TRY(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
                c->last_scan_ms = ms / nq;
            }
            std::vector<uint32_t> redo;
            for (uint32_t i = 0; i < nq; i++) {
                if (c->h_xvalid[i]) {
                    counts[q0 + i] = c->h_count[i];
                    memcpy(out + (size_t)(q0 + i) * out_stride, c->h_out + (size_t)i * p->k, (size_t)c->h_count[i] * sizeof(spl_gpu_result));
                } else {
                    redo.push_back(i);
                }
            }
            // near-ties crowd the boundary: widen the candidate set once, then score every row exactly
            for (uint32_t i : redo) {
                bool ok = false;
                if (kp0 < kExactMaxCand) {
                    if ((rc = enqueue_exact(c, p, c->d_q + (size_t)i * c->dim, 1, kExactMaxCand, c->d_out, p->k, c->d_count,
                                            c->m_xvalid, c->stream))) return rc;
                    CUDA_TRY(cudaMemcpyAsync(c->h_bounce, c->d_out, (size_t)p->k * sizeof(spl_gpu_result), cudaMemcpyDeviceToHost, c->stream));
                    CUDA_TRY(cudaMemcpyAsync(&c->h_bounce_count, c->d_count, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
                    CUDA_TRY(cudaStreamSynchronize(c->stream));
                    if (c->h_xvalid[0]) {
                        ok = true;
                        counts[q0 + i] = c->h_bounce_count;
                        memcpy(out + (size_t)(q0 + i) * out_stride, c->h_bounce, (size_t)c->h_bounce_count * sizeof(spl_gpu_result));
                    }
                }
                if (!ok) {
                    c->n_exact_unproven++;
                    if ((rc = search_one_full(c, p, q + (size_t)(q0 + i) * c->dim, true, out + (size_t)(q0 + i) * out_stride,
                                              out_stride, &counts[q0 + i], q0 + i))) return rc;
                }
            }
        }
        return 0;
    }
    if (exact) {  // k == 0 (everything, sorted) or k > kExactMaxK: exact score of every row + device-wide sort
        for (uint32_t qi = 0; qi < p->nq; qi++)
            if ((rc = search_one_full(c, p, q + (size_t)qi * c->dim, true, out + (size_t)qi * out_stride, out_stride, &counts[qi], qi))) return rc;
        return 0;
    }
    if (fused) {
        // process in chunks of q_cap queries through the pinned staging buffers
        for (uint32_t q0 = 0; q0 < p->nq; q0 += c->q_cap) {
            const uint32_t nq = (p->nq - q0 < c->q_cap) ? p->nq - q0 : c->q_cap;
            spl_gpu_search_params pp = *p;
            pp.nq = nq;
            memset(c->h_total + q0, 0, nq * sizeof(uint32_t));  // mapped memory; ordered before the launch below
            uint32_t fmode = p->mode == SPL_MODE_AUTO ? SPL_MODE_FAST : p->mode;
            const bool polled = nq == 1 && fmode == SPL_MODE_FAST && !c->force_ldg && scan_tma_qparam_supported(c->dim, p->k);
            if (polled) {
                // one query: it rides inside the launch (kernel parameter space) -- no host-to-device copy of its own
                ScanArgs a;
                fill_args(c, &pp, a);
                a.q = nullptr;
                a.out = c->m_out; a.out_count = c->m_count; a.total = c->m_total + q0;
                a.out_stride = p->k; a.count_stride = 1;
                a.done_seq_ptr = c->m_seq; a.done_seq = ++c->seq;
                c->scan_set ^= 1u;
                if (c->timing) CUDA_TRY(cudaEventRecord(c->ev0, c->stream));
                CUDA_TRY(launch_scan_tma_qparam(a, q + (size_t)q0 * c->dim, c->sm_count, c->stream));
                c->last_was_tma = true; c->last_tma_stream = c->stream;
                c->launches++;
            } else {
                memcpy(c->h_q, q + (size_t)q0 * c->dim, nq * qbytes);
                CUDA_TRY(cudaMemcpyAsync(c->d_q, c->h_q, nq * qbytes, cudaMemcpyHostToDevice, c->stream));
                if (c->timing) CUDA_TRY(cudaEventRecord(c->ev0, c->stream));
                if ((rc = enqueue_fused(c, &pp, c->d_q, c->m_out, p->k, c->m_count, c->stream, c->m_total + q0))) return rc;
            }
            if (c->timing) CUDA_TRY(cudaEventRecord(c->ev1, c->stream));
            if (polled && !c->timing) { if ((rc = wait_answer(c, c->seq))) return rc; }
            else CUDA_TRY(cudaStreamSynchronize(c->stream));
            if (c->timing) {
                float ms = 0.f;
                CUDA_TRY(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
                c->last_scan_ms = ms / nq;
            }
            for (uint32_t i = 0; i < nq; i++) {
                counts[q0 + i] = c->h_count[i];
                memcpy(out + (size_t)(q0 + i) * out_stride, c->h_out + (size_t)i * p->k, (size_t)c->h_count[i] * sizeof(spl_gpu_result));
            }
        }
        return 0;
    }
    // ---- unlimited / large k: key-per-row scan + device-wide sort (cold path) ----
    uint32_t mode = p->mode == SPL_MODE_AUTO ? SPL_MODE_FAST : p->mode;
    if (mode != SPL_MODE_FAST) return fail(-1, ENOSYS, "mode %u is not available for k = %u", mode, p->k);
    for (uint32_t qi = 0; qi < p->nq; qi++)
        if ((rc = search_one_full(c, p, q + (size_t)qi * c->dim, false, out + (size_t)qi * out_stride, out_stride, &counts[qi], qi))) return rc;
    return 0;
}

extern "C" int spl_gpu_merge_topk_device(spl_gpu_ctx *c, const spl_gpu_result *d_lists, const uint32_t *d_counts,
                                         uint32_t nlists, uint32_t list_stride, uint32_t order, uint32_t k,
                                         spl_gpu_result *d_out, uint32_t *d_out_count, void *stream) {
    if (!c || !d_lists || !d_counts || !d_out || !d_out_count) return fail(-2, EINVAL, "null argument");
    if (k == 0 || k > SPL_GPU_KMAX_FUSED) return fail(-2, EINVAL, "merge needs k in 1..%u", SPL_GPU_KMAX_FUSED);
    if (order > SPL_ORDER_L2) return fail(-2, EINVAL, "unknown order");
    cudaSetDevice(c->device);
    { std::lock_guard<std::mutex> lk(c->mu); c->last_was_tma = false; }
    CUDA_TRY(launch_merge_results(d_lists, d_counts, nlists, 1, list_stride, order, k, d_out, k, d_out_count,
                                  stream ? (cudaStream_t)stream : c->stream));
    c->launches++;
    return 0;
}

extern "C" int spl_gpu_merge_topk_batch_device(spl_gpu_ctx *c, const spl_gpu_result *d_lists, const uint32_t *d_counts,
                                               uint32_t nlists, uint32_t nq, uint32_t list_stride, uint32_t order, uint32_t k,
                                               spl_gpu_result *d_out, uint32_t out_stride, uint32_t *d_out_counts, void *stream) {
    if (!c || !d_lists || !d_counts || !d_out || !d_out_counts) return fail(-2, EINVAL, "null argument");
    if (k == 0 || k > SPL_GPU_KMAX_FUSED) return fail(-2, EINVAL, "merge needs k in 1..%u", SPL_GPU_KMAX_FUSED);
    if (order > SPL_ORDER_L2) return fail(-2, EINVAL, "unknown order");
    if (nq == 0 || out_stride < k) return fail(-2, EINVAL, "merge needs nq >= 1 and out_stride >= k");
    cudaSetDevice(c->device);
    { std::lock_guard<std::mutex> lk(c->mu); c->last_was_tma = false; }
    CUDA_TRY(launch_merge_results(d_lists, d_counts, nlists, nq, list_stride, order, k, d_out, out_stride, d_out_counts,
                                  stream ? (cudaStream_t)stream : c->stream));
    c->launches++;
    return 0;
}

extern "C" int spl_gpu_merge_topk_packed_device(spl_gpu_ctx *c, const void *d_packed, uint32_t nlists, uint32_t nq,
                                                uint32_t list_stride, uint32_t order, uint32_t k, spl_gpu_result *d_out,
                                                uint32_t out_stride, uint32_t *d_out_counts, void *stream) {
    if (!c || !d_packed || !d_out || !d_out_counts) return fail(-2, EINVAL, "null argument");
    if (k == 0 || k > SPL_GPU_KMAX_FUSED) return fail(-2, EINVAL, "merge needs k in 1..%u", SPL_GPU_KMAX_FUSED);
    if (order > SPL_ORDER_L2) return fail(-2, EINVAL, "unknown order");
    if (nq == 0 || out_stride < k || list_stride == 0) return fail(-2, EINVAL, "merge needs nq >= 1, list_stride >= 1 and out_stride >= k");
    const size_t block = SPL_GPU_PACKED_BLOCK_BYTES(nq, list_stride);
    const unsigned char *base = static_cast<const unsigned char *>(d_packed);
    cudaSetDevice(c->device);
    { std::lock_guard<std::mutex> lk(c->mu); c->last_was_tma = false; }
    CUDA_TRY(launch_merge_results(reinterpret_cast<const spl_gpu_result *>(base),
                                  reinterpret_cast<const uint32_t *>(base + SPL_GPU_PACKED_COUNTS_OFFSET(nq, list_stride)), nlists, nq,
                                  list_stride, order, k, d_out, out_stride, d_out_counts,
                                  stream ? (cudaStream_t)stream : c->stream, (uint32_t)(block / sizeof(spl_gpu_result)),
                                  (uint32_t)(block / sizeof(uint32_t))));
    c->launches++;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Several GPUs behind one call: one process, one context per GPU over contiguous slot ranges.
// Every consumer of splinter.h is a C host (splinter.h:21-23); this is what such a host calls instead of
// orchestrating ranks itself.  Per batch: the per-GPU scans are enqueued on every device's own stream and run side
// by side; each leaves ONE packed block ([nq][k] records + [nq] counts); the blocks are pulled onto the first device
// with one peer copy each (3.2 KB per GPU at k = 100 -- inside one process no collective library is needed), merged
// there with merge_results_kernel, and the merged records come back in one copy.
// ------------------------------------------------------------------------------------------------
struct spl_gpu_group {
    std::vector<spl_gpu_ctx *> ctx;
    std::vector<cudaEvent_t> done;           // per context: its shard's block is complete (recorded on its stream)
    std::vector<unsigned char *> d_local;    // per context, on its device: the packed block of the current batch
    unsigned char *d_blocks = nullptr;       // on device of ctx[0]: all blocks back to back
    spl_gpu_result *d_fin = nullptr, *h_fin = nullptr;
    uint32_t *d_fcnt = nullptr, *h_fcnt = nullptr;
    size_t block_cap = 0;                    // bytes per block the buffers above were sized for
    uint32_t fin_cap = 0, nq_cap = 0;        // records / queries
    std::mutex mu;
    std::atomic<uint64_t> reissued{0};
};

static void group_free_buffers(spl_gpu_group *g) {
    for (size_t i = 0; i < g->ctx.size(); i++) {
        if (i < g->d_local.size() && g->d_local[i]) { cudaSetDevice(g->ctx[i]->device); cudaFree(g->d_local[i]); g->d_local[i] = nullptr; }
    }
    if (!g->ctx.empty()) cudaSetDevice(g->ctx[0]->device);
    cudaFree(g->d_blocks); cudaFree(g->d_fin); cudaFree(g->d_fcnt); cudaFreeHost(g->h_fin); cudaFreeHost(g->h_fcnt);
    g->d_blocks = nullptr; g->d_fin = nullptr; g->d_fcnt = nullptr; g->h_fin = nullptr; g->h_fcnt = nullptr;
    g->block_cap = 0; g->fin_cap = 0; g->nq_cap = 0;
}

static int group_ensure_buffers(spl_gpu_group *g, uint32_t nq_need, uint32_t k) {
    const size_t bb_need = SPL_GPU_PACKED_BLOCK_BYTES(nq_need, k);
    const uint32_t fin_need = nq_need * k;
    if (bb_need <= g->block_cap && fin_need <= g->fin_cap && nq_need <= g->nq_cap) return 0;
    // grow every capacity to the largest seen so far: calls alternate between shapes (a batch, then single re-issues)
    const size_t bb = bb_need > g->block_cap ? bb_need : g->block_cap;
    const uint32_t fin = fin_need > g->fin_cap ? fin_need : g->fin_cap, nq = nq_need > g->nq_cap ? nq_need : g->nq_cap;
    for (spl_gpu_ctx *c : g->ctx) { cudaSetDevice(c->device); cudaStreamSynchronize(c->stream); }
    group_free_buffers(g);
    g->d_local.assign(g->ctx.size(), nullptr);
    for (size_t i = 0; i < g->ctx.size(); i++) {
        cudaSetDevice(g->ctx[i]->device);
        if (cudaMalloc((void **)&g->d_local[i], bb) != cudaSuccess) return fail(-1, ENOMEM, "group: cudaMalloc(%zu) failed", bb);
    }
    cudaSetDevice(g->ctx[0]->device);
    if (cudaMalloc((void **)&g->d_blocks, bb * g->ctx.size()) != cudaSuccess || cudaMalloc((void **)&g->d_fin, (size_t)fin * sizeof(spl_gpu_result)) != cudaSuccess ||
        cudaMalloc((void **)&g->d_fcnt, (size_t)nq * 4) != cudaSuccess)
        return fail(-1, ENOMEM, "group: cudaMalloc of the merge buffers failed");
    CUDA_TRY(cudaMallocHost((void **)&g->h_fin, (size_t)fin * sizeof(spl_gpu_result)));
    CUDA_TRY(cudaMallocHost((void **)&g->h_fcnt, (size_t)nq * 4));
    g->block_cap = bb; g->fin_cap = fin; g->nq_cap = nq;
    return 0;
}

static spl_gpu_group *group_build(uint32_t n_devices, const int32_t *devices, const spl_gpu_opts *opts, uint64_t total,
                                  const char *name, uint32_t dense_dim, uint64_t dense_row0) {
    if (n_devices == 0 || n_devices > 64) { fail(-2, EINVAL, "group needs 1..64 devices"); return nullptr; }
    spl_gpu_group *g = new spl_gpu_group();
    const uint64_t per = (total + n_devices - 1) / n_devices;   // contiguous, ceil-divided
    for (uint32_t i = 0; i < n_devices; i++) {
        const uint64_t lo = (uint64_t)i * per < total ? (uint64_t)i * per : total;
        const uint64_t cnt = lo + per <= total ? per : total - lo;
        if (cnt == 0) break;                                     // more devices than slots: the rest stay idle
        spl_gpu_opts o;
        memset(&o, 0, sizeof o);
        if (opts) o = *opts;
        o.struct_size = sizeof o;
        o.device = devices ? devices[i] : (int32_t)i;
        spl_gpu_ctx *c;
        if (name) {
            o.slot_begin = (uint32_t)lo; o.slot_count = (uint32_t)cnt;
            c = spl_gpu_attach(name, &o);
        } else {
            c = spl_gpu_open_dense(cnt, dense_dim, dense_row0 + lo, &o);
        }
        if (!c) { for (spl_gpu_ctx *x : g->ctx) destroy(x); delete g; return nullptr; }
        g->ctx.push_back(c);
    }
    g->done.resize(g->ctx.size());
    for (size_t i = 0; i < g->ctx.size(); i++) {
        cudaSetDevice(g->ctx[i]->device);
        if (cudaEventCreateWithFlags(&g->done[i], cudaEventDisableTiming) != cudaSuccess) {
            fail(-1, EIO, "group: cudaEventCreate failed");
            for (spl_gpu_ctx *x : g->ctx) destroy(x);
            delete g;
            return nullptr;
        }
    }
    return g;
}

extern "C" spl_gpu_group *spl_gpu_group_attach(const char *name_or_path, uint32_t n_devices, const int32_t *devices,
                                               const spl_gpu_opts *opts) {
    if (!name_or_path) { fail(-2, EINVAL, "null store name"); return nullptr; }
    // the store's slot count decides the shard ranges: read it from the header the way spl_gpu_attach will
    const bool persistent = opts && (opts->flags & SPL_OPT_PERSISTENT);
    int fd = persistent ? open(name_or_path, O_RDONLY) : shm_open(name_or_path, O_RDONLY, 0);
    if (fd < 0) { fail(-1, errno, "cannot open store '%s': %s", name_or_path, strerror(errno)); return nullptr; }
    uint8_t hdr[16];
    const ssize_t got = pread(fd, hdr, sizeof hdr, 0);
    close(fd);
    uint32_t magic = 0, version = 0, slots = 0;
    if (got == (ssize_t)sizeof hdr) { memcpy(&magic, hdr, 4); memcpy(&version, hdr + 4, 4); memcpy(&slots, hdr + 8, 4); }
    if (magic != layout::kMagic || version != layout::kVersion || slots == 0) {
        fail(-1, EINVAL, "'%s' is not a splinter v%u store", name_or_path, layout::kVersion);
        return nullptr;
    }
    return group_build(n_devices, devices, opts, slots, name_or_path, 0, 0);
}

extern "C" spl_gpu_group *spl_gpu_group_open_dense(uint64_t n_rows, uint32_t dim, uint32_t n_devices, const int32_t *devices,
                                                   const spl_gpu_opts *opts) {
    if (n_rows == 0) { fail(-2, EINVAL, "dense group needs rows"); return nullptr; }
    return group_build(n_devices, devices, opts, n_rows, nullptr, dim, 0);
}

extern "C" void spl_gpu_group_detach(spl_gpu_group *g) {
    if (!g) return;
    for (spl_gpu_ctx *c : g->ctx) { cudaSetDevice(c->device); cudaStreamSynchronize(c->stream); }
    group_free_buffers(g);
    for (size_t i = 0; i < g->done.size(); i++) { cudaSetDevice(g->ctx[i]->device); cudaEventDestroy(g->done[i]); }
    for (spl_gpu_ctx *c : g->ctx) destroy(c);
    delete g;
}

extern "C" uint32_t spl_gpu_group_size(const spl_gpu_group *g) { return g ? (uint32_t)g->ctx.size() : 0; }
extern "C" spl_gpu_ctx *spl_gpu_group_member(spl_gpu_group *g, uint32_t i) { return g && i < g->ctx.size() ? g->ctx[i] : nullptr; }
extern "C" uint64_t spl_gpu_group_reissued(const spl_gpu_group *g) { return g ? g->reissued.load() : 0; }

// every shard sweeps its own slot range, all at once (each context has its own sweep threads and staging)
extern "C" int spl_gpu_group_sync(spl_gpu_group *g, spl_gpu_sync_stats *total) {
    if (!g) return fail(-2, EINVAL, "null group");
    std::vector<spl_gpu_sync_stats> st(g->ctx.size());
    std::vector<int> rc(g->ctx.size(), 0);
    std::vector<std::string> err(g->ctx.size());
    std::vector<std::thread> th;
    for (size_t i = 0; i < g->ctx.size(); i++)
        th.emplace_back([&, i] { rc[i] = spl_gpu_sync(g->ctx[i], &st[i]); if (rc[i]) err[i] = g_err; });
    for (auto &t : th) t.join();
    for (size_t i = 0; i < g->ctx.size(); i++)
        if (rc[i]) return fail(rc[i], EIO, "group sync, shard %zu: %s", i, err[i].c_str());
    if (total) {
        memset(total, 0, sizeof *total);
        for (const spl_gpu_sync_stats &x : st) {
            total->slots_swept += x.slots_swept; total->rows_uploaded += x.rows_uploaded; total->meta_updated += x.meta_updated;
            total->skipped_busy += x.skipped_busy; total->aba_rechecked += x.aba_rechecked; total->bytes_h2d += x.bytes_h2d;
            total->rows_revalidated += x.rows_revalidated; total->stale_rows_found += x.stale_rows_found;
            if (x.sweep_ms > total->sweep_ms) total->sweep_ms = x.sweep_ms;
            if (x.upload_ms > total->upload_ms) total->upload_ms = x.upload_ms;
            if (x.generation > total->generation) total->generation = x.generation;
            total->header_epoch = x.header_epoch;
        }
    }
    return 0;
}

extern "C" int spl_gpu_group_fill_synthetic(spl_gpu_group *g, uint64_t seed_corpus, uint64_t seed_bloom) {
    if (!g) return fail(-2, EINVAL, "null group");
    for (spl_gpu_ctx *c : g->ctx) { int rc = spl_gpu_dense_fill_synthetic(c, seed_corpus, seed_bloom); if (rc) return rc; }
    return 0;
}

// one pass over all shards for nq <= q_cap queries already in host memory; merged records land in g->h_fin / h_fcnt
static int group_pass(spl_gpu_group *g, const spl_gpu_search_params *p, const float *q, uint32_t nq) {
    const size_t bb = SPL_GPU_PACKED_BLOCK_BYTES(nq, p->k);
    const size_t coff = SPL_GPU_PACKED_COUNTS_OFFSET(nq, p->k);
    spl_gpu_search_params pp = *p;
    pp.nq = nq;
    int rc;
    for (size_t i = 0; i < g->ctx.size(); i++) {
        spl_gpu_ctx *c = g->ctx[i];
        cudaSetDevice(c->device);
        std::lock_guard<std::mutex> lk(c->mu);
        const size_t qbytes = (size_t)nq * c->dim * sizeof(float);
        memcpy(c->h_q, q, qbytes);
        CUDA_TRY(cudaMemcpyAsync(c->d_q, c->h_q, qbytes, cudaMemcpyHostToDevice, c->stream));
        if ((rc = order_before_search(c, c->stream))) return rc;
        unsigned char *blk = i == 0 ? g->d_blocks : g->d_local[i];   // the first shard writes straight into the merge area
        rc = search_device_locked(c, &pp, c->d_q, reinterpret_cast<spl_gpu_result *>(blk), p->k,
                                  reinterpret_cast<uint32_t *>(blk + coff), c->stream);
        if (rc) { c->counters_dirty = true; order_after_search(c, c->stream); return rc; }
        if ((rc = order_after_search(c, c->stream))) return rc;
        CUDA_TRY(cudaEventRecord(g->done[i], c->stream));
    }
    spl_gpu_ctx *c0 = g->ctx[0];
    cudaSetDevice(c0->device);
    std::lock_guard<std::mutex> lk(c0->mu);
    for (size_t i = 1; i < g->ctx.size(); i++) {
        CUDA_TRY(cudaStreamWaitEvent(c0->stream, g->done[i], 0));
        CUDA_TRY(cudaMemcpyPeerAsync(g->d_blocks + i * bb, c0->device, g->d_local[i], g->ctx[i]->device, bb, c0->stream));
    }
    c0->last_was_tma = false;
    CUDA_TRY(launch_merge_results(reinterpret_cast<const spl_gpu_result *>(g->d_blocks),
                                  reinterpret_cast<const uint32_t *>(g->d_blocks + coff), (uint32_t)g->ctx.size(), nq, p->k, p->order,
                                  p->k, g->d_fin, p->k, g->d_fcnt, c0->stream, (uint32_t)(bb / sizeof(spl_gpu_result)),
                                  (uint32_t)(bb / sizeof(uint32_t))));
    c0->launches++;
    CUDA_TRY(cudaMemcpyAsync(g->h_fin, g->d_fin, (size_t)nq * p->k * sizeof(spl_gpu_result), cudaMemcpyDeviceToHost, c0->stream));
    CUDA_TRY(cudaMemcpyAsync(g->h_fcnt, g->d_fcnt, (size_t)nq * 4, cudaMemcpyDeviceToHost, c0->stream));
    CUDA_TRY(cudaStreamSynchronize(c0->stream));
    c0->search_recorded = false;   // (everything this pass enqueued on c0's stream has completed)
    return 0;
}

extern "C" int spl_gpu_group_search(spl_gpu_group *g, const spl_gpu_search_params *p, const float *q, spl_gpu_result *out,
                                    uint32_t out_stride, uint32_t *counts) {
    if (!g || g->ctx.empty()) return fail(-2, EINVAL, "null group");
    int rc = check_params(g->ctx[0], p);
    if (rc) return rc;
    if (!q || !out || !counts) return fail(-2, EINVAL, "null buffer");
    if (p->k == 0 || p->k > SPL_GPU_KMAX_FUSED) return fail(-2, EINVAL, "group search needs k in 1..%u", SPL_GPU_KMAX_FUSED);
    if (out_stride < p->k) return fail(-2, EINVAL, "out_stride < k");
    std::lock_guard<std::mutex> lk(g->mu);
    const uint32_t chunk = g->ctx[0]->q_cap;
    const uint32_t dim = g->ctx[0]->dim;
    if ((rc = group_ensure_buffers(g, p->nq < chunk ? p->nq : chunk, p->k))) return rc;
    for (uint32_t q0 = 0; q0 < p->nq; q0 += chunk) {
        const uint32_t nq = p->nq - q0 < chunk ? p->nq - q0 : chunk;
        if ((rc = group_pass(g, p, q + (size_t)q0 * dim, nq))) return rc;
        for (uint32_t i = 0; i < nq; i++) {
            counts[q0 + i] = g->h_fcnt[i];
            if (g->h_fcnt[i] != SPL_GPU_COUNT_UNANSWERED)
                memcpy(out + (size_t)(q0 + i) * out_stride, g->h_fin + (size_t)i * p->k, (size_t)g->h_fcnt[i] * sizeof(spl_gpu_result));
        }
        // what a shard could not answer in the requested mode goes to the fp32 scan, on every shard
        for (uint32_t i = 0; i < nq; i++) {
            if (counts[q0 + i] != SPL_GPU_COUNT_UNANSWERED) continue;
            spl_gpu_search_params one = *p;
            one.mode = SPL_MODE_FAST;
            if ((rc = group_pass(g, &one, q + (size_t)(q0 + i) * dim, 1))) return rc;
            counts[q0 + i] = g->h_fcnt[0];
            memcpy(out + (size_t)(q0 + i) * out_stride, g->h_fin, (size_t)g->h_fcnt[0] * sizeof(spl_gpu_result));
            g->reissued++;
        }
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------
// bulk embedding of ingested text; contract in include/splinter_gpu.h
// ------------------------------------------------------------------------------------------------
extern "C" int spl_gpu_ingest_embeddings(spl_gpu_ctx *c, uint64_t label_mask, uint32_t batch, spl_gpu_embed_fn embed, void *user,
                                         spl_gpu_set_embedding_fn set_embedding, spl_gpu_ingest_stats *stats) {
    using namespace layout;
    if (!c || !embed || !set_embedding) return fail(-2, EINVAL, "null argument");
    if (c->dense || !c->base) return fail(-2, EINVAL, "dense contexts have no store to ingest from");
    if (batch == 0) batch = 256;
    constexpr size_t S_VAL_OFF = 16, S_TYPE = 24;
    constexpr uint8_t kVartext = 1u << 7;                       // SPL_SLOT_TYPE_VARTEXT, splinter.h:89
    uint32_t maxval;
    memcpy(&maxval, c->base + H_MAXVAL, 4);
    const uint8_t *values = c->base + kHeaderSize + (size_t)c->total_slots * c->slot_size;   // splinter.c:107-109
    spl_gpu_ingest_stats st;
    memset(&st, 0, sizeof st);
    int rc = spl_gpu_sync(c, nullptr);
    if (rc) return rc;
    // candidates, in slot order (the order splinter_list hands keys to the daemon's backfill, splinter.c:466-477)
    std::vector<uint32_t> cand;
    std::unique_lock<std::mutex> caches(c->sweep_mu);           // the host caches belong to the sweep
    for (uint64_t r = 0; r < c->n; r++) {
        if (!(c->c_flags[r] & SPL_ROW_ENUM) || (c->c_bloom[r] & label_mask) != label_mask) continue;
        if (!(c->c_flags[r] & SPL_ROW_LIST)) continue;          // no value, nothing to embed
        if (c->c_state[r] & 2) continue;                        // the mirrored vector is not zero: embedded already
        const uint8_t *sl = c->base + kHeaderSize + (c->slot_begin + r) * c->slot_size;
        if (!(sl[S_TYPE] & kVartext)) continue;
        cand.push_back((uint32_t)r);
    }
    caches.unlock();
    st.candidates = cand.size();
    std::vector<char> text((size_t)batch * ((size_t)maxval + 1));
    std::vector<const char *> ptrs(batch);
    std::vector<uint32_t> lens(batch), rows(batch);
    std::vector<uint64_t> epochs(batch);
    std::vector<std::array<char, 64>> keys(batch);
    std::vector<float> vecs((size_t)batch * c->dim);
    for (size_t i0 = 0; i0 < cand.size();) {
        auto t0 = std::chrono::steady_clock::now();
        uint32_t nb = 0;
        for (; i0 < cand.size() && nb < batch; i0++) {
            const uint32_t r = cand[i0];
            const uint8_t *sl = c->base + kHeaderSize + (c->slot_begin + r) * c->slot_size;
            const uint64_t e1 = ld_acq64(sl + S_EPOCH);
            if (e1 & 1) { st.skipped_busy++; continue; }        // writer active (splinference.cpp:201)
            const uint32_t off = ld_acq32(sl + S_VAL_OFF), len = ld_acq32(sl + S_VAL_LEN);
            if (len == 0 || len > maxval || (uint64_t)off + len > (uint64_t)c->total_slots * maxval) { st.skipped_busy++; continue; }
            char *dst = text.data() + (size_t)nb * ((size_t)maxval + 1);
            memcpy(dst, values + off, len);
            dst[len] = '\0';
            memcpy(keys[nb].data(), sl + S_KEY, 64);
            keys[nb][63] = '\0';
            std::atomic_thread_fence(std::memory_order_acquire);
            if (ld_acq64(sl + S_EPOCH) != e1 || ld_acq64(sl + S_HASH) == 0) { st.skipped_busy++; continue; }
            ptrs[nb] = dst; lens[nb] = len; rows[nb] = r; epochs[nb] = e1;
            nb++;
        }
        auto t1 = std::chrono::steady_clock::now();
        st.gather_ms += std::chrono::duration<double, std::milli>(t1 - t0).count();
        if (nb == 0) continue;
        if (embed(user, nb, ptrs.data(), lens.data(), c->dim, vecs.data()) != 0)
            return fail(-1, EIO, "the embed callback failed on a batch of %u texts", nb);
        auto t2 = std::chrono::steady_clock::now();
        st.embed_ms += std::chrono::duration<double, std::milli>(t2 - t1).count();
        for (uint32_t i = 0; i < nb; i++) {
            if (set_embedding(keys[i].data(), vecs.data() + (size_t)i * c->dim) != 0) { st.write_failed++; continue; }
            const uint8_t *sl = c->base + kHeaderSize + (c->slot_begin + rows[i]) * c->slot_size;
            if (ld_acq64(sl + S_EPOCH) == epochs[i] + 2) st.embedded++; else st.raced++;   // splinference.cpp:282-286
        }
        st.write_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t2).count();
        st.batches++;
    }
    auto t3 = std::chrono::steady_clock::now();
    rc = spl_gpu_sync(c, nullptr);
    st.sync_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t3).count();
    if (stats) *stats = st;
    return rc;
}
