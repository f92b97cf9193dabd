// spl_noalloc.cu -- spl_gpu_scan_noalloc: the fused single-query scan for callers that must not
// allocate device memory.
//
// Who needs this: a host program that has handed ALL of its device memory to its own arena or
// pool (a database engine, an inference server with a caching allocator) and treats any further
// cudaMalloc as a bug; a latency-critical loop that cannot afford the device-wide stall cudaMalloc
// and cudaFree imply; code that runs where allocation is simply not possible any more (memory is
// exhausted by design).  Such a caller already owns the mirror columns, the query and the result
// buffer as device pointers; the only thing the normal path would still allocate is the kernel's
// scratch (one candidate list per CTA plus two counters).  Here that scratch is a fixed array
// baked into the library image instead.
//
// Cost: every GPU that loads libsplinter_gpu.so carries kNoAllocScratchBytes (257 KB) of device
// memory for this array from module load until the process exits, whether or not this entry point
// is ever called.
#include <errno.h>
#include <stdarg.h>
#include <stdio.h>

#include <mutex>

#include "spl_scan.cuh"

namespace spl { void set_last_error(const char *msg); }  // spl_host.cu: text behind spl_gpu_last_error()

using namespace spl;

namespace {

constexpr uint32_t kNoAllocMaxK = 64;      // largest k this entry point accepts
constexpr uint32_t kNoAllocMaxCtas = 256;  // one CTA per SM; no current part has more SMs than this
constexpr size_t kCountersBytes = 256;     // {work counter, done counter}, padded
constexpr size_t kNoAllocScratchBytes = kCountersBytes + (size_t)kNoAllocMaxCtas * kNoAllocMaxK * sizeof(Cand);

// The scratch itself.  In HOST code this name is only a handle the runtime knows the symbol by --
// it is not a device address and must never be handed to a kernel; set-up below asks the runtime
// for the real address once per device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
struct DeviceSetup {
    bool ready = false;
    unsigned char *scratch = nullptr;  // device address of g_noalloc_scratch ON THIS DEVICE
    int sm_count = 0;
};
DeviceSetup g_setup[kMaxDevices];
std::mutex g_setup_mu;

// failures are reported through the library's usual channel: errno + spl_gpu_last_error()
int fail(int rc, int err, const char *fmt, ...) {
    char msg[384];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(msg, sizeof msg, fmt, ap);
    va_end(ap);
    spl::set_last_error(msg);
    errno = err;
    return rc;
}

// Once per device, with that device current: resolve the array's address on this device and opt the
// kernel in to its shared-memory size.  Neither call allocates.
int setup_current_device(DeviceSetup **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(-1, EIO, "cudaGetDevice failed: %s (no CPU fallback exists)", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return fail(-2, EINVAL, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_setup_mu);
    DeviceSetup &s = g_setup[dev];
    if (!s.ready) {
        void *ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return fail(-1, EIO, "cudaGetSymbolAddress(scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return fail(-1, EIO, "cudaDeviceGetAttribute failed: %s", cudaGetErrorString(e));
        int major = 0;
        cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        if (major < 10) return fail(-1, EIO, "device %d is not sm_100: this library carries sm_100a code only", dev);
        e = scan_tma_prepare();
        if (e != cudaSuccess) return fail(-1, EIO, "kernel set-up failed on device %d: %s", dev, cudaGetErrorString(e));
        s.scratch = static_cast<unsigned char *>(ptr);
        s.sm_count = sms;
        s.ready = true;
    }
    *out = &s;
    return 0;
}

}  // namespace

extern "C" int spl_gpu_scan_noalloc(const float *d_E, const uint64_t *d_bloom, const uint8_t *d_flags,
                                    const uint64_t *d_epoch, uint64_t n_rows, uint64_t global_row0,
                                    const spl_gpu_search_params *p, const float *d_q, spl_gpu_result *d_out,
                                    uint32_t out_stride, uint32_t *d_counts, void *stream) {
    if (!d_E || !d_bloom || !d_flags || !d_epoch || !p || !d_q || !d_out || !d_counts)
        return fail(-2, EINVAL, "null pointer argument");
    if (p->struct_size != sizeof(spl_gpu_search_params)) return fail(-2, EINVAL, "spl_gpu_search_params size mismatch");
    if (p->k == 0 || p->k > kNoAllocMaxK)
        return fail(-2, EINVAL, "k = %u is outside 1..%u: the built-in scratch holds %u candidates per CTA at most "
                    "(use spl_gpu_search for larger k)", p->k, kNoAllocMaxK, kNoAllocMaxK);
    if (out_stride < p->k) return fail(-2, EINVAL, "out_stride < k");
    if (p->nq == 0 || p->nq > 4096) return fail(-2, EINVAL, "nq must be in 1..4096");
    // the kernel counts 32-row groups in 32 bits: (n + 31) must not wrap
    if (n_rows == 0 || n_rows > 0xFFFFFFE0ull) return fail(-2, EINVAL, "n_rows must be in 1..2^32-32");
    // a misaligned arena sub-allocation would otherwise surface as a sticky device fault: rows travel as 16-byte
    // bulk copies, the query is read as float4, records are written as 32-byte structs
    if (((uintptr_t)d_E | (uintptr_t)d_q) % 16 || ((uintptr_t)d_out | (uintptr_t)d_bloom | (uintptr_t)d_epoch) % 8 ||
        (uintptr_t)d_counts % 4)
        return fail(-2, EINVAL, "misaligned device pointer: E and q need 16-byte alignment, out / bloom / epoch 8, counts 4");
    if (p->order > SPL_ORDER_L2) return fail(-2, EINVAL, "unknown order %u", p->order);
    if (p->mode != SPL_MODE_AUTO && p->mode != SPL_MODE_FAST) return fail(-1, ENOSYS, "only SPL_MODE_FAST runs without allocation");
    if (!scan_tma_supported(p->dim, p->k))
        return fail(-2, EINVAL, "dim = %u is not served by the allocation-free kernel (128, 256, 512, 768, 1024 are)", p->dim);

    DeviceSetup *dev = nullptr;
    int rc = setup_current_device(&dev);
    if (rc) return rc;
    const int grid = dev->sm_count;
    if ((uint32_t)grid > kNoAllocMaxCtas)
        return fail(-2, EINVAL, "%d SMs exceed the %u CTAs the built-in scratch was sized for", grid, kNoAllocMaxCtas);

    cudaStream_t s = static_cast<cudaStream_t>(stream);
    ScanArgs a = {};
    a.E = d_E; a.bloom = d_bloom; a.flags = d_flags; a.epoch = d_epoch;
    a.n = (uint32_t)n_rows; a.dim = p->dim; a.k = p->k; a.order = p->order;
    a.mask = p->bloom_mask; a.min_sim = p->min_similarity; a.max_dist = p->max_distance;
    a.exclude = kIdxSentinel;
    if (p->exclude_slot != SPL_NO_SLOT && p->exclude_slot >= global_row0 && p->exclude_slot < global_row0 + n_rows)
        a.exclude = (uint32_t)(p->exclude_slot - global_row0);
    a.row0 = global_row0;
    a.counters = reinterpret_cast<uint32_t *>(dev->scratch);
    a.cta_cands = reinterpret_cast<Cand *>(dev->scratch + kCountersBytes);
    // The array keeps whatever the previous call (possibly an aborted one) left behind.  Only the two
    // counters have to start at zero; the candidate lists are fully rewritten before they are read.
    cudaError_t e = cudaMemsetAsync(dev->scratch, 0, kCountersBytes, s);
    if (e != cudaSuccess) return fail(-1, EIO, "clearing the scratch counters failed: %s", cudaGetErrorString(e));
    for (uint32_t qi = 0; qi < p->nq; qi++) {
        a.q = d_q + (size_t)qi * p->dim;
        a.out = d_out + (size_t)qi * out_stride;
        a.out_count = d_counts + qi;
        e = launch_scan_tma(a, grid, s);
        if (e != cudaSuccess) return fail(-1, EIO, "scan launch failed: %s", cudaGetErrorString(e));
    }
    return 0;
}
