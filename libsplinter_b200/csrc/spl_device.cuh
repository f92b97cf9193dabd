// spl_device.cuh -- device-side building blocks shared by the scan / merge kernels (sm_100a).
//
// Ordering keys.  The reference sorts hits with compare_results
// (/root/reference/splinter_cli_cmd_search.c:64-72): similarity descending, then distance
// ascending.  We fold (similarity, distance) into one 64-bit integer whose ascending order is
// exactly that comparator, and break the remaining ties (left unspecified by the reference's
// qsort) by ascending slot index.  The mapping is a bijection on non-NaN floats, so the two
// scores are recovered from the key bit-for-bit when results are written.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/splinter_gpu.h"

namespace spl {

constexpr uint32_t kWarp = 32;
constexpr uint64_t kKeySentinel = 0xFFFFFFFFFFFFFFFFull;
constexpr uint32_t kIdxSentinel = 0xFFFFFFFFu;

// monotone float -> uint32 (ascending); -0.0 is folded onto +0.0 first because the reference's
// float compares treat them as equal.
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t u = __float_as_uint(f + 0.0f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) {
    uint32_t u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
    return __uint_as_float(u);
}
__device__ __forceinline__ uint64_t make_key(uint32_t order, float sim, float dist) {
    uint32_t s = ~f2ord(sim), d = f2ord(dist);
    return order == SPL_ORDER_COSINE ? ((uint64_t)s << 32) | d : ((uint64_t)d << 32) | s;
}
__device__ __forceinline__ void split_key(uint32_t order, uint64_t key, float &sim, float &dist) {
    uint32_t hi = (uint32_t)(key >> 32), lo = (uint32_t)key;
    uint32_t s = order == SPL_ORDER_COSINE ? hi : lo, d = order == SPL_ORDER_COSINE ? lo : hi;
    sim = ord2f(~s);
    dist = ord2f(d);
}
__device__ __forceinline__ bool better(uint64_t k1, uint32_t i1, uint64_t k2, uint32_t i2) {
    return k1 < k2 || (k1 == k2 && i1 < i2);
}

// The reference's scalar tail of cosine_similarity (_search.c:50-52) with IEEE-rounded ops.
__device__ __forceinline__ float cosine_from_sums(float dot, float ma, float mb) {
    float den = __fmul_rn(__fsqrt_rn(ma), __fsqrt_rn(mb));
    return den < 1e-9f ? 0.0f : __fdiv_rn(dot, den);
}

struct alignas(16) Cand {
    uint64_t key;
    uint32_t idx;
    uint32_t pad;
};

// ---- per-warp bounded candidate list ("replace the worst") ---------------------------------
// k entries in shared memory, unsorted.  Until the list is full every candidate is appended;
// afterwards a candidate enters only if it beats the current worst entry, which it replaces.
// All arguments are warp-uniform; every lane of the warp must call these together.
struct WarpTopK {
    uint64_t *keys;
    uint32_t *idxs;
    uint32_t k, cnt;
    uint64_t thr_key;   // worst entry when cnt == k
    uint32_t thr_idx, thr_pos;

    __device__ __forceinline__ void init(uint64_t *kk, uint32_t *ii, uint32_t k_) {
        keys = kk; idxs = ii; k = k_; cnt = 0;
        thr_key = kKeySentinel; thr_idx = kIdxSentinel; thr_pos = 0;
    }
    __device__ __forceinline__ bool admits(uint64_t key, uint32_t idx) const {
        return cnt < k || better(key, idx, thr_key, thr_idx);
    }
    // thr := the worst entry (largest key, then largest slot index) and its position.  Each lane finds the worst of
    // its own entries; the warp-wide maximum is then taken word by word with three REDUX instructions (high word,
    // low word among the lanes that tie on it, slot index among those) instead of a five-step shuffle tree over
    // four registers -- this runs after every insertion, and small or selective scans insert a lot.
    __device__ void recompute(uint32_t lane) {
        uint64_t mk = 0; uint32_t mi = 0, mp = 0; bool has = false;
        for (uint32_t p = lane; p < cnt; p += kWarp) {
            uint64_t key = keys[p]; uint32_t id = idxs[p];
            if (!has || better(mk, mi, key, id)) { mk = key; mi = id; mp = p; has = true; }
        }
        const uint32_t hi = has ? (uint32_t)(mk >> 32) : 0u;
        const uint32_t mhi = __reduce_max_sync(0xffffffffu, hi);
        const bool c1 = has && hi == mhi;
        const uint32_t lo = c1 ? (uint32_t)mk : 0u;
        const uint32_t mlo = __reduce_max_sync(0xffffffffu, lo);
        const bool c2 = c1 && lo == mlo;
        const uint32_t mid = __reduce_max_sync(0xffffffffu, c2 ? mi : 0u);
        const unsigned win = __ballot_sync(0xffffffffu, c2 && mi == mid);   // slot indices are unique: one lane
        thr_key = ((uint64_t)mhi << 32) | mlo; thr_idx = mid;
        thr_pos = __shfl_sync(0xffffffffu, mp, win ? __ffs(win) - 1 : 0);
    }
    __device__ void insert(uint64_t key, uint32_t idx, uint32_t lane) {
        if (cnt < k) {
            if (lane == 0) { keys[cnt] = key; idxs[cnt] = idx; }
            cnt++;
            __syncwarp();
            if (cnt == k) recompute(lane);
        } else if (better(key, idx, thr_key, thr_idx)) {
            if (lane == 0) { keys[thr_pos] = key; idxs[thr_pos] = idx; }
            __syncwarp();
            recompute(lane);
        }
    }
    // feed `n` candidates held in (possibly global) memory through the list
    __device__ void absorb(const uint64_t *skeys, const uint32_t *sidx, uint32_t n, uint32_t lane) {
        for (uint32_t base = 0; base < n; base += kWarp) {
            uint32_t p = base + lane;
            uint64_t key = p < n ? skeys[p] : kKeySentinel;
            uint32_t idx = p < n ? sidx[p] : kIdxSentinel;
            unsigned m = __ballot_sync(0xffffffffu, idx != kIdxSentinel && admits(key, idx));
            while (m) {
                int l = __ffs(m) - 1;
                m &= m - 1;
                insert(__shfl_sync(0xffffffffu, key, l), __shfl_sync(0xffffffffu, idx, l), lane);
            }
        }
    }
    __device__ void absorb(const Cand *src, uint32_t n, uint32_t lane) {
        for (uint32_t base = 0; base < n; base += kWarp) {
            uint32_t p = base + lane;
            uint64_t key = kKeySentinel; uint32_t idx = kIdxSentinel;
            if (p < n) {
                int4 raw = __ldcg(reinterpret_cast<const int4 *>(src + p));
                key = ((uint64_t)(uint32_t)raw.y << 32) | (uint32_t)raw.x;
                idx = (uint32_t)raw.z;
            }
            unsigned m = __ballot_sync(0xffffffffu, idx != kIdxSentinel && admits(key, idx));
            while (m) {
                int l = __ffs(m) - 1;
                m &= m - 1;
                insert(__shfl_sync(0xffffffffu, key, l), __shfl_sync(0xffffffffu, idx, l), lane);
            }
        }
    }
};

__device__ __forceinline__ float4 ldg_stream(const float4 *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

}  // namespace spl
