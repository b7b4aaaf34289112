// common.cuh -- shared host/device helpers for libb200knn (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>
#include <atomic>
#include <string>

#include "../../include/b200knn.h"

namespace b2k {

typedef unsigned long long u64;

// ---- error plumbing (host) ---------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define B2K_CUDA(expr)                                                                      \
    do {                                                                                    \
        cudaError_t _e = (expr);                                                            \
        if (_e != cudaSuccess) {                                                            \
            b2k::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                           __LINE__);                                                       \
            return B2K_ECUDA;                                                               \
        }                                                                                   \
    } while (0)

// ---- selection keys -----------------------------------------------------------------
// A candidate is one 64-bit key: (float bits of the squared distance) << 32 | local row.
// Distances are >= 0, so unsigned order of the bits is numeric order; the low word makes
// the order lexicographic on (distance, id), the tie rule of the oracle.  ~0 is "empty".
static constexpr u64 KEY_EMPTY = ~0ull;
static constexpr uint32_t FLT_MAX_BITS = 0x7f7fffffu;

__device__ __forceinline__ u64 make_key(float dist, uint32_t row) {
    uint32_t b = __float_as_uint(dist);
    if (b == 0x80000000u) b = 0u;  // -0.0f
    return ((u64)b << 32) | row;
}
__device__ __forceinline__ float key_dist(u64 k) { return __uint_as_float((uint32_t)(k >> 32)); }
__device__ __forceinline__ uint32_t key_row(u64 k) { return (uint32_t)k; }

// 128-bit streaming load: read-only path, do not allocate in L1 (rows are read once).
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// Block-wide bitonic sort (ascending) of P = 2^m keys in shared memory.
// All threads of the block must call; ends with the data visible to all threads.
__device__ __forceinline__ void block_bitonic_sort(u64* s, int P) {
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = tid; i < (P >> 1); i += nt) {
                int lo = 2 * i - (i & (stride - 1));  // index with bit `stride` cleared
                int hi = lo + stride;
                bool up = ((lo & size) == 0);
                u64 a = s[lo], b = s[hi];
                if ((a > b) == up) { s[lo] = b; s[hi] = a; }
            }
            __syncthreads();
        }
    }
}

// Warp-wide bitonic sort of P = 2^m keys in shared memory private to the warp.
__device__ __forceinline__ void warp_bitonic_sort(u64* s, int P, int lane) {
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = lane; i < (P >> 1); i += 32) {
                int lo = 2 * i - (i & (stride - 1));
                int hi = lo + stride;
                bool up = ((lo & size) == 0);
                u64 a = s[lo], b = s[hi];
                if ((a > b) == up) { s[lo] = b; s[hi] = a; }
            }
            __syncwarp();
        }
    }
}

inline int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

}  // namespace b2k
