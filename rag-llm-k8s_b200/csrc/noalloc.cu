// noalloc.cu -- b2k_search_rows_noalloc: the exact scan search (kernel K2) for callers that
// must not allocate device memory.
//
// Who needs it.  `llm/rag.py` keeps an 8B-parameter LLM in the same process as the index; in a
// GPU deployment of that pattern the inference engine usually reserves nearly all of HBM up
// front, so a cudaMalloc made later by a search call can simply fail.  Other callers sit on a
// latency-critical thread where cudaMalloc / cudaFree are unwelcome because they synchronise
// the whole device and stall every other stream.  Both kinds of caller already own the rows,
// the queries and the result buffers; the only thing the scan kernel needs beyond those is its
// scratch: the per-CTA partial top-k lists and the "last CTA merges" ticket.  Here that scratch
// is a fixed array that is part of the library image, so the call allocates nothing: no
// cudaMalloc, no cudaFree, no handle, no workspace growth.
//
// Limits and sharing (read before use):
//   * the array is B2K_NOALLOC_SCRATCH_BYTES (4 MiB) on EVERY GPU that loads the library, from
//     load until the process exits;
//   * it holds at most (4 MiB - 256) / 8 = 524,256 keys, and one launch needs
//     grid * query_group * k keys with grid = 3 * SM count (444 on a B200): k up to
//     B2K_MAX_K = 1024 always fits (the query group shrinks to 1); a request that cannot fit
//     is refused with B2K_ELIMIT and a message -- nothing is truncated silently;
//   * all calls on a device share the one array and it keeps the previous call's contents.
//     The ticket word is cleared on the caller's stream before each launch and the lists are
//     fully rewritten by each launch, so calls that are ordered on the GPU (same stream, or
//     streams the caller has ordered) are independent; two calls that OVERLAP on the GPU race
//     on the array and give wrong results;
//   * the first call on each device looks up the array's device address
//     (cudaGetSymbolAddress, with that device current).  Make that first call during start-up,
//     before any stream capture.
#include <mutex>

#include "common.cuh"
#include "kernels.h"

namespace b2k {

// File-scope device array.  In HOST code this name is only a handle for the runtime: its value
// is not a device address and must never be handed to a kernel.  The address is obtained once
// per device below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B2K_NOALLOC_SCRATCH_BYTES];

static constexpr int NOALLOC_MAX_DEVICES = 64;
static constexpr size_t NOALLOC_HEADER = 256;  // ticket word (+ padding so the lists stay 256-aligned)
static std::mutex g_noalloc_mu;
static void* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

// Device address of the scratch array on `device`, which must be the current device.
static int noalloc_scratch_ptr(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) {
            set_error("b2k_search_rows_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                      cudaGetErrorString(e));
            return B2K_ECUDA;
        }
        g_noalloc_ptr[device] = ptr;
    }
    *out = static_cast<unsigned char*>(g_noalloc_ptr[device]);
    return B2K_OK;
}

}  // namespace b2k

using namespace b2k;

extern "C" int b2k_search_rows_noalloc(int device, const float* x_dev, int64_t n, int d, const float* q_dev,
                                        int64_t nq, int64_t k, float* D_dev, int64_t* I_dev, void* stream) {
    if (d <= 0 || n < 0 || nq < 0 || k <= 0 || (n > 0 && !x_dev) || (nq > 0 && (!q_dev || !D_dev || !I_dev))) {
        set_error("b2k_search_rows_noalloc: bad argument (n=%lld d=%d nq=%lld k=%lld)", (long long)n, d, (long long)nq,
                  (long long)k);
        return B2K_EINVAL;
    }
    if (k > B2K_MAX_K) {
        set_error("b2k_search_rows_noalloc: k=%lld exceeds B2K_MAX_K=%d", (long long)k, B2K_MAX_K);
        return B2K_ELIMIT;
    }
    if (n > 0xFFFFFFFEll) {
        set_error("b2k_search_rows_noalloc: n=%lld exceeds the 32-bit row index of a selection key", (long long)n);
        return B2K_ELIMIT;
    }
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) {
        set_error("b2k_search_rows_noalloc: device %d out of range", device);
        return B2K_EINVAL;
    }
    if (nq == 0) return B2K_OK;

    int prev = -1;
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        set_error("b2k_search_rows_noalloc: no usable CUDA device %d (%s); there is no CPU fallback", device,
                  cudaGetErrorString(e));
        return B2K_ECUDA;
    }
    int rc = B2K_OK;
    do {
        int sm_count = 0;
        e = cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess || sm_count <= 0) { set_error("b2k_search_rows_noalloc: %s", cudaGetErrorString(e)); rc = B2K_ECUDA; break; }
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const int kk = static_cast<int>(k);
        if (n == 0) { rc = launch_fill_empty(D_dev, I_dev, nq * kk, st); break; }

        unsigned char* scratch = nullptr;
        rc = noalloc_scratch_ptr(device, &scratch);  // first call on this device: symbol look-up
        if (rc) break;
        unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
        u64* lists = reinterpret_cast<u64*>(scratch + NOALLOC_HEADER);
        const size_t max_keys = (B2K_NOALLOC_SCRATCH_BYTES - NOALLOC_HEADER) / sizeof(u64);

        const int grid = scan_grid(n, sm_count, 3);
        if ((size_t)grid * kk > max_keys) {
            set_error("b2k_search_rows_noalloc: %d CTAs x k=%d keys do not fit the %u-byte static scratch", grid, kk,
                      (unsigned)B2K_NOALLOC_SCRATCH_BYTES);
            rc = B2K_ELIMIT;
            break;
        }
        for (int64_t q0 = 0; q0 < nq && !rc;) {
            int qt = scan_pick_qt(nq - q0, kk);
            while (qt > 1 && (size_t)grid * qt * kk > max_keys) qt >>= 1;
            // the ticket and the work counter must start at zero; the lists need no clearing (each CTA
            // rewrites its slots)
            e = cudaMemsetAsync(ticket, 0, 2 * sizeof(unsigned int), st);
            if (e != cudaSuccess) { set_error("b2k_search_rows_noalloc: %s", cudaGetErrorString(e)); rc = B2K_ECUDA; break; }
            rc = launch_scan(x_dev, q_dev + q0 * (int64_t)d, n, d, qt, kk, grid, lists, ticket, D_dev + q0 * kk,
                             I_dev + q0 * kk, st);
            q0 += qt;
        }
    } while (0);
    if (prev >= 0) cudaSetDevice(prev);
    return rc;
}
