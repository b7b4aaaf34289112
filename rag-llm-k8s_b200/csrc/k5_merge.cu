// k5_merge.cu -- K5: combine step of the row-sharded search.
//
// No faiss call in /root/reference/llm/rag.py corresponds to it (the reference is one
// process, one index).  After every GPU has produced its shard-local top-k for
// `index.search` (rag.py:116) and the lists have been all-gathered, this kernel keeps, per
// query, the k smallest (distance, id) of the nlists*k candidates -- the same
// lexicographic rule as the single-index search.
//
// One CTA per query; candidates sorted in shared memory by a bitonic network.  The payload
// is tiny (nq * nlists * k * 12 bytes), so this kernel is latency-bound by design.
#include "common.cuh"
#include "kernels.h"

namespace b2k {

static constexpr int MERGE_THREADS = 256;

__device__ __forceinline__ bool cand_less(u64 a, u64 b, const int64_t* ids) {
    uint32_t da = (uint32_t)(a >> 32), db = (uint32_t)(b >> 32);
    if (da != db) return da < db;
    if (a == KEY_EMPTY || b == KEY_EMPTY) return a < b;
    int64_t ia = ids[(uint32_t)a], ib = ids[(uint32_t)b];
    if (ia != ib) return ia < ib;
    return (uint32_t)a < (uint32_t)b;
}

// List l holds [nq][k] distances at Din + l*stride_d and ids at Iin + l*stride_i (strides in elements;
// contiguous [nlists][nq][k] arrays have stride nq*k, the packed all-gather records of the sharded
// index have their own).  Ids are translated to global row numbers on the way in: + id_offset[l]
// (contiguous shards), or, with cyc_block = B > 0, by the block-cyclic rule of the sharded indexes
// (global rows go to the shards in blocks of B, round-robin): local row r of shard l is global row
// ((r / B) * nlists + l) * B + r % B.
__global__ void __launch_bounds__(MERGE_THREADS)
merge_lists_kernel(int nlists, int64_t nq, int k, int P, const float* __restrict__ Din, int64_t stride_d,
                   const int64_t* __restrict__ Iin, int64_t stride_i, const int64_t* __restrict__ id_offset,
                   int64_t cyc_block, float* __restrict__ D, int64_t* __restrict__ I) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* keys = reinterpret_cast<u64*>(smem_raw);
    int64_t* ids = reinterpret_cast<int64_t*>(smem_raw + (size_t)P * 8);
    const int tid = threadIdx.x;
    const int64_t q = blockIdx.x;
    const int m = nlists * k;
    for (int s = tid; s < P; s += MERGE_THREADS) {
        u64 key = KEY_EMPTY;
        if (s < m) {
            int l = s / k, j = s - l * k;
            const int64_t o = q * k + j;
            int64_t id = Iin[(int64_t)l * stride_i + o];
            float dv = Din[(int64_t)l * stride_d + o];
            if (id >= 0 && dv < FLT_MAX) {
                if (cyc_block > 0) id = ((id / cyc_block) * nlists + l) * cyc_block + id % cyc_block;
                ids[s] = id + (id_offset ? id_offset[l] : 0);
                key = make_key(dv, (uint32_t)s);
            } else {
                ids[s] = -1;
            }
        }
        keys[s] = key;
    }
    __syncthreads();
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = tid; i < (P >> 1); i += MERGE_THREADS) {
                int lo = 2 * i - (i & (stride - 1));
                int hi = lo + stride;
                bool up = ((lo & size) == 0);
                u64 a = keys[lo], b = keys[hi];
                bool swap = up ? cand_less(b, a, ids) : cand_less(a, b, ids);
                if (swap) { keys[lo] = b; keys[hi] = a; }
            }
            __syncthreads();
        }
    }
    for (int j = tid; j < k; j += MERGE_THREADS) {
        u64 key = j < P ? keys[j] : KEY_EMPTY;
        if (key != KEY_EMPTY) {
            D[q * k + j] = key_dist(key);
            I[q * k + j] = ids[(uint32_t)key];
        } else {
            D[q * k + j] = FLT_MAX;
            I[q * k + j] = -1;
        }
    }
}

// ---- the sharded indexes' combine step: lists that are already sorted, over disjoint id sets --------
// Every shard's result is sorted by (distance, id) and the block-cyclic map keeps that order (a shard's
// global ids grow with its local rows), and no row lives on two shards.  So a candidate's place in the
// merged order is  (its index in its own list) + (for every other list: how many entries are smaller),
// the latter by binary search: nlists * k * (nlists - 1) * log2(k) steps per query instead of a sort of
// nlists * k keys with a block barrier per stage (8 lists of 100: 53 us -> a few us per batch).
__device__ __forceinline__ bool shard_less(uint32_t da, int64_t ia, uint32_t db, int64_t ib) {
    return da < db || (da == db && ia < ib);
}

__global__ void __launch_bounds__(MERGE_THREADS)
merge_sorted_kernel(int nlists, int64_t nq, int k, const float* __restrict__ Din, int64_t stride_d,
                    const int64_t* __restrict__ Iin, int64_t stride_i, int64_t cyc_block, float* __restrict__ D,
                    int64_t* __restrict__ I) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int m = nlists * k;
    int64_t* ids = reinterpret_cast<int64_t*>(smem_raw);                       // [m] global ids, -1 = padding
    uint32_t* dist = reinterpret_cast<uint32_t*>(smem_raw + (size_t)m * 8);    // [m] distance bits
    int* nvalid = reinterpret_cast<int*>(smem_raw + (size_t)m * 12);           // [nlists]
    const int tid = threadIdx.x;
    const int64_t q = blockIdx.x;
    for (int l = tid; l < nlists; l += MERGE_THREADS) nvalid[l] = 0;
    __syncthreads();
    for (int s = tid; s < m; s += MERGE_THREADS) {
        const int l = s / k, j = s - l * k;
        const int64_t o = q * k + j;
        int64_t id = Iin[(int64_t)l * stride_i + o];
        const float dv = Din[(int64_t)l * stride_d + o];
        if (id >= 0 && dv < FLT_MAX) {
            if (cyc_block > 0) id = ((id / cyc_block) * nlists + l) * cyc_block + id % cyc_block;
            uint32_t b = __float_as_uint(dv);
            if (b == 0x80000000u) b = 0u;
            dist[s] = b;
            atomicAdd(&nvalid[l], 1);
        } else {
            id = -1;
            dist[s] = 0xffffffffu;
        }
        ids[s] = id;
    }
    __syncthreads();
    int total = 0;
    for (int l = 0; l < nlists; ++l) total += nvalid[l];
    for (int s = tid; s < m; s += MERGE_THREADS) {
        const int64_t id = ids[s];
        if (id < 0) continue;
        const int l = s / k;
        int rank = s - l * k;
        const uint32_t dv = dist[s];
        for (int l2 = 0; l2 < nlists && rank < k; ++l2) {
            if (l2 == l) continue;
            // entries of list l2 smaller than this candidate (its valid prefix is sorted)
            int lo = 0, hi = nvalid[l2];
            const int base = l2 * k;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (shard_less(dist[base + mid], ids[base + mid], dv, id)) lo = mid + 1;
                else hi = mid;
            }
            rank += lo;
        }
        if (rank < k) {
            D[q * k + rank] = __uint_as_float(dv);
            I[q * k + rank] = id;
        }
    }
    for (int j = total + tid; j < k; j += MERGE_THREADS) {
        D[q * k + j] = FLT_MAX;
        I[q * k + j] = -1;
    }
}

static int launch_merge_any(int nlists, int64_t nq, int k, const float* Din, int64_t stride_d, const int64_t* Iin,
                            int64_t stride_i, const int64_t* id_offset_dev, int64_t cyc_block, float* D, int64_t* I,
                            cudaStream_t st) {
    if (nq <= 0) return B2K_OK;
    const int m = nlists * k;
    if (m > 8192) {
        set_error("merge: nlists*k = %d exceeds 8192", m);
        return B2K_ELIMIT;
    }
    const int P = next_pow2(m < 2 ? 2 : m);
    size_t smem = (size_t)P * 8 + (size_t)m * 8;
    if (smem > 48 * 1024)  // per-device attribute: set on every large launch (a handle may live on any GPU)
        B2K_CUDA(cudaFuncSetAttribute(merge_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    merge_lists_kernel<<<(unsigned)nq, MERGE_THREADS, smem, st>>>(nlists, nq, k, P, Din, stride_d, Iin, stride_i, id_offset_dev,
                                                                  cyc_block, D, I);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

int launch_merge_lists(int nlists, int64_t nq, int k, const float* Din, const int64_t* Iin,
                       const int64_t* id_offset_dev, float* D, int64_t* I, cudaStream_t st) {
    return launch_merge_any(nlists, nq, k, Din, nq * k, Iin, nq * k, id_offset_dev, 0, D, I, st);
}

int launch_merge_shards(int nlists, int64_t nq, int k, const float* Din, int64_t stride_d, const int64_t* Iin,
                        int64_t stride_i, int64_t cyc_block, float* D, int64_t* I, cudaStream_t st) {
    if (nq <= 0) return B2K_OK;
    const int m = nlists * k;
    if (m > 8192) { set_error("merge: nlists*k = %d exceeds 8192", m); return B2K_ELIMIT; }
    const size_t smem = (size_t)m * 12 + (size_t)nlists * 4;
    if (smem > 48 * 1024)
        B2K_CUDA(cudaFuncSetAttribute(merge_sorted_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    merge_sorted_kernel<<<(unsigned)nq, MERGE_THREADS, smem, st>>>(nlists, nq, k, Din, stride_d, Iin, stride_i, cyc_block, D, I);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

// ---- one-shot all-gather over peer memory (the exchange step of the one-process-per-GPU sharded index) ----
// Every rank stores its record (its shard's top-k + the screening word, a megabyte) straight into slot
// `my_rank` of every rank's receive buffer over NVLink -- plain 16-byte stores to peer addresses -- and then
// raises its sequence number in that rank's flag row.  No rendezvous, no staging, no reduction: for a payload
// this small the exchange is bound by launch and signalling latency, where a general collective pays ~100 us
// at 8 ranks and this pays one launch plus the store time.
// grid = (CTAs per destination, destinations).  The last CTA of a destination publishes the flag.
static constexpr int PUSH_THREADS = 256;
__global__ void __launch_bounds__(PUSH_THREADS)
push_record_kernel(const uint4* __restrict__ src, int64_t n16, const uint64_t* __restrict__ dst_ptrs,
                   const uint64_t* __restrict__ flag_ptrs, int my_rank, int64_t slot_bytes, int flag_index,
                   unsigned int seq, unsigned int* __restrict__ done) {
    const int p = blockIdx.y;
    uint4* dst = reinterpret_cast<uint4*>(dst_ptrs[p] + (uint64_t)my_rank * (uint64_t)slot_bytes);
    for (int64_t i = (int64_t)blockIdx.x * PUSH_THREADS + threadIdx.x; i < n16; i += (int64_t)gridDim.x * PUSH_THREADS) dst[i] = src[i];
    __threadfence_system();  // the stores above are visible system-wide before the flag below
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(&done[p], 1u) == gridDim.x - 1) {
            done[p] = 0;  // ready for the next call (launches on one stream do not overlap)
            volatile unsigned int* flag = reinterpret_cast<volatile unsigned int*>(flag_ptrs[p]) + flag_index;
            __threadfence_system();
            *flag = seq;
        }
    }
}

// Waits until every rank's record of call `seq` has landed here (flags[i] >= seq for i < n).  The wait is
// bounded by wall time -- two minutes, far beyond any skew between ranks that are alive (one of them may be
// busy on its host for seconds) --: a peer that never shows up (it died, or the ranks' calls went out of step)
// raises *err instead of hanging the GPU; the host then fails the search.
__global__ void wait_flags_kernel(const unsigned int* flags, int n, unsigned int seq, unsigned int* err) {
    const int i = threadIdx.x;
    if (i >= n) return;
    const volatile unsigned int* f = flags + i;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        if ((int)(*f - seq) >= 0) { __threadfence_system(); return; }
        __nanosleep(128);
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > 120ull * 1000000000ull) break;
    }
    atomicExch(err, 1u + (unsigned)i);
}

int launch_push_record(const void* src, int64_t bytes, const uint64_t* dst_ptrs_dev, const uint64_t* flag_ptrs_dev, int nranks,
                       int my_rank, int64_t slot_bytes, int flag_index, unsigned int seq, unsigned int* done_dev, cudaStream_t st) {
    if (bytes % 16 != 0 || (reinterpret_cast<uintptr_t>(src) & 15) || slot_bytes % 16 != 0) { set_error("push: record must be 16-byte aligned and sized"); return B2K_EINVAL; }
    const int64_t n16 = bytes / 16;
    int ctas = (int)((n16 + PUSH_THREADS * 4 - 1) / (PUSH_THREADS * 4));
    ctas = ctas < 1 ? 1 : (ctas > 16 ? 16 : ctas);
    push_record_kernel<<<dim3(ctas, nranks), PUSH_THREADS, 0, st>>>(static_cast<const uint4*>(src), n16, dst_ptrs_dev, flag_ptrs_dev,
                                                                   my_rank, slot_bytes, flag_index, seq, done_dev);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

int launch_wait_flags(const unsigned int* flags, int n, unsigned int seq, unsigned int* err, cudaStream_t st) {
    wait_flags_kernel<<<1, 64, 0, st>>>(flags, n, seq, err);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

}  // namespace b2k
