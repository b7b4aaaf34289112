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
    return launch_merge_any(nlists, nq, k, Din, stride_d, Iin, stride_i, nullptr, cyc_block, D, I, st);
}

}  // namespace b2k
