// kernels.h -- host-callable launchers of the CUDA kernels (internal to libb200knn).
#pragma once
#include "common.cuh"

namespace b2k {

// K1 (k1_prep.cu): per-row squared norms (+ optional TF32 residual rows) for rows [0,n) of X.
// max_norm_bits (nullable): float bits of the largest finite squared norm, updated with atomicMax.
int launch_row_prep(const float* X, int64_t n, int d, float* norms, float* xlo, unsigned int* max_norm_bits,
                    cudaStream_t st);
// bf16 (round-to-nearest) copies of rows [0,n) of X into xb, with the squared rounding-error norm per row
// (err2, nullable) and its running maximum (max_err_bits, nullable, float bits).  Needs d % 8 == 0.
int launch_row_bf16(const float* X, int64_t n, int d, void* xb, float* err2, unsigned int* max_err_bits, cudaStream_t st);
// -||x||^2/2 of rows [0,n) as three bf16 terms in a 32-byte record per row (the bf16 screening pass adds it by MMA)
int launch_norm_split(const float* norms, int64_t n, void* out, cudaStream_t st);
// outlier rows (k1_prep.cu): o = device words [0] count, [1] squared-norm cap bits, [2] squared rounding-error cap bits,
// [8..72) row ids; hist = 4096 words of scratch.  Rewrites the norm records of the rows it takes out of the screen.
static constexpr int B2K_OUTL_WORDS = 80;
static constexpr int B2K_OUTL_HIST_WORDS = 4096;
int launch_outlier_pass(const float* norms, int64_t n, void* rec, const unsigned int* max_norm_bits,
                        const unsigned int* max_err_bits, unsigned int* o, unsigned int* hist, cudaStream_t st);
int launch_synth(uint64_t seed, int64_t row0, int64_t nrows, int d, float* out, cudaStream_t st);

// K2 (k2_scan.cu)
int scan_pick_P(int k);
int scan_pick_qt(int64_t nq_left, int k);
int scan_grid(int64_t n, int sm_count, int ctas_per_sm);
size_t scan_partial_bytes(int grid, int qt, int k);
// bf16 = true: X holds bf16 rows and D receives screening distances; run_if (nullable): do nothing unless *run_if != 0
int launch_scan(const void* X, const float* Q, int64_t n, int d, int qt, int k, int grid,
                u64* partial, unsigned int* done, float* D, int64_t* I, cudaStream_t st, bool bf16 = false,
                const unsigned int* run_if = nullptr);
// exact fp32 re-rank of the bf16 scan's kp candidates per query -> k best; raises *flag when kp did not cover the slack zone
int launch_scan_rerank(const float* X, const float* Q, int64_t nq, int d, int k, int kp, const float* Da,
                       const int64_t* Ia, const unsigned int* max_err_bits, float* D, int64_t* I,
                       unsigned int* flag, cudaStream_t st);
int launch_fill_empty(float* D, int64_t* I, int64_t count, cudaStream_t st);

// K5 (k5_merge.cu): per query, k smallest (distance,id) of nlists*k candidates.
int launch_merge_lists(int nlists, int64_t nq, int k, const float* Din, const int64_t* Iin,
                       const int64_t* id_offset_dev, float* D, int64_t* I, cudaStream_t st);
// the same over lists with their own strides (elements) and block-cyclic shard-local ids (cyc_block rows per block; 0 = ids are global)
int launch_merge_shards(int nlists, int64_t nq, int k, const float* Din, int64_t stride_d, const int64_t* Iin,
                        int64_t stride_i, int64_t cyc_block, float* D, int64_t* I, cudaStream_t st);

// one-shot all-gather over peer memory: store `bytes` of src into slot my_rank of every rank's receive buffer
// (dst_ptrs_dev: device array of nranks base addresses), then publish seq in word flag_index of every rank's flag row
int launch_push_record(const void* src, int64_t bytes, const uint64_t* dst_ptrs_dev, const uint64_t* flag_ptrs_dev, int nranks,
                       int my_rank, int64_t slot_bytes, int flag_index, unsigned int seq, unsigned int* done_dev, cudaStream_t st);
// bounded wait for flags[0..n) >= seq; raises *err (1 + index of the missing rank) on time-out
int launch_wait_flags(const unsigned int* flags, int n, unsigned int seq, unsigned int* err, cudaStream_t st);

}  // namespace b2k
