// k3_api.h -- host interface of the tensor-core batched-search kernel (k3_tc.cu).
#pragma once
#include "common.cuh"

namespace b2k {

struct K3State;  // tensor maps, cached launch configuration

// true if the tensor-core path can serve this (d, k): needs d % 4 == 0 (TMA row stride).
bool k3_supported(int d, int k);
size_t k3_scratch_bytes(int sm_count, int d, int64_t nq, int k);
int k3_create(K3State** out, int device, int sm_count);
void k3_destroy(K3State* s);
void k3_invalidate_maps(K3State* s);  // row storage moved / resized
// D[nq][k], I[nq][k] (device) <- k nearest of rows [0,n) for each of the nq queries (device).
// exact = 1: 3xTF32 (needs xlo).  exact = 0: TF32 screening + exact fp32 re-rank (xlo unused;
// max_norm_bits = float bits of the largest squared row norm); synchronises the stream and sets
// *overflowed when the slack zone did not fit, in which case the caller must rerun with exact = 1.
int k3_search(K3State* s, const float* x, const float* xlo, const float* norms, const unsigned int* max_norm_bits,
              int64_t n, int64_t capacity, int d, const float* q, int64_t nq, int k, float* D, int64_t* I,
              unsigned char* scratch, cudaStream_t st, int exact, int* overflowed);

// optional CUDA-event timing of the main kernel (bench.py's roofline numbers)
void k3_set_timing(K3State* s, bool on);
void k3_set_cta_group(K3State* s, int cg);
void k3_set_sep_mode(K3State* s, int m);     // separate accumulator for the small cross terms: 0 auto, 1 on, 2 off  // 1 = one CTA per tile, 2 = CTA pairs (cta_group::2)
int k3_last_kernel_ms(K3State* s, float* ms);
int k3_last_mode(const K3State* s);  // 1 = 3xTF32, 2 = screening + exact re-rank

}  // namespace b2k
