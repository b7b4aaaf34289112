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
int k3_search(K3State* s, const float* x, const float* xlo, const float* norms, int64_t n, int64_t capacity,
              int d, const float* q, int64_t nq, int k, float* D, int64_t* I, unsigned char* scratch,
              cudaStream_t st);

// optional CUDA-event timing of the main kernel (bench.py's roofline numbers)
void k3_set_timing(K3State* s, bool on);
void k3_set_cta_group(K3State* s, int cg);
void k3_set_sep_mode(K3State* s, int m);     // separate accumulator for the small cross terms: 0 auto, 1 on, 2 off  // 1 = one CTA per tile, 2 = CTA pairs (cta_group::2)
int k3_last_kernel_ms(K3State* s, float* ms);

}  // namespace b2k
