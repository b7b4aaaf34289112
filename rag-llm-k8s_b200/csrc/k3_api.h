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
// D[nq][k], I[nq][k] (device) <- k nearest of rows [0,n) for each of the nq queries (device).
// mode 1: 3xTF32 (needs xlo).  mode 2: TF32 screening + exact fp32 re-rank (max_norm_bits = float
// bits of the largest squared row norm).  mode 3: bf16 screening + exact re-rank (needs xbf, the
// bf16 copies of the rows, max_err_bits = float bits of the largest squared rounding-error norm
// of a row, xnb = the 32-byte norm records that go with xbf, outl = the outlier words of launch_outlier_pass, and d % 8 == 0).  mode 4: the 3xTF32 product (needs xlo) used as a screen, ~10^4 times
// narrower slack than 2 / 3, + exact re-rank.  The screening modes synchronise the stream and set *overflowed when
// the slack zone did not fit, in which case the caller must rerun with mode 4 and, failing that,
// with the exact scan kernel.
int k3_search(K3State* s, const float* x, const float* xlo, const void* xbf, const void* xnb, const unsigned int* outl,
              const float* norms,
              const unsigned int* max_norm_bits, const unsigned int* max_err_bits, int64_t n, int d, const float* q,
              int64_t nq, int k, float* D, int64_t* I, unsigned char* scratch, cudaStream_t st, int mode,
              int* overflowed);

// optional CUDA-event timing of the main kernel (bench.py's roofline numbers)
void k3_set_timing(K3State* s, bool on);
void k3_set_cta_group(K3State* s, int cg);
void k3_set_sep_mode(K3State* s, int m);     // separate accumulator for the small cross terms: 0 auto, 1 on, 2 off  // 1 = one CTA per tile, 2 = CTA pairs (cta_group::2)
void k3_set_segments(K3State* s, int on);   // main pass in segments with bound refreshes in between (default on)
void k3_set_epi_halves(K3State* s, int nh); // epilogue warps: 0 auto, 1 four (one per scheduler), 2 eight
// flag != null: the screening modes raise *flag (device word, never reset by the library) instead of
// synchronising and reporting through *overflowed; the caller checks it and repeats the search with mode 1
void k3_set_ext_flag(K3State* s, unsigned int* flag);
// dense-rate probe: the main pass runs as usual (TMA, MMAs, TMEM loads, the filter's arithmetic) but no
// key ever passes, so its time is the kernel's ceiling for that product kind; results are all padding
void k3_set_probe(K3State* s, int on);
void k3_set_query_stationary(K3State* s, int on);  // 1: resident query blocks where they fit; 0 (default): both operands stream
void k3_set_prefetch_tiles(K3State* s, int n);      // L2 prefetch of the row tiles n tiles ahead of the scheduler (0 = off)
void k3_set_warm_keys(K3State* s, int n);           // warm-start sample per query in multiples of k (default 12)
int k3_last_kernel_ms(K3State* s, float* ms);
int k3_last_mode(const K3State* s);  // 1 = 3xTF32, 2 = TF32 screening + exact re-rank, 3 = bf16 screening + exact re-rank

}  // namespace b2k
