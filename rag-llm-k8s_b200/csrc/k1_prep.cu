// k1_prep.cu -- K1: per-row preparation done once at add() time, plus the synthetic-row
// generator used by the benchmarks.
//
// Stands in for the `fvec_norms_L2sqr` pass that faiss repeats over the whole database on
// EVERY batched search behind `index.search` (/root/reference/llm/rag.py:116): here the
// squared norms are computed when rows arrive (`index.add`, rag.py:80) and kept in HBM.
// The same pass can also emit the TF32 residual rows the tensor-core kernel (K3) needs:
//     x = hi + lo,  hi = x with the low 13 mantissa bits cleared (what a kind::tf32 MMA
//     reads from a raw fp32 operand), lo = x - hi (exact in fp32).
// Roofline: HBM; algorithmic bytes per row = d*4 read (+ d*4 written when lo is emitted).
#include <cuda_bf16.h>

#include "common.cuh"
#include "kernels.h"

namespace b2k {

__device__ __forceinline__ float tf32_residual(float x) {
    float hi = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
    return x - hi;  // exact: at most 13 significant bits
}

// Largest finite squared norm seen, kept as float bits (non-negative floats order like unsigned
// integers): one atomic per warp.  The tensor-core screening pass sizes its error slack with it.
__device__ __forceinline__ void publish_max(float vmax, unsigned int* max_norm_bits, int lane) {
    if (!max_norm_bits) return;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    if (lane == 0 && vmax > 0.f) atomicMax(max_norm_bits, __float_as_uint(vmax));
}

// 8 lanes per row (128 B per request), 4 rows per warp; d % 4 == 0 path.
__global__ void __launch_bounds__(256)
row_prep_vec_kernel(const float* __restrict__ X, int64_t n, int d, float* __restrict__ norms,
                    float* __restrict__ xlo, unsigned int* __restrict__ max_norm_bits) {
    const int lane = threadIdx.x & 31;
    float vmax = 0.f;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int g = lane >> 3, cl = (lane & 7) * 4;
    for (int64_t r0 = warp * 4; r0 < n; r0 += nwarps * 4) {
        const int64_t row = r0 + g;
        float acc = 0.f;
        if (row < n) {
            const float* xr = X + row * (int64_t)d;
            for (int c = cl; c < d; c += 32) {
                float4 v = ldg_stream(reinterpret_cast<const float4*>(xr + c));
                acc = fmaf(v.x, v.x, acc); acc = fmaf(v.y, v.y, acc);
                acc = fmaf(v.z, v.z, acc); acc = fmaf(v.w, v.w, acc);
                if (xlo) {
                    float4 l = make_float4(tf32_residual(v.x), tf32_residual(v.y), tf32_residual(v.z), tf32_residual(v.w));
                    *reinterpret_cast<float4*>(xlo + row * (int64_t)d + c) = l;
                }
            }
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        acc += __shfl_xor_sync(0xffffffffu, acc, 4);
        if (row < n && (lane & 7) == 0) norms[row] = acc;
        if (row < n && acc <= FLT_MAX) vmax = fmaxf(vmax, acc);  // finite norms only (NaN / inf rows never match)
    }
    publish_max(vmax, max_norm_bits, lane);
}

// any d: one warp per row.
__global__ void __launch_bounds__(256)
row_prep_any_kernel(const float* __restrict__ X, int64_t n, int d, float* __restrict__ norms,
                    float* __restrict__ xlo, unsigned int* __restrict__ max_norm_bits) {
    const int lane = threadIdx.x & 31;
    float vmax = 0.f;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = warp; row < n; row += nwarps) {
        const float* xr = X + row * (int64_t)d;
        float acc = 0.f;
        for (int c = lane; c < d; c += 32) {
            float v = __ldg(xr + c);
            acc = fmaf(v, v, acc);
            if (xlo) xlo[row * (int64_t)d + c] = tf32_residual(v);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) norms[row] = acc;
        if (acc <= FLT_MAX) vmax = fmaxf(vmax, acc);
    }
    publish_max(vmax, max_norm_bits, lane);
}

int launch_row_prep(const float* X, int64_t n, int d, float* norms, float* xlo, unsigned int* max_norm_bits,
                    cudaStream_t st) {
    if (n <= 0) return B2K_OK;
    const bool vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0) &&
                     (xlo == nullptr || (reinterpret_cast<uintptr_t>(xlo) & 15) == 0);
    int64_t rows_per_block = vec ? 32 : 8;
    int64_t blocks = (n + rows_per_block - 1) / rows_per_block;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (vec) row_prep_vec_kernel<<<(unsigned)blocks, 256, 0, st>>>(X, n, d, norms, xlo, max_norm_bits);
    else row_prep_any_kernel<<<(unsigned)blocks, 256, 0, st>>>(X, n, d, norms, xlo, max_norm_bits);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

// ---- bf16 copies for the bf16 screening mode of K3 ---------------------------------------------
// xb = x rounded to bf16 (round to nearest even), err2 = |x - xb|^2 per row.  The screening slack
// is built from these measured rounding errors (Cauchy-Schwarz), not from the worst case 2^-8.
// 8 lanes per row, 4 rows per warp; needs d % 8 == 0 (16-byte bf16 row pitch for TMA).
__global__ void __launch_bounds__(256)
row_bf16_kernel(const float* __restrict__ X, int64_t n, int d, __nv_bfloat16* __restrict__ xb,
                float* __restrict__ err2, unsigned int* __restrict__ max_err_bits) {
    const int lane = threadIdx.x & 31;
    float vmax = 0.f;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int g = lane >> 3, cl = (lane & 7) * 4;
    for (int64_t r0 = warp * 4; r0 < n; r0 += nwarps * 4) {
        const int64_t row = r0 + g;
        float acc = 0.f;
        if (row < n) {
            const float* xr = X + row * (int64_t)d;
            for (int c = cl; c < d; c += 32) {
                float4 v = ldg_stream(reinterpret_cast<const float4*>(xr + c));
                __nv_bfloat162 b01 = __floats2bfloat162_rn(v.x, v.y), b23 = __floats2bfloat162_rn(v.z, v.w);
                const float e0 = v.x - __low2float(b01), e1 = v.y - __high2float(b01);
                const float e2 = v.z - __low2float(b23), e3 = v.w - __high2float(b23);
                acc = fmaf(e0, e0, acc); acc = fmaf(e1, e1, acc); acc = fmaf(e2, e2, acc); acc = fmaf(e3, e3, acc);
                uint2 o;
                o.x = *reinterpret_cast<unsigned int*>(&b01);
                o.y = *reinterpret_cast<unsigned int*>(&b23);
                *reinterpret_cast<uint2*>(xb + row * (int64_t)d + c) = o;
            }
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        acc += __shfl_xor_sync(0xffffffffu, acc, 4);
        if (row < n && err2 && (lane & 7) == 0) err2[row] = acc;
        if (row < n && acc <= FLT_MAX) vmax = fmaxf(vmax, acc);  // rows with non-finite values never match anyway
    }
    publish_max(vmax, max_err_bits, lane);
}

int launch_row_bf16(const float* X, int64_t n, int d, void* xb, float* err2, unsigned int* max_err_bits, cudaStream_t st) {
    if (n <= 0) return B2K_OK;
    if (d % 8 != 0 || (reinterpret_cast<uintptr_t>(X) & 15) || (reinterpret_cast<uintptr_t>(xb) & 7)) {
        set_error("bf16 row copies need d %% 8 == 0 and aligned rows");
        return B2K_EINVAL;
    }
    int64_t blocks = (n + 31) / 32;
    if (blocks > 148 * 16) blocks = 148 * 16;
    row_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(X, n, d, static_cast<__nv_bfloat16*>(xb), err2, max_err_bits);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

// ---- -||x||^2 / 2 as three bf16 terms, one 32-byte record per row -----------------------------
// The bf16 screening pass lets the tensor core add the row norms: one more K = 16 MMA per tile with a
// block of ones on the query side and this record on the row side puts  q~.x~ - ||x||^2/2  in the
// accumulator, so the epilogue compares instead of computing.  h1 + h2 + h3 reproduces ||x||^2/2 to
// 2^-24 relative (three 8-bit mantissas); a non-finite norm gives -inf / NaN, which never passes.
__global__ void __launch_bounds__(256)
norm_split_kernel(const float* __restrict__ norms, int64_t n, uint4* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float h = 0.5f * norms[i];
    const __nv_bfloat16 b1 = __float2bfloat16_rn(h);
    const float r1 = h - __bfloat162float(b1);
    const __nv_bfloat16 b2 = __float2bfloat16_rn(r1);
    const float r2 = r1 - __bfloat162float(b2);
    const __nv_bfloat16 b3 = __float2bfloat16_rn(r2);
    const uint32_t u1 = __bfloat16_as_ushort(__hneg(b1)), u2 = __bfloat16_as_ushort(__hneg(b2)), u3 = __bfloat16_as_ushort(__hneg(b3));
    out[2 * i] = make_uint4(u1 | (u2 << 16), u3, 0u, 0u);
    out[2 * i + 1] = make_uint4(0u, 0u, 0u, 0u);
}

int launch_norm_split(const float* norms, int64_t n, void* out, cudaStream_t st) {
    if (n <= 0) return B2K_OK;
    norm_split_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(norms, n, static_cast<uint4*>(out));
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

// ---- outlier rows: a norm cap for the screening slack ------------------------------------------
// The screening slack is proportional to the largest row norm, so a handful of rows with 30 times the typical
// norm would widen every query's zone 30-fold (and push every batched search down the fallback chain).  This
// pass finds the few largest rows: if leaving out at most OUTL_MAX of them lowers the largest squared norm by
// OUTL_GAIN or more, those rows are taken out of the screen -- their norm records become +inf, so the
// norm-folded accumulator is -inf and they never pass -- and the re-rank kernel computes their exact distance
// for every query instead; the slack then uses the cap.  Otherwise nothing changes (count = 0, cap = max).
// `o` (device words): [0] count, [1] bits of the squared-norm cap, [2] bits of the squared rounding-error cap,
// [3] chosen histogram bin, [8 .. 8+OUTL_MAX) row ids.  hist: OUTL_BINS words.  Bins: float bits >> 19 (sixteen
// bins per octave), so the order of the bins is the order of the norms.
static constexpr int OUTL_MAX = 64;
static constexpr int OUTL_BINS = 4096;
static constexpr float OUTL_GAIN = 1.5f;

__global__ void outl_restore_kernel(unsigned int* o, const float* __restrict__ norms, uint4* __restrict__ rec, int64_t n) {
    // rows that were outliers get their finite norm record back (they may not be outliers any more)
    const int i = threadIdx.x;
    if (i < (int)o[0] && i < OUTL_MAX) {
        const int64_t r = o[8 + i];
        if (r < n) {
            const float h = 0.5f * norms[r];
            const __nv_bfloat16 b1 = __float2bfloat16_rn(h);
            const float r1 = h - __bfloat162float(b1);
            const __nv_bfloat16 b2 = __float2bfloat16_rn(r1);
            const __nv_bfloat16 b3 = __float2bfloat16_rn(r1 - __bfloat162float(b2));
            const uint32_t u1 = __bfloat16_as_ushort(__hneg(b1)), u2 = __bfloat16_as_ushort(__hneg(b2)), u3 = __bfloat16_as_ushort(__hneg(b3));
            rec[2 * r] = make_uint4(u1 | (u2 << 16), u3, 0u, 0u);
        }
    }
    __syncthreads();
    if (i == 0) o[0] = 0;
}

__global__ void __launch_bounds__(256)
outl_hist_kernel(const float* __restrict__ norms, int64_t n, unsigned int* __restrict__ hist) {
    __shared__ unsigned int sh[OUTL_BINS];
    for (int b = threadIdx.x; b < OUTL_BINS; b += 256) sh[b] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        const float v = norms[i];
        if (v <= FLT_MAX && v >= 0.f) atomicAdd(&sh[__float_as_uint(v) >> 19], 1u);  // non-finite rows never match anyway
    }
    __syncthreads();
    for (int b = threadIdx.x; b < OUTL_BINS; b += 256) if (sh[b]) atomicAdd(&hist[b], sh[b]);
}

__global__ void outl_pick_kernel(const unsigned int* __restrict__ hist, const unsigned int* __restrict__ max_norm_bits,
                                 const unsigned int* __restrict__ max_err_bits, unsigned int* o) {
    if (threadIdx.x != 0) return;
    // from the top: the lowest bin such that the rows in the bins ABOVE it number at most OUTL_MAX
    unsigned int above = 0;
    int b = OUTL_BINS - 1;
    while (b > 0 && above + hist[b] <= OUTL_MAX) { above += hist[b]; --b; }
    const float vmax = __uint_as_float(*max_norm_bits);
    const float cap = __uint_as_float(((unsigned int)(b + 1)) << 19);  // every row in the bins up to b has a squared norm below this
    if (above > 0 && cap > 0.f && vmax >= OUTL_GAIN * cap) {
        o[1] = __float_as_uint(cap);
        // bf16 (8 significant bits) rounds every element to within 2^-8 of its magnitude, so |x - x~| <= 2^-8 |x| for
        // the rows that stay; the measured maximum, where smaller, is kept
        const float ecap = cap * (1.0f / 65536.f) * 1.002f;
        o[2] = __float_as_uint(fminf(ecap, __uint_as_float(*max_err_bits)));
        o[3] = (unsigned int)b;
    } else {
        o[1] = *max_norm_bits;
        o[2] = *max_err_bits;
        o[3] = OUTL_BINS;  // nothing lies above this bin: no row is taken out
    }
}

__global__ void __launch_bounds__(256)
outl_collect_kernel(const float* __restrict__ norms, int64_t n, uint4* __restrict__ rec, unsigned int* o) {
    const unsigned int bsel = o[3];
    if (bsel >= OUTL_BINS) return;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        const float v = norms[i];
        if (v <= FLT_MAX && v >= 0.f && (__float_as_uint(v) >> 19) > bsel) {
            const unsigned int pos = atomicAdd(&o[0], 1u);
            if (pos < OUTL_MAX) {
                o[8 + pos] = (unsigned int)i;
                rec[2 * i] = make_uint4(0xFF80FF80u, 0xFF80u, 0u, 0u);  // -h = -inf in all three terms: the accumulator becomes -inf
            }
        }
    }
}

int launch_outlier_pass(const float* norms, int64_t n, void* rec, const unsigned int* max_norm_bits,
                        const unsigned int* max_err_bits, unsigned int* o, unsigned int* hist, cudaStream_t st) {
    outl_restore_kernel<<<1, OUTL_MAX, 0, st>>>(o, norms, static_cast<uint4*>(rec), n);
    B2K_CUDA(cudaMemsetAsync(hist, 0, OUTL_BINS * sizeof(unsigned int), st));
    if (n > 0) {
        int64_t blocks = (n + 256 * 16 - 1) / (256 * 16);
        blocks = blocks < 1 ? 1 : (blocks > 148 * 8 ? 148 * 8 : blocks);
        outl_hist_kernel<<<(unsigned)blocks, 256, 0, st>>>(norms, n, hist);
        outl_pick_kernel<<<1, 32, 0, st>>>(hist, max_norm_bits, max_err_bits, o);
        outl_collect_kernel<<<(unsigned)blocks, 256, 0, st>>>(norms, n, static_cast<uint4*>(rec), o);
        count_launch(4);
    } else {
        count_launch(1);
    }
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

// ---- synthetic rows (must stay bit-identical to oracle/flat_l2_oracle.c orc_synth_value) ----
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

__global__ void __launch_bounds__(256)
synth_kernel(uint64_t seed, int64_t row0, int64_t nrows, int d, float* __restrict__ out) {
    const int64_t total = nrows * (int64_t)d;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int64_t r = i / d;
        int c = (int)(i - r * d);
        uint64_t hr = mix64(seed * 0x9E3779B97F4A7C15ull + (uint64_t)(row0 + r) + 0x632BE59BD9B4E019ull);
        uint64_t h = mix64(hr + (uint64_t)c * 0xD1342543DE82EF95ull);
        int s = (int)(h & 0xFFFF) + (int)((h >> 16) & 0xFFFF) + (int)((h >> 32) & 0xFFFF) + (int)(h >> 48) - 131070;
        out[i] = __fmul_rn((float)s, 2.6429e-5f);
    }
}

int launch_synth(uint64_t seed, int64_t row0, int64_t nrows, int d, float* out, cudaStream_t st) {
    if (nrows <= 0) return B2K_OK;
    int64_t total = nrows * (int64_t)d;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    synth_kernel<<<(unsigned)blocks, 256, 0, st>>>(seed, row0, nrows, d, out);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

}  // namespace b2k
