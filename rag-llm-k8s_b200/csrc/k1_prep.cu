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
