// k2_scan.cu -- K2: the HBM-bound exact scan with fused top-k (nq < 20, and the generic
// fallback for any nq / any d).
//
// Stands in for faiss `exhaustive_L2sqr_seq` + heap result handler behind
// `index.search(xq, k)` at /root/reference/llm/rag.py:116 (nq = 1, k = 5 there).
// Numerics follow the oracle's rule (1): D = sum_t (q_t - x_t)^2 in fp32 (exact 0 for a
// query equal to a stored row).
//
// Design (B200): the database is read exactly once per group of <= 8 queries with 128-bit
// non-allocating loads, 8 lanes per row (one 128-byte line per row segment per request),
// 16 rows in flight per warp.  Selection never leaves the SM: each CTA keeps, per query,
// an append buffer of 64-bit (distance,row) keys in shared memory guarded by a running
// threshold; a full buffer is compacted by a block-wide bitonic sort.  Each CTA writes its
// k best keys, and the LAST CTA to finish (atomic ticket) merges all CTAs' lists and
// writes D / I -- so the whole search is ONE launch per query group and no distance ever
// reaches HBM.  Roofline: HBM; algorithmic bytes per launch = n * d * 4.
//
// BF variant (screening): the same kernel over the bf16 copies of the rows (half the bytes).
// d~ = |q - x~|^2 is computed in fp32 from the fp32 query and the bf16 row x~; by the triangle
// inequality | |q - x~| - |q - x| | <= |x - x~| <= E (E = largest rounding-error norm of a row,
// tracked by K1), so every exact top-k row satisfies sqrt(d~) <= sqrt(T~) + 2E with T~ the k-th
// smallest d~.  The scan therefore selects k' > k candidates by d~, and the re-rank kernel
// (scan_rerank_kernel) checks that the k'-th d~ lies strictly beyond that zone (else it raises a
// flag and the fp32 scan, launched right behind with "run only if flagged", redoes the search),
// computes the exact fp32 distances of the k' rows and writes the k best.  Results are exact.
#include "common.cuh"
#include "kernels.h"

namespace b2k {

static constexpr int SCAN_THREADS = 256;
static constexpr int SCAN_ROWS_PER_ITER = 128;  // 8 warps x 16 rows
static constexpr int SCAN_FINAL_P = 4096;       // final-merge buffer (keys)

struct ScanSmem {
    float* qs;   // [QT][dq]
    u64* buf;    // [max(QT*P, SCAN_FINAL_P)]
    u64* kth;    // [QT]  current k-th best key (strict upper bound for new keys)
    float* tau;  // [QT]  its distance
    int* cnt;    // [QT]
    u64* red;    // [2 * 8] block-reduction scratch
};

__device__ __forceinline__ ScanSmem carve(unsigned char* base, int QT, int dq, int P) {
    ScanSmem s;
    size_t off = 0;
    s.qs = (float*)(base + off); off += (size_t)QT * dq * 4;  // multiple of 16 bytes
    s.buf = (u64*)(base + off);
    size_t nbuf = (size_t)QT * P > (size_t)SCAN_FINAL_P ? (size_t)QT * P : (size_t)SCAN_FINAL_P;
    off += nbuf * 8;
    s.kth = (u64*)(base + off); off += (size_t)QT * 8;
    s.red = (u64*)(base + off); off += 16 * 8;
    s.tau = (float*)(base + off); off += (size_t)QT * 4;
    s.cnt = (int*)(base + off);
    return s;
}

size_t scan_smem_bytes(int QT, int d, int P) {
    int dq = (d + 3) & ~3;
    size_t nbuf = (size_t)QT * P > (size_t)SCAN_FINAL_P ? (size_t)QT * P : (size_t)SCAN_FINAL_P;
    return (size_t)QT * dq * 4 + nbuf * 8 + (size_t)QT * 8 + 16 * 8 + (size_t)QT * 4 + (size_t)QT * 4 + 16;
}

// Sort the first *cnt keys of buf (capacity P, power of two), keep the best k, refresh the
// threshold.  Block-wide; *cnt must be stable (call after a barrier).
__device__ __forceinline__ void compact(u64* buf, int P, int* cnt, u64* kth, float* tau, int k) {
    const int tid = threadIdx.x, nt = blockDim.x;
    const int c = *cnt;
    int Ps = 2;
    while (Ps < c) Ps <<= 1;
    if (Ps > P) Ps = P;
    for (int i = c + tid; i < Ps; i += nt) buf[i] = KEY_EMPTY;
    __syncthreads();
    block_bitonic_sort(buf, Ps);
    if (tid == 0) {
        int nc = c < k ? c : k;
        *cnt = nc;
        if (nc >= k) { *kth = buf[k - 1]; *tau = key_dist(buf[k - 1]); }
    }
    __syncthreads();
}

// 8 columns of a row as fp32: two float4 of an fp32 row, or one 16-byte group of 8 bf16
// (bf16 -> fp32 is a 16-bit shift, exact).
struct RowSeg8 { float v[8]; };
__device__ __forceinline__ RowSeg8 unpack_bf16x8(const uint4& w) {
    RowSeg8 r;
    r.v[0] = __uint_as_float(w.x << 16); r.v[1] = __uint_as_float(w.x & 0xffff0000u);
    r.v[2] = __uint_as_float(w.y << 16); r.v[3] = __uint_as_float(w.y & 0xffff0000u);
    r.v[4] = __uint_as_float(w.z << 16); r.v[5] = __uint_as_float(w.z & 0xffff0000u);
    r.v[6] = __uint_as_float(w.w << 16); r.v[7] = __uint_as_float(w.w & 0xffff0000u);
    return r;
}

// BF: X points to bf16 rows (d % 8 == 0), see the file header.  run_if (nullable): the launch is a
// conditional repeat and does nothing unless *run_if != 0.
template <int QT, bool VEC, bool BF>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_l2_topk_kernel(const void* __restrict__ Xv, const float* __restrict__ Q, int64_t n, int d,
                    int k, int P, u64* __restrict__ partial,
                    unsigned int* __restrict__ done, float* __restrict__ D, int64_t* __restrict__ I,
                    const unsigned int* __restrict__ run_if) {
    if (run_if && *run_if == 0u) return;
    const float* __restrict__ X = static_cast<const float*>(Xv);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int dq = BF ? ((d + 7) & ~7) : ((d + 3) & ~3);
    ScanSmem s = carve(smem_raw, QT, dq, P);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __shared__ int s_is_last;

    for (int i = tid; i < QT * dq; i += SCAN_THREADS) {
        int q = i / dq, c = i - q * dq;
        s.qs[i] = c < d ? Q[(int64_t)q * d + c] : 0.f;
    }
    if (tid < QT) {
        s.cnt[tid] = 0;
        s.kth[tid] = ((u64)FLT_MAX_BITS << 32);  // only distances < FLT_MAX are ever selected
        s.tau[tid] = FLT_MAX;
    }
    __syncthreads();

    // Rows are handed out one iteration (128 rows) at a time from a global counter (done[1]): with
    // a static split the CTAs finished up to 12 % apart (measured with %globaltimer at 10M x 384)
    // and the last-CTA merge has to wait for the slowest.  Thread 0 fetches the NEXT iteration's
    // index while the current one computes; the iteration's own barrier publishes it, so the
    // dynamic schedule costs no extra synchronisation.
    __shared__ unsigned int s_chunk[2];
    if (tid == 0) s_chunk[0] = atomicAdd(done + 1, 1u);
    __syncthreads();
    for (unsigned int it = 0;; ++it) {
        const int64_t base = (int64_t)s_chunk[it & 1] * SCAN_ROWS_PER_ITER;
        if (base >= n) break;
        const int64_t row_begin = base;
        const int64_t row_end = base + SCAN_ROWS_PER_ITER < n ? base + SCAN_ROWS_PER_ITER : n;
        if (tid == 0) s_chunk[(it + 1) & 1] = atomicAdd(done + 1, 1u);
        float tau_r[QT];
#pragma unroll
        for (int q = 0; q < QT; ++q) tau_r[q] = s.tau[q];
        int flag = 0;

        if (BF) {
            // bf16 rows: 8 lanes per row, 16 bytes = 8 columns per lane and request, 64 columns per step
            const unsigned short* Xb = static_cast<const unsigned short*>(Xv);
            const int g = lane >> 3, cl = (lane & 7) * 8;
            float acc[4][QT];
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int q = 0; q < QT; ++q) acc[u][q] = 0.f;
            int64_t rows[4];
            const unsigned short* rp[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                rows[u] = base + warp * 16 + u * 4 + g;
                int64_t rr = rows[u] < row_end ? rows[u] : row_begin;
                rp[u] = Xb + rr * (int64_t)d + cl;
            }
            uint4 xv[4];
            {
                const bool cok0 = cl < d;
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    xv[u] = cok0 ? ldg_stream(reinterpret_cast<const uint4*>(rp[u])) : make_uint4(0u, 0u, 0u, 0u);
            }
#pragma unroll 1
            for (int c0 = 0; c0 < d; c0 += 64) {
                const bool cok = (c0 + cl) < d;
                const bool nok = (c0 + 64 + cl) < d;
                uint4 xn[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    xn[u] = nok ? ldg_stream(reinterpret_cast<const uint4*>(rp[u] + c0 + 64)) : make_uint4(0u, 0u, 0u, 0u);
                RowSeg8 xs[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) xs[u] = unpack_bf16x8(xv[u]);
#pragma unroll
                for (int q = 0; q < QT; ++q) {
                    // beyond d both the query (zero padding) and the row (zero bits) are 0
                    const float4 qa = cok ? *reinterpret_cast<const float4*>(s.qs + q * dq + c0 + cl) : make_float4(0.f, 0.f, 0.f, 0.f);
                    const float4 qb = cok ? *reinterpret_cast<const float4*>(s.qs + q * dq + c0 + cl + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                    const float qv[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
#pragma unroll
                        for (int t = 0; t < 8; ++t) {
                            const float a = qv[t] - xs[u].v[t];
                            acc[u][q] = fmaf(a, a, acc[u][q]);
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) xv[u] = xn[u];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int q = 0; q < QT; ++q) {
                    float v = acc[u][q];
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    v += __shfl_xor_sync(0xffffffffu, v, 4);
                    acc[u][q] = v;
                }
            if ((lane & 7) == 0) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (rows[u] < row_end) {
#pragma unroll
                        for (int q = 0; q < QT; ++q) {
                            float dv = acc[u][q];
                            if (dv <= tau_r[q]) {
                                u64 key = make_key(dv, (uint32_t)rows[u]);
                                if (key < s.kth[q]) {
                                    int pos = atomicAdd(&s.cnt[q], 1);
                                    s.buf[q * P + pos] = key;
                                    if (pos + 1 > P - SCAN_ROWS_PER_ITER) flag = 1;
                                }
                            }
                        }
                    }
                }
            }
        } else if (VEC) {
            // 8 lanes per row, 4 rows per warp request, 4 requests (u) in flight.
            const int g = lane >> 3, cl = (lane & 7) * 4;
            float acc[4][QT];
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int q = 0; q < QT; ++q) acc[u][q] = 0.f;
            int64_t rows[4];
            const float* rp[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                rows[u] = base + warp * 16 + u * 4 + g;
                int64_t rr = rows[u] < row_end ? rows[u] : row_begin;  // clamp: load something valid
                rp[u] = X + rr * (int64_t)d + cl;
            }
            // software pipeline: the row segments of the next 32 columns are in flight while the
            // current ones are consumed (the loop is otherwise load -> wait -> compute per warp,
            // and the per-iteration barrier keeps the warps of a CTA in step)
            float4 xv[4];
            {
                const bool cok0 = cl < d;
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    xv[u] = cok0 ? ldg_stream(reinterpret_cast<const float4*>(rp[u])) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll 1
            for (int c0 = 0; c0 < d; c0 += 32) {
                const bool cok = (c0 + cl) < d;
                const bool nok = (c0 + 32 + cl) < d;
                float4 xn[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    xn[u] = nok ? ldg_stream(reinterpret_cast<const float4*>(rp[u] + c0 + 32)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int q = 0; q < QT; ++q) {
                    float4 qv = cok ? *reinterpret_cast<const float4*>(s.qs + q * dq + c0 + cl)
                                    : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        float a = qv.x - xv[u].x, b = qv.y - xv[u].y, c = qv.z - xv[u].z, e = qv.w - xv[u].w;
                        acc[u][q] = fmaf(a, a, acc[u][q]);
                        acc[u][q] = fmaf(b, b, acc[u][q]);
                        acc[u][q] = fmaf(c, c, acc[u][q]);
                        acc[u][q] = fmaf(e, e, acc[u][q]);
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) xv[u] = xn[u];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int q = 0; q < QT; ++q) {
                    float v = acc[u][q];
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    v += __shfl_xor_sync(0xffffffffu, v, 4);
                    acc[u][q] = v;
                }
            if ((lane & 7) == 0) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (rows[u] < row_end) {
#pragma unroll
                        for (int q = 0; q < QT; ++q) {
                            float dv = acc[u][q];
                            if (dv <= tau_r[q]) {
                                u64 key = make_key(dv, (uint32_t)rows[u]);
                                if (key < s.kth[q]) {
                                    int pos = atomicAdd(&s.cnt[q], 1);
                                    s.buf[q * P + pos] = key;
                                    if (pos + 1 > P - SCAN_ROWS_PER_ITER) flag = 1;
                                }
                            }
                        }
                    }
                }
            }
        } else {
            // any d: one row per warp at a time, 32 lanes stride the columns.
            for (int u = 0; u < 16; ++u) {
                const int64_t row = base + warp * 16 + u;
                if (row >= row_end) break;  // warp-uniform
                const float* xr = X + row * (int64_t)d;
                float acc[QT];
#pragma unroll
                for (int q = 0; q < QT; ++q) acc[q] = 0.f;
                for (int c = lane; c < d; c += 32) {
                    float xv = __ldg(xr + c);
#pragma unroll
                    for (int q = 0; q < QT; ++q) {
                        float a = s.qs[q * dq + c] - xv;
                        acc[q] = fmaf(a, a, acc[q]);
                    }
                }
#pragma unroll
                for (int q = 0; q < QT; ++q) {
                    float v = acc[q];
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                    if (lane == 0 && v <= tau_r[q]) {
                        u64 key = make_key(v, (uint32_t)row);
                        if (key < s.kth[q]) {
                            int pos = atomicAdd(&s.cnt[q], 1);
                            s.buf[q * P + pos] = key;
                            if (pos + 1 > P - SCAN_ROWS_PER_ITER) flag = 1;
                        }
                    }
                }
            }
        }

        if (__syncthreads_or(flag)) {
#pragma unroll 1
            for (int q = 0; q < QT; ++q)
                if (s.cnt[q] > k) compact(s.buf + q * P, P, &s.cnt[q], &s.kth[q], &s.tau[q], k);
        }
    }

    // ---- this CTA's k best per query -> global partial list -------------------------
    __syncthreads();
#pragma unroll 1
    for (int q = 0; q < QT; ++q) {
        compact(s.buf + q * P, P, &s.cnt[q], &s.kth[q], &s.tau[q], k);
        const int c = s.cnt[q];
        u64* out = partial + ((int64_t)blockIdx.x * QT + q) * k;
        for (int j = tid; j < k; j += SCAN_THREADS) out[j] = j < c ? s.buf[q * P + j] : KEY_EMPTY;
    }

    // ---- last CTA merges -------------------------------------------------------------
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        unsigned int t = atomicAdd(done, 1u);
        s_is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_is_last) return;
    __threadfence();

    const int G = gridDim.x;
    const int JH = (k + G - 1) / G;        // heads per list so that G * JH >= k
    const int M = G * JH;                  // <= 444 * 3 on a B200
    const int LPT = (G + SCAN_THREADS - 1) / SCAN_THREADS;              // lists per thread
    int KPT = (SCAN_FINAL_P / 2) / (SCAN_THREADS * LPT);                // keys per list per sweep
    KPT = KPT > 8 ? 8 : (KPT < 1 ? 1 : KPT);
#pragma unroll 1
    for (int q = 0; q < QT; ++q) {
        // Bound: every list is sorted, so the k-th smallest of the lists' first JH keys is an upper
        // bound of the global k-th key (k distinct keys lie at or below it) -- and a tight one: for
        // rows in random order its global rank is barely above k.
        u64 bound = KEY_EMPTY;
        if (M <= SCAN_FINAL_P) {
            for (int i = tid; i < M; i += SCAN_THREADS) {
                const int bb = i / JH, jj = i - bb * JH;
                s.buf[i] = jj < k ? ((const volatile u64*)partial)[((int64_t)bb * QT + q) * k + jj] : KEY_EMPTY;
            }
            int Ps = 2;
            while (Ps < M) Ps <<= 1;
            for (int i = M + tid; i < Ps; i += SCAN_THREADS) s.buf[i] = KEY_EMPTY;
            __syncthreads();
            block_bitonic_sort(s.buf, Ps);
            bound = s.buf[k - 1];       // KEY_EMPTY if fewer than k keys exist at all: take everything
            __syncthreads();
        }
        if (tid == 0) { s.cnt[0] = 0; s.kth[0] = KEY_EMPTY; s.tau[0] = FLT_MAX; }
        __syncthreads();

        // Sweep the lists front to back, KPT keys per list per sweep (independent loads: one L2 round
        // trip per sweep); a list is finished at its first key above the bound.  Random row order
        // ends after one sweep; k neighbours all inside one CTA's rows need k / KPT sweeps.
        bool live[4] = {true, true, true, true};
        for (int j0 = 0; j0 < k; j0 += KPT) {
            int flag = 0, any_live = 0;
            const u64 kth_now = s.kth[0];
            for (int l = 0; l < LPT && l < 4; ++l) {
                const int bb = tid + l * SCAN_THREADS;
                if (bb >= G || !live[l]) continue;
                const volatile u64* lst = partial + ((int64_t)bb * QT + q) * k;
                u64 keys[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) keys[j] = (j < KPT && j0 + j < k) ? lst[j0 + j] : KEY_EMPTY;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (j < KPT && keys[j] != KEY_EMPTY && keys[j] <= bound) {
                        if (keys[j] < kth_now) {
                            int pos = atomicAdd(&s.cnt[0], 1);
                            s.buf[pos] = keys[j];
                            if (pos + 1 > SCAN_FINAL_P / 2) flag = 1;
                        }
                    } else if (j < KPT) {
                        live[l] = false;
                    }
                }
                any_live |= live[l] ? 1 : 0;
            }
            // two block-wide votes per sweep: compaction needed? anything left to read?
            if (__syncthreads_or(flag)) compact(s.buf, SCAN_FINAL_P, &s.cnt[0], &s.kth[0], &s.tau[0], k);
            if (!__syncthreads_or(any_live)) break;
        }
        compact(s.buf, SCAN_FINAL_P, &s.cnt[0], &s.kth[0], &s.tau[0], k);
        const int c = s.cnt[0];
        for (int j = tid; j < k; j += SCAN_THREADS) {
            if (j < c) {
                u64 key = s.buf[j];
                D[(int64_t)q * k + j] = key_dist(key);
                I[(int64_t)q * k + j] = (int64_t)key_row(key);
            } else {
                D[(int64_t)q * k + j] = FLT_MAX;
                I[(int64_t)q * k + j] = -1;
            }
        }
        __syncthreads();
    }
    if (tid == 0) { done[0] = 0u; done[1] = 0u; }  // re-arm ticket and work counter for the next launch on this stream
}

__global__ void fill_empty_kernel(float* D, int64_t* I, int64_t count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) { D[i] = FLT_MAX; I[i] = -1; }
}

int launch_fill_empty(float* D, int64_t* I, int64_t count, cudaStream_t st) {
    if (count <= 0) return B2K_OK;
    fill_empty_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(D, I, count);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

int scan_pick_P(int k) { return next_pow2(k + SCAN_ROWS_PER_ITER); }

int scan_pick_qt(int64_t nq_left, int k) {
    int P = scan_pick_P(k);
    int qt = 8;
    while (qt > 1 && ((size_t)qt * P * 8 > 65536 || qt > nq_left)) qt >>= 1;
    return qt;
}

int scan_grid(int64_t n, int sm_count, int ctas_per_sm) {
    int64_t chunks = (n + SCAN_ROWS_PER_ITER - 1) / SCAN_ROWS_PER_ITER;
    int64_t g = (int64_t)sm_count * ctas_per_sm;
    if (g > chunks) g = chunks;
    if (g > 4 * SCAN_THREADS) g = 4 * SCAN_THREADS;  // the last-CTA merge tracks at most 4 lists per thread
    if (g < 1) g = 1;
    return (int)g;
}

size_t scan_partial_bytes(int grid, int qt, int k) { return (size_t)grid * qt * k * sizeof(u64); }

template <int QT, bool VEC, bool BF>
static int launch_scan_t(const void* X, const float* Q, int64_t n, int d, int k, int P, int grid,
                         u64* partial, unsigned int* done, float* D, int64_t* I, const unsigned int* run_if,
                         cudaStream_t st) {
    size_t smem = scan_smem_bytes(QT, d, P);
    if (smem > 48 * 1024)  // per-device attribute: set on every large launch (a handle may live on any GPU)
        B2K_CUDA(cudaFuncSetAttribute(scan_l2_topk_kernel<QT, VEC, BF>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // one wave: never launch more CTAs than are resident at once (register/smem limited per variant)
    int occ = 0, dev = 0, sms = 0;
    B2K_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, scan_l2_topk_kernel<QT, VEC, BF>, SCAN_THREADS, smem));
    B2K_CUDA(cudaGetDevice(&dev));
    B2K_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (occ > 0 && grid > occ * sms) grid = occ * sms;
    scan_l2_topk_kernel<QT, VEC, BF><<<grid, SCAN_THREADS, smem, st>>>(X, Q, n, d, k, P, partial, done, D, I, run_if);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

// One launch: QT (1,2,4,8) queries against rows [0,n).  Q, D, I already offset to the group.
// bf16: X holds bf16 rows (d % 8 == 0) and D receives the screening distances d~.
int launch_scan(const void* X, const float* Q, int64_t n, int d, int qt, int k, int grid,
                u64* partial, unsigned int* done, float* D, int64_t* I, cudaStream_t st, bool bf16,
                const unsigned int* run_if) {
    const int P = scan_pick_P(k);
    const bool vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
    if (scan_smem_bytes(qt, d, P) > 200 * 1024) {
        set_error("scan kernel: d=%d too large for shared memory", d);
        return B2K_ELIMIT;
    }
    if (bf16 && (d % 8 != 0 || !vec)) {
        set_error("scan kernel: bf16 rows need d %% 8 == 0 and 16-byte alignment");
        return B2K_EINVAL;
    }
#define B2K_SCAN_CASE(QT_)                                                                                        \
    case QT_:                                                                                                     \
        if (bf16) return launch_scan_t<QT_, true, true>(X, Q, n, d, k, P, grid, partial, done, D, I, run_if, st); \
        return vec ? launch_scan_t<QT_, true, false>(X, Q, n, d, k, P, grid, partial, done, D, I, run_if, st)     \
                   : launch_scan_t<QT_, false, false>(X, Q, n, d, k, P, grid, partial, done, D, I, run_if, st);
    switch (qt) {
        B2K_SCAN_CASE(1)
        B2K_SCAN_CASE(2)
        B2K_SCAN_CASE(4)
        B2K_SCAN_CASE(8)
    }
#undef B2K_SCAN_CASE
    set_error("scan kernel: bad query-group size %d", qt);
    return B2K_EINVAL;
}

// ---- re-rank of the bf16 scan's candidates ---------------------------------------------------
// One CTA per query: the kp candidates (ids Ia, screening distances Da, sorted by (d~, id)) ->
// exact fp32 distances (direct differences, one warp per row, fixed summation order: the
// arithmetic of the fp32 scan's rule (1)) -> the k best by (distance, id).  Coverage check first:
// every row that could be in the exact top-k has sqrt(d~) <= sqrt(T~) + 2E (file header), T~ the
// k-th candidate's d~; if the LAST candidate is not strictly beyond that zone, rows outside the
// candidate list could belong to the answer and *flag is raised (the caller's conditional fp32
// scan then overwrites D / I).  eta covers the fp32 rounding of the d~ sums themselves.
static constexpr int RERANK_THREADS = 1024;  // 32 warps: one query is one block, and a single query must not wait for 8 warps to walk 232 rows
__global__ void __launch_bounds__(RERANK_THREADS)
scan_rerank_kernel(const float* __restrict__ X, const float* __restrict__ Q, int d, int k, int kp,
                   const float* __restrict__ Da, const int64_t* __restrict__ Ia,
                   const unsigned int* __restrict__ max_err_bits, float* __restrict__ D, int64_t* __restrict__ I,
                   unsigned int* __restrict__ flag) {
    __shared__ u64 keys[1024];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t qi = blockIdx.x;
    const float* da = Da + qi * kp;
    const int64_t* ia = Ia + qi * kp;
    if (tid == 0 && ia[kp - 1] >= 0 && ia[k - 1] >= 0) {  // the list is full: rows exist beyond it
        const float E = sqrtf(__uint_as_float(*max_err_bits)) * 1.001f;
        const float eta = ((float)d * 0.125f + 8.f) * 5.97e-8f;
        const float zr = (sqrtf(da[k - 1]) * (1.f + 2.f * eta) + 2.f * E) * (1.f + 2.f * eta);
        const float zone = zr * zr * (1.f + 1e-6f);
        if (!(da[kp - 1] > zone)) atomicExch(flag, 1u);
    }
    int Ps = 2;
    while (Ps < kp) Ps <<= 1;
    for (int c = tid; c < Ps; c += RERANK_THREADS) keys[c] = KEY_EMPTY;
    __syncthreads();
    const float* qv = Q + qi * (int64_t)d;
    for (int c = warp; c < kp; c += RERANK_THREADS / 32) {
        const int64_t id = ia[c];
        if (id < 0) continue;  // warp-uniform
        const float* xv = X + id * (int64_t)d;
        float acc = 0.f;
        // up to four 512-byte slices of the row are requested before the first is used: the row is a random
        // HBM read, and one round trip per 512 columns beats one per 128 (same summation order per lane)
        for (int t0 = lane * 4; t0 < d; t0 += 512) {
            float4 a[4], b[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int t = t0 + j * 128;
                if (t < d) {
                    a[j] = __ldg(reinterpret_cast<const float4*>(qv + t));
                    b[j] = ldg_stream(reinterpret_cast<const float4*>(xv + t));
                } else {
                    a[j] = b[j] = make_float4(0.f, 0.f, 0.f, 0.f);  // adds exact zeros
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float e0 = a[j].x - b[j].x, e1 = a[j].y - b[j].y, e2 = a[j].z - b[j].z, e3 = a[j].w - b[j].w;
                acc = fmaf(e0, e0, acc); acc = fmaf(e1, e1, acc); acc = fmaf(e2, e2, acc); acc = fmaf(e3, e3, acc);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0 && acc < FLT_MAX) keys[c] = make_key(acc, (uint32_t)id);
    }
    __syncthreads();
    block_bitonic_sort(keys, Ps);
    for (int j = tid; j < k; j += RERANK_THREADS) {
        const u64 key = j < Ps ? keys[j] : KEY_EMPTY;
        if (key != KEY_EMPTY) {
            D[qi * k + j] = key_dist(key);
            I[qi * k + j] = (int64_t)key_row(key);
        } else {
            D[qi * k + j] = FLT_MAX;
            I[qi * k + j] = -1;
        }
    }
}

int launch_scan_rerank(const float* X, const float* Q, int64_t nq, int d, int k, int kp, const float* Da,
                       const int64_t* Ia, const unsigned int* max_err_bits, float* D, int64_t* I,
                       unsigned int* flag, cudaStream_t st) {
    if (nq <= 0) return B2K_OK;
    if (kp > 1024 || kp < k || d % 4 != 0) { set_error("scan re-rank: bad candidate count %d for k=%d", kp, k); return B2K_EINVAL; }
    scan_rerank_kernel<<<(unsigned)nq, RERANK_THREADS, 0, st>>>(X, Q, d, k, kp, Da, Ia, max_err_bits, D, I, flag);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    return B2K_OK;
}

}  // namespace b2k
