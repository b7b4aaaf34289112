// k3_tc.cu -- K3: batched exact search on the 5th-generation tensor cores, with the top-k
// selection fused into the accumulator epilogue; K4, the cross-CTA merge / bound refresh; K6,
// the exact fp32 re-rank behind the screening modes.
//
// Stands in for faiss `exhaustive_L2sqr_blas` (sgemm in 4096 x 1024 blocks, then
// ||q||^2 + ||x||^2 - 2<q,x>, clamp at 0) + heap/reservoir result handlers behind
// `index.search(xq, k)` (/root/reference/llm/rag.py:116) -- oracle rule (2).
//
// Numerics (template parameter MODE; chosen per search by csrc/index.cu):
//   3  one bf16 product (kind::f16) on bf16 copies of rows and queries       } screens: the lists keep
//   2  one TF32 product (kind::tf32 reads fp32 and ignores 13 mantissa bits) } every key within a
//   4  3xTF32 (as below) used as a screen                                    } rigorous error slack of
//      the k-th approximate distance, and K6 computes the exact fp32 distances (direct differences)
//      of those few hundred rows per query: the result is exact.  3 is the default, 2 serves when
//      there are no bf16 copies, 4 when the slack zone of 3 / 2 overflows its lists.
//   0 / 1  3xTF32: per 8 columns of K three products Q_lo*X_hi + Q_hi*X_lo + Q_hi*X_hi (hi = the fp32
//      value as the tensor core reads it, lo = the exact fp32 remainder), in one accumulator (0) or
//      with the two small terms in their own accumulator (1, long rows): faiss-like expansion
//      numerics, distances straight from the accumulator (option tensor_mode = 1).
//
// What runs where (per CTA; CG = 1, or CG = 2 for a CTA pair issuing cta_group::2 MMAs):
//   warp 0  scheduler + TMA producer: takes units from a global atomic counter (tile-major, so the
//           CTAs running at any moment share a few database tiles through L2 and every tile comes
//           from HBM once per batch), publishes them through a 4-deep shared-memory ring, and
//           issues cp.async.bulk.tensor 2-D boxes (one 128-byte swizzle row wide) of the 128 query
//           rows and the 256/CG database rows of the unit into an NS-stage ring (mbarrier tx).
//   warp 1  MMA issuer: one elected thread issues tcgen05.mma (M = 128*CG, N = 256), accumulators
//           in TMEM; two 256-column accumulators (all 512 TMEM columns), so the MMAs of unit u+1
//           overlap the epilogue of unit u; tcgen05.commit frees smem stages / publishes accumulators.
//   warps 2..5 (NH = 1) or 2..9 (NH = 2)  epilogue: tcgen05.ld 32 lanes x 32 columns at a time, two
//           register buffers.  M = queries, so a thread owns one query row (NH = 2: one column half
//           of it) and its pruning threshold lives in a register:
//                 s = ||x||^2 - 2*acc         (one FFMA; ||x||^2 staged in shared memory)
//                 any s <= tau - ||q||^2 ?    (half a 3-input FMNMX per element)
//           a thread that has a survivor in its 32 columns locates it (bit mask + register select
//           tree) and appends a 64-bit (distance,row) key to that query's private list in global
//           memory (L2-resident); a list that could overflow on the next tile is compacted by its
//           warp with a bitonic sort in shared memory, which also lowers the threshold and
//           publishes it (atomicMin) to the other CTAs.
// No distance matrix is ever written: per unit the only global traffic is the operand stream and
// the rare survivor keys.  The host (k3_search) runs one small warm-start launch, then the main
// pass in a few segments with a K4 bound refresh in between, then K4 / K6.
// Roofline: tensor; executed flops = (1 or 3) * 2 * nq * N * d.
#include <cuda.h>

#include "k3_api.h"
#include "kernels.h"
#include "sm100.cuh"

namespace b2k {

static constexpr int K3_BM = 128;       // query rows per CTA = TMEM lanes
static constexpr int K3_BN = 256;       // database rows per tile = TMEM columns per accumulator
static constexpr int K3_BK = 32;        // fp32 per k-chunk: one 128-byte swizzle row
// threads per CTA: warp 0 TMA, warp 1 MMA, then NH groups of four epilogue warps (NH = 1: warps
// 2..5 take all 256 columns of their 32 query rows; NH = 2: warps 2..5 take columns 0..127 and
// warps 6..9 columns 128..255 of the same rows, two warps per scheduler to hide each other's latencies)
static constexpr int k3_threads(int nh) { return 64 + 128 * nh; }
static constexpr int K3_A_BYTES = K3_BM * K3_BK * 4;  // 16 KB

struct K3Params {
    int64_t n;        // database rows
    int d;
    int64_t nq;
    int k;
    int P;            // per-(CTA,query) list capacity (power of two, >= k + 256)
    int nqb;          // query blocks of 128*CG
    int64_t nunits;   // ntiles * nqb
    int qpc;          // list slots per CTA = nqb * 128
    const float* norms;
    const float* qn;
    unsigned int* gts;  // [nq]   shared upper bound of the k-th distance (float bits)
    int* cnt;           // [ncta][qpc]
    float* tau;         // [ncta][qpc]
    u64* kth;           // [ncta][qpc]
    u64* cand;          // [ncta][qpc][P]
    u64* next_unit;     // dynamic scheduler: next unit to hand out
    const float* slack2;     // [nq] screening: 2 * (bound on |approximate - exact distance|) per query
    unsigned int* overflow;  // screening: set when a list cannot hold every key inside the slack
    int static_mode;    // 1: warm-start pass, CTA (pair) c takes tiles static_tile0 + c*static_tpc + [0,static_tpc), all query blocks
    int static_tpc;
    int static_qb_mod;  // warm-start pass: CTA (pair) c only takes the query blocks qb with qb % static_qb_mod == c % static_qb_mod
    int tile_cols;      // columns (rows of the tile) the epilogue looks at: 256, or fewer in the first warm-start pass
    int64_t static_tile0;
    int probe;          // 1: dense-rate probe -- every query row is treated as padding, so the filter never fires
    // query-stationary schedule (QS kernels): this launch covers tiles [qs_t0, qs_t1); work items are
    // (run of T consecutive tiles, query block), handed out run-major; runs are qs_T1 tiles long up to
    // tile qs_tA and qs_T2 (shorter: the tail balances better) from there on
    long long qs_t0, qs_tA, qs_t1;
    long long qs_itemsA;  // items of the first region = ceil((qs_tA - qs_t0) / qs_T1) * nqb
    int qs_T1, qs_T2;
    const unsigned int* outl;  // norm-folded screen: [0] number of outlier rows kept out of the screen, [8..) their ids (else null)
    int pf_tiles;         // > 0: the CTA (pair) that takes query block 0 of tile t asks L2 for the rows of tile t + pf_tiles
};

// A unit as it travels through the ring: (tile << 8) | query block, plus, for the query-stationary
// kernels, whether it is the first / last tile of its work item (the A operand is loaded with the
// first and released with the last).
static constexpr long long K3_U_FIRST = 1ll << 44, K3_U_LAST = 1ll << 45;
__device__ __forceinline__ int64_t unit_tile(long long u) { return (u >> 8) & 0xFFFFFFFFll; }
static constexpr int K3_QS_NA = 8;  // most k-chunks of a resident query block (d <= 512 in bf16, <= 256 in TF32)

static constexpr int K3_NSCHED = 4;  // depth of the unit-id ring between the scheduler and the other roles

// Shared memory: [QS: the resident query block, nkc chunks of 16 KB][NS operand stages][tail].
// Stage = [Q_hi][Q_lo?][X_hi][X_lo?], or just [X] for the query-stationary kernels.  The tail holds
// the tile norms, barriers, unit ring, TMEM slot, list lengths and the compaction scratch; its
// offsets are relative because the part before it depends on d at run time.
template <int CG, int NS, int MODE, bool QS>
struct K3Smem {
    static constexpr int B_ROWS = K3_BN / CG;
    static constexpr int B_BYTES = B_ROWS * K3_BK * 4;
    static constexpr int NPART = (MODE == 2 || MODE == 3) ? 1 : 2;  // one operand pair (one-product screening), or hi + lo (3xTF32)
    static constexpr int OFF_B = QS ? 0 : NPART * K3_A_BYTES;
    static constexpr int STAGE_BYTES = QS ? B_BYTES : NPART * (K3_A_BYTES + B_BYTES);
    // norm fold (bf16 screening with CTA pairs): two buffers of 32-byte norm records for the rows of a tile
    // and one block of "ones" rows for the queries, operands of one extra K = 16 MMA per tile
    static constexpr bool NF = (MODE == 3 && CG == 2);
    static constexpr int NREC_BYTES = B_ROWS * 32;
    static constexpr int ONES_BYTES = K3_BM * 32;
    static constexpr int REL_NREC = 0;
    static constexpr int REL_ONES = 2 * NREC_BYTES;
    static constexpr int REL_XN = NF ? 2 * NREC_BYTES + ONES_BYTES : 0;  // float[2][256] (unused with NF)
    static constexpr int REL_BAR = REL_XN + 2 * K3_BN * 4;      // full[NS] empty[NS] tfull[2] tempty[2] sfull[4] sempty[4] (afull[8] aempty[8]) (nfull[2] nempty[2] onefull)
    static constexpr int NBAR = 2 * NS + 4 + 2 * K3_NSCHED + (QS ? 2 * K3_QS_NA : 0) + (NF ? 5 : 0);
    static constexpr int REL_SCHED = REL_BAR + NBAR * 8;        // long long[K3_NSCHED]
    static constexpr int REL_TMEM = REL_SCHED + K3_NSCHED * 8;
    static constexpr int REL_CNT = (REL_TMEM + 16 + 15) & ~15;  // int[128]: list lengths of the unit in flight (NH = 2)
    static constexpr int REL_SCRATCH = REL_CNT + K3_BM * 4;     // 4 warps x P keys
    static size_t bytes(int P, int nkc) {
        return (size_t)(QS ? nkc * K3_A_BYTES : 0) + (size_t)NS * STAGE_BYTES + REL_SCRATCH + (size_t)4 * P * 8;
    }
};

// Consumer side of the unit-id ring: wait for slot `ss`, read the unit, release the slot to the
// scheduler (the leader CTA's producer thread).  Called by whole warps (warp-uniform result).
template <int CG>
__device__ __forceinline__ long long sched_take_warp(uint64_t* sfull, uint64_t* sempty, const long long* sched_u,
                                                     int ss, uint32_t sph, uint32_t rank, int lane) {
    ptx::mbar_wait(&sfull[ss], sph);  // the peer CTA's copy arrives by st.async on its own barrier: no cluster-scope acquire
    const long long u = *reinterpret_cast<const volatile long long*>(&sched_u[ss]);
    __syncwarp();
    // (u >= -1 always holds: the test makes the relaxed arrive wait for the value just read)
    if (lane == 0 && u >= -1) {
        if (CG == 1 || rank == 0) ptx::mbar_arrive(&sempty[ss]);
        else ptx::mbar_arrive_cluster(&sempty[ss], 0);
    }
    return u;
}

// Compiler fence on 32 registers: placed right after the NEXT chunk's tcgen05.ld has been issued, it
// pins every use of `v` below that point, so the filter of this chunk cannot be scheduled above
// the prefetch (nvcc otherwise sinks the prefetch under the filter and each chunk pays the full
// TMEM round trip, ~1.8k cycles measured).
__device__ __forceinline__ void reg_fence32(uint32_t (&v)[32]) {
    asm volatile("" : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                      "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]),
                      "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]),
                      "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
                 :: "memory");
}

// v[c] for a run-time c, without local memory: a binary tree of selects over SZ registers.
template <int LO, int N, int SZ>
__device__ __forceinline__ float reg_select(const float (&v)[SZ], int c) {
    if constexpr (N == 1) {
        return v[LO];
    } else {
        const float a = reg_select<LO, N / 2, SZ>(v, c), b = reg_select<LO + N / 2, N / 2, SZ>(v, c);
        return (c & (N / 2)) ? b : a;
    }
}

// One thread, 32 consecutive accumulator columns of its query row.  The dense part is two
// instructions per element or fewer: s = ||x||^2 - 2 ip (one FFMA) and "any s <= t" over four
// independent chains (nvcc turns the OR of compares into 3-input FMNMX: 1.5 per element).  With
// the bound refreshed between segments of the pass a warp meets a survivor in about one chunk
// out of ten; only then are the columns located (second, mask-building walk over the chains that fired).
// v = main accumulator, w = cross-term accumulator (SEP only).
// SHCNT: the list length lives in shared memory (*cnt_s) and is bumped atomically, because two
// threads (the two column halves of the row) append to the same list.
template <bool SEP, bool SHCNT>
__device__ __forceinline__ void k3_filter_chunk(const uint32_t (&v)[32], const uint32_t (&w)[32], const float* xn,
                                                float t_cmp, float qn, float tau_eff, u64 kth, uint32_t row0,
                                                u64* list, int& cnt, int* cnt_s) {
    bool any[4] = {false, false, false, false};
#pragma unroll
    for (int c = 0; c < 32; c += 4) {
        const float4 x4 = *reinterpret_cast<const float4*>(xn + c);
        const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float ip = SEP ? __uint_as_float(v[c + j]) + __uint_as_float(w[c + j]) : __uint_as_float(v[c + j]);
            any[j] |= fmaf(-2.f, ip, xs[j]) <= t_cmp;
        }
    }
    if (any[0] | any[1] | any[2] | any[3]) {
        // which columns: only the chains that reported a survivor are looked at again (chain j = the 8
        // columns c with c % 4 == j): an 8-bit mask built without branches, then a loop over its set
        // bits (usually one), the value picked out of the 8 registers by a 3-level select tree
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (any[j]) {
                uint32_t mask = 0;
                float ip[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int c = 4 * i + j;
                    ip[i] = SEP ? __uint_as_float(v[c]) + __uint_as_float(w[c]) : __uint_as_float(v[c]);
                    mask |= (fmaf(-2.f, ip[i], xn[c]) <= t_cmp) ? (1u << i) : 0u;
                }
                while (mask) {
                    const int i = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const int c = 4 * i + j;
                    const float ipc = reg_select<0, 8, 8>(ip, i);
                    const float xv = xn[c];
                    float dist = (qn + xv) - 2.f * ipc;  // oracle rule (2), then clamp
                    dist = dist < 0.f ? 0.f : dist;
                    if (dist <= tau_eff) {
                        const u64 key = make_key(dist, row0 + c);
                        if (key < kth) {
                            if constexpr (SHCNT) { list[atomicAdd(cnt_s, 1)] = key; }
                            else { list[cnt] = key; ++cnt; }
                        }
                    }
                }
            }
        }
    }
}

// The same for the norm-folded accumulator a = q~.x~ - ||x||^2/2 (bf16 screening with CTA pairs): a row passes
// iff a >= thr, so the dense part is a running maximum (3-input FMNMX: half an instruction per element, no norm
// loads); the approximate distance of a survivor is ||q||^2 - 2a.  ncols: columns of this chunk that are real rows.
template <bool SHCNT>
__device__ __forceinline__ void k3_filter_chunk_nf(const uint32_t (&v)[32], float thr, float qn, float tau_eff, u64 kth,
                                                   uint32_t row0, int ncols, u64* list, int& cnt, int* cnt_s) {
    float m[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) m[j] = __uint_as_float(v[j]);
#pragma unroll
    for (int c = 4; c < 32; c += 8) {
#pragma unroll
        for (int j = 0; j < 4; ++j) m[j] = fmaxf(fmaxf(m[j], __uint_as_float(v[c + j])), __uint_as_float(v[c + 4 + j]));
    }
    if (fmaxf(fmaxf(m[0], m[1]), fmaxf(m[2], m[3])) >= thr) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (m[j] >= thr) {
                uint32_t mask = 0;
                float a[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    a[i] = __uint_as_float(v[4 * i + j]);
                    mask |= (a[i] >= thr) ? (1u << i) : 0u;
                }
                while (mask) {
                    const int i = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const int c = 4 * i + j;
                    if (c < ncols) {
                        float dist = fmaf(-2.f, reg_select<0, 8, 8>(a, i), qn);
                        dist = dist < 0.f ? 0.f : dist;
                        if (dist <= tau_eff) {
                            const u64 key = make_key(dist, row0 + c);
                            if (key < kth) {
                                if constexpr (SHCNT) { list[atomicAdd(cnt_s, 1)] = key; }
                                else { list[cnt] = key; ++cnt; }
                            }
                        }
                    }
                }
            }
        }
    }
}

// SEP = true keeps the two small cross products (Q_lo*X_hi + Q_hi*X_lo) in their own TMEM
// accumulator (columns 256..511) and adds it to the main one in the epilogue.  The tensor core
// truncates on every accumulate (measured ~0.8 ulp of the accumulator per MMA, biased towards
// zero), so adding the 2^-11-sized terms into the large accumulator triples that error for
// nothing; with SEP the large accumulator sees d/8 accumulations instead of 3d/8.  It costs the
// double buffering (the MMAs of the next tile wait for the epilogue), so it is used for long
// rows (d >= 512) where the relative cost is a few percent and the error would otherwise
// exceed 1e-5 of ||q||^2 + ||x||^2 on near-duplicate unit-norm vectors.
template <int CG, int NS, int MODE, int NH, bool QS>
__global__ void __launch_bounds__(k3_threads(NH), 1)
k3_tf32x3_topk_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_ql,
                      const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_xl,
                      const __grid_constant__ CUtensorMap tm_n, const __grid_constant__ CUtensorMap tm_one,
                      const K3Params p) {
    // MODE: 0 3xTF32, one accumulator | 1 3xTF32, separate cross-term accumulator | 2 one TF32 product, screening
    //       3 one bf16 product, screening | 4 3xTF32 (as 1) used as a screen: second line of defence when the
    //       slack zone of 2 / 3 overflows (its error is ~2^-21, so its slack zone is ~10^4 times narrower)
    constexpr bool SEP = (MODE == 1 || MODE == 4);  // separate accumulator for the cross terms
    constexpr bool ONEP = (MODE == 2 || MODE == 3);  // one product per k-step: no residual operands
    constexpr bool SCREEN = (ONEP || MODE == 4);     // keep everything within the error slack; exact distances come from the re-rank kernel
    constexpr bool BF16 = (MODE == 3);                  // the screening product runs on bf16 copies of the rows (kind::f16)
    constexpr int BKE = BF16 ? 64 : 32;                 // elements per k-chunk: one 128-byte swizzle row either way
    static_assert(!QS || ONEP, "the query-stationary schedule is for the one-product modes");
    using L = K3Smem<CG, NS, MODE, QS>;
    constexpr bool NF = L::NF;  // the tensor core also adds -||x||^2/2 (one more MMA per tile): the epilogue only compares
    // No integer round-trip on this pointer: the compiler must keep seeing shared memory here, or
    // every access through it becomes a generic LD/ST (that cost 2k cycles per 32 accumulator
    // columns in the epilogue).  The kernel has no static shared memory, so the dynamic window
    // starts 1024-aligned (SWIZZLE_128B needs that); checked below.
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = smem_raw;
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = (CG == 2) ? ptx::cluster_ctarank() : 0u;
    const int nkc = (p.d + BKE - 1) / BKE;
    unsigned char* stages = smem + (QS ? nkc * K3_A_BYTES : 0);  // QS: the query block's nkc chunks stay in front
    unsigned char* tail = stages + NS * L::STAGE_BYTES;
    float* xn_s = reinterpret_cast<float*>(tail + L::REL_XN);
    uint64_t* full = reinterpret_cast<uint64_t*>(tail + L::REL_BAR);
    uint64_t* empty = full + NS;
    uint64_t* tfull = empty + NS;
    uint64_t* tempty = tfull + 2;
    uint64_t* sfull = tempty + 2;
    uint64_t* sempty = sfull + K3_NSCHED;
    uint64_t* afull = sempty + K3_NSCHED;   // QS only: chunk kc of the resident query block has landed
    uint64_t* aempty = afull + K3_QS_NA;    //          ... has been read by the last MMA of its work item
    uint64_t* nfull = sempty + K3_NSCHED + (QS ? 2 * K3_QS_NA : 0);  // NF only: the tile's norm records have landed
    uint64_t* nempty = nfull + 2;           //          ... have been read by the tile's norm MMA
    uint64_t* onefull = nempty + 2;         //          the block of ones has landed (once per kernel)
    unsigned char* nrec = tail + L::REL_NREC;
    unsigned char* ones = tail + L::REL_ONES;
    long long* sched_u = reinterpret_cast<long long*>(tail + L::REL_SCHED);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + L::REL_TMEM);
    u64* scratch = reinterpret_cast<u64*>(tail + L::REL_SCRATCH);
    int* cnt_s = reinterpret_cast<int*>(tail + L::REL_CNT);

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tm_q); ptx::prefetch_tmap(&tm_ql); ptx::prefetch_tmap(&tm_x); ptx::prefetch_tmap(&tm_xl);
        for (int s = 0; s < NS; ++s) { ptx::mbar_init(&full[s], CG); ptx::mbar_init(&empty[s], 1); }
        for (int s = 0; s < 2; ++s) { ptx::mbar_init(&tfull[s], 1); ptx::mbar_init(&tempty[s], 4 * NH * CG); }
        for (int s = 0; s < K3_NSCHED; ++s) { ptx::mbar_init(&sfull[s], 1); ptx::mbar_init(&sempty[s], (1 + 4 * NH) * CG); }
        if (QS) for (int s = 0; s < K3_QS_NA; ++s) { ptx::mbar_init(&afull[s], CG); ptx::mbar_init(&aempty[s], 1); }
        if (NF) {
            for (int s = 0; s < 2; ++s) { ptx::mbar_init(&nfull[s], CG); ptx::mbar_init(&nempty[s], 1); }
            ptx::mbar_init(onefull, CG);
            ptx::prefetch_tmap(&tm_n); ptx::prefetch_tmap(&tm_one);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) {
        ptx::tmem_alloc<CG>(tmem_slot, 512);
        ptx::tmem_relinquish<CG>();
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (CG == 2) ptx::cluster_sync_all();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== scheduler + TMA producer =====================
        if (lane == 0) {
            int stage = 0, ss = 0, static_i = 0;
            uint32_t phase = 0, sph = 0, aph = 0;
            long long item_tile = 0, item_end = 0;  // QS: tiles left of the current work item
            int item_qb = 0;
            int nb = 0;
            uint32_t nph = 0;
            if constexpr (NF) {
                if (rank == 0) ptx::mbar_arrive_expect_tx(onefull, 2 * L::ONES_BYTES);
                else ptx::mbar_arrive_cluster(onefull, 0);
                ptx::tma_load_2d_pair(ones, &tm_one, onefull, 0, 0);
            }
            for (;;) {
                long long u;
                if (rank == 0) {
                    // dynamic scheduling: units are handed out in order, so the CTAs running at any
                    // moment work on neighbouring database tiles and share them through L2
                    ptx::mbar_wait(&sempty[ss], sph ^ 1);
                    if constexpr (QS) {
                        // work item = (run of consecutive tiles, query block), run-major: the pairs working at
                        // any moment hold different query blocks but walk the same few runs of tiles
                        bool first = false;
                        if (item_tile >= item_end) {
                            const unsigned long long i = atomicAdd(p.next_unit, 1ull);
                            first = true;
                            if ((long long)i < p.qs_itemsA) {
                                const long long run = (long long)(i / (unsigned)p.nqb);
                                item_qb = (int)(i - (unsigned long long)run * (unsigned)p.nqb);
                                item_tile = p.qs_t0 + run * p.qs_T1;
                                item_end = item_tile + p.qs_T1 < p.qs_tA ? item_tile + p.qs_T1 : p.qs_tA;
                            } else {
                                const unsigned long long j = i - (unsigned long long)p.qs_itemsA;
                                const long long run = (long long)(j / (unsigned)p.nqb);
                                item_qb = (int)(j - (unsigned long long)run * (unsigned)p.nqb);
                                item_tile = p.qs_tA + run * p.qs_T2;
                                item_end = item_tile + p.qs_T2 < p.qs_t1 ? item_tile + p.qs_T2 : p.qs_t1;
                            }
                        }
                        if (item_tile >= item_end) {
                            u = -1;
                        } else {
                            u = (item_tile << 8) | (long long)item_qb | (first ? K3_U_FIRST : 0);
                            ++item_tile;
                            if (item_tile == item_end) u |= K3_U_LAST;
                        }
                    } else {
                    if (p.static_mode) {
                        const int c = blockIdx.x / CG, G = p.static_qb_mod;
                        const int mine = (p.nqb - (c % G) + G - 1) / G;  // query blocks of this CTA (pair)
                        if (static_i < mine * p.static_tpc) {
                            const int j = static_i / mine, qb = (c % G) + (static_i - j * mine) * G;
                            u = (p.static_tile0 + (long long)c * p.static_tpc + j) * p.nqb + qb;
                        } else {
                            u = -1;
                        }
                        ++static_i;
                    } else {
                        u = (long long)atomicAdd(p.next_unit, 1ull);
                    }
                    // consumers get the unit already split into (tile, query block): a 64-bit division
                    // per unit in each epilogue warp was ~350 cycles of dependent instructions, four times
                    if (u < 0 || u >= p.nunits) {
                        u = -1;
                    } else {
                        const uint32_t tt = (uint32_t)u / (uint32_t)p.nqb;  // nunits < 2^31 (2^24 tiles x 32 query blocks)
                        u = ((long long)tt << 8) | (long long)((uint32_t)u - tt * (uint32_t)p.nqb);
                    }
                    }
                    sched_u[ss] = u;
                    ptx::mbar_arrive(&sfull[ss]);
                    if (CG == 2) ptx::st_async_b64_cluster(&sched_u[ss], &sfull[ss], 1, (uint64_t)u);
                } else {
                    ptx::mbar_wait(&sfull[ss], sph);
                    u = *reinterpret_cast<const volatile long long*>(&sched_u[ss]);
                    if (u >= -1) ptx::mbar_arrive_cluster(&sempty[ss], 0);
                }
                if (++ss == K3_NSCHED) { ss = 0; sph ^= 1; }
                if (u < 0) break;
                const int64_t t = unit_tile(u);
                const int qb = (int)(u & 255);
                const bool load_a = QS && (u & K3_U_FIRST) != 0;
                const int qrow0 = qb * (K3_BM * CG) + (int)rank * K3_BM;
                const int64_t xrow0 = t * K3_BN + (int64_t)rank * L::B_ROWS;
                if (p.pf_tiles > 0 && qb == 0 && !p.static_mode) {
                    // Every row tile is read by nqb CTA pairs at about the same time, and the first of them waits for
                    // HBM.  One of them asks L2 for a tile a few dozen units ahead, so that by the time the scheduler
                    // hands that tile out its rows are an L2 hit for everybody.
                    const int64_t tp = t + p.pf_tiles;
                    if (tp * K3_BN < p.n) {
                        const int prow0 = (int)(tp * K3_BN + (int64_t)rank * L::B_ROWS);
                        for (int kc = 0; kc < nkc; ++kc) ptx::tma_prefetch_l2_2d(&tm_x, kc * BKE, prow0);
                        if (NF) ptx::tma_prefetch_l2_2d(&tm_n, 0, prow0);
                    }
                }
                if constexpr (NF) {
                    // this tile's norm records, into the buffer that goes with the tile's accumulator
                    ptx::mbar_wait(&nempty[nb], nph ^ 1);
                    if (rank == 0) ptx::mbar_arrive_expect_tx(&nfull[nb], 2 * L::NREC_BYTES);
                    else ptx::mbar_arrive_cluster(&nfull[nb], 0);
                    ptx::tma_load_2d_pair(nrec + nb * L::NREC_BYTES, &tm_n, &nfull[nb], 0, (int)xrow0);
                    nb ^= 1;
                    if (nb == 0) nph ^= 1;
                }
                for (int kc = 0; kc < nkc; ++kc) {
                    if (load_a) {
                        // chunk kc of the new query block replaces the old one as soon as the last MMA that read it has retired
                        ptx::mbar_wait(&aempty[kc], aph ^ 1);
                        unsigned char* sa = smem + kc * K3_A_BYTES;
                        if (CG == 1) {
                            ptx::mbar_arrive_expect_tx(&afull[kc], K3_A_BYTES);
                            ptx::tma_load_2d(sa, &tm_q, &afull[kc], kc * BKE, qrow0);
                        } else {
                            if (rank == 0) ptx::mbar_arrive_expect_tx(&afull[kc], 2 * K3_A_BYTES);
                            else ptx::mbar_arrive_cluster(&afull[kc], 0);
                            ptx::tma_load_2d_pair(sa, &tm_q, &afull[kc], kc * BKE, qrow0);
                        }
                    }
                    ptx::mbar_wait(&empty[stage], phase ^ 1);
                    unsigned char* sb = stages + stage * L::STAGE_BYTES;
                    if (CG == 1) {
                        ptx::mbar_arrive_expect_tx(&full[stage], L::STAGE_BYTES);
                        if (!QS) ptx::tma_load_2d(sb, &tm_q, &full[stage], kc * BKE, qrow0);
                        if (!ONEP) ptx::tma_load_2d(sb + K3_A_BYTES, &tm_ql, &full[stage], kc * BKE, qrow0);
                        ptx::tma_load_2d(sb + L::OFF_B, &tm_x, &full[stage], kc * BKE, (int)xrow0);
                        if (!ONEP) ptx::tma_load_2d(sb + L::OFF_B + L::B_BYTES, &tm_xl, &full[stage], kc * BKE, (int)xrow0);
                    } else {
                        if (rank == 0) ptx::mbar_arrive_expect_tx(&full[stage], 2 * L::STAGE_BYTES);
                        else ptx::mbar_arrive_cluster(&full[stage], 0);
                        if (!QS) ptx::tma_load_2d_pair(sb, &tm_q, &full[stage], kc * BKE, qrow0);
                        if (!ONEP) ptx::tma_load_2d_pair(sb + K3_A_BYTES, &tm_ql, &full[stage], kc * BKE, qrow0);
                        ptx::tma_load_2d_pair(sb + L::OFF_B, &tm_x, &full[stage], kc * BKE, (int)xrow0);
                        if (!ONEP) ptx::tma_load_2d_pair(sb + L::OFF_B + L::B_BYTES, &tm_xl, &full[stage], kc * BKE, (int)xrow0);
                    }
                    if (++stage == NS) { stage = 0; phase ^= 1; }
                }
                if (load_a) aph ^= 1;
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        if (rank == 0) {
            constexpr uint32_t idesc = BF16 ? ptx::umma_idesc_bf16(K3_BM * CG, K3_BN) : ptx::umma_idesc_tf32(K3_BM * CG, K3_BN);
            int stage = 0, as = 0, ss = 0;
            uint32_t phase = 0, aphase = 0, sph = 0, afph = 0;
            if constexpr (NF) ptx::mbar_wait(onefull, 0);
            for (;;) {
                const long long u = sched_take_warp<CG>(sfull, sempty, sched_u, ss, sph, rank, lane);
                if (++ss == K3_NSCHED) { ss = 0; sph ^= 1; }
                if (u < 0) break;
                const bool a_first = QS && (u & K3_U_FIRST) != 0, a_last = QS && (u & K3_U_LAST) != 0;
                ptx::mbar_wait(&tempty[as], aphase ^ 1);  // epilogue has drained this accumulator
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(as * K3_BN);
                const uint32_t d_corr = SEP ? tmem_base + (uint32_t)K3_BN : d_tmem;
                for (int kc = 0; kc < nkc; ++kc) {
                    if (a_first) ptx::mbar_wait(&afull[kc], afph);  // the new query block's chunk kc is in place
                    ptx::mbar_wait(&full[stage], phase);
                    ptx::tc_fence_after();
                    if (lane == 0) {
                        const uint32_t sb = ptx::smem_u32(stages + stage * L::STAGE_BYTES);
                        const uint64_t a_hi = ptx::umma_desc_k_sw128(QS ? ptx::smem_u32(smem + kc * K3_A_BYTES) : sb);
                        const uint64_t a_lo = ptx::umma_desc_k_sw128(sb + K3_A_BYTES);
                        const uint64_t b_hi = ptx::umma_desc_k_sw128(sb + L::OFF_B);
                        const uint64_t b_lo = ptx::umma_desc_k_sw128(sb + L::OFF_B + L::B_BYTES);
#pragma unroll
                        for (int ks = 0; ks < K3_BK / 8; ++ks) {
                            const uint64_t off = (uint64_t)(ks * 32 >> 4);  // 8 fp32 = 32 bytes along K
                            if constexpr (BF16) {
                                ptx::mma_bf16<CG>(d_tmem, a_hi + off, b_hi + off, idesc, (kc | ks) != 0);  // 16 bf16 = 32 bytes along K
                            } else if constexpr (ONEP) {
                                ptx::mma_tf32<CG>(d_tmem, a_hi + off, b_hi + off, idesc, (kc | ks) != 0);
                            } else {
                                ptx::mma_tf32<CG>(d_corr, a_lo + off, b_hi + off, idesc, (kc | ks) != 0);  // small terms first
                                ptx::mma_tf32<CG>(d_corr, a_hi + off, b_lo + off, idesc, 1u);
                                ptx::mma_tf32<CG>(d_tmem, a_hi + off, b_hi + off, idesc, SEP ? (uint32_t)((kc | ks) != 0) : 1u);
                            }
                        }
                        ptx::mma_commit<CG>(&empty[stage]);            // smem slot free once these MMAs retire
                        if (a_last) ptx::mma_commit<CG>(&aempty[kc]);  // ... and this chunk of the query block, for the next work item
                        if (!NF && kc == nkc - 1) ptx::mma_commit<CG>(&tfull[as]);  // accumulator complete
                    }
                    __syncwarp();
                    if (++stage == NS) { stage = 0; phase ^= 1; }
                }
                if constexpr (NF) {
                    // last, so that the large term meets the accumulator once: acc = q~.x~ - (h1 + h2 + h3)
                    ptx::mbar_wait(&nfull[as], aphase);
                    ptx::tc_fence_after();
                    if (lane == 0) {
                        ptx::mma_bf16<CG>(d_tmem, ptx::umma_desc_k_sw32(ptx::smem_u32(ones)),
                                          ptx::umma_desc_k_sw32(ptx::smem_u32(nrec + as * L::NREC_BYTES)), idesc, 1u);
                        ptx::mma_commit<CG>(&nempty[as]);
                        ptx::mma_commit<CG>(&tfull[as]);
                    }
                    __syncwarp();
                }
                if (a_first) afph ^= 1;
                if (SEP) { aphase ^= 1; }
                else { as ^= 1; if (as == 0) aphase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue: fused top-k selection =====================
        const int e = warp & 3;                 // TMEM lane quarter this warp may access
        const int h = (warp - 2) >> 2;          // column half this warp filters (always 0 when NH = 1)
        const int row = e * 32 + lane;          // query row inside the CTA's 128
        const int epi_tid = (warp - 2) * 32 + lane;
        constexpr int EPI_THREADS = 128 * NH;
        constexpr int HCOLS = K3_BN / NH;       // columns per thread and unit
        u64* scr = scratch + (size_t)e * p.P;   // compaction scratch (half 0 only)
        int as = 0, ss = 0, xb = 0;
        uint32_t aphase = 0, sph = 0;
        const float INF = __int_as_float(0x7f800000);

        // Per-unit inputs are fetched one unit ahead: the squared norms of the next tile go to the
        // other half of xn_s with cp.async while this tile is filtered, and the list state of the
        // next query block is loaded early (or simply kept in registers when the block repeats).
        // Without this the epilogue paid two global-memory round trips (~1.9k cycles) per unit.
        // With NH = 2 the thread of column half 0 owns the list state (length, bound, write-back,
        // compaction); the thread of half 1 only appends (through the shared length) and prunes with
        // the bound it last loaded -- possibly a unit or two old, which keeps a few more keys, never fewer.
        struct QState { int cnt; float tau; u64 kth; float qn, slack; unsigned int gts; };
        auto take = [&]() -> long long {
            const long long u = sched_take_warp<CG>(sfull, sempty, sched_u, ss, sph, rank, lane);
            if (++ss == K3_NSCHED) { ss = 0; sph ^= 1; }
            return u;
        };
        auto next_is_ready = [&]() -> bool {
            return ptx::mbar_try_wait(&sfull[ss], sph);
        };
        auto issue_norms = [&](long long u, float* dst) {
            if constexpr (NF) return;
            const int64_t t = unit_tile(u);
            for (int j = epi_tid; j < K3_BN; j += EPI_THREADS) {
                const int64_t r = t * K3_BN + j;
                if (r < p.n && j < p.tile_cols) ptx::cp_async_4(dst + j, p.norms + r);
                else dst[j] = INF;
            }
        };
        auto sidx_of = [&](int qb) -> int64_t { return (int64_t)blockIdx.x * p.qpc + (int64_t)qb * K3_BM + row; };
        auto qi_of = [&](int qb) -> int64_t { return (int64_t)qb * (K3_BM * CG) + (int64_t)rank * K3_BM + row; };
        auto load_state = [&](int qb, QState& S) {
            const int64_t sidx = sidx_of(qb), qi = qi_of(qb);
            const bool valid = qi < p.nq && !p.probe;
            S.cnt = __ldcg(p.cnt + sidx);
            S.tau = __ldcg(p.tau + sidx);
            S.kth = __ldcg(p.kth + sidx);
            S.qn = valid ? p.qn[qi] : 0.f;
            S.slack = (SCREEN && valid) ? p.slack2[qi] : 0.f;
            S.gts = valid ? __ldcg(p.gts + qi) : 0u;
        };

        long long cur = take();
        QState S;
        if (cur >= 0) {
            issue_norms(cur, xn_s);
            load_state((int)(cur & 255), S);
        }
        while (cur >= 0) {
            const int64_t t = unit_tile(cur);  // units arrive as (tile << 8) | query block (| work-item flags)
            const int qb = (int)(cur & 255);
            const int64_t qi = qi_of(qb);
            const bool valid = qi < p.nq && !p.probe;
            const int64_t sidx = sidx_of(qb);
            int cnt = S.cnt;
            float tau = S.tau;
            u64 kth = S.kth;
            const float qn = S.qn, slack = S.slack;
            // exact modes: tau is (an upper bound of) the k-th exact distance.  Screening: tau bounds the
            // k-th APPROXIMATE distance and every key within 2*sigma of it must be kept (see k3_search).
            const float tau_eff = valid ? fminf(tau, __uint_as_float(S.gts)) + slack : 0.f;
            const u64 kth_eff = SCREEN ? KEY_EMPTY : kth;
            // filter threshold on s = ||x||^2 - 2 ip, widened by a few ulps so it never rejects a
            // candidate the exact test below would accept
            const float t_cmp = valid ? (tau_eff - qn) + 2e-6f * (fabsf(tau_eff) + qn) : -INF;
            // norm fold: a = q~.x~ - ||x||^2/2 passes iff ||q||^2 - 2a <= tau_eff, i.e. a >= (||q||^2 - tau_eff)/2, widened likewise
            const float thr0 = 0.5f * (qn - tau_eff);
            const float thr = valid ? thr0 - 1e-6f * (fabsf(thr0) + qn) : INF;
            // columns of this tile that are real rows (the last tile is ragged; the warm-start pass looks at fewer)
            const int64_t rows_left = p.n - t * K3_BN;
            const int ncols_tile = (int)(rows_left < p.tile_cols ? rows_left : p.tile_cols) - h * HCOLS;
            float* xn = xn_s + xb * K3_BN;
            if (NH > 1 && h == 0) cnt_s[row] = cnt;
            ptx::cp_async_wait_all();
            // this tile's norms (and list lengths) are in place for all epilogue warps
            asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");

            // look one unit ahead if the scheduler has already published it
            long long nxt = -2;
            QState N;
            bool n_loaded = false;
            if (next_is_ready()) {
                nxt = take();
                if (nxt >= 0) {
                    issue_norms(nxt, xn_s + (xb ^ 1) * K3_BN);
                    if ((int)(nxt & 255) != qb || h != 0) { load_state((int)(nxt & 255), N); n_loaded = true; }
                }
            }

            ptx::mbar_wait(&tfull[as], aphase);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(e * 32) << 16) + (uint32_t)(as * K3_BN + h * HCOLS);
            u64* list = p.cand + sidx * p.P;
            const float* xnh = xn + h * HCOLS;
            // Two register buffers: the tcgen05.ld of the next 32 columns is in flight while this
            // chunk is filtered.
            uint32_t va[32], vb[32], wa[32], wb[32];
            const uint32_t row_base = (uint32_t)(t * K3_BN) + (uint32_t)(h * HCOLS);
            // probe levels (measurement only): 1 = the filter runs but nothing passes; 2 = the accumulators are
            // loaded but not filtered; 3 = the accumulators are handed back untouched (the MMA pipeline alone)
            if (p.probe < 3) {
            ptx::tmem_ld_32x32(taddr, va);
            if (SEP) ptx::tmem_ld_32x32(taddr + K3_BN, wa);
            if (p.probe == 2) {
#pragma unroll 1
                for (int c0 = 32; c0 < HCOLS; c0 += 32) {
                    ptx::tmem_ld_wait();
                    reg_fence32(va);
                    ptx::tmem_ld_32x32(taddr + c0, va);
                }
                ptx::tmem_ld_wait();
                reg_fence32(va);
            } else
#pragma unroll 1
            for (int c0 = 0; c0 < HCOLS; c0 += 64) {
                ptx::tmem_ld_wait();
                ptx::tmem_ld_32x32(taddr + c0 + 32, vb);
                if (SEP) ptx::tmem_ld_32x32(taddr + K3_BN + c0 + 32, wb);
                reg_fence32(va);
                if (SEP) reg_fence32(wa);
                if constexpr (NF) { if (ncols_tile > c0) k3_filter_chunk_nf<(NH > 1)>(va, thr, qn, tau_eff, kth_eff, row_base + c0, ncols_tile - c0, list, cnt, cnt_s + row); }
                else
                k3_filter_chunk<SEP, (NH > 1)>(va, wa, xnh + c0, t_cmp, qn, tau_eff, kth_eff, row_base + c0, list, cnt, cnt_s + row);
                ptx::tmem_ld_wait();
                if (c0 + 64 < HCOLS) {
                    ptx::tmem_ld_32x32(taddr + c0 + 64, va);
                    if (SEP) ptx::tmem_ld_32x32(taddr + K3_BN + c0 + 64, wa);
                }
                reg_fence32(vb);
                if (SEP) reg_fence32(wb);
                if constexpr (NF) { if (ncols_tile > c0 + 32) k3_filter_chunk_nf<(NH > 1)>(vb, thr, qn, tau_eff, kth_eff, row_base + c0 + 32, ncols_tile - c0 - 32, list, cnt, cnt_s + row); }
                else
                k3_filter_chunk<SEP, (NH > 1)>(vb, wb, xnh + c0 + 32, t_cmp, qn, tau_eff, kth_eff, row_base + c0 + 32, list, cnt, cnt_s + row);
            }
            }
            // accumulator drained: hand it back to the MMA warp (of the leader CTA)
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (CG == 1 || rank == 0) ptx::mbar_arrive(&tempty[as]);
                else ptx::mbar_arrive_cluster(&tempty[as], 0);
            }
            if (NH > 1) {
                // both column halves of this lane quarter have finished appending (their keys are
                // visible to each other: bar.sync orders the participants' memory accesses)
                asm volatile("bar.sync %0, 64;" ::"r"(2 + e) : "memory");
                if (h == 0) cnt = cnt_s[row];
            }

            // compaction of lists that could overflow on the next tile (warp-cooperative, half 0).
            // Half 1 cannot touch these lists meanwhile: its next append comes after the next bar.sync 1,
            // where half 0 only arrives once it is done here.
            unsigned need = h == 0 ? __ballot_sync(0xffffffffu, valid && cnt > p.P - K3_BN) : 0u;
            while (need) {
                const int Lq = __ffs(need) - 1;
                need &= need - 1;
                const int c = __shfl_sync(0xffffffffu, cnt, Lq);
                u64* lst = p.cand + (sidx - lane + Lq) * p.P;
                int Ps = 2;
                while (Ps < c) Ps <<= 1;
                __syncwarp();
                for (int i = lane; i < Ps; i += 32) scr[i] = i < c ? __ldcg(lst + i) : KEY_EMPTY;
                __syncwarp();
                warp_bitonic_sort(scr, Ps, lane);
                int nc = c < p.k ? c : p.k;
                if (SCREEN && c >= p.k) {
                    // keep every key within the slack of this list's k-th approximate distance
                    const float thr = key_dist(scr[p.k - 1]) + __shfl_sync(0xffffffffu, slack, Lq);
                    int le = 0;
                    for (int i = lane; i < c; i += 32) le += key_dist(scr[i]) <= thr ? 1 : 0;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) le += __shfl_xor_sync(0xffffffffu, le, o);
                    nc = le;
                    if (nc > p.P - K3_BN) {  // cannot make room for the next tile without dropping candidates
                        if (lane == 0) atomicExch(p.overflow, 1u);
                        nc = p.P - K3_BN;
                    }
                }
                for (int i = lane; i < nc; i += 32) lst[i] = scr[i];
                const u64 newkth = scr[p.k - 1 < Ps ? p.k - 1 : 0];
                __syncwarp();
                if (lane == Lq) {
                    cnt = nc;
                    if (nc >= p.k) {
                        kth = newkth;
                        tau = key_dist(newkth);
                        atomicMin(p.gts + qi, __float_as_uint(tau));
                    }
                }
            }

            // hand over to the next unit
            if (nxt == -2) {  // the scheduler had not published it yet: take it now (blocking)
                nxt = take();
                if (nxt >= 0) issue_norms(nxt, xn_s + (xb ^ 1) * K3_BN);
            }
            const int nqb_next = nxt >= 0 ? (int)(nxt & 255) : -1;
            if (h != 0) {
                // appending half: no state to write back; the next unit's bounds were (or are now) loaded
                if (nxt >= 0) {
                    if (n_loaded) S = N;
                    else load_state(nqb_next, S);
                }
            } else if (nqb_next == qb) {
                // same query block again: the list state stays in registers, only the shared bound is refreshed
                S.cnt = cnt; S.tau = tau; S.kth = kth;
                S.gts = valid ? __ldcg(p.gts + qi) : 0u;
                if (NH > 1) { p.tau[sidx] = tau; p.kth[sidx] = kth; }  // for the other half's next look
            } else {
                p.cnt[sidx] = cnt;
                p.tau[sidx] = tau;
                p.kth[sidx] = kth;
                if (nxt >= 0) {
                    if (n_loaded) S = N;
                    else load_state(nqb_next, S);
                }
            }
            cur = nxt;
            xb ^= 1;
            if (SEP) { aphase ^= 1; }
            else { as ^= 1; if (as == 0) aphase ^= 1; }
        }
        ptx::cp_async_wait_all();
    }

    // ---- teardown ----
    ptx::tc_fence_before();
    __syncthreads();
    if (CG == 2) ptx::cluster_sync_all();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<CG>(tmem_base, 512);
    }
}

// ---- state initialisation -------------------------------------------------------------------
__global__ void k3_init_kernel(int64_t nslots, int* cnt, float* tau, u64* kth, int64_t nq, unsigned int* gts,
                               u64* next_unit, unsigned int* overflow, unsigned short* ones) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ones && i < K3_BM * 16) ones[i] = (i & 15) < 3 ? 0x3F80 : 0;  // bf16 1.0 in the three columns that meet the norm terms
    if (i == 0) { *next_unit = 0; if (overflow) *overflow = 0; }
    if (i < nslots) { cnt[i] = 0; tau[i] = FLT_MAX; kth[i] = (u64)FLT_MAX_BITS << 32; }
    if (gts && i < nq) gts[i] = FLT_MAX_BITS;
}

// ---- K4: per query, merge the lists of every CTA that saw it ---------------------------------
// One CTA per query.  The lists hold every key under the warm-started shared bound: a few
// thousand per query, of which k are wanted.  A full sort of them is what made this kernel cost
// 0.2 ms, so selection is sample-based: sort 512 of the gathered keys, take the sample rank that
// (with a 4-sigma margin) leaves at least k keys of the whole set below it, keep only the keys
// under that value, and sort those.  The count is checked, so the result is exact: if the guess
// leaves fewer than k keys (or too many), the kernel falls back to sorting everything.
static constexpr int K4_THREADS = 256;
// K4_P (template parameter) = gathered keys kept in shared memory; 4096 is what runs.  A query
// with more live keys takes the sample-and-regather path, which is exact for any buffer size.
static constexpr int K4_SAMPLE = 512;
static constexpr int K4_DIRECT = 1024; // up to this many gathered keys are simply sorted

template <int K4_P>
struct K4Smem {
    static constexpr int OUT = K4_P / 2;  // keys under the refined bound (k <= 1024 plus the sampling margin)
    u64 buf[K4_P];
    u64 smp[K4_SAMPLE];
    u64 out[OUT];
    int cnt, cnt2;
    int sel[2];  // radix select: (bin, keys below it)
    u64 kth;
};

// every list_step-th list of query (qb,row), all keys <= kbound -> sm.buf (first K4_P of them); returns the total
template <int K4_P>
__device__ __forceinline__ int k4_gather(const K3Params& p, int cg, int ncta, int lstep, int r, int qb, int row,
                                         u64 kbound, K4Smem<K4_P>& sm) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __syncthreads();
    if (tid == 0) sm.cnt = 0;
    __syncthreads();
    constexpr int LB = 4;  // four lists per warp at a time: their counts, then their keys, are independent loads
    const int wstride = (K4_THREADS / 32) * LB * lstep;
    for (int cta0 = r + warp * LB * lstep; cta0 < ncta; cta0 += wstride) {
        int c[LB];
        const u64* lst[LB];
        int cmax = 0;
#pragma unroll
        for (int j = 0; j < LB; ++j) {
            const int cta = cta0 + j * lstep;
            const int64_t sidx = (int64_t)(cta < ncta ? cta : r) * p.qpc + (int64_t)qb * K3_BM + row;
            c[j] = cta < ncta ? p.cnt[sidx] : 0;
            lst[j] = p.cand + sidx * p.P;
        }
#pragma unroll
        for (int j = 0; j < LB; ++j) cmax = c[j] > cmax ? c[j] : cmax;
        for (int i = lane; i < cmax; i += 32) {
            u64 key[LB];
#pragma unroll
            for (int j = 0; j < LB; ++j) key[j] = i < c[j] ? lst[j][i] : KEY_EMPTY;
#pragma unroll
            for (int j = 0; j < LB; ++j) {
                if (key[j] <= kbound) {  // KEY_EMPTY (all ones) is above any bound
                    const int pos = atomicAdd(&sm.cnt, 1);
                    if (pos < K4_P) sm.buf[pos] = key[j];
                }
            }
        }
    }
    __syncthreads();
    return sm.cnt;
}

// sort the first n keys of a (capacity cap, power of two); block-wide
__device__ __forceinline__ void k4_sort(u64* a, int n, int cap) {
    int Ps = 2;
    while (Ps < n) Ps <<= 1;
    if (Ps > cap) Ps = cap;
    for (int j = n + threadIdx.x; j < Ps; j += K4_THREADS) a[j] = KEY_EMPTY;
    __syncthreads();
    block_bitonic_sort(a, Ps);
}

// Float bits of the k-th smallest distance (1 <= k <= n) among the n keys of `a`: a 3-pass radix
// select (11 + 11 + 10 bits) over the distance half of the keys, histogram in shared memory.  The
// bound refreshes only need this number, and a full bitonic sort of the gathered keys was half of
// their 57 us.  Distances are non-negative floats, so their bits order like unsigned integers.
// hist: 2048 words of scratch.  Block-wide; every thread returns the result.
__device__ __forceinline__ uint32_t k4_select_dist(const u64* a, int n, int k, uint32_t* hist, int* scratch2) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __shared__ int s_wsum[K4_THREADS / 32];
    uint32_t prefix = 0, mask = 0;
    int remaining = k;
#pragma unroll 1
    for (int pass = 0; pass < 3; ++pass) {
        const int shift = pass == 0 ? 21 : (pass == 1 ? 10 : 0);
        const int nb = pass == 2 ? 1024 : 2048;
        for (int b = tid; b < 2048; b += K4_THREADS) hist[b] = 0;
        __syncthreads();
        for (int i = tid; i < n; i += K4_THREADS) {
            const uint32_t dbits = (uint32_t)(a[i] >> 32);
            if ((dbits & mask) == prefix) atomicAdd(&hist[(dbits >> shift) & (uint32_t)(nb - 1)], 1u);
        }
        __syncthreads();
        // each thread owns 8 consecutive bins; block-wide exclusive scan of the per-thread sums
        int mine = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) mine += (int)hist[tid * 8 + j];
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_wsum[warp] = incl;
        __syncthreads();
        int before = incl - mine;
        for (int w = 0; w < warp; ++w) before += s_wsum[w];
        if (before < remaining && remaining <= before + mine) {  // exactly one thread: the k-th key is in its bins
            int c = before;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int hcount = (int)hist[tid * 8 + j];
                if (c < remaining && remaining <= c + hcount) { scratch2[0] = tid * 8 + j; scratch2[1] = c; }
                c += hcount;
            }
        }
        __syncthreads();
        prefix |= (uint32_t)scratch2[0] << shift;
        mask |= (uint32_t)(nb - 1) << shift;
        remaining -= scratch2[1];
        __syncthreads();
    }
    return prefix;
}

// An UPPER BOUND of the k-th smallest distance (1 <= k <= n) among the n keys of `a` that is itself one of the
// distances: a 2048-bin histogram over [min, max] of the keys, the bin in which the running count reaches k, and
// the largest key in the bins up to there.  The bound refreshes of the pass only prune with the number, so
// "k-th smallest or a hair above" is as good as exact, and it costs one histogram pass instead of three
// (9 block barriers instead of 15).  hist: 2048 words.  Block-wide; every thread returns the result.
__device__ __forceinline__ uint32_t k4_bound_linear(const u64* a, int n, int k, uint32_t* hist, int* scratch2) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __shared__ float s_lo[K4_THREADS / 32], s_hi[K4_THREADS / 32];
    __shared__ int s_ws[K4_THREADS / 32];
    float lo = __int_as_float(0x7f800000), hi = 0.f;
    for (int i = tid; i < n; i += K4_THREADS) {
        const float dv = key_dist(a[i]);
        lo = fminf(lo, dv); hi = fmaxf(hi, dv);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0) { s_lo[warp] = lo; s_hi[warp] = hi; }
    for (int b = tid; b < 2048; b += K4_THREADS) hist[b] = 0;
    __syncthreads();
#pragma unroll
    for (int w = 0; w < K4_THREADS / 32; ++w) { lo = fminf(lo, s_lo[w]); hi = fmaxf(hi, s_hi[w]); }
    if (!(hi > lo) || !(hi < FLT_MAX)) return __float_as_uint(hi);  // all keys equal (block-uniform exit)
    const float scale = 2047.f / (hi - lo);
    for (int i = tid; i < n; i += K4_THREADS) {
        int b = (int)((key_dist(a[i]) - lo) * scale);
        b = b < 0 ? 0 : (b > 2047 ? 2047 : b);
        atomicAdd(&hist[b], 1u);
    }
    __syncthreads();
    int mine = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) mine += (int)hist[tid * 8 + j];
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_ws[warp] = incl;
    __syncthreads();
    int before = incl - mine;
    for (int w = 0; w < warp; ++w) before += s_ws[w];
    if (before < k && k <= before + mine) {  // exactly one thread: the k-th key is in its bins
        int c = before;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int hcount = (int)hist[tid * 8 + j];
            if (c < k && k <= c + hcount) scratch2[0] = tid * 8 + j;
            c += hcount;
        }
    }
    __syncthreads();
    const int bsel = scratch2[0];
    float best = 0.f;  // largest key in bins 0..bsel (same bin arithmetic as above, so it covers at least k keys)
    for (int i = tid; i < n; i += K4_THREADS) {
        const float dv = key_dist(a[i]);
        int b = (int)((dv - lo) * scale);
        b = b < 0 ? 0 : (b > 2047 ? 2047 : b);
        if (b <= bsel) best = fmaxf(best, dv);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
    __syncthreads();  // s_hi is reused
    if (lane == 0) s_hi[warp] = best;
    __syncthreads();
#pragma unroll
    for (int w = 0; w < K4_THREADS / 32; ++w) best = fmaxf(best, s_hi[w]);
    return __float_as_uint(best);
}

template <int K4_P>
__global__ void __launch_bounds__(K4_THREADS)
k4_merge_kernel(const K3Params p, int cg, int ncta, int threshold_only, int list_step, float* __restrict__ D,
                int64_t* __restrict__ I, int set_counter, u64 counter_value) {
    extern __shared__ __align__(16) unsigned char k4_smem_raw[];
    K4Smem<K4_P>& sm = *reinterpret_cast<K4Smem<K4_P>*>(k4_smem_raw);
    const int tid = threadIdx.x;
    // a bound refresh between two segments of the pass also rewinds the scheduler's counter for the next one
    if (set_counter && blockIdx.x == 0 && tid == 0) *p.next_unit = counter_value;
    const int64_t qi = blockIdx.x;
    const int bq = K3_BM * cg;
    const int qb = (int)(qi / bq);
    const int r = (int)(qi - (int64_t)qb * bq) / K3_BM;
    const int row = (int)(qi % K3_BM);
    const int lstep = cg * list_step;
    const int k = p.k;
    u64 kbound = ((u64)p.gts[qi] << 32) | 0xffffffffull;
    const u64* res = nullptr;  // sorted keys, the k smallest first
    int nres = 0;

    for (int attempt = 0; attempt < 4 && !res; ++attempt) {
        const int total = k4_gather(p, cg, ncta, lstep, r, qb, row, kbound, sm);
        const int have = total < K4_P ? total : K4_P;
        if (threshold_only && attempt == 0 && total <= K4_P) {
            // bound refresh: only the k-th smallest distance is wanted, no order
            if (total >= k) {
                const uint32_t dk = k4_bound_linear(sm.buf, total, k, reinterpret_cast<uint32_t*>(sm.out), sm.sel);
                if (tid == 0) atomicMin(p.gts + qi, dk);
            }
            return;
        }
        if (total <= K4_DIRECT) {  // small enough: sort it all
            if (attempt > 0 && total < k) break;  // the tightened bound cut too deep: exhaustive path
            k4_sort(sm.buf, have, K4_P);
            res = sm.buf; nres = have;
            break;
        }
        // sample -> tighter bound
        for (int j = tid; j < K4_SAMPLE; j += K4_THREADS) sm.smp[j] = sm.buf[j];  // first arrivals: an arbitrary subset
        __syncthreads();
        block_bitonic_sort(sm.smp, K4_SAMPLE);
        const float mean = (float)k * (float)K4_SAMPLE / (float)total;      // expected sample rank of the k-th key
        const int rk = (int)(mean + 4.f * sqrtf(mean) + 3.f);
        if (rk >= K4_SAMPLE) break;                                          // cannot tighten: sort everything
        const u64 b2 = sm.smp[rk];
        if (total > K4_P) { kbound = b2; continue; }                         // did not fit: gather again under b2
        if (tid == 0) sm.cnt2 = 0;
        __syncthreads();
        for (int j = tid; j < have; j += K4_THREADS) {
            const u64 key = sm.buf[j];
            if (key <= b2) {
                const int pos = atomicAdd(&sm.cnt2, 1);
                if (pos < K4Smem<K4_P>::OUT) sm.out[pos] = key;
            }
        }
        __syncthreads();
        const int c2 = sm.cnt2;
        if (c2 >= k && c2 <= K4Smem<K4_P>::OUT) {
            k4_sort(sm.out, c2, K4Smem<K4_P>::OUT);
            res = sm.out; nres = c2;
        }
        break;  // unlucky guess: fall through to the exhaustive path
    }

    if (!res) {
        // exhaustive path: stream the lists under the original bound, compacting as needed
        const float bound = __uint_as_float(p.gts[qi]);
        __syncthreads();
        if (tid == 0) { sm.cnt = 0; sm.kth = KEY_EMPTY; }
        __syncthreads();
        for (int cta = r; cta < ncta; cta += lstep) {
            const int64_t sidx = (int64_t)cta * p.qpc + (int64_t)qb * K3_BM + row;
            const int c = p.cnt[sidx];
            const u64* lst = p.cand + sidx * p.P;
            for (int i0 = 0; i0 < c; i0 += K4_THREADS) {
                int flag = 0;
                const int i = i0 + tid;
                if (i < c) {
                    const u64 key = lst[i];
                    if (key_dist(key) <= bound && key < sm.kth) {
                        int pos = atomicAdd(&sm.cnt, 1);
                        sm.buf[pos] = key;
                        if (pos + 1 > K4_P - K4_THREADS) flag = 1;
                    }
                }
                if (__syncthreads_or(flag)) {
                    const int cc = sm.cnt;
                    k4_sort(sm.buf, cc, K4_P);
                    if (tid == 0) { sm.cnt = cc < k ? cc : k; if (cc >= k) sm.kth = sm.buf[k - 1]; }
                    __syncthreads();
                }
            }
        }
        __syncthreads();
        nres = sm.cnt;
        k4_sort(sm.buf, nres, K4_P);
        res = sm.buf;
    }

    if (threshold_only) {
        // warm start: the k-th smallest distance over a sample of the rows bounds the final one
        if (tid == 0 && nres >= k) atomicMin(p.gts + qi, (unsigned int)(res[k - 1] >> 32));
        return;
    }
    for (int j = tid; j < k; j += K4_THREADS) {
        if (j < nres) {
            D[qi * k + j] = key_dist(res[j]);
            I[qi * k + j] = (int64_t)key_row(res[j]);
        } else {
            D[qi * k + j] = FLT_MAX;
            I[qi * k + j] = -1;
        }
    }
}

// ---- screening support: per-query slack, and K6, the exact re-rank ---------------------------
// Screening mode runs ONE TF32 product per 8 columns instead of three.  A kind::tf32 MMA reads
// q and x with their low 13 mantissa bits ignored, so |q.x - (q.x)_tf32| <= 2^-9 * sum|q_t||x_t|
// <= 2^-9 |q||x| (Cauchy-Schwarz), and the approximate distance d~ = |q|^2 + |x|^2 - 2 (q.x)_tf32
// is within  sigma_q = c * 2^-8 * |q| * max_rows|x|  of the exact one (c = 1.02 + 8e-6 d also
// covers the accumulator truncation, ~d/8 * 2^-23).  If T is (an upper bound of) the k-th smallest
// d~, every exact top-k row has d~ <= T + 2 sigma_q.  The lists therefore keep every key with
// d~ <= T + 2 sigma_q, and K6 computes the exact fp32 distance (direct differences, as the scan
// kernel does) of those few hundred rows per query and selects the final k.  The result is exact;
// the tensor cores only decide which rows are worth an exact look.
static constexpr float K3_ROUND_SLACK = 8.f * 1.1920929e-7f;  // 2 * (4 ulps) of |q|^2 + max|x|^2
__global__ void k3_slack_kernel(int64_t nq, const float* __restrict__ qn, const unsigned int* __restrict__ max_norm_bits,
                                float coef, float* __restrict__ slack2) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const float mx2 = __uint_as_float(*max_norm_bits);
    // + the fp32 rounding of d~ = (|q|^2 + |x|^2) - 2 ip itself and of the two norms (a few ulps of
    // |q|^2 + |x|^2): it is all that is left when q = 0 or |q| >> max|x|, where the product term vanishes
    slack2[i] = coef * sqrtf(qn[i]) * sqrtf(mx2) + K3_ROUND_SLACK * (qn[i] + mx2);
}

// bf16 screening: q~, x~ are the bf16 roundings.  q.x - q~.x~ = (q - q~).x + q~.(x - x~), so by
// Cauchy-Schwarz |q.x - q~.x~| <= |q - q~| max|x| + |q~| max|x - x~|, with the rounding-error norms
// MEASURED (K1 bf16 pass: per query here, running maximum over the rows at add time) instead of the
// worst case 2^-8: typically 0.3 * 2^-8.  |q~| <= |q| (1 + 2^-8).  Products of two bf16 values are
// exact in the fp32 accumulator; the accumulation error term is the same as in the TF32 mode.
// On entry slack2[i] holds |q_i - q~_i|^2; on exit 2 * sigma_i.
__global__ void k3_slack_bf16_kernel(int64_t nq, const float* __restrict__ qn, const unsigned int* __restrict__ max_norm_bits,
                                     const unsigned int* __restrict__ max_err_bits, float acc_coef, float* __restrict__ slack2) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const float nqv = sqrtf(qn[i]), eq = sqrtf(slack2[i]);
    const float mx = sqrtf(__uint_as_float(*max_norm_bits)), ex = sqrtf(__uint_as_float(*max_err_bits));
    const float sigma = 2.f * 1.02f * (eq * mx + 1.004f * nqv * ex) + acc_coef * nqv * mx;
    slack2[i] = 2.f * sigma + K3_ROUND_SLACK * (qn[i] + mx * mx);  // rounding of d~ and of the norms, see k3_slack_kernel
}

template <int K4_P>
__global__ void __launch_bounds__(K4_THREADS)
k6_rerank_kernel(const K3Params p, int cg, int ncta, const float* __restrict__ X, const float* __restrict__ Q,
                 float* __restrict__ D, int64_t* __restrict__ I, int refresh_first) {
    extern __shared__ __align__(16) unsigned char k4_smem_raw[];
    K4Smem<K4_P>& sm = *reinterpret_cast<K4Smem<K4_P>*>(k4_smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t qi = blockIdx.x;
    const int bq = K3_BM * cg;
    const int qb = (int)(qi / bq);
    const int r = (int)(qi - (int64_t)qb * bq) / K3_BM;
    const int row = (int)(qi % K3_BM);
    const int k = p.k, d = p.d;
    // gts holds the k-th smallest approximate distance of the whole pass (K4 in threshold mode ran behind it), or,
    // with refresh_first, that of the rows scanned up to the last refresh: an upper bound of the final one
    const float slack = p.slack2[qi];
    const float bound = __uint_as_float(p.gts[qi]) + slack;
    const u64 kbound = bound < FLT_MAX ? (((u64)__float_as_uint(bound) << 32) | 0xffffffffull)
                                       : (((u64)FLT_MAX_BITS << 32) - 1);
    int total = k4_gather(p, cg, ncta, cg, r, qb, row, kbound, sm);
    if (total > K4_P && refresh_first) {
        // the bound of the last refresh is too loose for one gather: the refresh proper first (keys up to the old
        // bound, k-th smallest of those), then the keys within the slack of the result
        const u64 kb0 = ((u64)p.gts[qi] << 32) | 0xffffffffull;
        total = k4_gather(p, cg, ncta, cg, r, qb, row, kb0, sm);
        if (total <= K4_P) {
            float b2 = bound;
            if (total >= k) b2 = __uint_as_float(k4_select_dist(sm.buf, total, k, reinterpret_cast<uint32_t*>(sm.out), sm.sel)) + slack;
            const u64 kb2 = b2 < FLT_MAX ? (((u64)__float_as_uint(b2) << 32) | 0xffffffffull) : (((u64)FLT_MAX_BITS << 32) - 1);
            total = k4_gather(p, cg, ncta, cg, r, qb, row, kb2, sm);
        }
        refresh_first = 0;
    }
    if (total > K4_P) {  // more rows inside the slack than the buffer holds: the host reruns exactly
        if (tid == 0) atomicExch(p.overflow, 1u);
        return;
    }
    if (refresh_first && total >= k) {
        // The last bound refresh folded in: what was gathered is a superset of the k smallest approximate
        // distances, so its k-th smallest IS the final k-th; keep the keys within the slack of that.
        const uint32_t dk = k4_select_dist(sm.buf, total, k, reinterpret_cast<uint32_t*>(sm.out), sm.sel);
        const float b2 = __uint_as_float(dk) + slack;
        constexpr int PER = K4_P / K4_THREADS;
        u64 mine[PER];
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int i = tid + j * K4_THREADS;
            mine[j] = i < total ? sm.buf[i] : KEY_EMPTY;
        }
        __syncthreads();
        if (tid == 0) sm.cnt2 = 0;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < PER; ++j)
            if (mine[j] != KEY_EMPTY && key_dist(mine[j]) <= b2) sm.buf[atomicAdd(&sm.cnt2, 1)] = mine[j];
        __syncthreads();
        total = sm.cnt2;
    }
    if (p.outl) {
        // rows the screen never sees (their norms would have blown up every query's slack, k1_prep.cu): each of them
        // gets its exact distance for every query, like any other candidate
        const int nout = (int)p.outl[0] < 64 ? (int)p.outl[0] : 64;
        if (total + nout > K4_P) {
            if (tid == 0) atomicExch(p.overflow, 1u);
            return;
        }
        __syncthreads();
        if (tid == 0) sm.cnt2 = 0;
        __syncthreads();
        if (tid < nout) {
            const unsigned int id = p.outl[8 + tid];
            if ((int64_t)id < p.n) sm.buf[total + atomicAdd(&sm.cnt2, 1)] = make_key(0.f, id);
        }
        __syncthreads();
        total += sm.cnt2;
    }
    // exact distances: one warp per candidate (two candidates in flight per warp), 128-bit loads,
    // fixed summation order
    const float* qv = Q + qi * (int64_t)d;
    constexpr int NW = K4_THREADS / 32;
    for (int c = warp; c < total; c += 2 * NW) {
        const int c2 = c + NW;
        const bool has2 = c2 < total;
        const uint32_t xr0 = key_row(sm.buf[c]);
        const uint32_t xr1 = has2 ? key_row(sm.buf[c2]) : xr0;
        const float* xv0 = X + (int64_t)xr0 * d;
        const float* xv1 = X + (int64_t)xr1 * d;
        float acc0 = 0.f, acc1 = 0.f;
        // the rows are random 4*d-byte reads from HBM: up to four 512-byte slices of both rows are requested
        // before the first is used (one memory round trip per 512 columns instead of one per 128); the
        // summation order per lane is unchanged
        for (int t0 = lane * 4; t0 < d; t0 += 512) {
            float4 a[4], b0[4], b1[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int t = t0 + j * 128;
                if (t < d) {
                    a[j] = __ldg(reinterpret_cast<const float4*>(qv + t));
                    b0[j] = ldg_stream(reinterpret_cast<const float4*>(xv0 + t));
                    b1[j] = ldg_stream(reinterpret_cast<const float4*>(xv1 + t));
                } else {
                    a[j] = b0[j] = b1[j] = make_float4(0.f, 0.f, 0.f, 0.f);  // adds exact zeros
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float e0 = a[j].x - b0[j].x, e1 = a[j].y - b0[j].y, e2 = a[j].z - b0[j].z, e3 = a[j].w - b0[j].w;
                acc0 = fmaf(e0, e0, acc0); acc0 = fmaf(e1, e1, acc0); acc0 = fmaf(e2, e2, acc0); acc0 = fmaf(e3, e3, acc0);
                e0 = a[j].x - b1[j].x; e1 = a[j].y - b1[j].y; e2 = a[j].z - b1[j].z; e3 = a[j].w - b1[j].w;
                acc1 = fmaf(e0, e0, acc1); acc1 = fmaf(e1, e1, acc1); acc1 = fmaf(e2, e2, acc1); acc1 = fmaf(e3, e3, acc1);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            acc0 += __shfl_xor_sync(0xffffffffu, acc0, o);
            acc1 += __shfl_xor_sync(0xffffffffu, acc1, o);
        }
        __syncwarp();
        if (lane == 0) {
            sm.buf[c] = acc0 < FLT_MAX ? make_key(acc0, xr0) : KEY_EMPTY;
            if (has2) sm.buf[c2] = acc1 < FLT_MAX ? make_key(acc1, xr1) : KEY_EMPTY;
        }
    }
    __syncthreads();
    k4_sort(sm.buf, total, K4_P);
    for (int j = tid; j < k; j += K4_THREADS) {
        const u64 key = j < total ? sm.buf[j] : KEY_EMPTY;
        if (key != KEY_EMPTY) {
            D[qi * k + j] = key_dist(key);
            I[qi * k + j] = (int64_t)key_row(key);
        } else {
            D[qi * k + j] = FLT_MAX;
            I[qi * k + j] = -1;
        }
    }
}

// ---- host side ----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int g_default_cg = 0;  // 0 = choose per search (k3_cg_for)

struct K3State {
    int device = 0, sm_count = 148;
    int cg = g_default_cg;
    EncodeTiledFn encode = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    unsigned int* flag_host = nullptr;  // pinned: overflow flag of the screening mode
    bool timing = false;
    int last_mode = 0;  // 1 = 3xTF32, 2 = TF32 screening + re-rank, 3 = bf16 screening + re-rank (what the last search ran)
    int sep_mode = 0;  // long-row accuracy: 0 auto (d >= 512: separate accumulator), 1 separate accumulator, 2 off
    bool timed = false;
    unsigned int* ext_flag = nullptr;  // caller-owned overflow word (deferred screening check)
    int epi_halves = 0;  // 0 auto (2 where available), 1 four epilogue warps, 2 eight
    int segments = 1;  // cut the main pass into segments with a bound refresh in between (0 = one launch)
    int warm_keys = 12;  // warm-start sample per query, in multiples of k (option "tensor_warm_keys")
    int pf_tiles = 0;  // L2 prefetch distance in row tiles (option "tensor_prefetch_tiles"; 0 = off)
    int qs = 0;        // query-stationary schedule where it applies (option "tensor_query_stationary"; measured: no gain)
    int probe = 0;     // dense-rate probe (option "tensor_probe"): the pass runs with the filter shut, results are padding
};

static int k3_P(int k) { return next_pow2(k + K3_BN); }
// CG = 1 has shared memory for compaction scratch up to 512 keys per warp; CG = 2 up to 2048.
// 3xTF32: CTA pairs for long rows (measured faster at d = 1024: 815 vs 782 TFLOP/s) and for long
// lists.  Screening: the kernel is bound by the operand stream from L2 (~15 TB/s chip-wide at
// either precision), of which a pair needs a third less (each CTA loads half of the database
// tile), so pairs win once there is enough work to fill them: 6.15 vs 6.67 ms at C3 (bf16),
// 1.69 vs 2.21 ms at k = 1000, but 48 vs 35 us on 200k rows x 64 queries.
static int k3_cg_for(int k, int want, int d, bool screen, int64_t n, int64_t nq, int sm_count) {
    if (want == 1 || want == 2) return (want == 1 && !screen && k3_P(k) > 512) ? 2 : want;
    const int64_t units2 = ((n + K3_BN - 1) / K3_BN) * ((nq + 2 * K3_BM - 1) / (2 * K3_BM));
    // a pair serves 256 queries per unit: with <= 128 its second CTA would multiply padding
    const bool enough = nq > K3_BM && units2 >= 64LL * (sm_count / 2);
    if (screen) return enough ? 2 : 1;
    if (k3_P(k) > 512 || d >= 512) return 2;
    return enough ? 2 : 1;  // 3xTF32 at C3: 25.7 ms with pairs, 33.4 ms without
}

bool k3_supported(int d, int k) {
    // TMA needs a 16-byte row pitch
    return d % 4 == 0 && k3_P(k) <= 2048;
}

void k3_set_cta_group(K3State* s, int cg);

struct K3Layout {
    int cg, ncta, nqb, qpc, P;
    size_t off_qlo, off_qn, off_gts, off_slack, off_next, off_ones, off_cnt, off_tau, off_kth, off_cand, total;
};

static K3Layout k3_layout(int sm_count, int d, int64_t nq, int k, int want_cg, bool screen, int64_t n) {
    K3Layout L;
    L.cg = k3_cg_for(k, want_cg, d, screen, n, nq, sm_count);
    L.ncta = (sm_count / L.cg) * L.cg;
    L.nqb = (int)((nq + K3_BM * L.cg - 1) / (K3_BM * L.cg));
    L.qpc = L.nqb * K3_BM;
    L.P = k3_P(k);
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t o = 0;
    // query-side operand copy (residual rows, bf16 queries, or a padded fp32 copy): whole query blocks,
    // so the TMA box of the last block never leaves the tensor (see k3_search)
    L.off_qlo = o; o += al((size_t)L.nqb * K3_BM * L.cg * d * 4);
    L.off_qn = o; o += al((size_t)nq * 4);
    L.off_gts = o; o += al((size_t)nq * 4);
    L.off_slack = o; o += al((size_t)nq * 4);
    L.off_next = o; o += 256;  // [0] scheduler counter (u64), [16] overflow flag
    L.off_ones = o; o += K3_BM * 32;  // the block of ones of the norm fold: [128][16] bf16, columns 0..2 = 1
    size_t slots = (size_t)L.ncta * L.qpc;
    L.off_cnt = o; o += al(slots * 4);
    L.off_tau = o; o += al(slots * 4);
    L.off_kth = o; o += al(slots * 8);
    L.off_cand = o; o += al(slots * (size_t)L.P * 8);
    L.total = o;
    return L;
}

size_t k3_scratch_bytes(int sm_count, int d, int64_t nq, int k) {
    // sized for either CTA-group configuration
    size_t a = k3_layout(sm_count, d, nq, k, 1, true, 0).total, b = k3_layout(sm_count, d, nq, k, 2, false, 0).total;
    return a > b ? a : b;
}

int k3_create(K3State** out, int device, int sm_count) {
    K3State* s = new K3State();
    s->device = device;
    s->sm_count = sm_count;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
        set_error("cuTensorMapEncodeTiled not available from the driver (%s)", cudaGetErrorString(e));
        delete s;
        return B2K_ECUDA;
    }
    s->encode = (EncodeTiledFn)fn;
    e = cudaEventCreate(&s->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&s->ev1);
    if (e == cudaSuccess) e = cudaMallocHost(reinterpret_cast<void**>(&s->flag_host), 64);
    if (e != cudaSuccess) {
        set_error("tensor kernel state: %s", cudaGetErrorString(e));
        if (s->ev0) cudaEventDestroy(s->ev0);
        if (s->ev1) cudaEventDestroy(s->ev1);
        delete s;
        return e == cudaErrorMemoryAllocation ? B2K_ENOMEM : B2K_ECUDA;
    }
    *out = s;
    return B2K_OK;
}

void k3_destroy(K3State* s) {
    if (!s) return;
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->flag_host) cudaFreeHost(s->flag_host);
    delete s;
}

void k3_set_timing(K3State* s, bool on) { s->timing = on; }
void k3_set_cta_group(K3State* s, int cg) { s->cg = (cg == 1 || cg == 2) ? cg : 0; }
void k3_set_sep_mode(K3State* s, int m) { s->sep_mode = m; }
void k3_set_segments(K3State* s, int on) { s->segments = on; }
void k3_set_epi_halves(K3State* s, int nh) { s->epi_halves = nh; }
void k3_set_ext_flag(K3State* s, unsigned int* flag) { s->ext_flag = flag; }
void k3_set_probe(K3State* s, int on) { s->probe = on < 0 ? 0 : (on > 3 ? 3 : on); }
void k3_set_query_stationary(K3State* s, int on) { s->qs = on != 0; }
void k3_set_prefetch_tiles(K3State* s, int n) { s->pf_tiles = n < 0 ? 0 : n; }
void k3_set_warm_keys(K3State* s, int n) { s->warm_keys = n < 2 ? 2 : (n > 64 ? 64 : n); }
int k3_last_mode(const K3State* s) { return s->last_mode; }

int k3_last_kernel_ms(K3State* s, float* ms) {
    if (!s->timed) return B2K_EINVAL;
    B2K_CUDA(cudaEventSynchronize(s->ev1));
    B2K_CUDA(cudaEventElapsedTime(ms, s->ev0, s->ev1));
    return B2K_OK;
}

// rec32: the tensor is [rows][16] bf16 (32-byte records: norm terms / the ones block), boxes are whole records, 32-byte swizzle
static int make_map(K3State* s, CUtensorMap* m, const void* base, int64_t rows, int d, int box_rows, bool bf16 = false,
                    bool rec32 = false) {
    cuuint64_t gdim[2] = {(cuuint64_t)d, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)d * (bf16 ? 2 : 4)};
    cuuint32_t box[2] = {(cuuint32_t)(rec32 ? 16 : bf16 ? 64 : K3_BK), (cuuint32_t)box_rows};  // 128 bytes wide (32 for records)
    cuuint32_t estr[2] = {1, 1};
    CUresult r = s->encode(m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, rec32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d) for %lld x %d", (int)r, (long long)rows, d);
        return B2K_ECUDA;
    }
    return B2K_OK;
}

static constexpr size_t K3_SMEM_LIMIT = 227 * 1024;
struct K3Maps { CUtensorMap q, ql, x, xl, n, one; };

template <int CG, int NS, int MODE, int NH, bool QS = false>
static int launch_k3(const K3Maps& mm,
                     const K3Params& p, int ncta, cudaStream_t st) {
    const int nkc = (p.d + (MODE == 3 ? 64 : 32) - 1) / (MODE == 3 ? 64 : 32);
    size_t smem = K3Smem<CG, NS, MODE, QS>::bytes(p.P, nkc);
    B2K_CUDA(cudaFuncSetAttribute(k3_tf32x3_topk_kernel<CG, NS, MODE, NH, QS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ncta);
    cfg.blockDim = dim3(k3_threads(NH));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B2K_CUDA(cudaLaunchKernelEx(&cfg, k3_tf32x3_topk_kernel<CG, NS, MODE, NH, QS>, mm.q, mm.ql, mm.x, mm.xl, mm.n, mm.one, p));
    count_launch();
    return B2K_OK;
}

// nh: 1 = four epilogue warps, 2 = eight (two per scheduler; screening modes with one CTA per tile)
static int launch_k3_any(int cg, int P, int mode, int nh, const K3Maps& mm, const K3Params& p, int ncta, cudaStream_t st) {
#define B2K_K3_CASE(MODE_)                                                                        \
    case MODE_:                                                                                   \
        if (cg == 1) return launch_k3<1, 2, MODE_, 1>(mm, p, ncta, st);             \
        if (P <= 512) return launch_k3<2, 3, MODE_, 1>(mm, p, ncta, st);            \
        return launch_k3<2, 2, MODE_, 1>(mm, p, ncta, st);
    switch (mode) {
        B2K_K3_CASE(0)
        B2K_K3_CASE(1)
        B2K_K3_CASE(4)
    }
#undef B2K_K3_CASE
    // screening (2 = TF32, 3 = bf16): one operand pair per stage (32 KB per CTA, or 48 KB without a
    // partner), so the rings are deeper -- as deep as the compaction scratch of the list length leaves room for
    // (6 / 5 / 4 stages for lists of 512 / 1024 / 2048 keys: the pipeline alone loses a quarter of its rate
    // between 6 and 4 stages, scripts/stage_depth.py)
#define B2K_K3_SCREEN(MODE_, NH_)                                                                  \
    if (mode == MODE_ && nh == NH_) {                                                              \
        if (cg == 1) {                                                                             \
            if (P <= 512) return launch_k3<1, 4, MODE_, NH_>(mm, p, ncta, st);       \
            return launch_k3<1, 2, MODE_, NH_>(mm, p, ncta, st);                     \
        }                                                                                          \
        if (P <= 512) return launch_k3<2, 6, MODE_, NH_>(mm, p, ncta, st);           \
        if (P <= 1024) return launch_k3<2, 5, MODE_, NH_>(mm, p, ncta, st);          \
        return launch_k3<2, 4, MODE_, NH_>(mm, p, ncta, st);                         \
    }
    B2K_K3_SCREEN(2, 1)
    B2K_K3_SCREEN(2, 2)
    B2K_K3_SCREEN(3, 1)
    B2K_K3_SCREEN(3, 2)
#undef B2K_K3_SCREEN
    set_error("tensor kernel: bad accumulator mode %d", mode);
    return B2K_EINVAL;
}

// Query-stationary kernels: CTA pairs, eight epilogue warps, lists of <= 512 keys; nsb = operand stages that fit
// next to the resident query block (6 up to d = 384 in bf16, 4 up to d = 512).
static int k3_qs_stages(int mode, int d, int P) {
    if (mode != 2 && mode != 3) return 0;
    const int bke = mode == 3 ? 64 : 32;
    const int nkc = (d + bke - 1) / bke;
    if (nkc > K3_QS_NA || P > 512) return 0;
    if (K3Smem<2, 6, 3, true>::bytes(P, nkc) <= K3_SMEM_LIMIT) return 6;
    if (K3Smem<2, 4, 3, true>::bytes(P, nkc) <= K3_SMEM_LIMIT) return 4;
    return 0;
}

static int launch_k3_qs(int mode, int nsb, const K3Maps& mm, const K3Params& p, int ncta, cudaStream_t st) {
    if (mode == 3 && nsb == 6) return launch_k3<2, 6, 3, 2, true>(mm, p, ncta, st);
    if (mode == 3 && nsb == 4) return launch_k3<2, 4, 3, 2, true>(mm, p, ncta, st);
    if (mode == 2 && nsb == 6) return launch_k3<2, 6, 2, 2, true>(mm, p, ncta, st);
    if (mode == 2 && nsb == 4) return launch_k3<2, 4, 2, 2, true>(mm, p, ncta, st);
    set_error("tensor kernel: no query-stationary variant for mode %d with %d stages", mode, nsb);
    return B2K_EINVAL;
}

static void launch_k4(int64_t nq, cudaStream_t st, const K3Params& p, int cg, int ncta, int threshold_only,
                      int list_step, float* D, int64_t* I, int set_counter = 0, u64 counter_value = 0) {
    // bound refreshes of short lists (k <= 256) use half the key buffer: 28 KB instead of 52 KB of shared memory per
    // block, so that all nq blocks of a 1024-query batch are resident at once (the kernel is latency-bound on its
    // gather); a query with more live keys than the buffer takes the sample-and-regather path, exact for any size
    if (threshold_only && p.k <= 256)
        k4_merge_kernel<2048><<<(unsigned)nq, K4_THREADS, sizeof(K4Smem<2048>), st>>>(p, cg, ncta, threshold_only, list_step, D, I,
                                                                                      set_counter, counter_value);
    else
        k4_merge_kernel<4096><<<(unsigned)nq, K4_THREADS, sizeof(K4Smem<4096>), st>>>(p, cg, ncta, threshold_only, list_step, D, I,
                                                                                      set_counter, counter_value);
}

int k3_search(K3State* s, const float* x, const float* xlo, const void* xbf, const void* xnb, const unsigned int* outl,
              const float* norms,
              const unsigned int* max_norm_bits, const unsigned int* max_err_bits, int64_t n, int d, const float* q,
              int64_t nq, int k, float* D, int64_t* I, unsigned char* scratch, cudaStream_t st, int mode,
              int* overflowed) {
    const bool screen = mode != 1;            // modes 2, 3, 4: slack lists + exact re-rank
    const bool onep = mode == 2 || mode == 3;  // one product per k-step (no residual operands)
    const bool bf16 = mode == 3;
    if (mode < 1 || mode > 4) { set_error("tensor kernel: bad mode %d", mode); return B2K_EINVAL; }
    if (bf16 && (d % 8 != 0 || !xbf || !xnb || !max_err_bits)) { set_error("tensor kernel: bf16 screening needs d %% 8 == 0 and the bf16 rows"); return B2K_EINVAL; }
    if (overflowed) *overflowed = 0;
    s->last_mode = mode;
    K3Layout L = k3_layout(s->sm_count, d, nq, k, s->cg, onep, n);
    float* qlo = reinterpret_cast<float*>(scratch + L.off_qlo);  // 3xTF32: residual rows; bf16 screening: bf16 queries
    K3Params p;
    p.n = n; p.d = d; p.nq = nq; p.k = k; p.P = L.P; p.nqb = L.nqb; p.qpc = L.qpc;
    const int64_t ntiles = (n + K3_BN - 1) / K3_BN;
    p.nunits = ntiles * L.nqb;
    p.norms = norms;
    p.qn = reinterpret_cast<float*>(scratch + L.off_qn);
    p.gts = reinterpret_cast<unsigned int*>(scratch + L.off_gts);
    p.cnt = reinterpret_cast<int*>(scratch + L.off_cnt);
    p.tau = reinterpret_cast<float*>(scratch + L.off_tau);
    p.kth = reinterpret_cast<u64*>(scratch + L.off_kth);
    p.cand = reinterpret_cast<u64*>(scratch + L.off_cand);
    p.next_unit = reinterpret_cast<u64*>(scratch + L.off_next);
    // ext_flag (caller-owned device word, never reset here): the screening check is left to the caller and the call does not synchronise
    p.overflow = s->ext_flag ? s->ext_flag : reinterpret_cast<unsigned int*>(scratch + L.off_next + 16);
    p.slack2 = reinterpret_cast<float*>(scratch + L.off_slack);
    p.static_mode = 0;
    p.static_tpc = 1;
    p.static_qb_mod = 1;
    p.static_tile0 = 0;
    p.tile_cols = K3_BN;
    p.probe = s->probe;
    p.pf_tiles = s->pf_tiles;
    // the norm-folded kernels (bf16 screen, CTA pairs) run with the outlier rows out of the screen and the capped slack
    const bool nf = bf16 && L.cg == 2 && outl != nullptr;
    p.outl = nf ? outl : nullptr;
    p.qs_t0 = p.qs_tA = p.qs_t1 = 0; p.qs_itemsA = 0; p.qs_T1 = p.qs_T2 = 1;

    // queries: norms (+ TF32 residual rows for the exact mode; K1 reused), screening slack, list state reset
    int rc = launch_row_prep(q, nq, d, const_cast<float*>(p.qn), onep ? nullptr : qlo, nullptr, st);
    if (rc) return rc;
    // Query rows beyond nq: a TMA box that leaves the tensor is zero-filled, but slowly -- a pass with
    // 16 queries took 1.53 ms against 1.08 ms with 128 (same 7.5 GB from DRAM).  The one-product modes
    // therefore read their queries from a copy padded with zero rows to whole query blocks.
    const int64_t nq_pad = (int64_t)L.nqb * K3_BM * L.cg;
    if (bf16) {
        rc = launch_row_bf16(q, nq, d, qlo, const_cast<float*>(p.slack2), nullptr, st);
        if (rc) return rc;
        if (nq_pad > nq)
            B2K_CUDA(cudaMemsetAsync(reinterpret_cast<unsigned short*>(qlo) + nq * (int64_t)d, 0, (size_t)(nq_pad - nq) * d * 2, st));
        const float acc_coef = 8e-6f * (float)d / 256.f;
        k3_slack_bf16_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, st>>>(nq, p.qn, nf ? outl + 1 : max_norm_bits,
                                                                           nf ? outl + 2 : max_err_bits, acc_coef,
                                                                           const_cast<float*>(p.slack2));
        count_launch();
        B2K_CUDA(cudaGetLastError());
    } else if (mode == 4) {
        // 3xTF32 as a screen: the dropped lo*lo term is <= 2^-22 |q||x| and every accumulate truncates
        // (<= 1 ulp of a partial sum <= |q||x|; d/8 accumulates into the main accumulator, 2d/8 small
        // ones into the cross-term accumulator): |q.x - acc| <= (2^-22 + (3d/8 + 8) 2^-23) |q||x|, twice
        // that on the distance, and the slack is twice sigma
        const float coef = 2.f * 2.f * (2.4e-7f + ((float)d * 0.375f + 8.f) * 1.2e-7f);  // 3 ulp per accumulate for margin
        k3_slack_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, st>>>(nq, p.qn, max_norm_bits, coef, const_cast<float*>(p.slack2));
        count_launch();
        B2K_CUDA(cudaGetLastError());
    } else if (screen) {
        if (nq_pad > nq) {  // padded fp32 copy of the queries (TF32 screening reads the fp32 values as they are)
            B2K_CUDA(cudaMemcpyAsync(qlo, q, (size_t)nq * d * 4, cudaMemcpyDeviceToDevice, st));
            B2K_CUDA(cudaMemsetAsync(qlo + nq * (int64_t)d, 0, (size_t)(nq_pad - nq) * d * 4, st));
        }
        const float coef = 2.f * (1.02f + 8e-6f * (float)d) / 256.f;  // 2 * sigma_q / (|q| * max|x|), see k3_slack_kernel
        k3_slack_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, st>>>(nq, p.qn, max_norm_bits, coef, const_cast<float*>(p.slack2));
        count_launch();
        B2K_CUDA(cudaGetLastError());
    }
    const int64_t nslots = (int64_t)L.ncta * L.qpc;
    const int64_t ninit = nslots > nq ? nslots : nq;
    unsigned short* ones_dev = reinterpret_cast<unsigned short*>(scratch + L.off_ones);
    k3_init_kernel<<<(unsigned)(((ninit > K3_BM * 16 ? ninit : K3_BM * 16) + 255) / 256), 256, 0, st>>>(
        nslots, p.cnt, p.tau, p.kth, nq, p.gts, p.next_unit, s->ext_flag ? nullptr : p.overflow, bf16 ? ones_dev : nullptr);
    count_launch();
    B2K_CUDA(cudaGetLastError());

    // (1024-key buffers -- 16 KB per block, every query block resident at once -- were tried for k <= 256:
    // K4 did not get faster, it was bound by its sort, not by residency, and the smaller re-rank buffer
    // made the 3xTF32 screen overflow on data it otherwise handles)
    B2K_CUDA(cudaFuncSetAttribute(k4_merge_kernel<4096>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K4Smem<4096>)));
    CUtensorMap mq, mql, mx, mxl, mn, mone;
    if (bf16) {
        if ((rc = make_map(s, &mq, qlo, nq_pad, d, K3_BM, true))) return rc;
        if ((rc = make_map(s, &mx, xbf, n, d, K3_BN / L.cg, true))) return rc;
        mql = mq;
        mxl = mx;
        if (L.cg == 2) {  // the norm fold's operands: 32-byte records of the rows, and the block of ones
            if ((rc = make_map(s, &mn, xnb, n, 16, K3_BN / L.cg, true, true))) return rc;
            if ((rc = make_map(s, &mone, ones_dev, K3_BM, 16, K3_BM, true, true))) return rc;
        }
    } else if (mode == 2 && nq_pad > nq) {
        if ((rc = make_map(s, &mq, qlo, nq_pad, d, K3_BM))) return rc;
        if ((rc = make_map(s, &mx, x, n, d, K3_BN / L.cg))) return rc;
        mql = mq;
        mxl = mx;
    } else {
        if ((rc = make_map(s, &mq, q, nq, d, K3_BM))) return rc;
        if ((rc = make_map(s, &mql, onep ? q : qlo, nq, d, K3_BM))) return rc;
        if ((rc = make_map(s, &mx, x, n, d, K3_BN / L.cg))) return rc;
        if ((rc = make_map(s, &mxl, onep ? x : xlo, n, d, K3_BN / L.cg))) return rc;
    }

    if (!(bf16 && L.cg == 2)) { mn = mx; mone = mq; }  // unused by the other kernels
    const K3Maps mm{mq, mql, mx, mxl, mn, mone};

    // Warm start: the first 64 rows of one tile per CTA (pair) with open thresholds -- every key is
    // kept, so the pass is kept small; per query the k-th smallest distance of that sample becomes
    // the shared threshold (an upper bound of the final k-th distance).
    const int nclusters = L.ncta / L.cg;
    // kernel mode: 0 3xTF32 in one accumulator (double-buffered), 1 3xTF32 with a separate cross-term
    // accumulator, 2 TF32 screening, 3 bf16 screening (one product; exact distances from the re-rank kernel)
    const int sep = bf16 ? 3 : mode == 2 ? 2 : mode == 4 ? 4 : (s->sep_mode == 1 || (s->sep_mode == 0 && d >= 512)) ? 1 : 0;
    const int nh = s->epi_halves == 1 ? 1 : 2;  // eight epilogue warps where instantiated (screening modes)
    int cols_a = 64;
    int64_t warm_rows = 0;  // rows behind the bound the main pass starts with
    if (ntiles >= 2 * (int64_t)nclusters && (int64_t)nclusters * cols_a >= 2 * (int64_t)k) {
        K3Params pre = p;
        pre.static_mode = 1;
        pre.static_tpc = 1;
        pre.static_tile0 = 0;
        pre.tile_cols = cols_a;
        // each CTA (pair) only serves every G-th query block, leaving nclusters / G lists of 64 keys per query
        const int64_t wk = s->warm_keys;  // sample size per query, in multiples of k
        int G = (int)(((int64_t)nclusters * cols_a) / (wk * (int64_t)k));
        G = G < 1 ? 1 : (G > L.nqb ? L.nqb : G);
        // ... and of fewer than 64 columns when that is still more than the ~12k keys per query a usable
        // bound needs (few query blocks, small k): the pass keeps every key, so it costs what it keeps
        int want_cols = (int)((wk * (int64_t)k * G + nclusters - 1) / nclusters);
        want_cols = (want_cols + 7) & ~7;
        cols_a = want_cols < 8 ? 8 : (want_cols > 64 ? 64 : want_cols);
        pre.tile_cols = cols_a;
        pre.static_qb_mod = G;
        pre.nunits = (int64_t)nclusters * L.nqb;
        pre.n = (int64_t)nclusters * K3_BN;
        rc = launch_k3_any(L.cg, L.P, sep, nh, mm, pre, L.ncta, st);
        if (rc) return rc;
        launch_k4(nq, st, pre, L.cg, L.ncta, 1, 1, nullptr, nullptr);
        k3_init_kernel<<<(unsigned)((nslots + 255) / 256), 256, 0, st>>>(nslots, p.cnt, p.tau, p.kth, nq, nullptr, p.next_unit, nullptr, nullptr);
        count_launch(2);
        B2K_CUDA(cudaGetLastError());
        warm_rows = (int64_t)nclusters * cols_a / G;
    }

    // Main pass, in segments.  A bound that is the k-th distance of a sample of m rows lets about
    // k * n / m rows of the whole database through -- and inside one launch the bound barely moves,
    // because a list only learns a better one when it compacts.  So the pass is cut where the
    // scanned fraction has grown by a factor r, and between two segments K4 (threshold mode, ~50 us)
    // turns everything kept so far into the exact k-th distance of the rows scanned so far.  Each
    // segment then lets about k * (r - 1) keys per query through; r is chosen for max(400, 2k) keys
    // (measured best among 200..3000 on four shapes; without any refresh C3 takes 55 ms instead of 6.7).  The kept
    // keys stay in the lists across segments; the rows of every segment are scanned exactly once.
    long long seg_end[8];
    int nseg = 0;
    if (warm_rows > 0 && s->segments) {
        // one query block: the pass is bound by the row stream, the epilogue idles, survivors are nearly
        // free and every boundary (drain, K4, refill) is pure loss: 8x fewer, coarser steps
        const double target = s->segments > 1 ? (double)s->segments : (L.nqb == 1 ? 3200.0 : 400.0);
        const double r = 1.0 + (target > 2.0 * k ? target : 2.0 * k) / (double)k;
        const long long min_step = 4LL * nclusters;  // units: a shorter segment is all drain and refill
        double g = (double)warm_rows / (double)n;     // fraction of the rows behind the current bound
        long long prev = 0;
        while (nseg < 6) {
            g *= r;
            if (g >= 0.6) break;
            long long end = (long long)(g * (double)ntiles) * L.nqb;
            // a boundary that would come too early is moved out, never dropped: dropping it leaves the next
            // segment r times more permissive (measured: 2.0 instead of 0.94 ms on a 1.25M-row shard)
            if (end < prev + min_step) end = prev + min_step;
            end = (end + L.nqb - 1) / L.nqb * L.nqb;  // whole tiles
            if (end + min_step > p.nunits) break;
            seg_end[nseg++] = end;
            prev = end;
            g = (double)end / (double)p.nunits;       // what has really been scanned by then
        }
    }
    seg_end[nseg++] = p.nunits;
    // Query-stationary schedule (one-product modes, CTA pairs, d small enough for a CTA's 128 query rows to
    // stay in shared memory): a pair keeps its query block resident over a run of consecutive tiles and only
    // streams the rows, which halves the operand traffic from L2 -- the stated bound of the streaming form.
    const int qs_nsb = (s->qs && L.cg == 2 && nh == 2 && true) ? k3_qs_stages(sep, d, L.P) : 0;
    const bool qs = qs_nsb > 0 && L.nqb >= 2;
    if (s->timing) B2K_CUDA(cudaEventRecord(s->ev0, st));
    for (int sg = 0; sg < nseg; ++sg) {
        K3Params ps = p;
        ps.nunits = seg_end[sg];
        const long long u0 = sg > 0 ? seg_end[sg - 1] : 0;
        if (qs) {
            // segment boundaries are whole tiles (seg_end is a multiple of nqb); the counter restarts at 0
            ps.qs_t0 = u0 / L.nqb;
            ps.qs_t1 = seg_end[sg] / L.nqb;
            const long long nt = ps.qs_t1 - ps.qs_t0;
            const long long groups = nclusters / L.nqb > 0 ? nclusters / L.nqb : 1;  // pairs per query block
            long long T1 = nt / (groups * 4);                                       // >= 4 items per pair and segment
            T1 = T1 < 1 ? 1 : (T1 > 16 ? 16 : T1);
            long long T2 = T1 / 4 < 1 ? 1 : T1 / 4;                                  // the last eighth in short runs: a balanced tail
            long long nA = ((nt - nt / 8) / T1) * T1;
            ps.qs_T1 = (int)T1; ps.qs_T2 = (int)T2;
            ps.qs_tA = ps.qs_t0 + nA;
            ps.qs_itemsA = (nA / T1) * L.nqb;
        }
        if (sg > 0) {
            launch_k4(nq, st, p, L.cg, L.ncta, 1, 1, nullptr, nullptr, 1, qs ? 0ull : (u64)u0);
            count_launch();
            B2K_CUDA(cudaGetLastError());
        }
        rc = qs ? launch_k3_qs(sep, qs_nsb, mm, ps, L.ncta, st)
                : launch_k3_any(L.cg, L.P, sep, nh, mm, ps, L.ncta, st);
        if (rc) return rc;
    }
    if (s->timing) { B2K_CUDA(cudaEventRecord(s->ev1, st)); s->timed = true; }

    if (!screen) {
        launch_k4(nq, st, p, L.cg, L.ncta, 0, 1, D, I);
        count_launch();
        B2K_CUDA(cudaGetLastError());
        return B2K_OK;
    }
    // screening: the exact k-th smallest approximate distance per query, then the exact re-rank of
    // everything within the slack of it
    B2K_CUDA(cudaFuncSetAttribute(k6_rerank_kernel<4096>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K4Smem<4096>)));
    // with a refreshed bound from late in the pass (segments) and short lists, the last refresh happens inside the re-rank kernel
    // (and few enough keys under that bound: ~k / (fraction scanned at the last refresh), plus the slack zone)
    const double g_last = nseg >= 2 ? (double)seg_end[nseg - 2] / (double)p.nunits : 0.0;
    const int fold = (nseg >= 2 && k <= 256 && (double)k <= 1500.0 * g_last) ? 1 : 0;
    if (!fold) { launch_k4(nq, st, p, L.cg, L.ncta, 1, 1, nullptr, nullptr); count_launch(); }
    k6_rerank_kernel<4096><<<(unsigned)nq, K4_THREADS, sizeof(K4Smem<4096>), st>>>(p, L.cg, L.ncta, x, q, D, I, fold);
    count_launch();
    B2K_CUDA(cudaGetLastError());
    // a list or the re-rank buffer could not hold every row inside the slack (masses of near-
    // duplicates around the k-th neighbour): the caller repeats the search in the exact mode
    if (s->ext_flag) return B2K_OK;
    B2K_CUDA(cudaMemcpyAsync(s->flag_host, p.overflow, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
    B2K_CUDA(cudaStreamSynchronize(st));
    if (overflowed) *overflowed = (int)*s->flag_host;
    return B2K_OK;
}

}  // namespace b2k
