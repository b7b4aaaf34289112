/*
 * oracle/flat_l2_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C CPU restatement of what `faiss.IndexFlatL2` does for the calls made by
 * the reference's retrieval path:
 *     /root/reference/llm/rag.py:61,71   faiss.IndexFlatL2(dimension)
 *     /root/reference/llm/rag.py:80      index.add(vectors_np)
 *     /root/reference/llm/rag.py:116     D, I = index.search(q.reshape(1,-1), k)
 *     /root/reference/llm/rag.py:62,82   faiss.write_index     (see oracle.py: IxF2)
 *     /root/reference/llm/rag.py:74,153  faiss.read_index      (see oracle.py: IxF2)
 *
 * PARITY UNPINNED: the arithmetic of that path lives in the third-party wheel
 * `faiss-cpu` (unpinned in /root/reference/llm/requirements_rag.txt:4); its source
 * is not under /root/reference, the wheel is not installed and cannot be installed
 * here (no network), and the reference has no tests, golden vectors or fixtures
 * (SURVEY.md section 4, section 8c).  This file therefore restates faiss's PUBLISHED
 * algorithm (upstream faiss/utils/distances.cpp, faiss/impl/ResultHandler.h,
 * faiss/utils/Heap.h, as recorded in SURVEY.md section 8c):
 *
 *   (1) nq <  20 : D[i][j] = sum_t (q[i][t] - x[j][t])^2 in fp32      (exhaustive_L2sqr_seq)
 *   (2) nq >= 20 : D[i][j] = ||q_i||^2 + ||x_j||^2 - 2 <q_i,x_j>, fp32,
 *                  negative results clamped to 0                         (exhaustive_L2sqr_blas)
 *   (3) a candidate replaces the current worst only if it is strictly better
 *       (an equal later candidate does not displace an earlier one); rows are
 *       visited in increasing id, so among equal distances the lower ids stay;
 *   (4) output rows sorted by (distance, id) ascending;
 *   (5) if fewer than k rows qualify, the tail is I=-1, D=FLT_MAX (3.4028235e38);
 *       a row whose distance is not < FLT_MAX (inf, NaN) is never selected.
 *
 * (3)+(4) together are exactly "the k smallest (distance, id) pairs in
 * lexicographic order", which is what the bounded heap below keeps.  faiss's
 * reservoir handler (k >= 100) leaves boundary ties unspecified; the lexicographic
 * rule is one of its legal outcomes and is the one checked here.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library -- always as the checker or the timed CPU
 * baseline, never as part of the product path.
 *
 * Build: see oracle/build.py  (gcc -O3 -fopenmp -shared -fPIC; target_clones gives
 * AVX-512 / AVX2 / baseline variants so the .so built in one container runs on
 * another host CPU).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))
#if defined(__x86_64__)
#define ORC_CLONES __attribute__((target_clones("arch=x86-64-v4", "arch=x86-64-v3", "default")))
#else
#define ORC_CLONES
#endif

#define ORC_BLAS_THRESHOLD 20 /* faiss distance_compute_blas_threshold */
#define ORC_DB_BS 1024        /* faiss distance_compute_blas_database_bs */
#define ORC_LANES 16

/* ------------------------------------------------------------------------- */
/* Synthetic data: counter-based, integer-only up to one exact fp32 multiply, so
 * the CUDA generator (csrc/synth.cu) produces bit-identical values.           */

static inline uint64_t orc_mix64(uint64_t z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

static inline float orc_synth_value(uint64_t seed, uint64_t row, uint64_t col) {
    uint64_t hr = orc_mix64(seed * 0x9E3779B97F4A7C15ull + row + 0x632BE59BD9B4E019ull);
    uint64_t h = orc_mix64(hr + col * 0xD1342543DE82EF95ull);
    int32_t s = (int32_t)(h & 0xFFFF) + (int32_t)((h >> 16) & 0xFFFF) +
                (int32_t)((h >> 32) & 0xFFFF) + (int32_t)(h >> 48) - 131070;
    /* sum of four U{0..65535}: variance 4*(65536^2-1)/12; scale to unit variance */
    return (float)s * 2.6429e-5f;
}

ORC_API void orc_synth(uint64_t seed, int64_t row0, int64_t nrows, int d, float* out) {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < nrows; ++r)
        for (int c = 0; c < d; ++c)
            out[r * (int64_t)d + c] = orc_synth_value(seed, (uint64_t)(row0 + r), (uint64_t)c);
}

/* ------------------------------------------------------------------------- */
/* Bounded max-heap on (distance, id), lexicographic.                          */

typedef struct { float dis; int64_t id; } orc_pair;

static inline int orc_less(float da, int64_t ia, float db, int64_t ib) {
    return da < db || (da == db && ia < ib);
}

typedef struct { orc_pair* h; int64_t k; int64_t n; } orc_heap;

static inline void orc_heap_push(orc_heap* hp, float dis, int64_t id) {
    if (!(dis < FLT_MAX)) return; /* rule (5): inf / NaN / FLT_MAX never enter */
    orc_pair* h = hp->h;
    if (hp->n < hp->k) { /* sift up */
        int64_t i = hp->n++;
        while (i > 0) {
            int64_t p = (i - 1) >> 1;
            if (!orc_less(h[p].dis, h[p].id, dis, id)) break;
            h[i] = h[p]; i = p;
        }
        h[i].dis = dis; h[i].id = id;
        return;
    }
    /* full: replace the worst only if strictly better (rule 3) */
    if (!orc_less(dis, id, h[0].dis, h[0].id)) return;
    int64_t i = 0, n = hp->n;
    for (;;) {
        int64_t l = 2 * i + 1, r = l + 1, m;
        if (l >= n) break;
        m = (r < n && orc_less(h[l].dis, h[l].id, h[r].dis, h[r].id)) ? r : l;
        if (!orc_less(dis, id, h[m].dis, h[m].id)) break;
        h[i] = h[m]; i = m;
    }
    h[i].dis = dis; h[i].id = id;
}

static int orc_pair_cmp(const void* a, const void* b) {
    const orc_pair* pa = (const orc_pair*)a; const orc_pair* pb = (const orc_pair*)b;
    if (orc_less(pa->dis, pa->id, pb->dis, pb->id)) return -1;
    if (orc_less(pb->dis, pb->id, pa->dis, pa->id)) return 1;
    return 0;
}

static void orc_heap_emit(orc_heap* hp, float* D, int64_t* I) {
    qsort(hp->h, (size_t)hp->n, sizeof(orc_pair), orc_pair_cmp);
    for (int64_t j = 0; j < hp->k; ++j) {
        if (j < hp->n) { D[j] = hp->h[j].dis; I[j] = hp->h[j].id; }
        else { D[j] = FLT_MAX; I[j] = -1; }
    }
}

/* ------------------------------------------------------------------------- */
/* Arithmetic kernels.  Summation order is fixed by the source (ORC_LANES
 * partial sums, then a fixed-order horizontal add) so results do not depend on
 * compiler reassociation.                                                     */

ORC_CLONES
static float orc_l2sqr(const float* a, const float* b, int d) {
    float acc[ORC_LANES];
    for (int l = 0; l < ORC_LANES; ++l) acc[l] = 0.f;
    int t = 0;
    for (; t + ORC_LANES <= d; t += ORC_LANES)
        for (int l = 0; l < ORC_LANES; ++l) { float df = a[t + l] - b[t + l]; acc[l] += df * df; }
    for (int l = 0; t < d; ++t, ++l) { float df = a[t] - b[t]; acc[l] += df * df; }
    float s = 0.f;
    for (int l = 0; l < ORC_LANES; ++l) s += acc[l];
    return s;
}

ORC_CLONES
static float orc_norm2(const float* a, int d) {
    float acc[ORC_LANES];
    for (int l = 0; l < ORC_LANES; ++l) acc[l] = 0.f;
    int t = 0;
    for (; t + ORC_LANES <= d; t += ORC_LANES)
        for (int l = 0; l < ORC_LANES; ++l) acc[l] += a[t + l] * a[t + l];
    for (int l = 0; t < d; ++t, ++l) acc[l] += a[t] * a[t];
    float s = 0.f;
    for (int l = 0; l < ORC_LANES; ++l) s += acc[l];
    return s;
}

/* 4 queries x 2 database rows of inner products per call: the register-blocked
 * stand-in for the sgemm_ call in exhaustive_L2sqr_blas.                       */
ORC_CLONES
static void orc_ip_4x2(const float* q0, const float* q1, const float* q2, const float* q3,
                       const float* x0, const float* x1, int d, float* out /* [4][2] */) {
    float a[8][ORC_LANES];
    for (int i = 0; i < 8; ++i) for (int l = 0; l < ORC_LANES; ++l) a[i][l] = 0.f;
    int t = 0;
    for (; t + ORC_LANES <= d; t += ORC_LANES) {
        for (int l = 0; l < ORC_LANES; ++l) {
            float xa = x0[t + l], xb = x1[t + l];
            a[0][l] += q0[t + l] * xa; a[1][l] += q0[t + l] * xb;
            a[2][l] += q1[t + l] * xa; a[3][l] += q1[t + l] * xb;
            a[4][l] += q2[t + l] * xa; a[5][l] += q2[t + l] * xb;
            a[6][l] += q3[t + l] * xa; a[7][l] += q3[t + l] * xb;
        }
    }
    for (int l = 0; t < d; ++t, ++l) {
        float xa = x0[t], xb = x1[t];
        a[0][l] += q0[t] * xa; a[1][l] += q0[t] * xb;
        a[2][l] += q1[t] * xa; a[3][l] += q1[t] * xb;
        a[4][l] += q2[t] * xa; a[5][l] += q2[t] * xb;
        a[6][l] += q3[t] * xa; a[7][l] += q3[t] * xb;
    }
    for (int i = 0; i < 8; ++i) {
        float s = 0.f;
        for (int l = 0; l < ORC_LANES; ++l) s += a[i][l];
        out[i] = s;
    }
}

/* ------------------------------------------------------------------------- */

/* nq < 20: direct differences, one thread per query (what faiss does: OpenMP is
 * over queries, so nq == 1 runs on one core).  With row_parallel != 0 the scan
 * is split over rows instead and the per-thread heaps are merged -- a fairer CPU
 * number for the nq == 1 baseline; the result is identical.                    */
static void orc_search_seq(const float* x, int64_t n, int d, const float* q, int64_t nq,
                           int64_t k, float* D, int64_t* I, int row_parallel) {
    if (!row_parallel) {
#pragma omp parallel for schedule(dynamic, 1)
        for (int64_t i = 0; i < nq; ++i) {
            orc_heap hp; hp.h = (orc_pair*)malloc(sizeof(orc_pair) * (size_t)k); hp.k = k; hp.n = 0;
            const float* qi = q + i * (int64_t)d;
            for (int64_t j = 0; j < n; ++j) orc_heap_push(&hp, orc_l2sqr(qi, x + j * (int64_t)d, d), j);
            orc_heap_emit(&hp, D + i * k, I + i * k);
            free(hp.h);
        }
        return;
    }
    int nt = 1;
#ifdef _OPENMP
    nt = omp_get_max_threads();
#endif
    for (int64_t i = 0; i < nq; ++i) {
        const float* qi = q + i * (int64_t)d;
        orc_pair* all = (orc_pair*)malloc(sizeof(orc_pair) * (size_t)k * (size_t)nt);
        int64_t* cnt = (int64_t*)calloc((size_t)nt, sizeof(int64_t));
#pragma omp parallel num_threads(nt)
        {
            int tid = 0;
#ifdef _OPENMP
            tid = omp_get_thread_num();
#endif
            int64_t lo = n * tid / nt, hi = n * (tid + 1) / nt;
            orc_heap hp; hp.h = all + (int64_t)tid * k; hp.k = k; hp.n = 0;
            for (int64_t j = lo; j < hi; ++j) orc_heap_push(&hp, orc_l2sqr(qi, x + j * (int64_t)d, d), j);
            cnt[tid] = hp.n;
        }
        orc_heap hp; hp.h = (orc_pair*)malloc(sizeof(orc_pair) * (size_t)k); hp.k = k; hp.n = 0;
        for (int t = 0; t < nt; ++t)
            for (int64_t j = 0; j < cnt[t]; ++j) orc_heap_push(&hp, all[(int64_t)t * k + j].dis, all[(int64_t)t * k + j].id);
        orc_heap_emit(&hp, D + i * k, I + i * k);
        free(hp.h); free(all); free(cnt);
    }
}

/* nq >= 20: norms + inner products + clamp, database walked in blocks of 1024
 * rows (each block is reused from cache by every query), queries in groups of 4
 * spread over threads.                                                        */
static void orc_search_expand(const float* x, int64_t n, int d, const float* q, int64_t nq,
                              int64_t k, float* D, int64_t* I) {
    float* qn = (float*)malloc(sizeof(float) * (size_t)nq);
    float* xn = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
    orc_pair* heaps = (orc_pair*)malloc(sizeof(orc_pair) * (size_t)k * (size_t)nq);
    int64_t* hn = (int64_t*)calloc((size_t)nq, sizeof(int64_t));
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nq; ++i) qn[i] = orc_norm2(q + i * (int64_t)d, d);
#pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < n; ++j) xn[j] = orc_norm2(x + j * (int64_t)d, d);

    int64_t ngroups = (nq + 3) / 4;
    for (int64_t j0 = 0; j0 < n; j0 += ORC_DB_BS) {
        int64_t j1 = j0 + ORC_DB_BS < n ? j0 + ORC_DB_BS : n;
#pragma omp parallel for schedule(dynamic, 4)
        for (int64_t g = 0; g < ngroups; ++g) {
            int64_t i0 = g * 4;
            const float* qp[4]; int64_t qi[4];
            for (int a = 0; a < 4; ++a) { qi[a] = i0 + a < nq ? i0 + a : nq - 1; qp[a] = q + qi[a] * (int64_t)d; }
            int nvalid = (int)(nq - i0 < 4 ? nq - i0 : 4);
            orc_heap hp[4];
            for (int a = 0; a < nvalid; ++a) { hp[a].h = heaps + qi[a] * k; hp[a].k = k; hp[a].n = hn[qi[a]]; }
            for (int64_t j = j0; j < j1; j += 2) {
                int64_t jb = j + 1 < j1 ? j + 1 : j;
                float ip[8];
                orc_ip_4x2(qp[0], qp[1], qp[2], qp[3], x + j * (int64_t)d, x + jb * (int64_t)d, d, ip);
                for (int a = 0; a < nvalid; ++a) {
                    float d0 = qn[qi[a]] + xn[j] - 2.f * ip[a * 2 + 0];
                    if (d0 < 0.f) d0 = 0.f;
                    orc_heap_push(&hp[a], d0, j);
                    if (jb != j) {
                        float d1 = qn[qi[a]] + xn[jb] - 2.f * ip[a * 2 + 1];
                        if (d1 < 0.f) d1 = 0.f;
                        orc_heap_push(&hp[a], d1, jb);
                    }
                }
            }
            for (int a = 0; a < nvalid; ++a) hn[qi[a]] = hp[a].n;
        }
    }
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nq; ++i) {
        orc_heap hp; hp.h = heaps + i * k; hp.k = k; hp.n = hn[i];
        orc_heap_emit(&hp, D + i * k, I + i * k);
    }
    free(qn); free(xn); free(heaps); free(hn);
}

/* mode: 0 = faiss dispatch (nq < 20 -> direct, else expansion),
 *       1 = force direct, 2 = force expansion,
 *       3 = direct with the nq-loop serial and the row scan parallel.          */
ORC_API int orc_search(const float* x, int64_t n, int d, const float* q, int64_t nq, int64_t k,
                       float* D, int64_t* I, int mode) {
    if (d <= 0 || k <= 0 || nq < 0 || n < 0) return -1;
    if (nq == 0) return 0;
    if (mode == 1 || (mode == 0 && nq < ORC_BLAS_THRESHOLD)) orc_search_seq(x, n, d, q, nq, k, D, I, 0);
    else if (mode == 3) orc_search_seq(x, n, d, q, nq, k, D, I, 1);
    else orc_search_expand(x, n, d, q, nq, k, D, I);
    return 0;
}

/* The combine step of the row-sharded search (no faiss counterpart on the
 * reference's single process; it restates "concatenate the shards' lists, keep
 * the k smallest (distance, id)").  Din/Iin are [nlists][nq][k].               */
ORC_API int orc_merge(int nlists, int64_t nq, int64_t k, const float* Din, const int64_t* Iin,
                      float* D, int64_t* I) {
    for (int64_t i = 0; i < nq; ++i) {
        orc_heap hp; hp.h = (orc_pair*)malloc(sizeof(orc_pair) * (size_t)k); hp.k = k; hp.n = 0;
        for (int l = 0; l < nlists; ++l)
            for (int64_t j = 0; j < k; ++j) {
                int64_t o = ((int64_t)l * nq + i) * k + j;
                if (Iin[o] >= 0) orc_heap_push(&hp, Din[o], Iin[o]);
            }
        orc_heap_emit(&hp, D + i * k, I + i * k);
        free(hp.h);
    }
    return 0;
}

/* ---- BLAS-backed variant of rule (2) -----------------------------------------------------
 * faiss hands the inner-product block to sgemm_ (exhaustive_L2sqr_blas).  For the timed CPU
 * baseline the caller (oracle.py: search_blas) computes each block <Q, X_blk> with the best
 * BLAS present (MKL through torch.mm) and this code does what faiss does around the sgemm:
 * norms, dis = qn + xn - 2 ip, clamp, result-handler update, final reorder.               */
typedef struct { int64_t nq, k; orc_pair* heaps; int64_t* hn; } orc_state;

ORC_API void* orc_blas_begin(int64_t nq, int64_t k) {
    orc_state* st = (orc_state*)malloc(sizeof(orc_state));
    st->nq = nq; st->k = k;
    st->heaps = (orc_pair*)malloc(sizeof(orc_pair) * (size_t)k * (size_t)nq);
    st->hn = (int64_t*)calloc((size_t)nq, sizeof(int64_t));
    return st;
}

ORC_API void orc_norms(const float* x, int64_t n, int d, float* out) {
#pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < n; ++j) out[j] = orc_norm2(x + j * (int64_t)d, d);
}

/* ip is [nq][nb] row-major: inner products of every query with database rows j0 .. j0+nb-1 */
ORC_API void orc_blas_update(void* vst, int64_t nb, int64_t j0, const float* ip, const float* qn, const float* xn) {
    orc_state* st = (orc_state*)vst;
#pragma omp parallel for schedule(dynamic, 8)
    for (int64_t i = 0; i < st->nq; ++i) {
        orc_heap hp; hp.h = st->heaps + i * st->k; hp.k = st->k; hp.n = st->hn[i];
        const float* row = ip + i * nb;
        float qi = qn[i];
        /* fast reject against the current worst once the heap is full (what the heap handler's
         * threshold test does), exact (distance, id) test inside orc_heap_push */
        for (int64_t j = 0; j < nb; ++j) {
            float dv = qi + xn[j] - 2.f * row[j];
            if (dv < 0.f) dv = 0.f;
            if (hp.n == hp.k && dv > hp.h[0].dis) continue;
            orc_heap_push(&hp, dv, j0 + j);
        }
        st->hn[i] = hp.n;
    }
}

ORC_API void orc_blas_end(void* vst, float* D, int64_t* I) {
    orc_state* st = (orc_state*)vst;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < st->nq; ++i) {
        orc_heap hp; hp.h = st->heaps + i * st->k; hp.k = st->k; hp.n = st->hn[i];
        orc_heap_emit(&hp, D + i * st->k, I + i * st->k);
    }
    free(st->heaps); free(st->hn); free(st);
}

ORC_API int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU baseline and the parity
 * checker put the host cores back to work with this. */
ORC_API void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
