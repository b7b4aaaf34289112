// index.cu -- the C ABI of libb200knn.so (declared in include/b200knn.h): index lifecycle,
// row storage in HBM, IxF2 persistence and the search dispatch of the single-GPU handle; handles
// that span several GPUs are forwarded to multi.cu.
//
// Mirrors the `faiss.IndexFlatL2` surface used by /root/reference/llm/rag.py
// (:61/:71 construct, :80 add, :116 search, :86/:190-191 ntotal/d, :62/:82 write_index,
// :74/:153/:186 read_index).  HBM layout per row store (index_internal.h):
//     x     [capacity][d] float32  row-major, the rows as given            (K2, K3 "hi" operand)
//     norms [capacity]    float32  ||x||^2, written by K1 at add()         (K3 epilogue)
//     xlo   [capacity][d] float32  TF32 residual rows, only once a 3xTF32  (K3 "lo" operand)
//                                  search has been made on this store
//     xbf   [capacity][d] bf16     copies for the screening pass / bf16 scan, once one has run
// and per handle
//     ws    scratch: device copies of queries / results for host callers, per-CTA partial
//           top-k lists, K3 candidate lists.
#include <errno.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <mutex>
#include <new>
#include <string>

#include "common.cuh"
#include "kernels.h"
#include "k3_api.h"
#include "index_internal.h"

namespace b2k {

static thread_local std::string t_err;
std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    t_err = buf;
}

RowStore::~RowStore() {
    DeviceGuard g(device);
    cudaDeviceSynchronize();  // kernels of handles that shared this store may still be reading it
    cudaFree(x); cudaFree(norms); cudaFree(xlo); cudaFree(xbf); cudaFree(xnb); cudaFree(stat); cudaFree(outl); cudaFree(outl_hist);
    if (ev_rows) cudaEventDestroy(ev_rows);
}

static constexpr int64_t MAX_ROWS = 0xFFFFFFFEll;  // row index lives in the low 32 bits of a key
static constexpr size_t IO_CHUNK = (size_t)64 << 20;
static constexpr int SCAN_FLAG_SLOT = 4;  // h->done: [0] scan ticket, [1] scan work counter, [4] raised by the bf16 scan's re-rank when its candidates did not cover the slack zone
static constexpr int MAX_NORM_SLOT = 0;   // rs->stat: [0] bits of the largest squared row norm
static constexpr int MAX_ERR_SLOT = 1;    //           [1] bits of the largest squared bf16 rounding-error norm of a row
static constexpr int64_t TC_QUERY_CHUNK = 4096;  // queries per tensor-core pass (faiss: distance_compute_blas_query_bs)

// Validated devices are cached: cudaGetDeviceProperties costs about a millisecond, far too much
// for per-call use (the merge entry point is on the sharded search's critical path).
int check_device(int device, int* sm_count) {
    static std::mutex mu;
    static int cached_sm[64] = {0};
    if (device >= 0 && device < 64) {
        std::lock_guard<std::mutex> lk(mu);
        if (cached_sm[device] > 0) { if (sm_count) *sm_count = cached_sm[device]; return B2K_OK; }
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        set_error("no usable CUDA device (%s); libb200knn has no CPU fallback",
                  e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        return B2K_ECUDA;
    }
    if (device < 0 || device >= count) {
        set_error("device %d out of range (0..%d)", device, count - 1);
        return B2K_EINVAL;
    }
    int major = 0, minor = 0, sms = 0;
    B2K_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    B2K_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
    B2K_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    if (major != 10) {
        set_error("device %d is sm_%d%d; libb200knn is built for sm_100a (B200) only", device, major, minor);
        return B2K_ECUDA;
    }
    if (device < 64) { std::lock_guard<std::mutex> lk(mu); cached_sm[device] = sms; }
    if (sm_count) *sm_count = sms;
    return B2K_OK;
}

// true if p is page-locked host memory known to CUDA (cudaMallocHost / cudaHostRegister / torch pin_memory)
bool is_pinned_host(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

namespace {

// Spare buffers.  llm/rag.py builds a fresh index object per request (read_index, rag.py:153), i.e. a fresh view
// here, and a view's first search needs a workspace and a pinned staging buffer: cudaMalloc + cudaMallocHost cost
// ~2.5 ms, ten times the search itself at the reference's sizes.  A destroyed handle therefore leaves its (small)
// buffers in a per-device pool and the next handle picks them up.
struct Spare { void* p; size_t bytes; };
struct SparePool {
    std::mutex mu;
    std::vector<Spare> ws[64], pin[64];
};
SparePool g_spares;
constexpr size_t SPARE_MAX_BYTES = (size_t)64 << 20;  // larger buffers are not hoarded
constexpr size_t SPARE_MAX_COUNT = 8;

bool take_spare(std::vector<Spare>& v, size_t need, Spare* out) {
    std::lock_guard<std::mutex> lk(g_spares.mu);
    int best = -1;
    for (int i = 0; i < (int)v.size(); ++i)
        if (v[i].bytes >= need && (best < 0 || v[i].bytes < v[best].bytes)) best = i;
    if (best < 0) return false;
    *out = v[best];
    v.erase(v.begin() + best);
    return true;
}

// true if the pool took the buffer (the caller must have made sure no work still uses it)
bool give_spare(std::vector<Spare>& v, void* p, size_t bytes) {
    if (!p || bytes > SPARE_MAX_BYTES) return false;
    std::lock_guard<std::mutex> lk(g_spares.mu);
    if (v.size() >= SPARE_MAX_COUNT) return false;
    v.push_back(Spare{p, bytes});
    return true;
}

int ensure_ws(b2k_index* h, size_t bytes) {
    if (bytes <= h->ws_bytes) return B2K_OK;
    if (h->ws) {
        B2K_CUDA(cudaEventSynchronize(h->ev_ws));  // the last user of the old workspace is done
        if (!give_spare(g_spares.ws[h->device & 63], h->ws, h->ws_bytes)) B2K_CUDA(cudaFree(h->ws));
        h->ws = nullptr; h->ws_bytes = 0;
    }
    Spare sp;
    if (take_spare(g_spares.ws[h->device & 63], bytes, &sp)) { h->ws = sp.p; h->ws_bytes = sp.bytes; return B2K_OK; }
    size_t want = bytes + bytes / 4 + 4096;
    cudaError_t e = cudaMalloc(&h->ws, want);
    if (e != cudaSuccess) { cudaGetLastError(); set_error("workspace cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e)); return B2K_ENOMEM; }
    h->ws_bytes = want;
    return B2K_OK;
}

int ensure_pin(b2k_index* h, size_t bytes) {
    if (bytes <= h->pin_bytes) return B2K_OK;
    if (h->pin) {
        B2K_CUDA(cudaStreamSynchronize(h->stream));
        if (!give_spare(g_spares.pin[h->device & 63], h->pin, h->pin_bytes)) B2K_CUDA(cudaFreeHost(h->pin));
        h->pin = nullptr; h->pin_bytes = 0;
    }
    Spare sp;
    if (take_spare(g_spares.pin[h->device & 63], bytes, &sp)) { h->pin = sp.p; h->pin_bytes = sp.bytes; return B2K_OK; }
    size_t want = bytes + bytes / 4 + 4096;
    cudaError_t e = cudaMallocHost(&h->pin, want);
    if (e != cudaSuccess) { cudaGetLastError(); set_error("cudaMallocHost(%zu) failed: %s", want, cudaGetErrorString(e)); return B2K_ENOMEM; }
    h->pin_bytes = want;
    return B2K_OK;
}

int new_store(int d, int device, std::shared_ptr<RowStore>* out) {
    std::shared_ptr<RowStore> rs(new (std::nothrow) RowStore());
    if (!rs) { set_error("out of host memory"); return B2K_ENOMEM; }
    rs->d = d; rs->device = device;
    cudaError_t e = cudaMalloc(&rs->stat, 256);
    if (e == cudaSuccess) e = cudaMemset(rs->stat, 0, 256);
    if (e == cudaSuccess) e = cudaMalloc(&rs->outl, B2K_OUTL_WORDS * sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaMemset(rs->outl, 0, B2K_OUTL_WORDS * sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaMalloc(&rs->outl_hist, B2K_OUTL_HIST_WORDS * sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&rs->ev_rows, cudaEventDisableTiming);
    if (e != cudaSuccess) { cudaGetLastError(); set_error("row store: %s", cudaGetErrorString(e)); return B2K_ECUDA; }
    *out = rs;
    return B2K_OK;
}

// Make room for need_rows rows.  Caller holds rs->mu exclusively.  `st`: the stream the old rows are copied on.
int grow(RowStore* rs, int64_t need_rows, cudaStream_t st) {
    if (need_rows <= rs->capacity) return B2K_OK;
    if (need_rows > MAX_ROWS) { set_error("index limited to %lld rows per GPU", (long long)MAX_ROWS); return B2K_ELIMIT; }
    int64_t cap = rs->capacity * 2;
    if (cap < need_rows) cap = need_rows;
    if (cap < 1024) cap = 1024;
    if (cap > MAX_ROWS) cap = MAX_ROWS;
    float *nx = nullptr, *nn = nullptr, *nl = nullptr;
    void *nb = nullptr, *nnb = nullptr;
    size_t row_bytes = (size_t)rs->d * sizeof(float);
    cudaError_t e = cudaMalloc(&nx, (size_t)cap * row_bytes);
    if (e != cudaSuccess && cap > need_rows) {  // doubling did not fit: take exactly what is needed
        cudaGetLastError();
        cap = need_rows;
        e = cudaMalloc(&nx, (size_t)cap * row_bytes);
    }
    if (e == cudaSuccess) e = cudaMalloc(&nn, (size_t)cap * sizeof(float));
    if (e == cudaSuccess && rs->lo_enabled) e = cudaMalloc(&nl, (size_t)cap * row_bytes);
    if (e == cudaSuccess && rs->bf_enabled &&
        (cudaMalloc(&nb, (size_t)cap * row_bytes / 2) != cudaSuccess || cudaMalloc(&nnb, (size_t)cap * 32) != cudaSuccess)) {
        // no room for the bf16 copies next to the grown rows: drop them, TF32 screening needs none
        cudaGetLastError();
        cudaFree(nb);
        nb = nullptr;
        rs->bf_enabled = false;
        rs->bf_nomem = true;
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        cudaFree(nx); cudaFree(nn); cudaFree(nl); cudaFree(nb); cudaFree(nnb);
        set_error("cudaMalloc for %lld rows x %d failed: %s", (long long)cap, rs->d, cudaGetErrorString(e));
        return B2K_ENOMEM;
    }
    if (rs->filled > 0) {
        // rows may still be landing on a caller's stream (on-device add): order this copy behind them
        B2K_CUDA(cudaStreamWaitEvent(st, rs->ev_rows, 0));
        B2K_CUDA(cudaMemcpyAsync(nx, rs->x, (size_t)rs->filled * row_bytes, cudaMemcpyDeviceToDevice, st));
        B2K_CUDA(cudaMemcpyAsync(nn, rs->norms, (size_t)rs->filled * sizeof(float), cudaMemcpyDeviceToDevice, st));
        if (nl) B2K_CUDA(cudaMemcpyAsync(nl, rs->xlo, (size_t)rs->filled * row_bytes, cudaMemcpyDeviceToDevice, st));
        if (nb) B2K_CUDA(cudaMemcpyAsync(nb, rs->xbf, (size_t)rs->filled * row_bytes / 2, cudaMemcpyDeviceToDevice, st));
        if (nnb) B2K_CUDA(cudaMemcpyAsync(nnb, rs->xnb, (size_t)rs->filled * 32, cudaMemcpyDeviceToDevice, st));
        B2K_CUDA(cudaStreamSynchronize(st));
    }
    // cudaFree waits for everything already enqueued on the device, including searches of other
    // handles that read the old buffers (they enqueued under the shared lock, before this point)
    cudaFree(rs->x); cudaFree(rs->norms); cudaFree(rs->xlo); cudaFree(rs->xbf); cudaFree(rs->xnb);
    rs->x = nx; rs->norms = nn; rs->xlo = nl; rs->xbf = nb; rs->xnb = nnb; rs->capacity = cap;
    return B2K_OK;
}

// Rows [first, first+n) of rs->x are in place: compute their norms (and the derived copies that exist).
int prep_new_rows(RowStore* rs, int64_t first, int64_t n, cudaStream_t st) {
    rs->outl_dirty = true;
    int rc = launch_row_prep(rs->x + first * (int64_t)rs->d, n, rs->d, rs->norms + first,
                             rs->lo_enabled ? rs->xlo + first * (int64_t)rs->d : nullptr, rs->stat + MAX_NORM_SLOT, st);
    if (!rc && rs->bf_enabled)
        rc = launch_row_bf16(rs->x + first * (int64_t)rs->d, n, rs->d,
                             static_cast<unsigned short*>(rs->xbf) + first * (int64_t)rs->d, nullptr, rs->stat + MAX_ERR_SLOT, st);
    if (!rc && rs->bf_enabled) rc = launch_norm_split(rs->norms + first, n, static_cast<unsigned char*>(rs->xnb) + first * 32, st);
    return rc;
}

// The derived copies are created on first use, on the stream of the search that needs them, under
// the exclusive lock; ev_rows is re-recorded so that other handles' searches wait for the pass.
int enable_lo(RowStore* rs, cudaStream_t st) {
    if (rs->lo_enabled) return B2K_OK;
    std::unique_lock<std::shared_mutex> lk(rs->mu);
    if (rs->lo_enabled) return B2K_OK;
    if (rs->capacity > 0) {
        cudaError_t e = cudaMalloc(&rs->xlo, (size_t)rs->capacity * rs->d * sizeof(float));
        if (e != cudaSuccess) { cudaGetLastError(); rs->xlo = nullptr; set_error("cudaMalloc for TF32 residual rows failed: %s", cudaGetErrorString(e)); return B2K_ENOMEM; }
    }
    int rc = B2K_OK;
    if (rs->filled > 0) {
        B2K_CUDA(cudaStreamWaitEvent(st, rs->ev_rows, 0));
        rc = launch_row_prep(rs->x, rs->filled, rs->d, rs->norms, rs->xlo, nullptr, st);
        if (!rc) B2K_CUDA(cudaEventRecord(rs->ev_rows, st));
    }
    if (!rc) rs->lo_enabled = true;
    return rc;
}

// bf16 copies of the rows for the bf16 screening mode / bf16 scan (half the size of the rows themselves).
int enable_bf(RowStore* rs, cudaStream_t st) {
    if (rs->bf_enabled) return B2K_OK;
    std::unique_lock<std::shared_mutex> lk(rs->mu);
    if (rs->bf_enabled) return B2K_OK;
    if (rs->bf_nomem) { set_error("no room for the bf16 copies"); return B2K_ENOMEM; }
    if (rs->capacity > 0) {
        cudaError_t e = cudaMalloc(&rs->xbf, (size_t)rs->capacity * rs->d * 2);
        if (e == cudaSuccess) e = cudaMalloc(&rs->xnb, (size_t)rs->capacity * 32);
        if (e != cudaSuccess) {
            cudaGetLastError();
            cudaFree(rs->xbf);
            rs->xbf = nullptr; rs->xnb = nullptr; rs->bf_nomem = true;
            set_error("cudaMalloc for bf16 rows failed: %s", cudaGetErrorString(e));
            return B2K_ENOMEM;
        }
    }
    int rc = B2K_OK;
    if (rs->filled > 0) {
        B2K_CUDA(cudaStreamWaitEvent(st, rs->ev_rows, 0));
        rc = launch_row_bf16(rs->x, rs->filled, rs->d, rs->xbf, nullptr, rs->stat + MAX_ERR_SLOT, st);
        if (!rc) rc = launch_norm_split(rs->norms, rs->filled, rs->xnb, st);
        if (!rc) B2K_CUDA(cudaMemsetAsync(rs->outl, 0, B2K_OUTL_WORDS * sizeof(unsigned int), st));  // fresh records: no row is out
        rs->outl_dirty = true;
        if (!rc) B2K_CUDA(cudaEventRecord(rs->ev_rows, st));
    }
    if (!rc) rs->bf_enabled = true;
    return rc;
}

// Outlier rows for the norm-folded screen (see launch_outlier_pass): recomputed after the rows changed, under the
// exclusive lock and with the device idle -- the pass rewrites a few norm records that searches of other handles on
// this store may be reading.
int ensure_outliers(RowStore* rs, cudaStream_t st) {
    if (!rs->outl_dirty || !rs->bf_enabled) return B2K_OK;
    std::unique_lock<std::shared_mutex> lk(rs->mu);
    if (!rs->outl_dirty) return B2K_OK;
    B2K_CUDA(cudaDeviceSynchronize());
    B2K_CUDA(cudaStreamWaitEvent(st, rs->ev_rows, 0));
    int rc = launch_outlier_pass(rs->norms, rs->filled, rs->xnb, rs->stat + MAX_NORM_SLOT, rs->stat + MAX_ERR_SLOT, rs->outl,
                                 rs->outl_hist, st);
    if (rc) return rc;
    B2K_CUDA(cudaEventRecord(rs->ev_rows, st));
    rs->outl_dirty = false;
    return B2K_OK;
}

// Candidates the bf16 scan selects for a request of k (0: the bf16 scan cannot serve this k).
int scan_bf_candidates(int k) {
    const int kp = 2 * k + 32;
    return kp <= B2K_MAX_K ? kp : 0;
}

// The bf16 scan pays once the rows are large enough for HBM time to matter: below 128 MB of rows
// the extra re-rank launch costs more than the halved row stream saves.
bool scan_wants_bf(const b2k_index* h, int k) {
    if (h->opt_scan_rows == 1 || h->rs->bf_nomem || h->d % 8 != 0 || scan_bf_candidates(k) == 0) return false;
    if (h->opt_scan_rows == 2) return true;
    return (double)h->ntotal * h->d >= 32e6;
}

size_t scan_ws_bytes(const b2k_index* h, int64_t nq, int k, bool bf) {
    const int grid = scan_grid(h->ntotal, h->sm_count, h->opt_scan_ctas_per_sm);
    if (!bf) return scan_partial_bytes(grid, 8, k);
    const int kp = scan_bf_candidates(k);
    return align256(scan_partial_bytes(grid, 8, kp)) + align256((size_t)nq * kp * 4) + align256((size_t)nq * kp * 8);
}

int scan_search(b2k_index* h, int64_t nq, const float* qd, int k, float* Dd, int64_t* Id, unsigned char* ws_tail,
                cudaStream_t st, bool bf) {
    const RowStore* rs = h->rs.get();
    const int grid = scan_grid(h->ntotal, h->sm_count, h->opt_scan_ctas_per_sm);
    u64* partial = reinterpret_cast<u64*>(ws_tail);
    if (h->opt_timing) {
        if (!h->ev_t0) { B2K_CUDA(cudaEventCreate(&h->ev_t0)); B2K_CUDA(cudaEventCreate(&h->ev_t1)); }
        B2K_CUDA(cudaEventRecord(h->ev_t0, st));
    }
    h->scan_launches = 0;
    h->last_scan_bf = bf;
    if (bf) {
        // screening scan over the bf16 copies -> kp candidates per query -> exact re-rank; if the
        // candidates did not cover the slack zone (flag), the fp32 scan behind it redoes the search
        const int kp = scan_bf_candidates(k);
        unsigned char* w = ws_tail + align256(scan_partial_bytes(grid, 8, kp));
        float* Da = reinterpret_cast<float*>(w);
        int64_t* Ia = reinterpret_cast<int64_t*>(w + align256((size_t)nq * kp * 4));
        unsigned int* flag = h->done + SCAN_FLAG_SLOT;
        B2K_CUDA(cudaMemsetAsync(flag, 0, sizeof(unsigned int), st));
        for (int64_t q0 = 0; q0 < nq;) {
            int qt = scan_pick_qt(nq - q0, kp);
            if (qt > 4) qt = 4;  // 8 queries per pass would make the bf16 variant ALU-bound
            int rc = launch_scan(rs->xbf, qd + q0 * (int64_t)h->d, h->ntotal, h->d, qt, kp, grid, partial, h->done,
                                 Da + q0 * kp, Ia + q0 * kp, st, true, nullptr);
            if (rc) return rc;
            q0 += qt;
            ++h->scan_launches;
        }
        if (h->opt_timing) { B2K_CUDA(cudaEventRecord(h->ev_t1, st)); h->scan_timed = true; }
        int rc = launch_scan_rerank(rs->x, qd, nq, h->d, k, kp, Da, Ia, rs->stat + MAX_ERR_SLOT, Dd, Id, flag, st);
        if (rc) return rc;
        for (int64_t q0 = 0; q0 < nq;) {
            int qt = scan_pick_qt(nq - q0, k);
            rc = launch_scan(rs->x, qd + q0 * (int64_t)h->d, h->ntotal, h->d, qt, k, grid, partial, h->done,
                             Dd + q0 * k, Id + q0 * k, st, false, flag);
            if (rc) return rc;
            q0 += qt;
        }
        return B2K_OK;
    }
    for (int64_t q0 = 0; q0 < nq;) {
        int qt = scan_pick_qt(nq - q0, k);
        int rc = launch_scan(rs->x, qd + q0 * (int64_t)h->d, h->ntotal, h->d, qt, k, grid, partial, h->done,
                             Dd + q0 * k, Id + q0 * k, st);
        if (rc) return rc;
        q0 += qt;
        ++h->scan_launches;
    }
    if (h->opt_timing) { B2K_CUDA(cudaEventRecord(h->ev_t1, st)); h->scan_timed = true; }
    return B2K_OK;
}



// A handle over `rs` seeing its first n rows.
int make_handle(int d, int device, int sm, std::shared_ptr<RowStore> rs, int64_t n, b2k_index** out) {
    b2k_index* h = new (std::nothrow) b2k_index();
    if (!h) { set_error("out of host memory"); return B2K_ENOMEM; }
    h->d = d; h->device = device; h->sm_count = sm;
    h->rs = std::move(rs);
    h->ntotal = n;
    cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&h->done, 256);
    if (e == cudaSuccess) e = cudaMemset(h->done, 0, 256);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_ws, cudaEventDisableTiming);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("b2k_create: %s", cudaGetErrorString(e));
        if (h->stream) cudaStreamDestroy(h->stream);
        cudaFree(h->done);
        if (h->ev_ws) cudaEventDestroy(h->ev_ws);
        delete h;
        return B2K_ECUDA;
    }
    *out = h;
    return B2K_OK;
}

// Give h a private store holding exactly its own rows (it shared one that has since moved on, or
// it is about to drop its rows).  keep = false: start empty.
int detach(b2k_index* h, bool keep) {
    std::shared_ptr<RowStore> ns;
    int rc = new_store(h->d, h->device, &ns);
    if (rc) return rc;
    if (keep && h->ntotal > 0) {
        RowStore* os = h->rs.get();
        std::shared_lock<std::shared_mutex> lk(os->mu);
        rc = grow(ns.get(), h->ntotal, h->stream);
        if (rc) return rc;
        B2K_CUDA(cudaStreamWaitEvent(h->stream, os->ev_rows, 0));
        B2K_CUDA(cudaMemcpyAsync(ns->x, os->x, (size_t)h->ntotal * h->d * sizeof(float), cudaMemcpyDeviceToDevice, h->stream));
        ns->filled = h->ntotal;
        rc = prep_new_rows(ns.get(), 0, h->ntotal, h->stream);
        if (rc) return rc;
        B2K_CUDA(cudaEventRecord(ns->ev_rows, h->stream));
        B2K_CUDA(cudaStreamSynchronize(h->stream));
    }
    h->rs = ns;  // the old store goes away with its last handle
    h->screen_fail_streak = 0;
    return B2K_OK;
}

// Before rows are appended through h: h must sit at the end of its store.  Returns with rs->mu NOT held.
int own_the_tail(b2k_index* h) {
    bool at_end;
    {
        std::shared_lock<std::shared_mutex> lk(h->rs->mu);
        at_end = h->rs->filled == h->ntotal;
    }
    // (another view that appends between this check and the caller's exclusive lock is caught there)
    return at_end ? B2K_OK : detach(h, true);
}

// Appends n rows that `fill` writes at rs->x + first*d on stream st (a copy, or the generator).
template <class Fill>
int append_rows(b2k_index* h, int64_t n, cudaStream_t st, bool sync_after, Fill fill) {
    int rc = own_the_tail(h);
    if (rc) return rc;
    std::unique_lock<std::shared_mutex> sl(h->rs->mu);
    if (h->rs->filled != h->ntotal) {  // another view appended in the meantime
        sl.unlock();
        rc = detach(h, true);
        if (rc) return rc;
        sl = std::unique_lock<std::shared_mutex>(h->rs->mu);
    }
    RowStore* rs = h->rs.get();
    rc = grow(rs, rs->filled + n, h->stream);
    if (rc) return rc;
    // ev_rows is one event: before it is re-recorded on this stream, this stream takes over the
    // dependency on whatever stream recorded it last, so "all rows so far" stays one wait
    B2K_CUDA(cudaStreamWaitEvent(st, rs->ev_rows, 0));
    rc = fill(rs->x + rs->filled * (int64_t)rs->d, st);
    if (rc) return rc;
    rc = prep_new_rows(rs, rs->filled, n, st);
    if (rc) return rc;
    B2K_CUDA(cudaEventRecord(rs->ev_rows, st));
    rs->filled += n;
    h->ntotal += n;
    if (sync_after) {
        cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { h->sticky = (int)e; set_error("b2k_add: %s", cudaGetErrorString(e)); return B2K_ECUDA; }
    }
    return B2K_OK;
}

}  // namespace

void ixf2_put_header(unsigned char* b, int d, int64_t n) {
    memcpy(b, "IxF2", 4);
    int32_t d32 = d; memcpy(b + 4, &d32, 4);
    memcpy(b + 8, &n, 8);
    int64_t dummy = 1 << 20; memcpy(b + 16, &dummy, 8); memcpy(b + 24, &dummy, 8);
    b[32] = 1;
    int32_t metric = 1; memcpy(b + 33, &metric, 4);
    uint64_t count = (uint64_t)n * (uint64_t)d; memcpy(b + 37, &count, 8);
}

int ixf2_parse_header(const unsigned char* hdr, const char* path, int* d, int64_t* n) {
    if (memcmp(hdr, "IxF2", 4) != 0) {
        set_error("%s: index type '%.4s' not supported (only IxF2 = IndexFlatL2)", path, (const char*)hdr);
        return B2K_EIO;
    }
    int32_t d32, metric; int64_t nn; uint64_t count;
    memcpy(&d32, hdr + 4, 4); memcpy(&nn, hdr + 8, 8); memcpy(&metric, hdr + 33, 4); memcpy(&count, hdr + 37, 8);
    if (d32 <= 0 || nn < 0 || metric != 1 || count != (uint64_t)nn * (uint64_t)d32) {
        set_error("%s: inconsistent IxF2 header (d=%d ntotal=%lld metric=%d count=%llu)", path, d32, (long long)nn, metric, (unsigned long long)count);
        return B2K_EIO;
    }
    *d = d32; *n = nn;
    return B2K_OK;
}

}  // namespace b2k

using namespace b2k;

extern "C" {

int b2k_version(void) { return 200; }

const char* b2k_last_error(void) { return t_err.c_str(); }

int64_t b2k_launch_count(void) { return g_launches.load(); }

int b2k_create(int d, int device, b2k_index** out) {
    if (!out) { set_error("b2k_create: out is null"); return B2K_EINVAL; }
    *out = nullptr;
    if (d <= 0) { set_error("b2k_create: d must be positive, got %d", d); return B2K_EINVAL; }
    int sm = 148;
    int rc = check_device(device, &sm);
    if (rc) return rc;
    DeviceGuard g(device);
    std::shared_ptr<RowStore> rs;
    rc = new_store(d, device, &rs);
    if (rc) return rc;
    return make_handle(d, device, sm, rs, 0, out);
}

int b2k_create_multi(int d, const int* devices, int ndev, b2k_index** out) {
    if (!out) { set_error("b2k_create_multi: out is null"); return B2K_EINVAL; }
    *out = nullptr;
    if (!devices || ndev <= 0) { set_error("b2k_create_multi: empty device list"); return B2K_EINVAL; }
    if (ndev == 1) return b2k_create(d, devices[0], out);
    return multi_create(d, devices, ndev, out);
}

int b2k_view(const b2k_index* src, int64_t nrows, b2k_index** out) {
    if (!src || !out) { set_error("b2k_view: null argument"); return B2K_EINVAL; }
    *out = nullptr;
    if (src->multi) return multi_view(src, nrows, out);
    std::lock_guard<std::mutex> lk(src->mu);
    const int64_t n = (nrows < 0 || nrows > src->ntotal) ? src->ntotal : nrows;
    DeviceGuard g(src->device);
    return make_handle(src->d, src->device, src->sm_count, src->rs, n, out);
}

int b2k_destroy(b2k_index* h) {
    if (!h) return B2K_OK;
    if (h->multi) { multi_destroy(h); return B2K_OK; }
    DeviceGuard g(h->device);
    // work of this handle may still be in flight; work of other handles on a shared store is the store's business
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->ev_ws) cudaEventSynchronize(h->ev_ws);
    if (h->k3) k3_destroy(h->k3);
    if (!give_spare(g_spares.ws[h->device & 63], h->ws, h->ws_bytes)) cudaFree(h->ws);  // (all of this handle's work is done: synchronised above)
    cudaFree(h->done);
    if (h->pin && !give_spare(g_spares.pin[h->device & 63], h->pin, h->pin_bytes)) cudaFreeHost(h->pin);
    if (h->ev_ws) cudaEventDestroy(h->ev_ws);
    if (h->ev_t0) cudaEventDestroy(h->ev_t0);
    if (h->ev_t1) cudaEventDestroy(h->ev_t1);
    if (h->stream) cudaStreamDestroy(h->stream);
    h->rs.reset();
    delete h;
    return B2K_OK;
}

int64_t b2k_ntotal(const b2k_index* h) { return h ? h->ntotal : -1; }
int b2k_d(const b2k_index* h) { return h ? h->d : -1; }

int b2k_set_option(b2k_index* h, const char* key, int64_t value) {
    if (!h || !key) { set_error("b2k_set_option: null argument"); return B2K_EINVAL; }
    if (h->multi) return multi_set_option(h, key, value);
    std::lock_guard<std::mutex> lk(h->mu);
    if (!strcmp(key, "path")) {
        if (value < 0 || value > 2) { set_error("option path must be 0, 1 or 2"); return B2K_EINVAL; }
        h->opt_path = (int)value;
    } else if (!strcmp(key, "scan_ctas_per_sm")) {
        if (value < 1 || value > 8) { set_error("option scan_ctas_per_sm must be 1..8"); return B2K_EINVAL; }
        h->opt_scan_ctas_per_sm = (int)value;
    } else if (!strcmp(key, "tensor_cta_group")) {
        if (value < 0 || value > 2) { set_error("option tensor_cta_group must be 0 (auto), 1 or 2"); return B2K_EINVAL; }
        h->opt_cta_group = (int)value;
        if (h->k3) k3_set_cta_group(h->k3, h->opt_cta_group);
    } else if (!strcmp(key, "tensor_sep_accum")) {
        if (value < 0 || value > 2) { set_error("option tensor_sep_accum must be 0 (auto), 1 (on) or 2 (off)"); return B2K_EINVAL; }
        h->opt_sep = (int)value;
        if (h->k3) k3_set_sep_mode(h->k3, h->opt_sep);
    } else if (!strcmp(key, "tensor_mode")) {
        if (value < 0 || value > 4) { set_error("option tensor_mode must be 0 (auto), 1 (3xTF32), 2 (TF32 screening), 3 (bf16 screening) or 4 (3xTF32 screening); 2-4 end in the exact re-rank"); return B2K_EINVAL; }
        h->opt_tensor_mode = (int)value;
    } else if (!strcmp(key, "tensor_segments")) {
        if (value < 0 || value > 100000) { set_error("option tensor_segments must be 0 (one launch), 1 (segments with bound refreshes) or a target number of keys per query and segment"); return B2K_EINVAL; }
        h->opt_segments = (int)value;
        if (h->k3) k3_set_segments(h->k3, h->opt_segments);
    } else if (!strcmp(key, "tensor_epilogue_warps")) {
        if (value != 0 && value != 4 && value != 8) { set_error("option tensor_epilogue_warps must be 0 (auto), 4 or 8"); return B2K_EINVAL; }
        h->opt_epi = (int)value / 4;
        if (h->k3) k3_set_epi_halves(h->k3, h->opt_epi);
    } else if (!strcmp(key, "screen_flag_ptr")) {
        // device address of a caller-owned 32-bit word (0 = off): batched searches then do not
        // synchronise; they raise the word when the screening lists could not hold every candidate,
        // and the caller must check it (after its own synchronisation) and repeat that search with
        // tensor_mode = 1.  Used by the sharded indexes, which collect the words of all shards.
        h->opt_flag_ptr = reinterpret_cast<unsigned int*>(static_cast<uintptr_t>(value));
        if (h->k3) k3_set_ext_flag(h->k3, h->opt_flag_ptr);
    } else if (!strcmp(key, "scan_rows")) {
        if (value < 0 || value > 2) { set_error("option scan_rows must be 0 (auto), 1 (fp32 rows) or 2 (bf16 copies + exact re-rank)"); return B2K_EINVAL; }
        h->opt_scan_rows = (int)value;
    } else if (!strcmp(key, "prepare_copies")) {
        // build the bf16 copies now (start-up) instead of inside the first search that profits from them:
        // that search would otherwise allocate +50% of the row bytes and pay one conversion pass
        if (value && h->d % 8 == 0) {
            DeviceGuard g(h->device);
            int rc = enable_bf(h->rs.get(), h->stream);
            if (rc) return rc;
            B2K_CUDA(cudaStreamSynchronize(h->stream));
        }
    } else if (!strcmp(key, "tensor_query_stationary")) {
        h->opt_qs = value != 0;
        if (h->k3) k3_set_query_stationary(h->k3, h->opt_qs);
    } else if (!strcmp(key, "tensor_prefetch_tiles")) {
        if (value < 0 || value > 4096) { set_error("option tensor_prefetch_tiles must be 0..4096"); return B2K_EINVAL; }
        h->opt_pf = (int)value;
        if (h->k3) k3_set_prefetch_tiles(h->k3, h->opt_pf);
    } else if (!strcmp(key, "tensor_warm_keys")) {
        if (value < 2 || value > 64) { set_error("option tensor_warm_keys must be 2..64"); return B2K_EINVAL; }
        h->opt_warm = (int)value;
        if (h->k3) k3_set_warm_keys(h->k3, h->opt_warm);
    } else if (!strcmp(key, "tensor_probe")) {
        h->opt_probe = (int)(value < 0 ? 0 : (value > 3 ? 3 : value));
        if (h->k3) k3_set_probe(h->k3, h->opt_probe);
    } else if (!strcmp(key, "timing")) {
        h->opt_timing = value != 0;
        if (h->k3) k3_set_timing(h->k3, h->opt_timing != 0);
    } else {
        set_error("unknown option '%s'", key);
        return B2K_EINVAL;
    }
    return B2K_OK;
}

int b2k_get_stat(b2k_index* h, const char* key, double* out) {
    if (!h || !key || !out) { set_error("b2k_get_stat: null argument"); return B2K_EINVAL; }
    if (h->multi) return multi_get_stat(h, key, out);
    std::lock_guard<std::mutex> lk(h->mu);
    if (!strcmp(key, "last_path")) { *out = h->last_path; return B2K_OK; }
    if (!strcmp(key, "tensor_mode_used")) { *out = h->k3 ? k3_last_mode(h->k3) : 0; return B2K_OK; }
    if (!strcmp(key, "screen_fallbacks")) { *out = (double)h->stat_screen_fallbacks; return B2K_OK; }
    if (!strcmp(key, "scan_rows_used")) { *out = h->last_scan_bf ? 2 : 1; return B2K_OK; }
    if (!strcmp(key, "shares_rows")) { *out = h->rs.use_count() > 1 ? 1 : 0; return B2K_OK; }
    if (!strcmp(key, "devices")) { *out = 1; return B2K_OK; }
    if (!strcmp(key, "scan_fallback")) {  // 1 if the last bf16 scan had to be redone on the fp32 rows (synchronises)
        DeviceGuard g(h->device);
        unsigned int f = 0;
        if (h->last_scan_bf) {
            B2K_CUDA(cudaDeviceSynchronize());
            B2K_CUDA(cudaMemcpy(&f, h->done + SCAN_FLAG_SLOT, sizeof(f), cudaMemcpyDeviceToHost));
        }
        *out = f;
        return B2K_OK;
    }
    if (!strcmp(key, "tensor_kernel_ms")) {
        float ms = 0.f;
        DeviceGuard g(h->device);
        if (!h->k3 || k3_last_kernel_ms(h->k3, &ms)) { set_error("no timed tensor-core kernel yet (set option timing=1)"); return B2K_EINVAL; }
        *out = ms;
        return B2K_OK;
    }
    if (!strcmp(key, "scan_kernel_ms") || !strcmp(key, "scan_launches")) {
        if (!h->scan_timed) { set_error("no timed scan yet (set option timing=1)"); return B2K_EINVAL; }
        if (!strcmp(key, "scan_launches")) { *out = h->scan_launches; return B2K_OK; }
        DeviceGuard g(h->device);
        float ms = 0.f;
        B2K_CUDA(cudaEventSynchronize(h->ev_t1));
        B2K_CUDA(cudaEventElapsedTime(&ms, h->ev_t0, h->ev_t1));
        *out = ms;
        return B2K_OK;
    }
    set_error("unknown stat '%s'", key);
    return B2K_EINVAL;
}

int b2k_reserve(b2k_index* h, int64_t nrows) {
    if (!h || nrows < 0) { set_error("b2k_reserve: bad argument"); return B2K_EINVAL; }
    if (h->multi) return multi_reserve(h, nrows);
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    std::unique_lock<std::shared_mutex> sl(h->rs->mu);
    return grow(h->rs.get(), nrows, h->stream);
}

int b2k_reset(b2k_index* h) {
    if (!h) { set_error("b2k_reset: null handle"); return B2K_EINVAL; }
    if (h->multi) return multi_reset(h);
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    B2K_CUDA(cudaStreamSynchronize(h->stream));
    B2K_CUDA(cudaEventSynchronize(h->ev_ws));
    int rc = detach(h, false);  // a fresh empty store; the old one is freed with its last handle
    if (rc) return rc;
    h->ntotal = 0;
    B2K_CUDA(cudaMemset(h->done, 0, 256));
    return B2K_OK;
}

int b2k_add(b2k_index* h, int64_t n, const float* x, int on_device, void* stream) {
    if (!h) { set_error("b2k_add: null handle"); return B2K_EINVAL; }
    if (n < 0 || (n > 0 && !x)) { set_error("b2k_add: bad argument (n=%lld)", (long long)n); return B2K_EINVAL; }
    if (n == 0) return B2K_OK;
    if (h->multi) return multi_add(h, n, x, on_device, stream);
    std::lock_guard<std::mutex> lk(h->mu);
    if (h->sticky) { set_error("handle has a sticky CUDA error (%d)", h->sticky); return B2K_ECUDA; }
    DeviceGuard g(h->device);
    cudaStream_t st = on_device ? (cudaStream_t)stream : h->stream;
    const size_t bytes = (size_t)n * h->d * sizeof(float);
    return append_rows(h, n, st, !on_device, [&](float* dst, cudaStream_t s) -> int {
        // (cudaMemcpyDefault: a device source may live on another GPU of the process -- the multi-GPU handle's adds)
        B2K_CUDA(cudaMemcpyAsync(dst, x, bytes, on_device ? cudaMemcpyDefault : cudaMemcpyHostToDevice, s));
        return B2K_OK;
    });
}

int b2k_add_synthetic(b2k_index* h, int64_t n, uint64_t seed, int64_t first_row) {
    if (!h || n < 0) { set_error("b2k_add_synthetic: bad argument"); return B2K_EINVAL; }
    if (n == 0) return B2K_OK;
    if (h->multi) return multi_add_synthetic(h, n, seed, first_row);
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    const int d = h->d;
    return append_rows(h, n, h->stream, true, [&](float* dst, cudaStream_t s) -> int {
        return launch_synth(seed, first_row, n, d, dst, s);
    });
}

int b2k_synth(int device, uint64_t seed, int64_t row0, int64_t nrows, int d, float* out_dev, void* stream) {
    if (nrows < 0 || d <= 0 || (nrows > 0 && !out_dev)) { set_error("b2k_synth: bad argument"); return B2K_EINVAL; }
    int rc = check_device(device, nullptr);
    if (rc) return rc;
    DeviceGuard g(device);
    rc = launch_synth(seed, row0, nrows, d, out_dev, (cudaStream_t)stream);
    if (rc) return rc;
    if (!stream) B2K_CUDA(cudaStreamSynchronize(nullptr));
    return B2K_OK;
}

int b2k_get_rows(const b2k_index* hc, int64_t i0, int64_t n, float* out, int on_device) {
    b2k_index* h = const_cast<b2k_index*>(hc);
    if (!h || i0 < 0 || n < 0 || (n > 0 && !out)) { set_error("b2k_get_rows: bad argument"); return B2K_EINVAL; }
    if (h->multi) return multi_get_rows(h, i0, n, out, on_device);
    std::lock_guard<std::mutex> lk(h->mu);
    if (i0 + n > h->ntotal) { set_error("b2k_get_rows: rows [%lld,%lld) out of range (ntotal=%lld)", (long long)i0, (long long)(i0 + n), (long long)h->ntotal); return B2K_EINVAL; }
    if (n == 0) return B2K_OK;
    DeviceGuard g(h->device);
    std::shared_lock<std::shared_mutex> sl(h->rs->mu);
    B2K_CUDA(cudaStreamWaitEvent(h->stream, h->rs->ev_rows, 0));
    B2K_CUDA(cudaMemcpyAsync(out, h->rs->x + i0 * (int64_t)h->d, (size_t)n * h->d * sizeof(float),
                             on_device ? cudaMemcpyDefault : cudaMemcpyDeviceToHost, h->stream));
    B2K_CUDA(cudaStreamSynchronize(h->stream));
    return B2K_OK;
}

int b2k_search(b2k_index* h, int64_t nq, const float* xq, int64_t k, float* D, int64_t* I, int on_device,
               void* stream) {
    if (!h) { set_error("b2k_search: null handle"); return B2K_EINVAL; }
    if (k <= 0) { set_error("b2k_search: k must be positive, got %lld", (long long)k); return B2K_EINVAL; }
    if (nq < 0 || (nq > 0 && (!xq || !D || !I))) { set_error("b2k_search: bad argument"); return B2K_EINVAL; }
    if (k > B2K_MAX_K) { set_error("b2k_search: k=%lld exceeds B2K_MAX_K=%d", (long long)k, B2K_MAX_K); return B2K_ELIMIT; }
    if (nq == 0) return B2K_OK;
    if (h->multi) return multi_search(h, nq, xq, k, D, I, on_device, stream);
    std::lock_guard<std::mutex> lk(h->mu);
    if (h->sticky) { set_error("handle has a sticky CUDA error (%d)", h->sticky); return B2K_ECUDA; }
    DeviceGuard g(h->device);
    RowStore* rs = h->rs.get();
    const int kk = (int)k;
    cudaStream_t st = on_device ? (cudaStream_t)stream : h->stream;

    // faiss switches from direct differences to the sgemm expansion at nq = 20; here both kernels end
    // in direct fp32 differences (the tensor path re-ranks), so the switch is purely a matter of speed:
    // the scan costs one pass over the rows per group of <= 4 (bf16) or 8 (fp32) queries, the tensor
    // path ~0.3 ms of fixed work plus one cheap pass for any nq <= 128 (measured at 10M x 384: nq = 8
    // scan 4.9 ms, tensor 1.4 ms; nq = 16: 9.8 vs 1.4; nq = 2: 1.8 vs 1.35; at 100k rows nq = 4: 0.12 vs 0.16 ms).
    const double elems = (double)h->ntotal * h->d;
    const bool tc_pays = nq >= 20 || (h->opt_tensor_mode != 1 && ((nq >= 5 && elems >= 16e6) || (nq >= 3 && elems >= 256e6) ||
                                                                  (nq >= 2 && elems >= 2e9)));
    const bool want_tc = h->ntotal > 0 && (h->opt_path == 2 || (h->opt_path == 0 && tc_pays)) && k3_supported(h->d, kk);

    // workspace: [queries][D][I] for host callers, then the kernels' scratch
    size_t q_bytes = align256((size_t)nq * h->d * sizeof(float));
    size_t d_bytes = align256((size_t)nq * kk * sizeof(float));
    size_t i_bytes = align256((size_t)nq * kk * sizeof(int64_t));
    // device queries that are not 16-byte aligned (a torch view with an odd storage offset) are staged:
    // the re-rank kernels and the TMA maps read query rows with 128-bit accesses
    const bool stage_q = on_device && (reinterpret_cast<uintptr_t>(xq) & 15) != 0;
    size_t io_bytes = on_device ? (stage_q ? q_bytes : 0) : q_bytes + d_bytes + i_bytes;
    size_t scratch = 0;
    if (h->ntotal > 0) {
        if (want_tc) {
            scratch = k3_scratch_bytes(h->sm_count, h->d, nq < TC_QUERY_CHUNK ? nq : TC_QUERY_CHUNK, kk);
            const size_t sb = scan_ws_bytes(h, nq, kk, false);  // last line of defence of the screening modes
            if (sb > scratch) scratch = sb;
        }
        else scratch = scan_ws_bytes(h, nq, kk, scan_wants_bf(h, kk));
    }
    int rc = ensure_ws(h, io_bytes + align256(scratch));
    if (rc) return rc;
    B2K_CUDA(cudaStreamWaitEvent(st, rs->ev_rows, 0));
    B2K_CUDA(cudaStreamWaitEvent(st, h->ev_ws, 0));
    unsigned char* wsb = static_cast<unsigned char*>(h->ws);
    const float* qd = xq;
    float* Dd = D;
    int64_t* Id = I;
    bool q_pinned = false, out_pinned = false;
    if (!on_device) {
        // page-locked caller buffers are copied to / from directly; pageable ones go through the
        // handle's own pinned staging buffer (an extra host memcpy each way)
        q_pinned = is_pinned_host(xq);
        out_pinned = is_pinned_host(D) && is_pinned_host(I);
        if (!q_pinned || !out_pinned) {
            rc = ensure_pin(h, io_bytes);
            if (rc) return rc;
        }
        if (!q_pinned) memcpy(h->pin, xq, (size_t)nq * h->d * sizeof(float));
        B2K_CUDA(cudaMemcpyAsync(wsb, q_pinned ? (const void*)xq : h->pin, (size_t)nq * h->d * sizeof(float), cudaMemcpyHostToDevice, st));
        qd = reinterpret_cast<const float*>(wsb);
        Dd = reinterpret_cast<float*>(wsb + q_bytes);
        Id = reinterpret_cast<int64_t*>(wsb + q_bytes + d_bytes);
    }
    if (stage_q) {
        B2K_CUDA(cudaMemcpyAsync(wsb, xq, (size_t)nq * h->d * sizeof(float), cudaMemcpyDeviceToDevice, st));
        qd = reinterpret_cast<const float*>(wsb);
    }
    unsigned char* ws_tail = wsb + io_bytes;

    if (h->ntotal == 0) {
        rc = launch_fill_empty(Dd, Id, nq * kk, st);
    } else if (want_tc) {
        // 1 = 3xTF32 (faiss-like expansion numerics), 2 = TF32 screening, 3 = bf16 screening (rows need a
        // 16-byte bf16 pitch; if the bf16 copies do not fit in memory the TF32 screening serves
        // instead), 4 = 3xTF32 as a screen.  2, 3 and 4 end in the exact fp32 re-rank.
        int mode = h->opt_tensor_mode == 0 ? 3 : h->opt_tensor_mode;
        if (mode == 3 && (h->d % 8 != 0 || rs->bf_nomem)) mode = 2;
        // an index whose cheap screen keeps overflowing (outlier norms, offset data) starts at the
        // 3xTF32 screen; every 64th search tries the cheap one again (rows may have changed)
        if ((mode == 2 || mode == 3) && h->opt_tensor_mode == 0 && h->screen_fail_streak >= 2 && (++h->screen_skipped & 63) != 0) mode = 4;
        // (the derived copies are created before the shared lock is taken: creating them needs it exclusively)
        if (mode == 3 && enable_bf(rs, st) != B2K_OK) mode = 2;
        if (mode == 3) rc = ensure_outliers(rs, st);
        if (mode == 4 && h->opt_tensor_mode == 0 && enable_lo(rs, st) != B2K_OK) mode = 2;  // no room for the residual rows: cheap screen, then the scan
        if (mode == 1 || mode == 4) rc = enable_lo(rs, st);
        const bool cheap = mode == 2 || mode == 3;
        bool cheap_failed = false;
        if (!rc && !h->k3) {
            rc = k3_create(&h->k3, h->device, h->sm_count);
            if (!rc) k3_set_timing(h->k3, h->opt_timing != 0);
            if (!rc) k3_set_cta_group(h->k3, h->opt_cta_group);
            if (!rc) k3_set_sep_mode(h->k3, h->opt_sep);
            if (!rc) k3_set_segments(h->k3, h->opt_segments);
            if (!rc) k3_set_epi_halves(h->k3, h->opt_epi);
            if (!rc) k3_set_ext_flag(h->k3, h->opt_flag_ptr);
            if (!rc) k3_set_probe(h->k3, h->opt_probe);
            if (!rc) k3_set_query_stationary(h->k3, h->opt_qs);
            if (!rc) k3_set_prefetch_tiles(h->k3, h->opt_pf);
            if (!rc) k3_set_warm_keys(h->k3, h->opt_warm);
        }
        if (!rc) h->last_path = 2;
        // like faiss (query blocks of 4096), long batches go through in chunks: the candidate-list
        // scratch is sized per chunk and one pass over the rows per chunk is still compute-bound
        for (int64_t q0 = 0; q0 < nq && !rc; q0 += TC_QUERY_CHUNK) {
            const int64_t nqc = nq - q0 < TC_QUERY_CHUNK ? nq - q0 : TC_QUERY_CHUNK;
            const float* qc = qd + q0 * (int64_t)h->d;
            int overflowed = 0;
            {
                std::shared_lock<std::shared_mutex> sl(rs->mu);
                rc = k3_search(h->k3, rs->x, rs->xlo, rs->xbf, rs->xnb, rs->outl, rs->norms, rs->stat + MAX_NORM_SLOT, rs->stat + MAX_ERR_SLOT, h->ntotal,
                               h->d, qc, nqc, kk, Dd + q0 * kk, Id + q0 * kk, ws_tail, st, mode, &overflowed);
            }
            if (!rc && overflowed && mode != 4) {
                cheap_failed = true;
                // the slack zone of the cheap screen did not fit its lists (masses of near-duplicates around
                // the k-th neighbour, or a distribution whose norms dwarf its distances): repeat this chunk
                // with the 3xTF32 product as the screen, whose zone is ~10^4 times narrower
                ++h->stat_screen_fallbacks;
                if (enable_lo(rs, st) == B2K_OK) {  // (no room for the residual rows: straight to the scan, overflowed stays set)
                    std::shared_lock<std::shared_mutex> sl(rs->mu);
                    rc = k3_search(h->k3, rs->x, rs->xlo, rs->xbf, rs->xnb, rs->outl, rs->norms, rs->stat + MAX_NORM_SLOT, rs->stat + MAX_ERR_SLOT,
                                   h->ntotal, h->d, qc, nqc, kk, Dd + q0 * kk, Id + q0 * kk, ws_tail, st, 4, &overflowed);
                }
            }
            if (!rc && overflowed) {
                // even that zone is crowded (exact duplicates by the thousand): the scan kernel is exact whatever the data
                ++h->stat_screen_fallbacks;
                h->last_path = 1;
                std::shared_lock<std::shared_mutex> sl(rs->mu);
                rc = scan_search(h, nqc, qc, kk, Dd + q0 * kk, Id + q0 * kk, ws_tail, st, false);
            }
        }
        if (cheap) h->screen_fail_streak = cheap_failed ? h->screen_fail_streak + 1 : 0;
    } else {
        h->last_path = 1;
        bool bf = scan_wants_bf(h, kk);
        if (bf && enable_bf(rs, st) != B2K_OK) bf = false;  // no room for the copies: fp32 rows (the workspace is large enough for either)
        std::shared_lock<std::shared_mutex> sl(rs->mu);
        rc = scan_search(h, nq, qd, kk, Dd, Id, ws_tail, st, bf);
    }
    if (rc) return rc;
    B2K_CUDA(cudaEventRecord(h->ev_ws, st));

    if (!on_device) {
        unsigned char* pb = static_cast<unsigned char*>(h->pin);
        B2K_CUDA(cudaMemcpyAsync(out_pinned ? (void*)D : (void*)(pb + q_bytes), Dd, (size_t)nq * kk * sizeof(float), cudaMemcpyDeviceToHost, st));
        B2K_CUDA(cudaMemcpyAsync(out_pinned ? (void*)I : (void*)(pb + q_bytes + d_bytes), Id, (size_t)nq * kk * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { h->sticky = (int)e; set_error("b2k_search: %s", cudaGetErrorString(e)); return B2K_ECUDA; }
        if (!out_pinned) {
            memcpy(D, pb + q_bytes, (size_t)nq * kk * sizeof(float));
            memcpy(I, pb + q_bytes + d_bytes, (size_t)nq * kk * sizeof(int64_t));
        }
    }
    return B2K_OK;
}

// Device-side scratch of b2k_merge_topk for host callers / id offsets: one small buffer per device, grown on demand
// (no cudaMalloc / cudaFree per call), guarded by a mutex that also serialises its users.
namespace {
struct MergeScratch { std::mutex mu; void* p = nullptr; size_t bytes = 0; };
MergeScratch g_merge_scratch[64];
}  // namespace

int b2k_merge_topk(int device, int nlists, int64_t nq, int64_t k, const float* Din, const int64_t* Iin,
                   const int64_t* id_offset, float* D, int64_t* I, int on_device, void* stream) {
    if (nlists <= 0 || nq < 0 || k <= 0 || (nq > 0 && (!Din || !Iin || !D || !I))) {
        set_error("b2k_merge_topk: bad argument");
        return B2K_EINVAL;
    }
    if ((int64_t)nlists * k > 8192) { set_error("b2k_merge_topk: nlists*k = %lld exceeds 8192", (long long)nlists * k); return B2K_ELIMIT; }
    if (nq == 0) return B2K_OK;
    int rc = check_device(device, nullptr);
    if (rc) return rc;
    if (device >= 64) { set_error("b2k_merge_topk: device %d out of range", device); return B2K_EINVAL; }
    DeviceGuard g(device);
    cudaStream_t st = (cudaStream_t)stream;
    size_t in_n = (size_t)nlists * nq * k, out_n = (size_t)nq * k;
    int64_t* off_d = nullptr;
    float *Din_d = const_cast<float*>(Din), *D_d = D;
    int64_t *Iin_d = const_cast<int64_t*>(Iin), *I_d = I;
    size_t off_bytes = align256(id_offset ? (size_t)nlists * 8 : 0);
    size_t a = align256(in_n * 4), b = align256(in_n * 8), c = align256(out_n * 4), e2 = align256(out_n * 8);
    size_t need = off_bytes + (on_device ? 0 : a + b + c + e2);
    if (!need) return launch_merge_lists(nlists, nq, (int)k, Din_d, Iin_d, nullptr, D_d, I_d, st);
    MergeScratch& ms = g_merge_scratch[device];
    std::lock_guard<std::mutex> lk(ms.mu);
    if (need > ms.bytes) {
        if (ms.p) { cudaDeviceSynchronize(); cudaFree(ms.p); ms.p = nullptr; ms.bytes = 0; }
        cudaError_t e = cudaMalloc(&ms.p, need + need / 2);
        if (e != cudaSuccess) { cudaGetLastError(); set_error("b2k_merge_topk: cudaMalloc failed: %s", cudaGetErrorString(e)); return B2K_ENOMEM; }
        ms.bytes = need + need / 2;
    }
    unsigned char* tmp = static_cast<unsigned char*>(ms.p);
    cudaError_t e = cudaSuccess;
    if (id_offset) {
        off_d = reinterpret_cast<int64_t*>(tmp);
        e = cudaMemcpyAsync(off_d, id_offset, (size_t)nlists * 8, cudaMemcpyHostToDevice, st);
    }
    if (e == cudaSuccess && !on_device) {
        Din_d = reinterpret_cast<float*>(tmp + off_bytes);
        Iin_d = reinterpret_cast<int64_t*>(tmp + off_bytes + a);
        D_d = reinterpret_cast<float*>(tmp + off_bytes + a + b);
        I_d = reinterpret_cast<int64_t*>(tmp + off_bytes + a + b + c);
        e = cudaMemcpyAsync(Din_d, Din, in_n * 4, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(Iin_d, Iin, in_n * 8, cudaMemcpyHostToDevice, st);
    }
    if (e != cudaSuccess) { set_error("b2k_merge_topk: %s", cudaGetErrorString(e)); return B2K_ECUDA; }
    rc = launch_merge_lists(nlists, nq, (int)k, Din_d, Iin_d, off_d, D_d, I_d, st);
    if (rc) return rc;
    if (!on_device) {
        e = cudaMemcpyAsync(D, D_d, out_n * 4, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(I, I_d, out_n * 8, cudaMemcpyDeviceToHost, st);
    }
    // the scratch is shared: the next user may overwrite it, so this call's work is finished before the lock goes
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { set_error("b2k_merge_topk: %s", cudaGetErrorString(e)); return B2K_ECUDA; }
    return B2K_OK;
}

int b2k_merge_shards(int device, int nlists, int64_t nq, int64_t k, const float* Din, int64_t list_stride_d,
                     const int64_t* Iin, int64_t list_stride_i, int64_t cyclic_block, float* D, int64_t* I, void* stream) {
    if (nlists <= 0 || nq < 0 || k <= 0 || cyclic_block < 0 || (nq > 0 && (!Din || !Iin || !D || !I))) {
        set_error("b2k_merge_shards: bad argument");
        return B2K_EINVAL;
    }
    if ((int64_t)nlists * k > 8192) { set_error("b2k_merge_shards: nlists*k = %lld exceeds 8192", (long long)nlists * k); return B2K_ELIMIT; }
    if (nq == 0) return B2K_OK;
    int rc = check_device(device, nullptr);
    if (rc) return rc;
    DeviceGuard g(device);
    return launch_merge_shards(nlists, nq, (int)k, Din, list_stride_d, Iin, list_stride_i, cyclic_block, D, I, (cudaStream_t)stream);
}

int b2k_exchange_push(int device, const void* record, int64_t bytes, const uint64_t* dst_ptrs_dev, const uint64_t* flag_ptrs_dev,
                      int nranks, int my_rank, int64_t slot_bytes, int flag_index, uint32_t seq, uint32_t* done_dev, void* stream) {
    if (!record || bytes <= 0 || !dst_ptrs_dev || !flag_ptrs_dev || nranks <= 0 || nranks > 64 || my_rank < 0 || my_rank >= nranks || !done_dev) {
        set_error("b2k_exchange_push: bad argument");
        return B2K_EINVAL;
    }
    int rc = check_device(device, nullptr);
    if (rc) return rc;
    DeviceGuard g(device);
    return launch_push_record(record, bytes, dst_ptrs_dev, flag_ptrs_dev, nranks, my_rank, slot_bytes, flag_index, seq, done_dev, (cudaStream_t)stream);
}

int b2k_exchange_wait(int device, const uint32_t* flags_dev, int nranks, uint32_t seq, uint32_t* err_dev, void* stream) {
    if (!flags_dev || nranks <= 0 || nranks > 64 || !err_dev) { set_error("b2k_exchange_wait: bad argument"); return B2K_EINVAL; }
    int rc = check_device(device, nullptr);
    if (rc) return rc;
    DeviceGuard g(device);
    return launch_wait_flags(flags_dev, nranks, seq, err_dev, (cudaStream_t)stream);
}

// ---- IxF2 persistence ------------------------------------------------------------------
// Layout (little-endian, packed) [faiss-upstream index_write.cpp, SURVEY.md 8a-a5]:
//   "IxF2" | d int32 | ntotal int64 | 1<<20 int64 | 1<<20 int64 | is_trained u8 (1) |
//   metric_type int32 (1 = L2) | count u64 (= ntotal*d) | count float32

// rows [row0, row0 + nrows) of the handle -> f, through the pinned staging buffer
static int write_rows(b2k_index* h, FILE* f, int64_t row0, int64_t nrows, const char* what) {
    const size_t row_bytes = (size_t)h->d * sizeof(float);
    size_t total = (size_t)nrows * row_bytes, done = 0;
    if (!total) return B2K_OK;
    int rc = ensure_pin(h, total < IO_CHUNK ? total : IO_CHUNK);
    if (rc) return rc;
    std::shared_lock<std::shared_mutex> sl(h->rs->mu);
    B2K_CUDA(cudaStreamWaitEvent(h->stream, h->rs->ev_rows, 0));
    const unsigned char* src = reinterpret_cast<const unsigned char*>(h->rs->x) + (size_t)row0 * row_bytes;
    while (done < total) {
        size_t n = total - done < IO_CHUNK ? total - done : IO_CHUNK;
        cudaError_t e = cudaMemcpyAsync(h->pin, src + done, n, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { set_error("%s: %s", what, cudaGetErrorString(e)); return B2K_ECUDA; }
        if (fwrite(h->pin, 1, n, f) != n) { set_error("%s: write error: %s", what, strerror(errno)); return B2K_EIO; }
        done += n;
    }
    return B2K_OK;
}

int b2k_write(const b2k_index* hc, const char* path) {
    b2k_index* h = const_cast<b2k_index*>(hc);
    if (!h || !path) { set_error("b2k_write: null argument"); return B2K_EINVAL; }
    if (h->multi) return multi_write(h, path, 0, false);
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    std::string tmp = std::string(path) + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) { set_error("could not open %s for writing: %s", tmp.c_str(), strerror(errno)); return B2K_EIO; }
    unsigned char hdr[IXF2_HEADER];
    ixf2_put_header(hdr, h->d, h->ntotal);
    int rc = B2K_OK;
    if (fwrite(hdr, 1, IXF2_HEADER, f) != IXF2_HEADER) { set_error("error writing %s: %s", tmp.c_str(), strerror(errno)); rc = B2K_EIO; }
    if (!rc) rc = write_rows(h, f, 0, h->ntotal, "b2k_write");
    if (fclose(f) != 0 && !rc) { set_error("error writing %s: %s", tmp.c_str(), strerror(errno)); rc = B2K_EIO; }
    if (!rc && rename(tmp.c_str(), path) != 0) { set_error("could not rename %s to %s: %s", tmp.c_str(), path, strerror(errno)); rc = B2K_EIO; }
    if (rc) remove(tmp.c_str());
    return rc;
}

int b2k_write_append(const b2k_index* hc, const char* path, int64_t first_row) {
    b2k_index* h = const_cast<b2k_index*>(hc);
    if (!h || !path || first_row < 0) { set_error("b2k_write_append: bad argument"); return B2K_EINVAL; }
    if (h->multi) return multi_write(h, path, first_row, true);
    std::lock_guard<std::mutex> lk(h->mu);
    if (first_row > h->ntotal) { set_error("b2k_write_append: first_row %lld > ntotal %lld", (long long)first_row, (long long)h->ntotal); return B2K_EINVAL; }
    DeviceGuard g(h->device);
    FILE* f = fopen(path, "r+b");
    if (!f) { set_error("could not open %s for appending: %s", path, strerror(errno)); return B2K_EIO; }
    unsigned char hdr[IXF2_HEADER], want[IXF2_HEADER];
    ixf2_put_header(want, h->d, first_row);
    int rc = B2K_OK;
    if (fread(hdr, 1, IXF2_HEADER, f) != IXF2_HEADER || memcmp(hdr, want, IXF2_HEADER) != 0) {
        set_error("%s does not hold exactly the first %lld rows of a d=%d IxF2 index", path, (long long)first_row, h->d);
        rc = B2K_EINVAL;
    }
    const size_t row_bytes = (size_t)h->d * sizeof(float);
    if (!rc) {
        if (fseeko(f, 0, SEEK_END) != 0 || (size_t)ftello(f) != IXF2_HEADER + (size_t)first_row * row_bytes) {
            set_error("%s: size does not match its header", path);
            rc = B2K_EINVAL;
        }
    }
    if (!rc) rc = write_rows(h, f, first_row, h->ntotal - first_row, "b2k_write_append");
    if (!rc) {
        // rows first, header last: a reader that opens the file in between still sees a consistent
        // (old) header and ignores the extra bytes behind the payload it describes
        ixf2_put_header(want, h->d, h->ntotal);
        if (fflush(f) != 0 || fseeko(f, 0, SEEK_SET) != 0 || fwrite(want, 1, IXF2_HEADER, f) != IXF2_HEADER) {
            set_error("error updating the header of %s: %s", path, strerror(errno));
            rc = B2K_EIO;
        }
    }
    if (fclose(f) != 0 && !rc) { set_error("error closing %s: %s", path, strerror(errno)); rc = B2K_EIO; }
    return rc;
}

int b2k_read(const char* path, int device, b2k_index** out) {
    if (!path || !out) { set_error("b2k_read: null argument"); return B2K_EINVAL; }
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) { set_error("could not open %s for reading: %s", path, strerror(errno)); return B2K_EIO; }
    unsigned char hdr[IXF2_HEADER];
    if (fread(hdr, 1, IXF2_HEADER, f) != IXF2_HEADER) { fclose(f); set_error("%s: short read of the index header", path); return B2K_EIO; }
    int d32 = 0; int64_t n = 0;
    int rc = ixf2_parse_header(hdr, path, &d32, &n);
    if (rc) { fclose(f); return rc; }
    b2k_index* h = nullptr;
    rc = b2k_create(d32, device, &h);
    if (rc) { fclose(f); return rc; }
    {
        std::lock_guard<std::mutex> lk(h->mu);
        DeviceGuard g(h->device);
        RowStore* rs = h->rs.get();
        std::unique_lock<std::shared_mutex> sl(rs->mu);
        size_t total = (size_t)n * d32 * sizeof(float), done = 0;
        if (n > 0) rc = grow(rs, n, h->stream);
        if (!rc && total) {
            // two pinned halves: the H2D copy of one overlaps the fread of the other
            size_t half = total < IO_CHUNK ? total : IO_CHUNK;
            rc = ensure_pin(h, 2 * half);
            int which = 0;
            cudaEvent_t ev[2] = {nullptr, nullptr};
            if (!rc && (cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming) != cudaSuccess ||
                        cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming) != cudaSuccess)) {
                set_error("b2k_read: cudaEventCreate failed"); rc = B2K_ECUDA;
            }
            while (!rc && done < total) {
                size_t nb = total - done < half ? total - done : half;
                unsigned char* buf = static_cast<unsigned char*>(h->pin) + (size_t)which * half;
                cudaEventSynchronize(ev[which]);
                if (fread(buf, 1, nb, f) != nb) { set_error("%s: truncated IxF2 payload", path); rc = B2K_EIO; break; }
                cudaError_t e = cudaMemcpyAsync(reinterpret_cast<unsigned char*>(rs->x) + done, buf, nb, cudaMemcpyHostToDevice, h->stream);
                if (e == cudaSuccess) e = cudaEventRecord(ev[which], h->stream);
                if (e != cudaSuccess) { set_error("b2k_read: %s", cudaGetErrorString(e)); rc = B2K_ECUDA; break; }
                done += nb;
                which ^= 1;
            }
            if (!rc) rc = prep_new_rows(rs, 0, n, h->stream);
            if (!rc && cudaEventRecord(rs->ev_rows, h->stream) != cudaSuccess) { set_error("b2k_read: cudaEventRecord failed"); rc = B2K_ECUDA; }
            cudaError_t e = cudaStreamSynchronize(h->stream);
            if (!rc && e != cudaSuccess) { set_error("b2k_read: %s", cudaGetErrorString(e)); rc = B2K_ECUDA; }
            if (ev[0]) cudaEventDestroy(ev[0]);
            if (ev[1]) cudaEventDestroy(ev[1]);
        }
        if (!rc) { rs->filled = n; h->ntotal = n; }
    }
    fclose(f);
    if (rc) { b2k_destroy(h); return rc; }
    *out = h;
    return B2K_OK;
}

int b2k_read_multi(const char* path, const int* devices, int ndev, b2k_index** out) {
    if (!path || !out) { set_error("b2k_read_multi: null argument"); return B2K_EINVAL; }
    *out = nullptr;
    if (!devices || ndev <= 0) { set_error("b2k_read_multi: empty device list"); return B2K_EINVAL; }
    if (ndev == 1) return b2k_read(path, devices[0], out);
    return multi_read(path, devices, ndev, out);
}

}  // extern "C"
