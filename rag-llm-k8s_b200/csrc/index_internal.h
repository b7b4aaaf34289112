// index_internal.h -- what index.cu and multi.cu share: the row store, the handle, and the few
// internal entry points the multi-GPU handle is built from.
//
// Ownership model.  The rows of an index live in a RowStore (HBM buffers + the per-row data the
// kernels derive from them).  A handle (b2k_index) is a VIEW of the first `ntotal` rows of a store
// plus its own stream, workspace and options.  Several handles may share one store
// (b2k_view): llm/rag.py calls faiss.read_index on every request (rag.py:153, :186) and expects
// an independent index each time; with views each request gets its own handle for the price of a
// stream, while the rows stay resident once.  A handle that appends while it is the only one at
// the end of the store appends in place (other views keep seeing their own prefix); a handle that
// appends after the store has moved on first takes a private copy of its prefix.
#pragma once
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <vector>

#include "common.cuh"
#include "k3_api.h"

namespace b2k {

struct RowStore {
    int d = 0, device = 0;
    int64_t filled = 0, capacity = 0;  // rows present / allocated
    float* x = nullptr;                // [capacity][d] fp32, the rows as given       (K2, K3 "hi" operand)
    float* norms = nullptr;            // [capacity]    ||x||^2, K1 at add()          (K3 epilogue)
    float* xlo = nullptr;              // [capacity][d] TF32 residual rows, once a 3xTF32 mode has run
    void* xbf = nullptr;               // [capacity][d] bf16 copies, once a screening pass / bf16 scan has run
    void* xnb = nullptr;               // [capacity][16] bf16: -||x||^2/2 as three bf16 terms, next to xbf (K3 adds it by MMA)
    bool lo_enabled = false, bf_enabled = false, bf_nomem = false;
    unsigned int* stat = nullptr;      // device words: [0] bits of max ||x||^2, [1] bits of max ||x - bf16(x)||^2
    // outlier rows (k1_prep.cu: launch_outlier_pass): [0] count, [1] / [2] the norm / rounding-error caps the norm-folded
    // screen uses once those rows are out of it, [8..72) their ids; recomputed lazily after the rows changed
    unsigned int* outl = nullptr;
    unsigned int* outl_hist = nullptr;
    bool outl_dirty = true;
    cudaEvent_t ev_rows = nullptr;     // last write to any of the buffers above
    // shared: enqueueing work that reads the buffers; exclusive: add / grow / creating derived copies
    std::shared_mutex mu;
    ~RowStore();
};

struct Multi;

}  // namespace b2k

struct b2k_index {
    int d = 0, device = 0, sm_count = 148;
    int64_t ntotal = 0;                      // rows of the store this handle sees
    std::shared_ptr<b2k::RowStore> rs;
    b2k::Multi* multi = nullptr;             // set: this handle is a row-sharded index over several GPUs (multi.cu)
    cudaStream_t stream = nullptr;
    mutable std::mutex mu;
    void* ws = nullptr;
    size_t ws_bytes = 0;
    void* pin = nullptr;
    size_t pin_bytes = 0;
    unsigned int* done = nullptr;            // device words: [0] scan ticket, [1] scan work counter, [4] bf16-scan fallback flag
    int opt_path = 0;
    int opt_scan_ctas_per_sm = 4;  // upper bound; each variant is clamped to its own residency (64 regs x 256 threads -> 4)
    int opt_timing = 0;
    int opt_cta_group = 0;  // 0 = library default
    int opt_sep = 0;
    int opt_segments = 1;
    unsigned int* opt_flag_ptr = nullptr;
    int opt_scan_rows = 0;  // nq < 20 scan: 0 auto (bf16 copies + exact re-rank for large indexes), 1 fp32 rows, 2 bf16 copies
    bool last_scan_bf = false;
    int opt_epi = 0;
    int opt_probe = 0;
    int opt_qs = 0;
    int opt_pf = 0;
    int opt_warm = 12;
    int opt_tensor_mode = 0;  // 0 auto (bf16 screening + exact re-rank where d % 8 == 0 and memory allows, else TF32 screening), 1 3xTF32, 2 TF32 screening, 3 bf16 screening, 4 3xTF32 screening
    int64_t stat_screen_fallbacks = 0;
    int screen_fail_streak = 0;      // consecutive searches whose cheap screen overflowed
    unsigned int screen_skipped = 0;  // searches that started at the 3xTF32 screen because of that
    int last_path = 0;
    int sticky = 0;
    b2k::K3State* k3 = nullptr;
    cudaEvent_t ev_ws = nullptr;                   // workspace users -> next user (searches may come on different streams)
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;  // option timing: around the scan launches
    int scan_launches = 0;
    bool scan_timed = false;
};

namespace b2k {

struct DeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; }
        ok = (cudaSetDevice(dev) == cudaSuccess);
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

int check_device(int device, int* sm_count);
bool is_pinned_host(const void* p);
inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

// IxF2 header helpers (index.cu)
static const size_t IXF2_HEADER = 45;
void ixf2_put_header(unsigned char* b, int d, int64_t n);
// parses and validates; returns B2K_OK or B2K_EIO with the message set
int ixf2_parse_header(const unsigned char* hdr, const char* path, int* d, int64_t* n);

// multi.cu: the row-sharded handle behind the same C ABI
int multi_create(int d, const int* devices, int ndev, b2k_index** out);
void multi_destroy(b2k_index* h);
int multi_add(b2k_index* h, int64_t n, const float* x, int on_device, void* stream);
int multi_add_synthetic(b2k_index* h, int64_t n, uint64_t seed, int64_t first_row);
int multi_search(b2k_index* h, int64_t nq, const float* xq, int64_t k, float* D, int64_t* I, int on_device, void* stream);
int multi_reset(b2k_index* h);
int multi_reserve(b2k_index* h, int64_t nrows);
int multi_get_rows(const b2k_index* h, int64_t i0, int64_t n, float* out, int on_device);
int multi_write(const b2k_index* h, const char* path, int64_t first_row, bool append);
int multi_read(const char* path, const int* devices, int ndev, b2k_index** out);
int multi_view(const b2k_index* h, int64_t nrows, b2k_index** out);
int multi_set_option(b2k_index* h, const char* key, int64_t value);
int multi_get_stat(b2k_index* h, const char* key, double* out);

}  // namespace b2k
