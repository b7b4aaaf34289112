// multi.cu -- one IndexFlatL2 over several GPUs of the node, driven from ONE process and thread,
// behind the same C ABI (b2k_create_multi / b2k_read_multi; every other entry point takes the
// handle as it takes a single-GPU one).
//
// Why it exists: llm/rag.py is one Flask process (rag.py:204, replicas: 1 in llm/ragdeploy.yaml:6);
// an index that outgrows one B200 must still answer `index.search` (rag.py:116) from that process.
// The reference has no counterpart (one process, one faiss index in host memory); the behaviour to
// reproduce is that of ONE IndexFlatL2 holding all rows: same ids, same (distance, id) order.
//
// Layout.  Global rows go to the shards in blocks of CYC_BLOCK rows, round-robin: block b lives on
// shard b % S as its local block b / S.  The map is pure arithmetic in both directions, so there is
// no id table, incremental adds of any size keep the shards balanced to within one block, and the
// IxF2 file (global row order) can be streamed to and from the shards block by block.
//
// Search.  One worker thread per device enqueues that shard's search (the single-GPU path: K2 / K3
// + re-rank) on the shard's stream; the shard kernels write their k best STRAIGHT INTO the packed
// [S][nq][k] buffer on the first device through peer memory (plain stores over NVLink -- the
// exchange step costs no launch and no copy; without peer access a peer memcpy stands in), the
// first device waits for S events and runs the merge kernel (K5), which also turns shard-local row
// numbers into global ones.  Queries reach every device by its own H2D copy from pinned host memory
// (each GPU has its own PCIe link), or by a peer copy when they live on the first device.
// Every TMA tensor map the shard kernels build covers the shard's own memory only.
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <condition_variable>
#include <deque>
#include <functional>
#include <string.h>
#include <string>
#include <thread>

#include "index_internal.h"
#include "kernels.h"

namespace b2k {

static constexpr int64_t CYC_BLOCK = 4096;   // rows per block of the block-cyclic layout (a multiple of K3's 256-row tile)
static constexpr size_t MULTI_IO_CHUNK = (size_t)64 << 20;

// ---- one worker thread per device, for the whole process ---------------------------------
struct Task {
    std::function<int()> fn;
    int rc = 0;
    std::string err;
    bool done = false;
    std::mutex mu;
    std::condition_variable cv;
};

struct DevWorker {
    int device = 0;
    std::mutex mu;
    std::condition_variable cv;
    std::deque<Task*> q;
    void loop() {
        cudaSetDevice(device);
        for (;;) {
            Task* t;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return !q.empty(); });
                t = q.front();
                q.pop_front();
            }
            int rc = t->fn();
            std::string err = rc ? b2k_last_error() : "";
            {
                std::lock_guard<std::mutex> lk(t->mu);
                t->rc = rc;
                t->err = err;
                t->done = true;
            }
            t->cv.notify_one();
        }
    }
};

static DevWorker* worker_for(int device) {
    static std::mutex mu;
    static DevWorker* w[64] = {nullptr};
    std::lock_guard<std::mutex> lk(mu);
    if (!w[device]) {
        w[device] = new DevWorker();  // lives until the process exits, like the thread
        w[device]->device = device;
        std::thread(&DevWorker::loop, w[device]).detach();
    }
    return w[device];
}

struct Multi {
    int d = 0, ndev = 0;
    std::vector<int> dev;
    std::vector<b2k_index*> shard;
    std::mutex mu;
    bool peer_ok = false;       // every device can store into the first one's memory
    int opt_peer_writes = 1;    // 0: shards keep their results locally and a peer memcpy moves them (test hook, and the no-P2P path)
    int opt_tensor_mode = 0, opt_path = 0;
    int64_t redo_count = 0;
    // on the first device
    cudaStream_t st0 = nullptr;
    cudaEvent_t ev_q = nullptr;
    void* ws0 = nullptr;
    size_t ws0_bytes = 0;
    void* pin = nullptr;
    size_t pin_bytes = 0;
    unsigned int* flags_host = nullptr;  // pinned, one word per shard
    // per shard, on its device
    std::vector<unsigned int*> flag_dev;
    std::vector<void*> qbuf, obuf;
    std::vector<size_t> qbuf_bytes, obuf_bytes;
    std::vector<cudaEvent_t> ev_done;
};

namespace {

int64_t rows_of_shard(int s, int S, int64_t n) {
    const int64_t cyc = CYC_BLOCK * S;
    int64_t r = n % cyc - (int64_t)s * CYC_BLOCK;
    r = r < 0 ? 0 : (r > CYC_BLOCK ? CYC_BLOCK : r);
    return (n / cyc) * CYC_BLOCK + r;
}
inline int shard_of(int64_t g, int S) { return (int)((g / CYC_BLOCK) % S); }
inline int64_t local_row(int64_t g, int S) { return (g / CYC_BLOCK / S) * CYC_BLOCK + g % CYC_BLOCK; }

// fn(s) on the worker of every shard's device, concurrently; first failure wins
int run_on_all(Multi* m, const std::function<int(int)>& fn) {
    std::vector<Task> tasks(m->ndev);
    for (int s = 0; s < m->ndev; ++s) {
        tasks[s].fn = [s, &fn] { return fn(s); };
        DevWorker* w = worker_for(m->dev[s]);
        {
            std::lock_guard<std::mutex> lk(w->mu);
            w->q.push_back(&tasks[s]);
        }
        w->cv.notify_one();
    }
    int rc = B2K_OK;
    std::string err;
    for (int s = 0; s < m->ndev; ++s) {
        std::unique_lock<std::mutex> lk(tasks[s].mu);
        tasks[s].cv.wait(lk, [&] { return tasks[s].done; });
        if (tasks[s].rc && !rc) { rc = tasks[s].rc; err = tasks[s].err; }
    }
    if (rc) set_error("device %s", err.c_str());
    return rc;
}

int grow_dev_buffer(void** p, size_t* have, size_t need) {
    if (need <= *have) return B2K_OK;
    if (*p) { cudaFree(*p); *p = nullptr; *have = 0; }  // cudaFree synchronises
    const size_t want = need + need / 4 + 4096;
    cudaError_t e = cudaMalloc(p, want);
    if (e != cudaSuccess) { cudaGetLastError(); set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e)); return B2K_ENOMEM; }
    *have = want;
    return B2K_OK;
}

int ensure_multi_pin(Multi* m, size_t bytes) {
    if (bytes <= m->pin_bytes) return B2K_OK;
    if (m->pin) { cudaFreeHost(m->pin); m->pin = nullptr; m->pin_bytes = 0; }
    const size_t want = bytes + bytes / 4 + 4096;
    cudaError_t e = cudaMallocHost(&m->pin, want);
    if (e != cudaSuccess) { cudaGetLastError(); set_error("cudaMallocHost(%zu) failed: %s", want, cudaGetErrorString(e)); return B2K_ENOMEM; }
    m->pin_bytes = want;
    return B2K_OK;
}

void free_multi(Multi* m) {
    for (int s = 0; s < (int)m->shard.size(); ++s) {
        DeviceGuard g(m->dev[s]);
        if (m->shard[s]) b2k_destroy(m->shard[s]);
        if (s < (int)m->flag_dev.size()) cudaFree(m->flag_dev[s]);
        if (s < (int)m->qbuf.size()) cudaFree(m->qbuf[s]);
        if (s < (int)m->obuf.size()) cudaFree(m->obuf[s]);
        if (s < (int)m->ev_done.size() && m->ev_done[s]) cudaEventDestroy(m->ev_done[s]);
    }
    if (!m->dev.empty()) {
        DeviceGuard g(m->dev[0]);
        if (m->st0) { cudaStreamSynchronize(m->st0); cudaStreamDestroy(m->st0); }
        if (m->ev_q) cudaEventDestroy(m->ev_q);
        cudaFree(m->ws0);
        if (m->pin) cudaFreeHost(m->pin);
        if (m->flags_host) cudaFreeHost(m->flags_host);
    }
    delete m;
}

// Everything of a Multi except the shards' rows: per-shard flag words, events, peer access, first-device stream.
int setup_multi(Multi* m) {
    const int S = m->ndev;
    m->flag_dev.assign(S, nullptr);
    m->qbuf.assign(S, nullptr); m->obuf.assign(S, nullptr);
    m->qbuf_bytes.assign(S, 0); m->obuf_bytes.assign(S, 0);
    m->ev_done.assign(S, nullptr);
    m->peer_ok = true;
    for (int s = 0; s < S; ++s) {
        DeviceGuard g(m->dev[s]);
        B2K_CUDA(cudaMalloc(reinterpret_cast<void**>(&m->flag_dev[s]), 256));
        B2K_CUDA(cudaMemset(m->flag_dev[s], 0, 256));
        B2K_CUDA(cudaEventCreateWithFlags(&m->ev_done[s], cudaEventDisableTiming));
        int rc = b2k_set_option(m->shard[s], "screen_flag_ptr", (int64_t)reinterpret_cast<uintptr_t>(m->flag_dev[s]));
        if (rc) return rc;
        if (s > 0 && m->dev[s] != m->dev[0]) {
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, m->dev[s], m->dev[0]) != cudaSuccess) { cudaGetLastError(); can = 0; }
            if (can) {
                cudaError_t e = cudaDeviceEnablePeerAccess(m->dev[0], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
                if (e != cudaSuccess) { cudaGetLastError(); can = 0; }
            }
            if (!can) m->peer_ok = false;
        }
    }
    DeviceGuard g(m->dev[0]);
    B2K_CUDA(cudaStreamCreateWithFlags(&m->st0, cudaStreamNonBlocking));
    B2K_CUDA(cudaEventCreateWithFlags(&m->ev_q, cudaEventDisableTiming));
    B2K_CUDA(cudaMallocHost(reinterpret_cast<void**>(&m->flags_host), 64 * sizeof(unsigned int)));
    memset(m->flags_host, 0, 64 * sizeof(unsigned int));
    return B2K_OK;
}

int wrap_multi(Multi* m, int64_t ntotal, b2k_index** out) {
    b2k_index* h = new (std::nothrow) b2k_index();
    if (!h) { set_error("out of host memory"); return B2K_ENOMEM; }
    h->multi = m;
    h->d = m->d;
    h->device = m->dev[0];
    h->ntotal = ntotal;
    *out = h;
    return B2K_OK;
}

int check_devices(const int* devices, int ndev) {
    if (ndev > 64) { set_error("at most 64 devices"); return B2K_EINVAL; }
    for (int i = 0; i < ndev; ++i) {
        int rc = check_device(devices[i], nullptr);
        if (rc) return rc;
        if (devices[i] >= 64) { set_error("device %d out of range", devices[i]); return B2K_EINVAL; }
        // (test hook: B200KNN_TEST_DUP_DEVICES=1 lets a device appear twice, so that a one-GPU box can run the
        // sharded code path with several shards on the same GPU)
        const char* dup = getenv("B200KNN_TEST_DUP_DEVICES");
        for (int j = 0; j < i; ++j)
            if (devices[j] == devices[i] && !(dup && dup[0] == '1')) { set_error("device %d listed twice", devices[i]); return B2K_EINVAL; }
    }
    return B2K_OK;
}

// rows [N0, N0+n) of the global order, block by block: fn(shard, global first row, count)
template <class F>
int for_each_piece(int S, int64_t N0, int64_t n, int only_shard, F fn) {
    for (int64_t g = N0; g < N0 + n;) {
        const int64_t bend = (g / CYC_BLOCK + 1) * CYC_BLOCK;
        const int64_t e = bend < N0 + n ? bend : N0 + n;
        const int s = shard_of(g, S);
        if (only_shard < 0 || s == only_shard) {
            int rc = fn(s, g, e - g);
            if (rc) return rc;
        }
        g = e;
    }
    return B2K_OK;
}

// appends global rows [ntotal, ntotal+n): x (host, or device memory of the first GPU), or the synthetic stream if x == null
int add_rows(b2k_index* h, int64_t n, const float* x, int on_device, cudaStream_t caller, uint64_t seed, int64_t synth_first) {
    Multi* m = h->multi;
    const int S = m->ndev, d = m->d;
    const int64_t N0 = h->ntotal;
    if (x && on_device) {
        DeviceGuard g(m->dev[0]);
        B2K_CUDA(cudaEventRecord(m->ev_q, caller));  // the rows are ready on the caller's stream from here on
    }
    int rc = run_on_all(m, [&](int s) -> int {
        b2k_index* sh = m->shard[s];
        const int64_t cnt = rows_of_shard(s, S, N0 + n) - rows_of_shard(s, S, N0);
        if (cnt == 0) return B2K_OK;
        int r = b2k_reserve(sh, sh->ntotal + cnt);
        if (r) return r;
        if (x && on_device) B2K_CUDA(cudaStreamWaitEvent(sh->stream, m->ev_q, 0));
        r = for_each_piece(S, N0, n, s, [&](int, int64_t g, int64_t c) -> int {
            if (!x) return b2k_add_synthetic(sh, c, seed, synth_first + (g - N0));
            return b2k_add(sh, c, x + (g - N0) * (int64_t)d, on_device, on_device ? sh->stream : nullptr);
        });
        if (r) return r;
        if (x && on_device) B2K_CUDA(cudaStreamSynchronize(sh->stream));  // the caller may reuse x once we return
        return B2K_OK;
    });
    if (rc) return rc;
    h->ntotal = N0 + n;
    return B2K_OK;
}

int get_rows(Multi* m, int64_t i0, int64_t n, float* out, int on_device) {
    return for_each_piece(m->ndev, i0, n, -1, [&](int s, int64_t g, int64_t c) -> int {
        return b2k_get_rows(m->shard[s], local_row(g, m->ndev), c, out + (g - i0) * (int64_t)m->d, on_device);
    });
}

// one attempt of the sharded search; *flagged = some shard's screening lists overflowed (results then unusable)
int search_once(b2k_index* h, int64_t nq, const float* xq, int k, float* D, int64_t* I, int on_device, cudaStream_t caller,
                bool* flagged) {
    Multi* m = h->multi;
    const int S = m->ndev, d = m->d;
    const size_t nk = (size_t)nq * k;
    const size_t q_bytes = (size_t)nq * d * sizeof(float);
    const size_t Dp_bytes = align256((size_t)S * nk * 4), Ip_bytes = align256((size_t)S * nk * 8);
    const size_t Do_bytes = align256(nk * 4), Io_bytes = align256(nk * 8);
    DeviceGuard g0(m->dev[0]);
    int rc = grow_dev_buffer(&m->ws0, &m->ws0_bytes, Dp_bytes + Ip_bytes + Do_bytes + Io_bytes);
    if (rc) return rc;
    unsigned char* w = static_cast<unsigned char*>(m->ws0);
    float* Dp = reinterpret_cast<float*>(w);
    int64_t* Ip = reinterpret_cast<int64_t*>(w + Dp_bytes);
    float* Dd = on_device ? D : reinterpret_cast<float*>(w + Dp_bytes + Ip_bytes);
    int64_t* Id = on_device ? I : reinterpret_cast<int64_t*>(w + Dp_bytes + Ip_bytes + Do_bytes);
    cudaStream_t stm = on_device ? caller : m->st0;

    const void* q_src = xq;
    bool out_pinned = false;
    if (!on_device) {
        out_pinned = is_pinned_host(D) && is_pinned_host(I);
        const bool q_pinned = is_pinned_host(xq);
        rc = ensure_multi_pin(m, align256(q_bytes) + Do_bytes + Io_bytes);
        if (rc) return rc;
        if (!q_pinned) { memcpy(m->pin, xq, q_bytes); q_src = m->pin; }
    } else {
        B2K_CUDA(cudaEventRecord(m->ev_q, caller));
    }
    const bool direct = m->peer_ok && m->opt_peer_writes;

    rc = run_on_all(m, [&](int s) -> int {
        b2k_index* sh = m->shard[s];
        cudaStream_t st = sh->stream;
        int r = grow_dev_buffer(&m->qbuf[s], &m->qbuf_bytes[s], q_bytes);
        if (r) return r;
        if (on_device) B2K_CUDA(cudaStreamWaitEvent(st, m->ev_q, 0));
        B2K_CUDA(cudaMemcpyAsync(m->qbuf[s], q_src, q_bytes, on_device ? cudaMemcpyDefault : cudaMemcpyHostToDevice, st));
        B2K_CUDA(cudaMemsetAsync(m->flag_dev[s], 0, sizeof(unsigned int), st));
        float* Do = Dp + (size_t)s * nk;
        int64_t* Io = Ip + (size_t)s * nk;
        const bool local_out = !direct && s > 0;
        if (local_out) {
            r = grow_dev_buffer(&m->obuf[s], &m->obuf_bytes[s], Do_bytes + Io_bytes);
            if (r) return r;
            Do = reinterpret_cast<float*>(m->obuf[s]);
            Io = reinterpret_cast<int64_t*>(static_cast<unsigned char*>(m->obuf[s]) + Do_bytes);
        }
        // the shard's kernels write their k best where Do / Io point: with peer access that is the
        // packed buffer on the first GPU itself
        r = b2k_search(sh, nq, static_cast<const float*>(m->qbuf[s]), k, Do, Io, 1, st);
        if (r) return r;
        if (local_out) {
            B2K_CUDA(cudaMemcpyPeerAsync(Dp + (size_t)s * nk, m->dev[0], Do, m->dev[s], nk * 4, st));
            B2K_CUDA(cudaMemcpyPeerAsync(Ip + (size_t)s * nk, m->dev[0], Io, m->dev[s], nk * 8, st));
        }
        B2K_CUDA(cudaMemcpyAsync(m->flags_host + s, m->flag_dev[s], sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
        B2K_CUDA(cudaEventRecord(m->ev_done[s], st));
        return B2K_OK;
    });
    if (rc) return rc;
    for (int s = 0; s < S; ++s) B2K_CUDA(cudaStreamWaitEvent(stm, m->ev_done[s], 0));
    rc = launch_merge_shards(S, nq, k, Dp, (int64_t)nk, Ip, (int64_t)nk, CYC_BLOCK, Dd, Id, stm);
    if (rc) return rc;
    unsigned char* pb = static_cast<unsigned char*>(m->pin);
    if (!on_device) {
        B2K_CUDA(cudaMemcpyAsync(out_pinned ? (void*)D : (void*)(pb + align256(q_bytes)), Dd, nk * 4, cudaMemcpyDeviceToHost, stm));
        B2K_CUDA(cudaMemcpyAsync(out_pinned ? (void*)I : (void*)(pb + align256(q_bytes) + Do_bytes), Id, nk * 8, cudaMemcpyDeviceToHost, stm));
    }
    // the screening words of all shards are needed before the answer may be used: one synchronisation per search
    B2K_CUDA(cudaStreamSynchronize(stm));
    *flagged = false;
    for (int s = 0; s < S; ++s) *flagged |= m->flags_host[s] != 0;
    if (!on_device && !out_pinned && !*flagged) {
        memcpy(D, pb + align256(q_bytes), nk * 4);
        memcpy(I, pb + align256(q_bytes) + Do_bytes, nk * 8);
    }
    return B2K_OK;
}

int set_all(Multi* m, const char* key, int64_t v) {
    for (int s = 0; s < m->ndev; ++s) {
        int rc = b2k_set_option(m->shard[s], key, v);
        if (rc) return rc;
    }
    return B2K_OK;
}

}  // namespace

int multi_create(int d, const int* devices, int ndev, b2k_index** out) {
    if (d <= 0) { set_error("b2k_create_multi: d must be positive, got %d", d); return B2K_EINVAL; }
    int rc = check_devices(devices, ndev);
    if (rc) return rc;
    Multi* m = new (std::nothrow) Multi();
    if (!m) { set_error("out of host memory"); return B2K_ENOMEM; }
    m->d = d; m->ndev = ndev;
    m->dev.assign(devices, devices + ndev);
    m->shard.assign(ndev, nullptr);
    for (int s = 0; s < ndev && !rc; ++s) rc = b2k_create(d, devices[s], &m->shard[s]);
    if (!rc) rc = setup_multi(m);
    if (!rc) rc = wrap_multi(m, 0, out);
    if (rc) free_multi(m);
    return rc;
}

void multi_destroy(b2k_index* h) {
    free_multi(h->multi);
    delete h;
}

int multi_view(const b2k_index* h, int64_t nrows, b2k_index** out) {
    Multi* src = h->multi;
    std::lock_guard<std::mutex> lk(src->mu);
    const int64_t n = (nrows < 0 || nrows > h->ntotal) ? h->ntotal : nrows;
    Multi* m = new (std::nothrow) Multi();
    if (!m) { set_error("out of host memory"); return B2K_ENOMEM; }
    m->d = src->d; m->ndev = src->ndev; m->dev = src->dev;
    m->shard.assign(m->ndev, nullptr);
    int rc = B2K_OK;
    for (int s = 0; s < m->ndev && !rc; ++s) rc = b2k_view(src->shard[s], rows_of_shard(s, m->ndev, n), &m->shard[s]);
    if (!rc) rc = setup_multi(m);
    if (!rc) rc = wrap_multi(m, n, out);
    if (rc) free_multi(m);
    return rc;
}

int multi_add(b2k_index* h, int64_t n, const float* x, int on_device, void* stream) {
    std::lock_guard<std::mutex> lk(h->multi->mu);
    return add_rows(h, n, x, on_device, (cudaStream_t)stream, 0, 0);
}

int multi_add_synthetic(b2k_index* h, int64_t n, uint64_t seed, int64_t first_row) {
    std::lock_guard<std::mutex> lk(h->multi->mu);
    return add_rows(h, n, nullptr, 0, nullptr, seed, first_row);
}

int multi_reset(b2k_index* h) {
    Multi* m = h->multi;
    std::lock_guard<std::mutex> lk(m->mu);
    for (int s = 0; s < m->ndev; ++s) {
        int rc = b2k_reset(m->shard[s]);
        if (rc) return rc;
    }
    h->ntotal = 0;
    return B2K_OK;
}

int multi_reserve(b2k_index* h, int64_t nrows) {
    Multi* m = h->multi;
    std::lock_guard<std::mutex> lk(m->mu);
    for (int s = 0; s < m->ndev; ++s) {
        int rc = b2k_reserve(m->shard[s], rows_of_shard(s, m->ndev, nrows));
        if (rc) return rc;
    }
    return B2K_OK;
}

int multi_get_rows(const b2k_index* h, int64_t i0, int64_t n, float* out, int on_device) {
    Multi* m = h->multi;
    std::lock_guard<std::mutex> lk(m->mu);
    if (i0 + n > h->ntotal) { set_error("b2k_get_rows: rows [%lld,%lld) out of range (ntotal=%lld)", (long long)i0, (long long)(i0 + n), (long long)h->ntotal); return B2K_EINVAL; }
    return get_rows(m, i0, n, out, on_device);
}

int multi_search(b2k_index* h, int64_t nq, const float* xq, int64_t k, float* D, int64_t* I, int on_device, void* stream) {
    Multi* m = h->multi;
    std::lock_guard<std::mutex> lk(m->mu);
    bool flagged = false;
    int rc = search_once(h, nq, xq, (int)k, D, I, on_device, (cudaStream_t)stream, &flagged);
    // some shard's screening lists could not hold every candidate (masses of near-duplicates at the k-th
    // distance, norms that dwarf the distances): all shards repeat -- first with the 3xTF32 product as the
    // screen, then with the exact scan kernel
    if (!rc && flagged) {
        ++m->redo_count;
        rc = set_all(m, "tensor_mode", 4);
        if (!rc) rc = search_once(h, nq, xq, (int)k, D, I, on_device, (cudaStream_t)stream, &flagged);
        int rc2 = set_all(m, "tensor_mode", m->opt_tensor_mode);
        if (!rc) rc = rc2;
    }
    if (!rc && flagged) {
        ++m->redo_count;
        rc = set_all(m, "path", 1);
        if (!rc) rc = search_once(h, nq, xq, (int)k, D, I, on_device, (cudaStream_t)stream, &flagged);
        int rc2 = set_all(m, "path", m->opt_path);
        if (!rc) rc = rc2;
        if (!rc && flagged) { set_error("b2k_search: the exact scan raised the screening flag (internal error)"); rc = B2K_ECUDA; }
    }
    return rc;
}

int multi_set_option(b2k_index* h, const char* key, int64_t value) {
    Multi* m = h->multi;
    std::lock_guard<std::mutex> lk(m->mu);
    if (!strcmp(key, "screen_flag_ptr")) { set_error("option screen_flag_ptr belongs to the shards of a multi-GPU index"); return B2K_EINVAL; }
    if (!strcmp(key, "multi_peer_writes")) { m->opt_peer_writes = value != 0; return B2K_OK; }
    int rc = set_all(m, key, value);
    if (!rc && !strcmp(key, "tensor_mode")) m->opt_tensor_mode = (int)value;
    if (!rc && !strcmp(key, "path")) m->opt_path = (int)value;
    return rc;
}

int multi_get_stat(b2k_index* h, const char* key, double* out) {
    Multi* m = h->multi;
    std::lock_guard<std::mutex> lk(m->mu);
    if (!strcmp(key, "devices")) { *out = m->ndev; return B2K_OK; }
    if (!strcmp(key, "peer_writes")) { *out = (m->peer_ok && m->opt_peer_writes) ? 1 : 0; return B2K_OK; }
    if (!strcmp(key, "screen_fallbacks")) {
        double sum = (double)m->redo_count;
        for (int s = 0; s < m->ndev; ++s) { double v = 0; if (!b2k_get_stat(m->shard[s], key, &v)) sum += v; }
        *out = sum;
        return B2K_OK;
    }
    if (!strcmp(key, "tensor_kernel_ms") || !strcmp(key, "scan_kernel_ms")) {  // the slowest shard bounds the step
        double best = -1;
        int rc_last = B2K_EINVAL;
        for (int s = 0; s < m->ndev; ++s) { double v = 0; int rc = b2k_get_stat(m->shard[s], key, &v); if (!rc) { rc_last = B2K_OK; if (v > best) best = v; } }
        if (rc_last) return rc_last;
        *out = best;
        return B2K_OK;
    }
    return b2k_get_stat(m->shard[0], key, out);  // shard 0 holds the most rows
}

int multi_write(const b2k_index* hc, const char* path, int64_t first_row, bool append) {
    b2k_index* h = const_cast<b2k_index*>(hc);
    Multi* m = h->multi;
    std::lock_guard<std::mutex> lk(m->mu);
    if (first_row > h->ntotal) { set_error("b2k_write_append: first_row %lld > ntotal %lld", (long long)first_row, (long long)h->ntotal); return B2K_EINVAL; }
    const size_t row_bytes = (size_t)m->d * sizeof(float);
    std::string target = append ? std::string(path) : std::string(path) + ".tmp";
    FILE* f = fopen(target.c_str(), append ? "r+b" : "wb");
    if (!f) { set_error("could not open %s for writing: %s", target.c_str(), strerror(errno)); return B2K_EIO; }
    unsigned char hdr[IXF2_HEADER], want[IXF2_HEADER];
    int rc = B2K_OK;
    if (append) {
        ixf2_put_header(want, m->d, first_row);
        if (fread(hdr, 1, IXF2_HEADER, f) != IXF2_HEADER || memcmp(hdr, want, IXF2_HEADER) != 0) {
            set_error("%s does not hold exactly the first %lld rows of a d=%d IxF2 index", path, (long long)first_row, m->d);
            rc = B2K_EINVAL;
        }
        if (!rc && (fseeko(f, 0, SEEK_END) != 0 || (size_t)ftello(f) != IXF2_HEADER + (size_t)first_row * row_bytes)) {
            set_error("%s: size does not match its header", path);
            rc = B2K_EINVAL;
        }
    } else {
        ixf2_put_header(hdr, m->d, h->ntotal);
        if (fwrite(hdr, 1, IXF2_HEADER, f) != IXF2_HEADER) { set_error("error writing %s: %s", target.c_str(), strerror(errno)); rc = B2K_EIO; }
    }
    // the file is in global row order: gather it from the shards block by block
    int64_t chunk_rows = (int64_t)(MULTI_IO_CHUNK / row_bytes);
    if (chunk_rows < 1) chunk_rows = 1;
    if (!rc && h->ntotal > first_row) rc = ensure_multi_pin(m, (size_t)chunk_rows * row_bytes);
    for (int64_t r0 = first_row; !rc && r0 < h->ntotal; r0 += chunk_rows) {
        const int64_t c = h->ntotal - r0 < chunk_rows ? h->ntotal - r0 : chunk_rows;
        rc = get_rows(m, r0, c, static_cast<float*>(m->pin), 0);
        if (!rc && fwrite(m->pin, 1, (size_t)c * row_bytes, f) != (size_t)c * row_bytes) { set_error("error writing %s: %s", target.c_str(), strerror(errno)); rc = B2K_EIO; }
    }
    if (!rc && append) {
        // rows first, header last (see b2k_write_append)
        ixf2_put_header(want, m->d, h->ntotal);
        if (fflush(f) != 0 || fseeko(f, 0, SEEK_SET) != 0 || fwrite(want, 1, IXF2_HEADER, f) != IXF2_HEADER) {
            set_error("error updating the header of %s: %s", path, strerror(errno));
            rc = B2K_EIO;
        }
    }
    if (fclose(f) != 0 && !rc) { set_error("error closing %s: %s", target.c_str(), strerror(errno)); rc = B2K_EIO; }
    if (!append) {
        if (!rc && rename(target.c_str(), path) != 0) { set_error("could not rename %s to %s: %s", target.c_str(), path, strerror(errno)); rc = B2K_EIO; }
        if (rc) remove(target.c_str());
    }
    return rc;
}

int multi_read(const char* path, const int* devices, int ndev, b2k_index** out) {
    FILE* f = fopen(path, "rb");
    if (!f) { set_error("could not open %s for reading: %s", path, strerror(errno)); return B2K_EIO; }
    unsigned char hdr[IXF2_HEADER];
    if (fread(hdr, 1, IXF2_HEADER, f) != IXF2_HEADER) { fclose(f); set_error("%s: short read of the index header", path); return B2K_EIO; }
    int d = 0; int64_t n = 0;
    int rc = ixf2_parse_header(hdr, path, &d, &n);
    b2k_index* h = nullptr;
    if (!rc) rc = multi_create(d, devices, ndev, &h);
    if (rc) { fclose(f); return rc; }
    Multi* m = h->multi;
    {
        std::lock_guard<std::mutex> lk(m->mu);
        const size_t row_bytes = (size_t)d * sizeof(float);
        int64_t chunk_rows = (int64_t)(MULTI_IO_CHUNK / row_bytes);
        if (chunk_rows < 1) chunk_rows = 1;
        if (n > 0) rc = ensure_multi_pin(m, (size_t)chunk_rows * row_bytes);
        for (int s = 0; s < ndev && !rc; ++s) rc = b2k_reserve(m->shard[s], rows_of_shard(s, ndev, n));
        for (int64_t r0 = 0; !rc && r0 < n; r0 += chunk_rows) {
            const int64_t c = n - r0 < chunk_rows ? n - r0 : chunk_rows;
            if (fread(m->pin, 1, (size_t)c * row_bytes, f) != (size_t)c * row_bytes) { set_error("%s: truncated IxF2 payload", path); rc = B2K_EIO; break; }
            rc = add_rows(h, c, static_cast<const float*>(m->pin), 0, nullptr, 0, 0);  // every device copies its blocks over its own link
        }
    }
    fclose(f);
    if (rc) { multi_destroy(h); return rc; }
    *out = h;
    return B2K_OK;
}

}  // namespace b2k
