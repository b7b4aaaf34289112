/*
 * b200knn.h -- C ABI of libb200knn.so: an exact (brute-force) squared-L2 k-NN index for
 * NVIDIA B200 (sm_100a).  It is the drop-in boundary for the `faiss.IndexFlatL2` object
 * used by the retrieval path of oscka/rag-llm-k8s (`llm/rag.py`).  Each entry point cites
 * the reference call it stands in for (file:line under /root/reference).
 *
 * Conventions
 *   - plain C types only; no exception crosses the boundary;
 *   - every function returns 0 on success or a negative B2K_E* code; the message for the
 *     calling thread's last failure is b2k_last_error();
 *   - vectors are float32, row-major, C-contiguous; ids are int64 (row position, 0-based,
 *     sequential across adds); distances are SQUARED L2 in float32;
 *   - `on_device != 0` means the data pointers of that call are device pointers on the
 *     index's GPU and `stream` is the cudaStream_t to enqueue on (NULL = the legacy
 *     default stream, which is what torch calls its default stream); the call returns
 *     after enqueueing.  With `on_device == 0` the pointers
 *     are host memory and the call returns when the results are in the caller's buffers;
 *   - the handle owns all device memory; the library never keeps a caller pointer past
 *     return;
 *   - different handles are independent across threads; calls on ONE handle are
 *     serialised internally (search included: it uses the handle's workspace).
 *   - there is no CPU fallback: without a usable sm_100 GPU every compute call fails with
 *     B2K_ECUDA.
 */
#ifndef B200KNN_H_
#define B200KNN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2K_OK 0
#define B2K_EINVAL (-1)  /* bad argument (shape, k <= 0, null pointer)              */
#define B2K_ECUDA (-2)   /* CUDA error or no usable device (sticky on the handle)   */
#define B2K_EIO (-3)     /* file missing / short / not an IxF2 file                 */
#define B2K_ENOMEM (-4)  /* device or host allocation failed                        */
#define B2K_ELIMIT (-5)  /* size beyond a documented limit (k > B2K_MAX_K, ...)     */

#define B2K_MAX_K 1024   /* largest k of one search call (north_star: k <= 1024)     */

#if defined(__GNUC__)
#define B2K_API __attribute__((visibility("default")))
#else
#define B2K_API
#endif

typedef struct b2k_index b2k_index;

/* faiss.IndexFlatL2(dimension) -- llm/rag.py:61, :71.  Empty index, d fixed for life. */
B2K_API int b2k_create(int d, int device, b2k_index** out);
B2K_API int b2k_destroy(b2k_index* h);

/* index.add(vectors_np) -- llm/rag.py:80.  Appends n rows; also computes ||x||^2 per row
 * (kernel K1) and keeps the derived copies the search kernels use up to date once they exist
 * (bf16 copies for the screening passes, TF32 residual rows for the 3xTF32 modes). */
B2K_API int b2k_add(b2k_index* h, int64_t n, const float* x, int on_device, void* stream);

/* D, I = index.search(xq, k) -- llm/rag.py:116.  For each of nq queries the k smallest
 * squared distances ascending with their row ids; if ntotal < k the tail is I = -1,
 * D = FLT_MAX (3.4028235e38), which is what rag.py:118-119 then indexes metadata with.
 * Results are exact: whatever reduced-precision screen selects the candidates, the returned
 * distances are fp32 sums of (q - x)^2 over the fp32 rows, ties ordered by id (option
 * "tensor_mode" = 1 is the exception: it returns the 3xTF32 expansion values, faiss's sgemm
 * numerics).  A search served
 * by the tensor-core kernel (batches; see b2k_set_option) synchronises `stream` once even with
 * on_device != 0, unless option "screen_flag_ptr" is set. */
B2K_API int b2k_search(b2k_index* h, int64_t nq, const float* xq, int64_t k, float* D, int64_t* I,
               int on_device, void* stream);

/* index.ntotal, index.d -- llm/rag.py:86, :190-191. */
B2K_API int64_t b2k_ntotal(const b2k_index* h);
B2K_API int b2k_d(const b2k_index* h);

/* index.reset() -- named by the north_star; never called in llm/rag.py.  Drops all rows. */
B2K_API int b2k_reset(b2k_index* h);

/* faiss.write_index(index, path) -- llm/rag.py:62, :82.  Byte-exact "IxF2" file
 * (45-byte header + ntotal*d float32), written to path.tmp then renamed. */
B2K_API int b2k_write(const b2k_index* h, const char* path);

/* Incremental form of write_index for `update_index` (llm/rag.py:74-84: read the file, add
 * rows, rewrite the whole file).  If `path` is an IxF2 file of the same d that holds exactly
 * the first `first_row` rows of this index (header and size are checked; the row CONTENT is
 * the caller's claim -- the Python layer only makes it for an index it loaded from that file),
 * rows [first_row, ntotal) are appended and the 45-byte header is rewritten with one write.
 * The result is byte-identical to b2k_write.  Any mismatch -> B2K_EINVAL, file untouched. */
B2K_API int b2k_write_append(const b2k_index* h, const char* path, int64_t first_row);

/* faiss.read_index(path) -- llm/rag.py:74, :153, :186.  Missing file -> B2K_EIO. */
B2K_API int b2k_read(const char* path, int device, b2k_index** out);

/* The same index spread over several GPUs of the node, driven from this one process (llm/rag.py is
 * one Flask process: rag.py:204; replicas: 1 in llm/ragdeploy.yaml:6).  `devices` lists ndev distinct
 * CUDA devices; global rows go to them in blocks of 4096 rows, round-robin, ids stay global row
 * positions.  The handle is used with every other entry point exactly like a single-GPU one
 * (add / search / ntotal / reset / write / read / view); device pointers (`on_device != 0`) must live
 * on devices[0].  search: every GPU answers for its rows (its kernels write their k best straight into
 * a packed buffer on devices[0] through peer memory), then one merge kernel there.  The IxF2 file is
 * the same single file in global row order, so a file written by one device count loads on any other.
 * ndev == 1 gives the plain single-GPU handle. */
B2K_API int b2k_create_multi(int d, const int* devices, int ndev, b2k_index** out);
B2K_API int b2k_read_multi(const char* path, const int* devices, int ndev, b2k_index** out);

/* A second handle on the first `nrows` rows of `src` (nrows < 0: all of them) that SHARES their
 * storage in HBM -- what makes faiss.read_index affordable on every request (llm/rag.py:153, :186
 * reload the file per request and expect an independent index each time): the Python layer keeps the
 * loaded file resident and hands every caller its own view.  Views are independent indexes: a view
 * that appends while it sits at the end of the shared rows appends in place (other views keep seeing
 * their own prefix); one that appends after the rows have moved on first takes a private copy of its
 * prefix; reset drops only the view's own reference.  Each view has its own stream and workspace, so
 * views may be searched concurrently from different threads. */
B2K_API int b2k_view(const b2k_index* src, int64_t nrows, b2k_index** out);

/* Message of the calling thread's last failed call ("" if none). */
B2K_API const char* b2k_last_error(void);

/* ---- additions with no faiss call site in llm/rag.py (benchmarks, sharding, tests) ---- */

/* Pre-size the row storage so later adds do not reallocate (std::vector::reserve). */
B2K_API int b2k_reserve(b2k_index* h, int64_t nrows);

/* Copy rows [i0, i0+n) back out (faiss `reconstruct_n`; used by write_index and tests). */
B2K_API int b2k_get_rows(const b2k_index* h, int64_t i0, int64_t n, float* out, int on_device);

/* Counter-based synthetic rows generated on the GPU, bit-identical to oracle `orc_synth`:
 * value(seed, row, col) ~ unit-variance sum of four 16-bit uniforms.  b2k_add_synthetic
 * appends rows first_row .. first_row+n-1 of that stream without touching host memory
 * (100M x 768 does not fit in host RAM; SURVEY.md section 7, hard part 5). */
B2K_API int b2k_synth(int device, uint64_t seed, int64_t row0, int64_t nrows, int d, float* out_dev,
              void* stream);
B2K_API int b2k_add_synthetic(b2k_index* h, int64_t n, uint64_t seed, int64_t first_row);

/* Combine step of the row-sharded search: Din/Iin hold nlists
 * result sets [nlists][nq][k] (e.g. the all-gathered per-GPU top-k); writes the k
 * smallest (distance, id) per query.  id_offset (nullable, host array of nlists values)
 * is added to the non-negative ids of each list.  nlists * k <= 8192. */
B2K_API int b2k_merge_topk(int device, int nlists, int64_t nq, int64_t k, const float* Din,
                   const int64_t* Iin, const int64_t* id_offset, float* D, int64_t* I,
                   int on_device, void* stream);

/* The same combine step for the sharded indexes' own layouts, device pointers only: list l has its
 * distances at Din + l*list_stride_d and its ids at Iin + l*list_stride_i (strides in elements, so
 * the lists may sit inside packed all-gather records), and with cyclic_block = B > 0 the ids are
 * shard-local rows of the block-cyclic layout above (B rows per block, nlists shards) and come out as
 * global rows; B = 0 leaves them as they are.  Each list must be sorted ascending by (distance, id) with its
 * padding at the tail -- as b2k_search returns it -- and no id may occur in two lists (shards are disjoint): the
 * kernel merges by rank instead of sorting. */
B2K_API int b2k_merge_shards(int device, int nlists, int64_t nq, int64_t k, const float* Din,
                     int64_t list_stride_d, const int64_t* Iin, int64_t list_stride_i,
                     int64_t cyclic_block, float* D, int64_t* I, void* stream);
#define B2K_CYCLIC_BLOCK 4096  /* rows per block of the multi-GPU layout */

/* The exchange step of the one-process-per-GPU sharded index as a one-shot all-gather over peer memory (no
 * counterpart in the reference, which is one process).  b2k_exchange_push stores `bytes` of `record` (device memory,
 * 16-byte aligned and sized) into slot my_rank (slot_bytes apart) of every rank's receive buffer -- dst_ptrs_dev is a
 * DEVICE array of nranks base addresses, peer-mapped (e.g. torch symmetric memory) -- and then writes seq into word
 * flag_index of every rank's flag row (flag_ptrs_dev: device array of the rows' base addresses).  done_dev: nranks zeroed words of scratch on this device.  b2k_exchange_wait enqueues a
 * bounded wait until flags_dev[0..nranks) have all reached seq; if a rank never arrives it sets *err_dev (nonzero)
 * instead of hanging.  Sequence numbers must grow per flag word; callers alternate two buffer halves so that a rank
 * one call ahead never overwrites a record still being read. */
B2K_API int b2k_exchange_push(int device, const void* record, int64_t bytes, const uint64_t* dst_ptrs_dev,
                      const uint64_t* flag_ptrs_dev, int nranks, int my_rank, int64_t slot_bytes, int flag_index,
                      uint32_t seq, uint32_t* done_dev, void* stream);
B2K_API int b2k_exchange_wait(int device, const uint32_t* flags_dev, int nranks, uint32_t seq, uint32_t* err_dev, void* stream);

/* Search without allocating anything.  No counterpart in the reference (llm/rag.py never
 * manages device memory); it exists for callers that cannot call cudaMalloc -- a process
 * whose inference engine has already reserved HBM, or a latency-critical thread that must
 * avoid the device-wide synchronisation of cudaMalloc/cudaFree.  The caller owns everything:
 * x_dev (n x d float32 rows), q_dev (nq x d), D_dev (nq x k float32), I_dev (nq x k int64),
 * all device pointers on `device`; work is enqueued on `stream`.  Same results as b2k_search
 * on the scan path (direct differences, any nq).  The kernel's scratch (per-CTA partial
 * lists + a ticket) is a fixed B2K_NOALLOC_SCRATCH_BYTES array inside the library, held by
 * every GPU that loads it until the process exits.  Limits: k <= B2K_MAX_K and
 * (3 * SM count) * k keys must fit the array -- true for every k <= 1024 on a B200; anything
 * larger is refused with B2K_ELIMIT.  All calls on a device share that one array, which keeps
 * the previous call's contents: calls ordered on the GPU are independent (the ticket is
 * cleared on `stream` before each launch, the lists are rewritten), calls that overlap on the
 * GPU race.  The first call on each device resolves the array's address
 * (cudaGetSymbolAddress): make it at start-up, before any stream capture. */
#define B2K_NOALLOC_SCRATCH_BYTES (4u << 20)
B2K_API int b2k_search_rows_noalloc(int device, const float* x_dev, int64_t n, int d,
                                    const float* q_dev, int64_t nq, int64_t k, float* D_dev,
                                    int64_t* I_dev, void* stream);

/* Tuning / test hooks.  Keys: "path" (0 auto, 1 force the scan kernel, 2 force the
 * tensor-core kernel), "scan_ctas_per_sm", "timing" (0/1), "tensor_cta_group" (0 auto, 1, 2),
 * "tensor_sep_accum" (0 auto, 1 on, 2 off), "tensor_mode" (0 auto = 3 where it applies, else 2; 1 = 3xTF32 products,
 * 2 = one TF32 product as a screen + exact fp32 re-rank of the rows inside its error slack,
 * 3 = the same screen on bf16 copies of the rows (kept next to the fp32 rows, +50% memory;
 * needs d % 8 == 0, else 2 serves), 4 = the 3xTF32 product used as the screen (what 2 and 3 fall
 * back to when their slack zone overflows; if that overflows too, the exact scan kernel answers)), "tensor_segments" (1 default: the main pass runs in
 * segments with the k-th-distance bound refreshed in between; 0: one launch),
 * "tensor_epilogue_warps" (0 auto, 4, 8: warps filtering the accumulators in the screening modes),
 * "scan_rows" (what the nq < 20 scan reads: 0 auto = the bf16 copies + exact fp32 re-rank once the
 * rows exceed 128 MB, d % 8 == 0 and k <= 496; 1 = always the fp32 rows; 2 = bf16 whenever possible),
 * "screen_flag_ptr" (device address of a caller-owned 32-bit word, 0 = off: batched searches then
 * do not synchronise; they raise the word if the screening lists could not hold every candidate,
 * and the caller must check it after its own synchronisation and repeat that search with
 * "tensor_mode" = 1 -- the row-sharded index ships the word with its all-gather),
 * "tensor_probe" (measurement only: 1 = batched searches run their main pass with the selection filter
 * shut -- the dense rate of the MMA pipeline for the chosen "tensor_mode"; the results are all padding),
 * "tensor_query_stationary" (1: in the one-product screening modes a CTA pair keeps its query block resident in
 * shared memory over a run of row tiles when it fits -- d <= 512 in bf16 --; 0, the default: both operands stream.
 * Same results either way; measured equally fast),
 * "tensor_prefetch_tiles" (n > 0: the batched pass asks L2 for the row tile n tiles ahead of its scheduler; 0, the
 * default: off -- measured no faster at any distance, see profiles/r02_notes.md),
 * "tensor_warm_keys" (2..64, default 12: size of the warm-start sample per query in multiples of k; measured flat
 * between 12 and 64, worse at 6),
 * "prepare_copies" (1: build the bf16 copies of the rows now -- +50% of the row bytes, one conversion pass -- instead
 * of inside the first search that profits from them; a start-up call for servers that reserve HBM afterwards or
 * capture their searches),
 * "multi_peer_writes" (multi-GPU handles: 0 = shards keep their results locally and a peer memcpy moves
 * them to devices[0]; 1 = direct peer stores where the GPUs allow it, the default).
 * Unknown key -> B2K_EINVAL. */
B2K_API int b2k_set_option(b2k_index* h, const char* key, int64_t value);

/* Read-back of measurements / state.  Keys: "last_path" (1 = scan kernel, 2 = tensor-core
 * kernel served the last search), "tensor_kernel_ms" (CUDA-event duration of the last
 * tensor-core main kernel; needs option "timing" = 1; synchronises on that kernel),
 * "tensor_mode_used" (1 / 2 / 3 / 4), "screen_fallbacks" (query chunks repeated with a narrower screen, or by the scan kernel,
 * because the slack zone overflowed), "scan_kernel_ms" / "scan_launches" (the same for the scan kernel launches of the last
 * scan-path search), "scan_rows_used" (1 fp32 rows / 2 bf16 copies + re-rank), "scan_fallback"
 * (1 if the last bf16 scan was redone on the fp32 rows; synchronises the device), "shares_rows" (1 if
 * another handle shares this one's row storage), "devices" (GPUs the handle spans), "peer_writes"
 * (multi-GPU: 1 if the shards store their results directly into devices[0]'s memory). */
B2K_API int b2k_get_stat(b2k_index* h, const char* key, double* out);

/* Number of kernels this library has launched in this process (bench `gpu_launches`). */
B2K_API int64_t b2k_launch_count(void);

/* Library version: major*10000 + minor*100 + patch. */
B2K_API int b2k_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200KNN_H_ */
