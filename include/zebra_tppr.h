/*
 * zebra_tppr.h — C ABI of libzebra_tppr.so (sm_100a), the B200-native T-PPR hot path of Zebra.
 *
 * Each entry point is what a maintainer of the reference would bind (ctypes)
 * in place of the numba jitclass methods named beside it.  Reference = LuckyLYM/Zebra,
 * paths relative to the reference root.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on error; zebra_last_error() gives the text.
 *     No exceptions cross the ABI.
 *   - plain pointers and sizes only.  "host" pointers are ordinary CPU memory, "dev" pointers are
 *     CUDA device memory owned by the caller (e.g. torch tensors).  The library never frees
 *     caller memory.
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream); all device work of
 *     a call is enqueued on it.  A handle is not re-entrant.
 *   - node ids index rows of the tables: 0 <= id < num_nodes (the reference reserves id 0 as the
 *     padding node, train.py:124-125).
 */
#ifndef ZEBRA_TPPR_H
#define ZEBRA_TPPR_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct zebra_tppr zebra_tppr_t;
typedef struct zebra_plan zebra_plan_t;
typedef struct zebra_epoch zebra_epoch_t;
typedef struct zebra_csr zebra_csr_t;

/* ---- lifetime: replaces tppr_finder.__init__ (utils/util.py:393-400) ------------------------ */

/* k <= 31 (one candidate pair per warp lane); n_tppr <= 8. */
int zebra_tppr_create(zebra_tppr_t **out, int64_t num_nodes, int32_t k, int32_t n_tppr,
                      const double *alpha_list, const double *beta_list, int32_t device);
int zebra_tppr_destroy(zebra_tppr_t *h);

/* Bytes of ONE instance's table (num_nodes fixed-size records) and of one record.  Record
 * layout (little endian, 16-byte chunks each ending in a 32-bit version tag):
 * { f64 norm, int32 count|updates<<6, u32 tag } | { f64 score, int32 eidx, u32 tag }[k] |
 * { f64 ts, int32 node, u32 tag }[k] | pad to a multiple of 128 B.  A zero-filled table is the
 * empty state.  Read it with zebra_tppr_read_state rather than by hand. */
size_t zebra_tppr_table_bytes(const zebra_tppr_t *h);
size_t zebra_tppr_record_bytes(const zebra_tppr_t *h);

/* Point the handle at n_tppr caller-owned device tables (one per (alpha,beta) instance).
 * reset_tppr / backup_tppr / restore_tppr / restore_val_tppr (utils/util.py:402-444) are done by
 * the caller swapping or zero-filling tables and re-binding; the reference's aliasing semantics
 * therefore live in host code, not here. */
int zebra_tppr_bind_tables(zebra_tppr_t *h, void *const *dev_tables);

/* Zero the dependency counters (device + host mirror).  Optional; call only when no work of
 * this handle is in flight. */
int zebra_tppr_reset_tickets(zebra_tppr_t *h, void *stream);

/* ---- streaming top-k: replaces tppr_finder.streaming_topk / streaming_topk_no_fake /
 *      single_streaming_topk / compute_val_tppr (utils/util.py:473-576, 581-679, 682-782,
 *      787-873) -------------------------------------------------------------------------------
 *
 * One chronological span of n edges.  For every instance and every edge i, in order:
 * (1) the CURRENT vectors of src, dst (and neg) are written to output rows i, i+n (, i+2n)
 * exactly as extract_streaming_tppr does (utils/util.py:447-469), then (2) both endpoint vectors
 * are updated from their OLD values (decay, merge, new key, top-k) and committed.
 *
 *   host_nodes     int32 [rows_per_edge*n]  = [src ; dst ; neg]      (rows_per_edge 3)
 *                                           = [src ; dst]            (rows_per_edge 2 or 0)
 *   host_ts        f64   [>= n]             only the first n are read (utils/util.py:499)
 *   host_eidx      int64 [n]                must fit int32 (utils/util.py:450-451 narrows them)
 *   rows_per_edge  3 = with negative rows, 2 = no fake, 0 = update only (no output rows)
 *   inst           -1 = all instances, else only that instance is updated/extracted
 *   dev_out        device buffer of zebra_tppr_out_bytes(h, n, rows_per_edge) bytes, or NULL to
 *                  use an internal one.  Layout: for each instance j:
 *                  int32 nodes[R*k] | int32 eidx[R*k] | f32 dt[R*k] | f32 w[R*k],  R = rows*n.
 *                  Rows of instances excluded by `inst` are zero.
 *   host_out       if non-NULL, the same bytes are copied back to it and the call returns after
 *                  the copy completed (synchronous, like the reference).  If NULL the call only
 *                  enqueues work on `stream`.
 */
size_t zebra_tppr_out_bytes(const zebra_tppr_t *h, int64_t n, int32_t rows_per_edge);
int zebra_tppr_stream(zebra_tppr_t *h, const int32_t *host_nodes, const double *host_ts,
                      const int64_t *host_eidx, int64_t n, int32_t rows_per_edge, int32_t inst,
                      void *dev_out, void *host_out, void *stream);

/* compute_val_tppr's inputs are int64 pandas columns (train.py:243); same semantics as
 * zebra_tppr_stream(rows_per_edge = 0) over all instances, chunked and pipelined internally. */
int zebra_tppr_update_only(zebra_tppr_t *h, const int64_t *host_src, const int64_t *host_dst,
                           const double *host_ts, const int64_t *host_eidx, int64_t n_edges,
                           void *stream);

/* ---- resident spans ("plans") -----------------------------------------------------------------
 * T-PPR state depends only on the edge stream, never on the model (utils/util.py:495-574 reads
 * nothing but nodes, timestamps and edge ids), so a whole chronological span — an epoch's train
 * split, a validation pass with its seeded negatives (evaluation/evaluation.py:9-10) — can be
 * prepared once, kept resident in HBM and replayed batch by batch without any host->device
 * traffic.  Results are identical to calling zebra_tppr_stream once per batch of `batch_size`
 * edges: several batches may share one kernel launch because the dependency tickets, not the
 * batch boundary, order the updates.
 *
 *   host_src/dst/neg  int32 [n] each (neg may be NULL unless rows_per_edge == 3)
 *   plan_run          runs batches [first_batch, first_batch + n_batches) in order; a plan must be
 *                     started at batch 0 (which restarts the dependency counters) and run to its
 *                     end before other spans are streamed on the handle.  Batch b writes its rows,
 *                     in the per-batch layout of zebra_tppr_stream, to slot
 *                     (b - first_batch) % out_slots of dev_out; a slot holds
 *                     zebra_tppr_out_bytes(h, batch_size, rows_per_edge) bytes.  Asynchronous.
 */
int zebra_tppr_plan_create(zebra_tppr_t *h, const int32_t *host_src, const int32_t *host_dst,
                           const int32_t *host_neg, const double *host_ts, const int64_t *host_eidx,
                           int64_t n, int32_t rows_per_edge, int64_t batch_size, zebra_plan_t **out);
int zebra_tppr_plan_destroy(zebra_plan_t *plan);
int64_t zebra_tppr_plan_batches(const zebra_plan_t *plan);
/* Bytes the plan uploaded and keeps in HBM: 52 per staged edge record plus 48 per step of a chain node
 * (on a node-sharded handle only the edges this rank takes part in are staged, plus 4 per entry of its
 * work list). */
size_t zebra_tppr_plan_bytes(const zebra_plan_t *plan);
/* ---- raw messages of a batch in one kernel (model/tgn_model.py:204-226: four index operations, a time
 *      encoding and a torch.cat per batch, each behind its own small host-to-device copy) -----------------
 * For row r: dev_messages[r] = [ memory[u_r] | memory[v_r] | edge_feat[e_r] | cos((t_r - last_update[u_r]) * w + b) ],
 * float32, 2*dim_mem + dim_edge + dim_time columns.  dev_u: the distinct nodes of [src ; dst] whose last
 * interaction of the batch is (partner dev_v, edge dev_e, float32 time dev_t).  The time encoding is the
 * gather's (libdevice cosf, separate multiply and add).  An out-of-range id sets the sticky flag that
 * zebra_gather_check reports.  Asynchronous. */
int zebra_raw_messages(const int32_t *dev_u, const int32_t *dev_v, const int32_t *dev_e, const float *dev_t, int64_t n_rows,
                       const float *dev_memory, int64_t num_nodes, int32_t dim_mem, const float *dev_last_update,
                       const float *dev_edge_feat, int64_t n_edge_rows, int32_t dim_edge, const float *dev_time_w,
                       const float *dev_time_b, int32_t dim_time, float *dev_messages, void *stream);

/* ---- link-prediction metrics of a batch on the device (replaces the per-batch D2H + three scikit-learn
 *      calls of evaluation/evaluation.py:34-45 and train.py:218-227) -------------------------------------
 * dev_pos / dev_neg: float32[n_total] probabilities of the positive / negative edges, cut into batches of
 * batch_size (the last one may be shorter; batch_size <= 4096).  dev_out: float64[ceil(n_total/batch_size)][3]
 * = average precision, ROC AUC, accuracy of every batch, as scikit-learn defines them for binary labels
 * (distinct-threshold step integral / trapezoid; accuracy = mean(pos >= neg)).  Asynchronous. */
int zebra_edge_metrics(const float *dev_pos, const float *dev_neg, int64_t n_total, int64_t batch_size,
                       double *dev_out, void *stream);

/* Test / debug hook: copies the raw record of `node` in instance `inst` (zebra_tppr_record_bytes bytes:
 * header chunk and 2k entry chunks, each ending in its 32-bit tag) to host memory.  Synchronous. */
int zebra_debug_read_record(zebra_tppr_t *h, int32_t inst, int64_t node, void *host_out);
/* Test hook (GPU): the prune's exact-tie ordering in isolation.  host_values holds n_cases rows of
 * `stride` (>= 64) int32, row c uses its first host_n[c] (2..63) entries as stand-ins for scores; on return
 * host_ranks[c*stride + i] is the position of candidate i in numba's np.argsort of that row (reference:
 * utils/util.py:555 through numba.misc.quicksort). */
int zebra_debug_tie_order(const int32_t *host_values, const int32_t *host_n, int32_t n_cases, int32_t stride,
                          int32_t *host_ranks);
int zebra_tppr_plan_run(zebra_tppr_t *h, zebra_plan_t *plan, int64_t first_batch, int64_t n_batches,
                        int64_t batches_per_launch, void *dev_out, int64_t out_slots, void *stream);

/* ---- pipelined epochs: the train / validation / test loops of train.py:196-236 and
 *      evaluation/evaluation.py:17-33 call streaming_topk once per batch of a span that is known
 *      in full before the loop starts (the data split and the seeded negative sampler,
 *      utils/util.py:69-86).  An epoch takes that span from HOST memory, cuts it into windows of
 *      `batches_per_window` whole batches and pipelines, window by window: dependency pass on the
 *      host, host->device copy of the edge records, span kernel, device->host copy of the rows into
 *      pinned memory (second stream).  Two to three windows are in flight at any time.
 *
 *   epoch_begin   the five input arrays must stay valid until epoch_end.  Stages and launches the
 *                 first two windows, returns immediately.
 *   epoch_next    blocks until the next window's rows are in host memory and returns them:
 *                 *host_rows -> *n_batches consecutive slots of zebra_tppr_epoch_slot_bytes() bytes,
 *                 slot q = batch *first_batch + q in the per-batch layout of zebra_tppr_stream
 *                 (identical bytes to calling zebra_tppr_stream once per batch).  The memory belongs
 *                 to the library and is valid until the NEXT epoch_next / epoch_end call on the
 *                 handle.  *n_batches = 0 marks the end.  rows_per_edge = 0: *host_rows is NULL.
 *   epoch_end     waits for everything in flight and releases the epoch (the ring stays cached in
 *                 the handle for the next epoch).
 */
int zebra_tppr_epoch_begin(zebra_tppr_t *h, const int32_t *host_src, const int32_t *host_dst,
                           const int32_t *host_neg, const double *host_ts, const int64_t *host_eidx,
                           int64_t n, int32_t rows_per_edge, int64_t batch_size,
                           int64_t batches_per_window, void *stream, zebra_epoch_t **out);
int zebra_tppr_epoch_next(zebra_epoch_t *epoch, void **host_rows, int64_t *first_batch,
                          int64_t *n_batches);
int zebra_tppr_epoch_end(zebra_epoch_t *epoch);
int64_t zebra_tppr_epoch_slot_bytes(const zebra_epoch_t *epoch);

/* ---- node-sharded mode: one process (or handle) per GPU, T-PPR state partitioned by node id,
 *      cross-shard endpoint vectors exchanged inside the span kernel through peer-mapped mailboxes
 *      over NVLink (BASELINE.json north_star "partitioned by node id") --------
 * Node v lives on rank v & (world-1), local row v >> log2(world); world is 2, 4 or 8.  Every rank
 * creates its handle with the GLOBAL num_nodes, calls shard_init (which allocates the local table
 * slices and the arena = mailboxes + output region, and returns a 64-byte cudaIpcMemHandle_t),
 * exchanges the handles by any means (torch.distributed.all_gather) and calls shard_connect with
 * the world*64 bytes.  Within one process (several GPUs, or tests on one GPU) shard_connect_local
 * takes the handles of all ranks instead.
 *
 * Sharded handles stream through plans only (dependencies must be global to the span): every rank
 * builds the same plan from the same replicated edge arrays and calls zebra_tppr_plan_run with
 * the same arguments; dev_out is then a byte OFFSET into the shared output region (NULL = 0), not
 * a pointer.  A launch may cover at most max_window_edges edges.  Rows
 * are written by the owner of the node they describe into the output region of EVERY rank at the
 * plan-run slot offsets; after all ranks have synchronised their streams and met at a barrier, each
 * rank's zebra_tppr_shard_out() holds the complete result, byte-identical to the single-GPU one.
 * Records are touched by their owner only and the mailboxes carry their own cross-rank hand-shake (a rank
 * pushes into a peer's mailbox half only after that peer has finished the launch that used it before), so no
 * barrier is needed around shard_reset or between launches; the caller places a barrier only before READING
 * rows other ranks store (and before letting them overwrite slots it has not read yet).  shard_reset also
 * clears a recorded wait time-out.  Waits are bounded by wall time (ZEBRA_WAIT_TIMEOUT_MS, default 20 s).
 * backup/restore aliasing is a single-GPU feature; zebra_tppr_read_state reads owned nodes only.
 */
int zebra_tppr_shard_init(zebra_tppr_t *h, int32_t rank, int32_t world, int64_t max_window_edges,
                          size_t out_bytes, void *ipc_handle_out);
int zebra_tppr_shard_connect(zebra_tppr_t *h, const void *ipc_handles);
int zebra_tppr_shard_connect_local(zebra_tppr_t *h, zebra_tppr_t *const *ranks);
/* Unmaps the peers' arenas; every rank calls it (then meets at a barrier) BEFORE any rank destroys
 * its handle, so that no arena is freed while a peer still has it mapped. */
int zebra_tppr_shard_disconnect(zebra_tppr_t *h);
void *zebra_tppr_shard_out(zebra_tppr_t *h);
size_t zebra_tppr_shard_out_bytes(const zebra_tppr_t *h);
int zebra_tppr_shard_reset(zebra_tppr_t *h, void *stream);
/* Row delivery of plans with batch structure: 0 (default) every rank receives every row; 1 the rows
 * of batch b go to rank b & (world-1) only (same slot offset), the other ranks' slots of that batch
 * stay untouched.  Same setting on every rank. */
int zebra_tppr_shard_row_mode(zebra_tppr_t *h, int32_t mode);
/* How a prune orders candidates whose fp64 scores are exactly equal (utils/util.py:555, np.argsort):
 * 0 (default) = the order numba's quicksort gives, replayed exactly, so results are bit-identical to
 * the reference; 1 = insertion order (a stable sort).  Mode 1 differs from the reference only in which
 * of several exactly tied candidates is kept / in their slot order ("ties within tolerance" of the
 * acceptance criteria) and skips the replay, the most expensive part of a tie-heavy update. */
int zebra_tppr_set_tie_mode(zebra_tppr_t *h, int32_t mode);
/* Caps the persistent grid (blocks of 8 warps); lets two ranks share one GPU in tests. */
int zebra_tppr_set_max_blocks(zebra_tppr_t *h, int32_t blocks);

/* Debug / test access: copy the vectors of `n_ids` nodes of instance `inst` to host arrays
 * (counts int32[n_ids], eidx int32[n_ids*k], node int32[n_ids*k], ts f64[n_ids*k],
 * score f64[n_ids*k], norm f64[n_ids]).  Synchronous. */
int zebra_tppr_read_state(zebra_tppr_t *h, int32_t inst, const int64_t *host_ids, int64_t n_ids,
                          int32_t *counts, int32_t *eidx, int32_t *node, double *ts,
                          double *score, double *norm, void *stream);

/* Wait for the work enqueued on `stream` and report a device-side failure (dependency wait
 * time-out) of any span since the previous check.  zebra_tppr_stream with host_out == NULL is
 * asynchronous and does not check by itself. */
int zebra_tppr_sync_check(zebra_tppr_t *h, void *stream);

/* Counters (8 entries) of the last zebra_tppr_stream / update_only / plan_run / epoch: [0] work
 * items, [1] prunes that took the exact tie path, [2] kernel launches so far on this handle,
 * [3] spin time-outs (must be 0), [4] prunes ranked by the sorted-merge fast path, [5] edges with a chain-node endpoint,
 * [6] steps walked by resident chain warps (tppr_device.cuh "hub chains"), [7] zero. */
int zebra_tppr_counters(zebra_tppr_t *h, int64_t *out8);

/* Developer builds (-DZEBRA_PROFILE) accumulate clock64 cycles per kernel phase; this reads and
 * clears the 16 sums (all zero in a normal build).  Synchronises the device. */
int zebra_tppr_profile(zebra_tppr_t *h, uint64_t *out16);

/* Device time of the last span kernel(s) in milliseconds (CUDA events on `stream`); -1 if none. */
double zebra_tppr_last_kernel_ms(zebra_tppr_t *h);

/* ---- pruning strategy: replaces NeighborFinder / get_neighbor_finder / find_before /
 *      get_pruned_topk (utils/util.py:90-107, 144-154, 185-276) and the loop over instances in
 *      GraphDiffusionEmbedding.pruning_topk (modules/embedding_module.py:280-297) ----------------
 *
 * zebra_csr_create uploads the static time-sorted adjacency: node v's entries are
 * [indptr[v], indptr[v+1]) of adj_node / adj_eidx / adj_ts, sorted by timestamp (stable), exactly
 * the three per-node arrays the reference's NeighborFinder holds.  num_nodes = number of rows of
 * indptr minus one.
 *
 * zebra_prune_topk answers nq independent queries (node, time): pruned power iteration of `depth`
 * levels over the `width` most recent strictly-earlier neighbours, dictionary accumulation in the
 * reference's order, top-k by the reference's argsort.  k <= 32.
 *   host_q_nodes int32[nq], host_q_ts f64[nq]
 *   output (zebra_prune_out_bytes(nq,k) bytes): int32 nodes[nq*k] | int32 eidx[nq*k] | f32 dt[nq*k]
 *   | f32 w[nq*k] | int32 filled[nq].  filled[i] = 0 marks a query with no earlier neighbour, whose
 *   row the reference leaves untouched in the caller's arrays (utils/util.py:241-242); its row here
 *   is zero.  dev_out may be NULL (internal buffer); host_out non-NULL makes the call synchronous.
 */
int zebra_csr_create(zebra_csr_t **out, const int64_t *host_indptr, const int32_t *host_adj_node,
                     const int32_t *host_adj_eidx, const double *host_adj_ts, int64_t num_nodes,
                     int64_t nnz, int32_t device);
int zebra_csr_destroy(zebra_csr_t *csr);
size_t zebra_prune_out_bytes(int64_t nq, int32_t k);
int zebra_prune_topk(zebra_csr_t *csr, const int32_t *host_q_nodes, const double *host_q_ts, int64_t nq,
                     int32_t width, int32_t depth, double alpha, double beta, int32_t k,
                     void *dev_out, void *host_out, void *stream);
int64_t zebra_prune_launches(const zebra_csr_t *csr);

/* zebra_prune_topk_static: the pruning query for callers that cannot allocate.  It has NO counterpart in the
 * reference; it was added (code under CUDA graph capture, tight serving
 * loops and harnesses that forbid cudaMalloc).  Same kernel, same output layout and bits as zebra_prune_topk,
 * but every buffer is the caller's: the adjacency (the four arrays of zebra_csr_create), the queries and the
 * output (zebra_prune_out_bytes(nq,k) bytes) are DEVICE pointers on the current device, and the kernel's per-warp
 * work area plus the staged queries live in one file-scope device array of ZEBRA_PRUNE_STATIC_BYTES that every
 * GPU loading the library holds until exit.  The call allocates and frees nothing (no cudaMalloc / cudaFree /
 * cudaMallocAsync / pinned memory), does not synchronise, and enqueues three kernels on `stream`.
 * Limits (a clear error beyond them, nothing is launched): nq <= ZEBRA_PRUNE_STATIC_MAX_QUERIES per call, and
 * width + width^2 + ... + width^depth states per query must fit the work area once (44 bytes per state in
 * ZEBRA_PRUNE_STATIC_BYTES - 1 MiB: about 730,000 states; the usual width 10-20, depth 2 needs 5-19 KB per
 * warp and runs 1,700-6,000 warps at once).
 * Every call on a device SHARES the array and finds the previous call's contents in it; nothing in it has to
 * start at zero, so nothing is cleared.  Calls that overlap on the GPU (two streams, two threads) race: order
 * them on one stream or with events.
 * Query nodes cannot be checked on the host: an id outside [0, num_nodes) yields a zero row with filled = -1
 * (zebra_prune_topk rejects the call instead). */
#define ZEBRA_PRUNE_STATIC_BYTES (32u << 20)
#define ZEBRA_PRUNE_STATIC_MAX_QUERIES 65536
int zebra_prune_topk_static(const int64_t *dev_indptr, const int32_t *dev_adj_node, const int32_t *dev_adj_eidx,
                            const double *dev_adj_ts, int64_t num_nodes, const int32_t *dev_q_nodes,
                            const double *dev_q_ts, int64_t nq, int32_t width, int32_t depth, double alpha,
                            double beta, int32_t k, void *dev_out, void *stream);

/* ---- neighbour gather: replaces the gather / concat / weight-normalisation half of
 *      GraphDiffusionEmbedding.compute_embedding_tppr_ensemble (modules/embedding_module.py:236-271)
 *      and TimeEncode.forward (model/time_encoding.py:23-28) -----------------------------------------
 *
 * One launch for all instances of a batch.  dev_topk is the output buffer of zebra_tppr_stream /
 * zebra_tppr_plan_run for that batch ([n_tppr][4][rows*k] 4-byte planes: node, eidx, dt, w).
 *   dev_x         f32 [n_tppr, rows*k, dim_mem+dim_edge+dim_time] = memory[node] | edge_feat[eidx] |
 *                 cos(dt*time_w + time_b)               (torch.cat of :253,:257,:261 at :264)
 *   dev_wn        f32 [n_tppr, rows, k] = w / sum_k w, zero where the sum is zero        (:268-271)
 *   dev_src_rows  f32 [rows, dim_mem] = memory[dev_src_nodes] (:243-245), or both NULL
 * The two nn.Linear projections and the weighted sum stay in torch (north_star).
 * zebra_gather_bwd scatter-adds the memory-row part of the gradient (the only differentiable input:
 * edge features and time encoder are frozen) into dev_gmemory [num_nodes, dim_mem].
 * zebra_gather_check synchronises and reports an out-of-range id seen by any gather since the last
 * check.  Everything else is asynchronous on `stream`.
 */
int zebra_gather_fwd(const void *dev_topk, int32_t n_tppr, int64_t rows, int32_t k, const float *dev_memory,
                     int64_t num_nodes, int32_t dim_mem, const float *dev_edge_feat, int64_t n_edge_rows,
                     int32_t dim_edge, const float *dev_time_w, const float *dev_time_b, int32_t dim_time,
                     const int32_t *dev_src_nodes, float *dev_x, float *dev_wn, float *dev_src_rows,
                     void *stream);
int zebra_gather_bwd(const void *dev_topk, int32_t n_tppr, int64_t rows, int32_t k, const float *dev_gx,
                     const float *dev_wn, const float *dev_gsrc, const int32_t *dev_src_nodes,
                     float *dev_gmemory, int64_t num_nodes, int32_t dim_mem, int32_t dim_total,
                     void *stream);
/* Reproducible variant of the backward (no atomics): the caller passes the sources sorted by
 * destination node — sources [0, n_tppr*rows*k) are the gather slots in buffer order, the next
 * `rows` are the src rows (only if dev_gsrc != NULL); dev_order int64[total] is a STABLE sort
 * permutation of the source indices by node id, dev_sorted_nodes int32[total] the node ids in that
 * order.  Every node's sources are summed one after the other in that order and stored;
 * dev_gmemory must be zero-filled.  This mirrors the reference's backward, whose index_put with
 * accumulate sorts as well (torch IndexBackward), so a level-1 run is reproducible run to run. */
int zebra_gather_bwd_sorted(int32_t n_tppr, int64_t rows, int32_t k, const float *dev_gx, const float *dev_wn,
                            const float *dev_gsrc, const int32_t *dev_sorted_nodes, const int64_t *dev_order,
                            float *dev_gmemory, int64_t num_nodes, int32_t dim_mem, int32_t dim_total,
                            void *stream);
int zebra_gather_check(void *stream);

/* ---- lazy memory update feeding the gather: replaces, in the train branch of
 *      compute_embedding_tppr_ensemble (modules/embedding_module.py:227-230) and
 *      SequenceMemoryUpdater.get_updated_memory (modules/memory_updater.py:61-90), the D2H of the
 *      neighbour ids, the host-side unique + flag lookup, the clone of the whole [num_nodes, dim_mem]
 *      memory and the [num_nodes, dim_mem] gradient of that clone ------------------------------------
 *
 * zebra_pending_select: the PENDING SET = distinct node ids of dev_topk's id planes whose dev_flags byte
 * (Memory.nodes, modules/memory.py:22,27) is non-zero.  Outputs, all caller-owned device memory:
 *   dev_bitmap      u32 [(num_nodes+31)/32]  bit v set iff v is pending (zeroed by the call)
 *   dev_wordprefix  i32 [(num_nodes+31)/32]  number of pending nodes below word w's first bit
 *   dev_ids         i64 [>= min(num_nodes, n_tppr*rows*k)]  the pending ids in ascending order — exactly
 *                   the `unique_node_ids` of memory_updater.py:68-70, ready to index torch tensors with
 *   dev_count       i32 [1]  how many
 * The caller reads the count (4 bytes), runs the reference's own memory updater on those rows
 * (GRUCell(messages[ids], memory[ids]) -> upd_rows [count, dim_mem]) and hands the result to
 * zebra_gather_fwd_lazy, which is zebra_gather_fwd except that a pending node's row comes from
 * dev_upd_rows (row = its rank in dev_ids) instead of dev_memory.  zebra_gather_bwd_lazy is
 * zebra_gather_bwd_sorted except that it produces dev_gupd [count, dim_mem] (zero-filled by the caller),
 * the gradient of upd_rows: bit-identical to what autograd hands the updater in the reference
 * (grad of the clone, indexed by unique_node_ids), without materialising [num_nodes, dim_mem]. */
int zebra_pending_select(const void *dev_topk, int32_t n_tppr, int64_t rows, int32_t k, const uint8_t *dev_flags,
                         int64_t num_nodes, uint32_t *dev_bitmap, int32_t *dev_wordprefix, int64_t *dev_ids,
                         int32_t *dev_count, void *stream);
int zebra_gather_fwd_lazy(const void *dev_topk, int32_t n_tppr, int64_t rows, int32_t k, const float *dev_memory,
                          int64_t num_nodes, int32_t dim_mem, const float *dev_edge_feat, int64_t n_edge_rows,
                          int32_t dim_edge, const float *dev_time_w, const float *dev_time_b, int32_t dim_time,
                          const int32_t *dev_src_nodes, float *dev_x, float *dev_wn, float *dev_src_rows,
                          const float *dev_upd_rows, const uint32_t *dev_bitmap, const int32_t *dev_wordprefix,
                          void *stream);
int zebra_gather_bwd_lazy(int32_t n_tppr, int64_t rows, int32_t k, const float *dev_gx, const float *dev_wn,
                          const float *dev_gsrc, const int32_t *dev_sorted_nodes, const int64_t *dev_order,
                          const uint32_t *dev_bitmap, const int32_t *dev_wordprefix, float *dev_gupd,
                          int64_t num_nodes, int32_t dim_mem, int32_t dim_total, void *stream);

/* Test hook, host only (no CUDA call, works without a GPU): the staging records zebra_tppr_plan_create
 * would upload for this span on rank `rank` of `world` (world = 1: the single-GPU pass).  out_fields is
 * int32 [10][count]: src, dst, neg, eidx, prev_s, prev_d, prev_n, rthr_s, rthr_d, flags (bit 0: this
 * edge's negative read must signal, bit 1: update dst first); out_ts f64 [count]; out_list int32 [count]
 * span index of every record; out_batch_ofs int64 [n_batches+1]; all sized for count = n by the caller. */
int zebra_debug_host_pass(int64_t num_nodes, int32_t world, int32_t rank, const int32_t *host_src,
                          const int32_t *host_dst, const int32_t *host_neg, const double *host_ts,
                          const int64_t *host_eidx, int64_t n, int64_t batch_size, int32_t *out_fields,
                          double *out_ts, int32_t *out_list, int64_t *out_batch_ofs, int64_t *out_count);

const char *zebra_last_error(void);
const char *zebra_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* ZEBRA_TPPR_H */
