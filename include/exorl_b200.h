/*
 * exorl_b200 — C ABI of the B200-native particle-based kNN reward + n-step replay gather path.
 *
 * The reference (AOS55/exorl) has no FFI layer: the path is three Python call sites
 * (SURVEY.md §8b).  This header is the boundary a maintainer binds instead (ctypes stub in
 * INTEGRATION.md); every entry point cites the reference lines whose work it replaces.
 *
 * Conventions
 *   - plain pointers and sizes only; all pointers are DEVICE pointers unless named host_*;
 *   - row-major contiguous fp32 inputs, int32 neighbour indices;
 *   - every call only enqueues work on `stream` (a cudaStream_t) and never synchronises,
 *     except the explicitly named *_sync helpers;
 *   - return value: 0 on success, one of EXB_E* otherwise; exb_last_error() gives the text
 *     (thread-local, valid until the next failing call on that thread);
 *   - the caller owns every buffer; workspaces are sized by the *_workspace_bytes queries.
 */
#ifndef EXORL_B200_H_
#define EXORL_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EXB_OK 0
#define EXB_EINVAL 1       /* bad pointer / size / alignment                                   */
#define EXB_EKRANGE 2      /* k > number of bank rows: reference raises RuntimeError("selected
                              index k out of range") from topk, utils/utils.py:297            */
#define EXB_EUNSUPPORTED 3 /* shape outside what the kernel family supports (e.g. rep dim)     */
#define EXB_ECUDA 4        /* a CUDA runtime/driver call failed                                */
#define EXB_EFALLBACK 5    /* screening produced more undecided rows than the fallback capacity */

#define EXB_MAX_K 32       /* neighbours per query (reference uses 12 and 3)                   */
#define EXB_SCREEN_CAP 24  /* candidate slots per (query, split) kept by the screening kernel   */
#define EXB_SCREEN_MAX_K 16 /* largest k the screened path accepts (else use the exact path)      */
#define EXB_SCREEN_MAX_DIM 512

typedef void* exb_stream_t; /* cudaStream_t */

int exb_abi_version(void);
const char* exb_last_error(void);
/* Device properties the host side needs for grid sizing (SM count, cc major*10+minor). */
int exb_device_info(int* sm_count, int* cc);

/* --------------------------------------------------------------------------------------------
 * kNN, exact fp32 path (CUDA cores).  Replaces utils/utils.py:293-300 (PBE distance matrix +
 * topk) and agents/unsupervised_learning/proto.py:115-116: for every src row the k smallest
 * direct-difference distances sqrt(sum_c (src[i,c]-bank[j,c])^2), ascending, ties broken by the
 * smaller bank index, with their bank indices (+ idx_base).  No (nq, nb) matrix is materialised.
 * Also the fallback for rows the screening kernel cannot decide.
 * ------------------------------------------------------------------------------------------ */
size_t exb_knn_exact_workspace_bytes(int64_t nq, int64_t nb, int dim, int k);
int exb_knn_exact_f32(const float* src, int64_t nq, const float* bank, int64_t nb, int dim,
                      int k, int32_t idx_base, float* out_dist /* [nq,k] */,
                      int32_t* out_idx /* [nq,k] */, void* workspace, size_t workspace_bytes,
                      exb_stream_t stream);

/* --------------------------------------------------------------------------------------------
 * kNN, screened path (tcgen05 tensor cores + exact recompute) — same results as the exact path.
 *   1. exb_knn_prep_f16: fp32 rows -> fp16 K-major operand padded to a multiple of 64 columns,
 *      plus per row  half_norm = ||fp16(x)||^2 / 2  and  round_err = ||x - fp16(x)||_2 ; sets
 *      a range flag when a value does not fit fp16 (then every row goes to the exact kernel).
 *   2. exb_knn_screen_f16: TMA-fed tcgen05 GEMM  src16 . bank16^T  with the accumulator read
 *      back from TMEM and filtered per row against a running k-th-best-plus-margin threshold;
 *      keeps, per (query, bank split), every bank row whose screened distance is within the
 *      rigorous rounding margin of the k-th best (<= EXB_SCREEN_CAP of them, else the row is
 *      flagged undecided).  Nothing of size nq x nb is written to HBM.
 *   3. exb_knn_rerank_f32: exact fp32 direct-difference distance of the survivors, top-k.
 *   4. undecided rows are recomputed by the exact kernel.
 * exb_knn_screened_f32 runs 1-4.  Replaces the same reference lines as exb_knn_exact_f32.
 * ------------------------------------------------------------------------------------------ */
/* scalars: device struct {float err_max; float hn_max; int range_flag; int pad}, max-accumulated
 * over calls; the caller zeroes it. */
int exb_knn_prep_f16(const float* x, int64_t n, int dim, void* out_f16 /* [n, dim_pad] */,
                     float* half_norm /* [n] */, float* round_err /* [n] */,
                     float* scalars /* 16 bytes */, exb_stream_t stream);

/* Range-proof operands.  The screen may run on s * (x - mu) instead of x for any column shift mu
 * and power-of-two scale s shared by both operands: distances are translation invariant and
 * scale by s, the screening certificate is scale free, and the survivors are recomputed from the
 * caller's fp32 rows — the result is the exact path's, the transform only decides how much of
 * the fp16 format the rows use (reps with a large common offset, |x| > 65504 or |x| << 1e-4
 * would otherwise lose the screen and fall to the CUDA-core kernel).  exb_knn_screened_f32
 * derives (mu, s) itself for dataset-scale calls (one extra pass over the rows).  The pieces, for
 * callers that prepare operands themselves or shard the rows over ranks:
 *   exb_knn_colstats_f32       stats[3][dim] = per-column {sum, min, max} of x (deterministic);
 *                              combine shards with SUM / MIN / MAX
 *   exb_knn_affine_from_stats  affine[EXB_AFFINE_FLOATS] = {mu[0..dim), 0.., s at
 *                              [EXB_SCREEN_MAX_DIM], on-flag, spread}: mu = sum / n_rows,
 *                              s * max|x - mu| in [64, 128); non-finite statistics -> identity
 *   exb_knn_prep_affine_f16    exb_knn_prep_f16 of s * (x - mu); round_err includes the fp32
 *                              rounding of the subtraction.
 * affine records are device memory, 16-byte aligned.  Nothing here synchronises. */
#define EXB_AFFINE_FLOATS (EXB_SCREEN_MAX_DIM + 8)
size_t exb_knn_colstats_workspace_bytes(int64_t n, int dim);
int exb_knn_colstats_f32(const float* x, int64_t n, int dim, float* stats /* [3, dim] */,
                         void* workspace, size_t workspace_bytes, exb_stream_t stream);
int exb_knn_affine_from_stats(const float* stats /* [3, dim] */, double n_rows, int dim,
                              float* affine /* [EXB_AFFINE_FLOATS] */, exb_stream_t stream);
int exb_knn_prep_affine_f16(const float* x, int64_t n, int dim, const float* affine,
                            void* out_f16 /* [n, dim_pad] */, float* half_norm /* [n] */,
                            float* round_err /* [n] */, float* scalars /* 16 bytes */,
                            exb_stream_t stream);

/* Symmetric screen: kNN of a set against ITSELF (PBE / the relabel: utils/utils.py:290,
 * source = target) at dataset scale with HALF the tensor-core work of exb_knn_screened_f32 — the
 * product a^_i . a^_j serves the pairs (i, j) and (j, i), so only tile pairs I <= J are
 * multiplied.  d_sample[i * d_stride] = exact distance of row i to its k-th nearest row inside a
 * sample of the same set (any subset; e.g. exb_knn_screened_f32 of the set against every 32nd
 * row): it fixes every row's keep threshold before the pass.  log_cap / recv_cap: slots of the
 * per-(row, column half) hit logs and of the per-row received-hit list (about 3x the expected
 * k * n / sample_rows / 4 and k * n / sample_rows / 2 entries).  Same results as the other
 * paths; rows whose logs overflow are recomputed by the exact kernel (up to 4096; more -> status
 * flag bit 0 as for exb_knn_screened_f32), a transposed log that overflows sets status flag bit
 * 1 (the call must be repeated with exb_knn_screened_f32).  Status: exb_knn_screened_status_sync
 * on the workspace.  Needs n >= 1024, dim <= EXB_SCREEN_MAX_DIM, k <= 16. */
size_t exb_knn_sym_workspace_bytes(int64_t n, int dim, int k, int log_cap, int recv_cap);
int exb_knn_sym_f32(const float* rep, int64_t n, int dim, int k, int32_t idx_base,
                    const float* d_sample, int64_t d_stride, int log_cap, int recv_cap,
                    float* out_dist /* [n,k] */, int32_t* out_idx /* [n,k] */, void* workspace,
                    size_t workspace_bytes, exb_stream_t stream);

/* The same in three calls, for ranks that each own a block of the tile pairs (the multi-GPU
 * relabel, exorl_b200/relabel.py): every rank holds the whole set and passes the same d_sample;
 * rank r screens the query tile pairs [pair_lo, pair_hi) (tiles of 256 rows), pair u against the
 * bank tiles u, u+1, ... cyclically, (span_end - u) of them plus extra_tiles for u >= mid_pair
 * — so that every unordered tile pair is multiplied on exactly one rank.  The call leaves the
 * direct logs of its rows and the transposed hits (query row, score, bank row) it found for ANY
 * row in `hits` (device arrays inside the workspace; *count entries, read it after a sync).  The
 * ranks exchange the hits by owner of the query row, hand what they received (and their own) to
 * exb_knn_sym_scatter, and finish their rows with exb_knn_sym_rerank_f32 (outputs
 * [row_hi - row_lo, k]).  Status word as for exb_knn_sym_f32; the caller must also treat
 * *count > capacity as "invalid". */
typedef struct {
  int32_t* target;             /* query row the hit belongs to (global row id) */
  float* val;                  /* screened score */
  int32_t* src;                /* bank row (global row id) */
  unsigned long long* count;   /* device: entries appended (may exceed capacity: excess dropped) */
  long long capacity;
} exb_sym_hits;
int exb_knn_sym_screen_f32(const float* rep, int64_t n, int dim, int k, const float* d_sample,
                           int64_t d_stride, int log_cap, int recv_cap, int pair_lo, int pair_hi,
                           int span_end, int mid_pair, int extra_tiles, void* workspace,
                           size_t workspace_bytes, exb_sym_hits* hits, exb_stream_t stream);
int exb_knn_sym_scatter(int64_t n, int dim, int k, int log_cap, int recv_cap,
                        const int32_t* target, const float* val, const int32_t* src,
                        int64_t n_hits, void* workspace, exb_stream_t stream);

/* The exchange step inside the screen kernel (NVLink peer memory).  Every rank owns a
 * peer-visible hit buffer (exb_peer_buffer_alloc below, opened by the other ranks' processes with
 * exb_peer_buffer_open) described by an exb_sym_hits; peer_hits[r] (host array, n_ranks entries,
 * pointers as mapped in THIS process) is rank r's, tile_lo[r] .. tile_lo[r + 1] (host array,
 * n_ranks + 1 entries, tiles of 256 rows) the bank tiles whose rows rank r owns.  The kernel
 * appends each transposed hit straight to the buffer of the owner of its query row (plain stores
 * + one system-scope atomic per ~64 hits), so there is no all-to-all afterwards: every rank
 * zeroes its own *count before the pass (with a barrier in between: nobody may append before
 * the clear), and after a barrier behind the screen it hands its own buffer to
 * exb_knn_sym_scatter_hits (the count is read on the device; count > capacity sets status flag
 * bit 1) and finishes with exb_knn_sym_rerank_f32. */
#define EXB_SYM_MAX_RANKS 8
int exb_knn_sym_screen_peer_f32(const float* rep, int64_t n, int dim, int k,
                                const float* d_sample, int64_t d_stride, int log_cap,
                                int recv_cap, int pair_lo, int pair_hi, int span_end,
                                int mid_pair, int extra_tiles, int n_ranks,
                                const exb_sym_hits* peer_hits, const int32_t* tile_lo,
                                int operand_ready, void* workspace, size_t workspace_bytes,
                                exb_stream_t stream);
/* operand_ready != 0: the caller has filled the screen's operand inside the workspace — the
 * arrays exb_knn_sym_operand_slots returns (f16 [n, dim_pad], half_norm [n], round_err [n],
 * scalars 16 bytes zeroed before the first row is prepared, affine record EXB_AFFINE_FLOATS) —
 * so that G ranks prepare N/G rows each (exb_knn_prep_affine_f16 with one affine record from
 * all-reduced column statistics) and all-gather the shards instead of each preparing all N. */
struct exb_operand;   /* defined with the rep producers below */
int exb_knn_sym_operand_slots(int64_t n, int dim, int k, int log_cap, int recv_cap,
                              void* workspace, size_t workspace_bytes,
                              struct exb_operand* slots, float** affine_slot);
int exb_knn_sym_scatter_hits(int64_t n, int dim, int k, int log_cap, int recv_cap,
                             const exb_sym_hits* hits, void* workspace, exb_stream_t stream);

/* Peer-visible device memory: cudaMalloc + CUDA IPC.  alloc returns the device pointer and a
 * 64-byte handle to send to the other ranks' processes (any transport); open maps a handle of
 * ANOTHER process into this one (peer access over NVLink is enabled on first use; opening a
 * handle in the process that made it is an error — use the pointer); close unmaps, free releases
 * (after every peer closed).  clear = cudaMemsetAsync(ptr, 0, bytes) on `stream`. */
int exb_peer_buffer_alloc(size_t bytes, void** ptr, unsigned char* handle64);
int exb_peer_buffer_open(const unsigned char* handle64, void** ptr);
int exb_peer_buffer_close(void* ptr);
int exb_peer_buffer_free(void* ptr);
int exb_peer_buffer_clear(void* ptr, size_t bytes, exb_stream_t stream);

/* Rows src_host + i * src_pitch_bytes (i < rows, row_bytes each; pinned host memory) into the
 * dense device matrix dst [rows, row_bytes] with one asynchronous 2-D DMA on `stream`.  The
 * end-to-end relabel uploads the symmetric screen's sample (every stride-th row of the host-
 * resident reps, the rows utils/utils.py:290 would score) ahead of the row chunks with it. */
int exb_h2d_rows_2d(void* dst, const void* src_host, int64_t rows, int64_t row_bytes,
                    int64_t src_pitch_bytes, exb_stream_t stream);
int exb_knn_sym_rerank_f32(const float* rep, int64_t n, int dim, int k, int32_t idx_base,
                           int log_cap, int recv_cap, int64_t row_lo, int64_t row_hi,
                           float* out_dist, int32_t* out_idx, void* workspace,
                           exb_stream_t stream);

size_t exb_knn_screened_workspace_bytes(int64_t nq, int64_t nb, int dim, int k);
/* src_is_bank != 0 tells the call that src and bank are the same tensor (PBE: source = target,
 * utils/utils.py:290) so the fp16 operand is prepared once. */
int exb_knn_screened_f32(const float* src, int64_t nq, const float* bank, int64_t nb, int dim,
                         int k, int32_t idx_base, int src_is_bank, float* out_dist,
                         int32_t* out_idx, void* workspace, size_t workspace_bytes,
                         exb_stream_t stream);
/* Device-side status word of the last screened call that used `workspace`:
 *   [0] undecided rows sent to the exact fallback, [1] non-zero if they exceeded the
 *   fallback's capacity (4096 rows per call): exactly those rows are left uncomputed and
 *   marked with out_idx[row * k] = -1, every other row is final — recompute the marked rows
 *   with exb_knn_exact_f32 (exorl_b200/knn.py:_repair_overflow_rows does), [2] candidate
 *   pairs reranked.  Copies 4 ints to host and synchronises `stream`. */
int exb_knn_screened_status_sync(const void* workspace, int32_t host_status[4],
                                 exb_stream_t stream);

/* Debug: with EXB_SCREEN_TRACE=1 in the environment the per-update kernels stamp %globaltimer at
 * their phase boundaries (block 0 / row 0 only); this copies the 16 stamps (ns) to the host.
 *   [0] screen start, [1] first operand chunk landed, [2] MMAs issued, [3] accumulator ready,
 *   [4] k-th best known, [5] candidates collected, [6] lists written,
 *   [8] rerank start, [9] lists gathered, [10] survivors chosen, [11] top-k written.
 * For the streaming (dataset-scale) screen kernel the same 16 words are cycle counters summed
 * over the launch instead: [0] epilogue warps waiting for an accumulator, [1] epilogue warps
 * filtering, [2] slow-path (hit) chunk events, [3] cycles inside the slow path, [4] MMA issuer
 * waiting for a free accumulator (= stalled by the epilogue), [5] MMA issuer waiting for TMA
 * data, [6] MMA issuer total, [7] CTAs that issued MMAs. */
int exb_debug_trace_sync(const void* workspace, int64_t host_ns[16], exb_stream_t stream);

/* Per-update chain (exb_apt_reward_f32 / exb_proto_reward_f32 / exb_knn_screened_f32 on per-update
 * sizes): how rows the latency screen could not certify are recomputed inside the rerank kernel.
 * 0 (default): each row's own warp scans the bank ranges it lost.  1: the rows of a rerank block
 * scan the union of their ranges together through shared memory — 273 instead of 369 us when
 * EVERY row of a 1024-row batch is undecided (a whole trajectory as the batch, reps outside the
 * fp16 range), at 1.3 us more on ordinary batches (C1 32.0 instead of 30.7 us).  Same results.
 * Process-wide; returns the previous setting. */
int exb_knn_set_per_update_scan(int block_cooperative);

/* Measurement scripts only: set an A/B switch of the library (EXB_SCREEN_*, EXB_CHAIN_*, EXB_PDL,
 * ...; csrc/common.cuh) before the first call that consults it.  The product never reads the
 * environment. */
int exb_debug_knob_set(const char* name, int value);

/* Measurement hook (bench.py): when enabled, every screening-kernel launch is bracketed by CUDA
 * events on its own stream (up to 64 launches); collect returns their durations in ms. */
int exb_profile_begin(int enable);
int exb_profile_collect_sync(float* host_ms, int cap, int* host_n);

/* --------------------------------------------------------------------------------------------
 * Rep producers: the step in front of the kNN.  One pass computes the rep
 * the reference hands to PBE / the Proto kNN and writes BOTH the fp32 rep (out_rep, [n, rep_dim])
 * and the prepared fp16 operand (out_op; same bytes exb_knn_prep_f16 would produce from out_rep),
 * so the screen does not re-read the rep from HBM.  Either output may be NULL, not both.
 *   exb_rep_trunk_f32     tanh(LayerNorm(obs @ weight^T + bias)); replaces ICM.get_rep /
 *                         ICM.trunk, agents/unsupervised_learning/icm_apt.py:17-18,50-52.
 *                         weight [rep_dim, in_dim] row-major (nn.Linear layout); W^T is kept in
 *                         shared memory: exb_rep_trunk_smem_bytes(in_dim, rep_dim) <= 227 KB,
 *                         else EXB_EUNSUPPORTED (run the Linear as a GEMM, then the next call).
 *   exb_rep_ln_tanh_f32   tanh(LayerNorm(pre)) on given pre-activations (in place allowed).
 *   exb_rep_normalize_f32 F.normalize(x, dim=1, p=2, eps) = x / max(||x||, eps); replaces
 *                         agents/unsupervised_learning/proto.py:104 (in place allowed).
 * rep_dim % 4 == 0, 4 <= rep_dim <= EXB_SCREEN_MAX_DIM; rows 16-byte aligned.
 * exb_operand.scalars (16 device bytes) is max-accumulated: zero it before the first producer call.
 * exb_knn_screened_pre_f32 = exb_knn_screened_f32 with caller-prepared operands (bank_op == src_op
 * or NULL src_op with src == bank means source = target, utils/utils.py:290); skips step 1.
 * ------------------------------------------------------------------------------------------ */
typedef struct exb_operand {
  void* f16;        /* [n, dim_pad] fp16, dim_pad = dim rounded up to a multiple of 64 */
  float* half_norm; /* [n] */
  float* round_err; /* [n] */
  void* scalars;    /* 16 bytes {float err_max, float hn_max, int range_flag, int 0} */
} exb_operand;
size_t exb_rep_trunk_smem_bytes(int in_dim, int rep_dim);
int exb_rep_trunk_f32(const float* obs, int64_t n, int in_dim, const float* weight,
                      const float* bias, const float* ln_weight, const float* ln_bias,
                      float ln_eps, int rep_dim, float* out_rep, const exb_operand* out_op,
                      exb_stream_t stream);
int exb_rep_ln_tanh_f32(const float* pre, int64_t n, int rep_dim, const float* ln_weight,
                        const float* ln_bias, float ln_eps, float* out_rep,
                        const exb_operand* out_op, exb_stream_t stream);

/* exb_rep_trunk_f32 for input widths whose W^T does not fit shared memory (pixel APT: the
 * 39 200-wide encoder output, ddpg.py:23-24 -> icm_apt.py:17-18): the Linear runs on the tensor
 * cores as a 3xTF32 tcgen05 GEMM (fp32-class accuracy: operands split into tf32 hi + lo parts,
 * accumulation windows of 128 input columns promoted to fp32 registers; split-K partial tiles
 * summed in order), then bias + LayerNorm + Tanh + operand emit in one pass.  in_dim % 4 == 0,
 * rows 16-byte aligned; EXB_EUNSUPPORTED otherwise (run the Linear as a library GEMM and call
 * exb_rep_ln_tanh_f32).  workspace: exb_rep_trunk_wide_workspace_bytes (the weight's lo part and
 * the split-K partials). */
size_t exb_rep_trunk_wide_workspace_bytes(int64_t n, int in_dim, int rep_dim);
int exb_rep_trunk_wide_f32(const float* obs, int64_t n, int in_dim, const float* weight,
                           const float* bias, const float* ln_weight, const float* ln_bias,
                           float ln_eps, int rep_dim, float* out_rep,
                           const exb_operand* out_op, void* workspace, size_t workspace_bytes,
                           exb_stream_t stream);

int exb_rep_normalize_f32(const float* x, int64_t n, int dim, float eps, float* out_rep,
                          const exb_operand* out_op, exb_stream_t stream);
int exb_knn_screened_pre_f32(const float* src, int64_t nq, const exb_operand* src_op,
                             const float* bank, int64_t nb, const exb_operand* bank_op, int dim,
                             int k, int32_t idx_base, float* out_dist, int32_t* out_idx,
                             void* workspace, size_t workspace_bytes, exb_stream_t stream);

/* --------------------------------------------------------------------------------------------
 * Sharded kNN: merge G per-shard (dist, idx) lists (each [nq,k], ascending) into the global
 * k smallest per row; ties -> smaller index.  lists are laid out [G, nq, k] (the result of an
 * all-gather of per-rank results).  New composition; semantics of
 * utils/utils.py:297-300 applied to the union of shards.
 * ------------------------------------------------------------------------------------------ */
int exb_topk_merge(const float* dist_lists, const int32_t* idx_lists, int n_lists, int64_t nq,
                   int k, float* out_dist, int32_t* out_idx, exb_stream_t stream);

/* --------------------------------------------------------------------------------------------
 * Sharded kNN, "exchange the candidate lists" form (SURVEY.md §8e refinement): a rank screens all
 * queries against ITS bank shard and stops there (exb_knn_screen_lists_f32); ranks all-to-all the
 * lists by query block; each rank then reranks the lists of ALL shards for its own query block
 * against the whole fp32 bank (exb_knn_rerank_lists_f32) — the exact recompute and the top-k
 * selection are done once per query instead of once per (query, shard), and no merge is needed.
 *   lists per row and shard: exb_knn_screen_lists_count (0 = inputs too small for this form);
 *   out_val/out_idx [nq, lists, EXB_SCREEN_CAP] (indices local to `bank`), out_cnt [nq, lists]
 *   (-1 = the row is undecided), out_ha/out_ea [nq] query half norms / rounding errors,
 *   out_scalars 16 device bytes {float err_max, float hn_max, int range_flag, int 0}: combine
 *   across ranks with an element-wise MAX over the int32 view before the rerank.
 *   Rerank inputs are laid out [n_shards, nq, lists_per_shard, ...] (the all-to-all receive
 *   buffer), host_shard_lo[s] = first bank row of shard s; rows that stay undecided are
 *   recomputed by the exact kernel against the whole bank.  Same results as exb_knn_exact_f32.
 * ------------------------------------------------------------------------------------------ */
int exb_knn_screen_lists_count(int64_t nq, int64_t nb, int dim, int k);
int exb_knn_screen_lists_f32(const float* src, int64_t nq, const float* bank, int64_t nb, int dim,
                             int k, int src_is_bank, float* out_val, int32_t* out_idx,
                             int32_t* out_cnt, float* out_ha, float* out_ea, void* out_scalars,
                             void* workspace /* exb_knn_screened_workspace_bytes */,
                             size_t workspace_bytes, exb_stream_t stream);
/* exb_knn_screen_lists_f32 on caller-prepared operands (rep producers).  A bank shard's operand
 * is the row slice of the full operand; its scalars may be those of the whole set (maxima over a
 * superset keep the certificate rigorous). */
int exb_knn_screen_lists_pre_f32(const float* src, int64_t nq, const exb_operand* src_op,
                                 const float* bank, int64_t nb, const exb_operand* bank_op,
                                 int dim, int k, float* out_val, int32_t* out_idx,
                                 int32_t* out_cnt, float* out_ha, float* out_ea,
                                 void* out_scalars, void* workspace, size_t workspace_bytes,
                                 exb_stream_t stream);
size_t exb_knn_rerank_lists_workspace_bytes(int64_t nq, int64_t nb, int k);
int exb_knn_rerank_lists_f32(const float* src, int64_t nq, const float* bank, int64_t nb, int dim,
                             int k, int n_shards, int lists_per_shard, const float* cand_val,
                             const int32_t* cand_idx, const int32_t* cand_cnt,
                             const int32_t* host_shard_lo, const float* ha, const float* ea,
                             const void* scalars, float* out_dist, int32_t* out_idx,
                             void* workspace, size_t workspace_bytes, exb_stream_t stream);

/* --------------------------------------------------------------------------------------------
 * RMS state + PBE epilogue.  Replaces utils/utils.py:264-276 (RMS.__call__) and :301-318.
 *   knn_dist [nq,k] ascending.  avg != 0: all nq*k values enter the running moments and the
 *   reward is the mean over k; avg == 0: column k-1 only.  use_rms: update state (M,S,n) then
 *   divide by the UPDATED mean.  clip >= 0: max(r - clip, 0).  reward = log(r + 1), [nq].
 *   RMS state, all on the device and updated IN PLACE: rms_m / rms_s are the fp32 (1,) tensors
 *   RMS.M / RMS.S, rms_n is RMS.n (a python float in the reference) as one fp64 word; a call
 *   leaves M', S' and n + count (utils/utils.py:272-274) there.  No argument changes from one
 *   update to the next, so the per-update chain can be replayed from a CUDA graph.  When
 *   use_rms == 0 the state pointers may be NULL.
 * exb_pbe_moments exposes a shard's plain sums {count, sum, sum of squares} (fp64) so that ranks
 * can all-reduce(SUM) them before exb_pbe_finish; exb_pbe_reward does everything on one device.
 * The workspace carries a ticket word: zero it once with exb_pbe_workspace_init.
 * ------------------------------------------------------------------------------------------ */
size_t exb_pbe_workspace_bytes(int64_t nq, int k);
int exb_pbe_workspace_init(void* workspace, size_t workspace_bytes, exb_stream_t stream);
int exb_pbe_reward(const float* knn_dist, int64_t nq, int k, int avg, int use_rms, float clip,
                   float* rms_m, float* rms_s, double* rms_n, float* out_reward /* [nq] */, void* workspace,
                   size_t workspace_bytes, exb_stream_t stream);
int exb_pbe_moments(const float* knn_dist, int64_t nq, int k, int avg,
                    double* out_sums /* device double[3] = count, sum, sumsq */, void* workspace,
                    size_t workspace_bytes, exb_stream_t stream);
int exb_pbe_finish(const float* knn_dist, int64_t nq, int k, int avg, int use_rms, float clip,
                   const double* sums /* device double[3], already combined */,
                   float* rms_m, float* rms_s, double* rms_n, float* out_reward, exb_stream_t stream);
/* RMS.__call__ on an (n, 1) fp32 column (utils/utils.py:264-276), e.g. for RND (rnd.py:96). */
int exb_rms_update(const float* x, int64_t n, float* rms_m, float* rms_s, double* rms_n, void* workspace,
                   size_t workspace_bytes, exb_stream_t stream);

/* exb_pbe_reward without a workspace argument and without any allocation.  This entry point has
 * no counterpart in the reference: it was added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, results and RMS state convention as exb_pbe_reward;
 * the moments kernel's scratch (ticket word, sums, per-block partials) lives in a 16.5 KB
 * file-scope device array that every GPU loading the library holds until exit.  Limit:
 * nq * k <= EXB_PBE_NOALLOC_MAX_VALUES, beyond it the call fails with EXB_EINVAL and a message
 * in exb_last_error().  The call makes no cudaMalloc / cudaFree / allocating library call; the
 * first call on each device also looks up the array's address (cudaGetSymbolAddress), so make
 * that first call outside stream capture.  All calls on a device share the one array, which
 * keeps the previous call's contents (the ticket word is cleared on `stream` by every call):
 * calls that overlap on the GPU — different streams, or graphs replayed concurrently — race on
 * it; serialise them or use exb_pbe_reward with one workspace per stream. */
#define EXB_PBE_NOALLOC_MAX_VALUES (1LL << 24)
int exb_pbe_reward_noalloc(const float* knn_dist, int64_t nq, int k, int avg, int use_rms,
                           float clip, float* rms_m, float* rms_s, double* rms_n,
                           float* out_reward /* [nq] */, exb_stream_t stream);

/* --------------------------------------------------------------------------------------------
 * PBE.__call__ as ONE call: replaces utils/utils.py:289-319 end to end (distance + topk, RMS
 * update, divide by the updated mean, clip, mean over k, log(1 + r)); compute_apt_reward(source,
 * target, args) of the north star.  mode: 0 = auto, 1 = exact fp32 kernel, 2 = screened.
 * For the per-update sizes (a few thousand rows) the screened path is a 3-launch chain
 * prep -> tcgen05 screen -> rerank, with the RMS/epilogue executed by the last rerank block.
 * out_dist / out_idx [nq,k] are also produced (the neighbours the reference discards, :297).
 * The workspace must be zeroed ONCE after allocation (it carries the epilogue's ticket word);
 * RMS state convention as for exb_pbe_reward (in place, on the device).
 * ------------------------------------------------------------------------------------------ */
size_t exb_apt_reward_workspace_bytes(int64_t nq, int64_t nb, int dim, int k);
int exb_apt_reward_f32(const float* src, int64_t nq, const float* bank, int64_t nb, int dim, int k,
                       int src_is_bank, int mode, int avg, int use_rms, float clip,
                       float* rms_m, float* rms_s, double* rms_n, float* out_reward /* [nq] */,
                       float* out_dist /* [nq,k] */, int32_t* out_idx /* [nq,k] */,
                       void* workspace, size_t workspace_bytes, exb_stream_t stream);

/* exb_apt_reward_f32 on caller-prepared operands (rep producers above); screened mode only. */
int exb_apt_reward_pre_f32(const float* src, int64_t nq, const exb_operand* src_op,
                           const float* bank, int64_t nb, const exb_operand* bank_op, int dim,
                           int k, int avg, int use_rms, float clip, float* rms_m, float* rms_s, double* rms_n, float* out_reward, float* out_dist,
                           int32_t* out_idx, void* workspace, size_t workspace_bytes,
                           exb_stream_t stream);

/* RND intrinsic reward (the other consumer of the running moments):
 *   out_error[i]  = mean_c (target[i,c] - prediction[i,c])^2      (rnd.py:52-54)
 *   (M, S) <- RMS update with x = out_error, bs = n                (rnd.py:96, utils.py:264-276)
 *   out_reward[i] = scale * out_error[i] / (sqrt(S_updated) + 1e-8) (rnd.py:97-98)
 * workspace: exb_pbe_workspace_bytes, zeroed once; RMS state in place as above. */
int exb_rnd_reward_f32(const float* prediction, const float* target, int64_t n, int dim,
                       float scale, float* rms_m, float* rms_s, double* rms_n, float* out_error /* [n] */,
                       float* out_reward /* [n] */, void* workspace, size_t workspace_bytes,
                       exb_stream_t stream);

/* ProtoRL candidate sampling fused with the queue write.  Replaces proto.py:105-112:
 *   scores = protos(z).T; prob = softmax(scores, dim=1); candidates = Categorical(prob).sample();
 *   queue[ptr:ptr+P] = z[candidates]
 * z [batch, dim] (unit rows), protos [n_protos, dim] = the nn.Linear weight (no bias).
 * torch's one-sample multinomial is the exponential race argmax(probs / q), q ~ Exp(1): `noise`
 * [n_protos, batch] supplies q (parity with a recorded reference stream); noise == NULL draws q
 * from Philox4x32-10 keyed by (seed, call).  out_candidates [n_protos] int64 (batch row per
 * prototype); out_prob [n_protos] (optional) = the renormalised probability of the chosen row.
 * queue (optional) [queue_rows, dim]: rows [queue_ptr, queue_ptr + n_protos) are overwritten;
 * the slice must fit (the reference's slice assignment raises otherwise) — the caller advances
 * the pointer, (ptr + P) % queue_rows, like proto.py:112. */
int exb_proto_sample_f32(const float* z, int64_t batch, const float* protos, int n_protos, int dim,
                         const float* noise, uint64_t seed, uint64_t call, float* queue,
                         int64_t queue_rows, int64_t queue_ptr, int64_t* out_candidates,
                         float* out_prob, exb_stream_t stream);

/* ProtoRL reward as ONE call (proto.py:115-118): out_reward[i] = distance of z[i] to its k-th
 * nearest queue row; the neighbours are also returned.  mode and workspace as exb_apt_reward_f32
 * (exb_apt_reward_workspace_bytes). */
int exb_proto_reward_f32(const float* z, int64_t nq, const float* queue, int64_t nb, int dim, int k,
                         int mode, float* out_reward /* [nq] */, float* out_dist /* [nq,k] */,
                         int32_t* out_idx /* [nq,k] */, void* workspace, size_t workspace_bytes,
                         exb_stream_t stream);

/* ProtoRL reward: k-th smallest distance only, no RMS/clip/log (proto.py:115-118). */
int exb_take_kth(const float* knn_dist, int64_t nq, int k, float* out /* [nq] */,
                 exb_stream_t stream);

/* --------------------------------------------------------------------------------------------
 * Replay gather over a device-resident flat buffer.  Replaces the per-sample
 * python of ReplayBuffer._sample, utils/replay_buffer.py:223-235, and
 * OfflineReplayBuffer._sample, :87-96, plus collate/pin/H2D (:246-277).
 *   rows[b] = global row of the 1-based in-episode position idx (= episode start row + idx).
 *   obs      <- observation[rows-1]            (row_bytes_obs bytes per row, any dtype)
 *   action   <- action[rows]
 *   next_obs <- observation[rows+nstep-1]
 *   reward, discount: fp32 (1,) per sample, accumulated exactly in the reference's order:
 *       r = 0; g = 1; for i<nstep: r += g*reward[row+i]; g *= discount[row+i]*gamma
 *     offline (nstep_mode == 0): r = reward[row]; g = discount[row]*gamma   (:94-95)
 *   meta[m]  <- meta_src[m][rows-1]            (n_meta extra columns, :224-225)
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const void* src;      /* device base of the column                       */
  void* dst;            /* device output, batch rows                       */
  int64_t row_bytes;    /* bytes per row                                   */
  int32_t row_offset;   /* added to rows[b]: -1 obs/meta, 0 action, nstep-1 next_obs */
} exb_gather_col;

int exb_replay_gather(const int64_t* rows, int64_t batch, const exb_gather_col* host_cols,
                      int n_cols, const float* reward_src, const float* discount_src, int nstep,
                      int nstep_mode, float gamma, float* out_reward, float* out_discount,
                      exb_stream_t stream);

/* Device index draw (fast mode): Philox4x32-10, counter (sample, call_lo, call_hi, 0), key =
 * seed.  episode = mulhi(r0, n_episodes); idx = mulhi(r1, eps_len[e]-nstep+1)+1 — the range of
 * utils/replay_buffer.py:222 (nstep = 1 gives the offline range of :90) — rows = start+idx. */
int exb_replay_draw(uint64_t seed, uint64_t call, const int64_t* eps_start,
                    const int32_t* eps_len, int64_t n_episodes, int nstep, int64_t batch,
                    int64_t* out_rows, int32_t* out_episode /* may be NULL */,
                    exb_stream_t stream);

/* Draw + gather in ONE launch: every block recomputes the Philox draw of the samples it copies
 * (same arithmetic as exb_replay_draw with nstep_draw = nstep_mode ? nstep : 1), so a batch is a
 * single kernel.  Identical outputs to exb_replay_draw followed by exb_replay_gather; the drawn
 * rows are also written to out_rows [batch] when it is not NULL.  Replaces one whole
 * `next(replay_iter)` of utils/replay_buffer.py:214-244 (:77-99 offline). */
int exb_replay_sample(uint64_t seed, uint64_t call, const int64_t* eps_start,
                      const int32_t* eps_len, int64_t n_episodes, int64_t batch,
                      const exb_gather_col* host_cols, int n_cols, const float* reward_src,
                      const float* discount_src, int nstep, int nstep_mode, float gamma,
                      float* out_reward, float* out_discount, int64_t* out_rows /* may be NULL */,
                      exb_stream_t stream);

/* --------------------------------------------------------------------------------------------
 * Episode ingestion (host side, no GPU involved): reads the reference's episode files — ZIP
 * archives of .npy members written by np.savez_compressed, utils/replay_buffer.py:18-23 — so
 * that a pool of host threads can inflate episodes in parallel into pinned staging buffers.
 * Replaces np.load in load_episode (:26-29) as used by _try_fetch / _store_episode (:172-212)
 * and OfflineReplayBuffer._load (:57-75).  Stored and deflated members, little-endian scalar
 * dtypes, C order.  A handle may be used by one thread at a time; different handles are
 * independent.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  char name[64];       /* member name without ".npy" (the spec name: observation, action, ...) */
  char dtype[8];       /* numpy descr, e.g. "<f4", "|u1", "|b1"                                 */
  int32_t ndim;
  int32_t fortran_order;
  int64_t shape[8];
  int64_t nbytes;      /* payload bytes = prod(shape) * itemsize                                */
} exb_npz_array;

int exb_npz_open(const char* path, void** handle, int* n_arrays);
int exb_npz_info(void* handle, int i, exb_npz_array* out);
/* Inflate array i into host_dst (dst_bytes must equal nbytes). */
int exb_npz_read(void* handle, int i, void* host_dst, int64_t dst_bytes);
void exb_npz_close(void* handle);
/* Decoder used by exb_npz_read for deflated members, process-wide: 1 (default) = the library's
 * own raw-deflate decoder with zlib behind it (a stream it rejects or whose CRC-32 differs is
 * decoded again by zlib), 0 = zlib only, 2 = own decoder only (a stream it cannot decode is an
 * error; test use).  Returns the previous mode; other values leave the mode unchanged. */
int exb_npz_set_inflate_mode(int mode);

#ifdef __cplusplus
}
#endif
#endif /* EXORL_B200_H_ */
