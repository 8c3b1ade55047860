/*
 * sgc_b200.h -- C ABI of libsgc_b200.so: the B200 (sm_100a) implementation of SGC's
 * feature-propagation hot path.
 *
 * What this boundary replaces.  The reference (bellaj09/SGC) is pure Python; the arithmetic
 * of its hot path is reached through two third-party interfaces.  Every entry point below
 * names the reference call site whose work it takes over (paths relative to the reference
 * root):
 *
 *   normalization.py:5-12   aug_normalized_adjacency   -> sgc_csr_* / sgc_degree_* / sgc_norm_*
 *   utils.py:23-30          sparse_mx_to_torch_sparse_tensor -> sgc_norm_build (fp32 view), sgc_csr_to_coo
 *   utils.py:92-97          sgc_precompute (K x torch.spmm) -> sgc_spmm_plan_* / sgc_spmm_hop / sgc_precompute
 *   utils.py:118-119        Reddit feature standardisation   -> sgc_col_stats (+ fused into sgc_spmm_hop)
 *   utils.py:63,116-117     symmetrisation, induced sub-graph        -> sgc_sort_* / sgc_coo_reduce_* / sgc_subgraph_*
 *   normalization.py:21-28, utils.py:20,74  row_normalize, sparse first hop  -> sgc_row_normalize, sgc_spmm_hop_sparse_x
 *   downstream/TextSGC/utils.py:59-71,131-152  normalize_adj + precompute -> sgc_norm_values(order 1), sgc_gather_columns_dense, sgc_row_minmax, sgc_transpose_scale
 *   (no counterpart)        row-sharded multi-GPU propagation        -> sgc_symm_* / sgc_ipc_* / sgc_push_*
 *
 * Conventions
 *   - plain C types only; every pointer named *_d / documented "device" is a CUDA device
 *     pointer owned by the caller (PyTorch allocates, the library never does, except the small
 *     opaque host-side plan struct);
 *   - every call is stream-ordered on the cudaStream_t passed as `stream` (void* so that no CUDA
 *     header is needed to bind it) and does not synchronise unless its comment says so;
 *   - return value: 0 on success, a negative SGC_ERR_* code otherwise; a human-readable message
 *     for the calling thread's last failure is available from sgc_last_error();
 *   - index widths: row pointers are int64 (papers100M-shaped graphs exceed 2^31 bytes of
 *     offsets), column indices are int32 (N < 2^31);
 *   - there is no CPU fallback anywhere in this library.
 */
#ifndef SGC_B200_H
#define SGC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SGC_OK                0
#define SGC_ERR_INVALID      -1   /* bad argument (null pointer, negative size, misaligned buffer)  */
#define SGC_ERR_CUDA         -2   /* a CUDA runtime call failed; see sgc_last_error()               */
#define SGC_ERR_ARCH         -3   /* device is not compute capability 10.x                          */
#define SGC_ERR_WORKSPACE    -4   /* workspace smaller than the matching *_workspace_bytes() query  */
#define SGC_ERR_UNSUPPORTED  -5   /* flag combination not built                                      */

/* ---- housekeeping ------------------------------------------------------------------------ */
const char *sgc_version(void);
const char *sgc_last_error(void);
/* Fails with SGC_ERR_ARCH unless `device` is a compute-capability-10.x GPU (B200). */
int32_t sgc_device_check(int32_t device);
/* Number of kernels this library has launched in the calling process (bench.py gpu_launches). */
int64_t sgc_launch_count(void);

/* Sets the current device's L2 set-aside for persisting (evict_last) lines, clamped to the
 * device maximum; the granted size is written to granted_out (may be NULL).  Process-wide. */
int32_t sgc_l2_set_persisting_bytes(int64_t bytes, int64_t *granted_out);

/* Measurement aid (bench.py `roofline.l2`, tools/l2_peak.py): reads the `bytes` (multiple of 512) at buf_d
 * `reps` times from all SMs in 512-byte pieces (32 lanes x 16 bytes, the hop kernel's gather shape) with L1
 * allocation off -- with a buffer that stays resident in the 126 MB L2 the CUDA-event time of this launch is
 * the L2 -> SM read bandwidth of the box.  pattern bit 0: pieces in address order / in a pseudo-random
 * permutation, new per repetition (the gather's access order; every piece is read exactly once per repetition);
 * bit 1: L1::no_allocate loads / the hop kernel's own load flavour (L1-allocating, L2 evict_last hint).
 * piece_stride_bytes: distance between pieces (0 = 512, contiguous; 2432 = one 128-float slab of 608-float rows:
 * the Reddit-shaped hop's footprint -- bytes then spans the strided range, of which 512/stride is read).  sink_d: one
 * float the kernel may write (keeps the loads alive).  No counterpart in the reference. */
int32_t sgc_l2_read_probe(const void *buf_d, int64_t bytes, int64_t piece_stride_bytes, int32_t reps,
                          int32_t pattern, int32_t ctas_per_sm, float *sink_d, void *stream);

/* ---- column-slab transfers (utils.py:82-88,126-130: the reference's .cuda() / .cpu() copies) -------
 * Strided 2-D copy on the DMA engines (cudaMemcpy2DAsync, direction inferred from the pointers):
 * `rows` pieces of `width_bytes` each, `src_pitch_bytes` / `dst_pitch_bytes` apart.  Moves one
 * feature-column slab of a row-major [N, F] matrix between pinned host memory and padded device
 * rows, so that the hops of slab c (columns are independent under S X) run while slab c+1 is
 * still crossing PCIe.  Host memory must be pinned for the copy to be asynchronous. */
int32_t sgc_copy_2d(void *dst, int64_t dst_pitch_bytes, const void *src, int64_t src_pitch_bytes,
                    int64_t width_bytes, int64_t rows, void *stream);
/* Host-side: 1 if every one of the n doubles at v_host is bit-equal to `value`, 0 if not, 2 if the
 * scan gave up after budget_us microseconds (<= 0: no budget), -1 on bad arguments; `threads`
 * workers, memory-bound, early exit.  The stored values of the reference's adjacency matrices are
 * constant (1.0: utils.py:62 networkx adjacency; 2.0: utils.py:116 adj + adj.T of the symmetric 0/1
 * Reddit matrix); the host pipeline then fills the device value array instead of uploading nnz x 8
 * bytes.  No GPU work. */
int32_t sgc_host_all_equal_f64(const double *v_host, int64_t n, double value, int32_t threads,
                               int64_t budget_us);

/* ---- generic device primitives ------------------------------------------------------------ */
size_t  sgc_scan_workspace_bytes(int64_t n);
/* In-place exclusive prefix sum of data_d[0..n) (int64). */
int32_t sgc_scan_exclusive_i64(int64_t *data_d, int64_t n, void *ws_d, size_t ws_bytes, void *stream);

/* ---- K1: A -> canonical CSR of A + I            (normalization.py:6-7) -------------------- */
/* flag_d[0] <- 1 if every row has strictly ascending columns in [0, n_cols), else 0.            */
int32_t sgc_csr_check_canonical(int64_t n_rows, int64_t n_cols, const int64_t *row_ptr_d,
                                const int32_t *col_d, int32_t *flag_d, void *stream);
/* Sort columns inside each row and sum duplicate entries in stored order (scipy
 * sum_duplicates).  Rows are sorted by a stable per-row rank sort.
 * Outputs: row_ptr_out_d[n_rows+1], col_out_d / val_out_d sized nnz (upper bound); the new nnz is
 * row_ptr_out_d[n_rows]. */
size_t  sgc_csr_canonicalize_workspace_bytes(int64_t n_rows, int64_t nnz);
int32_t sgc_csr_canonicalize(int64_t n_rows, int64_t nnz, const int64_t *row_ptr_d,
                             const int32_t *col_d, const double *val_d, int64_t *row_ptr_out_d,
                             int32_t *col_out_d, double *val_out_d, void *ws_d, size_t ws_bytes,
                             void *stream);
/* Pass 1: per-row entry count of A+I (existing diagonal merged, results that are exactly zero
 * dropped, as scipy's csr_plus_csr does), exclusive-scanned into row_ptr_out_d[n+1]. Input must
 * be canonical.  The n rows may be a slice [diag_offset, diag_offset+n) of a larger square
 * matrix with global column ids (row-sharded multi-GPU): row i's diagonal is column
 * diag_offset+i.  diag_offset = 0 for a whole matrix. */
int32_t sgc_csr_selfloop_count(int64_t n, int64_t diag_offset, const int64_t *row_ptr_d,
                               const int32_t *col_d, const double *val_d, int64_t *row_ptr_out_d,
                               void *ws_d, size_t ws_bytes, void *stream);
/* Pass 2: sorted merge-insert of (i,i); writes col_out_d / aval_out_d (fp64 values of A+I). */
int32_t sgc_csr_selfloop_build(int64_t n, int64_t diag_offset, const int64_t *row_ptr_d,
                               const int32_t *col_d, const double *val_d,
                               const int64_t *row_ptr_out_d, int32_t *col_out_d,
                               double *aval_out_d, void *stream);

/* Fast path of K1 for canonical input (what every loader of the reference produces): element-
 * parallel, so hub rows cost no more than short ones.
 * count: one pass over (col, val) that (a) proves the input canonical -- row_ptr monotone from 0 to
 * nnz, strictly ascending columns in [0, n_cols) -- (b) detects entries A + I would drop (exact
 * zeros, a_ii == -1) and (c) records in info_d[n] where each row's self-loop goes; then
 * row_ptr_out_d[n+1] = scan(len_i + (row i lacks its diagonal)).  summary_d[0] (device int64) gets
 * flag bits: 1 = not canonical, 2 = has dropped entries -- if either is set the outputs are
 * unusable and the caller takes the general path above (sgc_csr_canonicalize / _selfloop_*);
 * summary_d[1] = nnz of A + I.  build: element-wise map (col, val) -> (col_out, aval_out). */
int32_t sgc_csr_augment_count(int64_t n, int64_t n_cols, int64_t diag_offset, int64_t nnz,
                              const int64_t *row_ptr_d, const int32_t *col_d, const double *val_d,
                              int32_t *info_d, int64_t *row_ptr_out_d, int64_t *summary_d, void *ws_d,
                              size_t ws_bytes, void *stream);
int32_t sgc_csr_augment_build(int64_t n, int64_t diag_offset, int64_t nnz, const int64_t *row_ptr_d,
                              const int32_t *col_d, const double *val_d, const int32_t *info_d,
                              const int64_t *row_ptr_out_d, int32_t *col_out_d, double *aval_out_d,
                              void *stream);

/* ---- K2: degree reduction and D^-1/2            (normalization.py:8-10) ------------------- */
/* rowsum_d[i] = sum of row i of A+I accumulated in stored order in fp64; dinv_d[i] =
 * rowsum^-1/2 (correctly rounded), +inf -> 0 (normalization.py:10); NaN for negative sums.
 * rowsum_d may be NULL. */
int32_t sgc_degree_inv_sqrt(int64_t n, const int64_t *row_ptr_d, const double *aval_d,
                            double *rowsum_d, double *dinv_d, void *stream);

/* ---- S = D^-1/2 (A+I) D^-1/2 structure + values (normalization.py:11-12, utils.py:25) ------ */
/* scipy's two CSR products drop entries whose product is exactly zero (rows/columns with
 * d == 0).  count: per-row kept entries, scanned into row_ptr_out_d[n+1]; also writes the number
 * of dropped entries to dropped_d[0] (when 0 the input structure can be reused as is).
 * dinv_rows_d[i] is D^-1/2 of row i, dinv_d[c] of column c (the same array for a whole square
 * matrix; dinv_d + row_begin for a row slice). */
int32_t sgc_norm_count(int64_t n, const int64_t *row_ptr_d, const int32_t *col_d,
                       const double *aval_d, const double *dinv_rows_d, const double *dinv_d,
                       int64_t *row_ptr_out_d, int64_t *dropped_d, void *ws_d, size_t ws_bytes,
                       void *stream);
/* build: compacts (col, aval) to the kept entries and optionally materialises the normalised
 * values: s64_out_d = (d_i*a_ij)*d_j in fp64 (what normalization.py:12 returns) and s32_out_d =
 * the same rounded once to fp32 (utils.py:25).  Any of col_out_d/aval_out_d/s64_out_d/s32_out_d
 * may be NULL.  row_ptr_out_d == row_ptr_d means "structure unchanged". */
int32_t sgc_norm_build(int64_t n, const int64_t *row_ptr_d, const int32_t *col_d,
                       const double *aval_d, const double *dinv_rows_d, const double *dinv_d,
                       const int64_t *row_ptr_out_d, int32_t *col_out_d, double *aval_out_d,
                       double *s64_out_d, float *s32_out_d, void *stream);
/* Fast path: values only, element-parallel.  s64_out_d / s32_out_d (either may be NULL) as in
 * sgc_norm_build; zero_flag_d[0] (device int64) <- 1 if any product is exactly zero, i.e. scipy
 * would prune the structure: the caller then runs sgc_norm_count + sgc_norm_build.  row_ptr_d[0]
 * must be 0 and row_ptr_d[n] == nnz.  order 0: (d_i * a_ij) * d_j, the association of
 * normalization.py:12; order 1: (a_ij * d_j) * d_i, the association of TextSGC's normalize_adj
 * (downstream/TextSGC/utils.py:66: adj.dot(D).transpose().dot(D).transpose()). */
int32_t sgc_norm_values(int64_t n, int64_t nnz, const int64_t *row_ptr_d, const int32_t *col_d,
                        const double *aval_d, const double *dinv_rows_d, const double *dinv_d,
                        int32_t order, double *s64_out_d, float *s32_out_d, int64_t *zero_flag_d,
                        void *stream);
/* CSR row pointers -> COO row indices (int64), and int32 -> int64 columns: the [2, nnz] index
 * tensor of utils.py:26-27.  idx_out_d has 2*nnz int64 (rows then cols). */
int32_t sgc_csr_to_coo(int64_t n, const int64_t *row_ptr_d, const int32_t *col_d, int64_t nnz,
                       int64_t *idx_out_d, void *stream);
/* Sorted COO row indices (int64) -> CSR row pointers; col int64 -> int32.  sorted_flag_d[0] is
 * set to 0 if the input is not in row-major strictly ascending order. */
int32_t sgc_coo_to_csr(int64_t n_rows, int64_t nnz, const int64_t *row_idx_d,
                       const int64_t *col_idx_d, int64_t *row_ptr_out_d, int32_t *col_out_d,
                       int32_t *sorted_flag_d, void *stream);

/* ---- K3: column mean / unbiased std             (utils.py:119) ----------------------------- */
size_t  sgc_col_stats_workspace_bytes(int64_t n, int64_t f);
/* x_d: fp32 [n, ld] row-major.  Outputs (device, f entries each; any may be NULL):
 * mean32/std32 (what the fused first hop consumes), mean64/std64. */
int32_t sgc_col_stats(const float *x_d, int64_t n, int64_t f, int64_t ld, float *mean32_d,
                      float *std32_d, double *mean64_d, double *std64_d, void *ws_d,
                      size_t ws_bytes, void *stream);

/* sgc_col_stats without a workspace argument.  (No counterpart in the reference: this exists
 * for hosts of the library, not for a reference call site.)
 *
 * Who needs it: callers that have no way to obtain device scratch -- code running where
 * cudaMalloc is off limits (inside another runtime's callback, or under a policy that gives the
 * whole device to an allocator the caller does not own), and loops that cannot afford the
 * device-wide synchronisation an allocation may imply.  They still pass x and the output
 * vectors; the per-block partial sums go to a static __device__ array inside the library.
 * This call performs no cudaMalloc/cudaFree and calls no allocating library routine.
 *
 * Limit: the partial sums take 16 * f * min(ceil(n/64), 4 * SM count) bytes and must fit in
 * SGC_COL_STATS_STATIC_MAX_BYTES (8 MiB); on a 148-SM B200 that is any n with f <= 885, and
 * wider matrices only below about 37,800 rows.  Beyond it the call returns SGC_ERR_WORKSPACE and touches nothing.
 * The 8 MiB are resident on every GPU that loads the library, until the process exits.
 *
 * First call on each device: looks up the array's device address (cudaGetSymbolAddress, with
 * that device current) and caches it.  Make that first call outside any stream capture.
 *
 * Sharing: all calls on a device use the same array.  Nothing in it needs clearing between
 * calls (each call overwrites every partial sum it later reads), but two calls whose kernels
 * can run at the same time on one GPU -- different streams, or different host threads -- race
 * on it and both results are garbage.  Order them on one stream, or with events.
 *
 * Results are bit-identical to sgc_col_stats on the same input (same blocks, same order). */
#define SGC_COL_STATS_STATIC_MAX_BYTES (8u << 20)
int32_t sgc_col_stats_static(const float *x_d, int64_t n, int64_t f, int64_t ld, float *mean32_d,
                             float *std32_d, double *mean64_d, double *std64_d, void *stream);

/* x <- (x - mean[c]) / std[c] in place, the fp32 expression of utils.py:119 evaluated exactly
 * (one subtraction, one IEEE division per element).  x_d: fp32 [n, ld] row-major. */
int32_t sgc_standardize_inplace(float *x_d, int64_t n, int64_t f, int64_t ld, const float *mean_d,
                                const float *std_d, void *stream);

/* ---- K4: load-balanced CSR SpMM hop             (utils.py:95, torch.spmm) ------------------ */
typedef struct sgc_plan sgc_plan_t;

/* Rows longer than `chunk` stored entries are split into chunks of `chunk` entries whose partial
 * sums are combined in chunk order by a second kernel (deterministic, no float atomics).
 * chunk <= 0: never split -- every output element is accumulated by sequential ascending-column
 * fp32 FMA, bit-identical to torch.spmm on CPU.  `slab_floats`: feature-slab width; all rows are
 * processed for slab 0, then slab 1, ... so that one slab of the gathered operand stays resident
 * in L2 (0 = library default for the shape). */
size_t  sgc_spmm_plan_workspace_bytes(int64_t n_rows, int64_t nnz, int32_t chunk);
/* Builds the split table on `stream`; SYNCHRONISES the stream once to read back the item count.
 * The plan keeps pointers into ws_d and to row_ptr_d/col_d: they must outlive the plan.  The first 256 bytes
 * of ws_d hold 16 pairs of ticket counters for the persistent hop grids: each ticket-scheduled launch takes the
 * next pair in turn and clears it on its own stream before the kernel, so up to 16 hops on one plan may be in
 * flight on different streams, and a launch that was aborted mid-run leaves nothing behind for later ones.
 * (The split-row scratch `partial_d` is the caller's: concurrent hops need one each.) */
int32_t sgc_spmm_plan_create(int64_t n_rows, int64_t n_cols, int64_t nnz,
                             const int64_t *row_ptr_d, const int32_t *col_d, int32_t chunk,
                             void *ws_d, size_t ws_bytes, sgc_plan_t **plan_out, void *stream);
void    sgc_spmm_plan_destroy(sgc_plan_t *plan);
/* Number of long-row chunk items (rows of the partial-sum buffer a hop needs). */
int64_t sgc_spmm_plan_partial_rows(const sgc_plan_t *plan);

/* Hot-column tagging (north_star (2): "L2 persistence hints keep the gathered X resident").  For
 * operands far larger than L2 only the most frequently gathered rows of X are worth keeping:
 * counts how often every column occurs in the plan's col array, chooses the count threshold above
 * which at most `hot_rows` columns lie, and writes col_tagged_d[nnz] = col with bit 31 set on
 * those columns.  Hops on this plan with hints bit 2 set then load tagged rows with an L2
 * evict_last policy and all others with evict_first.  hot_rows <= 0 switches tagging off.
 * SYNCHRONISES the stream (reads back a histogram).  col_tagged_d must outlive the plan.
 * The plan's row_ptr/col may describe a row slice: counts are local to the slice, which is what
 * this GPU's L2 sees. */
size_t  sgc_spmm_hot_workspace_bytes(int64_t n_cols);
int32_t sgc_spmm_plan_tag_hot(sgc_plan_t *plan, int64_t hot_rows, int32_t *col_tagged_d, void *ws_d,
                              size_t ws_bytes, int64_t *n_hot_out, void *stream);

#define SGC_MAX_PEERS        15   /* GPUs of one NVLink domain minus this one                       */

#define SGC_HOP_FUSED_NORM   1u   /* weights = fp32((d_i*a_ij)*d_j) on the fly from aval_d + dinv_d */
#define SGC_HOP_STD_DIV      2u   /* gathered x -> (x - mean[f]) / std[f]   (exact, utils.py:119)   */
#define SGC_HOP_STD_RCP      4u   /* gathered x -> (x - mean[f]) * (1/std[f])  (<= 1 ulp from DIV)   */
#define SGC_HOP_PAD_OWNED    8u   /* columns [f, round_up(f, 4)) of y_d (and of every peer_y_d) are padding
                                     that belongs to the operand (ldy >= round_up(f, 4)): the hop may cover
                                     them with whole-vector stores (they receive zeros when the pad columns
                                     of x are zero).  Without the flag nothing beyond column f-1 is written,
                                     so y_d may be a column view of a wider live matrix.                */
#define SGC_HOP_PITCH_OWNED 32u   /* stronger than PAD_OWNED (and implies it): ALL of columns [f, ldy) of y_d and
                                     of every peer_y_d are padding of the operand, i.e. y_d / peer_y_d are whole
                                     padded rows and not column views.  Only then may the peer variant send a
                                     tile of consecutive rows as ONE bulk store of rows*ldy floats (which
                                     zero-fills [f, ldy) on the peers).  Without it every pushed row is its own
                                     store of round_up(f, 4) floats (f floats when f % 4 == 0) and nothing
                                     else of the peers' rows is touched.                                */
#define SGC_HOP_ACCUMULATE  16u   /* y_d += S_b x instead of y_d = S x: the plan describes ONE BLOCK OF COLUMNS
                                     of S (a CSR holding only the entries whose column lies in the block) and
                                     every row continues its running sum from the value y_d holds -- same
                                     ascending-column FMA chain as the unblocked hop, so chunk = 0 results are
                                     bit-identical; rows without entries in the block are not touched.  Launch
                                     the blocks in ascending column order, the first one without the flag.
                                     Keeps the gathered part of x (one block of its rows) L2-resident when x
                                     is several times larger than L2.  Not with peer pushes / STD_*.       */

typedef struct sgc_hop_args {
    const double *aval_d;      /* FUSED_NORM: fp64 values of A+I (pruned structure)          */
    const double *dinv_d;      /* FUSED_NORM: D^-1/2 by column, n_cols entries                */
    const double *dinv_rows_d; /* FUSED_NORM: D^-1/2 by row, n_rows entries; NULL = dinv_d    */
    const float  *sval_d;      /* !FUSED_NORM: fp32 values of S                               */
    const float  *x_d;         /* gathered operand [n_cols, ldx] fp32                         */
    int64_t       ldx;
    float        *y_d;         /* output [n_rows, ldy] fp32                                   */
    int64_t       ldy;
    float        *partial_d;   /* [sgc_spmm_plan_partial_rows(plan), ldp] scratch (may be NULL if 0 rows) */
    int64_t       ldp;
    const float  *mean_d;      /* STD_*: f entries                                            */
    const float  *std_d;       /* STD_*: f entries                                            */
    int32_t       f;           /* feature columns                                             */
    int32_t       slab_floats; /* 0 = default                                                 */
    uint32_t      flags;
    int32_t       hints;       /* bit0: L2 evict_last policy on gathered x rows; bit1: evict_first on
                                  the streamed col/value reads and the y stores; bit2: use the
                                  plan's hot-column tags (sgc_spmm_plan_tag_hot) if it has them;
                                  measurement aids for the multi-GPU hop: bit3 run the peer variant
                                  but skip the pushes (results on peers are then stale), bit4 push
                                  with per-lane stores instead of cp.async.bulk, bit5 never merge the
                                  rows of a tile into one bulk store, bit6 end the kernel after the
                                  source reads of the bulk stores instead of their completion;
                                  bit7: one warp per task in a one-shot grid instead of the persistent,
                                  ticket-scheduled grid (A/B measurements); bit8: ticket grid even where
                                  the one-shot grid would be chosen by task count (A/B measurements) */
    /* Row-sharded multi-GPU hop (K5): every output row is ALSO stored, from the registers of the
     * SpMM epilogue, to the same (row, column) offset of n_peers other buffers -- the peers'
     * copies of the gathered operand of the next hop, mapped into this process with
     * sgc_ipc_open() and reached over NVLink.  peer_y_d[p] corresponds to y_d (same logical row
     * 0, same leading dimension ldy).  n_peers = 0 for a single-GPU hop. */
    int32_t       n_peers;
    int32_t       push_tile_tasks; /* consecutive tasks per group slot a warp draws per ticket (0 = default:
                                      with peers 4 where the rows of a tile are contiguous in y and travel
                                      as ONE bulk store, else 1; without peers 1-4); rows per tile <= 32 */
    float        *peer_y_d[SGC_MAX_PEERS];
    /* Optional halo mask, one uint32 per row of this hop (n_rows entries): bit p set = peer p's
     * shard gathers this row in the next hop, so it must be pushed to peer_y_d[p]; clear = peer p
     * never reads it and the push is skipped.  NULL = push every row to every peer. */
    const uint32_t *peer_mask_d;
} sgc_hop_args_t;

/* One hop Y = S X on `stream`. */
int32_t sgc_spmm_hop(const sgc_plan_t *plan, const sgc_hop_args_t *args, void *stream);

/* EXPERIMENT, not used by any product path (tools/sweep_tma_gather.py): the plain hop
 * (sval_d weights, 128-float slabs, 16-byte aligned padded operands) with the gathers of X staged through
 * shared memory by the TMA unit -- one cp.async.bulk global -> shared per neighbour row piece, `depth` (1..3)
 * batches of 8 pieces in flight per warp -- instead of per-lane vector loads.  Chunk items of split rows are
 * left as partial sums (no fix-up pass): timing aid; check results on a chunk = 0 plan. */
int32_t sgc_exp_spmm_hop_bulk_gather(const sgc_plan_t *plan, const sgc_hop_args_t *args, int32_t depth,
                                     int32_t ctas_per_sm, void *stream);
/* Same experiment with the tensor form of the TMA gather: cp.async.bulk.tensor.2d ... tile::gather4 over a 2-D
 * tensor map of X ([rows, ldx] floats, box 128 x 1) -- four neighbour rows per instruction. */
int32_t sgc_exp_spmm_hop_gather4(const sgc_plan_t *plan, const sgc_hop_args_t *args, int32_t depth,
                                 int32_t ctas_per_sm, void *stream);

/* K hops (utils.py:94-95).  Hop k reads the output of hop k-1; intermediates ping-pong between
 * buf_a_d and buf_b_d ([n, ld_buf] each; needed only when k_hops > 1 / > 2); the last hop writes
 * args->y_d.  The standardisation flags apply to the first hop only.  Square matrices only. */
int32_t sgc_precompute(const sgc_plan_t *plan, const sgc_hop_args_t *args, int32_t k_hops,
                       float *buf_a_d, float *buf_b_d, int64_t ld_buf, void *stream);

/* ---- graph preparation: the loader steps right before the normalisation ------
 *   adj + adj.T                         utils.py:116      sort the doubled COO stream, sum pairs
 *   max(adj, adj.T)                     utils.py:63       same, op = max
 *   adj[idx, :][:, idx]                 utils.py:117      sgc_subgraph_*
 *   unsorted / duplicated CSR input     (scipy sum_duplicates)   sort + reduce
 * Building block: stable LSD radix sort of (key, int32, fp64) triples by an int32 key in
 * [0, 2^key_bits), 8 bits per pass; payload arrays may be NULL (in/out together).  Sorting a COO
 * stream by column, then by row, leaves it row-major with ascending columns and equal pairs
 * adjacent in their original order. */
size_t  sgc_sort_workspace_bytes(int64_t n);
int32_t sgc_sort_by_key_i32(int64_t n, int32_t key_bits, const int32_t *key_in_d, const int32_t *p32_in_d,
                            const double *p64_in_d, int32_t *key_out_d, int32_t *p32_out_d,
                            double *p64_out_d, void *ws_d, size_t ws_bytes, void *stream);
/* rows_out_d[p] = row_offset + (row that stores entry p): COO row indices of a CSR, int32. */
int32_t sgc_csr_expand_rows(int64_t n, const int64_t *row_ptr_d, int32_t row_offset, int32_t *rows_out_d,
                            void *stream);
/* Row-major sorted COO stream (duplicates adjacent) -> canonical CSR.  op 0: a run of equal pairs
 * is summed in stored order (scipy sum_duplicates / csr_plus_csr); op 1: maximum of the run.
 * drop_zeros: results that compare equal to 0 are dropped (scipy's binary operators do).
 * Outputs sized for n entries; the new nnz is row_ptr_out_d[n_rows]. */
size_t  sgc_coo_reduce_workspace_bytes(int64_t n);
int32_t sgc_coo_reduce_to_csr(int64_t n, int64_t n_rows, const int32_t *row_d, const int32_t *col_d,
                              const double *val_d, int32_t op, int32_t drop_zeros, int64_t *row_ptr_out_d,
                              int32_t *col_out_d, double *val_out_d, void *ws_d, size_t ws_bytes, void *stream);
/* Induced sub-matrix M[idx, :][:, idx] for a list idx of m distinct row/column ids (int64, any
 * order): colmap_d[n_cols] <- position of each id in idx or -1; count -> scanned row_ptr_out_d[m+1];
 * fill -> (col, val) with new column ids, in the stored order of each selected row (ascending when
 * idx is ascending; otherwise sort + reduce afterwards). */
int32_t sgc_subgraph_colmap(int64_t n_cols, int64_t m, const int64_t *idx_d, int32_t *colmap_d, void *stream);
int32_t sgc_subgraph_count(int64_t m, const int64_t *idx_d, const int64_t *row_ptr_d, const int32_t *col_d,
                           const int32_t *colmap_d, int64_t *row_ptr_out_d, void *ws_d, size_t ws_bytes,
                           void *stream);
int32_t sgc_subgraph_fill(int64_t m, const int64_t *idx_d, const int64_t *row_ptr_d, const int32_t *col_d,
                          const double *val_d, const int32_t *colmap_d, const int64_t *row_ptr_out_d,
                          int32_t *col_out_d, double *val_out_d, void *stream);

/* ---- sparse feature matrices (normalization.py:21-28, utils.py:20,74) -----------
 * sgc_row_normalize: val_out[p] = r_inv[row(p)] * val[p] with r = row sum accumulated sequentially in
 *   stored order in the matrix dtype (fp32 when is_f64 == 0, as scipy does for the fp32 Planetoid
 *   features), r_inv = 1/r, inf -> 0 (normalization.py:24-25).  val_out_d may alias val_d.
 *   rinv_out_d (n entries of the same dtype) may be NULL.  zero_flag_d[0] <- 1 if some product is
 *   exactly zero (scipy's CSR product drops those entries: the caller then compacts the structure).
 * sgc_spmm_hop_sparse_x: Y = S X for X stored as canonical CSR (fp32): the first hop of
 *   sgc_precompute on citation-style sparse features without densifying X; bit-identical to the dense
 *   hop in its exact (chunk = 0) accumulation order.  Output row-major [n_rows, ldy], every column
 *   of every row written.  f * 4 bytes must fit the shared-memory row budget (200 KB). */
int32_t sgc_row_normalize(int64_t n, const int64_t *row_ptr_d, const void *val_d, int32_t is_f64,
                          void *val_out_d, void *rinv_out_d, int64_t *zero_flag_d, void *stream);
int32_t sgc_spmm_hop_sparse_x(int64_t n_rows, const int64_t *s_row_ptr_d, const int32_t *s_col_d,
                              const float *s_val_d, const int64_t *x_row_ptr_d, const int32_t *x_col_d,
                              const float *x_val_d, int32_t f, float *y_d, int64_t ldy, void *stream);

/* ---- TextSGC precompute (downstream/TextSGC/utils.py:131-152) -------------------
 *   train_feats = torch.spmm(adj, dense_adj[:, idx]).t(); min/max over documents; drop features
 *   whose range is 0; (feats - train_min) / train_range
 * sgc_gather_columns_dense: out_d[row, colmap_d[c]] = val_d[p] for every stored entry p = (row, c)
 *   with colmap_d[c] >= 0 (colmap from sgc_subgraph_colmap): the dense N x m operand
 *   dense_adj[:, idx] without ever forming the N x N dense matrix.  out_d must be zero-filled.
 * sgc_row_minmax: per row of the hop output y_d [n, m]: min_d[r] = min_t y[r,t], range_d[r] =
 *   max_t y[r,t] - min_d[r]  (fp32).
 * sgc_transpose_scale: out_d[t, j] = (y[keep_d[j], t] - min_d[keep_d[j]]) / range_d[keep_d[j]] for
 *   t < m, j < k; out_d is [m, ld_out] row-major.  One IEEE subtraction + division per element. */
int32_t sgc_gather_columns_dense(int64_t n, int64_t nnz, const int64_t *row_ptr_d, const int32_t *col_d,
                                 const float *val_d, const int32_t *colmap_d, float *out_d, int64_t ld,
                                 void *stream);
int32_t sgc_row_minmax(const float *y_d, int64_t n, int64_t m, int64_t ld, float *min_d, float *range_d,
                       void *stream);
int32_t sgc_transpose_scale(const float *y_d, int64_t m, int64_t ld, const int64_t *keep_d, int64_t k,
                            const float *min_d, const float *range_d, float *out_d, int64_t ld_out,
                            void *stream);

/* ---- K5: row-sharded multi-GPU propagation (north_star (3); no counterpart in the reference) ---
 * One process per GPU.  Every rank owns a contiguous row range of S and keeps two full-height
 * gathered operands (ping-pong).  The hop kernel itself pushes each output row to every peer
 * (sgc_hop_args.peer_y_d) while it computes, so there is no separate all-gather; the ranks only
 * meet at a barrier between hops.  The operands live in "symmetric" buffers allocated here with
 * cudaMalloc (the one place the library allocates device memory: a cudaIpc handle must cover a
 * whole allocation) and exchanged between the processes as 64-byte CUDA IPC handles. */
int32_t sgc_symm_alloc(size_t bytes, void **ptr_out);      /* zero-filled; synchronises          */
int32_t sgc_symm_free(void *ptr_d);
#define SGC_IPC_HANDLE_BYTES 64
int32_t sgc_ipc_export(void *ptr_d, void *handle_out /* 64 bytes, host */);
/* Maps a peer process's allocation into this process (enables peer access lazily). */
int32_t sgc_ipc_open(const void *handle /* 64 bytes, host */, void **ptr_out);
int32_t sgc_ipc_close(void *ptr_d);
/* dst_d[p] <- src_d for p < n_dst (row-major [rows, f] pieces with leading dimensions ld_src /
 * ld_dst): distributes this rank's rows of the initial features to every peer in one kernel. */
int32_t sgc_push_rows(const float *src_d, int64_t ld_src, int64_t rows, int32_t f,
                      float *const *dst_d /* host array of n_dst device pointers */, int32_t n_dst,
                      int64_t ld_dst, void *stream);

/* Contiguous variant on the TMA unit: src_d[0..bytes) -> every dst_d[p][0..bytes) with
 * cp.async.bulk (global -> shared -> global), tiles of tile_bytes, ctas_per_sm CTAs per SM. */
int32_t sgc_push_bytes_bulk(const void *src_d, int64_t bytes, void *const *dst_d, int32_t n_dst,
                            int32_t tile_bytes, int32_t ctas_per_sm, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SGC_B200_H */
