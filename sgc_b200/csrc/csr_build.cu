// K1 / K2: adjacency normalisation on the GPU.
//
// Restates /root/reference/normalization.py:5-12 as CSR kernels.  The reference goes through
// scipy (A + eye -> COO -> row sums -> pow -> two CSR x CSR products); here the same results come
// from one pass that merges the diagonal into every sorted row, one segmented reduction for the
// degrees and (only when some D^-1/2 entry is zero) one compaction pass:
//
//   selfloop_count / selfloop_build : row_ptr / col_idx of A+I, bit-exact against scipy's CSR
//                                     (existing diagonal merged, exact-zero results dropped)
//   (degrees: csr_fast.cu)
//   norm_count / norm_build         : structure after scipy's zero-product pruning + S values
//
// These are the GENERAL kernels: they handle entries that A + I drops and zero-product pruning.  For
// canonical input without such entries (every benchmark graph, every citation graph) the host takes
// the element-parallel fast path of csr_fast.cu and never launches them.
//
// Rows are handled by groups of 8 lanes (4 rows per warp): citation graphs have ~5 entries per
// row, Reddit ~100, hubs ~20k; 8 lanes keep loads coalesced in 32-/64-byte pieces without idling
// most of a warp on short rows.  All kernels are grid-stride with whole-wave grids.
#include "common.cuh"

namespace sgc {

constexpr int GRP = 8;                    // lanes per row
constexpr int ROW_THREADS = 256;          // 32 rows per block
constexpr int ROWS_PER_BLOCK = ROW_THREADS / GRP;

__device__ __forceinline__ unsigned group_mask(int lane) { return 0xffu << (lane & ~(GRP - 1)); }

// keep-rule of csr_plus_csr for A + I: the result value must compare != 0 (NaN is kept)
__device__ __forceinline__ bool keep_aug(int64_t row, int32_t c, double v)
{
    return (c == row) ? ((v + 1.0) != 0.0) : (v != 0.0);
}

__global__ void __launch_bounds__(ROW_THREADS)
check_canonical_kernel(int64_t n_rows, int64_t n_cols, const int64_t *__restrict__ row_ptr,
                       const int32_t *__restrict__ col, int32_t *__restrict__ flag)
{
    bool bad = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_rows;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        if (e < b) bad = true;
        int32_t prev = -1;
        for (int64_t p = b; p < e; ++p) {
            const int32_t c = col[p];
            if (c <= prev || c >= n_cols) bad = true;
            prev = c;
        }
    }
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicAnd(flag, 0);
}

__global__ void __launch_bounds__(ROW_THREADS)
selfloop_count_kernel(int64_t n, int64_t doff, const int64_t *__restrict__ row_ptr,
                      const int32_t *__restrict__ col, const double *__restrict__ val,
                      int64_t *__restrict__ counts)
{
    const int lane = threadIdx.x & 31, gl = lane & (GRP - 1);
    const unsigned gm = group_mask(lane);
    const int64_t groups = (int64_t)gridDim.x * ROWS_PER_BLOCK;
    for (int64_t i = (int64_t)blockIdx.x * ROWS_PER_BLOCK + threadIdx.x / GRP; i < n; i += groups) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        int cnt = 0;
        bool diag = false;
        for (int64_t p = b + gl; p < e; p += GRP) {
            const int32_t c = col[p];
            const double v = val[p];
            diag |= (c == i + doff);
            cnt += keep_aug(i + doff, c, v) ? 1 : 0;
        }
#pragma unroll
        for (int o = GRP / 2; o > 0; o >>= 1) cnt += __shfl_xor_sync(gm, cnt, o);
        const bool any_diag = (__ballot_sync(gm, diag) & gm) != 0u;
        if (gl == 0) counts[i] = cnt + (any_diag ? 0 : 1);   // absent diagonal: 0 + 1 is kept
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) counts[n] = 0;
}

__global__ void __launch_bounds__(ROW_THREADS)
selfloop_build_kernel(int64_t n, int64_t doff, const int64_t *__restrict__ row_ptr,
                      const int32_t *__restrict__ col, const double *__restrict__ val,
                      const int64_t *__restrict__ out_ptr, int32_t *__restrict__ col_out,
                      double *__restrict__ val_out)
{
    const int lane = threadIdx.x & 31, gl = lane & (GRP - 1);
    const unsigned gm = group_mask(lane);
    const int gshift = lane & ~(GRP - 1);
    const int64_t groups = (int64_t)gridDim.x * ROWS_PER_BLOCK;
    for (int64_t i = (int64_t)blockIdx.x * ROWS_PER_BLOCK + threadIdx.x / GRP; i < n; i += groups) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        const int32_t dc = (int32_t)(i + doff);           // column of this row's diagonal
        // does the row already store its diagonal?  (columns are unique, so at most once)
        bool diag = false;
        for (int64_t p = b + gl; p < e; p += GRP) diag |= (col[p] == dc);
        const bool has_diag = (__ballot_sync(gm, diag) & gm) != 0u;
        int64_t w = out_ptr[i];          // next free output slot of the row
        bool diag_written = has_diag;    // nothing to insert when the diagonal is merged in place
        for (int64_t p0 = b; p0 < e; p0 += GRP) {
            const int64_t p = p0 + gl;
            int32_t c = 0x7fffffff;
            double v = 0.0;
            bool keep = false;
            if (p < e) {
                c = col[p];
                v = val[p];
                if (c == dc) v += 1.0;                       // a_ii + 1 (normalization.py:6)
                keep = (v != 0.0);                            // NaN != 0 is true: kept
            }
            // first lane whose column is past the diagonal: the self-loop goes in front of it
            const unsigned past = (__ballot_sync(gm, p < e && c > dc) & gm) >> gshift;
            int ins_at = GRP;                                 // no insertion in this batch
            if (!diag_written && past) { ins_at = __ffs(past) - 1; diag_written = true; }
            const unsigned kept = (__ballot_sync(gm, keep) & gm) >> gshift;
            const int before = __popc(kept & ((1u << gl) - 1u)) + (gl >= ins_at ? 1 : 0);
            if (keep) { col_out[w + before] = c; val_out[w + before] = v; }
            if (ins_at < GRP && gl == ins_at) {
                const int slot = __popc(kept & ((1u << gl) - 1u));
                col_out[w + slot] = dc;
                val_out[w + slot] = 1.0;
            }
            w += __popc(kept) + (ins_at < GRP ? 1 : 0);
        }
        if (!diag_written && gl == 0) {                       // diagonal is the last entry of the row
            col_out[w] = dc;
            val_out[w] = 1.0;
        }
    }
}

// scipy's csr_matmat drops products that are exactly zero: first D*Ahat (d_i * a == 0), then
// (D*Ahat)*D ((d_i*a)*d_j == 0).  NaN products are kept.
__device__ __forceinline__ bool keep_norm(double di, double a, double dj)
{
    const double left = __dmul_rn(di, a);
    return (left != 0.0) && (__dmul_rn(left, dj) != 0.0);
}

__global__ void __launch_bounds__(ROW_THREADS)
norm_count_kernel(int64_t n, const int64_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                  const double *__restrict__ aval, const double *__restrict__ dinv_rows,
                  const double *__restrict__ dinv,
                  int64_t *__restrict__ counts, unsigned long long *__restrict__ dropped)
{
    const int lane = threadIdx.x & 31, gl = lane & (GRP - 1);
    const unsigned gm = group_mask(lane);
    const int64_t groups = (int64_t)gridDim.x * ROWS_PER_BLOCK;
    unsigned long long my_dropped = 0;
    for (int64_t i = (int64_t)blockIdx.x * ROWS_PER_BLOCK + threadIdx.x / GRP; i < n; i += groups) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        const double di = dinv_rows[i];
        int cnt = 0;
        for (int64_t p = b + gl; p < e; p += GRP)
            cnt += keep_norm(di, aval[p], dinv[col[p]]) ? 1 : 0;
#pragma unroll
        for (int o = GRP / 2; o > 0; o >>= 1) cnt += __shfl_xor_sync(gm, cnt, o);
        if (gl == 0) {
            counts[i] = cnt;
            my_dropped += (unsigned long long)((e - b) - cnt);
        }
    }
    if (my_dropped) atomicAdd(dropped, my_dropped);
    if (blockIdx.x == 0 && threadIdx.x == 0) counts[n] = 0;
}

__global__ void __launch_bounds__(ROW_THREADS)
norm_build_kernel(int64_t n, const int64_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                  const double *__restrict__ aval, const double *__restrict__ dinv_rows,
                  const double *__restrict__ dinv,
                  const int64_t *__restrict__ out_ptr, bool compact, int32_t *__restrict__ col_out,
                  double *__restrict__ aval_out, double *__restrict__ s64_out,
                  float *__restrict__ s32_out)
{
    const int lane = threadIdx.x & 31, gl = lane & (GRP - 1);
    const unsigned gm = group_mask(lane);
    const int gshift = lane & ~(GRP - 1);
    const int64_t groups = (int64_t)gridDim.x * ROWS_PER_BLOCK;
    for (int64_t i = (int64_t)blockIdx.x * ROWS_PER_BLOCK + threadIdx.x / GRP; i < n; i += groups) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        const double di = dinv_rows[i];
        int64_t w = compact ? out_ptr[i] : b;
        for (int64_t p0 = b; p0 < e; p0 += GRP) {
            const int64_t p = p0 + gl;
            bool keep = false;
            int32_t c = 0;
            double a = 0.0, s = 0.0;
            if (p < e) {
                c = col[p];
                a = aval[p];
                const double dj = dinv[c];
                s = norm_value64(di, a, dj);
                keep = compact ? keep_norm(di, a, dj) : true;
            }
            const unsigned kept = (__ballot_sync(gm, keep) & gm) >> gshift;
            if (keep) {
                const int64_t q = w + __popc(kept & ((1u << gl) - 1u));
                if (col_out) col_out[q] = c;
                if (aval_out) aval_out[q] = a;
                if (s64_out) s64_out[q] = s;
                if (s32_out) s32_out[q] = __double2float_rn(s);   // utils.py:25 astype(float32)
            }
            w += __popc(kept);
        }
    }
}

__global__ void __launch_bounds__(ROW_THREADS)
csr_to_coo_kernel(int64_t n, const int64_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                  int64_t nnz, int64_t *__restrict__ idx_out)
{
    const int lane = threadIdx.x & 31, gl = lane & (GRP - 1);
    const int64_t groups = (int64_t)gridDim.x * ROWS_PER_BLOCK;
    for (int64_t i = (int64_t)blockIdx.x * ROWS_PER_BLOCK + threadIdx.x / GRP; i < n; i += groups) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        for (int64_t p = b + gl; p < e; p += GRP) {
            idx_out[p] = i;
            idx_out[nnz + p] = (int64_t)col[p];
        }
    }
}

// row_ptr[r] = first entry whose row index is >= r: every entry p writes the pointers of the
// rows that start between its predecessor's row and its own.
__global__ void __launch_bounds__(256)
coo_to_csr_kernel(int64_t n_rows, int64_t nnz, const int64_t *__restrict__ row_idx,
                  const int64_t *__restrict__ col_idx, int64_t *__restrict__ row_ptr,
                  int32_t *__restrict__ col_out, int32_t *__restrict__ sorted_flag)
{
    bool bad = false;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p <= nnz;
         p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r_prev = (p == 0) ? -1 : row_idx[p - 1];
        const int64_t r_cur = (p == nnz) ? n_rows : row_idx[p];
        if (r_prev < -1 || r_prev >= n_rows) bad = true;
        if (p < nnz) {
            const int64_t c = col_idx[p];
            col_out[p] = (int32_t)c;
            if (r_cur < 0 || r_cur >= n_rows || r_cur < r_prev) bad = true;
            if (p > 0 && r_cur == r_prev && col_idx[p - 1] >= c) bad = true;
        }
        if (!bad)
            for (int64_t r = r_prev + 1; r <= r_cur && r <= n_rows; ++r) row_ptr[r] = p;
    }
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicAnd(sorted_flag, 0);
}

// ---- canonicalisation of unsorted / duplicated input (edge-case path) ----------------------------
// Stable rank sort inside each row (O(len^2) per row, one 8-lane group per row): the sorted
// position of entry k is the number of entries with a smaller column, or an equal column and an
// earlier stored position.  Duplicates therefore stay in stored order and are summed in that
// order by the fill pass, which is what scipy's sum_duplicates does.
__global__ void __launch_bounds__(ROW_THREADS)
rank_sort_rows_kernel(int64_t n, const int64_t *__restrict__ row_ptr,
                      const int32_t *__restrict__ col, const double *__restrict__ val,
                      int32_t *__restrict__ col_sorted, double *__restrict__ val_sorted)
{
    const int gl = threadIdx.x & (GRP - 1);
    const int64_t groups = (int64_t)gridDim.x * ROWS_PER_BLOCK;
    for (int64_t i = (int64_t)blockIdx.x * ROWS_PER_BLOCK + threadIdx.x / GRP; i < n; i += groups) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        for (int64_t p = b + gl; p < e; p += GRP) {
            const int32_t c = col[p];
            int64_t rank = 0;
            for (int64_t q = b; q < e; ++q) {
                const int32_t cq = col[q];
                rank += (cq < c || (cq == c && q < p)) ? 1 : 0;
            }
            col_sorted[b + rank] = c;
            val_sorted[b + rank] = val[p];
        }
    }
}

__global__ void __launch_bounds__(ROW_THREADS)
unique_count_kernel(int64_t n, const int64_t *__restrict__ row_ptr,
                    const int32_t *__restrict__ col_sorted, int64_t *__restrict__ counts)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        int64_t cnt = 0;
        for (int64_t p = b; p < e; ++p) cnt += (p == b || col_sorted[p] != col_sorted[p - 1]) ? 1 : 0;
        counts[i] = cnt;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) counts[n] = 0;
}

__global__ void __launch_bounds__(ROW_THREADS)
unique_fill_kernel(int64_t n, const int64_t *__restrict__ row_ptr,
                   const int32_t *__restrict__ col_sorted, const double *__restrict__ val_sorted,
                   const int64_t *__restrict__ out_ptr, int32_t *__restrict__ col_out,
                   double *__restrict__ val_out)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        int64_t w = out_ptr[i] - 1;
        double acc = 0.0;
        for (int64_t p = b; p < e; ++p) {
            const int32_t c = col_sorted[p];
            if (p == b || c != col_sorted[p - 1]) {
                if (p != b) val_out[w] = acc;
                ++w;
                col_out[w] = c;
                acc = val_sorted[p];
            } else {
                acc = __dadd_rn(acc, val_sorted[p]);
            }
        }
        if (e > b) val_out[w] = acc;
    }
}

}  // namespace sgc

using namespace sgc;

extern "C" {

int32_t sgc_csr_check_canonical(int64_t n_rows, int64_t n_cols, const int64_t *row_ptr_d,
                                const int32_t *col_d, int32_t *flag_d, void *stream)
{
    SGC_REQUIRE(n_rows >= 0 && row_ptr_d && flag_d, "csr_check_canonical: bad arguments");
    cudaStream_t s = as_stream(stream);
    const int32_t one = 1;
    SGC_CUDA_TRY(cudaMemcpyAsync(flag_d, &one, sizeof(one), cudaMemcpyHostToDevice, s));
    if (n_rows == 0) return SGC_OK;
    check_canonical_kernel<<<grid_for(ceil_div(n_rows, ROW_THREADS), 8), ROW_THREADS, 0, s>>>(
        n_rows, n_cols, row_ptr_d, col_d, flag_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

size_t sgc_csr_canonicalize_workspace_bytes(int64_t n_rows, int64_t nnz)
{
    const size_t a = ((size_t)nnz * sizeof(int32_t) + 255) / 256 * 256;
    const size_t b = ((size_t)nnz * sizeof(double) + 255) / 256 * 256;
    return a + b + scan_workspace_bytes(n_rows + 1) + 256;
}

int32_t sgc_csr_canonicalize(int64_t n_rows, int64_t nnz, const int64_t *row_ptr_d,
                             const int32_t *col_d, const double *val_d, int64_t *row_ptr_out_d,
                             int32_t *col_out_d, double *val_out_d, void *ws_d, size_t ws_bytes,
                             void *stream)
{
    SGC_REQUIRE(n_rows >= 0 && nnz >= 0 && row_ptr_d && row_ptr_out_d, "csr_canonicalize: bad arguments");
    SGC_REQUIRE(nnz == 0 || (col_d && val_d && col_out_d && val_out_d), "csr_canonicalize: null arrays");
    if (ws_bytes < sgc_csr_canonicalize_workspace_bytes(n_rows, nnz) || !ws_d) {
        set_error("csr_canonicalize: workspace too small");
        return SGC_ERR_WORKSPACE;
    }
    cudaStream_t s = as_stream(stream);
    char *ws = reinterpret_cast<char *>(ws_d);
    const size_t a = ((size_t)nnz * sizeof(int32_t) + 255) / 256 * 256;
    const size_t b = ((size_t)nnz * sizeof(double) + 255) / 256 * 256;
    int32_t *col_sorted = reinterpret_cast<int32_t *>(ws);
    double *val_sorted = reinterpret_cast<double *>(ws + a);
    void *scan_ws = ws + a + b;
    const int grid_g = grid_for(ceil_div(n_rows, ROWS_PER_BLOCK), 8);
    const int grid_t = grid_for(ceil_div(n_rows, ROW_THREADS), 8);
    if (n_rows == 0) {
        SGC_CUDA_TRY(cudaMemsetAsync(row_ptr_out_d, 0, sizeof(int64_t), s));
        return SGC_OK;
    }
    rank_sort_rows_kernel<<<grid_g, ROW_THREADS, 0, s>>>(n_rows, row_ptr_d, col_d, val_d, col_sorted, val_sorted);
    SGC_LAUNCH_CHECK();
    unique_count_kernel<<<grid_t, ROW_THREADS, 0, s>>>(n_rows, row_ptr_d, col_sorted, row_ptr_out_d);
    SGC_LAUNCH_CHECK();
    int32_t rc = scan_exclusive_i64(row_ptr_out_d, n_rows + 1, scan_ws, ws_bytes - a - b, s);
    if (rc != SGC_OK) return rc;
    unique_fill_kernel<<<grid_t, ROW_THREADS, 0, s>>>(n_rows, row_ptr_d, col_sorted, val_sorted,
                                                     row_ptr_out_d, col_out_d, val_out_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_csr_selfloop_count(int64_t n, int64_t diag_offset, const int64_t *row_ptr_d,
                               const int32_t *col_d, const double *val_d, int64_t *row_ptr_out_d,
                               void *ws_d, size_t ws_bytes, void *stream)
{
    SGC_REQUIRE(n >= 0 && row_ptr_d && row_ptr_out_d, "csr_selfloop_count: bad arguments");
    cudaStream_t s = as_stream(stream);
    selfloop_count_kernel<<<grid_for(ceil_div(n, ROWS_PER_BLOCK), 8), ROW_THREADS, 0, s>>>(
        n, diag_offset, row_ptr_d, col_d, val_d, row_ptr_out_d);
    SGC_LAUNCH_CHECK();
    return scan_exclusive_i64(row_ptr_out_d, n + 1, ws_d, ws_bytes, s);
}

int32_t sgc_csr_selfloop_build(int64_t n, int64_t diag_offset, const int64_t *row_ptr_d,
                               const int32_t *col_d, const double *val_d,
                               const int64_t *row_ptr_out_d, int32_t *col_out_d,
                               double *aval_out_d, void *stream)
{
    SGC_REQUIRE(n >= 0 && row_ptr_d && row_ptr_out_d, "csr_selfloop_build: bad arguments");
    if (n == 0) return SGC_OK;
    SGC_REQUIRE(col_out_d && aval_out_d, "csr_selfloop_build: null outputs");
    selfloop_build_kernel<<<grid_for(ceil_div(n, ROWS_PER_BLOCK), 8), ROW_THREADS, 0, as_stream(stream)>>>(
        n, diag_offset, row_ptr_d, col_d, val_d, row_ptr_out_d, col_out_d, aval_out_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_norm_count(int64_t n, const int64_t *row_ptr_d, const int32_t *col_d,
                       const double *aval_d, const double *dinv_rows_d, const double *dinv_d,
                       int64_t *row_ptr_out_d, int64_t *dropped_d, void *ws_d, size_t ws_bytes,
                       void *stream)
{
    SGC_REQUIRE(n >= 0 && row_ptr_d && dinv_d && dinv_rows_d && row_ptr_out_d && dropped_d, "norm_count: bad arguments");
    cudaStream_t s = as_stream(stream);
    SGC_CUDA_TRY(cudaMemsetAsync(dropped_d, 0, sizeof(int64_t), s));
    norm_count_kernel<<<grid_for(ceil_div(n, ROWS_PER_BLOCK), 8), ROW_THREADS, 0, s>>>(
        n, row_ptr_d, col_d, aval_d, dinv_rows_d, dinv_d, row_ptr_out_d,
        reinterpret_cast<unsigned long long *>(dropped_d));
    SGC_LAUNCH_CHECK();
    return scan_exclusive_i64(row_ptr_out_d, n + 1, ws_d, ws_bytes, s);
}

int32_t sgc_norm_build(int64_t n, const int64_t *row_ptr_d, const int32_t *col_d,
                       const double *aval_d, const double *dinv_rows_d, const double *dinv_d,
                       const int64_t *row_ptr_out_d, int32_t *col_out_d, double *aval_out_d,
                       double *s64_out_d, float *s32_out_d, void *stream)
{
    SGC_REQUIRE(n >= 0 && row_ptr_d && dinv_d && dinv_rows_d && row_ptr_out_d, "norm_build: bad arguments");
    if (n == 0) return SGC_OK;
    const bool compact = (row_ptr_out_d != row_ptr_d);
    norm_build_kernel<<<grid_for(ceil_div(n, ROWS_PER_BLOCK), 8), ROW_THREADS, 0, as_stream(stream)>>>(
        n, row_ptr_d, col_d, aval_d, dinv_rows_d, dinv_d, row_ptr_out_d, compact, col_out_d, aval_out_d,
        s64_out_d, s32_out_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_csr_to_coo(int64_t n, const int64_t *row_ptr_d, const int32_t *col_d, int64_t nnz,
                       int64_t *idx_out_d, void *stream)
{
    SGC_REQUIRE(n >= 0 && nnz >= 0 && row_ptr_d, "csr_to_coo: bad arguments");
    if (n == 0 || nnz == 0) return SGC_OK;
    SGC_REQUIRE(col_d && idx_out_d, "csr_to_coo: null arrays");
    csr_to_coo_kernel<<<grid_for(ceil_div(n, ROWS_PER_BLOCK), 8), ROW_THREADS, 0, as_stream(stream)>>>(
        n, row_ptr_d, col_d, nnz, idx_out_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_coo_to_csr(int64_t n_rows, int64_t nnz, const int64_t *row_idx_d,
                       const int64_t *col_idx_d, int64_t *row_ptr_out_d, int32_t *col_out_d,
                       int32_t *sorted_flag_d, void *stream)
{
    SGC_REQUIRE(n_rows >= 0 && nnz >= 0 && row_ptr_out_d && sorted_flag_d, "coo_to_csr: bad arguments");
    SGC_REQUIRE(nnz == 0 || (row_idx_d && col_idx_d && col_out_d), "coo_to_csr: null arrays");
    cudaStream_t s = as_stream(stream);
    const int32_t one = 1;
    SGC_CUDA_TRY(cudaMemcpyAsync(sorted_flag_d, &one, sizeof(one), cudaMemcpyHostToDevice, s));
    coo_to_csr_kernel<<<grid_for(ceil_div(nnz + 1, 256), 8), 256, 0, s>>>(
        n_rows, nnz, row_idx_d, col_idx_d, row_ptr_out_d, col_out_d, sorted_flag_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

}  // extern "C"
