// K1 / K2 fast path: element-parallel ("nonzero-split") adjacency normalisation for canonical input.
//
// Same results as the row-group kernels of csr_build.cu (which restate
// /root/reference/normalization.py:5-12 and stay as the general path), but the unit of work is a
// tile of EP_TILE stored entries instead of a row, so a 20 000-entry hub row costs as much as 200
// rows of 100 entries (the row-group kernels walk a hub with one 8-lane group: on the Reddit-shaped
// graph that tail was 80 % of their run time).  Every CTA finds the rows its tile touches with two
// binary searches over row_ptr, keeps that slice of row_ptr in shared memory, and every entry finds
// its row with a binary search there; all global accesses to col / val are coalesced.
//
//   aug_scan     : one pass over (col, val): proves the input canonical (strictly ascending columns
//                  in range), detects entries that A + I would drop (scipy drops exact zeros) and
//                  records per row where the self-loop goes: info[i] = -1 the row stores its
//                  diagonal (merged: a_ii + 1), k >= 0 insert before the row's k-th entry,
//                  INFO_END append.  Non-canonical input or drops -> flag, caller takes the general
//                  path.
//   aug_counts   : row_ptr of A + I = scan(len_i + (info_i != -1))
//   aug_build    : element-wise map to the output position; no compaction needed
//   degree       : r_i = sum in stored order (fp64, the order of scipy's coo matvec) with a warp per
//                  32 rows: long rows are loaded 32 entries at a time (coalesced) and added through
//                  a shuffle chain, so the order is sequential but the loads are not
//   norm_values  : S values fp64 / fp32 per entry + a flag if any product is exactly zero (scipy's
//                  CSR products would drop it: caller takes the compaction path)
#include "common.cuh"

namespace sgc {

constexpr int EP_THREADS = 256;
constexpr int EP_ITEMS = 16;
constexpr int EP_TILE = EP_THREADS * EP_ITEMS;     // stored entries per CTA
constexpr int EP_ROWS = EP_TILE;                   // row_ptr entries kept in shared memory
constexpr int32_t INFO_DIAG = -1;
constexpr int32_t INFO_END = 0x7f7f7f7f;           // cudaMemset(0x7f) pattern

constexpr int FLAG_NONCANONICAL = 1;
constexpr int FLAG_DROPS = 2;

// largest r in [lo, hi] with row_ptr[r] <= p (row_ptr[lo] <= p is the caller's invariant)
__device__ __forceinline__ int64_t last_row_at_or_before(const int64_t *__restrict__ row_ptr, int64_t lo,
                                                         int64_t hi, int64_t p)
{
    while (lo < hi) {
        const int64_t mid = lo + (hi - lo + 1) / 2;
        if (row_ptr[mid] <= p) lo = mid; else hi = mid - 1;
    }
    return lo;
}

struct RowTile {
    int64_t p0, p1;        // entries [p0, p1) of this CTA
    int64_t r_lo;          // row that owns entry p0
    int64_t nrows;         // rows r_lo .. r_lo + nrows - 1 intersect the tile (empty rows included)
    bool in_smem;          // their row_ptr slice (relative to p0) is in srp[0..nrows]
};

// srp: EP_ROWS + 1 ints of shared memory.  All threads of the CTA must call this.
__device__ __forceinline__ RowTile tile_setup(const int64_t *__restrict__ row_ptr, int64_t n, int64_t nnz,
                                              int32_t *srp)
{
    __shared__ int64_t s_r[2];
    RowTile t;
    t.p0 = (int64_t)blockIdx.x * EP_TILE;
    t.p1 = t.p0 + EP_TILE < nnz ? t.p0 + EP_TILE : nnz;
    if (threadIdx.x == 0) s_r[0] = last_row_at_or_before(row_ptr, 0, n - 1, t.p0);
    if (threadIdx.x == 32) s_r[1] = last_row_at_or_before(row_ptr, 0, n - 1, t.p1 - 1);
    __syncthreads();
    t.r_lo = s_r[0];
    t.nrows = s_r[1] - s_r[0] + 1;
    t.in_smem = t.nrows <= EP_ROWS;
    if (t.in_smem) {
        for (int i = threadIdx.x; i <= (int)t.nrows; i += EP_THREADS) {
            int64_t v = row_ptr[t.r_lo + i] - t.p0;
            srp[i] = (int32_t)(v < -1 ? -1 : (v > EP_TILE ? EP_TILE : v));
        }
    }
    __syncthreads();
    return t;
}

// row (global id) of entry p of the tile; *first = p is the first stored entry of that row
__device__ __forceinline__ int64_t tile_row_of(const RowTile &t, const int32_t *srp,
                                               const int64_t *__restrict__ row_ptr, int64_t p, bool *first)
{
    if (t.in_smem) {
        const int e = (int)(p - t.p0);
        int lo = 0, hi = (int)t.nrows - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (srp[mid] <= e) lo = mid; else hi = mid - 1;
        }
        *first = (srp[lo] == e);
        return t.r_lo + lo;
    }
    const int64_t r = last_row_at_or_before(row_ptr, t.r_lo, t.r_lo + t.nrows - 1, p);
    *first = (row_ptr[r] == p);
    return r;
}

__global__ void __launch_bounds__(EP_THREADS)
aug_scan_kernel(int64_t n, int64_t n_cols, int64_t doff, int64_t nnz, const int64_t *__restrict__ row_ptr,
                const int32_t *__restrict__ col, const double *__restrict__ val, int32_t *__restrict__ info,
                int64_t *__restrict__ flags)
{
    __shared__ int32_t srp[EP_ROWS + 1];
    const RowTile t = tile_setup(row_ptr, n, nnz, srp);
    int bad = 0;
    for (int64_t p = t.p0 + threadIdx.x; p < t.p1; p += EP_THREADS) {
        const int32_t c = col[p];
        const double v = val[p];
        bool first;
        const int64_t row = tile_row_of(t, srp, row_ptr, p, &first);
        const int64_t dc = row + doff;
        const int32_t prev = (first || p == 0) ? -1 : col[p - 1];
        if (c < 0 || c >= n_cols || (!first && prev >= c)) bad |= FLAG_NONCANONICAL;
        if (c == dc) {
            info[row] = INFO_DIAG;
            if (!((v + 1.0) != 0.0)) bad |= FLAG_DROPS;
        } else {
            if (!(v != 0.0)) bad |= FLAG_DROPS;
            if (c > dc && (first || prev < dc)) info[row] = (int32_t)(p - row_ptr[row]);
        }
    }
    bad = __reduce_or_sync(0xffffffffu, bad);
    if (bad && (threadIdx.x & 31) == 0) atomicOr(reinterpret_cast<unsigned long long *>(flags), (unsigned long long)bad);
}

__global__ void __launch_bounds__(256)
aug_counts_kernel(int64_t n, int64_t nnz, const int64_t *__restrict__ row_ptr, const int32_t *__restrict__ info,
                  int64_t *__restrict__ counts, int64_t *__restrict__ flags)
{
    bool bad = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t c = 0;
        if (i < n) {
            const int64_t b = row_ptr[i], e = row_ptr[i + 1];
            if (e < b || b < 0 || e > nnz) bad = true;
            c = (e - b) + (info[i] != INFO_DIAG ? 1 : 0);
        }
        counts[i] = c;
    }
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0)
        atomicOr(reinterpret_cast<unsigned long long *>(flags), (unsigned long long)FLAG_NONCANONICAL);
}

__global__ void store_total_kernel(const int64_t *__restrict__ scanned, int64_t n, int64_t *__restrict__ out)
{
    *out = scanned[n];
}

// rows whose row_ptr is not monotone were flagged by aug_counts; this kernel only checks the ends
__global__ void __launch_bounds__(256)
row_ptr_ends_kernel(int64_t n, int64_t nnz, const int64_t *__restrict__ row_ptr, int64_t *__restrict__ flags)
{
    if (threadIdx.x == 0 && blockIdx.x == 0 && (row_ptr[0] != 0 || row_ptr[n] != nnz))
        atomicOr(reinterpret_cast<unsigned long long *>(flags), (unsigned long long)FLAG_NONCANONICAL);
}

__global__ void __launch_bounds__(EP_THREADS)
aug_build_kernel(int64_t n, int64_t doff, int64_t nnz, const int64_t *__restrict__ row_ptr,
                 const int32_t *__restrict__ col, const double *__restrict__ val,
                 const int32_t *__restrict__ info, const int64_t *__restrict__ out_ptr,
                 int32_t *__restrict__ col_out, double *__restrict__ val_out)
{
    __shared__ int32_t srp[EP_ROWS + 1];
    const RowTile t = tile_setup(row_ptr, n, nnz, srp);
    for (int64_t p = t.p0 + threadIdx.x; p < t.p1; p += EP_THREADS) {
        const int32_t c = col[p];
        double v = val[p];
        bool first;
        const int64_t row = tile_row_of(t, srp, row_ptr, p, &first);
        const int32_t dc = (int32_t)(row + doff);
        const int32_t where = info[row];
        const int64_t k = p - row_ptr[row];
        int64_t o = out_ptr[row] + k;
        if (where >= 0 && k >= where) {
            if (k == where) { col_out[o] = dc; val_out[o] = 1.0; }     // the inserted self-loop
            o += 1;
        }
        if (c == dc) v += 1.0;                                          // a_ii + 1 (normalization.py:6)
        col_out[o] = c;
        val_out[o] = v;
    }
}

// self-loops appended after the last stored entry (and the only entry of empty rows)
__global__ void __launch_bounds__(256)
aug_append_kernel(int64_t n, int64_t doff, const int32_t *__restrict__ info, const int64_t *__restrict__ out_ptr,
                  int32_t *__restrict__ col_out, double *__restrict__ val_out)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (info[i] == INFO_END) {
            const int64_t o = out_ptr[i + 1] - 1;
            col_out[o] = (int32_t)(i + doff);
            val_out[o] = 1.0;
        }
    }
}

// One warp per 32 consecutive rows.  Rows of at most SHORT entries are summed by their own lane;
// longer rows by the whole warp, one after another: 32 entries per coalesced load; integer-valued rows
// are reduced in parallel (exact, see below), real-valued rows in stored order through a shuffle
// chain (all lanes carry the same running sum), next batch prefetched.
__global__ void __launch_bounds__(256)
degree_warp_kernel(int64_t n, const int64_t *__restrict__ row_ptr, const double *__restrict__ aval,
                   double *__restrict__ rowsum, double *__restrict__ dinv)
{
    constexpr int SHORT = 8;
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w * 32 < n; w += warps) {
        const int64_t i = w * 32 + lane;
        int64_t b = 0, e = 0;
        if (i < n) { b = row_ptr[i]; e = row_ptr[i + 1]; }
        double r = 0.0;
        if (e - b <= SHORT)
            for (int64_t p = b; p < e; ++p) r = __dadd_rn(r, aval[p]);
        unsigned todo = __ballot_sync(0xffffffffu, e - b > SHORT);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t rb = __shfl_sync(0xffffffffu, b, src), re = __shfl_sync(0xffffffffu, e, src);
            // Exact shortcut: if every entry of the row is an integer of magnitude <= 2^20 (binary and
            // count-weighted graphs: every BASELINE shape), all partial sums of fewer than 2^31 of them
            // are exact in fp64, so ANY summation order gives the bits of the sequential one -- the
            // lanes sum strided pieces with independent loads and combine by a tree.
            double part = 0.0;
            bool ints = true;
            for (int64_t q = rb + lane; q < re; q += 32) {
                const double v = aval[q];
                ints &= (v == rint(v)) && (fabs(v) <= 1048576.0);
                part += v;
            }
            double acc = 0.0;
            if (__all_sync(0xffffffffu, ints)) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
                acc = part;
            } else {
                // real-valued weights (TextSGC's PMI / TF-IDF graph): stored order, through a shuffle
                // chain; the row was just read, so these loads hit L1 / L2
                double nxt = (rb + lane < re) ? aval[rb + lane] : 0.0;
                for (int64_t q = rb; q < re; q += 32) {
                    const double cur = nxt;
                    if (q + 32 < re) nxt = (q + 32 + lane < re) ? aval[q + 32 + lane] : 0.0;
                    const int cnt = (re - q < 32) ? (int)(re - q) : 32;
                    if (cnt == 32) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) acc = __dadd_rn(acc, __shfl_sync(0xffffffffu, cur, j));
                    } else {
                        for (int j = 0; j < cnt; ++j) acc = __dadd_rn(acc, __shfl_sync(0xffffffffu, cur, j));
                    }
                }
            }
            if (lane == src) r = acc;
        }
        if (i < n) {
            if (rowsum) rowsum[i] = r;
            dinv[i] = rsqrt_refined(r);
        }
    }
}

__global__ void __launch_bounds__(EP_THREADS)
norm_values_kernel(int64_t n, int64_t nnz, const int64_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                   const double *__restrict__ aval, const double *__restrict__ dinv_rows,
                   const double *__restrict__ dinv, int order, double *__restrict__ s64_out,
                   float *__restrict__ s32_out, int64_t *__restrict__ zero_flag)
{
    __shared__ int32_t srp[EP_ROWS + 1];
    const RowTile t = tile_setup(row_ptr, n, nnz, srp);
    bool zero = false;
    for (int64_t p = t.p0 + threadIdx.x; p < t.p1; p += EP_THREADS) {
        const int32_t c = col[p];
        const double a = aval[p];
        bool first;
        const int64_t row = tile_row_of(t, srp, row_ptr, p, &first);
        const double di = dinv_rows[row], dj = dinv[c];
        // scipy's csr_matmat drops products that are exactly zero (first D*Ahat, then (D*Ahat)*D;
        // TextSGC's normalize_adj multiplies the column factor first: (Ahat*D)^T*D)
        const double left = order == 0 ? __dmul_rn(di, a) : __dmul_rn(a, dj);
        const double s = __dmul_rn(left, order == 0 ? dj : di);
        if (!(left != 0.0) || !(s != 0.0)) zero = true;
        if (s64_out) s64_out[p] = s;
        if (s32_out) s32_out[p] = __double2float_rn(s);                 // utils.py:25 astype(float32)
    }
    if (__any_sync(0xffffffffu, zero) && (threadIdx.x & 31) == 0)
        atomicOr(reinterpret_cast<unsigned long long *>(zero_flag), 1ull);
}

// Dense copy of selected columns of a CSR matrix: out[row, colmap[c]] = val for every stored entry
// whose column is selected (colmap[c] >= 0).  out must be zero-filled.  (TextSGC feeds
// dense_adj[:, idx] to torch.spmm, downstream/TextSGC/utils.py:135.)
__global__ void __launch_bounds__(EP_THREADS)
gather_columns_kernel(int64_t n, int64_t nnz, const int64_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                      const float *__restrict__ val, const int32_t *__restrict__ colmap, float *__restrict__ out,
                      int64_t ld)
{
    __shared__ int32_t srp[EP_ROWS + 1];
    const RowTile t = tile_setup(row_ptr, n, nnz, srp);
    for (int64_t p = t.p0 + threadIdx.x; p < t.p1; p += EP_THREADS) {
        const int32_t k = colmap[col[p]];
        if (k < 0) continue;
        bool first;
        const int64_t row = tile_row_of(t, srp, row_ptr, p, &first);
        out[row * ld + k] = val[p];
    }
}

}  // namespace sgc

using namespace sgc;

extern "C" {

int32_t sgc_csr_augment_count(int64_t n, int64_t n_cols, int64_t diag_offset, int64_t nnz,
                              const int64_t *row_ptr_d, const int32_t *col_d, const double *val_d,
                              int32_t *info_d, int64_t *row_ptr_out_d, int64_t *summary_d, void *ws_d,
                              size_t ws_bytes, void *stream)
{
    SGC_REQUIRE(n >= 0 && nnz >= 0 && row_ptr_d && row_ptr_out_d && summary_d, "csr_augment_count: bad arguments");
    SGC_REQUIRE(n == 0 || info_d, "csr_augment_count: info_d is null");
    SGC_REQUIRE(nnz == 0 || (col_d && val_d), "csr_augment_count: null arrays");
    cudaStream_t s = as_stream(stream);
    SGC_CUDA_TRY(cudaMemsetAsync(summary_d, 0, 2 * sizeof(int64_t), s));
    if (n > 0) SGC_CUDA_TRY(cudaMemsetAsync(info_d, 0x7f, (size_t)n * sizeof(int32_t), s));
    if (n > 0) {
        row_ptr_ends_kernel<<<1, 32, 0, s>>>(n, nnz, row_ptr_d, summary_d);
        SGC_LAUNCH_CHECK();
    }
    if (n > 0 && nnz > 0) {
        aug_scan_kernel<<<(unsigned)ceil_div(nnz, EP_TILE), EP_THREADS, 0, s>>>(
            n, n_cols, diag_offset, nnz, row_ptr_d, col_d, val_d, info_d, summary_d);
        SGC_LAUNCH_CHECK();
    }
    aug_counts_kernel<<<grid_for(ceil_div(n + 1, 256), 8), 256, 0, s>>>(n, nnz, row_ptr_d, info_d, row_ptr_out_d,
                                                                      summary_d);
    SGC_LAUNCH_CHECK();
    int32_t rc = scan_exclusive_i64(row_ptr_out_d, n + 1, ws_d, ws_bytes, s);
    if (rc != SGC_OK) return rc;
    store_total_kernel<<<1, 1, 0, s>>>(row_ptr_out_d, n, summary_d + 1);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_csr_augment_build(int64_t n, int64_t diag_offset, int64_t nnz, const int64_t *row_ptr_d,
                              const int32_t *col_d, const double *val_d, const int32_t *info_d,
                              const int64_t *row_ptr_out_d, int32_t *col_out_d, double *aval_out_d,
                              void *stream)
{
    SGC_REQUIRE(n >= 0 && nnz >= 0 && row_ptr_d && row_ptr_out_d, "csr_augment_build: bad arguments");
    if (n == 0) return SGC_OK;
    SGC_REQUIRE(info_d && col_out_d && aval_out_d, "csr_augment_build: null arrays");
    cudaStream_t s = as_stream(stream);
    if (nnz > 0) {
        aug_build_kernel<<<(unsigned)ceil_div(nnz, EP_TILE), EP_THREADS, 0, s>>>(
            n, diag_offset, nnz, row_ptr_d, col_d, val_d, info_d, row_ptr_out_d, col_out_d, aval_out_d);
        SGC_LAUNCH_CHECK();
    }
    aug_append_kernel<<<grid_for(ceil_div(n, 256), 8), 256, 0, s>>>(n, diag_offset, info_d, row_ptr_out_d,
                                                                   col_out_d, aval_out_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_degree_inv_sqrt(int64_t n, const int64_t *row_ptr_d, const double *aval_d,
                            double *rowsum_d, double *dinv_d, void *stream)
{
    SGC_REQUIRE(n >= 0 && row_ptr_d && dinv_d, "degree_inv_sqrt: bad arguments");
    if (n == 0) return SGC_OK;
    degree_warp_kernel<<<grid_for(ceil_div(n, 256), 8), 256, 0, as_stream(stream)>>>(n, row_ptr_d, aval_d, rowsum_d,
                                                                                     dinv_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_norm_values(int64_t n, int64_t nnz, const int64_t *row_ptr_d, const int32_t *col_d,
                        const double *aval_d, const double *dinv_rows_d, const double *dinv_d,
                        int32_t order, double *s64_out_d, float *s32_out_d, int64_t *zero_flag_d,
                        void *stream)
{
    SGC_REQUIRE(order == 0 || order == 1, "norm_values: order must be 0 or 1");
    SGC_REQUIRE(n >= 0 && nnz >= 0 && row_ptr_d && dinv_rows_d && dinv_d && zero_flag_d, "norm_values: bad arguments");
    cudaStream_t s = as_stream(stream);
    SGC_CUDA_TRY(cudaMemsetAsync(zero_flag_d, 0, sizeof(int64_t), s));
    if (n == 0 || nnz == 0) return SGC_OK;
    SGC_REQUIRE(col_d && aval_d, "norm_values: null arrays");
    norm_values_kernel<<<(unsigned)ceil_div(nnz, EP_TILE), EP_THREADS, 0, s>>>(
        n, nnz, row_ptr_d, col_d, aval_d, dinv_rows_d, dinv_d, order, s64_out_d, s32_out_d, zero_flag_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_gather_columns_dense(int64_t n, int64_t nnz, const int64_t *row_ptr_d, const int32_t *col_d,
                                 const float *val_d, const int32_t *colmap_d, float *out_d, int64_t ld,
                                 void *stream)
{
    SGC_REQUIRE(n >= 0 && nnz >= 0 && row_ptr_d && colmap_d, "gather_columns_dense: bad arguments");
    if (n == 0 || nnz == 0) return SGC_OK;
    SGC_REQUIRE(col_d && val_d && out_d && ld >= 1, "gather_columns_dense: null arrays");
    gather_columns_kernel<<<(unsigned)ceil_div(nnz, EP_TILE), EP_THREADS, 0, as_stream(stream)>>>(
        n, nnz, row_ptr_d, col_d, val_d, colmap_d, out_d, ld);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

}  // extern "C"
