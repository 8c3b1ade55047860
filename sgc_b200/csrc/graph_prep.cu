// Graph preparation on the GPU: the steps immediately before the normalisation in the reference's
// loaders:
//
//   adj + adj.T                                  /root/reference/utils.py:116   (Reddit)
//   adj + adj.T*(adj.T > adj) - adj*(adj.T > adj) /root/reference/utils.py:63    (citation: max(A, A^T))
//   adj[train_index, :][:, train_index]          /root/reference/utils.py:117   (inductive sub-graph)
//
// plus a fast canonicalisation of unsorted / duplicated CSR input.  All of them reduce to one
// primitive: a stable LSD radix sort of (key, int32, fp64) triples by an int32 key, 8 bits per
// pass.  Sorting a COO stream by column and then by row leaves it in row-major order with
// ascending columns and equal (row, col) pairs adjacent in their original order, after which one
// flag + scan + fill pass merges duplicates (sum in stored order = scipy's sum_duplicates; max for
// the citation symmetrisation) and rebuilds row_ptr.  Everything is HBM-bound streaming work:
// a pass reads and writes 16 bytes per element.
#include "common.cuh"

namespace sgc {

constexpr int RS_BITS = 8;
constexpr int RS_RADIX = 1 << RS_BITS;
constexpr int RS_THREADS = 256;
constexpr int RS_CHUNKS = 16;                          // 256-element chunks per block tile
constexpr int RS_TILE = RS_THREADS * RS_CHUNKS;

__global__ void __launch_bounds__(RS_THREADS)
radix_hist_kernel(const int32_t *__restrict__ key, int64_t n, int shift, int64_t n_blocks,
                  int64_t *__restrict__ counts)
{
    __shared__ unsigned int hist[RS_RADIX];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * RS_TILE;
#pragma unroll 4
    for (int c = 0; c < RS_CHUNKS; ++c) {
        const int64_t i = base + (int64_t)c * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&hist[((uint32_t)key[i] >> shift) & (RS_RADIX - 1)], 1u);
    }
    __syncthreads();
    counts[(int64_t)threadIdx.x * n_blocks + blockIdx.x] = hist[threadIdx.x];   // digit-major
}

__global__ void __launch_bounds__(RS_THREADS)
radix_scatter_kernel(const int32_t *__restrict__ key, const int32_t *__restrict__ p32,
                     const double *__restrict__ p64, int64_t n, int shift, int64_t n_blocks,
                     const int64_t *__restrict__ offsets, int32_t *__restrict__ key_out,
                     int32_t *__restrict__ p32_out, double *__restrict__ p64_out)
{
    __shared__ int64_t next[RS_RADIX];                              // next free slot per digit
    __shared__ unsigned short wcount[RS_THREADS / 32][RS_RADIX];    // per-warp digit counts of a chunk
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    next[threadIdx.x] = offsets[(int64_t)threadIdx.x * n_blocks + blockIdx.x];
#pragma unroll
    for (int w = 0; w < RS_THREADS / 32; ++w) wcount[w][threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * RS_TILE;
    for (int c = 0; c < RS_CHUNKS; ++c) {
        const int64_t i = base + (int64_t)c * RS_THREADS + threadIdx.x;
        const bool live = i < n;
        int32_t k = 0;
        uint32_t d = 0xffffffffu;                                   // dead lanes match each other only
        if (live) { k = key[i]; d = ((uint32_t)k >> shift) & (RS_RADIX - 1); }
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        if (live && rank == 0) wcount[warp][d] = (unsigned short)__popc(peers);
        __syncthreads();
        if (live) {
            int64_t pos = next[d] + rank;
            for (int w = 0; w < warp; ++w) pos += wcount[w][d];
            key_out[pos] = k;
            if (p32) p32_out[pos] = p32[i];
            if (p64) p64_out[pos] = p64[i];
        }
        __syncthreads();
        {
            unsigned int tot = 0;
#pragma unroll
            for (int w = 0; w < RS_THREADS / 32; ++w) { tot += wcount[w][threadIdx.x]; wcount[w][threadIdx.x] = 0; }
            next[threadIdx.x] += tot;
        }
        __syncthreads();
    }
}

// rows of a CSR as an int32 array (COO expansion), 8 lanes per row
__global__ void __launch_bounds__(256)
expand_rows_kernel(int64_t n, const int64_t *__restrict__ row_ptr, int32_t row_offset,
                   int32_t *__restrict__ rows_out)
{
    const int gl = threadIdx.x & 7;
    const int64_t groups = (int64_t)gridDim.x * 32;
    for (int64_t i = (int64_t)blockIdx.x * 32 + threadIdx.x / 8; i < n; i += groups) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        for (int64_t p = b + gl; p < e; p += 8) rows_out[p] = (int32_t)i + row_offset;
    }
}

// Sorted (row, col) stream -> run heads.  A head is the first element of a run of equal pairs; its
// combined value is the run's sum in stored order (op 0) or maximum (op 1).  keep[i] = 1 for heads
// whose combined value survives (drop_zeros: scipy's binary ops drop results that compare == 0).
__global__ void __launch_bounds__(256)
reduce_flag_kernel(int64_t n, const int32_t *__restrict__ row, const int32_t *__restrict__ col,
                   const double *__restrict__ val, int op, int drop_zeros,
                   int64_t *__restrict__ keep, double *__restrict__ combined)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n;
         i += (int64_t)gridDim.x * blockDim.x) {
        if (i == n) { keep[n] = 0; continue; }
        const int32_t r = row[i], c = col[i];
        const bool head = (i == 0) || row[i - 1] != r || col[i - 1] != c;
        int64_t k = 0;
        if (head) {
            double acc = val[i];
            for (int64_t j = i + 1; j < n && row[j] == r && col[j] == c; ++j)
                acc = (op == 1) ? fmax(acc, val[j]) : __dadd_rn(acc, val[j]);
            combined[i] = acc;
            k = (drop_zeros && acc == 0.0) ? 0 : 1;
        }
        keep[i] = k;
    }
}

// pos = exclusive scan of keep.  Kept heads write (col, value); every kept head also writes the
// row pointers of the rows that start between the previous kept entry's row and its own.
__global__ void __launch_bounds__(256)
reduce_fill_kernel(int64_t n, int64_t n_rows, const int32_t *__restrict__ row,
                   const int32_t *__restrict__ col, const double *__restrict__ combined,
                   const int64_t *__restrict__ pos, int64_t *__restrict__ row_ptr,
                   int32_t *__restrict__ col_out, double *__restrict__ val_out)
{
    const int64_t total = pos[n];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n;
         i += (int64_t)gridDim.x * blockDim.x) {
        if (i == n) {
            // rows after the last kept entry (and all rows when nothing is kept)
            int64_t j = n - 1;
            while (j >= 0 && pos[j + 1] == pos[j]) --j;
            const int64_t last_row = (j >= 0) ? row[j] : -1;
            for (int64_t r = last_row + 1; r <= n_rows; ++r) row_ptr[r] = total;
            continue;
        }
        if (pos[i + 1] == pos[i]) continue;                   // not a kept head
        const int64_t q = pos[i];
        col_out[q] = col[i];
        val_out[q] = combined[i];
        // previous kept head
        int64_t j = i - 1;
        while (j >= 0 && pos[j + 1] == pos[j]) --j;
        const int64_t prev_row = (j >= 0) ? row[j] : -1;
        for (int64_t r = prev_row + 1; r <= row[i]; ++r) row_ptr[r] = q;
    }
}

// new_col = map[old_col] (or -1); per-row count of surviving entries for the row selection idx
__global__ void __launch_bounds__(256)
subgraph_count_kernel(int64_t m, const int64_t *__restrict__ idx, const int64_t *__restrict__ row_ptr,
                      const int32_t *__restrict__ col, const int32_t *__restrict__ colmap,
                      int64_t *__restrict__ counts)
{
    const int lane = threadIdx.x & 31, gl = lane & 7;
    const unsigned gm = 0xffu << (lane & ~7);
    const int64_t groups = (int64_t)gridDim.x * 32;
    for (int64_t i = (int64_t)blockIdx.x * 32 + threadIdx.x / 8; i < m; i += groups) {
        const int64_t r = idx[i];
        const int64_t b = row_ptr[r], e = row_ptr[r + 1];
        int cnt = 0;
        for (int64_t p = b + gl; p < e; p += 8) cnt += (colmap[col[p]] >= 0) ? 1 : 0;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) cnt += __shfl_xor_sync(gm, cnt, o);
        if (gl == 0) counts[i] = cnt;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) counts[m] = 0;
}

__global__ void __launch_bounds__(256)
subgraph_fill_kernel(int64_t m, const int64_t *__restrict__ idx, const int64_t *__restrict__ row_ptr,
                     const int32_t *__restrict__ col, const double *__restrict__ val,
                     const int32_t *__restrict__ colmap, const int64_t *__restrict__ out_ptr,
                     int32_t *__restrict__ col_out, double *__restrict__ val_out)
{
    const int lane = threadIdx.x & 31, gl = lane & 7;
    const unsigned gm = 0xffu << (lane & ~7);
    const int gshift = lane & ~7;
    const int64_t groups = (int64_t)gridDim.x * 32;
    for (int64_t i = (int64_t)blockIdx.x * 32 + threadIdx.x / 8; i < m; i += groups) {
        const int64_t r = idx[i];
        const int64_t b = row_ptr[r], e = row_ptr[r + 1];
        int64_t w = out_ptr[i];
        for (int64_t p0 = b; p0 < e; p0 += 8) {
            const int64_t p = p0 + gl;
            int32_t nc = -1;
            double v = 0.0;
            if (p < e) { nc = colmap[col[p]]; v = val[p]; }
            const unsigned kept = (__ballot_sync(gm, nc >= 0) & gm) >> gshift;
            if (nc >= 0) {
                const int64_t q = w + __popc(kept & ((1u << gl) - 1u));
                col_out[q] = nc;
                val_out[q] = v;
            }
            w += __popc(kept);
        }
    }
}

__global__ void __launch_bounds__(256)
colmap_kernel(int64_t m, const int64_t *__restrict__ idx, int32_t *__restrict__ colmap)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m;
         i += (int64_t)gridDim.x * blockDim.x)
        colmap[idx[i]] = (int32_t)i;
}

static inline size_t al256(size_t v) { return (v + 255) / 256 * 256; }

}  // namespace sgc

using namespace sgc;

extern "C" {

size_t sgc_sort_workspace_bytes(int64_t n)
{
    if (n < 0) return 0;
    const int64_t n_blocks = ceil_div(n > 0 ? n : 1, RS_TILE);
    return al256((size_t)n * 4) * 2 + al256((size_t)n * 8) + al256((size_t)RS_RADIX * n_blocks * 8) +
           scan_workspace_bytes(RS_RADIX * n_blocks) + 256;
}

int32_t sgc_sort_by_key_i32(int64_t n, int32_t key_bits, const int32_t *key_in_d, const int32_t *p32_in_d,
                            const double *p64_in_d, int32_t *key_out_d, int32_t *p32_out_d,
                            double *p64_out_d, void *ws_d, size_t ws_bytes, void *stream)
{
    SGC_REQUIRE(n >= 0 && key_bits >= 1 && key_bits <= 31, "sort_by_key: bad arguments");
    if (n == 0) return SGC_OK;
    SGC_REQUIRE(key_in_d && key_out_d && key_in_d != key_out_d, "sort_by_key: keys must be out of place");
    SGC_REQUIRE((p32_in_d == nullptr) == (p32_out_d == nullptr) && (p64_in_d == nullptr) == (p64_out_d == nullptr),
                "sort_by_key: payload in/out arrays must come in pairs");
    if (!ws_d || ws_bytes < sgc_sort_workspace_bytes(n)) {
        set_error("sort_by_key: workspace too small");
        return SGC_ERR_WORKSPACE;
    }
    cudaStream_t s = as_stream(stream);
    const int64_t n_blocks = ceil_div(n, RS_TILE);
    SGC_REQUIRE(n_blocks < 0x7fffffffLL, "sort_by_key: too many elements");
    char *ws = reinterpret_cast<char *>(ws_d);
    int32_t *tmp_key = reinterpret_cast<int32_t *>(ws);          ws += al256((size_t)n * 4);
    int32_t *tmp_p32 = reinterpret_cast<int32_t *>(ws);          ws += al256((size_t)n * 4);
    double *tmp_p64 = reinterpret_cast<double *>(ws);            ws += al256((size_t)n * 8);
    int64_t *counts = reinterpret_cast<int64_t *>(ws);           ws += al256((size_t)RS_RADIX * n_blocks * 8);
    void *scan_ws = ws;
    const size_t scan_bytes = scan_workspace_bytes(RS_RADIX * n_blocks);
    const int passes = (key_bits + RS_BITS - 1) / RS_BITS;
    const int32_t *ksrc = key_in_d, *psrc = p32_in_d;
    const double *dsrc = p64_in_d;
    for (int p = 0; p < passes; ++p) {
        const bool to_out = ((passes - p) & 1) != 0;               // the last pass lands in the outputs
        int32_t *kdst = to_out ? key_out_d : tmp_key;
        int32_t *pdst = to_out ? p32_out_d : tmp_p32;
        double *ddst = to_out ? p64_out_d : tmp_p64;
        radix_hist_kernel<<<(unsigned)n_blocks, RS_THREADS, 0, s>>>(ksrc, n, p * RS_BITS, n_blocks, counts);
        SGC_LAUNCH_CHECK();
        int32_t rc = scan_exclusive_i64(counts, RS_RADIX * n_blocks, scan_ws, scan_bytes, s);
        if (rc != SGC_OK) return rc;
        radix_scatter_kernel<<<(unsigned)n_blocks, RS_THREADS, 0, s>>>(ksrc, psrc, dsrc, n, p * RS_BITS, n_blocks,
                                                                     counts, kdst, psrc ? pdst : nullptr,
                                                                     dsrc ? ddst : nullptr);
        SGC_LAUNCH_CHECK();
        ksrc = kdst; psrc = psrc ? pdst : nullptr; dsrc = dsrc ? ddst : nullptr;
    }
    return SGC_OK;
}

int32_t sgc_csr_expand_rows(int64_t n, const int64_t *row_ptr_d, int32_t row_offset, int32_t *rows_out_d,
                            void *stream)
{
    SGC_REQUIRE(n >= 0 && row_ptr_d, "csr_expand_rows: bad arguments");
    if (n == 0) return SGC_OK;
    if (!rows_out_d) return SGC_OK;              // legal when the matrix stores nothing
    expand_rows_kernel<<<grid_for(ceil_div(n, 32), 8), 256, 0, as_stream(stream)>>>(n, row_ptr_d, row_offset, rows_out_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

size_t sgc_coo_reduce_workspace_bytes(int64_t n)
{
    if (n < 0) return 0;
    return al256((size_t)(n + 1) * 8) + al256((size_t)n * 8) + scan_workspace_bytes(n + 1) + 256;
}

int32_t sgc_coo_reduce_to_csr(int64_t n, int64_t n_rows, const int32_t *row_d, const int32_t *col_d,
                              const double *val_d, int32_t op, int32_t drop_zeros, int64_t *row_ptr_out_d,
                              int32_t *col_out_d, double *val_out_d, void *ws_d, size_t ws_bytes, void *stream)
{
    SGC_REQUIRE(n >= 0 && n_rows >= 0 && row_ptr_out_d && (op == 0 || op == 1), "coo_reduce_to_csr: bad arguments");
    SGC_REQUIRE(n == 0 || (row_d && col_d && val_d && col_out_d && val_out_d), "coo_reduce_to_csr: null arrays");
    if (!ws_d || ws_bytes < sgc_coo_reduce_workspace_bytes(n)) {
        set_error("coo_reduce_to_csr: workspace too small");
        return SGC_ERR_WORKSPACE;
    }
    cudaStream_t s = as_stream(stream);
    char *ws = reinterpret_cast<char *>(ws_d);
    int64_t *keep = reinterpret_cast<int64_t *>(ws);             ws += al256((size_t)(n + 1) * 8);
    double *combined = reinterpret_cast<double *>(ws);           ws += al256((size_t)n * 8);
    const int grid = grid_for(ceil_div(n + 1, 256), 8);
    reduce_flag_kernel<<<grid, 256, 0, s>>>(n, row_d, col_d, val_d, op, drop_zeros, keep, combined);
    SGC_LAUNCH_CHECK();
    int32_t rc = scan_exclusive_i64(keep, n + 1, ws, scan_workspace_bytes(n + 1), s);
    if (rc != SGC_OK) return rc;
    reduce_fill_kernel<<<grid, 256, 0, s>>>(n, n_rows, row_d, col_d, combined, keep, row_ptr_out_d, col_out_d, val_out_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_subgraph_colmap(int64_t n_cols, int64_t m, const int64_t *idx_d, int32_t *colmap_d, void *stream)
{
    SGC_REQUIRE(n_cols >= 0 && m >= 0 && (n_cols == 0 || colmap_d), "subgraph_colmap: bad arguments");
    cudaStream_t s = as_stream(stream);
    if (n_cols > 0) SGC_CUDA_TRY(cudaMemsetAsync(colmap_d, 0xff, (size_t)n_cols * sizeof(int32_t), s));   // -1
    if (m == 0) return SGC_OK;
    SGC_REQUIRE(idx_d, "subgraph_colmap: idx_d is null");
    colmap_kernel<<<grid_for(ceil_div(m, 256), 8), 256, 0, s>>>(m, idx_d, colmap_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_subgraph_count(int64_t m, const int64_t *idx_d, const int64_t *row_ptr_d, const int32_t *col_d,
                           const int32_t *colmap_d, int64_t *row_ptr_out_d, void *ws_d, size_t ws_bytes,
                           void *stream)
{
    SGC_REQUIRE(m >= 0 && row_ptr_d && row_ptr_out_d && (m == 0 || (idx_d && colmap_d)), "subgraph_count: bad arguments");
    cudaStream_t s = as_stream(stream);
    subgraph_count_kernel<<<grid_for(ceil_div(m > 0 ? m : 1, 32), 8), 256, 0, s>>>(m, idx_d, row_ptr_d, col_d, colmap_d,
                                                                                  row_ptr_out_d);
    SGC_LAUNCH_CHECK();
    return scan_exclusive_i64(row_ptr_out_d, m + 1, ws_d, ws_bytes, s);
}

int32_t sgc_subgraph_fill(int64_t m, const int64_t *idx_d, const int64_t *row_ptr_d, const int32_t *col_d,
                          const double *val_d, const int32_t *colmap_d, const int64_t *row_ptr_out_d,
                          int32_t *col_out_d, double *val_out_d, void *stream)
{
    SGC_REQUIRE(m >= 0 && row_ptr_d && row_ptr_out_d, "subgraph_fill: bad arguments");
    if (m == 0) return SGC_OK;
    SGC_REQUIRE(idx_d && colmap_d, "subgraph_fill: null arrays");
    subgraph_fill_kernel<<<grid_for(ceil_div(m, 32), 8), 256, 0, as_stream(stream)>>>(
        m, idx_d, row_ptr_d, col_d, val_d, colmap_d, row_ptr_out_d, col_out_d, val_out_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

}  // extern "C"
