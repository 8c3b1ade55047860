// Feature-side siblings of the path for SPARSE feature matrices:
//   row_normalize            /root/reference/normalization.py:21-28   (citation loaders, utils.py:20)
//   first hop with sparse X  utils.py:95 on the densified matrix of utils.py:74
// Cora's feature matrix is 1.27 % dense, Citeseer's 0.85 %: the reference densifies it on the host
// (utils.py:74) and multiplies the dense matrix.  Here the row-normalised CSR stays on the GPU and the
// first hop accumulates only the stored entries, in the same per-element order (ascending
// neighbour index) as the dense kernel -- terms with x == 0 contribute exactly nothing there, so the
// results agree bit for bit.
#include "common.cuh"

namespace sgc {

// rowsum in the matrix dtype, sequential in stored order (scipy: csr @ ones), then r^-1 with
// inf -> 0 (normalization.py:24-25) and x * r^-1 in place.  One warp per 32 rows, long rows through
// the same coalesced-load + shuffle-chain scheme as the degree kernel (csr_fast.cu).
template <typename T> __device__ __forceinline__ T recip_rn(T r);
template <> __device__ __forceinline__ float recip_rn<float>(float r) { return __frcp_rn(r); }
template <> __device__ __forceinline__ double recip_rn<double>(double r) { return __drcp_rn(r); }
template <typename T> __device__ __forceinline__ T add_rn(T a, T b);
template <> __device__ __forceinline__ float add_rn<float>(float a, float b) { return __fadd_rn(a, b); }
template <> __device__ __forceinline__ double add_rn<double>(double a, double b) { return __dadd_rn(a, b); }
template <typename T> __device__ __forceinline__ T mul_rn(T a, T b);
template <> __device__ __forceinline__ float mul_rn<float>(float a, float b) { return __fmul_rn(a, b); }
template <> __device__ __forceinline__ double mul_rn<double>(double a, double b) { return __dmul_rn(a, b); }

template <typename T>
__global__ void __launch_bounds__(256)
row_normalize_kernel(int64_t n, const int64_t *__restrict__ row_ptr, const T *__restrict__ val,
                     T *__restrict__ out, T *__restrict__ rinv_out, int64_t *__restrict__ zero_flag)
{
    constexpr int SHORT = 8;
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    bool zero = false;
    for (int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w * 32 < n; w += warps) {
        const int64_t i = w * 32 + lane;
        int64_t b = 0, e = 0;
        if (i < n) { b = row_ptr[i]; e = row_ptr[i + 1]; }
        T r = 0;
        if (e - b <= SHORT)
            for (int64_t p = b; p < e; ++p) r = add_rn<T>(r, val[p]);
        unsigned todo = __ballot_sync(0xffffffffu, e - b > SHORT);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t rb = __shfl_sync(0xffffffffu, b, src), re = __shfl_sync(0xffffffffu, e, src);
            T acc = 0;
            for (int64_t q = rb; q < re; q += 32) {
                const T cur = (q + lane < re) ? val[q + lane] : (T)0;
                const int cnt = (re - q < 32) ? (int)(re - q) : 32;
                for (int j = 0; j < cnt; ++j) acc = add_rn<T>(acc, __shfl_sync(0xffffffffu, cur, j));
            }
            if (lane == src) r = acc;
        }
        T rinv = recip_rn<T>(r);                 // np.power(rowsum, -1)
        if (isinf(rinv)) rinv = 0;               // r_inv[np.isinf(r_inv)] = 0.
        if (i < n && rinv_out) rinv_out[i] = rinv;
        // scale: short rows by their lane, long rows by the warp
        if (e - b <= SHORT)
            for (int64_t p = b; p < e; ++p) {
                const T s = mul_rn<T>(rinv, val[p]);
                zero |= !(s != (T)0);
                out[p] = s;
            }
        todo = __ballot_sync(0xffffffffu, e - b > SHORT);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t rb = __shfl_sync(0xffffffffu, b, src), re = __shfl_sync(0xffffffffu, e, src);
            const T ri = __shfl_sync(0xffffffffu, rinv, src);
            for (int64_t q = rb + lane; q < re; q += 32) {
                const T s = mul_rn<T>(ri, val[q]);
                zero |= !(s != (T)0);
                out[q] = s;
            }
        }
    }
    // scipy's CSR product drops results that are exactly zero: the caller compacts if this fires
    if (__any_sync(0xffffffffu, zero) && lane == 0) atomicOr(reinterpret_cast<unsigned long long *>(zero_flag), 1ull);
}

// Y[i, :] = sum_j S_ij X[j, :] with X in CSR.  One warp per output row, the dense output row lives
// in shared memory (f floats per warp); the entries of S's row are walked in stored (ascending
// column) order and the lanes split the stored entries of X's row j, so every output element is
// accumulated by sequential FMA in ascending j -- the order of the dense kernel and of torch.spmm
// on the CPU.  X must be canonical (no duplicate column inside a row).
__global__ void
spmm_sparse_x_kernel(int64_t n_rows, const int64_t *__restrict__ s_ptr, const int32_t *__restrict__ s_col,
                     const float *__restrict__ s_val, const int64_t *__restrict__ x_ptr,
                     const int32_t *__restrict__ x_col, const float *__restrict__ x_val, int32_t f,
                     float *__restrict__ y, int64_t ldy)
{
    extern __shared__ float acc_all[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    float *acc = acc_all + (size_t)wib * f;
    for (int64_t i = (int64_t)blockIdx.x * wpb + wib; i < n_rows; i += (int64_t)gridDim.x * wpb) {
        for (int c = lane; c < f; c += 32) acc[c] = 0.f;
        __syncwarp();
        const int64_t b = s_ptr[i], e = s_ptr[i + 1];
        for (int64_t p = b; p < e; ++p) {
            const int32_t j = s_col[p];
            const float w = s_val[p];
            const int64_t xb = x_ptr[j], xe = x_ptr[j + 1];
            for (int64_t q = xb + lane; q < xe; q += 32) {
                const int32_t c = x_col[q];
                acc[c] = __fmaf_rn(w, x_val[q], acc[c]);
            }
            __syncwarp();
        }
        float *dst = y + i * ldy;
        for (int c = lane; c < f; c += 32) dst[c] = acc[c];
        __syncwarp();
    }
}

}  // namespace sgc

using namespace sgc;

extern "C" {

int32_t sgc_row_normalize(int64_t n, const int64_t *row_ptr_d, const void *val_d, int32_t is_f64,
                          void *val_out_d, void *rinv_out_d, int64_t *zero_flag_d, void *stream)
{
    SGC_REQUIRE(n >= 0 && row_ptr_d && zero_flag_d, "row_normalize: bad arguments");
    cudaStream_t s = as_stream(stream);
    SGC_CUDA_TRY(cudaMemsetAsync(zero_flag_d, 0, sizeof(int64_t), s));
    if (n == 0) return SGC_OK;
    const int grid = grid_for(ceil_div(n, 256), 8);
    if (is_f64)
        row_normalize_kernel<double><<<grid, 256, 0, s>>>(n, row_ptr_d, (const double *)val_d, (double *)val_out_d,
                                                          (double *)rinv_out_d, zero_flag_d);
    else
        row_normalize_kernel<float><<<grid, 256, 0, s>>>(n, row_ptr_d, (const float *)val_d, (float *)val_out_d,
                                                         (float *)rinv_out_d, zero_flag_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_spmm_hop_sparse_x(int64_t n_rows, const int64_t *s_row_ptr_d, const int32_t *s_col_d,
                              const float *s_val_d, const int64_t *x_row_ptr_d, const int32_t *x_col_d,
                              const float *x_val_d, int32_t f, float *y_d, int64_t ldy, void *stream)
{
    SGC_REQUIRE(n_rows >= 0 && f >= 1 && ldy >= f && s_row_ptr_d && x_row_ptr_d && y_d, "spmm_hop_sparse_x: bad arguments");
    if (n_rows == 0) return SGC_OK;
    const size_t row_bytes = (size_t)f * sizeof(float);
    const size_t budget = 200 * 1024;
    if (row_bytes > budget) {
        set_error("spmm_hop_sparse_x: %d feature columns exceed the shared-memory row (%zu KB); densify instead", f,
                  budget / 1024);
        return SGC_ERR_UNSUPPORTED;
    }
    int wpb = (int)(budget / row_bytes);
    if (wpb > 8) wpb = 8;
    // keep at least two CTAs per SM resident when the row is small enough
    while (wpb > 1 && (size_t)wpb * row_bytes > 100 * 1024) wpb /= 2;
    const size_t smem = (size_t)wpb * row_bytes;
    static thread_local size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        SGC_CUDA_TRY(cudaFuncSetAttribute(spmm_sparse_x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)budget));
        configured = budget;
    }
    const int per_sm = (int)((220 * 1024) / (smem > 1024 ? smem : 1024));
    const int grid = grid_for(ceil_div(n_rows, wpb), per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm));
    spmm_sparse_x_kernel<<<grid, wpb * 32, smem, as_stream(stream)>>>(n_rows, s_row_ptr_d, s_col_d, s_val_d,
                                                                     x_row_ptr_d, x_col_d, x_val_d, f, y_d, ldy);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

}  // extern "C"
