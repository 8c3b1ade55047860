// K3: column mean and unbiased standard deviation of a row-major fp32 matrix.
//
// Restates the statistics of /root/reference/utils.py:119
//     features = (features - features.mean(dim=0)) / features.std(dim=0)
// (torch's std is the unbiased one, divisor n-1).  One pass over X: every thread owns one column
// of a 256-column tile and walks a contiguous block of rows, so a warp reads 128 contiguous bytes
// per row; sums are kept in fp64 and shifted by the column's first value so that
// sum((x-k)^2) - sum(x-k)^2/n does not cancel when |mean| >> std.  Block partials are reduced in
// block order by a second kernel (deterministic; no atomics).  HBM-bound: reads 4*n*f bytes once.
#include "common.cuh"

#include <mutex>

namespace sgc {

constexpr int CS_THREADS = 256;

__global__ void __launch_bounds__(CS_THREADS)
col_stats_partial_kernel(const float *__restrict__ x, int64_t n, int64_t f, int64_t ld,
                         int64_t rows_per_block, double *__restrict__ part)
{
    const int64_t c = (int64_t)blockIdx.y * CS_THREADS + threadIdx.x;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = (r0 + rows_per_block < n) ? r0 + rows_per_block : n;
    double s1 = 0.0, s2 = 0.0;
    if (c < f) {
        const double k = (double)x[c];             // shift: first row of the column
        const float *px = x + c;
        int64_t r = r0;
        for (; r + 4 <= r1; r += 4) {
            const float a0 = __ldcs(px + (r + 0) * ld), a1 = __ldcs(px + (r + 1) * ld);
            const float a2 = __ldcs(px + (r + 2) * ld), a3 = __ldcs(px + (r + 3) * ld);
            const double d0 = (double)a0 - k, d1 = (double)a1 - k, d2 = (double)a2 - k, d3 = (double)a3 - k;
            s1 += d0; s2 = fma(d0, d0, s2);
            s1 += d1; s2 = fma(d1, d1, s2);
            s1 += d2; s2 = fma(d2, d2, s2);
            s1 += d3; s2 = fma(d3, d3, s2);
        }
        for (; r < r1; ++r) {
            const double d = (double)__ldcs(px + r * ld) - k;
            s1 += d; s2 = fma(d, d, s2);
        }
        part[((int64_t)blockIdx.x * 2 + 0) * f + c] = s1;
        part[((int64_t)blockIdx.x * 2 + 1) * f + c] = s2;
    }
}

__global__ void __launch_bounds__(CS_THREADS)
col_stats_final_kernel(const float *__restrict__ x, int64_t n, int64_t f, int n_parts,
                       const double *__restrict__ part, float *__restrict__ mean32,
                       float *__restrict__ std32, double *__restrict__ mean64,
                       double *__restrict__ std64)
{
    const int64_t c = (int64_t)blockIdx.x * CS_THREADS + threadIdx.x;
    if (c >= f) return;
    double s1 = 0.0, s2 = 0.0;
    for (int b = 0; b < n_parts; ++b) {
        s1 += part[((int64_t)b * 2 + 0) * f + c];
        s2 += part[((int64_t)b * 2 + 1) * f + c];
    }
    const double k = (double)x[c];
    const double dn = (double)n;
    const double mean = k + s1 / dn;
    double var = (s2 - s1 * s1 / dn) / (dn - 1.0);      // n == 1 -> 0/0 = NaN, like torch
    if (var < 0.0) var = 0.0;
    const double sd = sqrt(var);
    if (mean32) mean32[c] = (float)mean;
    if (std32) std32[c] = (float)sd;
    if (mean64) mean64[c] = mean;
    if (std64) std64[c] = sd;
}

// x <- (x - mean[c]) / std[c] in place: the reference's own fp32 expression (utils.py:119), one
// subtraction and one IEEE division per element, so the result is bit-identical to evaluating
// it with torch.  Streaming pass: 8 bytes of HBM traffic per element; float4 path when rows are
// 16-byte aligned.
__global__ void __launch_bounds__(256)
standardize_inplace_kernel(float *__restrict__ x, int64_t n, int64_t f, int64_t ld,
                           const float *__restrict__ mean, const float *__restrict__ stdv, bool vec4)
{
    if (vec4) {
        const int64_t f4 = (f + 3) / 4;                 // vectors per row (row tail is inside ld)
        const int64_t total = n * f4;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
             i += (int64_t)gridDim.x * blockDim.x) {
            const int64_t r = i / f4, c = (i - r * f4) * 4;
            float4 *p = reinterpret_cast<float4 *>(x + r * ld + c);
            float4 v = *p;
            float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (c + k < f) e[k] = __fdiv_rn(__fsub_rn(e[k], __ldg(mean + c + k)), __ldg(stdv + c + k));
            *p = make_float4(e[0], e[1], e[2], e[3]);
        }
    } else {
        const int64_t total = n * f;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
             i += (int64_t)gridDim.x * blockDim.x) {
            const int64_t r = i / f, c = i - r * f;
            float *p = x + r * ld + c;
            *p = __fdiv_rn(__fsub_rn(*p, __ldg(mean + c)), __ldg(stdv + c));
        }
    }
}

static int stats_blocks(int64_t n)
{
    int64_t b = ceil_div(n, 64);                 // at least 64 rows per block
    const int64_t cap = (int64_t)sm_count() * 4;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace sgc

using namespace sgc;

// ---- the same statistics for callers that must not allocate ---------------------------------
//
// sgc_col_stats needs a few MB of device scratch for its per-block partial sums, and the caller
// has to get that from somewhere.  Some callers cannot: a plugin called back from inside another
// framework's op or stream callback where cudaMalloc is forbidden, a process whose whole device
// budget was handed to a pool it does not own, or a real-time loop where cudaMalloc's implicit
// device synchronisation is not acceptable.  For them the partial sums live in the array below,
// which is part of the library image: every GPU on which the library's module is loaded carries
// these 8 MiB from then until the process exits, used or not.
//
// Host code must not treat `cs_static_scratch` as a pointer: on the host that name is only a
// handle for the symbol.  static_scratch_ptr() asks the runtime for the real device address,
// once per device and with that device current, and the kernels are handed that address.
constexpr size_t CS_STATIC_DOUBLES = SGC_COL_STATS_STATIC_MAX_BYTES / sizeof(double);
static __device__ __align__(256) double cs_static_scratch[CS_STATIC_DOUBLES];

constexpr int CS_MAX_DEVICES = 64;
static std::mutex cs_static_mutex;
static double *cs_static_addr[CS_MAX_DEVICES];     // device address of the array, per device ordinal

// Device address of cs_static_scratch on the current device (cached after the first call there).
static int32_t static_scratch_ptr(double **out)
{
    int dev = -1;
    SGC_CUDA_TRY(cudaGetDevice(&dev));
    SGC_REQUIRE(dev >= 0 && dev < CS_MAX_DEVICES, "col_stats_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(cs_static_mutex);
    if (!cs_static_addr[dev]) {
        double *p = nullptr;
        SGC_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), cs_static_scratch));
        SGC_REQUIRE(p != nullptr, "col_stats_static: cudaGetSymbolAddress returned a null address");
        cs_static_addr[dev] = p;
    }
    *out = cs_static_addr[dev];
    return SGC_OK;
}

extern "C" {

size_t sgc_col_stats_workspace_bytes(int64_t n, int64_t f)
{
    return (size_t)stats_blocks(n) * 2 * (size_t)(f > 0 ? f : 1) * sizeof(double) + 256;
}

int32_t sgc_col_stats(const float *x_d, int64_t n, int64_t f, int64_t ld, float *mean32_d,
                      float *std32_d, double *mean64_d, double *std64_d, void *ws_d,
                      size_t ws_bytes, void *stream)
{
    SGC_REQUIRE(x_d && n >= 1 && f >= 1 && ld >= f, "col_stats: bad arguments (n=%lld f=%lld ld=%lld)",
                (long long)n, (long long)f, (long long)ld);
    if (!ws_d || ws_bytes < sgc_col_stats_workspace_bytes(n, f)) {
        set_error("col_stats: workspace too small");
        return SGC_ERR_WORKSPACE;
    }
    cudaStream_t s = as_stream(stream);
    const int nb = stats_blocks(n);
    const int64_t rows_per_block = ceil_div(n, nb);
    const unsigned tiles = (unsigned)ceil_div(f, CS_THREADS);
    double *part = reinterpret_cast<double *>(ws_d);
    col_stats_partial_kernel<<<dim3((unsigned)nb, tiles), CS_THREADS, 0, s>>>(x_d, n, f, ld, rows_per_block, part);
    SGC_LAUNCH_CHECK();
    col_stats_final_kernel<<<tiles, CS_THREADS, 0, s>>>(x_d, n, f, nb, part, mean32_d, std32_d, mean64_d, std64_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

// Column statistics with the partial sums in the library's own static array; see the block
// comment above cs_static_scratch and the declaration in include/sgc_b200.h.
int32_t sgc_col_stats_static(const float *x_d, int64_t n, int64_t f, int64_t ld, float *mean32_d,
                             float *std32_d, double *mean64_d, double *std64_d, void *stream)
{
    SGC_REQUIRE(x_d && n >= 1 && f >= 1 && ld >= f, "col_stats_static: bad arguments (n=%lld f=%lld ld=%lld)",
                (long long)n, (long long)f, (long long)ld);
    // same block count as sgc_col_stats, so the partial sums are added in the same order and the
    // two entry points agree bit for bit
    const int nb = stats_blocks(n);
    const size_t need = (size_t)nb * 2 * (size_t)f * sizeof(double);
    if (need > SGC_COL_STATS_STATIC_MAX_BYTES) {
        set_error("col_stats_static: n=%lld f=%lld needs %zu bytes of partial sums, the built-in scratch holds "
                  "%zu; use sgc_col_stats with a caller-provided workspace",
                  (long long)n, (long long)f, need, (size_t)SGC_COL_STATS_STATIC_MAX_BYTES);
        return SGC_ERR_WORKSPACE;
    }
    double *part = nullptr;
    const int32_t rc = static_scratch_ptr(&part);
    if (rc != SGC_OK) return rc;
    cudaStream_t s = as_stream(stream);
    const int64_t rows_per_block = ceil_div(n, nb);
    const unsigned tiles = (unsigned)ceil_div(f, CS_THREADS);
    // Nothing in the array has to start at zero: the first kernel writes every partial sum the
    // second one reads (nb blocks x 2 x f columns), so whatever the previous call left is dead.
    col_stats_partial_kernel<<<dim3((unsigned)nb, tiles), CS_THREADS, 0, s>>>(x_d, n, f, ld, rows_per_block, part);
    SGC_LAUNCH_CHECK();
    col_stats_final_kernel<<<tiles, CS_THREADS, 0, s>>>(x_d, n, f, nb, part, mean32_d, std32_d, mean64_d, std64_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_standardize_inplace(float *x_d, int64_t n, int64_t f, int64_t ld, const float *mean_d,
                                const float *std_d, void *stream)
{
    SGC_REQUIRE(x_d && mean_d && std_d && n >= 0 && f >= 1 && ld >= f, "standardize_inplace: bad arguments");
    if (n == 0) return SGC_OK;
    const bool vec4 = (ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(x_d) & 15) == 0);
    const int64_t work = vec4 ? n * ((f + 3) / 4) : n * f;
    standardize_inplace_kernel<<<grid_for(ceil_div(work, 256), 8), 256, 0, as_stream(stream)>>>(
        x_d, n, f, ld, mean_d, std_d, vec4);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

}  // extern "C"
