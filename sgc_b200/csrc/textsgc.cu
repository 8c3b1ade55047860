// Epilogue of the TextSGC precompute: /root/reference/downstream/TextSGC/utils.py:131-152
//     feats = torch.spmm(adj, dense_adj[:, idx]).t()          -> the hop kernel, then the transpose here
//     max / min over the training documents (dim 0 of feats = along each ROW of the hop output)
//     (feats - train_min) / train_range on the features with range > 0
// Hop output Y is [N, m] row-major (m = documents of the phase); the consumer wants [m, k] with
// k = number of kept features (rows of Y).
#include <cfloat>

#include "common.cuh"

namespace sgc {

// One warp per row of Y: min, max (torch.max/min over the training documents) and range = max - min
// in fp32, the reference's own expression (utils.py:137-139).  NaNs are not propagated (fminf/fmaxf);
// TextSGC inputs are finite.
__global__ void __launch_bounds__(256)
row_minmax_kernel(const float *__restrict__ y, int64_t n, int64_t m, int64_t ld, float *__restrict__ mn,
                  float *__restrict__ range)
{
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < n; r += warps) {
        const float *row = y + r * ld;
        float lo = FLT_MAX, hi = -FLT_MAX;
        for (int64_t c = lane; c < m; c += 32) {
            const float v = row[c];
            lo = fminf(lo, v);
            hi = fmaxf(hi, v);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if (lane == 0) {
            mn[r] = lo;
            range[r] = __fsub_rn(hi, lo);
        }
    }
}

// out[t, j] = (y[keep[j], t] - mn[keep[j]]) / range[keep[j]]   (utils.py:140-144,149-150): a 32 x 32
// tile transposed through shared memory so that both the reads of Y and the writes of out are
// coalesced; one IEEE subtraction and one IEEE division per element, as torch does.
__global__ void __launch_bounds__(256)
transpose_scale_kernel(const float *__restrict__ y, int64_t m, int64_t ld, const int64_t *__restrict__ keep,
                       int64_t k, const float *__restrict__ mn, const float *__restrict__ range,
                       float *__restrict__ out, int64_t ld_out)
{
    __shared__ float tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8
    const int64_t j0 = (int64_t)blockIdx.y * 32, t0 = (int64_t)blockIdx.x * 32;
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
        const int64_t j = j0 + i, t = t0 + tx;
        float v = 0.f;
        if (j < k && t < m) {
            const int64_t r = keep[j];
            v = __fdiv_rn(__fsub_rn(y[r * ld + t], mn[r]), range[r]);
        }
        tile[i][tx] = v;
    }
    __syncthreads();
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
        const int64_t t = t0 + i, j = j0 + tx;
        if (t < m && j < k) out[t * ld_out + j] = tile[tx][i];
    }
}

}  // namespace sgc

using namespace sgc;

extern "C" {

int32_t sgc_row_minmax(const float *y_d, int64_t n, int64_t m, int64_t ld, float *min_d, float *range_d,
                       void *stream)
{
    SGC_REQUIRE(n >= 0 && m >= 1 && ld >= m, "row_minmax: bad shape");
    if (n == 0) return SGC_OK;
    SGC_REQUIRE(y_d && min_d && range_d, "row_minmax: null arrays");
    row_minmax_kernel<<<grid_for(ceil_div(n, 8), 8), 256, 0, as_stream(stream)>>>(y_d, n, m, ld, min_d, range_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

int32_t sgc_transpose_scale(const float *y_d, int64_t m, int64_t ld, const int64_t *keep_d, int64_t k,
                            const float *min_d, const float *range_d, float *out_d, int64_t ld_out,
                            void *stream)
{
    SGC_REQUIRE(m >= 0 && k >= 0 && ld >= m && ld_out >= k, "transpose_scale: bad shape");
    if (m == 0 || k == 0) return SGC_OK;
    SGC_REQUIRE(y_d && keep_d && min_d && range_d && out_d, "transpose_scale: null arrays");
    const dim3 grid((unsigned)ceil_div(m, 32), (unsigned)ceil_div(k, 32));
    SGC_REQUIRE(grid.y <= 65535u, "transpose_scale: more than 2M kept features");
    transpose_scale_kernel<<<grid, 256, 0, as_stream(stream)>>>(y_d, m, ld, keep_d, k, min_d, range_d, out_d,
                                                               ld_out);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

}  // extern "C"
