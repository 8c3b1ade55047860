// Internal helpers shared by the kernels of libsgc_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#include "../../include/sgc_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libsgc_b200 is written for sm_100a (B200) only"
#endif

namespace sgc {

void set_error(const char *fmt, ...);
int32_t cuda_fail(cudaError_t e, const char *what, const char *file, int line);
void count_launch(int n = 1);
int sm_count();

#define SGC_CUDA_TRY(expr)                                                         \
    do {                                                                           \
        cudaError_t _e = (expr);                                                   \
        if (_e != cudaSuccess) return sgc::cuda_fail(_e, #expr, __FILE__, __LINE__); \
    } while (0)

#define SGC_REQUIRE(cond, ...)                 \
    do {                                       \
        if (!(cond)) {                         \
            sgc::set_error(__VA_ARGS__);       \
            return SGC_ERR_INVALID;            \
        }                                      \
    } while (0)

// Checks the launch itself (configuration errors); execution errors surface at the next sync.
#define SGC_LAUNCH_CHECK()                                                         \
    do {                                                                           \
        sgc::count_launch();                                                       \
        cudaError_t _e = cudaPeekAtLastError();                                    \
        if (_e != cudaSuccess) return sgc::cuda_fail(_e, "kernel launch", __FILE__, __LINE__); \
    } while (0)

static inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Grid size for a grid-stride kernel: enough blocks to cover `work_blocks`, capped at
// `per_sm` resident blocks on each of the SMs (148 on B200) so the grid is whole waves.
static inline int grid_for(int64_t work_blocks, int per_sm)
{
    int64_t cap = (int64_t)sm_count() * per_sm;
    if (work_blocks < 1) work_blocks = 1;
    return (int)(work_blocks < cap ? work_blocks : cap);
}

size_t scan_workspace_bytes(int64_t n);
int32_t scan_exclusive_i64(int64_t *data, int64_t n, void *ws, size_t ws_bytes, cudaStream_t s);

// ---- device helpers ------------------------------------------------------------------------
#ifdef __CUDACC__

// r^-1/2 in fp64, correctly rounded: 1/sqrt(r) then one Newton step evaluated with exact
// (FMA) residuals.  0 -> 0 (normalization.py:10 maps +inf to 0), negative/NaN -> NaN,
// +inf -> 0 (pow(inf, -0.5) == 0).
__device__ __forceinline__ double rsqrt_refined(double r)
{
    if (!(r > 0.0)) return (r == 0.0) ? 0.0 : __longlong_as_double(0x7ff8000000000000LL);
    if (isinf(r)) return 0.0;
    double y0 = __ddiv_rn(1.0, __dsqrt_rn(r));
    double t = __dmul_rn(y0, y0);
    double h = __fma_rn(y0, y0, -t);
    double e = __dadd_rn(__fma_rn(r, t, -1.0), __dmul_rn(r, h));
    return __fma_rn(__dmul_rn(-0.5, y0), e, y0);
}

// The single rounding of S to fp32 (utils.py:25): fp32((d_i * a_ij) * d_j), products in fp64.
__device__ __forceinline__ double norm_value64(double di, double a, double dj)
{
    return __dmul_rn(__dmul_rn(di, a), dj);
}

#endif  // __CUDACC__

}  // namespace sgc
