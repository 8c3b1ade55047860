// K5 support: symmetric buffers, CUDA IPC mapping of the peers' buffers, and the kernel that
// distributes a rank's rows of the initial features to every peer.
//
// The exchange of the row-sharded propagation is not a separate collective: spmm_hop_kernel
// (spmm.cuh) stores every output row to the local buffer and to the NVLink-mapped buffers of all
// peers from the same registers.  What is left for this file is the plumbing that makes a peer's
// buffer addressable (cudaIpc*), and push_rows_kernel for the initial distribution of X.
#include <cstring>

#include "common.cuh"

namespace sgc {

struct PushDst { int32_t n; float *p[SGC_MAX_PEERS + 1]; };

// One 16-byte vector per thread per destination; rows are [rows, f] pieces of wider matrices.
// Source reads are streamed once (evict-first); destination stores go to peer memory.
template <int VEC>
__global__ void __launch_bounds__(256)
push_rows_kernel(const float *__restrict__ src, int64_t ld_src, int64_t rows, int32_t f,
                 PushDst dst, int64_t ld_dst)
{
    const int64_t per_row = (f + VEC - 1) / VEC;
    const int64_t total = rows * per_row;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / per_row;
        const int c = (int)(i - r * per_row) * VEC;
        const float *s = src + r * ld_src + c;
        float v[VEC];
        if (VEC == 4) {
            const float4 t = __ldcs(reinterpret_cast<const float4 *>(s));
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
#pragma unroll
            for (int k = 0; k < VEC; ++k) v[k] = (c + k < f) ? __ldcs(s + k) : 0.f;
        }
        for (int d = 0; d < dst.n; ++d) {
            float *q = dst.p[d] + r * ld_dst + c;
            if (VEC == 4) {
                *reinterpret_cast<float4 *>(q) = make_float4(v[0], v[1], v[2], v[3]);
            } else {
#pragma unroll
                for (int k = 0; k < VEC; ++k)
                    if (c + k < f) q[k] = v[k];
            }
        }
    }
}

}  // namespace sgc

using namespace sgc;

extern "C" {

int32_t sgc_symm_alloc(size_t bytes, void **ptr_out)
{
    SGC_REQUIRE(ptr_out != nullptr, "symm_alloc: ptr_out is null");
    *ptr_out = nullptr;
    if (bytes == 0) bytes = 256;
    void *p = nullptr;
    SGC_CUDA_TRY(cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemset(p, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        cudaFree(p);
        return cuda_fail(e, "symm_alloc memset", __FILE__, __LINE__);
    }
    *ptr_out = p;
    return SGC_OK;
}

int32_t sgc_symm_free(void *ptr_d)
{
    if (ptr_d) SGC_CUDA_TRY(cudaFree(ptr_d));
    return SGC_OK;
}

int32_t sgc_ipc_export(void *ptr_d, void *handle_out)
{
    SGC_REQUIRE(ptr_d && handle_out, "ipc_export: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == SGC_IPC_HANDLE_BYTES, "unexpected IPC handle size");
    cudaIpcMemHandle_t h;
    SGC_CUDA_TRY(cudaIpcGetMemHandle(&h, ptr_d));
    memcpy(handle_out, &h, sizeof(h));
    return SGC_OK;
}

int32_t sgc_ipc_open(const void *handle, void **ptr_out)
{
    SGC_REQUIRE(handle && ptr_out, "ipc_open: null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void *p = nullptr;
    SGC_CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *ptr_out = p;
    return SGC_OK;
}

int32_t sgc_ipc_close(void *ptr_d)
{
    if (ptr_d) SGC_CUDA_TRY(cudaIpcCloseMemHandle(ptr_d));
    return SGC_OK;
}

int32_t sgc_push_rows(const float *src_d, int64_t ld_src, int64_t rows, int32_t f,
                      float *const *dst_d, int32_t n_dst, int64_t ld_dst, void *stream)
{
    SGC_REQUIRE(rows >= 0 && f >= 1 && ld_src >= f && ld_dst >= f, "push_rows: bad shape");
    SGC_REQUIRE(n_dst >= 0 && n_dst <= SGC_MAX_PEERS + 1, "push_rows: n_dst out of range");
    if (rows == 0 || n_dst == 0) return SGC_OK;
    SGC_REQUIRE(src_d && dst_d, "push_rows: null pointer");
    PushDst d{};
    d.n = n_dst;
    bool vec4 = (ld_src % 4 == 0) && (ld_dst % 4 == 0) && ((reinterpret_cast<uintptr_t>(src_d) & 15) == 0);
    for (int i = 0; i < n_dst; ++i) {
        SGC_REQUIRE(dst_d[i] != nullptr, "push_rows: dst_d[%d] is null", i);
        d.p[i] = dst_d[i];
        vec4 = vec4 && ((reinterpret_cast<uintptr_t>(dst_d[i]) & 15) == 0);
    }
    // the vector path writes whole 4-float groups: the row tail must lie inside both strides
    if (vec4 && ((f + 3) / 4 * 4 > ld_src || (f + 3) / 4 * 4 > ld_dst)) vec4 = false;
    cudaStream_t s = as_stream(stream);
    if (vec4) {
        const int64_t work = rows * ((f + 3) / 4);
        push_rows_kernel<4><<<grid_for(ceil_div(work, 256), 8), 256, 0, s>>>(src_d, ld_src, rows, f, d, ld_dst);
    } else {
        const int64_t work = rows * (int64_t)f;
        push_rows_kernel<1><<<grid_for(ceil_div(work, 256), 8), 256, 0, s>>>(src_d, ld_src, rows, f, d, ld_dst);
    }
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

}  // extern "C"
