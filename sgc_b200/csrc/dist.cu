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

// ---- bulk (TMA, non-tensor) variant: contiguous bytes -> every destination -----------------------
// One elected thread per CTA streams tiles global -> shared (cp.async.bulk + mbarrier) and fans
// each tile out to all destinations with cp.async.bulk shared -> global; two stages so that the
// load of tile i+1 overlaps the stores of tile i.  The copies run on the TMA unit, not the LSU.
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(32)
push_bulk_kernel(const char *__restrict__ src, int64_t bytes, PushDst dst, int tile)
{
    extern __shared__ __align__(128) char stage[];          // 2 * tile
    __shared__ __align__(8) uint64_t bar[2];
    if (threadIdx.x != 0) return;
    const uint32_t b0 = smem_u32(&bar[0]), b1 = smem_u32(&bar[1]);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(b0));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(b1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const int64_t n_tiles = (bytes + tile - 1) / tile;
    int it = 0;
    for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        const int st = it & 1;
        const uint32_t phase = (uint32_t)(it >> 1) & 1u;
        const uint32_t bar_a = st ? b1 : b0;
        const uint32_t buf = smem_u32(stage + (size_t)st * tile);
        // the stores issued two iterations ago have finished reading this stage
        asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        const int64_t off = t * (int64_t)tile;
        const uint32_t size = (uint32_t)((bytes - off < tile) ? (bytes - off) : tile);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_a), "r"(size) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(buf), "l"(src + off), "r"(size), "r"(bar_a) : "memory");
        uint32_t done = 0;
        while (!done) {
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                         "selp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(bar_a), "r"(phase) : "memory");
        }
        for (int d = 0; d < dst.n; ++d)
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                         :: "l"(reinterpret_cast<char *>(dst.p[d]) + off), "r"(buf), "r"(size) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
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

int32_t sgc_push_bytes_bulk(const void *src_d, int64_t bytes, void *const *dst_d, int32_t n_dst,
                            int32_t tile_bytes, int32_t ctas_per_sm, void *stream)
{
    SGC_REQUIRE(bytes >= 0 && bytes % 16 == 0, "push_bytes_bulk: byte count must be a multiple of 16");
    SGC_REQUIRE(n_dst >= 0 && n_dst <= SGC_MAX_PEERS + 1, "push_bytes_bulk: n_dst out of range");
    SGC_REQUIRE(tile_bytes >= 16 && tile_bytes % 16 == 0 && tile_bytes <= 100 * 1024, "push_bytes_bulk: bad tile size");
    if (bytes == 0 || n_dst == 0) return SGC_OK;
    SGC_REQUIRE(src_d && dst_d && (reinterpret_cast<uintptr_t>(src_d) & 15) == 0, "push_bytes_bulk: bad source");
    PushDst d{};
    d.n = n_dst;
    for (int i = 0; i < n_dst; ++i) {
        SGC_REQUIRE(dst_d[i] && (reinterpret_cast<uintptr_t>(dst_d[i]) & 15) == 0, "push_bytes_bulk: bad destination %d", i);
        d.p[i] = reinterpret_cast<float *>(dst_d[i]);
    }
    if (ctas_per_sm < 1) ctas_per_sm = 4;
    const size_t smem = 2 * (size_t)tile_bytes;
    SGC_CUDA_TRY(cudaFuncSetAttribute(push_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = grid_for(ceil_div(bytes, tile_bytes), ctas_per_sm);
    push_bulk_kernel<<<grid, 32, smem, as_stream(stream)>>>(reinterpret_cast<const char *>(src_d), bytes, d, tile_bytes);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

}  // extern "C"
