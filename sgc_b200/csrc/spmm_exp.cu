// EXPERIMENT (kept out of the product path; tools/sweep_tma_gather.py drives it): the hop with the gathers of
// X staged through shared memory by the TMA unit instead of per-lane LDG.128 -- SURVEY section 7 step 5 /
// north_star (2) "col_idx and values staged through shared memory or TMA".
//
// One warp = one task (row or chunk item x 128-float slab), ticket-scheduled like the product kernel.  For a
// batch of up to BATCH neighbours one lane hands the TMA unit one cp.async.bulk global -> shared per neighbour
// (the 512-byte slab piece of that row of X) completing on an mbarrier; DEPTH batches are in flight per warp,
// so a warp keeps DEPTH * BATCH * 512 bytes outstanding without holding them in registers (the LDG kernel
// holds 8 x 16 bytes per lane = 4 KB per warp).  The lanes then read their float4 of every staged piece from
// shared memory and accumulate in the same ascending order as the product kernel, so results are bit-identical
// and the experiment is checked, not just timed.
//
// Second form (G4 = true, sgc_exp_spmm_hop_gather4): the tensor instruction cp.async.bulk.tensor.2d ...
// tile::gather4 -- ONE instruction names four row indices of X (a 2-D CUtensorMap over [rows, ldx], box 128 x 1)
// and the TMA unit lands the four 512-byte slab pieces in consecutive shared-memory rows: a batch of 8
// neighbours is 2 instructions instead of 8 bulk copies.  Columns past ldx (the last slab of a 602-column
// matrix) are zero-filled by the unit.
#include <cuda.h>
#include <cstring>
#include "spmm.cuh"

namespace sgc {

constexpr int EXP_THREADS = 128;
constexpr int EXP_BATCH = 8;        // neighbours per mbarrier phase
constexpr int EXP_SW = 128;         // slab width in floats (512-byte pieces)

__device__ __forceinline__ uint32_t exp_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int DEPTH, bool G4>
__global__ void __launch_bounds__(EXP_THREADS)
spmm_hop_bulk_gather_kernel(const HopDev p, const int64_t n_tasks, const __grid_constant__ CUtensorMap tmap)
{
    constexpr int WARPS = EXP_THREADS / 32;
    extern __shared__ __align__(128) float exp_smem[];
    typedef float StageT[DEPTH][EXP_BATCH][EXP_SW];
    StageT *stage = reinterpret_cast<StageT *>(exp_smem);                              // [WARPS]
    __shared__ __align__(8) uint64_t bars[WARPS][DEPTH];
    __shared__ float wsm[WARPS][DEPTH][EXP_BATCH];
    const int lane = threadIdx.x & 31;
    const int wic = threadIdx.x >> 5;
    const uint64_t pol_s = l2_policy((p.hints & 2) ? 2 : 0);
    if (lane == 0) {
        for (int d = 0; d < DEPTH; ++d)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(exp_smem_u32(&bars[wic][d])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    uint32_t phase_bits = 0;                 // bit d = parity the next wait on barrier d expects
    const int64_t per_slab = p.n_long_items + p.n_rows;

#pragma unroll 1
    for (;;) {
        const int64_t task = next_ticket(p.tile_counter, lane);
        if (task >= n_tasks) break;
        const int64_t slab = task / per_slab;
        const int64_t t = task - slab * per_slab;
        int64_t row, beg, end;
        float *dst;
        bool valid = true;
        if (t < p.n_long_items) {
            row = p.li_row[t]; beg = p.li_beg[t]; end = beg + p.li_len[t];
            dst = p.partial + t * p.ldp;
        } else {
            row = t - p.n_long_items; beg = p.row_ptr[row]; end = p.row_ptr[row + 1];
            if (p.chunk > 0 && end - beg > p.chunk) valid = false;
            dst = p.y + row * p.ldy;
        }
        if (!valid) continue;
        const int f0 = (int)slab * EXP_SW;
        int nfl = p.f - f0;
        nfl = nfl > EXP_SW ? EXP_SW : nfl;
        const uint32_t bytes = (uint32_t)((nfl + 3) & ~3) * 4u;
        const int fo = f0 + lane * 4;
        const bool act = fo < p.f;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        const int len = (int)(end - beg);
        const int n_batches = (len + EXP_BATCH - 1) / EXP_BATCH;

        // (col, weight) of 32 entries per coalesced load, broadcast by shuffles as in the product kernel
        int32_t c_reg = 0;
        float w_reg = 0.f;
        int loaded_base = -32;
        auto issue = [&](int b) {
            const int e0 = b * EXP_BATCH;                        // first entry of the batch, relative to beg
            if (e0 >= loaded_base + 32) {                        // next 32 (col, weight) pairs
                loaded_base = e0;
                const int64_t idx = beg + e0 + lane;
                c_reg = 0; w_reg = 0.f;
                if (idx < end) {
                    c_reg = ld_stream_i32(p.col + idx, pol_s);
                    w_reg = ld_stream_f32(p.sval + idx, pol_s);
                }
            }
            const int cnt = (len - e0) < EXP_BATCH ? (len - e0) : EXP_BATCH;
            const int d = b % DEPTH;
            const uint32_t bar = exp_smem_u32(&bars[wic][d]);
            // every lane learns the batch's columns (the shuffles are warp-wide); lane 0 issues the copies
            int32_t cj[EXP_BATCH];
#pragma unroll
            for (int j = 0; j < EXP_BATCH; ++j) cj[j] = __shfl_sync(0xffffffffu, c_reg, (e0 - loaded_base + j) & 31);
            if (G4) {
                if (lane == 0) {
                    // whole boxes always: 4 rows x 128 floats per instruction (rows past the batch repeat its
                    // first neighbour and are never read back; columns past ldx arrive as zeros)
                    const int n4 = (cnt + 3) >> 2;
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                                 :: "r"(bar), "r"((uint32_t)n4 * 4u * EXP_SW * 4u) : "memory");
#pragma unroll
                    for (int q = 0; q < EXP_BATCH / 4; ++q) {
                        if (q < n4) {
                            const int32_t r0 = cj[4 * q] & 0x7fffffff;
                            const int32_t r1 = (4 * q + 1 < cnt) ? (cj[4 * q + 1] & 0x7fffffff) : r0;
                            const int32_t r2 = (4 * q + 2 < cnt) ? (cj[4 * q + 2] & 0x7fffffff) : r0;
                            const int32_t r3 = (4 * q + 3 < cnt) ? (cj[4 * q + 3] & 0x7fffffff) : r0;
                            asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes "
                                         "[%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                                         :: "r"(exp_smem_u32(&stage[wic][d][4 * q][0])), "l"(&tmap), "r"(f0), "r"(r0), "r"(r1),
                                            "r"(r2), "r"(r3), "r"(bar)
                                         : "memory");
                        }
                    }
                }
            } else if (lane == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                             :: "r"(bar), "r"(bytes * (uint32_t)cnt) : "memory");
#pragma unroll
                for (int j = 0; j < EXP_BATCH; ++j)
                    if (j < cnt)
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     :: "r"(exp_smem_u32(&stage[wic][d][j][0])),
                                        "l"(p.x + (int64_t)(cj[j] & 0x7fffffff) * p.ldx + f0), "r"(bytes), "r"(bar)
                                     : "memory");
            }
            // the weights of a batch must outlive the (col, weight) reload of a later batch: park them in smem
            const float wv = __shfl_sync(0xffffffffu, w_reg, (e0 - loaded_base + (lane & (EXP_BATCH - 1))) & 31);
            if (lane < EXP_BATCH) wsm[wic][d][lane] = wv;
            __syncwarp();
        };

        const int pre = n_batches < DEPTH ? n_batches : DEPTH;
        for (int b = 0; b < pre; ++b) issue(b);
#pragma unroll 1
        for (int b = 0; b < n_batches; ++b) {
            const int d = b % DEPTH;
            const uint32_t bar = exp_smem_u32(&bars[wic][d]);
            const uint32_t parity = (phase_bits >> d) & 1u;
            uint32_t done = 0;
            while (!done)
                asm volatile("{\n.reg .pred q;\nmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n"
                             "selp.u32 %0, 1, 0, q;\n}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
            phase_bits ^= (1u << d);
            const int e0 = b * EXP_BATCH;
            const int cnt = (len - e0) < EXP_BATCH ? (len - e0) : EXP_BATCH;
            float4 v[EXP_BATCH];
#pragma unroll
            for (int j = 0; j < EXP_BATCH; ++j)
                v[j] = (j < cnt && act) ? *reinterpret_cast<const float4 *>(&stage[wic][d][j][lane * 4])
                                        : make_float4(0.f, 0.f, 0.f, 0.f);
            float wj[EXP_BATCH];
#pragma unroll
            for (int j = 0; j < EXP_BATCH; ++j) wj[j] = wsm[wic][d][j];
            __syncwarp();                                   // every lane has read the stage: it may be refilled
            if (b + DEPTH < n_batches) issue(b + DEPTH);
#pragma unroll
            for (int j = 0; j < EXP_BATCH; ++j) {
                if (j < cnt) {
                    acc[0] = __fmaf_rn(wj[j], v[j].x, acc[0]);
                    acc[1] = __fmaf_rn(wj[j], v[j].y, acc[1]);
                    acc[2] = __fmaf_rn(wj[j], v[j].z, acc[2]);
                    acc[3] = __fmaf_rn(wj[j], v[j].w, acc[3]);
                }
            }
        }
        if (act) {
            if (fo + 4 <= p.f_store) {
                st_hint_v4(dst + fo, acc, pol_s);
            } else {
                for (int v = 0; v < 4; ++v)
                    if (fo + v < p.f) st_hint_v1(dst + fo + v, acc[v], pol_s);
            }
        }
    }
    retire_warp(p.tile_counter, lane, (unsigned long long)gridDim.x * WARPS);
}

}  // namespace sgc

using namespace sgc;

// cuTensorMapEncodeTiled through the runtime's driver entry point (the library does not link libcuda)
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int32_t make_gather_map(CUtensorMap *map, const float *x, int64_t ldx, int64_t rows)
{
    static encode_tiled_fn encode = nullptr;
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        SGC_CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
        SGC_REQUIRE(fn && q == cudaDriverEntryPointSuccess, "exp_hop: cuTensorMapEncodeTiled is not available");
        encode = reinterpret_cast<encode_tiled_fn>(fn);
    }
    const cuuint64_t dims[2] = {(cuuint64_t)ldx, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ldx * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)EXP_SW, 1u};
    const cuuint32_t estr[2] = {1u, 1u};
    const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(x), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("exp_hop: cuTensorMapEncodeTiled failed (%d)", (int)r);
        return SGC_ERR_CUDA;
    }
    return SGC_OK;
}

static int32_t exp_hop(const sgc_plan_t *plan, const sgc_hop_args_t *a, int32_t depth, int32_t ctas_per_sm, bool g4,
                       void *stream);

extern "C" {

int32_t sgc_exp_spmm_hop_bulk_gather(const sgc_plan_t *plan, const sgc_hop_args_t *a, int32_t depth,
                                     int32_t ctas_per_sm, void *stream)
{
    return exp_hop(plan, a, depth, ctas_per_sm, false, stream);
}

int32_t sgc_exp_spmm_hop_gather4(const sgc_plan_t *plan, const sgc_hop_args_t *a, int32_t depth,
                                 int32_t ctas_per_sm, void *stream)
{
    return exp_hop(plan, a, depth, ctas_per_sm, true, stream);
}

}  // extern "C"

static int32_t exp_hop(const sgc_plan_t *plan, const sgc_hop_args_t *a, int32_t depth, int32_t ctas_per_sm, bool g4,
                       void *stream)
{
    SGC_REQUIRE(plan && a, "exp_hop: null plan or args");
    const sgc_plan_impl *pl = &plan->impl;
    SGC_REQUIRE(a->x_d && a->y_d && a->sval_d && a->f >= 1, "exp_hop: needs x_d, y_d, sval_d");
    SGC_REQUIRE(a->ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(a->x_d) & 15) == 0 && a->ldx >= (a->f + 3) / 4 * 4,
                "exp_hop: x must be 16-byte aligned with ldx %% 4 == 0 and padded rows");
    SGC_REQUIRE(a->ldy % 4 == 0 && (reinterpret_cast<uintptr_t>(a->y_d) & 15) == 0, "exp_hop: y must be 16-byte aligned");
    SGC_REQUIRE(pl->tile_counter != nullptr, "exp_hop: the plan has no workspace");
    SGC_REQUIRE(a->flags == SGC_HOP_PAD_OWNED || a->flags == 0, "exp_hop: plain hops only");
    if (pl->n_long_items > 0) SGC_REQUIRE(a->partial_d && a->ldp % 4 == 0, "exp_hop: partial_d needed");
    if (pl->n_rows == 0) return SGC_OK;
    HopDev p{};
    p.row_ptr = pl->row_ptr; p.col = pl->col; p.sval = a->sval_d;
    p.x = a->x_d; p.ldx = a->ldx; p.y = a->y_d; p.ldy = a->ldy;
    p.partial = a->partial_d; p.ldp = a->ldp;
    p.li_row = pl->li_row; p.li_beg = pl->li_beg; p.li_len = pl->li_len;
    p.n_rows = pl->n_rows; p.n_long_items = pl->n_long_items;
    p.f = a->f; p.chunk = pl->chunk;
    p.n_slabs = (a->f + EXP_SW - 1) / EXP_SW;
    p.y_vec = 4;
    const bool pad_owned = (a->flags & SGC_HOP_PAD_OWNED) != 0 && (a->f + 3) / 4 * 4 <= a->ldy;
    p.f_store = pad_owned ? (a->f + 3) / 4 * 4 : a->f;
    p.hints = a->hints & 3;
    p.tile_counter = pl->tile_counter;
    const int64_t n_tasks = (int64_t)p.n_slabs * (p.n_long_items + p.n_rows);
    cudaStream_t s = as_stream(stream);
    if (ctas_per_sm < 1) ctas_per_sm = 4;
    const int64_t want = ceil_div(n_tasks, EXP_THREADS / 32);
    const int64_t cap = (int64_t)sm_count() * ctas_per_sm;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    if (depth < 1) depth = 1;
    if (depth > 3) depth = 3;
    const size_t smem = (size_t)(EXP_THREADS / 32) * depth * EXP_BATCH * EXP_SW * sizeof(float);
    CUtensorMap tmap;
    std::memset(&tmap, 0, sizeof(tmap));
    if (g4) {
        SGC_REQUIRE(a->ldx % 4 == 0, "exp_hop: gather4 needs a row pitch that is a multiple of 16 bytes");
        int32_t rc = make_gather_map(&tmap, a->x_d, a->ldx, pl->n_cols);
        if (rc != SGC_OK) return rc;
    }
#define SGC_EXP_LAUNCH(D, G)                                                                                     \
    do {                                                                                                         \
        SGC_CUDA_TRY(cudaFuncSetAttribute(spmm_hop_bulk_gather_kernel<D, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                          (int)smem));                                                           \
        spmm_hop_bulk_gather_kernel<D, G><<<grid, EXP_THREADS, smem, s>>>(p, n_tasks, tmap);                     \
    } while (0)
    if (g4) {
        if (depth == 1) SGC_EXP_LAUNCH(1, true);
        else if (depth == 2) SGC_EXP_LAUNCH(2, true);
        else SGC_EXP_LAUNCH(3, true);
    } else {
        if (depth == 1) SGC_EXP_LAUNCH(1, false);
        else if (depth == 2) SGC_EXP_LAUNCH(2, false);
        else SGC_EXP_LAUNCH(3, false);
    }
#undef SGC_EXP_LAUNCH
    SGC_LAUNCH_CHECK();
    // Rows split into chunk items are left as partial sums (the product's fix-up kernel is not run): the
    // experiment times the gather kernel alone; tools/sweep_tma_gather.py checks results on a chunk = 0 plan.
    return SGC_OK;
}
