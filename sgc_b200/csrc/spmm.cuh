// K4: the propagation hop  Y = S X  (what torch.spmm does at /root/reference/utils.py:95) as a
// load-balanced CSR SpMM for sm_100a.
//
// Work decomposition
//   task  = (feature slab, row-or-chunk).  Tasks are numbered slab-major, so the grid sweeps all
//           rows for feature columns [0, SW), then [SW, 2SW), ...: the gathered operand touched
//           at any time is N x SW floats, which the caller sizes to stay resident in the 126 MB
//           L2 (two dies: ~half of it is usable for data every SM reads) while col/val stream
//           through with an evict-first policy.
//   group = G lanes own one task; a warp runs 32/G tasks.  Each lane keeps U vectors of VEC
//           floats of the output row in registers (slab width SW = G*VEC*U).
//   rows longer than `chunk` entries are split at plan time into chunk items whose partial sums
//           go to a scratch matrix and are combined in chunk order by spmm_fixup_kernel
//           (deterministic; no float atomics).  Unsplit rows are accumulated by sequential
//           ascending-column fp32 FMA, the order torch.spmm uses on the CPU.
//
// Inner loop: the G lanes of a group load G (col, value) pairs with one coalesced access, turn
// them into fp32 weights (fused mode: fp32((d_i*a_ij)*d_j) from the fp64 A+I value and D^-1/2,
// so S is never materialised), then broadcast them lane by lane with shuffles while every lane
// issues UNR independent 16-byte gathers of its piece of the neighbour rows before the FMAs that
// consume them.  Optional first-hop standardisation (x-mean)/std is applied to the gathered
// value in registers.
#pragma once
#include "common.cuh"
#include <atomic>

namespace sgc {

constexpr int HOP_THREADS = 128;
#ifndef SGC_HOP_MIN_CTAS
#define SGC_HOP_MIN_CTAS 8      // resident CTAs per SM the plain single-GPU hop kernels are compiled for
#endif
// Register cap by variant.  The plain variants (one vector per lane, no fused standardisation) fit 56 registers
// without spilling and run with 8-9 CTAs per SM instead of 7: Reddit shape K = 2 8.03 -> 7.68 ms, products hop
// 7.09 -> 6.96 ms, bit-identical (tools/sweep_occupancy.py, profiles/r02_sweep_occupancy.log; 10 CTAs / 48
// registers: no further gain).  The variants with two vectors per lane or the standardisation in the gather
// would spill under that cap and keep the looser one.
template <int U, int STD> constexpr int hop_min_ctas() { return (U == 1 && STD == 0) ? SGC_HOP_MIN_CTAS : 5; }
constexpr int64_t TICKET_MIN_WORK = 2 * 1000 * 1000;      // gathers per launch above which the ticket grid pays

struct HopDev {
    const int64_t *row_ptr;
    const int32_t *col;
    const double *aval;
    const double *dinv;
    const double *dinv_rows;
    const float *sval;
    const float *x;
    int64_t ldx;
    float *y;
    int64_t ldy;
    float *partial;
    int64_t ldp;
    const float *mean;
    const float *stdv;
    const int32_t *li_row;
    const int64_t *li_beg;
    const int32_t *li_len;
    int64_t n_rows;
    int64_t n_long_items;
    int32_t f;
    int32_t chunk;
    int32_t n_slabs;
    int32_t y_vec;       // widest legal vector store into y (1, 2 or 4 floats)
    int32_t hints;       // bit0: L2 evict_last on gathered rows; bit1: evict_first on streams
    int32_t f_store;     // columns [0, f_store) of y may be written with whole vectors (f rounded up only when
                         // the pad columns belong to the operand; else == f and tails go element-wise)
    int32_t tpw;         // K5: tasks per tile and group slot of the peer variant (rows per tile = tpw * 32/G)
    int32_t stage_stride;// K5: floats between staged rows (ldy when tiles are contiguous in y, else the slab width)
    unsigned long long *tile_counter;   // [2]: next ticket, finished warps (zero between launches); null: one-shot grid
    int64_t work;        // stored entries x feature slabs of this launch (grid policy: ticket grid vs one-shot)
    int32_t accumulate;  // column-blocked hop: rows ADD onto what y holds (acc starts from y, rows without entries
                         // in this block are left alone); chunk items still start from 0, the fix-up adds them
    int32_t n_peers;     // K5: copies of y in other GPUs' memory that receive every output row
    float *peer_y[SGC_MAX_PEERS];
    const uint32_t *peer_mask;   // per-row bit set of peers that need the row (null: all)
};

// The first 256 bytes of a plan workspace hold 16 independent ticket-counter pairs.  Every ticket-scheduled launch
// takes the next pair in turn and clears it on its stream first, so hops on one plan from different streams do
// not share counters (up to 16 in flight), and a launch that died mid-run cannot leave stale tickets behind.
constexpr int TICKET_SLOTS = 16;

struct sgc_plan_impl {
    int64_t n_rows, n_cols, nnz;
    const int64_t *row_ptr;
    const int32_t *col;
    int32_t chunk;
    int64_t n_long_items, n_long_rows;
    const int32_t *li_row;
    const int64_t *li_beg;
    const int32_t *li_len;
    const int32_t *lr_row;
    const int64_t *lr_first;
    const int32_t *lr_cnt;
    unsigned long long *tile_counter;   // head of the plan workspace: TICKET_SLOTS pairs (next ticket, retired warps)
    mutable std::atomic<uint32_t> ticket_seq{0};   // launches so far: launch i uses pair i % TICKET_SLOTS
    const int32_t *col_tagged;   // col with bit 31 set on "hot" columns (sgc_spmm_plan_tag_hot), or null
    int64_t n_hot_cols;
    int32_t hot_threshold;
};

}  // namespace sgc

struct sgc_plan { sgc::sgc_plan_impl impl; };     // the opaque sgc_plan_t of the C ABI

namespace sgc {

typedef int32_t (*hop_launch_fn)(const HopDev &, int64_t n_tasks, cudaStream_t);

// Filled by the instantiation units (spmm_inst_*.cu).
hop_launch_fn find_hop_kernel(int g, int vec, int u, bool fused, int stdmode, bool peers, bool hot);

#ifdef __CUDACC__

__device__ __forceinline__ uint64_t l2_policy(int kind)
{
    uint64_t p;
    if (kind == 1)
        asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    else if (kind == 2)
        asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    else
        asm("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
    return p;
}

template <int VEC> struct Vec;
template <> struct Vec<4> {
    float v[4];
    __device__ __forceinline__ void load(const float *p, uint64_t pol)
    {
        asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                     : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "l"(p), "l"(pol));
    }
    // Zero-instruction scheduling fence: volatile asm statements keep their program order, so
    // every gather of a batch is issued before the first FMA that consumes one.
    __device__ __forceinline__ void pin() { asm volatile("" : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3])); }
};
template <> struct Vec<2> {
    float v[2];
    __device__ __forceinline__ void load(const float *p, uint64_t pol)
    {
        asm volatile("ld.global.nc.L2::cache_hint.v2.f32 {%0,%1}, [%2], %3;"
                     : "=f"(v[0]), "=f"(v[1]) : "l"(p), "l"(pol));
    }
    __device__ __forceinline__ void pin() { asm volatile("" : "+f"(v[0]), "+f"(v[1])); }
};
template <> struct Vec<1> {
    float v[1];
    __device__ __forceinline__ void load(const float *p, uint64_t pol)
    {
        asm volatile("ld.global.nc.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(v[0]) : "l"(p), "l"(pol));
    }
    __device__ __forceinline__ void pin() { asm volatile("" : "+f"(v[0])); }
};

__device__ __forceinline__ int32_t ld_stream_i32(const int32_t *p, uint64_t pol)
{
    int32_t r;
    asm("ld.global.nc.L1::no_allocate.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(r) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ float ld_stream_f32(const float *p, uint64_t pol)
{
    float r;
    asm("ld.global.nc.L1::no_allocate.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(r) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ double ld_stream_f64(const double *p, uint64_t pol)
{
    double r;
    asm("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(r) : "l"(p), "l"(pol));
    return r;
}

__device__ __forceinline__ void st_hint_v4(float *p, const float *v, uint64_t pol)
{
    asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;"
                 :: "l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_hint_v2(float *p, const float *v, uint64_t pol)
{
    asm volatile("st.global.L2::cache_hint.v2.f32 [%0], {%1,%2}, %3;"
                 :: "l"(p), "f"(v[0]), "f"(v[1]), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_hint_v1(float *p, float v, uint64_t pol)
{
    asm volatile("st.global.L2::cache_hint.f32 [%0], %1, %2;" :: "l"(p), "f"(v), "l"(pol) : "memory");
}

// Plain stores for peer (NVLink-mapped) addresses: cache policies do not apply to them.
__device__ __forceinline__ void st_peer_v4(float *p, const float *v)
{
    asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
}
__device__ __forceinline__ void st_peer_v2(float *p, const float *v)
{
    asm volatile("st.global.v2.f32 [%0], {%1,%2};" :: "l"(p), "f"(v[0]), "f"(v[1]) : "memory");
}
__device__ __forceinline__ void st_peer_v1(float *p, float v)
{
    asm volatile("st.global.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
}

// STD: 0 none, 1 exact division, 2 multiply by reciprocal
template <int STD>
__device__ __forceinline__ float standardize(float x, float mu, float s)
{
    if (STD == 1) return __fdiv_rn(__fsub_rn(x, mu), s);
    if (STD == 2) return __fmul_rn(__fsub_rn(x, mu), s);
    return x;
}

// One task of the hop: a group of G lanes accumulates one (row or chunk item) x (feature slab) piece in
// registers and stores it to y / the partial-sum scratch.  Returns what the peer epilogue needs.
template <int G, int VEC, int U>
struct TaskOut {
    float acc[U][VEC];
    int fo[U];
    bool act[U];
    int64_t row;
    int f0;
    int sv;
    bool valid, to_y;
};

// HOT: p.col carries a tag in bit 31 ("this column is among the most frequently gathered ones");
// tagged rows of X are loaded with an L2 evict_last policy, all others with evict_first, which
// pins the hub rows in L2 when X as a whole is far larger than L2.
template <int G, int VEC, int U, bool FUSED, int STD, bool HOT>
__device__ __forceinline__ void hop_task(const HopDev &p, const int64_t task, const int lane, const uint64_t pol_x,
                                         const uint64_t pol_hot, const uint64_t pol_s, TaskOut<G, VEC, U> &o)
{
    constexpr int GPW = 32 / G;                       // tasks per warp
    constexpr int SW = G * VEC * U;                   // slab width in floats
    constexpr int UNR = (U * VEC >= 8) ? 4 : 8;       // gathers in flight per lane
    static_assert(G % UNR == 0, "unroll must divide the group size");
    const int gl = lane % G;
    const int64_t per_slab = p.n_long_items + p.n_rows;
    const int64_t slab = task / per_slab;
    const int64_t t = task - slab * per_slab;

    bool valid = slab < p.n_slabs;
    int64_t row = 0, beg = 0, end = 0;
    float *dst = nullptr;
    bool to_y = true;
    if (valid) {
        if (t < p.n_long_items) {
            row = p.li_row[t];
            beg = p.li_beg[t];
            end = beg + p.li_len[t];
            dst = p.partial + t * p.ldp;
            to_y = false;
        } else {
            row = t - p.n_long_items;
            beg = p.row_ptr[row];
            end = p.row_ptr[row + 1];
            if (p.chunk > 0 && end - beg > p.chunk) valid = false;   // owned by chunk items
            dst = p.y + row * p.ldy;
        }
    }
    if (!valid) beg = end = 0;
    // column-blocked hop (sgc_b200/propagate.py: hop_col_blocked): this launch covers one block of S's columns
    // and continues the row's sum where the previous block's launch left it -- same FMA chain, same order
    const bool accum = p.accumulate != 0 && to_y;
    if (accum && valid && beg == end) valid = false;     // nothing in this block: y keeps its value

    const int f0 = (int)slab * SW;
    int fo_ld[U];
    float mu[U][VEC], sg[U][VEC];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        o.fo[u] = f0 + (u * G + gl) * VEC;
        o.act[u] = valid && o.fo[u] < p.f;
        fo_ld[u] = o.act[u] ? o.fo[u] : 0;
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
            mu[u][v] = 0.f;
            sg[u][v] = 1.f;
            if (STD != 0 && o.act[u] && o.fo[u] + v < p.f) {
                mu[u][v] = p.mean[o.fo[u] + v];
                const float s = p.stdv[o.fo[u] + v];
                sg[u][v] = (STD == 2) ? __frcp_rn(s) : s;
            }
        }
    }

#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
        for (int v = 0; v < VEC; ++v) o.acc[u][v] = 0.f;
    if (accum && valid) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (!o.act[u]) continue;
            const float *q = dst + o.fo[u];
            if (VEC == 4 && p.y_vec == 4 && o.fo[u] + 4 <= p.f_store) {
                const float4 t = *reinterpret_cast<const float4 *>(q);
                o.acc[u][0] = t.x; o.acc[u][1 % VEC] = t.y; o.acc[u][2 % VEC] = t.z; o.acc[u][3 % VEC] = t.w;
            } else {
#pragma unroll
                for (int v = 0; v < VEC; ++v)
                    if (o.fo[u] + v < p.f) o.acc[u][v] = q[v];
            }
        }
    }

    const double di = (FUSED && valid) ? p.dinv_rows[row] : 0.0;
    const int len = (int)(end - beg);
    const int nb = (len + G - 1) / G;
    const int nb_max = (GPW == 1) ? nb : __reduce_max_sync(0xffffffffu, nb);

    for (int bt = 0; bt < nb_max; ++bt) {
        const int64_t base = beg + (int64_t)bt * G;
        const int64_t idx = base + gl;
        int32_t c = 0;
        float w = 0.f;
        if (idx < end) {
            c = ld_stream_i32(p.col + idx, pol_s);
            if (FUSED) {
                const double a = ld_stream_f64(p.aval + idx, pol_s);
                const double dj = __ldg(p.dinv + (HOT ? (c & 0x7fffffff) : c));
                w = __double2float_rn(norm_value64(di, a, dj));
            } else {
                w = ld_stream_f32(p.sval + idx, pol_s);
            }
        }
        int cnt = (int)(end - base);
        cnt = cnt < 0 ? 0 : (cnt > G ? G : cnt);
        const int cnt_max = (GPW == 1) ? cnt : __reduce_max_sync(0xffffffffu, cnt);

        for (int j0 = 0; j0 < cnt_max; j0 += UNR) {
            int32_t cj[UNR];
            float wj[UNR];
            Vec<VEC> xv[UNR][U];
#pragma unroll
            for (int jj = 0; jj < UNR; ++jj) {
                cj[jj] = __shfl_sync(0xffffffffu, c, j0 + jj, G);
                wj[jj] = __shfl_sync(0xffffffffu, w, j0 + jj, G);
            }
            // Gathers are unconditional: lanes past the end of their row hold column 0 and lanes
            // past the last feature read offset 0, so every address is valid; only the FMAs are
            // predicated.  This keeps the UNR loads back to back in the instruction stream.
#pragma unroll
            for (int jj = 0; jj < UNR; ++jj) {
                const float *xr = p.x + (int64_t)(HOT ? (cj[jj] & 0x7fffffff) : cj[jj]) * p.ldx;
                const uint64_t pol = (HOT && cj[jj] < 0) ? pol_hot : pol_x;
#pragma unroll
                for (int u = 0; u < U; ++u) xv[jj][u].load(xr + fo_ld[u], pol);
            }
#pragma unroll
            for (int jj = 0; jj < UNR; ++jj)
#pragma unroll
                for (int u = 0; u < U; ++u) xv[jj][u].pin();
#pragma unroll
            for (int jj = 0; jj < UNR; ++jj) {
                if (j0 + jj < cnt) {
#pragma unroll
                    for (int u = 0; u < U; ++u)
#pragma unroll
                        for (int v = 0; v < VEC; ++v)
                            o.acc[u][v] = __fmaf_rn(wj[jj],
                                                    standardize<STD>(xv[jj][u].v[v], mu[u][v], sg[u][v]),
                                                    o.acc[u][v]);
                }
            }
        }
    }

    const int sv = to_y ? (p.y_vec < VEC ? p.y_vec : VEC) : VEC;   // partial rows are padded and aligned
    const uint64_t pol_y = pol_s;
    if (valid) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (!o.act[u]) continue;
            float *q = dst + o.fo[u];
            if (VEC == 4 && sv == 4 && o.fo[u] + 4 <= p.f_store) {
                st_hint_v4(q, o.acc[u], pol_y);
            } else if (VEC >= 2 && sv >= 2 && o.fo[u] + VEC <= p.f_store) {
#pragma unroll
                for (int v = 0; v < VEC; v += 2) st_hint_v2(q + v, &o.acc[u][v], pol_y);
            } else {
#pragma unroll
                for (int v = 0; v < VEC; ++v)
                    if (o.fo[u] + v < p.f) st_hint_v1(q + v, o.acc[u][v], pol_y);
            }
        }
    }
    o.row = row; o.f0 = f0; o.sv = sv; o.valid = valid; o.to_y = to_y;
}

template <int G, int VEC, int U, bool FUSED, int STD, bool HOT>
__global__ void __launch_bounds__(HOP_THREADS, (hop_min_ctas<U, STD>()))
spmm_hop_kernel(const HopDev p)
{
    constexpr int GPW = 32 / G;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (HOP_THREADS / 32) + (threadIdx.x >> 5);
    const uint64_t pol_x = l2_policy(HOT ? 2 : ((p.hints & 1) ? 1 : 0));    // HOT: policy of untagged rows
    const uint64_t pol_hot = l2_policy(1);
    const uint64_t pol_s = l2_policy((p.hints & 2) ? 2 : 0);
    TaskOut<G, VEC, U> o;
    hop_task<G, VEC, U, FUSED, STD, HOT>(p, warp * GPW + lane / G, lane, pol_x, pol_hot, pol_s, o);
}

// ---- ticket scheduling ------------------------------------------------------------------------------
// Persistent grids draw their work from a global counter: rows of a power-law graph differ in cost by orders of
// magnitude, and a warp that is handed its next piece of work when it asks for it leaves no tail (measured on
// B200: Reddit shape 4.70 -> 4.52 ms per hop, products 7.85 -> 7.52 ms against one warp per task in a one-shot
// grid; bit-identical results).  Tickets are handed out in task order, so the rows being gathered at any time
// are neighbours in the CSR and share a feature slab, exactly as in the one-shot grid.
// counters[0] = next ticket, counters[1] = warps that have finished.  The last warp to finish resets both, so
// the next launch on this plan (stream-ordered) starts from zero without a memset.
__device__ __forceinline__ int64_t next_ticket(unsigned long long *counters, int lane)
{
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(counters, 1ull);
    return (int64_t)__shfl_sync(0xffffffffu, t, 0);
}

__device__ __forceinline__ void retire_warp(unsigned long long *counters, int lane, unsigned long long n_warps)
{
    if (lane == 0) {
        const unsigned long long done = atomicAdd(counters + 1, 1ull);
        if (done == n_warps - 1) {
            counters[0] = 0;
            counters[1] = 0;
            __threadfence();
        }
    }
}

// The single-GPU hop on a persistent grid: every warp draws tiles of `tpw` x 32/G consecutive tasks.
template <int G, int VEC, int U, bool FUSED, int STD, bool HOT>
__global__ void __launch_bounds__(HOP_THREADS, (hop_min_ctas<U, STD>()))
spmm_hop_ticket_kernel(const HopDev p, const int64_t n_tiles)
{
    constexpr int GPW = 32 / G;
    const int lane = threadIdx.x & 31;
    const uint64_t pol_x = l2_policy(HOT ? 2 : ((p.hints & 1) ? 1 : 0));
    const uint64_t pol_hot = l2_policy(1);
    const uint64_t pol_s = l2_policy((p.hints & 2) ? 2 : 0);
    const int R = p.tpw;
#pragma unroll 1
    for (;;) {
        const int64_t tile = next_ticket(p.tile_counter, lane);
        if (tile >= n_tiles) break;
        const int64_t t0 = tile * R * GPW;
#pragma unroll 1
        for (int it = 0; it < R; ++it) {
            TaskOut<G, VEC, U> o;
            hop_task<G, VEC, U, FUSED, STD, HOT>(p, t0 + it * GPW + lane / G, lane, pol_x, pol_hot, pol_s, o);
        }
    }
    retire_warp(p.tile_counter, lane, (unsigned long long)gridDim.x * (HOP_THREADS / 32));
}

// K5: the hop with the NVLink exchange fused in.  Every output row also goes to each peer's copy of the next
// hop's gathered operand, inside this kernel, while other warps are still gathering.
//
// Persistent grid (one wave of CTAs); every warp draws TILES of RT = tpw * GPW consecutive tasks from a global
// ticket counter (rows of a power-law graph differ in cost by orders of magnitude: a static split of the tiles
// over the warps left a 15 % tail, measured at 2 GPUs; tickets are handed out in task order, so the rows being
// gathered at any time are still neighbours in the CSR).  The output rows of a tile are staged in shared
// memory and handed to the TMA unit as cp.async.bulk shared -> global stores, so the LSU never waits on NVLink:
//   * tile store: when the tile is RT consecutive rows of y, all computed here, and a peer needs all of
//     them, ONE bulk store of RT * ldy * 4 bytes goes to that peer (rows are contiguous in the operand,
//     pad columns included).  Measured 8-GPU all-to-all: 416-byte stores 578 GB/s per GPU, 1.6 KB 641,
//     8 KB 670 (profiles/r01_nvlink_probe_n8_products.log) -- row-sized stores were the limiter of the
//     first version of this kernel (491 GB/s inside the hop);
//   * row store: otherwise each lane l < RT pushes row l of the tile, piece by piece, to the peers whose
//     halo mask asks for it.
// Two staging buffers per warp: the stores of tile i keep reading buffer i&1 while tile i+1 is gathered;
// before a buffer is overwritten every lane waits until at most the newer group is still being read
// (wait_group.read 1).  Before the warp exits every lane waits for full COMPLETION of its stores
// (cp.async.bulk.wait_group 0, not .read): the writes have been performed at the peers when the grid
// completes, and the stream-ordered barrier that follows the kernel (dist.py) publishes them.
#ifndef SGC_PEERS_MIN_CTAS
#define SGC_PEERS_MIN_CTAS 6
#endif
template <int G, int VEC, int U, bool FUSED, bool HOT>
__global__ void __launch_bounds__(HOP_THREADS, SGC_PEERS_MIN_CTAS)
spmm_hop_peers_kernel(const HopDev p, const int64_t n_tiles)
{
    constexpr int GPW = 32 / G;
    constexpr int SW = G * VEC * U;
    constexpr int WARPS = HOP_THREADS / 32;
    const int lane = threadIdx.x & 31;
    const int gl = lane % G;
    const int wic = threadIdx.x >> 5;
    const uint64_t pol_x = l2_policy(HOT ? 2 : ((p.hints & 1) ? 1 : 0));
    const uint64_t pol_hot = l2_policy(1);
    const uint64_t pol_s = l2_policy((p.hints & 2) ? 2 : 0);
    const int R = p.tpw;                       // tasks per tile and group slot
    const int RT = R * GPW;                    // rows (tasks) per tile, <= 32
    const int stride = p.stage_stride;         // floats between staged rows: ldy (tile-capable) or SW
    const bool bulk = p.y_vec == 4 && !(p.hints & 16);
    const bool tile_capable = bulk && stride == (int)p.ldy && !(p.hints & 32);

    extern __shared__ __align__(128) float stage_all[];
    __shared__ int32_t row_s[WARPS][32];       // local row of each staged slot, -1: nothing to push
    float *stage_w = stage_all + (size_t)wic * 2 * RT * stride;
    const int64_t per_slab = p.n_long_items + p.n_rows;

#pragma unroll 1
    for (int iter = 0;; ++iter) {
        const int64_t tile = next_ticket(p.tile_counter, lane);
        if (tile >= n_tiles) break;
        float *buf = stage_w + (size_t)(iter & 1) * RT * stride;
        if (bulk) {
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncwarp();
        }
        const int64_t t0 = tile * RT;
        // RT consecutive rows of y: all tasks are row tasks of one slab (long-row chunk items come first)
        bool whole = tile_capable && (t0 % per_slab) >= p.n_long_items && (t0 % per_slab) + RT <= per_slab;
        uint32_t need_and = 0xffffffffu;
        int64_t first_row = 0;
        int f0_tile = 0;
#pragma unroll 1
        for (int it = 0; it < R; ++it) {
            TaskOut<G, VEC, U> o;
            hop_task<G, VEC, U, FUSED, 0, HOT>(p, t0 + it * GPW + lane / G, lane, pol_x, pol_hot, pol_s, o);
            if (p.n_peers <= 0 || (p.hints & 8)) continue;     // hints bit 3: measurement aid, no push
            const bool push = o.valid && o.to_y;
            if (bulk) {
                const int slot = it * GPW + lane / G;
                float *mine = buf + (size_t)slot * stride;
                if (push) {
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int c = (u * G + gl) * VEC;
                        if (c < stride) {
#pragma unroll
                            for (int v = 0; v < VEC; ++v) mine[c + v] = o.act[u] ? o.acc[u][v] : 0.f;
                        }
                    }
                }
                if (gl == 0) row_s[wic][slot] = push ? (int32_t)o.row : -1;
                whole = whole && __all_sync(0xffffffffu, push);
                if (it == 0) { first_row = __shfl_sync(0xffffffffu, o.row, 0); f0_tile = __shfl_sync(0xffffffffu, o.f0, 0); }
                if (p.peer_mask && push && gl == 0) need_and &= __ldg(p.peer_mask + o.row);
            } else if (push) {
                // unaligned outputs: fire-and-forget per-lane stores from the registers
                const int64_t off = o.row * p.ldy;
                const uint32_t need = p.peer_mask ? __ldg(p.peer_mask + o.row) : 0xffffffffu;
                for (int pp = 0; pp < p.n_peers; ++pp) {
                    if (!((need >> pp) & 1u)) continue;
                    float *base = p.peer_y[pp] + off;
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        if (!o.act[u]) continue;
                        float *q = base + o.fo[u];
                        if (VEC == 4 && o.sv == 4 && o.fo[u] + 4 <= p.f_store) {
                            st_peer_v4(q, o.acc[u]);
                        } else if (VEC >= 2 && o.sv >= 2 && o.fo[u] + VEC <= p.f_store) {
#pragma unroll
                            for (int v = 0; v < VEC; v += 2) st_peer_v2(q + v, &o.acc[u][v]);
                        } else {
#pragma unroll
                            for (int v = 0; v < VEC; ++v)
                                if (o.fo[u] + v < p.f) st_peer_v1(q + v, o.acc[u][v]);
                        }
                    }
                }
            }
        }
        if (!bulk || p.n_peers <= 0 || (p.hints & 8)) continue;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        need_and = __reduce_and_sync(0xffffffffu, need_and);
        const uint32_t all_peers = (p.n_peers >= 32) ? 0xffffffffu : ((1u << p.n_peers) - 1u);
        const uint32_t tile_peers = whole ? (need_and & all_peers) : 0u;
        if (lane == 0 && tile_peers) {
            const uint32_t src = (uint32_t)__cvta_generic_to_shared(buf);
            const uint32_t bytes = (uint32_t)RT * (uint32_t)stride * 4u;
            const int64_t off = first_row * p.ldy;
            for (int pp = 0; pp < p.n_peers; ++pp)
                if ((tile_peers >> pp) & 1u)
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                                 :: "l"(p.peer_y[pp] + off), "r"(src), "r"(bytes) : "memory");
        }
        if (lane < RT && tile_peers != all_peers) {
            const int32_t r = row_s[wic][lane];
            if (r >= 0) {
                uint32_t need = (p.peer_mask ? __ldg(p.peer_mask + r) : 0xffffffffu) & all_peers & ~tile_peers;
                // the slab of this slot's task (a tile may straddle two slabs when it is not `whole`)
                const int64_t task = t0 + lane;
                const int f0 = whole ? f0_tile : (int)(task / per_slab) * SW;
                int nfl = p.f - f0;
                nfl = nfl > SW ? SW : nfl;
                const uint32_t bytes = (uint32_t)((nfl + 3) & ~3) * 4u;       // pad columns lie inside ldy
                const uint32_t src = (uint32_t)__cvta_generic_to_shared(buf + (size_t)lane * stride);
                const int64_t off = (int64_t)r * p.ldy + f0;
                for (int pp = 0; need; ++pp, need >>= 1)
                    if (need & 1u)
                        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                                     :: "l"(p.peer_y[pp] + off), "r"(src), "r"(bytes) : "memory");
            }
        }
        // every lane commits one (possibly empty) group per tile, so "1 pending" always means "the previous
        // tile's stores"
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    // completion (not just source reads): the peers' memory holds every row before this grid completes
    if (p.hints & 64) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // measurement aid
    else asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    retire_warp(p.tile_counter, lane, (unsigned long long)gridDim.x * WARPS);
}

template <int G, int VEC, int U, bool FUSED, int STD, bool PEERS, bool HOT>
int32_t launch_hop(const HopDev &p, int64_t n_tasks, cudaStream_t s)
{
    constexpr int GPW = 32 / G;
    if constexpr (PEERS) {
        static_assert(STD == 0, "the peer variant has no fused standardisation");
        const int RT = p.tpw * GPW;
        const int64_t n_tiles = ceil_div(n_tasks, RT);
        if (n_tiles <= 0) return SGC_OK;
        const size_t smem = (size_t)(HOP_THREADS / 32) * 2 * RT * p.stage_stride * sizeof(float);
        auto kern = spmm_hop_peers_kernel<G, VEC, U, FUSED, HOT>;
        SGC_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        SGC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, HOP_THREADS, smem));
        if (occ < 1) {
            set_error("spmm_hop: peer variant does not fit an SM (%zu bytes of shared memory)", smem);
            return SGC_ERR_UNSUPPORTED;
        }
        const int64_t want = ceil_div(n_tiles, HOP_THREADS / 32);
        const int64_t cap = (int64_t)sm_count() * occ;
        SGC_CUDA_TRY(cudaMemsetAsync(p.tile_counter, 0, 2 * sizeof(unsigned long long), s));
        kern<<<(unsigned)(want < cap ? want : cap), HOP_THREADS, smem, s>>>(p, n_tiles);
        SGC_LAUNCH_CHECK();
        return SGC_OK;
    } else {
        if (p.tile_counter != nullptr) {
            auto kern = spmm_hop_ticket_kernel<G, VEC, U, FUSED, STD, HOT>;
            static int occ = 0;                         // per instantiation; the device type is fixed (sm_100a)
            if (occ == 0) SGC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, HOP_THREADS, 0));
            const int64_t cap = (int64_t)sm_count() * (occ > 0 ? occ : 1);
            // small problems (Cora, Pubmed: a few tasks per resident warp, a few 100 k gathers) are launch-bound and
            // have no tail to speak of: one warp per task in a one-shot grid (Pubmed K=2: 0.092 ms vs 0.129 ms with
            // tickets).  Otherwise tickets: by task count, or by work -- a row part of a sharded graph has few tasks
            // per resident warp but millions of gathers (Reddit shape, a quarter of the rows x 2 slabs: 121 k tasks,
            // 11.7 M gathers).  hints bit 8 forces tickets, bit 7 the one-shot grid (A/B measurements).
            if (n_tasks > 32 * cap * (HOP_THREADS / 32) * GPW || p.work > TICKET_MIN_WORK || (p.hints & 256)) {
                const int RT = p.tpw * GPW;
                const int64_t n_tiles = ceil_div(n_tasks, RT);
                const int64_t want = ceil_div(n_tiles, HOP_THREADS / 32);
                SGC_CUDA_TRY(cudaMemsetAsync(p.tile_counter, 0, 2 * sizeof(unsigned long long), s));
                kern<<<(unsigned)(want < cap ? want : cap), HOP_THREADS, 0, s>>>(p, n_tiles);
                SGC_LAUNCH_CHECK();
                return SGC_OK;
            }
        }
        const int64_t warps = ceil_div(n_tasks, GPW);
        const int64_t blocks = ceil_div(warps, HOP_THREADS / 32);
        if (blocks <= 0) return SGC_OK;
        if (blocks > 0x7fffffffLL) {
            set_error("spmm_hop: grid too large (%lld blocks)", (long long)blocks);
            return SGC_ERR_INVALID;
        }
        spmm_hop_kernel<G, VEC, U, FUSED, STD, HOT><<<(unsigned)blocks, HOP_THREADS, 0, s>>>(p);
        SGC_LAUNCH_CHECK();
        return SGC_OK;
    }
}

#endif  // __CUDACC__

}  // namespace sgc
