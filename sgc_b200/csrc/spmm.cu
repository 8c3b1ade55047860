// K4 host side: split-table ("plan") construction, kernel selection, the hop and K-hop entry
// points.  Kernels are in spmm.cuh; see there for the work decomposition.
#include <new>

#include "spmm.cuh"

namespace sgc {

hop_launch_fn find_hop_kernel_v4(int g, int u, bool fused, int stdmode, bool peers, bool hot);
hop_launch_fn find_hop_kernel_v2(int g, int u, bool fused, int stdmode, bool peers, bool hot);
hop_launch_fn find_hop_kernel_v1(int g, int u, bool fused, int stdmode, bool peers, bool hot);

hop_launch_fn find_hop_kernel(int g, int vec, int u, bool fused, int stdmode, bool peers, bool hot)
{
    if (vec == 4) return find_hop_kernel_v4(g, u, fused, stdmode, peers, hot);
    if (vec == 2) return find_hop_kernel_v2(g, u, fused, stdmode, peers, hot);
    if (vec == 1) return find_hop_kernel_v1(g, u, fused, stdmode, peers, hot);
    return nullptr;
}

// ---- plan construction ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
plan_count_kernel(int64_t n, const int64_t *__restrict__ row_ptr, int32_t chunk,
                  int64_t *__restrict__ n_chunks, int64_t *__restrict__ is_long)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t c = 0;
        if (i < n) {
            const int64_t len = row_ptr[i + 1] - row_ptr[i];
            if (len > chunk) c = (len + chunk - 1) / chunk;
        }
        n_chunks[i] = c;
        is_long[i] = c > 0 ? 1 : 0;
    }
}

__global__ void __launch_bounds__(256)
plan_fill_kernel(int64_t n, const int64_t *__restrict__ row_ptr, int32_t chunk,
                 const int64_t *__restrict__ chunk_off, const int64_t *__restrict__ long_off,
                 int32_t *__restrict__ li_row, int64_t *__restrict__ li_beg,
                 int32_t *__restrict__ li_len, int32_t *__restrict__ lr_row,
                 int64_t *__restrict__ lr_first, int32_t *__restrict__ lr_cnt)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = row_ptr[i], e = row_ptr[i + 1];
        if (e - b <= chunk) continue;
        const int64_t first = chunk_off[i];
        const int64_t cnt = chunk_off[i + 1] - first;
        const int64_t k = long_off[i];
        lr_row[k] = (int32_t)i;
        lr_first[k] = first;
        lr_cnt[k] = (int32_t)cnt;
        for (int64_t c = 0; c < cnt; ++c) {
            const int64_t cb = b + c * chunk;
            li_row[first + c] = (int32_t)i;
            li_beg[first + c] = cb;
            li_len[first + c] = (int32_t)((e - cb < chunk) ? (e - cb) : chunk);
        }
    }
}

struct PeerList { int32_t n; float *y[SGC_MAX_PEERS]; const uint32_t *mask; };

// y[row] = partial[first] + partial[first+1] + ... in chunk order; one thread per
// (long row, feature), consecutive threads on consecutive features.
__global__ void __launch_bounds__(256)
spmm_fixup_kernel(int64_t n_long_rows, const int32_t *__restrict__ lr_row,
                  const int64_t *__restrict__ lr_first, const int32_t *__restrict__ lr_cnt,
                  const float *__restrict__ partial, int64_t ldp, float *__restrict__ y,
                  int64_t ldy, int32_t f, PeerList peers, int32_t accumulate)
{
    const int64_t k = blockIdx.x;
    if (k >= n_long_rows) return;
    const int64_t row = lr_row[k], first = lr_first[k];
    const int cnt = lr_cnt[k];
    for (int c = threadIdx.x; c < f; c += blockDim.x) {
        float acc = partial[first * ldp + c];
        if (accumulate) acc = __fadd_rn(y[row * ldy + c], acc);       // column-blocked hop: onto the earlier blocks' sum
        for (int j = 1; j < cnt; ++j) acc = __fadd_rn(acc, partial[(first + j) * ldp + c]);
        y[row * ldy + c] = acc;
        const uint32_t need = peers.mask ? peers.mask[row] : 0xffffffffu;
        for (int pp = 0; pp < peers.n; ++pp)
            if ((need >> pp) & 1u) peers.y[pp][row * ldy + c] = acc;             // K5 push
    }
}

// ---- hot-column tagging ----------------------------------------------------------------------------
// When X is far larger than L2 (products / papers100M shapes) the hub rows are the only part of it
// worth keeping in L2.  One-off, at plan time: count how often every column is gathered, pick the
// count threshold above which at most `hot_rows` columns lie (a 64 Ki-bin histogram read back to
// the host), and write a private copy of col with bit 31 set on those columns.
constexpr int HOT_BINS = 65536;

__global__ void __launch_bounds__(256)
hot_count_kernel(int64_t nnz, const int32_t *__restrict__ col, uint32_t *__restrict__ cnt)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz;
         i += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(cnt + col[i], 1u);
}

__global__ void __launch_bounds__(256)
hot_hist_kernel(int64_t n_cols, const uint32_t *__restrict__ cnt, unsigned long long *__restrict__ hist)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_cols;
         i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t c = cnt[i];
        atomicAdd(hist + (c < HOT_BINS ? c : HOT_BINS - 1), 1ull);
    }
}

__global__ void __launch_bounds__(256)
hot_tag_kernel(int64_t nnz, const int32_t *__restrict__ col, const uint32_t *__restrict__ cnt,
               uint32_t threshold, int32_t *__restrict__ col_tagged)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int32_t c = col[i];
        col_tagged[i] = (cnt[c] >= threshold) ? (c | (int32_t)0x80000000) : c;
    }
}

static inline size_t align256(size_t v) { return (v + 255) / 256 * 256; }

struct PlanLayout {
    size_t off_counter;          // 256 bytes: the peer variant's tile ticket counter
    size_t off_chunks, off_long, off_scan, scan_bytes;
    size_t off_li_row, off_li_len, off_li_beg, off_lr_row, off_lr_cnt, off_lr_first, total;
    int64_t max_items, max_long;
};

static PlanLayout plan_layout(int64_t n_rows, int64_t nnz, int32_t chunk)
{
    PlanLayout L{};
    L.off_counter = 0;
    if (chunk <= 0) { L.total = 512; return L; }
    L.max_long = nnz / chunk + 1;
    L.max_items = 2 * (nnz / chunk) + 2;
    size_t o = 256;
    L.off_chunks = o; o += align256((size_t)(n_rows + 1) * 8);
    L.off_long = o;   o += align256((size_t)(n_rows + 1) * 8);
    L.scan_bytes = scan_workspace_bytes(n_rows + 1);
    L.off_scan = o;   o += align256(L.scan_bytes);
    L.off_li_row = o; o += align256((size_t)L.max_items * 4);
    L.off_li_len = o; o += align256((size_t)L.max_items * 4);
    L.off_li_beg = o; o += align256((size_t)L.max_items * 8);
    L.off_lr_row = o; o += align256((size_t)L.max_long * 4);
    L.off_lr_cnt = o; o += align256((size_t)L.max_long * 4);
    L.off_lr_first = o; o += align256((size_t)L.max_long * 8);
    L.total = o + 256;
    return L;
}

// Default slab width (floats) for a gathered operand of n_cols rows.
//   * everything fits in L2, or nothing does: whole rows (one pass over col/val, the longest
//     contiguous gathers).  Measured on B200, products shape (N = 2.45 M, F = 100, X = 980 MB):
//     whole rows 9.3 ms per hop, 64-float slabs 13.8 ms, 32-float slabs 20 ms.
//   * otherwise the widest slab whose N x SW x 4-byte working set fits the L2 budget, so the
//     slab stays resident while all rows sweep over it (Reddit shape: 128 floats = 119 MB,
//     5.0 ms per hop; see profiles/ for the sweep over 32..256).
constexpr double L2_SLAB_BUDGET = 128.0 * 1024 * 1024;

static bool slab_fits_l2(int64_t n_cols, int slab) { return (double)n_cols * slab * 4.0 <= L2_SLAB_BUDGET; }

static int default_slab(int64_t n_cols, int f, int vec)
{
    const int widths4[] = {256, 128, 64, 32};
    const int widths2[] = {128, 64, 32};
    const int widths1[] = {64, 32};
    const int *w = vec == 4 ? widths4 : (vec == 2 ? widths2 : widths1);
    const int nw = vec == 4 ? 4 : (vec == 2 ? 3 : 2);
    // narrowest built width that covers f in one slab (or the widest one if none does)
    int whole = w[0];
    for (int i = 0; i < nw; ++i)
        if (w[i] >= f) whole = w[i];
    int best = 0;
    for (int i = 0; i < nw; ++i)               // widest first
        if (w[i] <= whole && slab_fits_l2(n_cols, w[i])) { best = w[i]; break; }
    return best ? best : whole;
}

static bool slab_shape(int slab, int vec, int *g, int *u)
{
    const int lanes = slab / vec;     // G * U
    if (lanes == 64) { *g = 32; *u = 2; return true; }
    if (lanes == 32 || lanes == 16 || lanes == 8) { *g = lanes; *u = 1; return true; }
    return false;
}

static int32_t run_hop(const sgc_plan_impl *pl, const sgc_hop_args_t *a, const float *x,
                       int64_t ldx, float *y, int64_t ldy, uint32_t flags, bool push_peers,
                       cudaStream_t s)
{
    const bool fused = (flags & SGC_HOP_FUSED_NORM) != 0;
    const int stdmode = (flags & SGC_HOP_STD_DIV) ? 1 : ((flags & SGC_HOP_STD_RCP) ? 2 : 0);
    SGC_REQUIRE(x && y && a->f >= 1 && ldx >= a->f && ldy >= a->f, "spmm_hop: bad operand shape");
    SGC_REQUIRE(x != y, "spmm_hop: in-place hop is not supported");
    if (fused) SGC_REQUIRE(a->aval_d && a->dinv_d, "spmm_hop: FUSED_NORM needs aval_d and dinv_d");
    else SGC_REQUIRE(a->sval_d || pl->nnz == 0, "spmm_hop: sval_d is null");
    if (stdmode) SGC_REQUIRE(a->mean_d && a->std_d, "spmm_hop: standardisation needs mean_d and std_d");
    if (pl->n_long_items > 0)
        SGC_REQUIRE(a->partial_d && a->ldp >= a->f && a->ldp % 4 == 0 &&
                        (reinterpret_cast<uintptr_t>(a->partial_d) & 15) == 0,
                    "spmm_hop: partial_d must be 16-byte aligned with ldp %% 4 == 0 and ldp >= f");
    if (pl->n_rows == 0) return SGC_OK;

    int vec = 1;
    if (ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) vec = 4;
    else if (ldx % 2 == 0 && (reinterpret_cast<uintptr_t>(x) & 7) == 0) vec = 2;
    int y_vec = 1;
    if (ldy % 4 == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0) y_vec = 4;
    else if (ldy % 2 == 0 && (reinterpret_cast<uintptr_t>(y) & 7) == 0) y_vec = 2;

    int slab = a->slab_floats > 0 ? a->slab_floats : default_slab(pl->n_cols, a->f, vec);
    int g = 0, u = 0;
    if (!slab_shape(slab, vec, &g, &u)) {
        // requested width not built for this alignment: fall to the nearest built one
        slab = default_slab(pl->n_cols, a->f, vec);
        if (!slab_shape(slab, vec, &g, &u)) {
            set_error("spmm_hop: no kernel for slab %d vec %d", slab, vec);
            return SGC_ERR_UNSUPPORTED;
        }
    }
    const bool peers_on = push_peers && a->n_peers > 0;
    // hot-column tags are used when the plan has them, the caller asks (hints bit 2) and the
    // variant exists (no fused standardisation)
    const bool hot_on = pl->col_tagged != nullptr && (a->hints & 4) != 0 && stdmode == 0;
    hop_launch_fn fn = find_hop_kernel(g, vec, u, fused, stdmode, peers_on, hot_on);
    if (!fn) {
        set_error("spmm_hop: kernel variant g=%d vec=%d u=%d fused=%d std=%d peers=%d not built", g, vec, u,
                  (int)fused, stdmode, (int)peers_on);
        return SGC_ERR_UNSUPPORTED;
    }

    HopDev p{};
    p.row_ptr = pl->row_ptr; p.col = hot_on ? pl->col_tagged : pl->col;
    p.aval = a->aval_d; p.dinv = a->dinv_d; p.sval = a->sval_d;
    p.dinv_rows = a->dinv_rows_d ? a->dinv_rows_d : a->dinv_d;
    p.x = x; p.ldx = ldx; p.y = y; p.ldy = ldy;
    p.partial = a->partial_d; p.ldp = a->ldp;
    p.mean = a->mean_d; p.stdv = a->std_d;
    p.li_row = pl->li_row; p.li_beg = pl->li_beg; p.li_len = pl->li_len;
    p.n_rows = pl->n_rows; p.n_long_items = pl->n_long_items;
    p.f = a->f; p.chunk = pl->chunk;
    p.n_slabs = (a->f + slab - 1) / slab;
    p.work = pl->nnz * (int64_t)p.n_slabs;
    p.y_vec = y_vec;
    // whole-vector stores may spill into columns [f, round4(f)) only when the caller owns them as padding
    const bool pitch_owned = (flags & SGC_HOP_PITCH_OWNED) != 0;      // [f, ldy) is padding: y is not a column view
    const bool pad_owned = (flags & (SGC_HOP_PAD_OWNED | SGC_HOP_PITCH_OWNED)) != 0 && (a->f + 3) / 4 * 4 <= ldy;
    p.f_store = pad_owned ? (a->f + 3) / 4 * 4 : a->f;
    p.accumulate = (flags & SGC_HOP_ACCUMULATE) ? 1 : 0;
    SGC_REQUIRE(!p.accumulate || (!peers_on && stdmode == 0),
                "spmm_hop: SGC_HOP_ACCUMULATE excludes peer pushes and fused standardisation");
    p.hints = a->hints & (3 | 8 | 16 | 32 | 64 | 256);
    // keeping gathered rows with evict_last only helps when the slab can actually stay in L2
    if (!slab_fits_l2(pl->n_cols, slab < a->f ? slab : a->f)) p.hints &= ~1;
    SGC_REQUIRE(a->n_peers >= 0 && a->n_peers <= SGC_MAX_PEERS, "spmm_hop: n_peers out of range");
    p.n_peers = push_peers ? a->n_peers : 0;
    PeerList peers{};
    peers.n = p.n_peers;
    p.peer_mask = peers.mask = (p.n_peers > 0) ? a->peer_mask_d : nullptr;
    for (int i = 0; i < p.n_peers; ++i) {
        SGC_REQUIRE(a->peer_y_d[i] != nullptr, "spmm_hop: peer_y_d[%d] is null", i);
        p.peer_y[i] = peers.y[i] = a->peer_y_d[i];
        // the bulk-store path assumes the peers' copies are laid out like y (same alignment)
        SGC_REQUIRE(y_vec != 4 || (reinterpret_cast<uintptr_t>(a->peer_y_d[i]) & 15) == 0,
                    "spmm_hop: peer_y_d[%d] must be 16-byte aligned like y_d", i);
    }
    // ticket-scheduled persistent grid unless the plan has no counters or the caller asks for the one-shot grid
    // (hints bit 7, A/B measurements)
    // (the pair of counters this launch owns: see TICKET_SLOTS)
    unsigned long long *const tickets = pl->tile_counter
        ? pl->tile_counter + 2 * (pl->ticket_seq.fetch_add(1, std::memory_order_relaxed) % TICKET_SLOTS) : nullptr;
    p.tile_counter = (a->hints & 128) ? nullptr : tickets;
    // tasks per ticket, measured on B200 (profiles/r02_sweep_persistent.log): Reddit shape (5 slabs) 3.89 ms with
    // 1, 4.05 with 2, 4.26 with 4; products shape (whole rows) 7.51 / 7.39 / 7.38 ms
    p.tpw = a->push_tile_tasks > 0 ? a->push_tile_tasks : (p.n_slabs > 1 ? 1 : 4);
    if (p.tpw * (32 / g) > 32) p.tpw = g;                  // at most 32 tasks per ticket
    if (peers_on) {
        SGC_REQUIRE(y_vec != 4 || a->f % 4 == 0 || pad_owned,
                    "spmm_hop: peer pushes of a width that is not a multiple of 4 floats need SGC_HOP_PAD_OWNED "
                    "(the bulk stores cover the pad columns)");
        // tiles are contiguous in y when one slab covers the whole padded row; a tile store covers the columns
        // [f, ldy) of every row on the peers, so the caller must have declared them padding (PITCH_OWNED) --
        // otherwise rows leave one by one and only round4(f) floats of each are written
        p.stage_stride = (pitch_owned && p.n_slabs == 1 && ldy <= slab && y_vec == 4) ? (int32_t)ldy : slab;
        // rows per tile: push_tile_tasks tasks per group slot, at most 32 rows per warp tile.  Default: 4 where a
        // tile can leave as ONE bulk store (rows contiguous in the operand: bigger NVLink transactions -- fused
        // products hop on 8 GPUs 1.66 ms with single rows, 1.57 with 2, 1.52 with 4, 1.53 with 8, 1.87 with 16 --
        // and half the staging memory of 8, which comes out of L1: compute side 4.10 -> 3.98 ms on a half of the
        // rows, profiles/r02_sweep_peer_tiles_1gpu.log); 1 where every row piece is its own store anyway (feature
        // slabs narrower than the row: Reddit shape) -- there a tile is only a scheduling quantum, and 8 tasks per
        // ticket left a row part of 121 k tasks with 4 tickets per warp (1.62 ms per hop against 0.77 ms for the
        // plain kernel, profiles/r02_scale_n8_grid2x4_tile8.json)
        const int gpw = 32 / g;
        const bool tiles = p.stage_stride == (int32_t)ldy;
        int tpw = a->push_tile_tasks > 0 ? a->push_tile_tasks : (tiles ? 4 : 1);
        if (tpw * gpw > 32) tpw = 32 / gpw;
        p.tpw = tpw;
        SGC_REQUIRE(pl->tile_counter != nullptr, "spmm_hop: the plan has no workspace (needed by the peer variant)");
        p.tile_counter = tickets;
    }
    const int64_t n_tasks = (int64_t)p.n_slabs * (p.n_long_items + p.n_rows);
    int32_t rc = fn(p, n_tasks, s);
    if (rc != SGC_OK) return rc;
    if (pl->n_long_rows > 0) {
        spmm_fixup_kernel<<<(unsigned)pl->n_long_rows, 256, 0, s>>>(
            pl->n_long_rows, pl->lr_row, pl->lr_first, pl->lr_cnt, a->partial_d, a->ldp, y, ldy, a->f, peers,
            p.accumulate);
        SGC_LAUNCH_CHECK();
    }
    return SGC_OK;
}

// When the feature width is not a multiple of the slab and the remainder is narrow (at most half a
// slab), the last slab would run with most lanes of every group idle (e.g. 152 columns = one
// 128-float slab + 24 columns on 32-lane groups).  The remainder is then a second launch on the
// shifted column range with the narrowest group that covers it (24 columns -> 8 lanes, 4 rows per
// warp).  Columns are independent, so results are unchanged.
static int32_t run_hop_auto(const sgc_plan_impl *pl, const sgc_hop_args_t *a, const float *x,
                            int64_t ldx, float *y, int64_t ldy, uint32_t flags, bool push_peers,
                            cudaStream_t s)
{
    if (a->slab_floats == 0 && x && y && a->f > 32 && pl->n_rows > 0) {
        int vec = 1;
        if (ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) vec = 4;
        else if (ldx % 2 == 0 && (reinterpret_cast<uintptr_t>(x) & 7) == 0) vec = 2;
        const int S = default_slab(pl->n_cols, a->f, vec);
        const int tail = a->f % S, main = a->f - tail;
        if (main > 0 && tail > 0 && tail <= S / 2) {
            int St = 32;
            while (St < tail) St *= 2;
            if (St < S) {
                sgc_hop_args_t a1 = *a, a2 = *a;
                a1.f = main; a1.slab_floats = S;
                a2.f = tail; a2.slab_floats = St;
                if (a->mean_d) a2.mean_d = a->mean_d + main;
                if (a->std_d) a2.std_d = a->std_d + main;
                if (a->partial_d) a2.partial_d = a->partial_d + main;
                for (int i = 0; i < SGC_MAX_PEERS; ++i)
                    if (a->peer_y_d[i]) a2.peer_y_d[i] = a->peer_y_d[i] + main;
                int32_t rc = run_hop(pl, &a1, x, ldx, y, ldy, flags, push_peers, s);
                if (rc != SGC_OK) return rc;
                return run_hop(pl, &a2, x + main, ldx, y + main, ldy, flags, push_peers, s);
            }
        }
    }
    return run_hop(pl, a, x, ldx, y, ldy, flags, push_peers, s);
}

}  // namespace sgc

using namespace sgc;

extern "C" {

size_t sgc_spmm_plan_workspace_bytes(int64_t n_rows, int64_t nnz, int32_t chunk)
{
    if (n_rows < 0 || nnz < 0) return 0;
    return plan_layout(n_rows, nnz, chunk).total;
}

int32_t sgc_spmm_plan_create(int64_t n_rows, int64_t n_cols, int64_t nnz,
                             const int64_t *row_ptr_d, const int32_t *col_d, int32_t chunk,
                             void *ws_d, size_t ws_bytes, sgc_plan_t **plan_out, void *stream)
{
    SGC_REQUIRE(plan_out, "spmm_plan_create: plan_out is null");
    *plan_out = nullptr;
    SGC_REQUIRE(n_rows >= 0 && n_cols >= 0 && nnz >= 0 && row_ptr_d, "spmm_plan_create: bad arguments");
    SGC_REQUIRE(n_rows < 0x7fffffffLL && n_cols < 0x7fffffffLL, "spmm_plan_create: more than 2^31-1 rows/cols");
    SGC_REQUIRE(nnz == 0 || col_d, "spmm_plan_create: col_d is null");
    if (chunk < 0) chunk = 0;
    if (chunk > 0 && chunk < 32) chunk = 32;
    const PlanLayout L = plan_layout(n_rows, nnz, chunk);
    if (chunk > 0 && (!ws_d || ws_bytes < L.total)) {
        set_error("spmm_plan_create: workspace too small (%zu < %zu)", ws_bytes, L.total);
        return SGC_ERR_WORKSPACE;
    }
    // the first 256 bytes of the workspace hold the peer variant's ticket counter (when there is a workspace)
    const bool has_counter = ws_d != nullptr && ws_bytes >= 256;
    cudaStream_t s = as_stream(stream);
    sgc_plan *pl = new (std::nothrow) sgc_plan();
    SGC_REQUIRE(pl, "spmm_plan_create: out of host memory");
    sgc_plan_impl &P = pl->impl;
    P.n_rows = n_rows; P.n_cols = n_cols; P.nnz = nnz;
    P.row_ptr = row_ptr_d; P.col = col_d; P.chunk = chunk;
    P.tile_counter = has_counter ? reinterpret_cast<unsigned long long *>(ws_d) : nullptr;
    if (has_counter) {
        cudaError_t e0 = cudaMemsetAsync(ws_d, 0, 256, s);
        if (e0 != cudaSuccess) { delete pl; return cuda_fail(e0, "plan counters", __FILE__, __LINE__); }
    }
    if (chunk > 0 && n_rows > 0) {
        char *ws = reinterpret_cast<char *>(ws_d);
        int64_t *n_chunks = reinterpret_cast<int64_t *>(ws + L.off_chunks);
        int64_t *is_long = reinterpret_cast<int64_t *>(ws + L.off_long);
        const int grid = grid_for(ceil_div(n_rows + 1, 256), 8);
        plan_count_kernel<<<grid, 256, 0, s>>>(n_rows, row_ptr_d, chunk, n_chunks, is_long);
        count_launch();
        int32_t rc = scan_exclusive_i64(n_chunks, n_rows + 1, ws + L.off_scan, L.scan_bytes, s);
        if (rc == SGC_OK) rc = scan_exclusive_i64(is_long, n_rows + 1, ws + L.off_scan, L.scan_bytes, s);
        int64_t totals[2] = {0, 0};
        cudaError_t e = cudaSuccess;
        if (rc == SGC_OK) {
            e = cudaMemcpyAsync(&totals[0], n_chunks + n_rows, 8, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaMemcpyAsync(&totals[1], is_long + n_rows, 8, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            if (e != cudaSuccess) rc = cuda_fail(e, "plan readback", __FILE__, __LINE__);
        }
        if (rc == SGC_OK && (totals[0] > L.max_items || totals[1] > L.max_long)) {
            set_error("spmm_plan_create: internal bound exceeded");
            rc = SGC_ERR_INVALID;
        }
        if (rc != SGC_OK) { delete pl; return rc; }
        P.n_long_items = totals[0];
        P.n_long_rows = totals[1];
        int32_t *li_row = reinterpret_cast<int32_t *>(ws + L.off_li_row);
        int32_t *li_len = reinterpret_cast<int32_t *>(ws + L.off_li_len);
        int64_t *li_beg = reinterpret_cast<int64_t *>(ws + L.off_li_beg);
        int32_t *lr_row = reinterpret_cast<int32_t *>(ws + L.off_lr_row);
        int32_t *lr_cnt = reinterpret_cast<int32_t *>(ws + L.off_lr_cnt);
        int64_t *lr_first = reinterpret_cast<int64_t *>(ws + L.off_lr_first);
        if (P.n_long_rows > 0) {
            plan_fill_kernel<<<grid, 256, 0, s>>>(n_rows, row_ptr_d, chunk, n_chunks, is_long, li_row,
                                                 li_beg, li_len, lr_row, lr_first, lr_cnt);
            count_launch();
        }
        P.li_row = li_row; P.li_len = li_len; P.li_beg = li_beg;
        P.lr_row = lr_row; P.lr_cnt = lr_cnt; P.lr_first = lr_first;
        e = cudaPeekAtLastError();
        if (e != cudaSuccess) { delete pl; return cuda_fail(e, "plan kernels", __FILE__, __LINE__); }
    }
    *plan_out = pl;
    return SGC_OK;
}

void sgc_spmm_plan_destroy(sgc_plan_t *plan) { delete plan; }

size_t sgc_spmm_hot_workspace_bytes(int64_t n_cols)
{
    if (n_cols < 0) return 0;
    return align256((size_t)n_cols * sizeof(uint32_t)) + align256((size_t)HOT_BINS * sizeof(unsigned long long)) + 256;
}

int32_t sgc_spmm_plan_tag_hot(sgc_plan_t *plan, int64_t hot_rows, int32_t *col_tagged_d, void *ws_d,
                              size_t ws_bytes, int64_t *n_hot_out, void *stream)
{
    SGC_REQUIRE(plan, "plan_tag_hot: null plan");
    sgc_plan_impl &P = plan->impl;
    if (n_hot_out) *n_hot_out = 0;
    if (hot_rows <= 0 || P.nnz == 0 || P.n_cols == 0) {          // switch tagging off
        P.col_tagged = nullptr; P.n_hot_cols = 0; P.hot_threshold = 0;
        return SGC_OK;
    }
    SGC_REQUIRE(col_tagged_d, "plan_tag_hot: col_tagged_d is null");
    if (!ws_d || ws_bytes < sgc_spmm_hot_workspace_bytes(P.n_cols)) {
        set_error("plan_tag_hot: workspace too small");
        return SGC_ERR_WORKSPACE;
    }
    cudaStream_t s = as_stream(stream);
    char *ws = reinterpret_cast<char *>(ws_d);
    uint32_t *cnt = reinterpret_cast<uint32_t *>(ws);
    const size_t cnt_bytes = align256((size_t)P.n_cols * sizeof(uint32_t));
    unsigned long long *hist = reinterpret_cast<unsigned long long *>(ws + cnt_bytes);
    SGC_CUDA_TRY(cudaMemsetAsync(ws, 0, cnt_bytes + (size_t)HOT_BINS * sizeof(unsigned long long), s));
    hot_count_kernel<<<grid_for(ceil_div(P.nnz, 256), 8), 256, 0, s>>>(P.nnz, P.col, cnt);
    SGC_LAUNCH_CHECK();
    hot_hist_kernel<<<grid_for(ceil_div(P.n_cols, 256), 8), 256, 0, s>>>(P.n_cols, cnt, hist);
    SGC_LAUNCH_CHECK();
    unsigned long long *h = new (std::nothrow) unsigned long long[HOT_BINS];
    SGC_REQUIRE(h, "plan_tag_hot: out of host memory");
    cudaError_t e = cudaMemcpyAsync(h, hist, (size_t)HOT_BINS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { delete[] h; return cuda_fail(e, "plan_tag_hot readback", __FILE__, __LINE__); }
    // smallest threshold (>= 2: a column gathered once gains nothing from staying in L2) whose
    // column set fits the budget
    unsigned long long above = 0;
    int thr = HOT_BINS;             // "nothing is hot"
    for (int b = HOT_BINS - 1; b >= 2; --b) {
        if (above + h[b] > (unsigned long long)hot_rows) break;
        above += h[b];
        thr = b;
    }
    delete[] h;
    if (thr >= HOT_BINS || above == 0) {
        P.col_tagged = nullptr; P.n_hot_cols = 0; P.hot_threshold = 0;
        return SGC_OK;
    }
    hot_tag_kernel<<<grid_for(ceil_div(P.nnz, 256), 8), 256, 0, s>>>(P.nnz, P.col, cnt, (uint32_t)thr, col_tagged_d);
    SGC_LAUNCH_CHECK();
    P.col_tagged = col_tagged_d;
    P.n_hot_cols = (int64_t)above;
    P.hot_threshold = thr;
    if (n_hot_out) *n_hot_out = (int64_t)above;
    return SGC_OK;
}

int64_t sgc_spmm_plan_partial_rows(const sgc_plan_t *plan) { return plan ? plan->impl.n_long_items : 0; }

int32_t sgc_spmm_hop(const sgc_plan_t *plan, const sgc_hop_args_t *args, void *stream)
{
    SGC_REQUIRE(plan && args, "spmm_hop: null plan or args");
    return run_hop_auto(&plan->impl, args, args->x_d, args->ldx, args->y_d, args->ldy, args->flags, true,
                        as_stream(stream));
}

int32_t sgc_precompute(const sgc_plan_t *plan, const sgc_hop_args_t *args, int32_t k_hops,
                       float *buf_a_d, float *buf_b_d, int64_t ld_buf, void *stream)
{
    SGC_REQUIRE(plan && args, "precompute: null plan or args");
    SGC_REQUIRE(k_hops >= 1, "precompute: k_hops must be >= 1 (degree 0 returns the input on the host side)");
    SGC_REQUIRE(plan->impl.n_rows == plan->impl.n_cols, "precompute: K-hop propagation needs a square matrix");
    if (k_hops >= 2) SGC_REQUIRE(buf_a_d && ld_buf >= args->f, "precompute: buf_a_d required for k_hops >= 2");
    if (k_hops >= 3) SGC_REQUIRE(buf_b_d, "precompute: buf_b_d required for k_hops >= 3");
    cudaStream_t s = as_stream(stream);
    const uint32_t std_bits = SGC_HOP_STD_DIV | SGC_HOP_STD_RCP;
    const float *src = args->x_d;
    int64_t ld_src = args->ldx;
    for (int k = 0; k < k_hops; ++k) {
        const bool last = (k == k_hops - 1);
        float *dst = last ? args->y_d : ((k & 1) ? buf_b_d : buf_a_d);
        const int64_t ld_dst = last ? args->ldy : ld_buf;
        uint32_t flags = (k == 0) ? args->flags : (args->flags & ~std_bits);
        // the ping-pong buffers are this call's own padded storage; the last hop writes the caller's y
        flags = last ? flags : (flags | SGC_HOP_PAD_OWNED);
        int32_t rc = run_hop_auto(&plan->impl, args, src, ld_src, dst, ld_dst, flags, false, s);
        if (rc != SGC_OK) return rc;
        src = dst;
        ld_src = ld_dst;
    }
    return SGC_OK;
}

}  // extern "C"
