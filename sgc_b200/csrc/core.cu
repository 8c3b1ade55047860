// Housekeeping entry points and the device-wide exclusive scan used by every compaction step
// (self-loop insertion, zero-product pruning, long-row split table, transpose, sub-graph).
#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "common.cuh"

namespace sgc {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int32_t cuda_fail(cudaError_t e, const char *what, const char *file, int line)
{
    const char *base = strrchr(file, '/');
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e),
              base ? base + 1 : file, line, what);
    return SGC_ERR_CUDA;
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count()
{
    static thread_local int cached_dev = -1, cached = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
            cached = v;
        cached_dev = dev;
    }
    return cached;
}

// ---- exclusive scan (int64) --------------------------------------------------------------------
// Three-phase, recursive on the tile totals: each 256-thread block scans a tile of 2048 values
// held in registers (8 per thread), writes the tile-local exclusive scan and its total; the totals
// are scanned by the same routine; a last pass adds the tile offsets.  Purely HBM-bound: reads n,
// writes n, twice.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_kernel(int64_t *__restrict__ data, int64_t *__restrict__ tile_sums, int64_t n)
{
    __shared__ int64_t warp_tot[SCAN_THREADS / 32];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int64_t v[SCAN_ITEMS];
    int64_t sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        v[k] = (base + k < n) ? data[base + k] : 0;
        sum += v[k];
    }
    // inclusive scan of the per-thread sums across the warp, then across warps
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int64_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    int64_t warp_off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; ++w) {
        int64_t t = warp_tot[w];
        if (w < warp) warp_off += t;
        total += t;
    }
    int64_t run = warp_off + incl - sum;  // exclusive prefix of this thread inside the tile
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        if (base + k < n) data[base + k] = run;
        run += v[k];
    }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_add_offsets_kernel(int64_t *__restrict__ data, const int64_t *__restrict__ tile_offs, int64_t n)
{
    const int64_t off = tile_offs[blockIdx.x];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k)
        if (base + k < n) data[base + k] += off;
}

size_t scan_workspace_bytes(int64_t n)
{
    size_t total = 0;
    int64_t m = n;
    while (m > 1) {
        m = ceil_div(m, SCAN_TILE);
        total += (size_t)((m + 31) / 32 * 32) * sizeof(int64_t);
        if (m == 1) break;
    }
    return total + 256;
}

int32_t scan_exclusive_i64(int64_t *data, int64_t n, void *ws, size_t ws_bytes, cudaStream_t s)
{
    if (n <= 0) return SGC_OK;
    if (ws_bytes < scan_workspace_bytes(n) || ws == nullptr) {
        set_error("scan: workspace too small (%zu < %zu)", ws_bytes, scan_workspace_bytes(n));
        return SGC_ERR_WORKSPACE;
    }
    const int64_t tiles = ceil_div(n, SCAN_TILE);
    int64_t *sums = reinterpret_cast<int64_t *>(ws);
    scan_tiles_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, s>>>(data, sums, n);
    SGC_LAUNCH_CHECK();
    if (tiles > 1) {
        const size_t used = (size_t)((tiles + 31) / 32 * 32) * sizeof(int64_t);
        int32_t rc = scan_exclusive_i64(sums, tiles, reinterpret_cast<char *>(ws) + used,
                                        ws_bytes - used, s);
        if (rc != SGC_OK) return rc;
        scan_add_offsets_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, s>>>(data, sums, n);
        SGC_LAUNCH_CHECK();
    }
    return SGC_OK;
}

}  // namespace sgc

extern "C" {

const char *sgc_version(void) { return "sgc_b200 0.1.0 (sm_100a)"; }

const char *sgc_last_error(void) { return sgc::g_err; }

int64_t sgc_launch_count(void) { return sgc::g_launches.load(std::memory_order_relaxed); }

int32_t sgc_device_check(int32_t device)
{
    int n = 0;
    SGC_CUDA_TRY(cudaGetDeviceCount(&n));
    SGC_REQUIRE(device >= 0 && device < n, "device %d out of range (%d visible)", device, n);
    int major = 0, minor = 0;
    SGC_CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    SGC_CUDA_TRY(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
    if (major != 10) {
        sgc::set_error("device %d is compute capability %d.%d; libsgc_b200 contains sm_100a code only",
                       device, major, minor);
        return SGC_ERR_ARCH;
    }
    return SGC_OK;
}

int32_t sgc_l2_set_persisting_bytes(int64_t bytes, int64_t *granted_out)
{
    int dev = 0;
    SGC_CUDA_TRY(cudaGetDevice(&dev));
    int max_bytes = 0;
    SGC_CUDA_TRY(cudaDeviceGetAttribute(&max_bytes, cudaDevAttrMaxPersistingL2CacheSize, dev));
    if (bytes < 0) bytes = 0;
    if (bytes > (int64_t)max_bytes) bytes = max_bytes;
    SGC_CUDA_TRY(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)bytes));
    size_t got = 0;
    SGC_CUDA_TRY(cudaDeviceGetLimit(&got, cudaLimitPersistingL2CacheSize));
    if (granted_out) *granted_out = (int64_t)got;
    return SGC_OK;
}

size_t sgc_scan_workspace_bytes(int64_t n) { return sgc::scan_workspace_bytes(n); }

int32_t sgc_scan_exclusive_i64(int64_t *data_d, int64_t n, void *ws_d, size_t ws_bytes, void *stream)
{
    SGC_REQUIRE(n >= 0 && (n == 0 || data_d != nullptr), "scan: bad arguments");
    return sgc::scan_exclusive_i64(data_d, n, ws_d, ws_bytes, sgc::as_stream(stream));
}

int32_t sgc_copy_2d(void *dst, int64_t dst_pitch_bytes, const void *src, int64_t src_pitch_bytes,
                    int64_t width_bytes, int64_t rows, void *stream)
{
    if (rows == 0 || width_bytes == 0) return SGC_OK;
    SGC_REQUIRE(dst && src && rows > 0 && width_bytes > 0 && dst_pitch_bytes >= width_bytes &&
                    src_pitch_bytes >= width_bytes,
                "copy_2d: bad arguments");
    SGC_CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)dst_pitch_bytes, src, (size_t)src_pitch_bytes,
                                   (size_t)width_bytes, (size_t)rows, cudaMemcpyDefault,
                                   sgc::as_stream(stream)));
    return SGC_OK;
}


// Host-side helper of the host-to-host pipeline (pipeline.py): is every one of the n doubles at v (host
// memory) bit-equal to `value`?  Adjacency matrices of the reference's loaders are constant-valued (1.0:
// utils.py:62 networkx adjacency; 2.0: utils.py:116, adj + adj.T of the symmetric 0/1 Reddit matrix); the
// pipeline then fills the device array instead of sending nnz x 8 bytes over PCIe.  Memory-bound: `threads` workers
// over contiguous pieces, early exit once any piece has seen a mismatch or the time budget is spent
// (budget_us <= 0: none) -- the scan only pays while it is cheaper than the upload it saves.
// Returns 1 all equal / 0 not / 2 gave up on the budget / -1 bad input.
int32_t sgc_host_all_equal_f64(const double *v, int64_t n, double value, int32_t threads, int64_t budget_us)
{
    if (n < 0 || (n > 0 && v == nullptr)) return -1;
    if (n == 0) return 1;
    uint64_t want;
    memcpy(&want, &value, sizeof(want));
    const uint64_t *w = reinterpret_cast<const uint64_t *>(v);
    using clock = std::chrono::steady_clock;
    const clock::time_point deadline = clock::now() + std::chrono::microseconds(budget_us > 0 ? budget_us : 0);
    std::atomic<int> state{1};                         // 1 equal so far, 0 differs, 2 out of time
    auto scan = [&](int64_t lo, int64_t hi) {
        constexpr int64_t STEP = 1 << 15;              // 256 KB between looks at the flag and the clock
        for (int64_t b = lo; b < hi; b += STEP) {
            if (state.load(std::memory_order_relaxed) != 1) return;
            if (budget_us > 0 && clock::now() > deadline) {
                int expect = 1;
                state.compare_exchange_strong(expect, 2, std::memory_order_relaxed);
                return;
            }
            const int64_t e = b + STEP < hi ? b + STEP : hi;
            uint64_t acc = 0;
            for (int64_t i = b; i < e; ++i) acc |= w[i] ^ want;
            if (acc) { state.store(0, std::memory_order_relaxed); return; }
        }
    };
    int64_t t = threads < 1 ? 1 : threads;
    if (t > 64) t = 64;
    if (n < t * (int64_t)(1 << 16)) t = 1 + (n >> 16);  // at least 512 KB per worker
    if (t <= 1) { scan(0, n); return state.load(); }
    const int64_t per = (n + t - 1) / t;
    std::vector<std::thread> pool;
    try {
        pool.reserve((size_t)t - 1);
    } catch (...) {
        scan(0, n);
        return state.load();
    }
    for (int64_t k = 1; k < t; ++k) {
        const int64_t lo = k * per, hi = lo + per < n ? lo + per : n;
        if (lo >= hi) continue;
        try {
            pool.emplace_back(scan, lo, hi);
        } catch (...) {                                // no more threads to be had: this piece inline
            scan(lo, hi);
        }
    }
    scan(0, per < n ? per : n);
    for (auto &th : pool) th.join();
    return state.load();
}

}  // extern "C"
