// Measurement aid: L2 -> SM read bandwidth of this GPU (bench.py `roofline.l2.peak`, tools/l2_peak.py).
//
// The Reddit-shaped hop gathers every row of a 119 MB feature slab ~100 times per pass: its traffic is
// served by L2, not DRAM, so the roofline that bounds it is the L2 -> SM path.  MEASURED_PEAKS.json only
// holds a DRAM copy figure; this kernel measures the other one on the box the bench runs on: every warp
// reads 512-byte pieces (32 lanes x 16 bytes, the hop kernel's gather shape) of a buffer small enough to
// stay L2-resident, with L1 allocation switched off so that every load is served by L2.
#include "common.cuh"

namespace sgc {

constexpr int PROBE_THREADS = 256;
constexpr int PROBE_UNROLL = 8;

__device__ __forceinline__ float4 ld_l2_v4(const float4 *p)
{
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ float4 ld_hint_v4(const float4 *p, uint64_t pol)
{
    float4 r;
    asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p), "l"(pol));
    return r;
}

// A permutation of [0, n): odd multiply + constant, xor-shift and odd multiply are each invertible modulo
// 2^bits (the smallest power of two >= n), and values that land in [n, 2^bits) are walked round the cycle until
// they fall inside (fewer than two steps on average).  Every piece is therefore read exactly once per repetition
// -- the working set is the whole buffer and the byte count bytes-per-rep is exact.
__device__ __forceinline__ int64_t permute_piece(uint64_t p, uint64_t n, int bits, uint64_t salt)
{
    const uint64_t mask = (bits >= 64) ? ~0ull : ((1ull << bits) - 1ull);
    const int sh = bits / 2 > 0 ? bits / 2 : 1;
    do {
        p = (p * 0x9E3779B97F4A7C15ull + salt) & mask;
        p ^= p >> sh;
        p = (p * 0xD1B54A32D192ED03ull) & mask;
        p ^= p >> sh;
    } while (p >= n);
    return (int64_t)p;
}

// pattern bit 0: 0 = pieces in address order (grid-stride), 1 = pieces in a pseudo-random order without
// repetition (the gather's: a permutation of the pieces, new for every repetition);
// pattern bit 1: 0 = L1::no_allocate loads, 1 = the hop kernel's own load (L1-allocating, L2 evict_last hint).
__global__ void __launch_bounds__(PROBE_THREADS)
l2_read_probe_kernel(const float4 *__restrict__ buf, int64_t n_pieces, int64_t stride_v4, int32_t reps,
                     int32_t pattern, float *__restrict__ sink)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (PROBE_THREADS / 32) + (threadIdx.x >> 5);
    const int64_t n_warps = (int64_t)gridDim.x * (PROBE_THREADS / 32);
    float acc = 0.f;
    uint64_t pol;
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    const bool hashed = (pattern & 1) != 0, alloc = (pattern & 2) != 0;
    int bits = 1;
    while (bits < 63 && (1ll << bits) < n_pieces) ++bits;
    for (int r = 0; r < reps; ++r) {
        for (int64_t p0 = warp; p0 < n_pieces; p0 += n_warps * PROBE_UNROLL) {
            float4 v[PROBE_UNROLL];
#pragma unroll
            for (int u = 0; u < PROBE_UNROLL; ++u) {
                int64_t p = p0 + (int64_t)u * n_warps;
                if (p >= n_pieces) {                        // tail of the last round: nothing to read, nothing counted
                    v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                    continue;
                }
                if (hashed) p = permute_piece((uint64_t)p, (uint64_t)n_pieces, bits, (uint64_t)r * 0xD1B54A32D192ED03ull);
                v[u] = alloc ? ld_hint_v4(buf + p * stride_v4 + lane, pol) : ld_l2_v4(buf + p * stride_v4 + lane);
            }
#pragma unroll
            for (int u = 0; u < PROBE_UNROLL; ++u) acc += v[u].x + v[u].y + v[u].z + v[u].w;
        }
    }
    if (acc == 1.2345e-30f) sink[0] = acc;      // keeps the loads alive; practically never true
}

}  // namespace sgc

using namespace sgc;

extern "C" {

int32_t sgc_l2_read_probe(const void *buf_d, int64_t bytes, int64_t piece_stride_bytes, int32_t reps,
                          int32_t pattern, int32_t ctas_per_sm, float *sink_d, void *stream)
{
    if (piece_stride_bytes <= 0) piece_stride_bytes = 512;
    SGC_REQUIRE(piece_stride_bytes >= 512 && piece_stride_bytes % 16 == 0, "l2_read_probe: bad piece stride");
    SGC_REQUIRE(buf_d && sink_d, "l2_read_probe: null pointer");
    SGC_REQUIRE((reinterpret_cast<uintptr_t>(buf_d) & 15) == 0 && bytes >= 512,
                "l2_read_probe: buffer must be 16-byte aligned and hold at least one 512-byte piece");
    SGC_REQUIRE(reps >= 1 && pattern >= 0 && pattern <= 3, "l2_read_probe: bad reps / pattern");
    if (ctas_per_sm < 1) ctas_per_sm = 8;
    const int grid = sm_count() * ctas_per_sm;
    l2_read_probe_kernel<<<grid, PROBE_THREADS, 0, as_stream(stream)>>>(
        reinterpret_cast<const float4 *>(buf_d), (bytes - 512) / piece_stride_bytes + 1, piece_stride_bytes / 16, reps,
        pattern, sink_d);
    SGC_LAUNCH_CHECK();
    return SGC_OK;
}

}  // extern "C"
