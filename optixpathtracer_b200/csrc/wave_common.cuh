// wave_common.cuh — what the kernels of one wave share (wavefront.cu, shade_layered.cu):
// the per-wave parameter block (the reference's LaunchParams, LaunchParams.h:9-29, cut to
// what the stages read), the device-side counters, and the warp-aggregated queue append.
#pragma once

#include "integrator.cuh"
#include "wavefront.cuh"

struct B2ptRenderer;

namespace b2pt {

constexpr int kWaveThreads = 128;

struct WaveParams {
    CameraParams cam;
    const B2ptPointLight* lights;
    int32_t n_lights;
    int32_t max_bounces;
    int32_t mode;          // B2PT_BSDF_*
    uint64_t seed;
    uint32_t p0, np;       // pixel tile [p0, p0 + np)
    uint32_t spw;          // samples in this wave
    uint32_t frame0;       // frame.id of the wave's first sample
    uint32_t capacity;     // slots per SoA array
    int32_t cull_nee;      // 1: K6 drops shadow rays with a provably zero contribution
    int32_t cull_layered;  // 1: ... also for the layered BSDF (only where skipping f leaves the random stream
                           //    of the rest untouched: the Philox engine, shade_layered.cu)
};

// counters of one wave, zeroed before K4:  active[b] = rays traced at bounce b,
// queued[b][k] = hits shaded with BSDF k at bounce b
struct WaveCounts {
    uint32_t* active;   // [max_bounces + 1]
    uint32_t* queued;   // [max_bounces][4]
};

// layout of B2ptRenderer::d_counts (uint32 words); max_bounces <= 99 is enforced by b2pt_create / the setter
constexpr int kCountsActive = 0;          // [100]
constexpr int kCountsQueued = 100;        // [100][4]
constexpr int kCountsPartition = 501;     // [100][4][2] K7p cursors (LCG modes)
constexpr int kCountsLayered = 1301;      // [100][4]: f-hit count, task cursor, spare, spare (Philox engine)
constexpr int kCountsShadow = 1701;        // [100][4]: shadow rays per light bucket (K6 -> K8, ShadowRays::bucket_n)
constexpr int kCountsWords = 1701 + 400;

// ---- the layered walk engine (shade_layered.cu) ----------------------------------------------------
// B2PT_LAYERED_ENGINE=0 sends the Philox mode through the serial k_shade<LAYERED> too (A/B, tests)
bool layered_engine_enabled();
size_t layered_scratch_bytes_per_path();
// K7d-p / K7d-e / K7d-f of bounce `bounce` on r->stream; lcounts = 4 zeroed words (f-list length, task cursor)
void launch_layered_prep(B2ptRenderer* r, const WaveParams& wp, const PathState& cur, const WaveCounts& wc,
                         uint32_t* lcounts, uint32_t bounce);
void launch_layered_walks(B2ptRenderer* r, const WaveParams& wp, const WaveCounts& wc, uint32_t* lcounts, uint32_t bounce);
void launch_layered_finish(B2ptRenderer* r, const WaveParams& wp, const PathState& cur, const PathState& nxt,
                           const WaveCounts& wc, const uint32_t* lcounts, uint32_t bounce);
// parity-hook rows of the engine's per-lane logic (device pointers, default stream)
void launch_fast_layered_rows_f(int64_t n, const float* albedo, const float* roughness, const float* wo, const float* wi,
                                const uint32_t* seed, float* f_out);
void launch_fast_layered_rows_sample(int64_t n, const float* albedo, const float* roughness, const float* wo,
                                     const uint32_t* seed, uint8_t* ok, float* f_out, float* wi_out, float* pdf_out,
                                     uint8_t* flags_out);

#if defined(__CUDACC__)
__device__ __forceinline__ void path_identity(const WaveParams& wp, uint32_t id, uint32_t& pixel, uint32_t& sample) {
    const uint32_t s_local = id / wp.np;
    pixel = wp.p0 + (id - s_local * wp.np);
    sample = wp.frame0 + s_local;
}

// warp-aggregated append: every lane of the warp must call this
__device__ __forceinline__ uint32_t warp_append(bool want, uint32_t* counter) {
    const uint32_t mask = __ballot_sync(0xffffffffu, want);
    if (mask == 0u) return 0u;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + (uint32_t)__popc(mask & ((1u << lane) - 1u));
}

// block-aggregated append: ONE atomic per block and call instead of one per warp (the queue counters are single
// addresses every warp of the grid hits; a same-address atomic costs about half a nanosecond of L2 time and its
// round trip sits between a warp's loads and its stores).  Every thread of the block must call this, the same number
// of times; `trip` (the caller's loop counter) alternates the two copies of the shared scratch so that a fast warp's
// next call cannot overwrite what a slow warp still reads.  Positions are deterministic inside the block.
constexpr int kAppendWarps = kWaveThreads / 32;
struct BlockAppendScratch {
    uint32_t cnt[2][kAppendWarps];
    uint32_t base[2];
};
__device__ __forceinline__ uint32_t block_append(bool want, uint32_t* counter, BlockAppendScratch& sh, uint32_t trip) {
    const uint32_t mask = __ballot_sync(0xffffffffu, want);
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, b = trip & 1u;
    if (lane == 0) sh.cnt[b][warp] = (uint32_t)__popc(mask);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
#pragma unroll
        for (int w = 0; w < kAppendWarps; w++) tot += sh.cnt[b][w];
        sh.base[b] = tot ? atomicAdd(counter, tot) : 0u;
    }
    __syncthreads();
    uint32_t pos = sh.base[b] + (uint32_t)__popc(mask & ((1u << lane) - 1u));
    for (uint32_t w = 0; w < warp; w++) pos += sh.cnt[b][w];
    return pos;
}
#endif

}  // namespace b2pt
