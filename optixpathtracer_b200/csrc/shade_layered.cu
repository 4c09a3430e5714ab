// shade_layered.cu — K7d for the production (Philox) generator: the layered BSDF as a walk engine.
//
// Replaces, for rng_mode PHILOX, the serial k_shade<LAYERED> of wavefront.cu (which stays the kernel of the
// bit-exact LCG modes): closest-hit part 2 of the reference (devicePrograms.cu:453-509) with
// GlossyDiffuse::f / Sample_f (PBRT/GlossyDiffuse.h:141-367, :372-514) inside.  Three kernels per bounce:
//
//   K7d-p k_layered_prep    one thread per layered hit: local wo and light direction; hits whose light is visible
//                           AND on wo's side of the shading normal are appended to the f list (f is exactly 0 for
//                           the others)
//   K7d-e k_layered_engine  persistent warps over WALK tasks (5 per f-list hit + 1 sample walk per hit), one
//                           task per lane, refilled at step boundaries; per-lane logic in fast_layered.cuh
//   K7d-f k_layered_finish  one thread per layered hit: entrance reflection + the five partial sums in walk
//                           order (deterministic), NEE add, throughput, next ray, survivor compaction
//
// Built with --use_fast_math (see fast_layered.cuh for what that means for parity).
#include <algorithm>
#include <cstdlib>

#include "bsdf_rows.cuh"
#include "render.h"
#include "wave_common.cuh"

namespace b2pt {

namespace {

// B2PT_WARP_APPEND (build-time A/B): one atomic per warp instead of one per block
#if defined(B2PT_WARP_APPEND)
#define B2PT_APPEND(want, counter) warp_append(want, counter)
#else
#define B2PT_APPEND(want, counter) block_append(want, counter, s_app, it)
#endif

constexpr uint32_t kNoF = 0xffffffffu;
// K7d-p / K7d-f are gather / scatter kernels.  What held them at 3 TB/s was the queue atomic (one same-address atomic per warp
// and trip; B2PT_APPEND below: one per block, finish 77 % of the DRAM peak, profiles/r02_ab_atomics.txt) — not occupancy:
// built for 16 blocks per SM (32 registers, 28-40 bytes of spills) they were 10-30 % SLOWER
// (profiles/r02_ab_prep_finish_occupancy.txt), so the grid stays at 8 blocks per SM with the registers the compiler wants
constexpr int kPrepGrid = 8;

// scratch of one bounce, `capacity` slots per plane (B2ptRenderer::d_layered)
// task record of one hit, 64 bytes, written by K7d-p and read by every walk task of the hit (the five f walks of
// a hit sit in neighbouring lanes and share the loads):
//   q[0] = local wo, alpha      q[1] = albedo, Philox key word 0
//   q[2] = local light direction, Lambda(wl)  (f-list hits only)      q[3] = Philox key word 1
struct alignas(16) HitTask {
    float4 q[4];
};
struct LayeredScratch {
    HitTask* task;                // by slot
    uint32_t* fidx;               // slot -> position in the f list, or kNoF
    uint32_t* flist;              // f list: slots
    float* fpart;                 // [walk][channel][f-list position]: 15 planes
    float *swx, *swy, *swz;       // sample walk: direction (0,0,0 = no sample), by slot
    float *sbx, *sby, *sbz;       // ... and weight f |cos| / pdf
    uint32_t capacity;
};
constexpr int kLayeredScratchWords = 16 + 2 + 15 + 6;

__device__ __forceinline__ fastlayer::WalkKey task_key(const WaveParams& wp, uint32_t id, uint32_t bounce, uint32_t walk) {
    uint32_t pixel, sample;
    path_identity(wp, id, pixel, sample);
    fastlayer::WalkKey k;
    k.k0 = (uint32_t)wp.seed ^ pixel;
    k.k1 = (uint32_t)(wp.seed >> 32) ^ sample;
    k.c1 = bounce | (fastlayer::kStageWalk << 16);
    k.c2 = walk;
    return k;
}

// ---- K7d-p -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWaveThreads)
k_layered_prep(WaveParams wp, PathState st, SurfaceState sf, LayeredScratch ls, const uint32_t* __restrict__ lq,
               const uint32_t* __restrict__ queued, uint32_t* __restrict__ lcounts, uint32_t bounce) {
    const uint32_t n = queued[4 * bounce + kBsdfLayered];
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t iters = (n + stride - 1) / stride;
    __shared__ BlockAppendScratch s_app;
    for (uint32_t it = 0; it < iters; it++) {
        const uint32_t q = it * stride + blockIdx.x * blockDim.x + threadIdx.x;
        bool want_f = false;
        uint32_t i = 0;
        float3 wl = v3(0.0f);
        float3 wo = v3(0.0f);
        float alpha = 0.0f;
        if (q < n) {
            i = lq[q];
            // every load of the hit before the first store: the kernel is bound by memory latency, and a load
            // that follows a store through a pointer the compiler cannot tell apart waits for it
            const float3 ns = v3(sf.nx[i], sf.ny[i], sf.nz[i]);
            const float3 rd = v3(st.dx[i], st.dy[i], st.dz[i]);
            const float rough = sf.rough[i];
            const uint32_t id = st.id[i];
            const float3 albedo = v3(sf.ar[i], sf.ag[i], sf.ab[i]);
            const uint32_t info = sf.info[i];
            const float3 p = v3(sf.px[i], sf.py[i], sf.pz[i]);
            float3 t, b;
            tangent_frame(ns, t, b);
            wo = to_local(t, b, ns, normalize(-rd));
            alpha = rough * rough;                       // Surface.h:26
            const fastlayer::WalkKey key = task_key(wp, id, bounce, 0u);
            if ((info & kInfoHasLight) && (info & kInfoVisible)) {
                const B2ptPointLight& L = wp.lights[info & 0xffu];
                wl = to_local(t, b, ns, normalize(v3(L.position[0], L.position[1], L.position[2]) - p));
                want_f = wo.z * wl.z > 0.0f;
            }
            ls.task[i].q[0] = make_float4(wo.x, wo.y, wo.z, alpha);
            ls.task[i].q[1] = make_float4(albedo.x, albedo.y, albedo.z, __uint_as_float(key.k0));
            ls.task[i].q[3] = make_float4(__uint_as_float(key.k1), 0.0f, 0.0f, 0.0f);
        }
        const uint32_t j = B2PT_APPEND(want_f, lcounts);
        if (want_f) {
            ls.flist[j] = i;
            ls.task[i].q[2] = make_float4(wl.x, wl.y, wl.z, fastlayer::walk_lambda_wi(wo.z < 0.0f ? -wl : wl, alpha));
        }
        if (q < n) ls.fidx[i] = want_f ? j : kNoF;
    }
}

// ---- K7d-e -------------------------------------------------------------------------------------
// REFILL = idle lanes that trigger a task fetch.  A fetched lane starts at step 0 while its neighbours are
// mid-walk; that costs nothing extra because every step has the same shape (fast_layered.cuh).
template <int REFILL, int MINB, int RNG = 10>
__global__ void __launch_bounds__(kWaveThreads, MINB)
k_layered_engine(LayeredScratch ls, const uint32_t* __restrict__ lq, const uint32_t* __restrict__ queued,
                 uint32_t* __restrict__ lcounts, uint32_t bounce) {
    using namespace fastlayer;
    const uint32_t n = queued[4 * bounce + kBsdfLayered];
    const uint32_t n_f = lcounts[0];
    const uint32_t n_tasks = (uint32_t)kWalks * n_f + n;
    uint32_t* cursor = lcounts + 1;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t below = (1u << lane) - 1u;

    Walk k;
    k.step = -1;
    WalkKey key = {0u, 0u, 0u, 0u};
    uint32_t out = 0;            // f walk: f-list position; sample walk: slot
    bool exhausted = false;      // warp-uniform
    // (one atomic on the task cursor per fetch; reserving 128 tasks per atomic and handing them out over the next
    // fetches was measured slower — engine 93.9 -> 98.5 ms per c4 step, profiles/r02_ab_atomics.txt — the cursor is
    // not what the engine waits for)
    for (;;) {
        const uint32_t idle = __ballot_sync(0xffffffffu, k.step < 0);
        if (!exhausted && __popc(idle) >= REFILL) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(cursor, (uint32_t)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, 0);
            exhausted = base + (uint32_t)__popc(idle) >= n_tasks;
            if (k.step < 0) {
                const uint32_t t = base + (uint32_t)__popc(idle & below);
                if (t < n_tasks) {
                    const bool is_f = t < (uint32_t)kWalks * n_f;
                    uint32_t i, walk;
                    if (is_f) {                 // walk-minor: the five walks of a hit sit in neighbouring lanes
                        out = t / (uint32_t)kWalks;
                        walk = t - out * (uint32_t)kWalks;
                        i = ls.flist[out];
                    } else {
                        walk = (uint32_t)kWalks;
                        i = lq[t - (uint32_t)kWalks * n_f];
                        out = i;
                    }
                    const float4* rec = ls.task[i].q;
                    const float4 q0 = __ldg(rec), q1 = __ldg(rec + 1), q3 = __ldg(rec + 3);
                    key.k0 = __float_as_uint(q1.w);
                    key.k1 = __float_as_uint(q3.x);
                    key.c1 = bounce | (kStageWalk << 16);
                    key.c2 = walk;
                    if (is_f) {
                        const float4 q2 = __ldg(rec + 2);
                        walk_begin_f(k, v3(q0.x, q0.y, q0.z), v3(q2.x, q2.y, q2.z), v3(q1.x, q1.y, q1.z), q0.w, q2.w);
                    } else {
                        walk_begin_s(k, v3(q0.x, q0.y, q0.z), v3(q1.x, q1.y, q1.z), q0.w);
                    }
                }
            }
        }
        if (__ballot_sync(0xffffffffu, k.step >= 0) == 0u) {
            if (exhausted) break;
            continue;
        }
        if (k.step >= 0 && !walk_step<RNG>(k, key)) {
            if (k.is_f) {
                float* fp = ls.fpart + (size_t)key.c2 * 3u * ls.capacity + out;
                fp[0] = k.f.x;
                fp[ls.capacity] = k.f.y;
                fp[2 * (size_t)ls.capacity] = k.f.z;
            } else {
                const float3 w = k.ok ? k.w : v3(0.0f);
                ls.swx[out] = w.x; ls.swy[out] = w.y; ls.swz[out] = w.z;
                ls.sbx[out] = k.beta.x; ls.sby[out] = k.beta.y; ls.sbz[out] = k.beta.z;
            }
        }
    }
}

// ---- K7d-f -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWaveThreads)
k_layered_finish(WaveParams wp, PathState st, PathState nx, SurfaceState sf, LayeredScratch ls,
                 const uint32_t* __restrict__ lq, WaveCounts wc, const uint32_t* __restrict__ lcounts, uint32_t bounce,
                 float* __restrict__ wave_out, TraceCounters* __restrict__ counters) {
    const uint32_t n = wc.queued[4 * bounce + kBsdfLayered];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(&counters->walk_f_hits, (unsigned long long)lcounts[0]);
        atomicAdd(&counters->walk_s_hits, (unsigned long long)n);
    }
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t iters = (n + stride - 1) / stride;
    __shared__ BlockAppendScratch s_app;
    for (uint32_t it = 0; it < iters; it++) {
        const uint32_t q = it * stride + blockIdx.x * blockDim.x + threadIdx.x;
        bool alive = false;
        float3 beta = v3(0.0f), radiance = v3(0.0f), next_o = v3(0.0f), next_d = v3(0.0f);
        uint32_t id = 0;
        if (q < n) {
            const uint32_t i = lq[q];
            // all loads by slot first, then the partial sums by f-list position: two dependent rounds of memory
            // latency per hit instead of six (the conditions that used to guard the loads guard only the math now)
            id = st.id[i];
            beta = v3(st.bx[i], st.by[i], st.bz[i]);
            radiance = v3(st.rx[i], st.ry[i], st.rz[i]);
            const float3 p = v3(sf.px[i], sf.py[i], sf.pz[i]);
            const uint32_t j = ls.fidx[i];
            const float4 q0 = ls.task[i].q[0], q2 = ls.task[i].q[2];
            const uint32_t info = sf.info[i];
            const float3 wi = v3(ls.swx[i], ls.swy[i], ls.swz[i]);
            const float3 sb = v3(ls.sbx[i], ls.sby[i], ls.sbz[i]);
            const float3 ns = v3(sf.nx[i], sf.ny[i], sf.nz[i]);
            const float3 ng = v3(sf.gx[i], sf.gy[i], sf.gz[i]);
            const bool has_f = j != kNoF;
            const uint32_t jj = has_f ? j : 0u;      // a hit without f walks reads position 0 and ignores it
            float3 sum = v3(0.0f);
#pragma unroll
            for (int w = 0; w < fastlayer::kWalks; w++) {
                const float* fp = ls.fpart + (size_t)w * 3u * ls.capacity + jj;
                sum += v3(fp[0], fp[ls.capacity], fp[2 * (size_t)ls.capacity]);
            }
            if (has_f) {   // NEE (devicePrograms.cu:453-472) with f = entrance reflection + mean of the five walks
                float3 wo = v3(q0.x, q0.y, q0.z), wl = v3(q2.x, q2.y, q2.z);
                const float cos_l = fabsf(wl.z);
                if (wo.z < 0.0f) { wo = -wo; wl = -wl; }
                const float3 f = v3(fastlayer::entrance_reflection(wo, wl, q0.w)) + sum * (1.0f / (float)fastlayer::kWalks);
                const float3 spectrum = f * cos_l;
                if (nonzero(spectrum)) {
                    const B2ptPointLight& L = wp.lights[info & 0xffu];
                    const float p_light = wp.n_lights == 1 ? 1.0f : 1.0f / (float)wp.n_lights;
                    const float3 dl = p - v3(L.position[0], L.position[1], L.position[2]);
                    const float3 li = v3(L.color[0], L.color[1], L.color[2]) / dot(dl, dl);
                    radiance += (beta * spectrum * li) / p_light;
                }
            }
            if (wi.z != 0.0f) {   // GetNewRayDirection + throughput + next ray (:474-509)
                beta *= sb;
                float3 t, b;
                tangent_frame(ns, t, b);
                float3 offset = 1e-3f * ng;
                if (wi.z < 0.0f) offset = -offset;
                next_o = p + offset;
                next_d = normalize(to_world(t, b, ns, wi));
                alive = path_continues(beta, (int)bounce + 1, wp.max_bounces);
            }
            if (!alive) {
                wave_out[id] = radiance.x;
                wave_out[wp.capacity + id] = radiance.y;
                wave_out[2 * wp.capacity + id] = radiance.z;
            }
        }
        const uint32_t jn = B2PT_APPEND(alive, wc.active + bounce + 1);
        if (alive) {
            nx.ox[jn] = next_o.x; nx.oy[jn] = next_o.y; nx.oz[jn] = next_o.z;
            nx.dx[jn] = next_d.x; nx.dy[jn] = next_d.y; nx.dz[jn] = next_d.z;
            nx.bx[jn] = beta.x; nx.by[jn] = beta.y; nx.bz[jn] = beta.z;
            nx.rx[jn] = radiance.x; nx.ry[jn] = radiance.y; nx.rz[jn] = radiance.z;
            nx.id[jn] = id;
            nx.seed[jn] = 0u;
        }
    }
}

LayeredScratch scratch_view(B2ptRenderer* r, uint32_t capacity) {
    r->d_layered.alloc((size_t)kLayeredScratchWords * capacity);
    float* w = r->d_layered.get();
    auto take = [&]() { float* p = w; w += capacity; return p; };
    LayeredScratch ls;
    ls.task = reinterpret_cast<HitTask*>(w); w += (size_t)16 * capacity;
    ls.fidx = reinterpret_cast<uint32_t*>(take());
    ls.flist = reinterpret_cast<uint32_t*>(take());
    ls.fpart = w; w += (size_t)15 * capacity;
    ls.swx = take(); ls.swy = take(); ls.swz = take();
    ls.sbx = take(); ls.sby = take(); ls.sbz = take();
    ls.capacity = capacity;
    return ls;
}

// ---- rows for the parity hooks (b2pt_bsdf_f / b2pt_bsdf_sample with rng PHILOX) ---------------
__global__ void k_fast_rows_f(int64_t n, const float* albedo, const float* roughness, const float* wo, const float* wi,
                              const uint32_t* seed, float* f_out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) fast_layered_row_f(i, albedo, roughness, wo, wi, seed, f_out);
}
__global__ void k_fast_rows_sample(int64_t n, const float* albedo, const float* roughness, const float* wo,
                                   const uint32_t* seed, uint8_t* ok, float* f_out, float* wi_out, float* pdf_out,
                                   uint8_t* flags_out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) fast_layered_row_sample(i, albedo, roughness, wo, seed, ok, f_out, wi_out, pdf_out, flags_out);
}

}  // namespace

bool layered_engine_enabled() {
    static const bool on = [] { const char* e = getenv("B2PT_LAYERED_ENGINE"); return !e || atoi(e) != 0; }();
    return on;
}

size_t layered_scratch_bytes_per_path() { return (size_t)kLayeredScratchWords * 4; }

// K7d-p, K7d-e, K7d-f of bounce `bounce`, one call each so that the scheduler can time them apart
void launch_layered_prep(B2ptRenderer* r, const WaveParams& wp, const PathState& cur, const WaveCounts& wc,
                         uint32_t* lcounts, uint32_t bounce) {
    const LayeredScratch ls = scratch_view(r, wp.capacity);
    const uint32_t* lq = r->d_queues.get() + (size_t)kBsdfLayered * wp.capacity;
    k_layered_prep<<<r->sm_count * kPrepGrid, kWaveThreads, 0, r->stream>>>(wp, cur, r->surf, ls, lq, wc.queued, lcounts, bounce);
}

void launch_layered_walks(B2ptRenderer* r, const WaveParams& wp, const WaveCounts& wc, uint32_t* lcounts, uint32_t bounce) {
    cudaStream_t s = r->stream;
    const LayeredScratch ls = scratch_view(r, wp.capacity);
    const uint32_t* lq = r->d_queues.get() + (size_t)kBsdfLayered * wp.capacity;
    // A/B (profiles/r02_ab_layered_engine_c4.txt): B2PT_ENGINE_REFILL = idle lanes that trigger a fetch
    // (1 | 4 | 8 | 16 | 32; 16: c4 573 Msamples/s against 507 at 1 and 542 at 32 — a fetch runs with few lanes, a
    // whole idle half-warp does not cost more than that), B2PT_ENGINE_OCC = blocks per SM (4 | 6 | 8)
    static const int refill = [] { const char* e = getenv("B2PT_ENGINE_REFILL"); return e ? atoi(e) : 16; }();
    static const int occ = [] { const char* e = getenv("B2PT_ENGINE_OCC"); return e ? atoi(e) : 6; }();
    // The Philox block of a walk step is inlined in the engine (RNG = 10: engine 94.7 -> 91.8 ms per c4 step against the
    // out-of-line call the other draw sites use).  B2PT_WALK_RNG=7 (A/B, default refill / occupancy only): seven rounds
    // instead of ten, the smallest Crush-resistant count of the Philox paper: 88.6 ms, c4 603 -> 613 Msamples/s
    // (profiles/r02_ab_walk_rng.txt); B2PT_WALK_RNG=0: the out-of-line call.  Ten rounds stay the default.
    static const int walk_rng = [] { const char* e = getenv("B2PT_WALK_RNG"); return e ? atoi(e) : 10; }();
    if (walk_rng == 0 || walk_rng == 7) {
        if (walk_rng == 0) k_layered_engine<16, 6, 0><<<r->sm_count * 6, kWaveThreads, 0, s>>>(ls, lq, wc.queued, lcounts, bounce);
        else k_layered_engine<16, 6, 7><<<r->sm_count * 6, kWaveThreads, 0, s>>>(ls, lq, wc.queued, lcounts, bounce);
        return;
    }
#define B2PT_ENGINE(R, O) k_layered_engine<R, O><<<r->sm_count * O, kWaveThreads, 0, s>>>(ls, lq, wc.queued, lcounts, bounce)
    if (occ >= 8) {
        if (refill >= 32) B2PT_ENGINE(32, 8); else if (refill >= 16) B2PT_ENGINE(16, 8); else if (refill >= 8) B2PT_ENGINE(8, 8);
        else if (refill >= 4) B2PT_ENGINE(4, 8); else B2PT_ENGINE(1, 8);
    } else if (occ >= 6) {
        if (refill >= 32) B2PT_ENGINE(32, 6); else if (refill >= 16) B2PT_ENGINE(16, 6); else if (refill >= 8) B2PT_ENGINE(8, 6);
        else if (refill >= 4) B2PT_ENGINE(4, 6); else B2PT_ENGINE(1, 6);
    } else {
        if (refill >= 32) B2PT_ENGINE(32, 4); else if (refill >= 16) B2PT_ENGINE(16, 4); else if (refill >= 8) B2PT_ENGINE(8, 4);
        else if (refill >= 4) B2PT_ENGINE(4, 4); else B2PT_ENGINE(1, 4);
    }
#undef B2PT_ENGINE
}

void launch_layered_finish(B2ptRenderer* r, const WaveParams& wp, const PathState& cur, const PathState& nxt,
                           const WaveCounts& wc, const uint32_t* lcounts, uint32_t bounce) {
    const LayeredScratch ls = scratch_view(r, wp.capacity);
    const uint32_t* lq = r->d_queues.get() + (size_t)kBsdfLayered * wp.capacity;
    k_layered_finish<<<r->sm_count * kPrepGrid, kWaveThreads, 0, r->stream>>>(wp, cur, nxt, r->surf, ls, lq, wc, lcounts, bounce,
                                                                       r->d_wave_out.get(), r->d_counters.get());
}

void launch_fast_layered_rows_f(int64_t n, const float* albedo, const float* roughness, const float* wo, const float* wi,
                                const uint32_t* seed, float* f_out) {
    k_fast_rows_f<<<(unsigned)((n + 127) / 128), 128>>>(n, albedo, roughness, wo, wi, seed, f_out);
}
void launch_fast_layered_rows_sample(int64_t n, const float* albedo, const float* roughness, const float* wo,
                                     const uint32_t* seed, uint8_t* ok, float* f_out, float* wi_out, float* pdf_out,
                                     uint8_t* flags_out) {
    k_fast_rows_sample<<<(unsigned)((n + 127) / 128), 128>>>(n, albedo, roughness, wo, seed, ok, f_out, wi_out, pdf_out, flags_out);
}

}  // namespace b2pt
