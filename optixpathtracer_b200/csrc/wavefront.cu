// wavefront.cu — the wavefront integrator: kernels K4, K6, K7a-d, K9 and the
// host-side scheduler that replaces the reference's megakernel launch
// (OptixRenderer::Render, OptixRenderer.cpp:617-647 -> optixLaunch(W, H, 1) of
// __raygen__renderFrame, devicePrograms.cu:667-706).
//
// One WAVE = a tile of `np` pixels x `spw` consecutive samples, all paths in
// flight at once in structure-of-arrays buffers.  Per bounce b:
//
//   K5 k_closest   rays of st[b&1][0 .. n_b)            -> hits          (trace_kernels.cu)
//   K6 k_surface   hit -> surface record, class draw, light pick, shadow ray;
//                  misses retire; hits are appended to the queue of their BSDF
//   K8 k_any       shadow rays -> visible bit                            (trace_kernels.cu)
//   K7 k_shade<B>  one launch per BSDF present: NEE add, BSDF sample, throughput,
//                  next ray; survivors are compacted into st[(b+1)&1]
//
// and once per wave K4 k_raygen before, K9 k_accumulate after.  All counts
// (n_b, queue lengths) live in device memory; the host enqueues the whole wave
// without ever synchronising.
//
// Determinism: a path's arithmetic depends only on (pixel, sample index, seed);
// where a path sits in the buffers (atomics order) never enters the result, and
// K9 adds the samples of a pixel in sample order.  Same inputs => same bits.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "render.h"
#include "traverse.cuh"
#include "wave_common.cuh"

// This file is compiled twice.  As wavefront.cu (-fmad=false, host-library sin / cos / exp: the arithmetic that
// reproduces the reference bit for bit) it holds the kernels of all three generators and the scheduler.  Through
// wavefront_relaxed.cu (B2PT_RELAXED_MATH, --use_fast_math) it holds only the PHILOX kernels, in relaxed arithmetic —
// what the production generator runs (B2PT_RELAXED=0 switches back): its images are an independent estimate of the
// reference's anyway, so the bit-parity arithmetic only cost it instructions.
namespace b2pt {

#if !defined(B2PT_RELAXED_MATH)
void enqueue_wave_philox_relaxed(B2ptRenderer* r, const WaveParams& wp, float* dst, bool overwrite);
#endif

namespace {

constexpr int kThreads = kWaveThreads;

// ---- K4 ----------------------------------------------------------------------
template <class Rng>
__global__ void __launch_bounds__(kThreads)
k_raygen(WaveParams wp, PathState st, uint32_t* __restrict__ active0) {
    const uint32_t n = wp.np * wp.spw;
    if (blockIdx.x == 0 && threadIdx.x == 0) *active0 = n;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint32_t pixel, sample;
        path_identity(wp, i, pixel, sample);
        const int y = (int)(pixel / (uint32_t)wp.cam.width), x = (int)(pixel - (uint32_t)y * wp.cam.width);
        float3 o, d;
        camera_ray(wp.cam, x, y, o, d);
        st.ox[i] = o.x; st.oy[i] = o.y; st.oz[i] = o.z;
        st.dx[i] = d.x; st.dy[i] = d.y; st.dz[i] = d.z;
        st.bx[i] = 1.0f; st.by[i] = 1.0f; st.bz[i] = 1.0f;
        st.rx[i] = 0.0f; st.ry[i] = 0.0f; st.rz[i] = 0.0f;
        st.id[i] = i;
        st.seed[i] = RngIO<Rng>::first_state(pixel, sample);
    }
}

// ---- K6 ----------------------------------------------------------------------
template <class Rng, int MINB = 8>
__global__ void __launch_bounds__(kThreads, MINB)
k_surface(WaveParams wp, DeviceScene sc, PathState st, const HitRecord* __restrict__ hits, SurfaceState sf,
          ShadowRays sh, uint32_t* __restrict__ queues, WaveCounts wc, uint32_t bounce, float* __restrict__ wave_out,
          TraceCounters* __restrict__ counters) {
    const uint32_t n = wc.active[bounce];
    uint32_t n_culled = 0;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t iters = (n + stride - 1) / stride;
    const DeviceTexFetch tex{sc.textures};
    uint32_t* qn = wc.queued + 4 * bounce;
    __shared__ BlockAppendScratch s_app[2];
    // block-aggregated append to the light buckets; two copies, used by alternate trips of the loop
    __shared__ uint32_t s_cnt[2][kThreads / 32][kShadowBuckets];
    __shared__ uint32_t s_base[2][kShadowBuckets];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    // the hit of the NEXT trip is fetched a trip ahead: the kernel is a chain of dependent loads (hit -> triangle -> normals,
    // uvs, material -> texels) at half occupancy, and this takes the first link off it
    const uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t tri_next = i0 < n ? hits[i0].tri : kHitMiss;
    for (uint32_t it = 0; it < iters; it++) {
        const uint32_t i = it * stride + i0;
        int kind = -1;
        int bucket = -1;   // light bucket of this slot's shadow ray, -1: none
        bool culled = false;
        const uint32_t tri_now = tri_next;
        if (i + stride < n) tri_next = hits[i + stride].tri;
        if (i < n) {
            HitRecord hit;
            hit.tri = tri_now;
            hit.t = hit.u = hit.v = 0.0f;
            // K5 decided WHICH triangle is hit (trace_deferred.cuh); its t, u, v and back-face flag — the
            // fp64 evaluation of that one triangle — are computed here, where every lane has a hit
            if (hit.tri != kHitMiss)
                hit = finish_hit(sc.tris, hit.tri & ~kHitBackface, v3(st.ox[i], st.oy[i], st.oz[i]),
                                 v3(st.dx[i], st.dy[i], st.dz[i]));
            if (hit.tri == kHitMiss) {   // __miss__radiance (:576-583): the path ends with what it has
                const uint32_t id = st.id[i];
                wave_out[id] = st.rx[i];
                wave_out[wp.capacity + id] = st.ry[i];
                wave_out[2 * wp.capacity + id] = st.rz[i];
                sh.tmax[i] = -1.0f;
            } else {
                RngCtx ctx;
                ctx.seed = wp.seed;
                path_identity(wp, st.id[i], ctx.pixel, ctx.sample);
                ctx.bounce = bounce;
                ctx.lcg_state = st.seed[i];
                Rng rng = RngIO<Rng>::open(ctx, 0);
                SurfacePoint s;
                surface_stage(sc, tex, hit, v3(st.dx[i], st.dy[i], st.dz[i]), wp.lights, wp.n_lights, rng, s);
                st.seed[i] = RngIO<Rng>::close(rng);
                sf.px[i] = s.p.x; sf.py[i] = s.p.y; sf.pz[i] = s.p.z;
                sf.gx[i] = s.ng.x; sf.gy[i] = s.ng.y; sf.gz[i] = s.ng.z;
                sf.nx[i] = s.ns.x; sf.ny[i] = s.ns.y; sf.nz[i] = s.ns.z;
                sf.ar[i] = s.sp.albedo.x; sf.ag[i] = s.sp.albedo.y; sf.ab[i] = s.sp.albedo.z;
                sf.rough[i] = s.sp.roughness;
                sf.info[i] = (s.light >= 0 ? ((uint32_t)s.light | kInfoHasLight) : 0u) |
                             (s.conductor ? kInfoConductor : 0u);
                kind = wp.mode == B2PT_BSDF_AUTO ? (s.conductor ? kBsdfConductor : kBsdfLayered) : wp.mode - 1;
                // Shadow rays whose answer cannot matter are not traced.  The reference traces every one
                // (LightVisibility, :216-241) and then multiplies by f(wo, wl); for the three BSDFs whose f
                // draws no random numbers, f is EXACTLY zero when the material is a perfect mirror /
                // smooth glass (alpha < 1e-3) or, for the opaque ones, when wo and wl lie on opposite
                // sides of the shading normal — the first tests of conductor_f / diffuse_f /
                // dielectric_f (bsdf.cuh), evaluated here with the same operands shade_stage will use
                // (to_local(...).z = dot(n, v), wl = the shadow direction).  Leaving `visible` unset makes
                // shade_stage skip the same zero.  The layered BSDF is excluded: its f consumes draws,
                // so whether it runs must follow the reference.  Output bits are unchanged; the counter
                // `shadow_rays` keeps the reference's count, `shadow_rays_traced` is what K8 really ran.
                bool trace_shadow = s.sh_tmax >= 0.0f;
                if (trace_shadow && (kind != kBsdfLayered || wp.cull_layered)) {
                    const float3 ray_d = v3(st.dx[i], st.dy[i], st.dz[i]);
                    const float wo_z = dot(s.ns, normalize(-ray_d)), wl_z = dot(s.ns, s.sh_d);
                    const bool smooth = effectively_smooth(roughness_to_alpha(s.sp.roughness));
                    const bool opposite = !(wo_z * wl_z > 0.0f);   // !same_hemisphere(wo, wl)
                    // layered (walk engine only, where skipping f does not shift the random stream of the sample): every
                    // walk of f ends at the opaque base when the light is on the other side — exactly 0 (fast_layered.cuh)
                    const bool zero_f = kind == kBsdfConductor ? (smooth || opposite)
                                      : kind == kBsdfDiffuse   ? opposite
                                      : kind == kBsdfLayered   ? opposite
                                                               : smooth;   // dielectric
                    if (zero_f && wp.cull_nee) {
                        trace_shadow = false;
                        culled = true;
                    }
                }
                sh.ox[i] = s.sh_o.x; sh.oy[i] = s.sh_o.y; sh.oz[i] = s.sh_o.z;
                sh.dx[i] = s.sh_d.x; sh.dy[i] = s.sh_d.y; sh.dz[i] = s.sh_d.z;
                sh.tmax[i] = trace_shadow ? s.sh_tmax : -1.0f;
                if (trace_shadow) bucket = s.light & (kShadowBuckets - 1);
            }
        }
        if (sh.order) {
            // K8's work list, ordered by light (ShadowRays, wavefront.cuh).  The append is aggregated over the BLOCK:
            // the rays that the block's 128 consecutive slots send to one light land next to each other in that
            // light's bucket — a warp of K8 then traces rays from neighbouring paths to the same light — and the
            // four bucket counters see one atomic per block and trip instead of one per warp (per-warp appends made
            // K6 twice as slow and left K8's warps as mixed as before: profiles/r02_ab_sort_shadow.txt).
            uint32_t* cnt = s_cnt[it & 1u][0];
            uint32_t mine = 0;
#pragma unroll
            for (int k = 0; k < kShadowBuckets; k++) {
                const uint32_t m = __ballot_sync(0xffffffffu, bucket == k);
                if (lane == 0) cnt[warp * kShadowBuckets + k] = (uint32_t)__popc(m);
                if (bucket == k) mine = (uint32_t)__popc(m & ((1u << lane) - 1u));
            }
            __syncthreads();
            if (threadIdx.x < kShadowBuckets) {
                uint32_t tot = 0;
#pragma unroll
                for (int w = 0; w < kThreads / 32; w++) tot += cnt[w * kShadowBuckets + threadIdx.x];
                s_base[it & 1u][threadIdx.x] = tot ? atomicAdd(sh.bucket_n + threadIdx.x, tot) : 0u;
            }
            __syncthreads();
            if (bucket >= 0) {
                uint32_t pos = s_base[it & 1u][bucket] + mine;
                for (uint32_t w = 0; w < warp; w++) pos += cnt[w * kShadowBuckets + bucket];
                sh.order[(size_t)bucket * sh.capacity + pos] = i;
            }
        }
        if (wp.mode == B2PT_BSDF_AUTO) {   // two queues can be non-empty: conductor and layered
#pragma unroll
            for (int k = 0; k < 4; k += 3) {
                const uint32_t pos = block_append(kind == k, qn + k, s_app[k ? 1 : 0], it);
                if (kind == k) queues[(size_t)k * wp.capacity + pos] = i;
            }
        } else {
            const int k = wp.mode - 1;
            const uint32_t pos = block_append(kind == k, qn + k, s_app[0], it);
            if (kind == k) queues[(size_t)k * wp.capacity + pos] = i;
        }
        n_culled += culled ? 1u : 0u;
    }
    for (int off = 16; off > 0; off >>= 1) n_culled += __shfl_xor_sync(0xffffffffu, n_culled, off);
    if ((threadIdx.x & 31) == 0 && n_culled) atomicAdd(&counters->shadow_culled, (unsigned long long)n_culled);
}

// ---- K7p: order each shade queue by light visibility ----------------------------------------
// A hit whose shadow ray reached the light evaluates f(wo, wl) before it samples the BSDF; one whose ray was
// blocked (or never sent) only samples.  For the layered BSDF that is six random walks against one, and a warp
// with both kinds runs at the pace of the first: (5 v + 1) / 6 of its lanes busy for a visible fraction v.  This
// pass rewrites each queue with the visible hits at the front and the others at the back (warp-aggregated
// atomics on two cursors per queue), so all warps but one per queue are uniform.  Positions inside a queue never
// enter a result (every output is addressed by the path's id), so the image bits do not change.  Only the
// layered queue is reordered: c4 74 -> 89, c5 393 -> 493 Msamples/s.  For the three light BSDFs f costs about as
// much as a sample, and reading the path state in scrambled order costs more than the idle lanes did
// (c2 -5 %, c1 -9 % with all four queues reordered; B2PT_PARTITION=2 does that, 0 none).
__global__ void __launch_bounds__(256)
k_partition_queues(const uint32_t* __restrict__ queues, uint32_t* __restrict__ queues2, const uint32_t* __restrict__ info,
                   WaveCounts wc, uint32_t bounce, uint32_t capacity, uint32_t* __restrict__ cursors /* [4][2], zeroed */,
                   uint32_t first_kind) {
    const uint32_t kind = blockIdx.y + first_kind;
    const uint32_t n = wc.queued[4 * bounce + kind];
    const uint32_t* q = queues + (size_t)kind * capacity;
    uint32_t* q2 = queues2 + (size_t)kind * capacity;
    uint32_t* cur = cursors + 2 * kind;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t iters = (n + stride - 1) / stride;
    for (uint32_t it = 0; it < iters; it++) {
        const uint32_t k = it * stride + blockIdx.x * blockDim.x + threadIdx.x;
        const bool valid = k < n;
        const uint32_t i = valid ? q[k] : 0u;
        const bool vis = valid && (info[i] & kInfoVisible) != 0;
        const uint32_t mv = __ballot_sync(0xffffffffu, vis), mn = __ballot_sync(0xffffffffu, valid && !vis);
        uint32_t bv = 0, bn = 0;
        if (lane == 0) {
            if (mv) bv = atomicAdd(cur, (uint32_t)__popc(mv));
            if (mn) bn = atomicAdd(cur + 1, (uint32_t)__popc(mn));
        }
        bv = __shfl_sync(0xffffffffu, bv, 0);
        bn = __shfl_sync(0xffffffffu, bn, 0);
        const uint32_t below = (1u << lane) - 1u;
        if (vis) q2[bv + (uint32_t)__popc(mv & below)] = i;
        else if (valid) q2[n - 1u - (bn + (uint32_t)__popc(mn & below))] = i;
    }
}

// ---- K7a-d -------------------------------------------------------------------
// One launch per BSDF: every thread of the kernel runs the same BSDF code.
template <int KIND, class Rng, int MINB = 6>
__global__ void __launch_bounds__(kThreads, MINB)
k_shade(WaveParams wp, PathState st, PathState nx, SurfaceState sf, const uint32_t* __restrict__ queue,
        WaveCounts wc, uint32_t bounce, float* __restrict__ wave_out) {
    const uint32_t n = wc.queued[4 * bounce + KIND];
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t iters = (n + stride - 1) / stride;
    __shared__ BlockAppendScratch s_app;
    for (uint32_t it = 0; it < iters; it++) {
        const uint32_t q = it * stride + blockIdx.x * blockDim.x + threadIdx.x;
        bool alive = false;
        PathVertex pv;
        uint32_t id = 0, seed = 0;
        if (q < n) {
            const uint32_t i = queue[q];
            SurfacePoint s;
            s.p = v3(sf.px[i], sf.py[i], sf.pz[i]);
            s.ng = v3(sf.gx[i], sf.gy[i], sf.gz[i]);
            s.ns = v3(sf.nx[i], sf.ny[i], sf.nz[i]);
            s.sp.albedo = v3(sf.ar[i], sf.ag[i], sf.ab[i]);
            s.sp.roughness = sf.rough[i];
            const uint32_t info = sf.info[i];
            s.light = (info & kInfoHasLight) ? (int)(info & 0xffu) : -1;
            s.conductor = (info & kInfoConductor) != 0;
            pv.beta = v3(st.bx[i], st.by[i], st.bz[i]);
            pv.radiance = v3(st.rx[i], st.ry[i], st.rz[i]);
            pv.next_o = pv.next_d = v3(0.0f);
            id = st.id[i];
            RngCtx ctx;
            ctx.seed = wp.seed;
            path_identity(wp, id, ctx.pixel, ctx.sample);
            ctx.bounce = bounce;
            ctx.lcg_state = st.seed[i];
            Rng rng = RngIO<Rng>::open(ctx, 1);
            alive = shade_stage<KIND>(s, (info & kInfoVisible) != 0, v3(st.dx[i], st.dy[i], st.dz[i]), wp.lights,
                                      wp.n_lights, rng, pv);
            seed = RngIO<Rng>::close(rng);
            alive = alive && path_continues(pv.beta, (int)bounce + 1, wp.max_bounces);
            if (!alive) {   // radiance is final (:658)
                wave_out[id] = pv.radiance.x;
                wave_out[wp.capacity + id] = pv.radiance.y;
                wave_out[2 * wp.capacity + id] = pv.radiance.z;
            }
        }
        const uint32_t j = block_append(alive, wc.active + bounce + 1, s_app, it);
        if (alive) {
            nx.ox[j] = pv.next_o.x; nx.oy[j] = pv.next_o.y; nx.oz[j] = pv.next_o.z;
            nx.dx[j] = pv.next_d.x; nx.dy[j] = pv.next_d.y; nx.dz[j] = pv.next_d.z;
            nx.bx[j] = pv.beta.x; nx.by[j] = pv.beta.y; nx.bz[j] = pv.beta.z;
            nx.rx[j] = pv.radiance.x; nx.ry[j] = pv.radiance.y; nx.rz[j] = pv.radiance.z;
            nx.id[j] = id;
            nx.seed[j] = seed;
        }
    }
}

// ---- K9 ----------------------------------------------------------------------
// Replaces the GL accumulate pass (AddPathtracedFrame.frag:13-25 driven by
// OptixView.cpp:212-255): the samples of a pixel are added in sample order to
// the fp32 sum (accumulate) or the single sample overwrites the frame (Render,
// devicePrograms.cu:704-705).  Non-finite samples are zeroed and counted when
// accumulating (the reference only repairs them when writing the EXR,
// WriteImage.cpp:52-55); Render passes them through untouched.
__global__ void __launch_bounds__(256)
k_accumulate(WaveParams wp, const float* __restrict__ wave_out, float* __restrict__ dst, int overwrite,
             TraceCounters* counters) {
    uint32_t bad = 0;
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < wp.np; p += gridDim.x * blockDim.x) {
        float* o = dst + 3 * (size_t)(wp.p0 + p);
        if (overwrite) {
            o[0] = wave_out[p];
            o[1] = wave_out[wp.capacity + p];
            o[2] = wave_out[2 * wp.capacity + p];
            continue;
        }
        float r = o[0], g = o[1], b = o[2];
        for (uint32_t s = 0; s < wp.spw; s++) {
            const uint32_t id = s * wp.np + p;
            const float sr = wave_out[id], sg = wave_out[wp.capacity + id], sb = wave_out[2 * wp.capacity + id];
            if (isfinite(sr) && isfinite(sg) && isfinite(sb)) {
                r += sr; g += sg; b += sb;
            } else {
                bad++;
            }
        }
        o[0] = r; o[1] = g; o[2] = b;
    }
    if (bad) atomicAdd(&counters->nan_samples, (unsigned long long)bad);
}

// ray census of a finished wave (device-side, no host round trip)
__global__ void k_tally(WaveCounts wc, int max_bounces, int has_lights, TraceCounters* counters) {
    unsigned long long rr = 0, sr = 0;
    for (int b = threadIdx.x; b < max_bounces; b += blockDim.x) {
        rr += wc.active[b];
        if (has_lights) sr += (unsigned long long)wc.queued[4 * b] + wc.queued[4 * b + 1] + wc.queued[4 * b + 2] + wc.queued[4 * b + 3];
    }
    for (int off = 16; off > 0; off >>= 1) {
        rr += __shfl_xor_sync(0xffffffffu, rr, off);
        sr += __shfl_xor_sync(0xffffffffu, sr, off);
    }
    if (threadIdx.x == 0) {
        atomicAdd(&counters->radiance_rays, rr);
        atomicAdd(&counters->shadow_rays, sr);
    }
}

__global__ void __launch_bounds__(256)
k_resolve(size_t n, const float* __restrict__ sum, float inv_spp, float* __restrict__ out) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = sum[i] * inv_spp;
}

// ---- host side -------------------------------------------------------------------
struct WaveBuffers {
    PathState st[2];
    SurfaceState sf;
    ShadowRays sh;
};

// lay the SoA arrays out in r->d_words (one allocation, `capacity` slots each)
void layout_wave(B2ptRenderer* r, uint32_t capacity) {
    if (r->wave_capacity == capacity && r->d_words.get()) return;
    const size_t words = (size_t)(2 * kPathStateWords + kSurfaceStateWords + kShadowWords) * capacity;
    r->d_words.alloc(words);
    r->d_hits.alloc(capacity);
    r->d_queues.alloc((size_t)4 * capacity);
    r->d_queues2.alloc((size_t)4 * capacity);
    r->d_wave_out.alloc((size_t)3 * capacity);
    r->d_counts.alloc((size_t)kCountsWords);   // wave_common.cuh: active, queued, K7p cursors, walk-engine counters
    float* w = r->d_words.get();
    auto take = [&]() { float* p = w; w += capacity; return p; };
    for (int k = 0; k < 2; k++) {
        PathState& s = r->st[k];
        s.ox = take(); s.oy = take(); s.oz = take();
        s.dx = take(); s.dy = take(); s.dz = take();
        s.bx = take(); s.by = take(); s.bz = take();
        s.rx = take(); s.ry = take(); s.rz = take();
        s.id = reinterpret_cast<uint32_t*>(take());
        s.seed = reinterpret_cast<uint32_t*>(take());
    }
    SurfaceState& f = r->surf;
    f.px = take(); f.py = take(); f.pz = take();
    f.gx = take(); f.gy = take(); f.gz = take();
    f.nx = take(); f.ny = take(); f.nz = take();
    f.ar = take(); f.ag = take(); f.ab = take();
    f.rough = take();
    f.info = reinterpret_cast<uint32_t*>(take());
    ShadowRays& h = r->shadow;
    h.ox = take(); h.oy = take(); h.oz = take();
    h.dx = take(); h.dy = take(); h.dz = take();
    h.tmax = take();
    r->wave_capacity = capacity;
}

template <class Rng>
void enqueue_wave(B2ptRenderer* r, const WaveParams& wp, float* dst, bool overwrite) {
    cudaStream_t s = r->stream;
    const int grid = r->sm_count * 8;
    const DeviceScene sc = r->scene.view();
    WaveCounts wc;
    wc.active = r->d_counts.get() + kCountsActive;
    wc.queued = r->d_counts.get() + kCountsQueued;
    LaunchCfg lc;
    lc.stream = s;
    lc.grid = r->sm_count * trace_grid_blocks_per_sm();
    lc.stats = r->cfg.collect_stats != 0;
    lc.launch_counter = &r->kernel_launches;
    auto count = [&](int k = 1) { r->kernel_launches += (uint64_t)k; };
    // with kernel timing on, an event closes every kernel; elapsed(i, i+1) belongs to class tag[i]
    auto mark = [&](int cls, bool is_launch = true) {
        if (is_launch) r->launches_kernel_cls[cls]++;
        if (!r->kernel_timing) return;
        if (r->timing_used == r->timing_events.size()) {
            cudaEvent_t e;
            B2PT_CUDA(cudaEventCreate(&e));
            r->timing_events.push_back(e);
        }
        B2PT_CUDA(cudaEventRecord(r->timing_events[r->timing_used++], s));
        r->timing_tags.push_back(cls);
    };
    mark(kClsOther, false);   // opening event of the wave

    r->d_counts.zero(s);
    if (wp.max_bounces == 0) r->d_wave_out.zero(s);
    k_raygen<Rng><<<grid, kThreads, 0, s>>>(wp, r->st[0], wc.active);
    count();
    mark(kClsOther);
    const bool kinds[4] = {wp.mode == B2PT_BSDF_AUTO || wp.mode == B2PT_BSDF_CONDUCTOR, wp.mode == B2PT_BSDF_DIFFUSE,
                           wp.mode == B2PT_BSDF_DIELECTRIC, wp.mode == B2PT_BSDF_AUTO || wp.mode == B2PT_BSDF_LAYERED};
    // the layered BSDF with the production generator runs as a walk engine (shade_layered.cu); the LCG modes — whose
    // draws must come in the reference's order — run the serial kernel below, visible hits first (K7p)
    const bool engine = std::is_same<Rng, PhiloxRng>::value && kinds[3] && layered_engine_enabled();
    // A/B, off by default: c2 +1.9 %, c5 +1.9 %, c4 -0.6 % (profiles/r02_ab_sort_shadow.txt)
    static const bool sort_shadow = [] { const char* e = getenv("B2PT_SORT_SHADOW"); return e && atoi(e) != 0; }();
    for (int b = 0; b < wp.max_bounces; b++) {
        const PathState& cur = r->st[b & 1];
        const PathState& nxt = r->st[(b + 1) & 1];
        launch_closest(lc, sc, cur, wc.active + b, 0.0f, 100.0f, r->d_hits.get(), r->d_counters.get(), /*finish=*/false);   // tmin 0, tmax 100 (:652)
        mark(kClsClosest);
        // B2PT_SORT_SHADOW=1: K8 takes its rays from a light-ordered list instead of scanning the slots.  The list lives in d_queues2,
        // which the serial layered kernel's partition pass (K7p) only overwrites after K8 has read it.
        ShadowRays shadow = r->shadow;
        if (sort_shadow && wp.n_lights > 0) {
            shadow.order = r->d_queues2.get();
            shadow.bucket_n = r->d_counts.get() + kCountsShadow + kShadowBuckets * b;
            shadow.capacity = wp.capacity;
        }
        k_surface<Rng><<<grid, kThreads, 0, s>>>(wp, sc, cur, r->d_hits.get(), r->surf, shadow, r->d_queues.get(),
                                                 wc, (uint32_t)b, r->d_wave_out.get(), r->d_counters.get());
        count();
        mark(kClsSurface);
        if (wp.n_lights > 0) {
            launch_any(lc, sc, shadow, wc.active + b, 0.0f, r->surf.info, nullptr, r->d_counters.get());
            mark(kClsAny);
        }
        const uint32_t* q = r->d_queues.get();
        const size_t cap = wp.capacity;
        // 8 resident blocks per SM (64 registers, a few spilled words) beat 6 (80 registers) for all shade kernels:
        // they are bound by instruction issue / fetch, not by registers (profiles/r01_ab_shade_occupancy.txt)
        const int g3 = r->sm_count * 16;
        if (kinds[0]) { k_shade<kBsdfConductor, Rng, 8><<<g3, kThreads, 0, s>>>(wp, cur, nxt, r->surf, q + 0 * cap, wc, (uint32_t)b, r->d_wave_out.get()); count(); mark(kClsShade); }
        if (kinds[1]) { k_shade<kBsdfDiffuse, Rng, 8><<<g3, kThreads, 0, s>>>(wp, cur, nxt, r->surf, q + 1 * cap, wc, (uint32_t)b, r->d_wave_out.get()); count(); mark(kClsShade); }
        if (kinds[2]) { k_shade<kBsdfDielectric, Rng, 8><<<g3, kThreads, 0, s>>>(wp, cur, nxt, r->surf, q + 2 * cap, wc, (uint32_t)b, r->d_wave_out.get()); count(); mark(kClsShade); }
        if (kinds[3] && engine) {
            uint32_t* lcounts = r->d_counts.get() + kCountsLayered + 4 * b;
            launch_layered_prep(r, wp, cur, wc, lcounts, (uint32_t)b);
            count(); mark(kClsShade);
            launch_layered_walks(r, wp, wc, lcounts, (uint32_t)b);
            count(); mark(kClsEngine);
            launch_layered_finish(r, wp, cur, nxt, wc, lcounts, (uint32_t)b);
            count(); mark(kClsShade);
        } else if (kinds[3]) {
            const uint32_t* q3 = q;
            if (wp.n_lights > 0) {   // K7p: visible hits first in the layered queue
                k_partition_queues<<<dim3((unsigned)(r->sm_count * 4), 1), 256, 0, s>>>(
                    r->d_queues.get(), r->d_queues2.get(), r->surf.info, wc, (uint32_t)b, wp.capacity,
                    r->d_counts.get() + kCountsPartition + 8 * b, 3u);
                count();
                mark(kClsShade);
                q3 = r->d_queues2.get();
            }
            k_shade<kBsdfLayered, Rng, 8><<<g3, kThreads, 0, s>>>(wp, cur, nxt, r->surf, q3 + 3 * cap, wc, (uint32_t)b, r->d_wave_out.get());
            count();
            mark(kClsShade);
        }
    }
    k_accumulate<<<r->sm_count * 4, 256, 0, s>>>(wp, r->d_wave_out.get(), dst, overwrite ? 1 : 0, r->d_counters.get());
    k_tally<<<1, 32, 0, s>>>(wc, wp.max_bounces, wp.n_lights > 0 ? 1 : 0, r->d_counters.get());
    count(2);
    mark(kClsOther);
    B2PT_CUDA(cudaGetLastError());
}

}  // namespace

#if defined(B2PT_RELAXED_MATH)
void enqueue_wave_philox_relaxed(B2ptRenderer* r, const WaveParams& wp, float* dst, bool overwrite) {
    enqueue_wave<PhiloxRng>(r, wp, dst, overwrite);
}
#else
void render_frames(B2ptRenderer* r, int spp, bool accumulate) {
    B2PT_REQUIRE(r->width > 0 && r->height > 0, B2PT_E_STATE, "resize first");
    B2PT_REQUIRE(r->have_camera, B2PT_E_STATE, "set_camera first");
    B2PT_REQUIRE(accumulate || spp == 1, B2PT_E_INVALID, "Render traces exactly one sample per pixel");
    const uint64_t n_pixels = (uint64_t)r->width * r->height;
    int log2cap = r->cfg.wave_paths_log2;
    if (log2cap <= 0) {
        // Default: the whole call in ONE wave when it fits (up to 2^26 paths = 16 GB of wave state; a B200 has
        // 180 GB).  The late bounces of a wave are latency-bound launches — a few thousand rays, each launch
        // lasting as long as its slowest ray, ~0.1 ms whatever the count (profiles/r01_launches_wide.csv) — so
        // their cost is per WAVE, not per path: c2 374 -> 452 -> 480 Msamples/s for 2^22 -> 2^24 -> 2^26.
        // Grow-only: a later, smaller call (Render after RenderAccumulate) keeps the buffers.
        const uint64_t want = n_pixels * (uint64_t)std::max(spp, 1);
        log2cap = 20;
        while ((1ull << log2cap) < want && log2cap < 26) log2cap++;
        const size_t per_path = (size_t)(2 * kPathStateWords + kSurfaceStateWords + kShadowWords) * 4 + sizeof(HitRecord) + 32 + 12 +
                                layered_scratch_bytes_per_path();
        const size_t held = r->d_hits.size();                      // paths the buffers hold now
        if (((size_t)1 << log2cap) > held) {                       // growing: never more than half of the free memory
            size_t free_b = 0, total_b = 0;                        // (asked only now: cudaMemGetInfo can take tens of ms
            B2PT_CUDA(cudaMemGetInfo(&free_b, &total_b));          //  when a monitoring tool holds the driver)
            const size_t have = held * per_path;                   // released before a larger allocation
            while (log2cap > 20 && (per_path << log2cap) > (free_b + have) / 2) log2cap--;
        }
        while (log2cap < 27 && ((size_t)1 << log2cap) < held) log2cap++;
    }
    B2PT_REQUIRE(log2cap >= 10 && log2cap <= 27, B2PT_E_INVALID, "wave_paths_log2 must be in [10, 27]");
    const uint32_t capacity = 1u << log2cap;
    layout_wave(r, capacity);

    WaveParams wp;
    wp.cam.position = v3(r->cam_pos[0], r->cam_pos[1], r->cam_pos[2]);
    memcpy(wp.cam.inv_view, r->cam_inv_view, 64);
    memcpy(wp.cam.inv_proj, r->cam_inv_proj, 64);
    wp.cam.width = r->width;
    wp.cam.height = r->height;
    wp.lights = r->d_lights.get();
    wp.n_lights = (int32_t)r->lights.size();
    wp.max_bounces = r->cfg.max_bounces;
    wp.mode = r->cfg.bsdf_mode;
    wp.seed = r->cfg.seed;
    wp.capacity = capacity;
    static const int cull_nee = [] { const char* e = getenv("B2PT_CULL_NEE"); return e ? atoi(e) : 1; }();
    // the Philox kernels in relaxed arithmetic (wavefront_relaxed.cu) unless B2PT_RELAXED=0: c2 +10 %, c3 +9 %, c5 +2 %
    // (the light shade kernels run twice as fast; profiles/r02_ab_relaxed.txt), gated by tests/test_relaxed_gpu.py
    static const bool relaxed = [] { const char* e = getenv("B2PT_RELAXED"); return !e || atoi(e) != 0; }();
    wp.cull_nee = cull_nee;
    wp.cull_layered = cull_nee && r->cfg.rng_mode == B2PT_RNG_PHILOX && layered_engine_enabled();
    float* dst = accumulate ? r->d_accum.get() : r->d_frame.get();

    r->timing_used = 0;
    r->timing_tags.clear();
    B2PT_CUDA(cudaEventRecord(r->ev0, r->stream));
    const uint32_t tile = (uint32_t)std::min<uint64_t>(n_pixels, capacity);
    const uint32_t spw_max = std::max<uint32_t>(1u, capacity / tile);
    for (int s0 = 0; s0 < spp; s0 += (int)spw_max) {
        wp.spw = (uint32_t)std::min<int>((int)spw_max, spp - s0);
        wp.frame0 = r->frame_id + 1u + (uint32_t)s0;   // frame.id is pre-incremented (OptixRenderer.cpp:623)
        for (uint64_t p0 = 0; p0 < n_pixels; p0 += tile) {
            wp.p0 = (uint32_t)p0;
            wp.np = (uint32_t)std::min<uint64_t>(tile, n_pixels - p0);
            switch (r->cfg.rng_mode) {
                case B2PT_RNG_LCG_DEVICE: enqueue_wave<LcgRng<false>>(r, wp, dst, !accumulate); break;
                case B2PT_RNG_LCG_HOST: enqueue_wave<LcgRng<true>>(r, wp, dst, !accumulate); break;
                default:
                    if (relaxed) enqueue_wave_philox_relaxed(r, wp, dst, !accumulate);
                    else enqueue_wave<PhiloxRng>(r, wp, dst, !accumulate);
                    break;
            }
        }
    }
    B2PT_CUDA(cudaEventRecord(r->ev1, r->stream));
    B2PT_CUDA(cudaEventSynchronize(r->ev1));
    float ms = 0.0f;
    B2PT_CUDA(cudaEventElapsedTime(&ms, r->ev0, r->ev1));
    r->device_ms_render += ms;
    for (size_t i = 1; i < r->timing_used; i++) {
        // tag[i] is the class of the kernel that ended at event i; the opening mark of a wave is
        // tagged kClsOther and collects the memset before raygen
        float dt = 0.0f;
        B2PT_CUDA(cudaEventElapsedTime(&dt, r->timing_events[i - 1], r->timing_events[i]));
        r->ms_kernel[r->timing_tags[i]] += dt;
    }
    r->frame_id += (uint32_t)spp;
    r->samples += n_pixels * (uint64_t)spp;
    if (accumulate) r->spp_accumulated += (uint32_t)spp;
}

void resolve_mean(B2ptRenderer* r) {
    const size_t n = (size_t)r->width * r->height * 3;
    const float inv = r->spp_accumulated ? 1.0f / (float)r->spp_accumulated : 0.0f;
    k_resolve<<<r->sm_count * 4, 256, 0, r->stream>>>(n, r->d_accum.get(), inv, r->d_frame.get());
    r->kernel_launches++;
    B2PT_CUDA(cudaGetLastError());
}
#endif  // !B2PT_RELAXED_MATH

}  // namespace b2pt
