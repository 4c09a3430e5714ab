// bvh_build.cu — scene flattening + GPU LBVH build (kernels K1-K3).
//
// Stands in for OptixRenderer::BuildAccel (Renderer/OptiX/OptixRenderer.cpp:
// 306-456: per-mesh uploads :328-340, triangle build inputs with a FLOAT12
// pre-transform :342-373, optixAccelBuild :416, optixAccelCompact :440),
// CreateTextures (:562-612) and the material half of BuildSBT (:486-560).
//
//   k_flatten      one thread per triangle: world-space vertices/normals
//                  (glm's mat4*vec4 association, devicePrograms.cu:77-81),
//                  centroid bounds by block reduce + ordered-int atomics
//   k_morton       63-bit Morton key of the box centre
//   radix sort     stable LSD, 8 x 8-bit passes: k_rs_hist / k_rs_scan /
//                  k_rs_scatter (warp match_any ranking; hand-written, no CUB)
//   k_karras       Karras 2012 topology, one thread per internal node
//   k_refit        bottom-up AABB refit with arrival counters
//   k_emit_*       leaf-order triangle / normal / uv records and 64-byte nodes
#include <algorithm>
#include <cstdlib>
#include <vector>

#include "scene.cuh"

namespace b2pt {

namespace {

constexpr int kThreads = 256;
inline int blocks_for(int64_t n, int threads = kThreads) { return (int)((n + threads - 1) / threads); }

// ---- ordered-int encoding so float min/max can use integer atomics --------
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(uint32_t u) {
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

struct BoundsOrd {
    uint32_t lo[3], hi[3];      // centroid bounds
    uint32_t vlo[3], vhi[3];    // vertex bounds
};

__global__ void k_init_bounds(BoundsOrd* b) {
    for (int i = 0; i < 3; i++) {
        b->lo[i] = b->vlo[i] = 0xffffffffu;
        b->hi[i] = b->vhi[i] = 0u;
    }
}

// K1a: world-space triangle soup + bounds.
__global__ void __launch_bounds__(kThreads)
k_flatten(int n_tris, const int32_t* __restrict__ indices, const uint32_t* __restrict__ tri_mesh,
          const MeshInfo* __restrict__ meshes, const float* __restrict__ positions,
          float3* __restrict__ wv, BoundsOrd* bounds) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    float3 clo = v3(INFINITY), chi = v3(-INFINITY), vlo = v3(INFINITY), vhi = v3(-INFINITY);
    if (g < n_tris) {
        const MeshInfo& mi = meshes[tri_mesh[g]];
        float3 p[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const int vi = indices[3 * g + k] + mi.vtx_offset;
            const float4 w = mat4_mul(mi.model, positions[3 * vi], positions[3 * vi + 1], positions[3 * vi + 2], 1.0f);
            p[k] = v3(w.x, w.y, w.z);
            wv[3 * g + k] = p[k];
        }
        const Aabb bb = aabb_of_tri(p[0], p[1], p[2]);
        vlo = bb.lo;
        vhi = bb.hi;
        clo = chi = (bb.lo + bb.hi) * 0.5f;
    }
    // warp reduce, then one atomic per warp and component
    float r[12] = {clo.x, clo.y, clo.z, vlo.x, vlo.y, vlo.z, chi.x, chi.y, chi.z, vhi.x, vhi.y, vhi.z};
#pragma unroll
    for (int i = 0; i < 12; i++) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const float o = __shfl_xor_sync(0xffffffffu, r[i], off);
            r[i] = i < 6 ? fminf(r[i], o) : fmaxf(r[i], o);
        }
    }
    if ((threadIdx.x & 31) == 0) {
        for (int i = 0; i < 3; i++) {
            if (r[i] <= r[6 + i]) {  // at least one valid triangle in the warp
                atomicMin(&bounds->lo[i], f2ord(r[i]));
                atomicMax(&bounds->hi[i], f2ord(r[6 + i]));
                atomicMin(&bounds->vlo[i], f2ord(r[3 + i]));
                atomicMax(&bounds->vhi[i], f2ord(r[9 + i]));
            }
        }
    }
}

// K1b: Morton keys.
__global__ void __launch_bounds__(kThreads)
k_morton(int n_tris, const float3* __restrict__ wv, const BoundsOrd* __restrict__ bounds,
         uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_tris) return;
    const float3 lo = v3(ord2f(bounds->lo[0]), ord2f(bounds->lo[1]), ord2f(bounds->lo[2]));
    const float3 hi = v3(ord2f(bounds->hi[0]), ord2f(bounds->hi[1]), ord2f(bounds->hi[2]));
    const float3 ext = hi - lo;
    const float3 inv = v3(ext.x > 0.0f ? 1.0f / ext.x : 0.0f, ext.y > 0.0f ? 1.0f / ext.y : 0.0f,
                          ext.z > 0.0f ? 1.0f / ext.z : 0.0f);
    const Aabb bb = aabb_of_tri(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2]);
    keys[g] = morton63((bb.lo + bb.hi) * 0.5f, lo, inv);
    vals[g] = (uint32_t)g;
}

// ---- K2: stable LSD radix sort, 8 bits per pass ---------------------------
constexpr int kRsWarps = 8;
constexpr int kRsItems = 8;                                // 32-key steps per warp
constexpr int kRsTile = kRsWarps * 32 * kRsItems;          // 2048 keys per block
constexpr uint32_t kRsNoDigit = 0xffffu;

__global__ void __launch_bounds__(kRsWarps * 32)
k_rs_hist(int n, const uint64_t* __restrict__ keys, int shift, int n_tiles, uint32_t* __restrict__ hist) {
    __shared__ uint32_t cnt[256];
    cnt[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * kRsTile;
    for (int i = threadIdx.x; i < kRsTile; i += blockDim.x) {
        const int k = base + i;
        if (k < n) atomicAdd(&cnt[(uint32_t)(keys[k] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * n_tiles + blockIdx.x] = cnt[threadIdx.x];  // digit-major
}

// exclusive scan of `m` counters by ONE block (m = 256 * n_tiles, small)
__global__ void __launch_bounds__(1024)
k_rs_scan(int m, uint32_t* __restrict__ data) {
    __shared__ uint32_t part[1024];
    const int per = (m + 1023) / 1024;
    const int lo = threadIdx.x * per, hi = min(lo + per, m);
    uint32_t s = 0;
    for (int i = lo; i < hi; i++) s += data[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {  // Hillis-Steele inclusive scan
        uint32_t v = threadIdx.x >= off ? part[threadIdx.x - off] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = threadIdx.x ? part[threadIdx.x - 1] : 0;
    for (int i = lo; i < hi; i++) {
        const uint32_t v = data[i];
        data[i] = run;
        run += v;
    }
}

__global__ void __launch_bounds__(kRsWarps * 32)
k_rs_scatter(int n, const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
             uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int shift, int n_tiles,
             const uint32_t* __restrict__ hist) {
    __shared__ uint32_t cnt[kRsWarps][256];   // per-warp digit counts -> running offsets
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < kRsWarps * 256; i += blockDim.x) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const int wbase = blockIdx.x * kRsTile + warp * (32 * kRsItems);
    uint64_t key[kRsItems];
    uint32_t dig[kRsItems];
#pragma unroll
    for (int s = 0; s < kRsItems; s++) {
        const int k = wbase + s * 32 + lane;
        key[s] = k < n ? keys_in[k] : 0;
        dig[s] = k < n ? ((uint32_t)(key[s] >> shift) & 255u) : kRsNoDigit;
        const uint32_t peers = __match_any_sync(0xffffffffu, dig[s]);
        if (dig[s] != kRsNoDigit && lane == __ffs(peers) - 1) cnt[warp][dig[s]] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    {   // thread d: turn per-warp counts of digit d into exclusive offsets + global base
        const int d = threadIdx.x;
        uint32_t run = hist[d * n_tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kRsWarps; w++) {
            const uint32_t c = cnt[w][d];
            cnt[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int s = 0; s < kRsItems; s++) {
        const int k = wbase + s * 32 + lane;
        const uint32_t peers = __match_any_sync(0xffffffffu, dig[s]);
        uint32_t pos = 0;
        if (dig[s] != kRsNoDigit) pos = cnt[warp][dig[s]] + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
        if (dig[s] != kRsNoDigit && lane == __ffs(peers) - 1) cnt[warp][dig[s]] += __popc(peers);
        __syncwarp();
        if (dig[s] != kRsNoDigit) {
            keys_out[pos] = key[s];
            vals_out[pos] = vals_in[k];
        }
    }
}

// ---- K3a: Karras topology --------------------------------------------------
// node ids: internal i in [0, n-2]; leaf k is encoded as ~k.  parent[] is
// indexed by internal id for internal nodes and by (n-1)+k for leaves.
__global__ void __launch_bounds__(kThreads)
k_karras(int n, const uint64_t* __restrict__ keys, int2* __restrict__ children, int* __restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int first, last, split;
    karras_node(keys, n, i, first, last, split);
    const int left = (split == first) ? ~split : split;
    const int right = (split + 1 == last) ? ~(split + 1) : split + 1;
    children[i] = make_int2(left, right);
    parent[left >= 0 ? left : (n - 1) + ~left] = i;
    parent[right >= 0 ? right : (n - 1) + ~right] = i;
    if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ Aabb leaf_box(const float3* __restrict__ wv, uint32_t prim, float pad) {
    return aabb_pad(aabb_of_tri(wv[3 * prim], wv[3 * prim + 1], wv[3 * prim + 2]), pad);
}

// ---- K3b: bottom-up refit --------------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_refit(int n, const float3* __restrict__ wv, const uint32_t* __restrict__ order,
        const int2* __restrict__ children, const int* __restrict__ parent, int* __restrict__ arrivals,
        volatile float* node_box /* 6 floats per internal node */, float pad) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int cur = parent[(n - 1) + k];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&arrivals[cur], 1) == 0) return;  // sibling subtree not finished yet
        __threadfence();
        const int2 ch = children[cur];
        Aabb b[2];
#pragma unroll
        for (int s = 0; s < 2; s++) {
            const int c = s ? ch.y : ch.x;
            if (c < 0) {
                b[s] = leaf_box(wv, order[~c], pad);
            } else {
                volatile float* p = node_box + 6 * (size_t)c;
                b[s].lo = v3(p[0], p[1], p[2]);
                b[s].hi = v3(p[3], p[4], p[5]);
            }
        }
        const Aabb u = aabb_union(b[0], b[1]);
        volatile float* q = node_box + 6 * (size_t)cur;
        q[0] = u.lo.x; q[1] = u.lo.y; q[2] = u.lo.z;
        q[3] = u.hi.x; q[4] = u.hi.y; q[5] = u.hi.z;
        cur = parent[cur];
    }
}

// ---- K3c: emit traversal layout -------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_emit_nodes(int n, const float3* __restrict__ wv, const uint32_t* __restrict__ order,
             const int2* __restrict__ children, const float* __restrict__ node_box, float pad,
             BvhNode2* __restrict__ nodes) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int2 ch = children[i];
    Aabb b[2];
#pragma unroll
    for (int s = 0; s < 2; s++) {
        const int c = s ? ch.y : ch.x;
        if (c < 0) {
            b[s] = leaf_box(wv, order[~c], pad);
        } else {
            const float* p = node_box + 6 * (size_t)c;
            b[s].lo = v3(p[0], p[1], p[2]);
            b[s].hi = v3(p[3], p[4], p[5]);
        }
    }
    BvhNode2 nd;
    nd.c0xy = make_float4(b[0].lo.x, b[0].hi.x, b[0].lo.y, b[0].hi.y);
    nd.c1xy = make_float4(b[1].lo.x, b[1].hi.x, b[1].lo.y, b[1].hi.y);
    nd.cz = make_float4(b[0].lo.z, b[0].hi.z, b[1].lo.z, b[1].hi.z);
    nd.child0 = ch.x;
    nd.child1 = ch.y;
    nd.pad0 = nd.pad1 = 0;
    nodes[i] = nd;
}

// A/B switch: B2PT_COLLAPSE=fixed opens both children of every node (two binary levels per wide node)
static int collapse_greedy() {
    static const int v = [] { const char* e = getenv("B2PT_COLLAPSE"); return !(e && e[0] == 'f') ? 1 : 0; }();
    return v;
}

// triangles per leaf of the wide tree: B2ptConfig::bvh_leaf_size, 0 = default; B2PT_LEAF overrides (A/B)
static int wide_leaf_size(int cfg_leaf_size) {
    static const int env = [] { const char* e = getenv("B2PT_LEAF"); return e ? atoi(e) : 0; }();
    const int v = env > 0 ? env : (cfg_leaf_size > 0 ? cfg_leaf_size : kDefaultLeafSize);
    return v < 1 ? 1 : (v > 8 ? 8 : v);
}

// ---- K3d: collapse binary nodes into 4-wide nodes ---------------------------------------
// Wide node i describes the subtree of binary node i.  Its slots start as the two children of i; while
// a slot is free, the internal slot with the largest surface area is replaced by its own two children
// (the usual surface-area-greedy collapse).  greedy = 0 keeps the fixed two-level collapse: both
// children are opened, whatever their size.  Either way every descent crosses at least one binary level
// and pushes at most three entries (kWideStackSize, trace_wide.cuh).
struct Slot4 {
    float lx, hx, ly, hy, lz, hz;
    int32_t child;
};
__device__ __forceinline__ Slot4 slot_of(const BvhNode2& nd, int s) {
    const float4 xy = s ? nd.c1xy : nd.c0xy;
    Slot4 r;
    r.lx = xy.x; r.hx = xy.y; r.ly = xy.z; r.hy = xy.w;
    r.lz = s ? nd.cz.z : nd.cz.x;
    r.hz = s ? nd.cz.w : nd.cz.y;
    r.child = s ? nd.child1 : nd.child0;
    return r;
}
__device__ __forceinline__ float slot_area(const Slot4& b) {
    const float dx = b.hx - b.lx, dy = b.hy - b.ly, dz = b.hz - b.lz;
    return dx * dy + dy * dz + dz * dx;
}

// BvhNode4P: the float nearest to v on the outward side (down: <= v, else >= v) whose low mantissa byte is `payload`.
__device__ __forceinline__ float embed_byte(float v, uint32_t payload, bool down) {
    const uint32_t u = __float_as_uint(v);
    const uint32_t sign = u & 0x80000000u, m = u & 0x7fffffffu;
    const bool grow = (sign != 0u) == down;   // the outward side is the larger magnitude
    uint32_t w = (m & ~0xffu) | payload;
    if (grow) {
        if (w < m) w += 0x100u;
        return __uint_as_float(sign | w);
    }
    if (w > m) {
        if (m < 0x100u) return __uint_as_float((sign ^ 0x80000000u) | payload);   // cross zero: a denormal of the other sign
        w -= 0x100u;
    }
    return __uint_as_float(sign | w);
}
__device__ __forceinline__ BvhNode4P pack4(const Slot4 sl[4]) {
    float lx[4], hx[4], ly[4], hy[4], lz[4], hz[4];
    for (int k = 0; k < 4; k++) {
        const bool used = sl[k].child != kEmptyChild;
        const uint32_t c = (uint32_t)sl[k].child;
        const float big = 3.0e38f;
        lx[k] = embed_byte(used ? sl[k].lx : big, c & 0xffu, true);
        hx[k] = embed_byte(used ? sl[k].hx : big, (c >> 8) & 0xffu, false);
        ly[k] = embed_byte(used ? sl[k].ly : big, (c >> 16) & 0xffu, true);
        hy[k] = embed_byte(used ? sl[k].hy : big, (c >> 24) & 0xffu, false);
        lz[k] = used ? sl[k].lz : big;
        hz[k] = used ? sl[k].hz : big;
    }
    BvhNode4P p;
    p.lox = make_float4(lx[0], lx[1], lx[2], lx[3]); p.hix = make_float4(hx[0], hx[1], hx[2], hx[3]);
    p.loy = make_float4(ly[0], ly[1], ly[2], ly[3]); p.hiy = make_float4(hy[0], hy[1], hy[2], hy[3]);
    p.loz = make_float4(lz[0], lz[1], lz[2], lz[3]); p.hiz = make_float4(hz[0], hz[1], hz[2], hz[3]);
    return p;
}

// Multi-triangle leaves (B2ptConfig::bvh_leaf_size = L > 1): a slot whose binary subtree holds at most L triangles
// becomes ONE leaf that names the run of triangle records — a Karras subtree covers a contiguous range of the
// Morton-sorted triangles, and the records are stored in that order — child = ~(first | (count - 1) << 28)
// (kLeafCountShift, bvh.cuh; count <= 8, first < 2^28).  A ray that enters the slot's box tests the whole run
// instead of visiting one more node.  The size of a subtree is found by walking to its leftmost and rightmost
// leaf, at most L - 1 levels each way.
__device__ __forceinline__ void leafify(Slot4& b, const BvhNode2* __restrict__ nodes, int leaf_size) {
    if (leaf_size <= 1 || b.child < 0) return;
    int lo = b.child, hi = b.child;
    for (int s = 0; s < leaf_size && lo >= 0; s++) lo = nodes[lo].child0;
    for (int s = 0; s < leaf_size && hi >= 0; s++) hi = nodes[hi].child1;
    if (lo >= 0 || hi >= 0) return;
    const int first = ~lo, count = ~hi - first + 1;
    if (count <= leaf_size && first < (1 << kLeafCountShift)) b.child = ~(first | ((count - 1) << kLeafCountShift));
}

__global__ void __launch_bounds__(kThreads)
k_collapse4(int n_nodes, const BvhNode2* __restrict__ nodes, BvhNode4* __restrict__ wide, int greedy, int leaf_size) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const BvhNode2 nd = nodes[i];
    Slot4 sl[4];
    int k = 0;
#pragma unroll
    for (int s = 0; s < 2; s++) {
        Slot4 b = slot_of(nd, s);
        leafify(b, nodes, leaf_size);
        if (b.lx <= b.hx) sl[k++] = b;             // drops the never-hit slot of a one-triangle scene
    }
    if (greedy) {
        while (k < 4) {
            int pick = -1;
            float best = -1.0f;
            for (int j = 0; j < k; j++) {
                if (sl[j].child < 0) continue;
                const float a = slot_area(sl[j]);
                if (a > best) { best = a; pick = j; }
            }
            if (pick < 0) break;
            const BvhNode2 g = nodes[sl[pick].child];
            sl[pick] = slot_of(g, 0);
            sl[k] = slot_of(g, 1);
            leafify(sl[pick], nodes, leaf_size);
            leafify(sl[k], nodes, leaf_size);
            k++;
        }
    } else {
        const int k0 = k;
        for (int j = 0; j < k0; j++) {
            if (sl[j].child < 0) continue;
            const BvhNode2 g = nodes[sl[j].child];
            sl[j] = slot_of(g, 0);
            sl[k] = slot_of(g, 1);
            leafify(sl[j], nodes, leaf_size);
            leafify(sl[k], nodes, leaf_size);
            k++;
        }
    }
    // empty slot: lo = hi = +inf fails the slab test for every direction (an INVERTED box would pass it:
    // min(+inf, -inf) = -inf <= max = +inf)
    for (; k < 4; k++) {
        sl[k].lx = sl[k].hx = sl[k].ly = sl[k].hy = sl[k].lz = sl[k].hz = INFINITY;
        sl[k].child = kEmptyChild;
    }
    BvhNode4 w;
    w.lox = make_float4(sl[0].lx, sl[1].lx, sl[2].lx, sl[3].lx); w.hix = make_float4(sl[0].hx, sl[1].hx, sl[2].hx, sl[3].hx);
    w.loy = make_float4(sl[0].ly, sl[1].ly, sl[2].ly, sl[3].ly); w.hiy = make_float4(sl[0].hy, sl[1].hy, sl[2].hy, sl[3].hy);
    w.loz = make_float4(sl[0].lz, sl[1].lz, sl[2].lz, sl[3].lz); w.hiz = make_float4(sl[0].hz, sl[1].hz, sl[2].hz, sl[3].hz);
#pragma unroll
    for (int j = 0; j < 4; j++) { w.child[j] = sl[j].child; w.pad[j] = 0; }
    wide[i] = w;
}


// ---- K3e: compaction of the wide nodes ("optixAccelCompact", OptixRenderer.cpp:440) ---------------------------
// K3d writes one wide node per BINARY node, but a descent from the root crosses one to three binary levels per
// step, so only a part of them can ever be reached (README scene: 58 k of 143 k).  The reachable ones are
// numbered breadth-first — the levels every ray walks through end up next to each other at the front of the
// array, whatever the triangle count — the child indices are rewritten, and only the format the traversal
// kernels read is kept.  The breadth-first sweep also measures the depth of the tree, which the build checks
// against the traversal stacks (kStackSize, kWideStackSize) instead of trusting it.
constexpr int kMaxLevels = 128;   // > 63 + 32 (Morton-63 prefix + index tie-break levels, bvh.cuh)

// one level of the sweep: children of the nodes in `fin` are appended to `fout` (order irrelevant: sorted below),
// one atomic per warp
template <bool WIDE>
__global__ void __launch_bounds__(kThreads)
k_bfs_level(const BvhNode4* __restrict__ wide, const BvhNode2* __restrict__ nodes, const int* __restrict__ fin,
            int* __restrict__ fout, uint32_t* __restrict__ counts, int L, int* __restrict__ level) {
    const uint32_t n_in = counts[L];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t iters = (n_in + stride - 1) / stride;
    for (uint32_t it = 0; it < iters; it++) {
        const uint32_t j = it * stride + blockIdx.x * blockDim.x + threadIdx.x;
        int c[4] = {kEmptyChild, kEmptyChild, kEmptyChild, kEmptyChild};
        if (j < n_in) {
            const int i = fin[j];
            if (WIDE) {
                const int4 ch = *reinterpret_cast<const int4*>(wide[i].child);
                c[0] = ch.x; c[1] = ch.y; c[2] = ch.z; c[3] = ch.w;
            } else {
                c[0] = nodes[i].child0; c[1] = nodes[i].child1;
            }
        }
        uint32_t mine = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) mine += (c[k] >= 0 && c[k] != kEmptyChild) ? 1u : 0u;
        uint32_t incl = mine;   // inclusive scan over the warp
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= (uint32_t)d) incl += up;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        uint32_t base = 0;
        if (lane == 31 && total) base = atomicAdd(&counts[L + 1], total);
        base = __shfl_sync(0xffffffffu, base, 31);
        uint32_t at = base + incl - mine;
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (c[k] >= 0 && c[k] != kEmptyChild) {
                if (level) level[c[k]] = L + 1;
                fout[at++] = c[k];
            }
    }
}
__global__ void k_bfs_init(int* fin, uint32_t* counts, int* level) {
    fin[0] = 0;
    counts[0] = 1u;
    if (level) level[0] = 0;
}
// sort key = (level, old index): breadth-first, and inside a level in the order of the Morton-sorted binary nodes
__global__ void __launch_bounds__(kThreads)
k_compact_keys(int n_nodes, const int* __restrict__ level, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    keys[i] = level[i] >= 0 ? (((uint64_t)(uint32_t)level[i] << 32) | (uint32_t)i) : ((uint64_t)0xffull << 32 | (uint32_t)i);
    vals[i] = (uint32_t)i;
}
__global__ void __launch_bounds__(kThreads)
k_compact_remap(int n_reach, const uint32_t* __restrict__ order, int* __restrict__ remap) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n_reach) remap[order[j]] = j;
}
__global__ void __launch_bounds__(kThreads)
k_compact_emit(int n_reach, const uint32_t* __restrict__ order, const int* __restrict__ remap,
               const BvhNode4* __restrict__ wide, BvhNode4* __restrict__ out4, BvhNode4P* __restrict__ outp) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_reach) return;
    const BvhNode4 w = wide[order[j]];
    Slot4 sl[4];
    const float lx[4] = {w.lox.x, w.lox.y, w.lox.z, w.lox.w}, hx[4] = {w.hix.x, w.hix.y, w.hix.z, w.hix.w};
    const float ly[4] = {w.loy.x, w.loy.y, w.loy.z, w.loy.w}, hy[4] = {w.hiy.x, w.hiy.y, w.hiy.z, w.hiy.w};
    const float lz[4] = {w.loz.x, w.loz.y, w.loz.z, w.loz.w}, hz[4] = {w.hiz.x, w.hiz.y, w.hiz.z, w.hiz.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
        sl[k].lx = lx[k]; sl[k].hx = hx[k]; sl[k].ly = ly[k]; sl[k].hy = hy[k]; sl[k].lz = lz[k]; sl[k].hz = hz[k];
        const int c = w.child[k];
        sl[k].child = (c >= 0 && c != kEmptyChild) ? remap[c] : c;
    }
    if (outp) outp[j] = pack4(sl);
    if (out4) {
        BvhNode4 o = w;
#pragma unroll
        for (int k = 0; k < 4; k++) o.child[k] = sl[k].child;
        out4[j] = o;
    }
}

// single-triangle scene: one node, second child never hit
__global__ void k_emit_single(const float3* __restrict__ wv, float pad, BvhNode2* __restrict__ nodes) {
    const Aabb b = leaf_box(wv, 0, pad);
    BvhNode2 nd;
    nd.c0xy = make_float4(b.lo.x, b.hi.x, b.lo.y, b.hi.y);
    nd.c1xy = make_float4(INFINITY, -INFINITY, INFINITY, -INFINITY);
    nd.cz = make_float4(b.lo.z, b.hi.z, INFINITY, -INFINITY);
    nd.child0 = ~0;
    nd.child1 = ~0;
    nd.pad0 = nd.pad1 = 0;
    nodes[0] = nd;
}

__global__ void __launch_bounds__(kThreads)
k_emit_tris(int n, const float3* __restrict__ wv, const float* __restrict__ normals,
            const uint32_t* __restrict__ order, const uint32_t* __restrict__ tri_mesh,
            const MeshInfo* __restrict__ meshes, const int32_t* __restrict__ indices,
            const float* __restrict__ texcoords, TriRecord* __restrict__ tris,
            NrmRecord* __restrict__ nrms, UvRecord* __restrict__ uvs) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint32_t g = order[s];
    const uint32_t m = tri_mesh[g];
    const MeshInfo& mi = meshes[m];
    tris[s] = tri_pack(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2], g, m, mi.cutout ? 1u : 0u);
    const int i0 = indices[3 * g] + mi.vtx_offset, i1 = indices[3 * g + 1] + mi.vtx_offset,
              i2 = indices[3 * g + 2] + mi.vtx_offset;
    float3 n0 = v3(0.0f), n1 = v3(0.0f), n2 = v3(0.0f);
    if (mi.has_normals) {
        n0 = v3(normals[3 * i0], normals[3 * i0 + 1], normals[3 * i0 + 2]);
        n1 = v3(normals[3 * i1], normals[3 * i1 + 1], normals[3 * i1 + 2]);
        n2 = v3(normals[3 * i2], normals[3 * i2 + 1], normals[3 * i2 + 2]);
    }
    NrmRecord nr;
    nr.a = make_float4(n0.x, n0.y, n0.z, n1.x);
    nr.b = make_float4(n1.y, n1.z, n2.x, n2.y);
    nr.c = make_float4(n2.z, __uint_as_float((uint32_t)mi.has_normals), 0.0f, 0.0f);
    nrms[s] = nr;
    if (uvs) {
        UvRecord ur;
        ur.t0 = ur.t1 = ur.t2 = make_float2(0.0f, 0.0f);
        if (mi.has_uv) {
            ur.t0 = make_float2(texcoords[2 * i0], texcoords[2 * i0 + 1]);
            ur.t1 = make_float2(texcoords[2 * i1], texcoords[2 * i1 + 1]);
            ur.t2 = make_float2(texcoords[2 * i2], texcoords[2 * i2 + 1]);
        }
        uvs[s] = ur;
    }
}

}  // namespace

DeviceScene SceneBuffers::view() const {
    DeviceScene s;
    s.nodes = nodes.get();
    s.nodes4 = nodes4.get();
    s.nodes4p = nodes4p.get();
    s.tris = tris.get();
    s.nrms = nrms.get();
    s.uvs = uvs.get();
    s.materials = materials.get();
    s.textures = textures.get();
    s.n_nodes = n_nodes;
    s.n_wide_nodes = n_wide_nodes;
    s.n_tris = n_tris;
    s.n_meshes = n_meshes;
    s.n_textures = n_textures;
    s.any_alpha_texture = any_alpha_texture;
    s.max_abs = fmaxf(fmaxf(fmaxf(fabsf(bounds_lo.x), fabsf(bounds_lo.y)), fabsf(bounds_lo.z)),
                      fmaxf(fmaxf(fabsf(bounds_hi.x), fabsf(bounds_hi.y)), fabsf(bounds_hi.z)));
    return s;
}

int64_t SceneBuffers::device_bytes() const {
    return (int64_t)(nodes.bytes() + nodes4.bytes() + nodes4p.bytes() + tris.bytes() + nrms.bytes() + uvs.bytes() + materials.bytes() +
                     textures.bytes() + tri_order.bytes());
}

void SceneBuffers::free_textures() {
    for (auto t : texture_objects) cudaDestroyTextureObject(t);
    for (auto a : texture_arrays) cudaFreeArray(a);
    texture_objects.clear();
    texture_arrays.clear();
}

// CreateTextures (OptixRenderer.cpp:562-612): uchar4 array, wrap, bilinear,
// normalised float reads, normalised coordinates, sRGB = 0.
static void create_textures(const B2ptSceneDesc& desc, SceneBuffers& out, cudaStream_t stream) {
    out.free_textures();
    out.n_textures = desc.n_textures;
    for (int i = 0; i < desc.n_textures; i++) {
        const B2ptTexture& t = desc.textures[i];
        B2PT_REQUIRE(t.rgba && t.width > 0 && t.height > 0, B2PT_E_INVALID, "texture with no pixels");
        cudaChannelFormatDesc cd = cudaCreateChannelDesc<uchar4>();
        cudaArray_t arr = nullptr;
        B2PT_CUDA(cudaMallocArray(&arr, &cd, t.width, t.height));
        out.texture_arrays.push_back(arr);
        const size_t pitch = (size_t)t.width * 4;
        B2PT_CUDA(cudaMemcpy2DToArray(arr, 0, 0, t.rgba, pitch, pitch, t.height, cudaMemcpyHostToDevice));
        cudaResourceDesc rd = {};
        rd.resType = cudaResourceTypeArray;
        rd.res.array.array = arr;
        cudaTextureDesc td = {};
        td.addressMode[0] = td.addressMode[1] = cudaAddressModeWrap;
        td.filterMode = cudaFilterModeLinear;
        td.readMode = cudaReadModeNormalizedFloat;
        td.normalizedCoords = 1;
        td.maxAnisotropy = 1;
        td.maxMipmapLevelClamp = 99;
        td.minMipmapLevelClamp = 0;
        td.mipmapFilterMode = cudaFilterModePoint;
        td.borderColor[0] = 1.0f;
        td.sRGB = 0;
        cudaTextureObject_t obj = 0;
        B2PT_CUDA(cudaCreateTextureObject(&obj, &rd, &td, nullptr));
        out.texture_objects.push_back(obj);
    }
    out.textures.upload(out.texture_objects.data(), out.texture_objects.size(), stream);
}

void build_scene(const B2ptSceneDesc& desc, SceneBuffers& out, cudaStream_t stream, int leaf_size) {
    B2PT_REQUIRE(desc.n_meshes >= 0 && (desc.n_meshes == 0 || desc.meshes), B2PT_E_INVALID, "bad mesh array");
    B2PT_REQUIRE(desc.n_textures >= 0 && (desc.n_textures == 0 || desc.textures), B2PT_E_INVALID,
                 "bad texture array");

    // ---- host: concatenate the per-mesh arrays (memcpy only) --------------
    int64_t n_vtx = 0, n_tri = 0;
    bool any_uv = false;
    for (int m = 0; m < desc.n_meshes; m++) {
        const B2ptMesh& me = desc.meshes[m];
        B2PT_REQUIRE(me.n_vertices >= 0 && me.n_triangles >= 0, B2PT_E_INVALID, "negative mesh size");
        B2PT_REQUIRE(me.n_triangles == 0 || (me.positions && me.indices), B2PT_E_INVALID,
                     "mesh without positions/indices");
        n_vtx += me.n_vertices;
        n_tri += me.n_triangles;
        any_uv |= me.texcoords != nullptr;
    }
    B2PT_REQUIRE(n_tri < (1ll << 30), B2PT_E_UNSUPPORTED, "more than 2^30 triangles");
    std::vector<float> pos((size_t)n_vtx * 3), nrm((size_t)n_vtx * 3, 0.0f), uv(any_uv ? (size_t)n_vtx * 2 : 0, 0.0f);
    std::vector<int32_t> idx((size_t)n_tri * 3);
    std::vector<uint32_t> tri_mesh((size_t)n_tri);
    std::vector<MeshInfo> infos((size_t)desc.n_meshes);
    std::vector<MaterialRecord> mats((size_t)desc.n_meshes);
    std::vector<char> texture_opaque((size_t)std::max(desc.n_textures, 0), 0);
    for (int t = 0; t < desc.n_textures; t++) {
        const B2ptTexture& tx = desc.textures[t];
        if (!tx.rgba || tx.width <= 0 || tx.height <= 0) continue;
        bool opaque = true;
        const size_t n_px = (size_t)tx.width * (size_t)tx.height;
        for (size_t i = 0; i < n_px && opaque; i++) opaque = tx.rgba[4 * i + 3] == 255;
        texture_opaque[(size_t)t] = opaque ? 1 : 0;
    }
    int64_t vo = 0, to = 0;
    int any_alpha = 0;
    for (int m = 0; m < desc.n_meshes; m++) {
        const B2ptMesh& me = desc.meshes[m];
        MeshInfo& mi = infos[m];
        memcpy(mi.model, me.model, sizeof mi.model);
        mi.vtx_offset = (int32_t)vo;
        mi.tri_offset = (int32_t)to;
        mi.has_normals = me.normals != nullptr;
        mi.has_uv = me.texcoords != nullptr;
        mi.cutout = 0;
        if (me.n_vertices) {
            memcpy(&pos[vo * 3], me.positions, (size_t)me.n_vertices * 12);
            if (me.normals) memcpy(&nrm[vo * 3], me.normals, (size_t)me.n_vertices * 12);
            if (me.texcoords) memcpy(&uv[vo * 2], me.texcoords, (size_t)me.n_vertices * 8);
        }
        for (int64_t t = 0; t < me.n_triangles; t++) {
            for (int k = 0; k < 3; k++) {
                const int32_t v = me.indices[3 * t + k];
                B2PT_REQUIRE(v >= 0 && v < me.n_vertices, B2PT_E_INVALID, "vertex index out of range");
                idx[(to + t) * 3 + k] = v;
            }
            tri_mesh[to + t] = (uint32_t)m;
        }
        MaterialRecord& mr = mats[m];
        mr.albedo[0] = me.albedo[0]; mr.albedo[1] = me.albedo[1]; mr.albedo[2] = me.albedo[2];
        mr.metallic = me.metallic;
        mr.roughness = me.roughness;
        auto tex_ok = [&](int32_t t) { return (t >= 0 && t < desc.n_textures) ? t : -1; };
        mr.tex_albedo = tex_ok(me.tex_albedo);
        mr.tex_normal = tex_ok(me.tex_normal);
        mr.tex_metal_rough = tex_ok(me.tex_metal_rough);
        memcpy(mr.mcol, me.model, sizeof mr.mcol);
        // The any-hit predicate (alpha.cuh) can only reject where the albedo texture's alpha is below 0.9 after the sRGB
        // decode.  A texture whose texels all carry alpha 255 filters to exactly 1.0 (oracle/texunit.h: 8-bit weights that
        // sum to 256 over 16-bit unorm texels) and accepts every hit: its meshes — the usual case of an opaque albedo
        // map — trace like untextured ones.  Meshes without texcoords read uv = (0,0).
        if (mr.tex_albedo >= 0 && !texture_opaque[(size_t)mr.tex_albedo]) {
            mi.cutout = 1;
            any_alpha = 1;
        }
        vo += me.n_vertices;
        to += me.n_triangles;
    }

    const int n = (int)n_tri;
    out.n_tris = n;
    out.n_meshes = desc.n_meshes;
    out.any_alpha_texture = any_alpha;
    out.materials.upload(mats.data(), mats.size(), stream);
    create_textures(desc, out, stream);
    out.n_nodes = n == 0 ? 0 : (n == 1 ? 1 : n - 1);
    out.nodes.alloc((size_t)out.n_nodes);
    out.nodes4.release();
    out.nodes4p.release();
    out.n_wide_nodes = 0;
    out.depth = out.wide_depth = 0;
    out.tris.alloc((size_t)n);
    out.nrms.alloc((size_t)n);
    out.uvs.alloc(any_uv ? (size_t)n : 0);
    out.tri_order.alloc((size_t)n);
    if (n == 0) {
        B2PT_CUDA(cudaStreamSynchronize(stream));
        return;
    }

    // ---- device: temporaries ----------------------------------------------
    DeviceBuffer<float> d_pos, d_nrm, d_uv;
    DeviceBuffer<int32_t> d_idx;
    DeviceBuffer<uint32_t> d_tri_mesh, d_vals[2], d_hist;
    DeviceBuffer<MeshInfo> d_infos;
    DeviceBuffer<float3> d_wv;
    DeviceBuffer<uint64_t> d_keys[2];
    DeviceBuffer<BoundsOrd> d_bounds;
    DeviceBuffer<int2> d_children;
    DeviceBuffer<int> d_parent, d_arrivals;
    DeviceBuffer<float> d_node_box;
    d_pos.upload(pos.data(), pos.size(), stream);
    d_nrm.upload(nrm.data(), nrm.size(), stream);
    if (any_uv) d_uv.upload(uv.data(), uv.size(), stream);
    d_idx.upload(idx.data(), idx.size(), stream);
    d_tri_mesh.upload(tri_mesh.data(), tri_mesh.size(), stream);
    d_infos.upload(infos.data(), infos.size(), stream);
    d_wv.alloc((size_t)n * 3);
    d_keys[0].alloc(n); d_keys[1].alloc(n);
    d_vals[0].alloc(n); d_vals[1].alloc(n);
    d_bounds.alloc(1);
    const int n_tiles = (n + kRsTile - 1) / kRsTile;
    d_hist.alloc((size_t)256 * n_tiles);
    d_children.alloc(n > 1 ? n - 1 : 1);
    d_parent.alloc((size_t)2 * n);
    d_arrivals.alloc(n > 1 ? n - 1 : 1);
    d_node_box.alloc((size_t)6 * (n > 1 ? n - 1 : 1));

    cudaEvent_t ev0, ev1;
    B2PT_CUDA(cudaEventCreate(&ev0));
    B2PT_CUDA(cudaEventCreate(&ev1));
    B2PT_CUDA(cudaEventRecord(ev0, stream));

    k_init_bounds<<<1, 1, 0, stream>>>(d_bounds.get());
    k_flatten<<<blocks_for(n), kThreads, 0, stream>>>(n, d_idx.get(), d_tri_mesh.get(), d_infos.get(),
                                                       d_pos.get(), d_wv.get(), d_bounds.get());
    B2PT_CUDA(cudaGetLastError());
    BoundsOrd hb;
    d_bounds.download(&hb, 1, stream);
    B2PT_CUDA(cudaStreamSynchronize(stream));
    out.bounds_lo = v3(ord2f(hb.vlo[0]), ord2f(hb.vlo[1]), ord2f(hb.vlo[2]));
    out.bounds_hi = v3(ord2f(hb.vhi[0]), ord2f(hb.vhi[1]), ord2f(hb.vhi[2]));
    float maxabs = 0.0f;
    for (int i = 0; i < 3; i++) {
        maxabs = fmaxf(maxabs, fabsf(ord2f(hb.vlo[i])));
        maxabs = fmaxf(maxabs, fabsf(ord2f(hb.vhi[i])));
    }
    B2PT_REQUIRE(std::isfinite(maxabs), B2PT_E_INVALID, "non-finite vertex positions");
    // outward box padding: 2^-17 of the largest coordinate (see traverse.cuh, trace_wide.cuh)
    const float pad = fmaxf(maxabs, 1e-30f) * 7.6293945e-6f;

    k_morton<<<blocks_for(n), kThreads, 0, stream>>>(n, d_wv.get(), d_bounds.get(), d_keys[0].get(),
                                                      d_vals[0].get());
    int cur = 0;
    for (int pass = 0; pass < 8; pass++) {
        const int shift = pass * 8;
        k_rs_hist<<<n_tiles, kRsWarps * 32, 0, stream>>>(n, d_keys[cur].get(), shift, n_tiles, d_hist.get());
        k_rs_scan<<<1, 1024, 0, stream>>>(256 * n_tiles, d_hist.get());
        k_rs_scatter<<<n_tiles, kRsWarps * 32, 0, stream>>>(n, d_keys[cur].get(), d_vals[cur].get(),
                                                            d_keys[cur ^ 1].get(), d_vals[cur ^ 1].get(),
                                                            shift, n_tiles, d_hist.get());
        cur ^= 1;
    }
    B2PT_CUDA(cudaGetLastError());
    const uint64_t* keys = d_keys[cur].get();
    const uint32_t* order = d_vals[cur].get();

    if (n > 1) {
        d_arrivals.zero(stream);
        k_karras<<<blocks_for(n - 1), kThreads, 0, stream>>>(n, keys, d_children.get(), d_parent.get());
        k_refit<<<blocks_for(n), kThreads, 0, stream>>>(n, d_wv.get(), order, d_children.get(),
                                                         d_parent.get(), d_arrivals.get(),
                                                         d_node_box.get(), pad);
        k_emit_nodes<<<blocks_for(n - 1), kThreads, 0, stream>>>(n, d_wv.get(), order, d_children.get(),
                                                                  d_node_box.get(), pad, out.nodes.get());
    } else {
        k_emit_single<<<1, 1, 0, stream>>>(d_wv.get(), pad, out.nodes.get());
    }
    k_emit_tris<<<blocks_for(n), kThreads, 0, stream>>>(n, d_wv.get(), d_nrm.get(), order, d_tri_mesh.get(),
                                                         d_infos.get(), d_idx.get(), d_uv.get(),
                                                         out.tris.get(), out.nrms.get(), out.uvs.get());
    B2PT_CUDA(cudaGetLastError());
    B2PT_CUDA(cudaMemcpyAsync(out.tri_order.get(), order, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
    // (the compaction below reuses the sort buffers `order` lives in: the triangles are emitted first)
    if (out.n_nodes > 0) {
        // K3d into a temporary, one wide node per binary node; K3e keeps the reachable ones, breadth-first
        const int nn = out.n_nodes;
        DeviceBuffer<BvhNode4> d_wide;
        DeviceBuffer<int> d_level, d_front[2], d_remap;
        DeviceBuffer<uint32_t> d_counts;
        d_wide.alloc((size_t)nn);
        d_level.alloc((size_t)nn);
        d_front[0].alloc((size_t)nn); d_front[1].alloc((size_t)nn);
        d_remap.alloc((size_t)nn);
        d_counts.alloc(2 * (kMaxLevels + 1));
        k_collapse4<<<blocks_for(nn), kThreads, 0, stream>>>(nn, out.nodes.get(), d_wide.get(), collapse_greedy(), wide_leaf_size(leaf_size));
        B2PT_CUDA(cudaMemsetAsync(d_level.get(), 0xff, (size_t)nn * 4, stream));
        d_counts.zero(stream);
        uint32_t* cw = d_counts.get();                     // per-level node counts of the wide tree ...
        uint32_t* cb = d_counts.get() + kMaxLevels + 1;    // ... and of the binary tree (depth only)
        // levels are enqueued in batches; a batch whose last level came out empty ends the sweep (a balanced tree
        // of a few million triangles is 12-25 levels deep, kMaxLevels is for the degenerate ones)
        const int sweep_grid = 148 * 2;
        std::vector<uint32_t> h_counts(2 * (kMaxLevels + 1), 0u);
        auto sweep = [&](bool is_wide, uint32_t* cnt, uint32_t* h_cnt, int batch) {
            k_bfs_init<<<1, 1, 0, stream>>>(d_front[0].get(), cnt, is_wide ? d_level.get() : nullptr);
            for (int L0 = 0; L0 < kMaxLevels; L0 += batch) {
                const int L1 = std::min(L0 + batch, kMaxLevels);
                for (int L = L0; L < L1; L++) {
                    if (is_wide)
                        k_bfs_level<true><<<sweep_grid, kThreads, 0, stream>>>(d_wide.get(), nullptr, d_front[L & 1].get(),
                                                                                d_front[(L + 1) & 1].get(), cnt, L, d_level.get());
                    else
                        k_bfs_level<false><<<sweep_grid, kThreads, 0, stream>>>(nullptr, out.nodes.get(), d_front[L & 1].get(),
                                                                                 d_front[(L + 1) & 1].get(), cnt, L, nullptr);
                }
                B2PT_CUDA(cudaMemcpyAsync(h_cnt, cnt, (kMaxLevels + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
                B2PT_CUDA(cudaStreamSynchronize(stream));
                if (h_cnt[L1] == 0) break;
            }
        };
        sweep(true, cw, h_counts.data(), 16);
        if (n > 1) sweep(false, cb, h_counts.data() + kMaxLevels + 1, 32);
        int n_reach = 0;
        for (int L = 0; L <= kMaxLevels; L++) {
            if (h_counts[L]) out.wide_depth = L + 1;
            n_reach += (int)h_counts[L];
            if (h_counts[kMaxLevels + 1 + L]) out.depth = L + 1;
        }
        if (n == 1) out.depth = 1;
        B2PT_REQUIRE(h_counts[kMaxLevels] == 0 && h_counts[2 * kMaxLevels + 1] == 0, B2PT_E_UNSUPPORTED,
                     "acceleration structure deeper than 128 levels");
        out.n_wide_nodes = n_reach;
        // order = reachable nodes by (level, old index): five 8-bit passes cover level (8 bits) + index (32 bits)
        k_compact_keys<<<blocks_for(nn), kThreads, 0, stream>>>(nn, d_level.get(), d_keys[0].get(), d_vals[0].get());
        const int nt2 = (nn + kRsTile - 1) / kRsTile;
        int cur2 = 0;
        for (int pass = 0; pass < 5; pass++) {
            const int shift = pass * 8;
            k_rs_hist<<<nt2, kRsWarps * 32, 0, stream>>>(nn, d_keys[cur2].get(), shift, nt2, d_hist.get());
            k_rs_scan<<<1, 1024, 0, stream>>>(256 * nt2, d_hist.get());
            k_rs_scatter<<<nt2, kRsWarps * 32, 0, stream>>>(nn, d_keys[cur2].get(), d_vals[cur2].get(), d_keys[cur2 ^ 1].get(),
                                                            d_vals[cur2 ^ 1].get(), shift, nt2, d_hist.get());
            cur2 ^= 1;
        }
        const uint32_t* worder = d_vals[cur2].get();
        k_compact_remap<<<blocks_for(n_reach), kThreads, 0, stream>>>(n_reach, worder, d_remap.get());
        const int fmt = trace_wide_formats();   // bit 0: BvhNode4, bit 2: BvhNode4P
        if (fmt & 1) out.nodes4.alloc((size_t)n_reach);
        if (fmt & 4) out.nodes4p.alloc((size_t)n_reach);
        k_compact_emit<<<blocks_for(n_reach), kThreads, 0, stream>>>(n_reach, worder, d_remap.get(), d_wide.get(),
                                                                      out.nodes4.get(), out.nodes4p.get());
        B2PT_CUDA(cudaGetLastError());
        B2PT_CUDA(cudaStreamSynchronize(stream));   // the temporaries above die with this scope
    }
    // the stack the scene's traversal kernel has must hold the deepest descent (bvh.cuh)
    {
        char msg[160];
        if (trace_uses_wide(any_alpha)) {
            snprintf(msg, sizeof msg, "acceleration structure too deep for the traversal stack: %d wide levels, at most %d",
                     out.wide_depth, (kWideStackSize - 1) / 3);
            B2PT_REQUIRE(3 * out.wide_depth + 1 <= kWideStackSize, B2PT_E_UNSUPPORTED, msg);
        } else {
            snprintf(msg, sizeof msg, "acceleration structure too deep for the traversal stack: %d levels, at most %d "
                     "(binary-node kernels, B2PT_BVH4=0)", out.depth, kStackSize - 1);
            B2PT_REQUIRE(out.depth + 1 <= kStackSize, B2PT_E_UNSUPPORTED, msg);
        }
    }
    B2PT_CUDA(cudaEventRecord(ev1, stream));
    B2PT_CUDA(cudaStreamSynchronize(stream));
    B2PT_CUDA(cudaEventElapsedTime(&out.build_ms, ev0, ev1));
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
}

}  // namespace b2pt
