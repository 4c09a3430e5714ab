// bvh.cuh — acceleration-structure layout + per-element build math.
//
// Replaces what the reference delegates to OptiX: optixAccelBuild/Compact over
// one triangle build input per mesh with a FLOAT12 pre-transform
// (Renderer/OptiX/OptixRenderer.cpp:306-456) and RT-core traversal behind
// optixTrace (devicePrograms.cu:224,248).  B200 has no RT cores; this is a
// GPU-built LBVH (Morton codes -> radix sort -> Karras 2012 topology ->
// atomic bottom-up refit) emitted into 16-byte-aligned nodes read with
// 128-bit loads.
//
// Layout in HBM
//   BvhNode2 (64 B, 4 x LDG.128): both children's boxes live in the parent so
//   one node fetch decides both children (Aila-Laine layout).
//   TriRecord (48 B, 3 x LDG.128): world-space v0,v1,v2 + global primitive id,
//   stored in leaf (Morton) order.
//   NrmRecord (48 B): object-space vertex normals + has-normals flag, same order.
#pragma once

#include "vecmath.cuh"

// 128-bit read-only loads of the 16-byte-aligned records below
#if defined(__CUDA_ARCH__)
#define B2PT_LDG4(p) __ldg(reinterpret_cast<const float4*>(p))
#else
#define B2PT_LDG4(p) (*reinterpret_cast<const float4*>(p))
#endif

namespace b2pt {

struct alignas(16) BvhNode2 {
    float4 c0xy;    // child0: lo.x, hi.x, lo.y, hi.y
    float4 c1xy;    // child1: lo.x, hi.x, lo.y, hi.y
    float4 cz;      // child0 lo.z, hi.z, child1 lo.z, hi.z
    int32_t child0; // >= 0: internal node index; < 0: ~(leaf triangle slot)
    int32_t child1;
    int32_t pad0, pad1;
};
static_assert(sizeof(BvhNode2) == 64, "node must be 64 bytes");

// 4-wide node (K5/K8 read this one): the four grandchildren of binary node i — or a child itself where
// that child is a leaf — stored per axis across the four slots, so a visit is 7 x LDG.128 and covers two
// levels of the binary tree in one dependent step.  Entry i exists for every binary node (same index), the
// traversal from the root only ever reaches the even-depth ones.  Empty slots have the box
// lo = hi = +inf (never hit) and child = kEmptyChild.
struct alignas(16) BvhNode4 {
    float4 lox, hix, loy, hiy, loz, hiz;   // component k = slot k
    int32_t child[4];                      // >= 0: wide node index; < 0: leaf, ~(first triangle slot | (count - 1) << 28)
    int32_t pad[4];
};
static_assert(sizeof(BvhNode4) == 128, "wide node must be 128 bytes");
constexpr int32_t kEmptyChild = 0x7ffffffe;
// A leaf child of a wide node names a RUN of triangle records: ~(first | (count - 1) << kLeafCountShift), count <= 8.
// One triangle per leaf (the default) is count = 1: ~slot, as in the binary nodes.
constexpr int kLeafCountShift = 28;
constexpr int kDefaultLeafSize = 2;   // 1 | 2 | 3 | 4 measured: profiles/r02_ab_leaf_size.txt
constexpr uint32_t kLeafSlotMask = (1u << kLeafCountShift) - 1u;

// Traversal stacks (per ray): the binary kernels push one entry per level below the root plus a sentinel, the wide
// kernels at most three per level plus the sentinel.  A Karras tree over 63-bit Morton keys with index tie-breaks
// can be 63 + log2(n) levels deep (many triangles in one Morton cell, a teapot in a stadium), so the build MEASURES
// the depth of both trees (K3e) and refuses a scene the stack of its traversal kernel could not hold
// (B2PT_E_UNSUPPORTED) — the kernels themselves push without a bound check.
constexpr int kStackSize = 64;
constexpr int kWideStackSize = 3 * kStackSize;

// The 4-wide node in 96 bytes, three 256-bit loads per visit (what K5/K8 read by default, B2PT_STEP=6): the same fp32 boxes, with the four
// child indices hidden in the low mantissa byte of the sixteen x / y planes (child k = the low bytes of
// lox[k], hix[k], loy[k], hiy[k], least significant first).  K3d rounds each of those planes OUTWARD to a value
// whose low byte is the payload — at most 511 ulp, 2^-14 relative — so the traversal uses the planes as they
// are, payload included, and still only prunes.  Empty slots: planes of +3e38 (finite: the payload must not
// turn them into NaNs), child = kEmptyChild.
struct alignas(32) BvhNode4P {
    float4 lox, hix, loy, hiy, loz, hiz;   // component k = slot k
};
static_assert(sizeof(BvhNode4P) == 96, "packed wide node must be 96 bytes");

struct alignas(16) TriRecord {
    float4 a;   // v0.x v0.y v0.z v1.x
    float4 b;   // v1.y v1.z v2.x v2.y
    float4 c;   // v2.z, prim id (bits), mesh id (bits), unused
};
static_assert(sizeof(TriRecord) == 48, "triangle record must be 48 bytes");

struct alignas(16) NrmRecord {
    float4 a;   // n0.x n0.y n0.z n1.x      (OBJECT space, as uploaded: the reference
                //  interpolates first and transforms after, devicePrograms.cu:95-100)
    float4 b;   // n1.y n1.z n2.x n2.y
    float4 c;   // n2.z, has_normals (bits), unused, unused
};

struct alignas(8) UvRecord {
    float2 t0, t1, t2;
};

constexpr uint32_t kHitMiss = 0xFFFFFFFFu;
constexpr uint32_t kHitBackface = 0x80000000u;

// 16-byte hit record written by the traversal kernels.
struct alignas(16) HitRecord {
    uint32_t tri;   // leaf slot of the triangle | kHitBackface, or kHitMiss
    float t, u, v;  // fp32 roundings of the fp64 decision
};

// ---------------------------------------------------------------------------
// Morton codes (K1)
// ---------------------------------------------------------------------------
B2PT_HD uint64_t expand_bits_21(uint32_t v) {
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// 63-bit Morton code of a point normalised to the scene's centroid bounds.
B2PT_HD uint64_t morton63(float3 c, float3 lo, float3 inv_extent) {
    const float s = 2097152.0f;  // 2^21
    float fx = fminf(fmaxf((c.x - lo.x) * inv_extent.x * s, 0.0f), s - 1.0f);
    float fy = fminf(fmaxf((c.y - lo.y) * inv_extent.y * s, 0.0f), s - 1.0f);
    float fz = fminf(fmaxf((c.z - lo.z) * inv_extent.z * s, 0.0f), s - 1.0f);
    return (expand_bits_21((uint32_t)fx) << 2) | (expand_bits_21((uint32_t)fy) << 1) |
           expand_bits_21((uint32_t)fz);
}

// ---------------------------------------------------------------------------
// Karras 2012 "Maximizing Parallelism in the Construction of BVHs" (K3)
// ---------------------------------------------------------------------------
B2PT_HD int clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
B2PT_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}

// length of the common prefix of sorted keys i and j; duplicates are
// disambiguated by the index (Karras §4).  Returns -1 outside [0,n).
B2PT_HD int karras_delta(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + clz32((uint32_t)i ^ (uint32_t)j);
    return clz64(a ^ b);
}

// internal node i covers sorted leaves [first,last] and splits after `split`.
B2PT_HD void karras_node(const uint64_t* keys, int n, int i, int& first, int& last, int& split) {
    const int d = (karras_delta(keys, n, i, i + 1) - karras_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = karras_delta(keys, n, i, i - d);
    int lmax = 2;
    while (karras_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (karras_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = karras_delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {  // t = ceil(l/2), ceil(l/4), ..., 1
        t = (t + 1) >> 1;
        if (karras_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    split = i + s * d + (d < 0 ? -1 : 0);
    first = d > 0 ? i : j;
    last = d > 0 ? j : i;
}

struct Aabb {
    float3 lo, hi;
};
B2PT_HD Aabb aabb_empty() {
    Aabb b;
    b.lo = v3(INFINITY);
    b.hi = v3(-INFINITY);
    return b;
}
B2PT_HD Aabb aabb_union(Aabb a, Aabb b) {
    Aabb r;
    r.lo = v3(fminf(a.lo.x, b.lo.x), fminf(a.lo.y, b.lo.y), fminf(a.lo.z, b.lo.z));
    r.hi = v3(fmaxf(a.hi.x, b.hi.x), fmaxf(a.hi.y, b.hi.y), fmaxf(a.hi.z, b.hi.z));
    return r;
}
B2PT_HD Aabb aabb_of_tri(float3 a, float3 b, float3 c) {
    Aabb r;
    r.lo = v3(fminf(fminf(a.x, b.x), c.x), fminf(fminf(a.y, b.y), c.y), fminf(fminf(a.z, b.z), c.z));
    r.hi = v3(fmaxf(fmaxf(a.x, b.x), c.x), fmaxf(fmaxf(a.y, b.y), c.y), fmaxf(fmaxf(a.z, b.z), c.z));
    return r;
}
// Outward padding so that the fp32 slab test stays conservative with respect
// to the fp64 triangle decision (DESIGN.md "exact hits").
B2PT_HD Aabb aabb_pad(Aabb b, float pad) {
    b.lo = b.lo - v3(pad);
    b.hi = b.hi + v3(pad);
    return b;
}

B2PT_HD void tri_unpack(const TriRecord& r, float3& v0, float3& v1, float3& v2) {
    v0 = v3(r.a.x, r.a.y, r.a.z);
    v1 = v3(r.a.w, r.b.x, r.b.y);
    v2 = v3(r.b.z, r.b.w, r.c.x);
}
// flags (c.w, as bits): 1 = the triangle's mesh has an alpha cut-out texture — the any-hit predicate must see its hits
// (the wide kernels read the flag with the record instead of chasing materials[mesh] for every triangle they test)
B2PT_HD TriRecord tri_pack(float3 v0, float3 v1, float3 v2, uint32_t prim, uint32_t mesh, uint32_t flags = 0u) {
    TriRecord r;
    r.a = make_float4(v0.x, v0.y, v0.z, v1.x);
    r.b = make_float4(v1.y, v1.z, v2.x, v2.y);
#if defined(__CUDA_ARCH__)
    r.c = make_float4(v2.z, __uint_as_float(prim), __uint_as_float(mesh), __uint_as_float(flags));
#else
    float fp, fm, ff;
    __builtin_memcpy(&fp, &prim, 4);
    __builtin_memcpy(&fm, &mesh, 4);
    __builtin_memcpy(&ff, &flags, 4);
    r.c = make_float4(v2.z, fp, fm, ff);
#endif
    return r;
}
B2PT_HD uint32_t f2u_bits(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u;
    __builtin_memcpy(&u, &f, 4);
    return u;
#endif
}
B2PT_HD float u2f_bits(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    __builtin_memcpy(&f, &u, 4);
    return f;
#endif
}

}  // namespace b2pt
