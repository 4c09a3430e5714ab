// hostsim.cpp — TEST HARNESS: runs the per-thread logic of the CUDA kernels
// (the B2PT_HD functions in optixpathtracer_b200/csrc/*.cuh) on the host so
// that traversal / BVH-topology / BSDF logic can be checked against the oracle
// in the CPU-only test tier.  It is never part of the product: libb2pt.so does
// not contain it and nothing in the package loads it.
//
// The LBVH here is built by a sequential loop over the SAME per-element
// functions the build kernels call (morton63, karras_node, aabb_*), so a
// topology bug shows up without a GPU.
#include <algorithm>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../optixpathtracer_b200/csrc/traverse.cuh"

using namespace b2pt;

namespace {

struct HostBvh {
    std::vector<BvhNode2> nodes;
    std::vector<TriRecord> tris;
    std::vector<uint32_t> order;
};

Aabb box_of(const std::vector<int2>& children, const std::vector<Aabb>& node_box,
            const std::vector<Aabb>& leaf_box, int c) {
    return c < 0 ? leaf_box[~c] : node_box[c];
}

void build(const float* soup, int n, HostBvh& out) {
    out.nodes.clear();
    out.tris.clear();
    out.order.clear();
    if (n == 0) return;
    std::vector<float3> wv(3 * (size_t)n);
    memcpy(wv.data(), soup, sizeof(float) * 9 * (size_t)n);
    float3 clo = v3(INFINITY), chi = v3(-INFINITY);
    float maxabs = 0.f;
    for (int g = 0; g < n; g++) {
        Aabb b = aabb_of_tri(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2]);
        float3 c = (b.lo + b.hi) * 0.5f;
        clo = v3(fminf(clo.x, c.x), fminf(clo.y, c.y), fminf(clo.z, c.z));
        chi = v3(fmaxf(chi.x, c.x), fmaxf(chi.y, c.y), fmaxf(chi.z, c.z));
        const float vals[6] = {b.lo.x, b.lo.y, b.lo.z, b.hi.x, b.hi.y, b.hi.z};
        for (float f : vals) maxabs = fmaxf(maxabs, fabsf(f));
    }
    const float pad = fmaxf(maxabs, 1e-30f) * 3.8146973e-6f;
    const float3 ext = chi - clo;
    const float3 inv = v3(ext.x > 0 ? 1.0f / ext.x : 0.f, ext.y > 0 ? 1.0f / ext.y : 0.f, ext.z > 0 ? 1.0f / ext.z : 0.f);
    std::vector<uint64_t> keys(n);
    for (int g = 0; g < n; g++) {
        Aabb b = aabb_of_tri(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2]);
        keys[g] = morton63((b.lo + b.hi) * 0.5f, clo, inv);
    }
    out.order.resize(n);
    std::iota(out.order.begin(), out.order.end(), 0u);
    std::stable_sort(out.order.begin(), out.order.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
    std::vector<uint64_t> sk(n);
    for (int s = 0; s < n; s++) sk[s] = keys[out.order[s]];
    std::vector<Aabb> leaf_box(n);
    out.tris.resize(n);
    for (int s = 0; s < n; s++) {
        const uint32_t g = out.order[s];
        leaf_box[s] = aabb_pad(aabb_of_tri(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2]), pad);
        out.tris[s] = tri_pack(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2], g, 0);
    }
    if (n == 1) {
        BvhNode2 nd;
        const Aabb b = leaf_box[0];
        nd.c0xy = make_float4(b.lo.x, b.hi.x, b.lo.y, b.hi.y);
        nd.c1xy = make_float4(INFINITY, -INFINITY, INFINITY, -INFINITY);
        nd.cz = make_float4(b.lo.z, b.hi.z, INFINITY, -INFINITY);
        nd.child0 = ~0; nd.child1 = ~0; nd.pad0 = nd.pad1 = 0;
        out.nodes.push_back(nd);
        return;
    }
    std::vector<int2> children(n - 1);
    std::vector<int> first(n - 1), last(n - 1);
    for (int i = 0; i < n - 1; i++) {
        int f, l, sp;
        karras_node(sk.data(), n, i, f, l, sp);
        first[i] = f; last[i] = l;
        children[i] = make_int2(sp == f ? ~sp : sp, sp + 1 == l ? ~(sp + 1) : sp + 1);
    }
    // refit: process nodes by decreasing range length is not needed — recurse.
    std::vector<Aabb> node_box(n - 1);
    std::vector<char> done(n - 1, 0);
    std::vector<int> stack;
    stack.push_back(0);
    while (!stack.empty()) {
        const int i = stack.back();
        const int2 ch = children[i];
        bool ready = true;
        if (ch.x >= 0 && !done[ch.x]) { stack.push_back(ch.x); ready = false; }
        if (ch.y >= 0 && !done[ch.y]) { stack.push_back(ch.y); ready = false; }
        if (!ready) continue;
        node_box[i] = aabb_union(box_of(children, node_box, leaf_box, ch.x), box_of(children, node_box, leaf_box, ch.y));
        done[i] = 1;
        stack.pop_back();
    }
    out.nodes.resize(n - 1);
    for (int i = 0; i < n - 1; i++) {
        const int2 ch = children[i];
        const Aabb b0 = box_of(children, node_box, leaf_box, ch.x), b1 = box_of(children, node_box, leaf_box, ch.y);
        BvhNode2 nd;
        nd.c0xy = make_float4(b0.lo.x, b0.hi.x, b0.lo.y, b0.hi.y);
        nd.c1xy = make_float4(b1.lo.x, b1.hi.x, b1.lo.y, b1.hi.y);
        nd.cz = make_float4(b0.lo.z, b0.hi.z, b1.lo.z, b1.hi.z);
        nd.child0 = ch.x; nd.child1 = ch.y; nd.pad0 = nd.pad1 = 0;
        out.nodes[i] = nd;
    }
}

HostBvh g_bvh;

}  // namespace

extern "C" {

// Build the LBVH over a world-space triangle soup (nt*9 floats). Returns the node count.
int hostsim_build(const float* soup, int nt) {
    build(soup, nt, g_bvh);
    return (int)g_bvh.nodes.size();
}

// Structural self-check: every leaf slot reachable exactly once, child boxes
// contain their triangles. Returns 0 when consistent.
int hostsim_check_topology() {
    const int n = (int)g_bvh.tris.size();
    if (n <= 1) return 0;
    std::vector<int> seen(n, 0);
    std::vector<int> st{0};
    int visited = 0;
    while (!st.empty()) {
        int c = st.back(); st.pop_back();
        if (c < 0) { if (~c >= n) return 2; seen[~c]++; continue; }
        if (c >= n - 1) return 3;
        visited++;
        st.push_back(g_bvh.nodes[c].child0);
        st.push_back(g_bvh.nodes[c].child1);
    }
    if (visited != n - 1) return 4;
    for (int s = 0; s < n; s++) if (seen[s] != 1) return 5;
    return 0;
}

void hostsim_trace_closest(const float* org, const float* dir, int64_t n, float tmin, float tmax,
                           int32_t* prim, float* t, float* u, float* v, uint8_t* backface,
                           uint64_t* counters /* nodes, tris, exact */) {
    uint64_t cn = 0, ct = 0, ce = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : cn, ct, ce)
    for (int64_t i = 0; i < n; i++) {
        HitRecord h;
        TraceStats ts = {0, 0, 0};
        traverse<false, true>(g_bvh.nodes.data(), g_bvh.tris.data(), (int)g_bvh.nodes.size(),
                              v3(org[3 * i], org[3 * i + 1], org[3 * i + 2]),
                              v3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]), tmin, tmax, h, &ts);
        const bool hit = h.tri != kHitMiss;
        prim[i] = hit ? (int32_t)g_bvh.order[h.tri & ~kHitBackface] : -1;
        t[i] = hit ? h.t : 0.f; u[i] = hit ? h.u : 0.f; v[i] = hit ? h.v : 0.f;
        if (backface) backface[i] = hit && (h.tri & kHitBackface);
        cn += ts.nodes; ct += ts.tris; ce += ts.exact;
    }
    if (counters) { counters[0] = cn; counters[1] = ct; counters[2] = ce; }
}

void hostsim_trace_any(const float* org, const float* dir, int64_t n, float tmin, const float* tmax,
                       uint8_t* occluded) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < n; i++) {
        HitRecord h;
        occluded[i] = traverse<true, false>(g_bvh.nodes.data(), g_bvh.tris.data(), (int)g_bvh.nodes.size(),
                                            v3(org[3 * i], org[3 * i + 1], org[3 * i + 2]),
                                            v3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]), tmin, tmax[i], h,
                                            nullptr) ? 1 : 0;
    }
}

}  // extern "C"
