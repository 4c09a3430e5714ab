// host_bvh.h — TEST INFRASTRUCTURE ONLY.  A plain host BVH2 so that the CPU
// oracle and the host-compiled reference can trace rays in reasonable time.
//
// NOT reference code: the reference has no CPU traversal (it calls OptiX,
// devicePrograms.cu:224,248).  The tree only prunes; every leaf decision is
// the fp64 exact_triangle() of exact_tri.h and the slab test is done in double
// with padded boxes, so results are identical to oracle_closest_bruteforce
// (tests/test_trace_cpu.py::test_host_bvh_equals_bruteforce checks that).
// Independent of the product's LBVH: binned-SAH top-down build, median fallback.
#pragma once
#include <algorithm>
#include <cfloat>
#include <vector>

#include "exact_tri.h"

namespace oracle {

struct ClosestHit {
    int64_t prim = -1;
    double t = 0, u = 0, v = 0;
    bool backface = false;
};

class HostBvh {
  public:
    void build(const float* soup, int64_t n) {
        tris_ = soup;
        n_ = n;
        nodes_.clear();
        order_.resize((size_t)n);
        for (int64_t i = 0; i < n; i++) order_[i] = i;
        lo_.resize((size_t)n * 3);
        hi_.resize((size_t)n * 3);
        double maxabs = 0;
        for (int64_t i = 0; i < n; i++)
            for (int a = 0; a < 3; a++) {
                const float* t = soup + 9 * i;
                lo_[3 * i + a] = std::min({t[a], t[3 + a], t[6 + a]});
                hi_[3 * i + a] = std::max({t[a], t[3 + a], t[6 + a]});
                maxabs = std::max({maxabs, std::fabs((double)lo_[3 * i + a]), std::fabs((double)hi_[3 * i + a])});
            }
        pad_ = std::max(maxabs, 1e-30) * 1e-9;
        if (n > 0) {
            nodes_.reserve((size_t)(2 * n));
            nodes_.emplace_back();
            subdivide(0, 0, n);
        }
    }

    struct AcceptAll {
        bool operator()(int64_t, double, double) const { return true; }
    };
    ClosestHit closest(const float* o, const float* d, double tmin, double tmax) const {
        return closest(o, d, tmin, tmax, AcceptAll());
    }
    bool any(const float* o, const float* d, double tmin, double tmax) const {
        return any(o, d, tmin, tmax, AcceptAll());
    }

    // `accept(prim, u, v)` plays the any-hit program: returning false ignores
    // the intersection (optixIgnoreIntersection, devicePrograms.cu:549-561).
    template <class Accept>
    ClosestHit closest(const float* o, const float* d, double tmin, double tmax, const Accept& accept) const {
        ClosestHit best;
        if (nodes_.empty()) return best;
        double bt = tmax;
        const double inv[3] = {1.0 / safe(d[0]), 1.0 / safe(d[1]), 1.0 / safe(d[2])};
        int stack[128];
        int sp = 0;
        stack[sp++] = 0;
        while (sp) {
            const Node& nd = nodes_[stack[--sp]];
            if (!slab(nd, o, inv, tmin, bt)) continue;
            if (nd.count) {
                for (int64_t k = nd.first; k < nd.first + nd.count; k++) {
                    const int64_t p = order_[k];
                    const Hit h = exact_triangle(o, d, tris_ + 9 * p);
                    if (!h.hit || !(h.t > tmin) || !(h.t < tmax)) continue;
                    if (h.t < bt || (h.t == bt && best.prim >= 0 && p < best.prim)) {
                        if (!accept(p, h.u, h.v)) continue;
                        bt = h.t;
                        best.prim = p; best.t = h.t; best.u = h.u; best.v = h.v; best.backface = h.backface;
                    }
                }
            } else {
                stack[sp++] = nd.left;
                stack[sp++] = nd.left + 1;
            }
        }
        return best;
    }

    template <class Accept>
    bool any(const float* o, const float* d, double tmin, double tmax, const Accept& accept) const {
        if (nodes_.empty()) return false;
        const double inv[3] = {1.0 / safe(d[0]), 1.0 / safe(d[1]), 1.0 / safe(d[2])};
        int stack[128];
        int sp = 0;
        stack[sp++] = 0;
        while (sp) {
            const Node& nd = nodes_[stack[--sp]];
            if (!slab(nd, o, inv, tmin, tmax)) continue;
            if (nd.count) {
                for (int64_t k = nd.first; k < nd.first + nd.count; k++) {
                    const Hit h = exact_triangle(o, d, tris_ + 9 * order_[k]);
                    if (h.hit && h.t > tmin && h.t < tmax && accept(order_[k], h.u, h.v)) return true;
                }
            } else {
                stack[sp++] = nd.left;
                stack[sp++] = nd.left + 1;
            }
        }
        return false;
    }

  private:
    struct Node {
        double lo[3], hi[3];
        int left = 0;          // children at left, left+1
        int64_t first = 0;
        int count = 0;         // > 0: leaf
    };
    static double safe(float x) { return std::fabs(x) > 1e-30f ? (double)x : (x < 0 || std::signbit(x) ? -1e-30 : 1e-30); }

    bool slab(const Node& nd, const float* o, const double* inv, double tmin, double tmax) const {
        double t0 = tmin, t1 = tmax;
        for (int a = 0; a < 3; a++) {
            double x0 = (nd.lo[a] - o[a]) * inv[a], x1 = (nd.hi[a] - o[a]) * inv[a];
            if (x0 > x1) std::swap(x0, x1);
            t0 = std::max(t0, x0);
            t1 = std::min(t1, x1);
        }
        return t0 <= t1 * (1.0 + 1e-12) + 1e-300;
    }

    void bounds(int64_t first, int64_t count, double* lo, double* hi) const {
        for (int a = 0; a < 3; a++) { lo[a] = DBL_MAX; hi[a] = -DBL_MAX; }
        for (int64_t k = first; k < first + count; k++)
            for (int a = 0; a < 3; a++) {
                lo[a] = std::min(lo[a], (double)lo_[3 * order_[k] + a]);
                hi[a] = std::max(hi[a], (double)hi_[3 * order_[k] + a]);
            }
    }

    void subdivide(int ni, int64_t first, int64_t count) {
        double lo[3], hi[3];
        bounds(first, count, lo, hi);
        for (int a = 0; a < 3; a++) { nodes_[ni].lo[a] = lo[a] - pad_; nodes_[ni].hi[a] = hi[a] + pad_; }
        if (count <= 4) { nodes_[ni].first = first; nodes_[ni].count = (int)count; return; }
        // binned SAH over centroids
        double clo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, chi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
        auto cen = [&](int64_t p, int a) { return 0.5 * ((double)lo_[3 * p + a] + (double)hi_[3 * p + a]); };
        for (int64_t k = first; k < first + count; k++)
            for (int a = 0; a < 3; a++) { clo[a] = std::min(clo[a], cen(order_[k], a)); chi[a] = std::max(chi[a], cen(order_[k], a)); }
        constexpr int B = 16;
        int best_axis = -1, best_bin = 0;
        double best_cost = DBL_MAX;
        for (int a = 0; a < 3; a++) {
            if (!(chi[a] > clo[a])) continue;
            double blo[B][3], bhi[B][3];
            int64_t bc[B] = {0};
            for (int b = 0; b < B; b++) for (int c = 0; c < 3; c++) { blo[b][c] = DBL_MAX; bhi[b][c] = -DBL_MAX; }
            const double scale = B / (chi[a] - clo[a]);
            for (int64_t k = first; k < first + count; k++) {
                const int64_t p = order_[k];
                int b = std::min(B - 1, (int)((cen(p, a) - clo[a]) * scale));
                bc[b]++;
                for (int c = 0; c < 3; c++) { blo[b][c] = std::min(blo[b][c], (double)lo_[3 * p + c]); bhi[b][c] = std::max(bhi[b][c], (double)hi_[3 * p + c]); }
            }
            double la[B], ra[B];
            int64_t ln[B], rn[B];
            double l0[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, l1[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
            int64_t c = 0;
            auto area = [](const double* x0, const double* x1) {
                const double e0 = x1[0] - x0[0], e1 = x1[1] - x0[1], e2 = x1[2] - x0[2];
                return e0 * e1 + e1 * e2 + e2 * e0;
            };
            for (int b = 0; b < B - 1; b++) {
                c += bc[b];
                for (int k = 0; k < 3; k++) { l0[k] = std::min(l0[k], blo[b][k]); l1[k] = std::max(l1[k], bhi[b][k]); }
                ln[b] = c; la[b] = c ? area(l0, l1) : 0;
            }
            double r0[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, r1[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
            c = 0;
            for (int b = B - 1; b > 0; b--) {
                c += bc[b];
                for (int k = 0; k < 3; k++) { r0[k] = std::min(r0[k], blo[b][k]); r1[k] = std::max(r1[k], bhi[b][k]); }
                rn[b - 1] = c; ra[b - 1] = c ? area(r0, r1) : 0;
            }
            for (int b = 0; b < B - 1; b++) {
                if (!ln[b] || !rn[b]) continue;
                const double cost = la[b] * ln[b] + ra[b] * rn[b];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = b; }
            }
        }
        int64_t mid;
        if (best_axis < 0) {
            mid = first + count / 2;  // all centroids equal: split by index
        } else {
            const int a = best_axis;
            const double scale = B / (chi[a] - clo[a]);
            auto it = std::partition(order_.begin() + first, order_.begin() + first + count, [&](int64_t p) {
                return std::min(B - 1, (int)((cen(p, a) - clo[a]) * scale)) <= best_bin;
            });
            mid = it - order_.begin();
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        const int left = (int)nodes_.size();
        nodes_[ni].left = left;
        nodes_.emplace_back();
        nodes_.emplace_back();
        subdivide(left, first, mid - first);
        subdivide(left + 1, mid, first + count - mid);
    }

    const float* tris_ = nullptr;
    int64_t n_ = 0;
    double pad_ = 0;
    std::vector<Node> nodes_;
    std::vector<int64_t> order_;
    std::vector<float> lo_, hi_;
};

}  // namespace oracle
