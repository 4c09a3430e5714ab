"""CPU tier: the ray/triangle contract.

1. known-answer tests pin the oracle (oracle/oracle_trace.cpp) to hand-computed
   hits — the reference holds no vectors at this boundary (OptiX is closed), so
   these analytic cases are what pins it;
2. the kernels' per-thread traversal logic (traverse.cuh, via tests/hostsim)
   must equal the brute-force oracle BIT FOR BIT: primitive ids, t, u, v,
   back-face flag — no tolerance, no tie exceptions.
"""
import numpy as np

from tests import cases
from tests.util import bits, random_rays
from optixpathtracer_b200 import scenes

f32 = np.float32


def test_oracle_known_answers(oracle):
    soup = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], dtype=f32)
    o = np.array([[0.25, 0.25, 1.0], [0.25, 0.25, -2.0], [2, 2, 1], [0.25, 0.25, 1.0], [0.25, 0.25, 1.0]], dtype=f32)
    d = np.array([[0, 0, -1], [0, 0, 1], [0, 0, -1], [0, 0, 1], [1, 0, 0]], dtype=f32)
    prim, t, u, v, bf = oracle.closest(soup, o, d)
    assert prim.tolist() == [0, 0, -1, -1, -1]
    assert t[0] == f32(1.0) and t[1] == f32(2.0)
    assert u[0] == f32(0.25) and v[0] == f32(0.25)
    # CCW seen from +z: ray coming from +z hits the front face
    assert bf.tolist()[:2] == [0, 1]
    # t range is open: tmax = 1 excludes t == 1, tmin = 1 excludes it too
    assert oracle.closest(soup, o[:1], d[:1], 0.0, 1.0)[0][0] == -1
    assert oracle.closest(soup, o[:1], d[:1], 1.0, 100.0)[0][0] == -1
    assert oracle.closest(soup, o[:1], d[:1], 0.0, 1.0001)[0][0] == 0
    assert oracle.any(soup, o, d, 100.0).tolist() == [1, 1, 0, 0, 0]
    assert oracle.any(soup, o, d, [0.5, 2.5, 100, 100, 100]).tolist() == [0, 1, 0, 0, 0]


def test_oracle_tie_rule(oracle):
    soup = cases.edge_soup()
    o = np.array([[0.75, 0.25, 2.0]], dtype=f32)
    d = np.array([[0, 0, -1]], dtype=f32)
    prim, t, *_ = oracle.closest(soup, o, d)
    assert prim[0] == 0 and t[0] == f32(2.0)          # duplicate triangle 2 loses the tie
    assert oracle.tie_count(soup, o, d, prim)[0] == 2


def _compare(hostsim, oracle, soup, o, d, tmin=0.0, tmax=100.0):
    n_nodes, topo = hostsim.build(soup)
    assert topo == 0
    hp, ht, hu, hv, hbf, _ = hostsim.closest(o, d, tmin, tmax)
    op, ot, ou, ov, obf = oracle.closest(soup, o, d, tmin, tmax)
    assert np.array_equal(hp, op)
    assert np.array_equal(bits(ht), bits(ot))
    assert np.array_equal(bits(hu), bits(ou))
    assert np.array_equal(bits(hv), bits(ov))
    assert np.array_equal(hbf, obf)
    return op, ot


def test_traversal_logic_scene1(hostsim, oracle):
    sc = cases.small_spheres()
    soup, _ = scenes.scene_triangle_soup(sc)
    o, d = cases.scene_rays(sc, 64, 48, 3000, seed=7)
    prim, t = _compare(hostsim, oracle, soup, o, d)
    assert (prim >= 0).mean() > 0.5
    # shadow-style any-hit queries with per-ray tmax around the closest hit
    tm = np.where(prim >= 0, t * np.where(np.arange(len(t)) % 2 == 0, f32(0.999), f32(1.001)), f32(5.0)).astype(f32)
    assert np.array_equal(hostsim.any(o, d, tm), oracle.any(soup, o, d, tm))


def test_traversal_logic_edge_cases(hostsim, oracle):
    soup = cases.edge_soup()
    o, d = cases.edge_rays()
    _compare(hostsim, oracle, soup, o, d)
    _compare(hostsim, oracle, soup, o, d, tmin=0.5, tmax=2.5)
    assert np.array_equal(hostsim.any(o, d, 100.0), oracle.any(soup, o, d, 100.0))


def test_traversal_logic_tiny_scenes(hostsim, oracle):
    rng = np.random.default_rng(3)
    o, d = random_rays(rng, 500, [-2, -2, -2], [2, 2, 2])
    for nt in (1, 2, 3, 7):
        soup = rng.uniform(-1, 1, (nt, 3, 3)).astype(f32)
        _compare(hostsim, oracle, soup, o, d)
    # empty scene: every ray misses
    n_nodes, topo = hostsim.build(np.zeros((0, 3, 3), dtype=f32))
    assert n_nodes == 0 and topo == 0
    assert (hostsim.closest(o, d)[0] == -1).all()
    assert (hostsim.any(o, d, 10.0) == 0).all()


def test_traversal_logic_random_soup_far_origins(hostsim, oracle):
    """Small triangles seen from far away: the regime where an fp32-only
    Moeller-Trumbore misclassifies edge hits; the fp64 decision must not."""
    rng = np.random.default_rng(11)
    c = rng.uniform(-1, 1, (4000, 1, 3))
    soup = (c + rng.normal(scale=0.01, size=(4000, 3, 3))).astype(f32)
    tgt = soup.reshape(-1, 3)[rng.integers(0, 12000, 4000)]          # aim at vertices: edge-on hits
    o = (tgt + rng.normal(size=(4000, 3)) * 30).astype(f32)
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    _compare(hostsim, oracle, soup, o, d, 0.0, 1000.0)


def test_host_bvh_equals_bruteforce(oracle):
    """oracle/host_bvh.h (used to make the CPU oracle render in finite time) must
    return exactly what the exhaustive search returns."""
    import ctypes as C
    from tests.util import fp, vp
    sc = cases.small_spheres()
    soup, _ = scenes.scene_triangle_soup(sc)
    o, d = cases.scene_rays(sc, 48, 32, 2000, seed=9)
    n = len(o)
    op, ot, ou, ov, obf = oracle.closest(soup, o, d)
    prim = np.empty(n, np.int32); t = np.empty(n, f32); u = np.empty(n, f32); v = np.empty(n, f32)
    bf = np.empty(n, np.uint8)
    s9 = np.ascontiguousarray(soup.reshape(-1, 9))
    oracle.lib.oracle_closest_bvh(fp(s9), C.c_int64(len(s9)), fp(o), fp(d), C.c_int64(n), C.c_float(0),
                                  C.c_float(100), vp(prim), fp(t), fp(u), fp(v), vp(bf))
    assert np.array_equal(prim, op) and np.array_equal(bits(t), bits(ot))
    assert np.array_equal(bits(u), bits(ou)) and np.array_equal(bits(v), bits(ov)) and np.array_equal(bf, obf)
    tm = np.where(op >= 0, ot * f32(0.999), f32(3.0)).astype(f32)
    occ = np.empty(n, np.uint8)
    oracle.lib.oracle_any_bvh(fp(s9), C.c_int64(len(s9)), fp(o), fp(d), C.c_int64(n), C.c_float(0), fp(tm), vp(occ))
    assert np.array_equal(occ, oracle.any(soup, o, d, tm))
