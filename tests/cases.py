"""Seeded ray/scene cases shared by the CPU-tier and GPU-tier traversal tests."""
from __future__ import annotations

import numpy as np

from optixpathtracer_b200 import camera_rays, scenes

f32 = np.float32


def small_spheres(segments=24, rings=12):
    return scenes.spheres_scene(segments=segments, rings=rings)


def scene_rays(scene, w, h, n_random, seed):
    """Primary rays of the scene camera + random interior rays + axis-parallel rays."""
    pos, iv, ip = scene["camera"].launch_matrices(w, h)
    o, d = camera_rays(pos, iv, ip, w, h)
    rng = np.random.default_rng(seed)
    o2 = rng.uniform([-0.5, -0.5, -2.0], [3.0, 2.5, 2.0], (n_random, 3)).astype(f32)
    d2 = rng.normal(size=(n_random, 3))
    d2 = (d2 / np.linalg.norm(d2, axis=1, keepdims=True)).astype(f32)
    # rays with exact zero direction components (slab test edge case)
    axes = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1],
                     [0.6, 0.8, 0], [0, 0.6, -0.8]], dtype=f32)
    o3 = rng.uniform([0.5, 0.0, -1.5], [3.0, 2.0, 1.5], (len(axes) * 40, 3)).astype(f32)
    d3 = np.tile(axes, (40, 1))
    return np.concatenate([o, o2, o3]), np.concatenate([d, d2, d3])


def edge_soup():
    """Hand-made soup exercising ties, shared edges, degenerate and tiny triangles."""
    t = []
    # two coplanar triangles sharing the diagonal of a unit quad at z = 0
    t.append([[0, 0, 0], [1, 0, 0], [1, 1, 0]])
    t.append([[0, 0, 0], [1, 1, 0], [0, 1, 0]])
    # an exact duplicate of triangle 0 (tie on t: lower primitive id must win)
    t.append([[0, 0, 0], [1, 0, 0], [1, 1, 0]])
    # a degenerate (zero-area) triangle
    t.append([[0.2, 0.2, 0.5], [0.4, 0.4, 0.5], [0.6, 0.6, 0.5]])
    # a tiny triangle far from the origin
    t.append([[50, 50, 50], [50.001, 50, 50], [50, 50.001, 50]])
    # back-facing quad at z = -1 (opposite winding)
    t.append([[0, 0, -1], [1, 1, -1], [1, 0, -1]])
    t.append([[0, 0, -1], [0, 1, -1], [1, 1, -1]])
    return np.array(t, dtype=f32)


def edge_rays():
    o, d = [], []
    # through the shared diagonal, vertices, and edges of the quad at z=0
    for x, y in [(0.5, 0.5), (0.25, 0.25), (0.0, 0.0), (1.0, 1.0), (1.0, 0.0), (0.5, 0.0),
                 (0.0, 0.5), (0.3, 0.7), (0.7, 0.3), (1.0, 0.5), (0.5, 1.0), (1.5, 0.5), (-0.1, 0.2)]:
        o.append([x, y, 2.0]); d.append([0, 0, -1])
        o.append([x, y, -3.0]); d.append([0, 0, 1])
    # grazing, parallel to the plane
    o.append([-1, 0.5, 0.0]); d.append([1, 0, 0])
    o.append([-1, 0.5, 1e-7]); d.append([1, 0, 0])
    # origin exactly on the surface (t = 0 must not count)
    o.append([0.5, 0.25, 0.0]); d.append([0, 0, -1])
    o.append([0.5, 0.25, 0.0]); d.append([0, 0, 1])
    # towards the tiny far triangle
    o.append([0, 0, 0]); d.append([50.0003, 50.0003, 50])
    o.append([50.0003, 50.0003, 60]); d.append([0, 0, -1])
    o = np.array(o, dtype=f32)
    d = np.array(d, dtype=f32)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    return o, d


# ---------------------------------------------------------------------------
# BSDF rows: roughness x wo lattice (incl. grazing and back side)
# x albedo set x LCG seeds.
# ---------------------------------------------------------------------------
ROUGHNESS = (0.0, 0.02, 0.1, 0.3, 0.5, 0.8, 1.0)
ALBEDOS = ((0.8, 0.05, 0.05), (1.0, 1.0, 1.0), (0.2, 0.5, 0.9), (0.0, 0.3, 0.9999))


def hemisphere_lattice(n_theta=7, n_phi=6, both_sides=True):
    out = []
    for it in range(n_theta):
        ct = 1.0 - (it + 0.35) / n_theta * 0.999        # from ~1 down to grazing
        st = np.sqrt(max(0.0, 1 - ct * ct))
        for ip in range(n_phi):
            ph = 2 * np.pi * (ip + 0.21 * it) / n_phi
            out.append((st * np.cos(ph), st * np.sin(ph), ct))
            if both_sides:
                out.append((st * np.cos(ph + 0.3), st * np.sin(ph + 0.3), -ct))
    out.append((0.0, 0.0, 1.0))
    out.append((0.0, 0.0, -1.0))
    d = np.array(out, dtype=np.float64)
    return (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)


def bsdf_rows(seed=1, seeds_per_case=3, with_wi=True):
    """Rows for the f / sample / pdf hooks: dict of flat arrays."""
    rng = np.random.default_rng(seed)
    wos = hemisphere_lattice()
    rows = []
    for r in ROUGHNESS:
        for a in ALBEDOS:
            for wo in wos:
                for _ in range(seeds_per_case):
                    rows.append((a, r, wo))
    n = len(rows)
    albedo = np.array([x[0] for x in rows], dtype=f32)
    rough = np.array([x[1] for x in rows], dtype=f32)
    wo = np.array([x[2] for x in rows], dtype=f32)
    seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    out = {"albedo": albedo, "roughness": rough, "wo": wo, "seed": seeds}
    if with_wi:
        wi = rng.normal(size=(n, 3))
        wi /= np.linalg.norm(wi, axis=1, keepdims=True)
        # a third of the rows near the mirror direction (where rough lobes are non-trivial)
        k = np.arange(n) % 3 == 0
        mirror = wo * np.array([-1, -1, 1], dtype=f32)
        wi[k] = mirror[k] + rng.normal(scale=0.15, size=(k.sum(), 3))
        wi /= np.linalg.norm(wi, axis=1, keepdims=True)
        out["wi"] = wi.astype(f32)
    return out


# ---------------------------------------------------------------------------
# whole-frame cases (reference-rendered goldens: tools/make_golden.py)
# ---------------------------------------------------------------------------
FRAME_W, FRAME_H, FRAME_BOUNCES, FRAME_N = 64, 36, 6, 2


def checker_texture(n=16, alpha_holes=False):
    t = np.zeros((n, n, 4), dtype=np.uint8)
    yy, xx = np.mgrid[0:n, 0:n]
    c = ((xx // 2 + yy // 2) % 2).astype(np.uint8)
    t[..., 0] = 60 + 180 * c
    t[..., 1] = 200 - 120 * c
    t[..., 2] = 90 + (xx * 9) % 160
    t[..., 3] = 255
    if alpha_holes:
        t[..., 3] = np.where((xx % 4 == 0) & (yy % 4 == 0), 40, 255)
    return t


def frame_scenes():
    """name -> scene: the plain spheres scene, a half-metal variant (AUTO mixes
    conductor and layered), and a textured variant (albedo + alpha cut-out,
    normal map, metal-rough map)."""
    plain = scenes.spheres_scene(segments=16, rings=8)
    mixed = scenes.spheres_scene(segments=16, rings=8)
    for k, me in enumerate(mixed["meshes"]):
        me["metallic"] = (k % 3) * 0.5
    tex = scenes.spheres_scene(segments=16, rings=8)
    n = 16
    yy, xx = np.mgrid[0:n, 0:n]
    nm = np.zeros((n, n, 4), dtype=np.uint8)
    nm[..., 0] = 128 + (20 * np.sin(xx * 0.8)).astype(np.int32)
    nm[..., 1] = 128 + (20 * np.cos(yy * 0.8)).astype(np.int32)
    nm[..., 2] = 235
    nm[..., 3] = 255
    mr = np.zeros((n, n, 4), dtype=np.uint8)
    mr[..., 0] = np.where((xx + yy) % 2 == 0, 255, 0)
    mr[..., 1] = 40 + 10 * ((xx * 3 + yy) % 16)
    mr[..., 3] = 255
    tex["textures"] = [checker_texture(n, alpha_holes=True), nm, mr]
    for k, me in enumerate(tex["meshes"]):
        if k % 2 == 0:
            me["tex_albedo"] = 0
        if k % 3 == 0:
            me["tex_normal"] = 1
        if k % 4 == 0:
            me["tex_metal_rough"] = 2
    return {"plain": plain, "mixed": mixed, "textured": tex}
