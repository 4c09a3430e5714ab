"""GPU tier: K1-K3 (LBVH build) + K5/K8 (traversal) against the oracle.

Everything goes through the C-ABI (b2pt_create / b2pt_trace_closest /
b2pt_trace_any).  Bar: primitive ids, t, u, v bit-exact; occlusion exact.
"""
import os
import numpy as np
import pytest

from tests import cases
from tests.util import bits, random_rays
from optixpathtracer_b200 import scenes

pytestmark = pytest.mark.gpu
f32 = np.float32


def _renderer(scene, **kw):
    from optixpathtracer_b200 import OptixRenderer
    return OptixRenderer(None, scene, **kw)


def _soup_scene(soup):
    soup = np.asarray(soup, dtype=f32).reshape(-1, 3, 3)
    return {"meshes": [{"positions": soup.reshape(-1, 3), "normals": None, "texcoords": None,
                        "indices": np.arange(len(soup) * 3, dtype=np.int32).reshape(-1, 3),
                        "model": np.eye(4, dtype=f32), "albedo": (1, 1, 1), "metallic": 0.0,
                        "roughness": 0.5}] if len(soup) else [], "textures": []}


def _compare(r, oracle, soup, o, d, tmin=0.0, tmax=100.0):
    gp, gm, gt, gu, gv = r.TraceClosest(o, d, tmin, tmax)
    op, ot, ou, ov, obf = oracle.closest(soup, o, d, tmin, tmax)
    assert np.array_equal(gp, op), f"{(gp != op).sum()} primitive ids differ"
    assert np.array_equal(bits(gt), bits(ot))
    assert np.array_equal(bits(gu), bits(ou))
    assert np.array_equal(bits(gv), bits(ov))
    return op, ot, gm


def test_bvh_structure():
    sc = cases.small_spheres()
    with _renderer(sc) as r:
        st = r.Stats()
        nt = st["n_triangles"]
        assert nt == sum(len(m["indices"]) for m in sc["meshes"])
        assert st["n_nodes"] == nt - 1 and st["node_bytes"] in (64, 96, 128) and st["triangle_bytes"] == 48   # 96: the packed 4-wide nodes (default)
        nodes, order = r.DownloadBvh()
        assert sorted(order.tolist()) == list(range(nt))          # a permutation
        child = nodes.view(np.int32)[:, 12:14]
        seen = np.zeros(nt, dtype=np.int32)
        visited = 0
        stack = [0]
        while stack:
            assert visited < nt, "the child links do not form a tree"
            c = stack.pop()
            if c < 0:
                seen[~c] += 1
                continue
            visited += 1
            stack += [int(child[c, 0]), int(child[c, 1])]
        assert visited == nt - 1 and (seen == 1).all()
        # child boxes contain their subtree's triangles: check the root's two boxes cover the scene
        soup, _ = scenes.scene_triangle_soup(sc)
        lo = soup.reshape(-1, 3).min(0)
        hi = soup.reshape(-1, 3).max(0)
        root = nodes[0]
        blo = np.array([min(root[0], root[4]), min(root[2], root[6]), min(root[8], root[10])])
        bhi = np.array([max(root[1], root[5]), max(root[3], root[7]), max(root[9], root[11])])
        assert (blo <= lo).all() and (bhi >= hi).all()


def test_scene1_closest_and_any(oracle):
    sc = cases.small_spheres()
    soup, mesh_ids = scenes.scene_triangle_soup(sc)
    o, d = cases.scene_rays(sc, 96, 64, 6000, seed=5)
    with _renderer(sc) as r:
        prim, t, gm = _compare(r, oracle, soup, o, d)
        hit = prim >= 0
        assert np.array_equal(gm[hit], mesh_ids[prim[hit]]) and (gm[~hit] == -1).all()
        tm = np.where(hit, t * np.where(np.arange(len(t)) % 2 == 0, f32(0.999), f32(1.001)), f32(5.0)).astype(f32)
        assert np.array_equal(r.TraceAny(o, d, tm), oracle.any(soup, o, d, tm))
        _compare(r, oracle, soup, o, d, tmin=1.0, tmax=4.0)


def test_edge_cases(oracle):
    soup = cases.edge_soup()
    o, d = cases.edge_rays()
    with _renderer(_soup_scene(soup)) as r:
        _compare(r, oracle, soup, o, d)
        _compare(r, oracle, soup, o, d, tmin=0.5, tmax=2.5)
        assert np.array_equal(r.TraceAny(o, d, 100.0), oracle.any(soup, o, d, 100.0))


def test_tiny_and_empty_scenes(oracle):
    rng = np.random.default_rng(3)
    o, d = random_rays(rng, 500, [-2, -2, -2], [2, 2, 2])
    for nt in (1, 2, 3, 7):
        soup = rng.uniform(-1, 1, (nt, 3, 3)).astype(f32)
        with _renderer(_soup_scene(soup)) as r:
            _compare(r, oracle, soup, o, d)
    with _renderer(_soup_scene(np.zeros((0, 3, 3)))) as r:
        prim, *_ = r.TraceClosest(o, d)
        assert (prim == -1).all()
        assert (r.TraceAny(o, d, 10.0) == 0).all()
        assert r.Stats()["n_nodes"] == 0


def test_rays_through_mesh_vertices_and_edges(oracle):
    """The deferred K5 orders hits by fp32 intervals and falls back to fp64 when intervals overlap: aim
    rays at the shared vertices and edge midpoints of a closed tessellated mesh (and a hair beside them),
    where two to six triangles are hit at the same distance, and through two coincident shells."""
    from optixpathtracer_b200 import scenes as S
    pos, _, _, idx = S.uv_sphere(24, 12)
    tri = pos[idx].astype(f32)                                   # (n, 3, 3)
    soup = np.concatenate([tri, tri * f32(1.0), tri[::-1] * f32(2.0)]).astype(f32)   # duplicate shell + outer shell
    rng = np.random.default_rng(21)
    verts = pos.astype(np.float64)
    mids = 0.5 * (tri[:, 0].astype(np.float64) + tri[:, 1].astype(np.float64))
    targets = np.concatenate([verts, mids, verts + rng.normal(scale=1e-7, size=verts.shape),
                              mids + rng.normal(scale=1e-7, size=mids.shape)])
    org = rng.normal(size=targets.shape)
    org = 5.0 * org / np.linalg.norm(org, axis=1, keepdims=True)
    inside = rng.random(len(targets)) < 0.3                     # some rays start inside (back faces first)
    org[inside] *= 0.02
    d = targets - org
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o, d = org.astype(f32), d.astype(f32)
    with _renderer(_soup_scene(soup)) as r:
        _compare(r, oracle, soup, o, d, 0.0, 100.0)
        _compare(r, oracle, soup, o, d, 0.5, 5.2)
        tm = np.full(len(o), 4.5, f32)
        assert np.array_equal(r.TraceAny(o, d, tm), oracle.any(soup, o, d, tm))


@pytest.mark.parametrize("leaf", [1, 2, 3, 5, 8])
def test_multi_triangle_leaves(oracle, leaf):
    """B2ptConfig::bvh_leaf_size: a leaf of the wide tree names a run of up to `leaf` triangle records.  Same hits,
    same bits, whatever the run length — on a closed mesh with coincident shells (several triangles of one run hit at
    the same distance) and on random soup; fewer wide nodes the longer the runs."""
    if os.environ.get("B2PT_LEAF"):
        pytest.skip("B2PT_LEAF overrides the config field")
    from optixpathtracer_b200 import scenes as S
    pos, _, _, idx = S.uv_sphere(20, 10)
    tri = pos[idx].astype(f32)
    rng = np.random.default_rng(77)
    soup = np.concatenate([tri, tri, tri[::-1] * f32(1.7), rng.uniform(-2, 2, (300, 3, 3)).astype(f32)]).astype(f32)
    o, d = random_rays(rng, 5000, [-3, -3, -3], [3, 3, 3])
    verts = pos.astype(np.float64)
    org = rng.normal(size=verts.shape)
    org = 4.0 * org / np.linalg.norm(org, axis=1, keepdims=True)
    dv = verts - org
    dv /= np.linalg.norm(dv, axis=1, keepdims=True)
    o = np.concatenate([o, org.astype(f32)])
    d = np.concatenate([d, dv.astype(f32)])
    with _renderer(_soup_scene(soup), bvh_leaf_size=leaf) as r:
        _compare(r, oracle, soup, o, d)
        _compare(r, oracle, soup, o, d, tmin=0.8, tmax=4.1)
        tm = np.full(len(o), 3.9, f32)
        assert np.array_equal(r.TraceAny(o, d, tm), oracle.any(soup, o, d, tm))
        n_wide = r.Stats()["n_wide_nodes"]
    with _renderer(_soup_scene(soup), bvh_leaf_size=1) as r1:
        assert n_wide <= r1.Stats()["n_wide_nodes"] and (leaf == 1 or n_wide < r1.Stats()["n_wide_nodes"])


def test_opaque_albedo_texture_traces_like_an_untextured_mesh(oracle):
    """The any-hit predicate (devicePrograms.cu:518-561) can only reject where the albedo texture's alpha is below 0.9:
    a texture whose texels all carry alpha 255 accepts every hit, so its meshes take the traversal path of untextured
    ones (TriRecord flag, bvh_build.cu) — same hits as the brute force WITHOUT a predicate, bit for bit.  One texel
    below 255 and the mesh is a cut-out mesh again: rays through that texel's area change their hit."""
    sc = cases.small_spheres()
    soup, _ = scenes.scene_triangle_soup(sc)
    o, d = cases.scene_rays(sc, 64, 48, 3000, seed=9)
    tex = np.full((8, 8, 4), 255, np.uint8)
    tex[..., :3] = 120
    for me in sc["meshes"]:
        me["tex_albedo"] = 0
        if me.get("texcoords") is None:
            me["texcoords"] = np.zeros((len(me["positions"]), 2), f32)
    sc["textures"] = [tex]
    with _renderer(sc) as r:
        prim_opaque, *_ = _compare(r, oracle, soup, o, d)
        tm = np.full(len(o), 6.0, f32)
        assert np.array_equal(r.TraceAny(o, d, tm), oracle.any(soup, o, d, tm))
    holes = tex.copy()
    holes[..., 3] = 0                          # fully transparent: every hit of a textured mesh is ignored
    sc["textures"] = [holes]
    with _renderer(sc) as r:
        gp, *_ = r.TraceClosest(o, d, 0.0, 100.0)
        assert (gp == -1).all() and (prim_opaque >= 0).any()


def test_far_origins_small_triangles(oracle):
    rng = np.random.default_rng(11)
    c = rng.uniform(-1, 1, (4000, 1, 3))
    soup = (c + rng.normal(scale=0.01, size=(4000, 3, 3))).astype(f32)
    tgt = soup.reshape(-1, 3)[rng.integers(0, 12000, 4000)]
    o = (tgt + rng.normal(size=(4000, 3)) * 30).astype(f32)
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    with _renderer(_soup_scene(soup)) as r:
        _compare(r, oracle, soup, o, d, 0.0, 1000.0)


def test_model_matrix_and_mesh_offsets(oracle):
    """Several meshes with non-trivial transforms: the GPU flatten (k_flatten)
    must produce the same fp32 world-space soup as the numpy statement."""
    rng = np.random.default_rng(21)
    meshes = []
    for k in range(5):
        pos = rng.uniform(-1, 1, (30, 3)).astype(f32)
        idx = rng.integers(0, 30, (40, 3)).astype(np.int32)
        M = np.eye(4, dtype=f32)
        M[:3, :3] = (np.linalg.qr(rng.normal(size=(3, 3)))[0] * rng.uniform(0.3, 2.0)).astype(f32)
        M[:3, 3] = rng.uniform(-2, 2, 3)
        meshes.append({"positions": pos, "normals": None, "texcoords": None, "indices": idx,
                       "model": M, "albedo": (1, 1, 1), "metallic": 0.0, "roughness": 0.5})
    sc = {"meshes": meshes, "textures": []}
    soup, _ = scenes.scene_triangle_soup(sc)
    o, d = random_rays(rng, 4000, [-4, -4, -4], [4, 4, 4])
    with _renderer(sc) as r:
        _compare(r, oracle, soup, o, d)


def test_stats_counters():
    sc = cases.small_spheres()
    o, d = cases.scene_rays(sc, 64, 64, 0, seed=1)
    with _renderer(sc, collect_stats=True) as r:
        r.TraceClosest(o, d)
        st = r.Stats()
        assert st["radiance_rays"] == len(o)
        assert st["nodes_visited"] > len(o) and st["triangles_tested"] > 0
        assert st["exact_tests"] <= st["triangles_tested"]


# ---- deep and degenerate trees: the build measures the depth and checks it against the traversal stacks ------------
def _depth_checked(r, n_tris):
    st = r.Stats()
    assert 1 <= st["bvh_wide_depth"] <= st["bvh_depth"] <= 127
    assert 3 * st["bvh_wide_depth"] + 1 <= 192, "wide traversal stack (kWideStackSize)"
    assert 1 <= st["n_wide_nodes"] <= max(st["n_nodes"], 1)
    return st


def test_many_coincident_triangles(oracle):
    """10^5 triangles with ONE centroid cell (exact duplicates and near-duplicates): the Karras build splits them on
    the index bits — log2(n) extra levels under the deepest Morton prefix.  Exact ties: the lowest primitive id wins."""
    rng = np.random.default_rng(5)
    base = np.array([[0.0, 0.0, 0.0], [1e-3, 0.0, 0.0], [0.0, 1e-3, 0.0]], f32)
    n = 100_000
    soup = np.tile(base, (n, 1, 1)).astype(f32)
    soup[n // 2:] += rng.normal(scale=1e-9, size=(n - n // 2, 3, 3)).astype(f32)      # near-duplicates
    far = rng.uniform(-1, 1, (64, 3, 3)).astype(f32)                                  # a few ordinary triangles around
    soup = np.concatenate([soup, far])
    o = np.concatenate([rng.uniform(0, 1e-3, (256, 3)) * [1, 1, 0] + [0, 0, 1.0], rng.uniform(-1, 1, (256, 3))]).astype(f32)
    d = np.concatenate([np.tile([0, 0, -1.0], (256, 1)), rng.normal(size=(256, 3))])
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    with _renderer(_soup_scene(soup)) as r:
        st = _depth_checked(r, len(soup))
        assert st["bvh_depth"] >= 17                  # log2(10^5) levels of index tie-breaks
        prim, _, _ = _compare(r, oracle, soup, o, d)
        assert (prim[:256] >= 0).sum() > 50
        tm = np.full(len(o), 50.0, f32)
        assert np.array_equal(r.TraceAny(o, d, tm), oracle.any(soup, o, d, tm))


def test_teapot_in_a_stadium(oracle):
    """A 10^-6-scale tessellated sphere inside a unit room: all of its triangles share the last Morton cells, the
    room's triangles sit 20 octaves above."""
    pos, _, _, idx = scenes.uv_sphere(32, 16)
    tiny = (pos[idx].astype(np.float64) * 1e-6 + [0.3, 0.3, 0.3]).astype(f32)
    room = []
    for axis in range(3):
        for side in (0.0, 1.0):
            q = np.zeros((4, 3))
            a, b = [k for k in range(3) if k != axis]
            q[:, axis] = side
            q[:, a] = [0, 1, 1, 0]
            q[:, b] = [0, 0, 1, 1]
            room += [q[[0, 1, 2]], q[[0, 2, 3]]]
    soup = np.concatenate([tiny, np.array(room, f32)]).astype(f32)
    rng = np.random.default_rng(9)
    o1 = rng.uniform(0.05, 0.95, (512, 3))
    tgt = tiny.reshape(-1, 3)[rng.integers(0, tiny.shape[0] * 3, 512)].astype(np.float64)
    d1 = tgt + rng.normal(scale=3e-7, size=tgt.shape) - o1
    o2, d2 = random_rays(rng, 512, [0.05] * 3, [0.95] * 3)
    o = np.concatenate([o1, o2]).astype(f32)
    d = np.concatenate([d1 / np.linalg.norm(d1, axis=1, keepdims=True), d2]).astype(f32)
    with _renderer(_soup_scene(soup)) as r:
        _depth_checked(r, len(soup))
        prim, _, _ = _compare(r, oracle, soup, o, d)
        assert (prim[:512] < len(tiny)).sum() > 100      # the rays aimed at the tiny sphere do find it


def test_sliver_chain(oracle):
    """A geometric chain of slivers: centroids at 2^-k, every Morton level splits one triangle off."""
    t = []
    for k in range(1, 40):
        x = 2.0 ** -k
        t.append([[x, 0.0, 0.0], [x * 1.5, 1e-4 * x, 0.0], [x, 0.0, 1.0]])       # long thin triangles
        t.append([[0.0, x, 0.0], [1e-4 * x, x * 1.5, 0.0], [0.0, x, 1.0]])
    soup = np.array(t, f32)
    rng = np.random.default_rng(2)
    o, d = random_rays(rng, 2000, [-0.2, -0.2, -0.2], [1.2, 1.2, 1.2])
    tgt = soup.reshape(-1, 3)[rng.integers(0, soup.shape[0] * 3, 1000)]
    o2 = (tgt + rng.normal(scale=0.3, size=tgt.shape)).astype(f32)
    d2 = tgt - o2
    d2 = (d2 / np.linalg.norm(d2, axis=1, keepdims=True)).astype(f32)
    o, d = np.concatenate([o, o2]), np.concatenate([d, d2])
    with _renderer(_soup_scene(soup)) as r:
        _depth_checked(r, len(soup))
        _compare(r, oracle, soup, o, d)


DEEP_PROG = r'''
import sys, numpy as np
sys.path.insert(0, %r)
from tests.test_trace_gpu import _deep_textured_scene, _renderer
from optixpathtracer_b200._lib import B2ptError
try:
    with _renderer(_deep_textured_scene()) as r:
        st = r.Stats()
        assert st["bvh_depth"] + 1 <= 64, st["bvh_depth"]
        print("DEEP built", st["bvh_depth"])
except B2ptError as e:
    assert e.code == -5 and "too deep" in str(e), str(e)
    print("DEEP refused")
'''


def _deep_textured_scene():
    """A sliver chain down the Morton levels with 2^10 coincident triangles at its end, on a textured mesh."""
    t = []
    for k in range(1, 22):
        x = 2.0 ** -k
        for axis in range(3):
            p = np.zeros(3)
            p[axis] = x
            t.append([p, p + [1e-9, 0, 0], p + [0, 1e-9, 0]])
    t += [[[0, 0, 0], [1e-9, 0, 0], [0, 1e-9, 0]]] * 1024
    soup = np.array(t, f32)
    sc = _soup_scene(soup)
    sc["meshes"][0]["texcoords"] = np.zeros((len(soup) * 3, 2), f32)
    sc["meshes"][0]["tex_albedo"] = 0
    tex = np.full((2, 2, 4), 255, np.uint8)
    tex[1, 1, 3] = 0          # one transparent texel: a cut-out texture (an opaque one would trace like no texture at all)
    sc["textures"] = [tex]
    return sc


def test_too_deep_for_the_binary_kernels_is_refused_not_overrun():
    """The binary-node kernels (B2PT_BVH4=0) have a 64-entry stack; the binary tree of this scene is deeper than that:
    b2pt_create must say so (B2PT_E_UNSUPPORTED) — or build a tree that fits; it must never build one that does not.
    The default (wide) kernels hold it: 3 entries per wide level, 192 entries."""
    import subprocess, sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    env = {k: v for k, v in os.environ.items() if not k.startswith("B2PT_")}
    env.update({"B2PT_BVH4": "0", "B2PT_NO_REBUILD": "1"})
    p = subprocess.run([sys.executable, "-c", DEEP_PROG % str(root)], env=env, capture_output=True, text=True, timeout=600, cwd=str(root))
    assert p.returncode == 0 and "DEEP " in p.stdout, p.stderr[-2000:]
    with _renderer(_deep_textured_scene()) as r:
        st = r.Stats()
        assert 3 * st["bvh_wide_depth"] + 1 <= 192
