"""Deterministic synthetic scenes.

The README scene assets are not in the reference repository (SURVEY F6:
main.cpp:6-19 points at assets/Models/TestScenes/1/untitled.gltf, which is
absent), so Scene1 "Spheres" is re-authored here from the camera/light
constants in main.cpp:6-19 and the published PNGs: a 6x6 grid of UV-spheres
whose roughness grows left to right, a grey back wall, a green left wall and
a blue right wall, four point lights.  No RNG is involved.

A scene is a plain dict:
    {"meshes": [ {positions (nv,3) f32, normals (nv,3) f32 | None,
                  texcoords (nv,2) f32 | None, indices (nt,3) i32,
                  model (4,4) f32 M[row, col], albedo (3,), metallic, roughness,
                  tex_albedo, tex_normal, tex_metal_rough, name}, ... ],
     "textures": [ (h,w,4) uint8, ... ],
     "lights": [ (position(3), color(3)), ... ],
     "camera": Camera}
"""
from __future__ import annotations

import numpy as np

from .camera import Camera, blender_to_engine_position

f32 = np.float32


def uv_sphere(segments: int = 64, rings: int = 32):
    """Unit UV-sphere: (rings+1)*(segments+1) vertices, 2*segments*(rings-1)
    triangles (pole quads degenerate to one triangle), CCW seen from outside."""
    v = np.arange(rings + 1, dtype=np.float64) / rings
    u = np.arange(segments + 1, dtype=np.float64) / segments
    theta = v[:, None] * np.pi                  # 0 = +y pole
    phi = u[None, :] * 2.0 * np.pi
    x = np.sin(theta) * np.cos(phi)
    y = np.cos(theta) * np.ones_like(phi)
    z = np.sin(theta) * np.sin(phi)
    pos = np.stack([x, y, z], -1).reshape(-1, 3).astype(f32)
    uv = np.stack([np.broadcast_to(u[None, :], x.shape), np.broadcast_to(v[:, None], x.shape)],
                  -1).reshape(-1, 2).astype(f32)
    tris = []
    stride = segments + 1
    for r in range(rings):
        for s in range(segments):
            a = r * stride + s
            b = a + 1
            c = a + stride
            d = c + 1
            if r != 0:
                tris.append((a, b, c))
            if r != rings - 1:
                tris.append((b, d, c))
    idx = np.array(tris, dtype=np.int32)
    # make the winding counter-clockwise seen from outside
    p0, p1, p2 = pos[idx[:, 0]], pos[idx[:, 1]], pos[idx[:, 2]]
    n = np.cross(p1 - p0, p2 - p0)
    flip = (n * (p0 + p1 + p2)).sum(-1) < 0
    idx[flip] = idx[flip][:, [0, 2, 1]]
    return pos, pos.copy(), uv, idx


def _trs(translation=(0, 0, 0), scale=(1, 1, 1)) -> np.ndarray:
    m = np.eye(4, dtype=f32)
    m[0, 0], m[1, 1], m[2, 2] = scale
    m[:3, 3] = translation
    return m


def _mesh(pos, nrm, uv, idx, model, albedo, metallic, roughness, name):
    return {"positions": np.ascontiguousarray(pos, dtype=f32),
            "normals": None if nrm is None else np.ascontiguousarray(nrm, dtype=f32),
            "texcoords": None if uv is None else np.ascontiguousarray(uv, dtype=f32),
            "indices": np.ascontiguousarray(idx, dtype=np.int32),
            "model": np.asarray(model, dtype=f32),
            "albedo": np.asarray(albedo, dtype=f32), "metallic": float(metallic),
            "roughness": float(roughness), "tex_albedo": -1, "tex_normal": -1,
            "tex_metal_rough": -1, "name": name}


def quad(p0, p1, p2, p3, normal):
    """Two triangles p0-p1-p2, p0-p2-p3, wound CCW seen from `normal`'s side."""
    pos = np.array([p0, p1, p2, p3], dtype=f32)
    n = np.cross(pos[1] - pos[0], pos[2] - pos[0])
    idx = np.array([[0, 1, 2], [0, 2, 3]], dtype=np.int32)
    if np.dot(n, normal) < 0:
        idx = idx[:, [0, 2, 1]]
    nrm = np.tile(np.asarray(normal, dtype=f32), (4, 1))
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], dtype=f32)
    return pos, nrm, uv, idx


SPHERE_RADIUS = 0.205
SPHERE_PITCH = 0.467
BACK_WALL_X = -0.6
SIDE_WALL_Z = 2.376


def scene1_camera() -> Camera:
    cam = Camera()
    cam.set_blender_position((3.85382, 0.0, 1.0))      # main.cpp:10
    cam.set_blender_rotation((90.0, 0.0, 90.0))        # main.cpp:11
    return cam


def scene1_lights(color: float = 1.0):
    c = np.array([color] * 3, dtype=f32)
    return [(blender_to_engine_position((1.33906, -0.7, 0.299367)), c.copy()),   # main.cpp:14-17
            (blender_to_engine_position((1.33906, 0.7, 0.299367)), c.copy()),
            (blender_to_engine_position((1.33906, 0.7, 1.69937)), c.copy()),
            (blender_to_engine_position((1.33906, -0.7, 1.69937)), c.copy())]


def spheres_scene(metallic: float = 0.0, segments: int = 64, rings: int = 32,
                  light_color: float = 1.0, sphere_albedo=(0.8, 0.05, 0.05)):
    """Scene1 "Spheres": 36 spheres + 3 walls = 39 meshes, 142 854 triangles at
    the default tessellation."""
    pos, nrm, uv, idx = uv_sphere(segments, rings)
    meshes = []
    for r in range(6):          # rows, bottom to top
        for c in range(6):      # columns, image left (+z) to right (-z)
            y = 1.0 + (r - 2.5) * SPHERE_PITCH
            z = (2.5 - c) * SPHERE_PITCH
            meshes.append(_mesh(pos, nrm, uv, idx, _trs((0.0, y, z), (SPHERE_RADIUS,) * 3),
                                sphere_albedo, metallic, c / 5.0, f"sphere_r{r}_c{c}"))
    ylo, yhi, xfront = -2.0, 4.0, 5.0
    bx, sz = BACK_WALL_X, SIDE_WALL_Z
    walls = [
        (quad((bx, ylo, -sz), (bx, ylo, sz), (bx, yhi, sz), (bx, yhi, -sz), (1, 0, 0)),
         (0.8, 0.8, 0.8), "wall_back"),
        (quad((bx, ylo, sz), (xfront, ylo, sz), (xfront, yhi, sz), (bx, yhi, sz), (0, 0, -1)),
         (0.05, 0.8, 0.05), "wall_left_green"),
        (quad((bx, ylo, -sz), (xfront, ylo, -sz), (xfront, yhi, -sz), (bx, yhi, -sz), (0, 0, 1)),
         (0.05, 0.05, 0.8), "wall_right_blue"),
    ]
    for (wp, wn, wuv, wi), alb, name in walls:
        meshes.append(_mesh(wp, wn, wuv, wi, np.eye(4, dtype=f32), alb, metallic, 0.5, name))
    return {"meshes": meshes, "textures": [], "lights": scene1_lights(light_color),
            "camera": scene1_camera()}


def textured_spheres_scene(n: int = 64, **kw):
    """The README scene with what a glTF export of it usually carries: every other sphere has an albedo texture with
    alpha cut-out holes (the any-hit predicate, devicePrograms.cu:518-561), every third a normal map, every fourth
    a metal-rough map; half of the spheres are metallic, so AUTO mode mixes conductor and layered."""
    sc = spheres_scene(**kw)
    yy, xx = np.mgrid[0:n, 0:n]
    alb = np.zeros((n, n, 4), dtype=np.uint8)
    c = ((xx // 4 + yy // 4) % 2).astype(np.uint8)
    alb[..., 0] = 60 + 180 * c
    alb[..., 1] = 200 - 120 * c
    alb[..., 2] = 90 + (xx * 3) % 160
    alb[..., 3] = np.where((xx % 8 < 2) & (yy % 8 < 2), 40, 255)   # 1/16 of the surface is cut out
    nm = np.zeros((n, n, 4), dtype=np.uint8)
    nm[..., 0] = 128 + (20 * np.sin(xx * 0.4)).astype(np.int32)
    nm[..., 1] = 128 + (20 * np.cos(yy * 0.4)).astype(np.int32)
    nm[..., 2] = 235
    nm[..., 3] = 255
    mr = np.zeros((n, n, 4), dtype=np.uint8)
    mr[..., 0] = np.where((xx // 4 + yy // 4) % 2 == 0, 255, 0)
    mr[..., 1] = 40 + 10 * ((xx * 3 + yy) % 16)
    mr[..., 3] = 255
    sc["textures"] = [alb, nm, mr]
    for k, me in enumerate(sc["meshes"][:36]):
        me["metallic"] = float(k % 2)
        if k % 2 == 0:
            me["tex_albedo"] = 0
        if k % 3 == 0:
            me["tex_normal"] = 1
        if k % 4 == 0:
            me["tex_metal_rough"] = 2
    return sc


def halfmetal_spheres_scene(**kw):
    """textured_spheres_scene without its textures: the untextured AUTO-mode comparison point."""
    sc = spheres_scene(**kw)
    for k, me in enumerate(sc["meshes"][:36]):
        me["metallic"] = float(k % 2)
    return sc


def _value_noise(nx, ny, seed, octaves=3):
    """SplitMix64-seeded value noise on an (ny, nx) grid, deterministic."""
    def splitmix(n, s):
        out = np.empty(n, dtype=np.uint64)
        x = np.uint64(s)
        with np.errstate(over="ignore"):
            for i in range(n):
                x = x + np.uint64(0x9E3779B97F4A7C15)
                z = x
                z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
                z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
                out[i] = z ^ (z >> np.uint64(31))
        return (out >> np.uint64(11)).astype(np.float64) / float(1 << 53)
    h = np.zeros((ny, nx), dtype=np.float64)
    amp, cells = 1.0, 8
    for o in range(octaves):
        g = splitmix((cells + 1) * (cells + 1), seed + o).reshape(cells + 1, cells + 1)
        xs = np.linspace(0, cells, nx, endpoint=True)
        ys = np.linspace(0, cells, ny, endpoint=True)
        x0 = np.minimum(xs.astype(int), cells - 1)
        y0 = np.minimum(ys.astype(int), cells - 1)
        fx = xs - x0
        fy = ys - y0
        fx = fx * fx * (3 - 2 * fx)
        fy = fy * fy * (3 - 2 * fy)
        a = g[y0][:, x0]
        b = g[y0][:, x0 + 1]
        c = g[y0 + 1][:, x0]
        d = g[y0 + 1][:, x0 + 1]
        h += amp * ((a * (1 - fx) + b * fx) * (1 - fy)[:, None] + (c * (1 - fx) + d * fx) * fy[:, None])
        amp *= 0.5
        cells *= 2
    return h


def terrain_scene(patches: int = 4, n: int = 256, seed: int = 1234, with_spheres: bool = True):
    """C5: patches x patches displaced height-field patches of n x n vertices
    (default 16 x 130 050 = 2 080 800 triangles), checkerboard metallic,
    roughness 0.1 + 0.05*patch, 36 spheres floating above, 4 lights of colour 50."""
    meshes = []
    size = 8.0 / patches
    full = _value_noise(patches * (n - 1) + 1, patches * (n - 1) + 1, seed)
    ii, jj = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
    a = (ii * n + jj).ravel()
    idx = np.concatenate([np.stack([a, a + 1, a + n], -1), np.stack([a + 1, a + n + 1, a + n], -1)]).astype(np.int32)
    for pi in range(patches):
        for pj in range(patches):
            hh = full[pi * (n - 1): pi * (n - 1) + n, pj * (n - 1): pj * (n - 1) + n]
            xs = -4.0 + pi * size + np.arange(n) * (size / (n - 1))
            zs = -4.0 + pj * size + np.arange(n) * (size / (n - 1))
            X, Z = np.meshgrid(xs, zs, indexing="ij")
            Y = -1.0 + 0.8 * hh
            pos = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(f32)
            gy_x, gy_z = np.gradient(Y, xs, zs)
            nrm = np.stack([-gy_x, np.ones_like(Y), -gy_z], -1)
            nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
            tri = idx.copy()
            p0, p1, p2 = pos[tri[:, 0]], pos[tri[:, 1]], pos[tri[:, 2]]
            flip = np.cross(p1 - p0, p2 - p0)[:, 1] < 0
            tri[flip] = tri[flip][:, [0, 2, 1]]
            pid = pi * patches + pj
            meshes.append(_mesh(pos, nrm.reshape(-1, 3), None, tri, np.eye(4, dtype=f32),
                                (0.7, 0.7, 0.65), float((pi + pj) % 2), 0.1 + 0.05 * pid,
                                f"terrain_{pi}_{pj}"))
    if with_spheres:
        spos, snrm, suv, sidx = uv_sphere(64, 32)
        for r in range(6):
            for c in range(6):
                meshes.append(_mesh(spos, snrm, suv, sidx,
                                    _trs((-2.5 + r, 1.0 + 0.3 * ((r + c) % 3), -2.5 + c), (0.35,) * 3),
                                    (0.8, 0.3, 0.1), float((r + c) % 2), c / 5.0, f"sphere_{r}_{c}"))
    cam = Camera()
    d = 11.0
    cam.position = np.array([d * np.cos(np.radians(30.0)), d * np.sin(np.radians(30.0)), 0.0], dtype=f32)
    cam.rotation = np.array([30.0, 270.0, 0.0], dtype=f32)
    c = np.array([50.0] * 3, dtype=f32)
    lights = [(np.array(p, dtype=f32), c.copy()) for p in
              [(3.0, 4.0, 3.0), (-3.0, 4.0, 3.0), (-3.0, 4.0, -3.0), (3.0, 4.0, -3.0)]]
    return {"meshes": meshes, "textures": [], "lights": lights, "camera": cam}


def scene_triangle_soup(scene) -> tuple[np.ndarray, np.ndarray]:
    """World-space triangle soup (nt,3,3) fp32 in global primitive order, with
    glm's mat4*vec4 association ((m0*x + m1*y) + (m2*z + m3*1)), plus the mesh
    id of every triangle — the same arithmetic as k_flatten."""
    tris, mesh_ids = [], []
    for m, me in enumerate(scene["meshes"]):
        M = me["model"].astype(f32)
        p = me["positions"].astype(f32)
        w = (M[None, :3, 0] * p[:, 0:1] + M[None, :3, 1] * p[:, 1:2]) + \
            (M[None, :3, 2] * p[:, 2:3] + M[None, :3, 3] * f32(1.0))
        w = w.astype(f32)
        tris.append(w[me["indices"]])
        mesh_ids.append(np.full(len(me["indices"]), m, dtype=np.int32))
    if not tris:
        return np.zeros((0, 3, 3), dtype=f32), np.zeros(0, dtype=np.int32)
    return np.ascontiguousarray(np.concatenate(tris)), np.concatenate(mesh_ids)
