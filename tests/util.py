"""ctypes wrappers of the CPU checkers used by the tests.

Oracle  -> oracle/liboracle.so           (plain C++ restatement)
RefOracle -> oracle/_ref/liboracle_ref.so (the reference's own code, host build)
HostSim -> tests/hostsim/libhostsim.so   (the kernels' per-thread logic on the host)
All three are test infrastructure; the product never loads them.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
f32 = np.float32


def fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def _build_oracle():
    so = ROOT / "oracle" / "liboracle.so"
    srcs = list((ROOT / "oracle").glob("oracle_*.cpp")) + list((ROOT / "oracle").glob("*.h"))
    if not so.exists() or any(s.stat().st_mtime > so.stat().st_mtime for s in srcs):
        r = subprocess.run(["make", "-C", str(ROOT / "oracle"), "liboracle.so"], capture_output=True, text=True)
        if r.returncode != 0 and not so.exists():
            raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    return so


def _build_hostsim():
    d = ROOT / "tests" / "hostsim"
    so = d / "libhostsim.so"
    srcs = list(d.glob("*.cpp")) + list((ROOT / "optixpathtracer_b200" / "csrc").glob("*.cuh"))
    if not so.exists() or any(s.stat().st_mtime > so.stat().st_mtime for s in srcs):
        cmd = ["g++", "-std=c++17", "-O2", "-fPIC", "-fopenmp", "-ffp-contract=off", "-shared",
               "-I/usr/local/cuda/include", "-o", str(so)] + [str(p) for p in sorted(d.glob("*.cpp"))]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0 and not so.exists():
            raise RuntimeError(r.stderr)
    return so


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(str(_build_oracle()))

    def closest(self, soup, org, dirs, tmin=0.0, tmax=100.0):
        soup = np.ascontiguousarray(soup, dtype=f32).reshape(-1, 9)
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = len(o)
        prim = np.empty(n, np.int32)
        t, u, v = (np.empty(n, f32) for _ in range(3))
        bf = np.empty(n, np.uint8)
        self.lib.oracle_closest_bruteforce(fp(soup), C.c_int64(len(soup)), fp(o), fp(d), C.c_int64(n),
                                           C.c_float(tmin), C.c_float(tmax), vp(prim), fp(t), fp(u),
                                           fp(v), vp(bf))
        return prim, t, u, v, bf

    def any(self, soup, org, dirs, tmax, tmin=0.0):
        soup = np.ascontiguousarray(soup, dtype=f32).reshape(-1, 9)
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = len(o)
        tm = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=f32), (n,)), dtype=f32)
        occ = np.empty(n, np.uint8)
        self.lib.oracle_any_bruteforce(fp(soup), C.c_int64(len(soup)), fp(o), fp(d), C.c_int64(n),
                                       C.c_float(tmin), fp(tm), vp(occ))
        return occ

    def tie_count(self, soup, org, dirs, prim, tmin=0.0, tmax=100.0):
        soup = np.ascontiguousarray(soup, dtype=f32).reshape(-1, 9)
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        prim = np.ascontiguousarray(prim, dtype=np.int32)
        ties = np.empty(len(o), np.int32)
        self.lib.oracle_tie_count(fp(soup), C.c_int64(len(soup)), fp(o), fp(d), C.c_int64(len(o)),
                                  C.c_float(tmin), C.c_float(tmax), vp(prim), vp(ties))
        return ties


class HostSim:
    def __init__(self):
        self.lib = C.CDLL(str(_build_hostsim()))

    def build(self, soup):
        soup = np.ascontiguousarray(soup, dtype=f32).reshape(-1, 9)
        self._soup = soup
        n = self.lib.hostsim_build(fp(soup), len(soup))
        return n, self.lib.hostsim_check_topology()

    def closest(self, org, dirs, tmin=0.0, tmax=100.0):
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = len(o)
        prim = np.empty(n, np.int32)
        t, u, v = (np.empty(n, f32) for _ in range(3))
        bf = np.empty(n, np.uint8)
        cnt = (C.c_uint64 * 3)()
        self.lib.hostsim_trace_closest(fp(o), fp(d), C.c_int64(n), C.c_float(tmin), C.c_float(tmax),
                                       vp(prim), fp(t), fp(u), fp(v), vp(bf), cnt)
        return prim, t, u, v, bf, tuple(cnt)

    def render(self, scene, width, height, *, bsdf_mode=0, rng_mode=2, seed=0x5EED, max_bounces=8,
               frame_first=1, n_frames=1, lights=None):
        """The wavefront stages (integrator.cuh) run per pixel on the host.
        Returns (frames (n,H,W,3), {"radiance_rays", "shadow_rays"})."""
        from optixpathtracer_b200.api import _SceneMarshal
        from optixpathtracer_b200._lib import B2ptPointLight
        self.lib.hostsim_scene_create.restype = C.c_void_p
        m = _SceneMarshal(scene)
        h = C.c_void_p(self.lib.hostsim_scene_create(C.byref(m.desc)))
        try:
            pos, iv, ip = scene["camera"].launch_matrices(width, height)
            lights = scene["lights"] if lights is None else lights
            arr = (B2ptPointLight * max(len(lights), 1))()
            for i, (p, c) in enumerate(lights):
                arr[i].position = (C.c_float * 3)(*np.asarray(p, dtype=f32))
                arr[i].color = (C.c_float * 3)(*np.asarray(c, dtype=f32))
            frames = np.empty((n_frames, height, width, 3), f32)
            rays = (C.c_uint64 * 2)()
            self.lib.hostsim_render(h, C.c_int(width), C.c_int(height), fp(pos), fp(iv), fp(ip), arr,
                                    C.c_int(len(lights)), C.c_int(max_bounces), C.c_int(bsdf_mode),
                                    C.c_int(rng_mode), C.c_uint64(seed), C.c_uint32(frame_first),
                                    C.c_int(n_frames), fp(frames), rays)
        finally:
            self.lib.hostsim_scene_destroy(h)
        return frames, {"radiance_rays": rays[0], "shadow_rays": rays[1]}

    # BSDF rows: same signatures as RefOracle.bsdf_* plus the LCG flavour
    def bsdf_f(self, bsdf, albedo, roughness, wo, wi, seed=None, lcg_mode=2):
        wo = np.ascontiguousarray(wo, dtype=f32).reshape(-1, 3)
        wi = np.ascontiguousarray(wi, dtype=f32).reshape(-1, 3)
        n = len(wo)
        a, r = RefOracle._prep(n, albedo, roughness)
        sd = np.ascontiguousarray(np.broadcast_to(np.asarray(0 if seed is None else seed, dtype=np.uint32), (n,)), dtype=np.uint32)
        out = np.empty((n, 3), f32)
        so = np.empty(n, np.uint32)
        self.lib.hostsim_bsdf_f(C.c_int(bsdf), C.c_int(lcg_mode), C.c_int64(n), fp(a), fp(r), fp(wo), fp(wi),
                                vp(sd), fp(out), vp(so))
        return out, so

    def bsdf_sample(self, bsdf, albedo, roughness, wo, seed, lcg_mode=2):
        wo = np.ascontiguousarray(wo, dtype=f32).reshape(-1, 3)
        n = len(wo)
        a, r = RefOracle._prep(n, albedo, roughness)
        sd = np.ascontiguousarray(np.broadcast_to(np.asarray(seed, dtype=np.uint32), (n,)), dtype=np.uint32)
        ok = np.empty(n, np.uint8)
        f = np.empty((n, 3), f32)
        wi = np.empty((n, 3), f32)
        pdf = np.empty(n, f32)
        flags = np.empty(n, np.uint8)
        so = np.empty(n, np.uint32)
        self.lib.hostsim_bsdf_sample(C.c_int(bsdf), C.c_int(lcg_mode), C.c_int64(n), fp(a), fp(r), fp(wo), vp(sd),
                                     vp(ok), fp(f), fp(wi), fp(pdf), vp(flags), vp(so))
        return ok, f, wi, pdf, flags, so

    def bsdf_pdf(self, bsdf, roughness, wo, wi, reflection=True, transmission=True):
        wo = np.ascontiguousarray(wo, dtype=f32).reshape(-1, 3)
        wi = np.ascontiguousarray(wi, dtype=f32).reshape(-1, 3)
        n = len(wo)
        _, r = RefOracle._prep(n, (1, 1, 1), roughness)
        out = np.empty(n, f32)
        self.lib.hostsim_bsdf_pdf(C.c_int(bsdf), C.c_int64(n), fp(r), fp(wo), fp(wi), C.c_int(int(reflection)),
                                  C.c_int(int(transmission)), fp(out))
        return out

    def any(self, org, dirs, tmax, tmin=0.0):
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = len(o)
        tm = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=f32), (n,)), dtype=f32)
        occ = np.empty(n, np.uint8)
        self.lib.hostsim_trace_any(fp(o), fp(d), C.c_int64(n), C.c_float(tmin), fp(tm), vp(occ))
        return occ


def bits(a):
    return np.ascontiguousarray(a, dtype=f32).view(np.uint32)


def random_rays(rng, n, lo, hi):
    o = rng.uniform(lo, hi, (n, 3)).astype(f32)
    d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    return o, d


class RefOracle:
    """oracle/_ref/liboracle_ref.so: the reference's own device code compiled for
    the host (oracle/ref_harness.cpp).  Takes the same scene dicts / B2ptSceneDesc
    as the product."""

    PATH = ROOT / "oracle" / "_ref" / "liboracle_ref.so"

    @classmethod
    def available(cls) -> bool:
        return cls.PATH.exists()

    def __init__(self):
        self.lib = C.CDLL(str(self.PATH))
        self.lib.ref_scene_create.restype = C.c_void_p
        self.lib.ref_scene_triangles.restype = C.c_int64
        self.lib.ref_tea16.restype = C.c_uint32
        self.lib.ref_rnd.restype = C.c_float

    # -- BSDF level ---------------------------------------------------------
    @staticmethod
    def _prep(n, albedo, roughness):
        a = np.ascontiguousarray(np.broadcast_to(np.asarray(albedo, dtype=f32), (n, 3)), dtype=f32)
        r = np.ascontiguousarray(np.broadcast_to(np.asarray(roughness, dtype=f32), (n,)), dtype=f32)
        return a, r

    def bsdf_f(self, bsdf, albedo, roughness, wo, wi, seed=None):
        wo = np.ascontiguousarray(wo, dtype=f32).reshape(-1, 3)
        wi = np.ascontiguousarray(wi, dtype=f32).reshape(-1, 3)
        n = len(wo)
        a, r = self._prep(n, albedo, roughness)
        sd = np.ascontiguousarray(np.broadcast_to(np.asarray(0 if seed is None else seed, dtype=np.uint32), (n,)), dtype=np.uint32)
        out = np.empty((n, 3), f32)
        so = np.empty(n, np.uint32)
        self.lib.ref_bsdf_f(C.c_int(bsdf), C.c_int64(n), fp(a), fp(r), fp(wo), fp(wi), vp(sd), fp(out), vp(so))
        return out, so

    def bsdf_sample(self, bsdf, albedo, roughness, wo, seed):
        wo = np.ascontiguousarray(wo, dtype=f32).reshape(-1, 3)
        n = len(wo)
        a, r = self._prep(n, albedo, roughness)
        sd = np.ascontiguousarray(np.broadcast_to(np.asarray(seed, dtype=np.uint32), (n,)), dtype=np.uint32)
        ok = np.empty(n, np.uint8)
        f = np.empty((n, 3), f32)
        wi = np.empty((n, 3), f32)
        pdf = np.empty(n, f32)
        flags = np.empty(n, np.uint8)
        so = np.empty(n, np.uint32)
        self.lib.ref_bsdf_sample(C.c_int(bsdf), C.c_int64(n), fp(a), fp(r), fp(wo), vp(sd), vp(ok), fp(f), fp(wi),
                                 fp(pdf), vp(flags), vp(so))
        return ok, f, wi, pdf, flags, so

    def bsdf_pdf(self, bsdf, roughness, wo, wi, reflection=True, transmission=True):
        wo = np.ascontiguousarray(wo, dtype=f32).reshape(-1, 3)
        wi = np.ascontiguousarray(wi, dtype=f32).reshape(-1, 3)
        n = len(wo)
        _, r = self._prep(n, (1, 1, 1), roughness)
        out = np.empty(n, f32)
        self.lib.ref_bsdf_pdf(C.c_int(bsdf), C.c_int64(n), fp(r), fp(wo), fp(wi), C.c_int(int(reflection)),
                              C.c_int(int(transmission)), fp(out))
        return out

    # -- frame level --------------------------------------------------------
    def render(self, scene, width, height, *, bsdf_mode=0, max_bounces=8, frame_first=1, n_frames=1,
               lights=None, restated=False, want_frames=False):
        """Returns (running mean (H,W,3), frames (n,H,W,3) | None, counters dict)."""
        from optixpathtracer_b200.api import _SceneMarshal
        from optixpathtracer_b200._lib import B2ptPointLight
        m = _SceneMarshal(scene)
        h = C.c_void_p(self.lib.ref_scene_create(C.byref(m.desc)))
        try:
            pos, iv, ip = scene["camera"].launch_matrices(width, height)
            lights = scene["lights"] if lights is None else lights
            arr = (B2ptPointLight * max(len(lights), 1))()
            for i, (p, c) in enumerate(lights):
                arr[i].position = (C.c_float * 3)(*np.asarray(p, dtype=f32))
                arr[i].color = (C.c_float * 3)(*np.asarray(c, dtype=f32))
            mean = np.empty((height, width, 3), f32)
            frames = np.empty((n_frames, height, width, 3), f32) if want_frames else None
            cnt = (C.c_uint64 * 4)()
            self.lib.ref_render(h, C.c_int(width), C.c_int(height), fp(pos), fp(iv), fp(ip), arr,
                                C.c_int(len(lights)), C.c_int(max_bounces), C.c_int(bsdf_mode),
                                C.c_int(int(restated)), C.c_uint32(frame_first), C.c_int(n_frames), fp(mean),
                                fp(frames) if want_frames else None, cnt)
        finally:
            self.lib.ref_scene_destroy(h)
        return mean, frames, {"radiance_rays": cnt[0], "shadow_rays": cnt[1], "nan_samples": cnt[2],
                              "threads": cnt[3]}

    def camera_rays(self, camera, width, height):
        pos, iv, ip = camera.launch_matrices(width, height)
        o = np.empty((width * height, 3), f32)
        d = np.empty((width * height, 3), f32)
        self.lib.ref_camera_rays(C.c_int(width), C.c_int(height), fp(pos), fp(iv), fp(ip), fp(o), fp(d))
        return o, d


class RefHost:
    """oracle/_ref/libref_host.so: the reference's own HOST code compiled in place (oracle/ref_host.cpp) —
    ModelLoader.cpp + Mesh.cpp (glTF ingestion, N1), WriteImage.cpp (BMP / EXR writers, N3), Camera.cpp."""

    PATH = ROOT / "oracle" / "_ref" / "libref_host.so"

    @classmethod
    def available(cls) -> bool:
        return cls.PATH.exists()

    def __init__(self):
        self.lib = C.CDLL(str(self.PATH))
        self.lib.refhost_load_model.restype = C.c_void_p

    def load_model(self, path) -> dict:
        """ModelLoader::LoadModel(folder, file) -> the scene dict the product's loaders return."""
        from pathlib import Path as P
        p = P(path)
        h = C.c_void_p(self.lib.refhost_load_model(str(p.parent).encode() + b"/", p.name.encode()))
        try:
            nm, nt = C.c_int(), C.c_int()
            self.lib.refhost_model_counts(h, C.byref(nm), C.byref(nt))
            meshes, textures = [], []
            for i in range(nm.value):
                nv, ntri, nuv = C.c_int(), C.c_int(), C.c_int()
                model = np.empty(16, f32); albedo = np.empty(3, f32)
                met, rough = C.c_float(), C.c_float()
                tex = (C.c_int * 3)()
                self.lib.refhost_mesh_info(h, i, C.byref(nv), C.byref(ntri), C.byref(nuv), fp(model), fp(albedo),
                                           C.byref(met), C.byref(rough), tex)
                pos = np.empty((nv.value, 3), f32); nrm = np.empty((nv.value, 3), f32)
                uv = np.empty((nuv.value, 2), f32); idx = np.empty((ntri.value, 3), np.int32)
                self.lib.refhost_mesh_arrays(h, i, fp(pos), fp(nrm), fp(uv), vp(idx))
                meshes.append({"positions": pos, "normals": nrm, "texcoords": uv, "indices": idx,
                               "model": model.reshape(4, 4).T.copy(), "albedo": albedo, "metallic": met.value,
                               "roughness": rough.value, "tex_albedo": tex[0], "tex_normal": tex[1],
                               "tex_metal_rough": tex[2]})
            for i in range(nt.value):
                w, hh = C.c_int(), C.c_int()
                self.lib.refhost_texture_info(h, i, C.byref(w), C.byref(hh))
                px = np.empty((hh.value, w.value, 4), np.uint8)
                self.lib.refhost_texture_pixels(h, i, vp(px))
                textures.append(px)
        finally:
            self.lib.refhost_model_free(h)
        return {"meshes": meshes, "textures": textures}

    def write_exr(self, rgb, path):
        a = np.ascontiguousarray(rgb, dtype=f32)
        self.lib.refhost_write_exr(fp(a), a.shape[1], a.shape[0], str(path).encode())

    def write_bmp(self, rgb, path):
        a = np.ascontiguousarray(rgb, dtype=f32)
        self.lib.refhost_write_bmp(fp(a), a.shape[1], a.shape[0], str(path).encode())

    def camera(self, blender_pos, blender_rot, fov_deg, aspect):
        bp = np.ascontiguousarray(blender_pos, f32); br = np.ascontiguousarray(blender_rot, f32)
        pos = np.empty(3, f32); view = np.empty(16, f32); proj = np.empty(16, f32)
        self.lib.refhost_camera(fp(bp), fp(br), C.c_float(fov_deg), C.c_float(aspect), fp(pos), fp(view), fp(proj))
        return pos, view, proj

    def camera_engine(self, pos, rot, fov_deg, aspect):
        p = np.ascontiguousarray(pos, f32); r = np.ascontiguousarray(rot, f32)
        view = np.empty(16, f32); proj = np.empty(16, f32)
        self.lib.refhost_camera_engine(fp(p), fp(r), C.c_float(fov_deg), C.c_float(aspect), fp(view), fp(proj))
        return view, proj
