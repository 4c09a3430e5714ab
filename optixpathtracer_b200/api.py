"""Python mirror of the reference's renderer surface, over the C-ABI.

`OptixRenderer` keeps the reference's method names and argument meaning
(Renderer/OptiX/OptixRenderer.h:63-76: ctor(ptxPath, model), Resize, Render,
SetCamera, SetLights, SetMaxBounces) so tests read like calls into the
reference; every method is a thin ctypes call into libb2pt.so.  Nothing here
computes: without the CUDA library the import of `_lib` fails loudly.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import (B2ptConfig, B2ptMesh, B2ptPointLight, B2ptSceneDesc, B2ptStats, B2ptTexture,
                   BSDF_NAMES, RNG_NAMES, check)

f32 = np.float32


def _fp(a):
    return a.ctypes.data_as(_lib.c_float_p)


class _SceneMarshal:
    """Keeps the numpy arrays alive while the C structs point into them."""

    def __init__(self, scene: dict):
        meshes = scene.get("meshes", [])
        textures = scene.get("textures", [])
        self.keep = []
        self.c_meshes = (B2ptMesh * max(len(meshes), 1))()
        for i, me in enumerate(meshes):
            pos = np.ascontiguousarray(me["positions"], dtype=f32).reshape(-1, 3)
            idx = np.ascontiguousarray(me["indices"], dtype=np.int32).reshape(-1, 3)
            nrm = me.get("normals")
            uv = me.get("texcoords")
            cm = self.c_meshes[i]
            cm.positions = _fp(pos)
            self.keep += [pos, idx]
            if nrm is not None:
                nrm = np.ascontiguousarray(nrm, dtype=f32).reshape(-1, 3)
                cm.normals = _fp(nrm)
                self.keep.append(nrm)
            if uv is not None:
                uv = np.ascontiguousarray(uv, dtype=f32).reshape(-1, 2)
                cm.texcoords = _fp(uv)
                self.keep.append(uv)
            cm.indices = idx.ctypes.data_as(_lib.c_int32_p)
            cm.n_vertices = pos.shape[0]
            cm.n_triangles = idx.shape[0]
            model = np.asarray(me.get("model", np.eye(4)), dtype=f32).reshape(4, 4)
            cm.model = (C.c_float * 16)(*model.T.reshape(16))       # column-major
            cm.albedo = (C.c_float * 3)(*np.asarray(me.get("albedo", (1, 1, 1)), dtype=f32))
            cm.metallic = float(me.get("metallic", 0.0))
            cm.roughness = float(me.get("roughness", 0.5))
            cm.tex_albedo = int(me.get("tex_albedo", -1))
            cm.tex_normal = int(me.get("tex_normal", -1))
            cm.tex_metal_rough = int(me.get("tex_metal_rough", -1))
        self.c_textures = (B2ptTexture * max(len(textures), 1))()
        for i, t in enumerate(textures):
            t = np.ascontiguousarray(t, dtype=np.uint8)
            assert t.ndim == 3 and t.shape[2] == 4
            self.keep.append(t)
            self.c_textures[i].rgba = t.ctypes.data_as(_lib.c_uint8_p)
            self.c_textures[i].height, self.c_textures[i].width = t.shape[0], t.shape[1]
        self.desc = B2ptSceneDesc(self.c_meshes, len(meshes), self.c_textures, len(textures))


class OptixRenderer:
    """Drop-in for the reference class of the same name.

    `model` is a scene dict (see scenes.py) standing for the reference's
    `Model*`; `ptx_path` is accepted and ignored (there is no PTX to JIT).
    """

    def __init__(self, ptx_path: str | None, model: dict, *, device: int = 0,
                 bsdf_mode="auto", rng_mode="philox", seed: int = 0x5EED, max_bounces: int = 8,
                 wave_paths_log2: int = 0, collect_stats: bool = False, bvh_leaf_size: int = 0):
        self.lib = _lib.load()
        self.model = model
        cfg = B2ptConfig()
        self.lib.b2pt_default_config(C.byref(cfg))
        cfg.device = device
        cfg.bsdf_mode = BSDF_NAMES[bsdf_mode] if isinstance(bsdf_mode, str) else int(bsdf_mode)
        cfg.rng_mode = RNG_NAMES[rng_mode] if isinstance(rng_mode, str) else int(rng_mode)
        cfg.seed = seed
        cfg.max_bounces = max_bounces
        cfg.wave_paths_log2 = wave_paths_log2
        cfg.collect_stats = int(collect_stats)
        cfg.bvh_leaf_size = bvh_leaf_size      # triangles per leaf of the wide tree, 0 = default (2)
        marshal = _SceneMarshal(model)
        h = C.c_void_p()
        rc = self.lib.b2pt_create(C.byref(marshal.desc), C.byref(cfg), C.byref(h))
        if rc != 0:
            check(None, rc)
        self._h = h
        self.width = self.height = 0

    # -- life cycle ---------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self.lib.b2pt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- the reference surface ---------------------------------------------
    def Resize(self, new_size):
        w, h = int(new_size[0]), int(new_size[1])
        check(self._h, self.lib.b2pt_resize(self._h, w, h))
        if w and h:
            self.width, self.height = w, h

    def SetCamera(self, camera):
        pos, iv, ip = camera.launch_matrices(self.width, self.height)
        self.SetCameraRaw(pos, iv, ip)

    def SetCameraRaw(self, position, inverse_view_cm, inverse_proj_cm):
        p = np.ascontiguousarray(position, dtype=f32)
        iv = np.ascontiguousarray(inverse_view_cm, dtype=f32).reshape(16)
        ip = np.ascontiguousarray(inverse_proj_cm, dtype=f32).reshape(16)
        check(self._h, self.lib.b2pt_set_camera(self._h, _fp(p), _fp(iv), _fp(ip)))

    def SetLights(self, lights):
        arr = (B2ptPointLight * max(len(lights), 1))()
        for i, (p, c) in enumerate(lights):
            arr[i].position = (C.c_float * 3)(*np.asarray(p, dtype=f32))
            arr[i].color = (C.c_float * 3)(*np.asarray(c, dtype=f32))
        check(self._h, self.lib.b2pt_set_lights(self._h, arr, len(lights)))

    def SetMaxBounces(self, max_bounces: int):
        check(self._h, self.lib.b2pt_set_max_bounces(self._h, int(max_bounces)))

    def Render(self, h_pixels: np.ndarray | None = None) -> np.ndarray:
        """One sample per pixel, overwrite (OptixRenderer.cpp:617-647)."""
        if h_pixels is None:
            h_pixels = np.empty((self.height, self.width, 3), dtype=f32)
        # before the first Resize the call is a silent no-op like the reference's (:621)
        assert h_pixels.dtype == f32 and (self.width == 0 or h_pixels.size == self.width * self.height * 3)
        check(self._h, self.lib.b2pt_render(self._h, _fp(h_pixels)))
        return h_pixels

    # -- extensions ---------------------------------------------------------
    def SetBsdfMode(self, mode):
        m = BSDF_NAMES[mode] if isinstance(mode, str) else int(mode)
        check(self._h, self.lib.b2pt_set_bsdf_mode(self._h, m))

    def SetRng(self, mode, seed: int = 0x5EED):
        m = RNG_NAMES[mode] if isinstance(mode, str) else int(mode)
        check(self._h, self.lib.b2pt_set_rng(self._h, m, seed))

    def SetFrameId(self, frame_id: int):
        check(self._h, self.lib.b2pt_set_frame_id(self._h, frame_id))

    def GetFrameId(self) -> int:
        v = C.c_uint32()
        check(self._h, self.lib.b2pt_get_frame_id(self._h, C.byref(v)))
        return v.value

    def RenderAccumulate(self, spp: int, out: np.ndarray | None = None, download: bool = True):
        if not download:
            check(self._h, self.lib.b2pt_render_accumulate_device(self._h, int(spp)))
            return None
        if out is None:
            out = np.empty((self.height, self.width, 3), dtype=f32)
        check(self._h, self.lib.b2pt_render_accumulate(self._h, int(spp), _fp(out)))
        return out

    def ClearAccumulation(self):
        check(self._h, self.lib.b2pt_clear_accumulation(self._h))

    def AccumulationDevice(self):
        p = C.c_void_p()
        n = C.c_uint32()
        check(self._h, self.lib.b2pt_get_accumulation_device(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def DownloadAccumulation(self):
        out = np.empty((self.height, self.width, 3), dtype=f32)
        n = C.c_uint32()
        check(self._h, self.lib.b2pt_download_accumulation(self._h, _fp(out), C.byref(n)))
        return out, n.value

    def UploadAccumulation(self, rgb_sum: np.ndarray, spp: int):
        a = np.ascontiguousarray(rgb_sum, dtype=f32)
        check(self._h, self.lib.b2pt_upload_accumulation(self._h, _fp(a), int(spp)))

    # -- image output (Final.frag display transform, WriteImage.cpp) -----------------
    def Display(self, apply_oetf: bool = True, top_row_first: bool = True) -> np.ndarray:
        """8-bit image of the running mean: (H, W, 3) uint8."""
        out = np.empty((self.height, self.width, 3), dtype=np.uint8)
        check(self._h, self.lib.b2pt_display_rgb8(self._h, int(apply_oetf), int(top_row_first),
                                                  out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def WriteBMP(self, path, apply_oetf: bool = False):
        """WriteImage::WriteBMP (the reference writes the linear pixels, clamped: apply_oetf False)."""
        check(self._h, self.lib.b2pt_write_bmp(self._h, str(path).encode(), int(apply_oetf)))

    def WritePNG(self, path, apply_oetf: bool = True):
        check(self._h, self.lib.b2pt_write_png(self._h, str(path).encode(), int(apply_oetf)))

    def WriteEXR(self, path):
        check(self._h, self.lib.b2pt_write_exr(self._h, str(path).encode()))

    def Stats(self) -> dict:
        s = B2ptStats()
        check(self._h, self.lib.b2pt_get_stats(self._h, C.byref(s)))
        out = {}
        for k, _ in B2ptStats._fields_:
            v = getattr(s, k)
            out[k] = list(v) if hasattr(v, "__len__") else v
        return out

    def SetKernelTiming(self, on: bool = True):
        check(self._h, self.lib.b2pt_set_kernel_timing(self._h, int(on)))

    def ResetStats(self):
        check(self._h, self.lib.b2pt_reset_stats(self._h))

    # -- one process per GPU: NCCL reduce inside the library (csrc/multi.cu) -------------
    @staticmethod
    def DistUniqueId() -> np.ndarray:
        """128 bytes from rank 0 that every rank passes to DistInit (broadcast them with whatever the host has)."""
        idb = np.zeros(128, dtype=np.uint8)
        rc = _lib.load().b2pt_dist_unique_id(idb.ctypes.data_as(_lib.c_uint8_p))
        if rc != 0:
            check(None, rc)
        return idb

    def DistInit(self, unique_id: np.ndarray, rank: int, world: int):
        idb = np.ascontiguousarray(unique_id, dtype=np.uint8)
        assert idb.size == 128
        check(self._h, self.lib.b2pt_dist_init(self._h, idb.ctypes.data_as(_lib.c_uint8_p), int(rank), int(world)))

    def DistRenderAccumulate(self, spp: int, out: np.ndarray | None = None):
        """Collective: every rank traces `spp` samples from its own frame id; rank 0's `out` receives the mean."""
        check(self._h, self.lib.b2pt_dist_render_accumulate(self._h, int(spp), _fp(out) if out is not None else None))
        return out

    def DistShutdown(self):
        check(self._h, self.lib.b2pt_dist_shutdown(self._h))

    # -- parity hooks -------------------------------------------------------
    def TraceClosest(self, org, dirs, tmin: float = 0.0, tmax: float = 100.0):
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = o.shape[0]
        prim = np.empty(n, dtype=np.int32)
        mesh = np.empty(n, dtype=np.int32)
        t = np.empty(n, dtype=f32)
        u = np.empty(n, dtype=f32)
        v = np.empty(n, dtype=f32)
        i32 = _lib.c_int32_p
        check(self._h, self.lib.b2pt_trace_closest(self._h, n, _fp(o), _fp(d), tmin, tmax,
                                                   prim.ctypes.data_as(i32), mesh.ctypes.data_as(i32),
                                                   _fp(t), _fp(u), _fp(v)))
        return prim, mesh, t, u, v

    def TraceAny(self, org, dirs, tmax, tmin: float = 0.0):
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = o.shape[0]
        tm = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=f32), (n,)), dtype=f32)
        occ = np.empty(n, dtype=np.uint8)
        check(self._h, self.lib.b2pt_trace_any(self._h, n, _fp(o), _fp(d), tmin, _fp(tm),
                                               occ.ctypes.data_as(_lib.c_uint8_p)))
        return occ

    NOALLOC_MAX_RAYS = 32768   # B2PT_NOALLOC_MAX_RAYS (include/b2pt.h)

    def TraceAnyNoAlloc(self, org, dirs, tmax, tmin: float = 0.0):
        """b2pt_trace_any_noalloc: TraceAny through the library's static device scratch — no cudaMalloc in the
        call, at most NOALLOC_MAX_RAYS rays, calls on one GPU must not overlap."""
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = o.shape[0]
        tm = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=f32), (n,)), dtype=f32)
        occ = np.empty(n, dtype=np.uint8)
        check(self._h, self.lib.b2pt_trace_any_noalloc(self._h, n, _fp(o), _fp(d), tmin, _fp(tm),
                                                       occ.ctypes.data_as(_lib.c_uint8_p)))
        return occ

    def DownloadBvh(self):
        st = self.Stats()
        # the binary (64-byte) nodes the builder emits; the 4-wide nodes K5/K8 walk are derived from them
        nodes = np.empty((st["n_nodes"], 16), dtype=f32)
        order = np.empty(st["n_triangles"], dtype=np.int32)
        check(self._h, self.lib.b2pt_download_bvh(self._h, nodes.ctypes.data_as(C.c_void_p),
                                                  order.ctypes.data_as(_lib.c_int32_p)))
        return nodes, order


class MultiRenderer:
    """One process, several GPUs (b2pt_create_multi): the samples of a call are split over the devices and the
    sums are combined by one kernel on devices[0] reading its peers over NVLink.  Same method names as
    OptixRenderer for the reference surface; `renderer(i)` wraps device i's handle for everything else."""

    def __init__(self, model: dict, devices, *, bsdf_mode="auto", rng_mode="philox", seed: int = 0x5EED,
                 max_bounces: int = 8, wave_paths_log2: int = 0):
        self.lib = _lib.load()
        cfg = B2ptConfig()
        self.lib.b2pt_default_config(C.byref(cfg))
        cfg.bsdf_mode = BSDF_NAMES[bsdf_mode] if isinstance(bsdf_mode, str) else int(bsdf_mode)
        cfg.rng_mode = RNG_NAMES[rng_mode] if isinstance(rng_mode, str) else int(rng_mode)
        cfg.seed = seed
        cfg.max_bounces = max_bounces
        cfg.wave_paths_log2 = wave_paths_log2
        marshal = _SceneMarshal(model)
        dev = np.ascontiguousarray(devices, dtype=np.int32)
        h = C.c_void_p()
        rc = self.lib.b2pt_create_multi(dev.ctypes.data_as(_lib.c_int32_p), len(dev), C.byref(marshal.desc), C.byref(cfg), C.byref(h))
        if rc != 0:
            msg = self.lib.b2pt_multi_last_error(None)
            raise _lib.B2ptError(rc, msg.decode() if msg else "")
        self._h = h
        self.devices = [int(d) for d in dev]
        self.width = self.height = 0

    def _check(self, rc):
        if rc != 0:
            msg = self.lib.b2pt_multi_last_error(self._h)
            raise _lib.B2ptError(rc, msg.decode() if msg else "")

    def close(self):
        if getattr(self, "_h", None):
            self.lib.b2pt_destroy_multi(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def Resize(self, new_size):
        w, h = int(new_size[0]), int(new_size[1])
        self._check(self.lib.b2pt_multi_resize(self._h, w, h))
        if w and h:
            self.width, self.height = w, h

    def SetCamera(self, camera):
        pos, iv, ip = camera.launch_matrices(self.width, self.height)
        self._check(self.lib.b2pt_multi_set_camera(self._h, _fp(np.ascontiguousarray(pos, f32)), _fp(np.ascontiguousarray(iv, f32)),
                                                   _fp(np.ascontiguousarray(ip, f32))))

    def SetLights(self, lights):
        arr = (B2ptPointLight * max(len(lights), 1))()
        for i, (p, c) in enumerate(lights):
            arr[i].position = (C.c_float * 3)(*np.asarray(p, dtype=f32))
            arr[i].color = (C.c_float * 3)(*np.asarray(c, dtype=f32))
        self._check(self.lib.b2pt_multi_set_lights(self._h, arr, len(lights)))

    def SetMaxBounces(self, n: int):
        self._check(self.lib.b2pt_multi_set_max_bounces(self._h, int(n)))

    def SetFrameId(self, frame_id: int):
        self._check(self.lib.b2pt_multi_set_frame_id(self._h, int(frame_id)))

    def ClearAccumulation(self):
        self._check(self.lib.b2pt_multi_clear_accumulation(self._h))

    def RenderAccumulate(self, spp_total: int, out: np.ndarray | None = None, download: bool = True):
        if download and out is None:
            out = np.empty((self.height, self.width, 3), dtype=f32)
        self._check(self.lib.b2pt_multi_render_accumulate(self._h, int(spp_total), _fp(out) if download else None))
        return out if download else None

    def renderer(self, i: int) -> "OptixRenderer":
        """Device i's renderer as a borrowed OptixRenderer (stats, per-device settings); do not close it."""
        r = OptixRenderer.__new__(OptixRenderer)
        r.lib = self.lib
        r.model = None
        r._h = C.c_void_p(self.lib.b2pt_multi_renderer(self._h, int(i)))
        r.close = lambda: None          # the handle belongs to the MultiRenderer
        r.width, r.height = self.width, self.height
        return r
