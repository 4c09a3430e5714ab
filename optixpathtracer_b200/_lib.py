"""ctypes binding of include/b2pt.h (the C-ABI of libb2pt.so).

The library is the product; there is no Python or CPU fallback.  If the shared
object is missing this module raises at load time, and every compute call on a
box without a CUDA device returns B2PT_E_CUDA, which `check` turns into an
exception.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libb2pt.so"

B2PT_OK = 0
BSDF_AUTO, BSDF_CONDUCTOR, BSDF_DIFFUSE, BSDF_DIELECTRIC, BSDF_LAYERED = range(5)
RNG_PHILOX, RNG_LCG_DEVICE, RNG_LCG_HOST = range(3)
BSDF_NAMES = {"auto": BSDF_AUTO, "conductor": BSDF_CONDUCTOR, "diffuse": BSDF_DIFFUSE,
              "dielectric": BSDF_DIELECTRIC, "layered": BSDF_LAYERED}
RNG_NAMES = {"philox": RNG_PHILOX, "lcg_device": RNG_LCG_DEVICE, "lcg_host": RNG_LCG_HOST}

c_float_p = C.POINTER(C.c_float)
c_int32_p = C.POINTER(C.c_int32)
c_uint32_p = C.POINTER(C.c_uint32)
c_uint8_p = C.POINTER(C.c_uint8)


class B2ptMesh(C.Structure):
    _fields_ = [
        ("positions", c_float_p), ("normals", c_float_p), ("texcoords", c_float_p),
        ("indices", c_int32_p), ("n_vertices", C.c_int32), ("n_triangles", C.c_int32),
        ("model", C.c_float * 16), ("albedo", C.c_float * 3),
        ("metallic", C.c_float), ("roughness", C.c_float),
        ("tex_albedo", C.c_int32), ("tex_normal", C.c_int32), ("tex_metal_rough", C.c_int32),
    ]


class B2ptTexture(C.Structure):
    _fields_ = [("rgba", c_uint8_p), ("width", C.c_int32), ("height", C.c_int32)]


class B2ptSceneDesc(C.Structure):
    _fields_ = [("meshes", C.POINTER(B2ptMesh)), ("n_meshes", C.c_int32),
                ("textures", C.POINTER(B2ptTexture)), ("n_textures", C.c_int32)]


class B2ptPointLight(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("color", C.c_float * 3)]


class B2ptConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("bsdf_mode", C.c_int32), ("rng_mode", C.c_int32),
                ("seed", C.c_uint64), ("max_bounces", C.c_int32), ("wave_paths_log2", C.c_int32),
                ("bvh_leaf_size", C.c_int32), ("collect_stats", C.c_int32)]


class B2ptStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("radiance_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("nodes_visited", C.c_uint64), ("triangles_tested", C.c_uint64),
                ("exact_tests", C.c_uint64), ("nan_pixels", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("device_ms_render", C.c_double),
                ("device_ms_build", C.c_double), ("n_triangles", C.c_int32), ("n_nodes", C.c_int32),
                ("node_bytes", C.c_int32), ("triangle_bytes", C.c_int32),
                ("scene_device_bytes", C.c_int64),
                ("nodes_visited_shadow", C.c_uint64), ("triangles_tested_shadow", C.c_uint64),
                ("exact_tests_shadow", C.c_uint64),
                ("device_ms_kernel", C.c_double * 5), ("launches_kernel", C.c_uint64 * 5),
                ("shadow_rays_traced", C.c_uint64), ("warp_slots_node", C.c_uint64),
                ("warp_slots_tri", C.c_uint64),
                ("device_ms_engine", C.c_double), ("launches_engine", C.c_uint64),
                ("walk_f_hits", C.c_uint64), ("walk_s_hits", C.c_uint64),
                ("n_wide_nodes", C.c_int32), ("bvh_depth", C.c_int32), ("bvh_wide_depth", C.c_int32),
                ("reserved0", C.c_int32)]


# name -> (restype, argtypes); exactly the symbols include/b2pt.h declares.
_H = C.c_void_p
SIGNATURES = {
    "b2pt_abi_version": (C.c_int, []),
    "b2pt_default_config": (None, [C.POINTER(B2ptConfig)]),
    "b2pt_create": (C.c_int, [C.POINTER(B2ptSceneDesc), C.POINTER(B2ptConfig), C.POINTER(_H)]),
    "b2pt_destroy": (None, [_H]),
    "b2pt_last_error": (C.c_char_p, [_H]),
    "b2pt_resize": (C.c_int, [_H, C.c_int32, C.c_int32]),
    "b2pt_set_camera": (C.c_int, [_H, c_float_p, c_float_p, c_float_p]),
    "b2pt_set_lights": (C.c_int, [_H, C.POINTER(B2ptPointLight), C.c_int32]),
    "b2pt_set_max_bounces": (C.c_int, [_H, C.c_int32]),
    "b2pt_set_bsdf_mode": (C.c_int, [_H, C.c_int32]),
    "b2pt_set_rng": (C.c_int, [_H, C.c_int32, C.c_uint64]),
    "b2pt_set_frame_id": (C.c_int, [_H, C.c_uint32]),
    "b2pt_get_frame_id": (C.c_int, [_H, c_uint32_p]),
    "b2pt_render": (C.c_int, [_H, c_float_p]),
    "b2pt_render_accumulate": (C.c_int, [_H, C.c_int32, c_float_p]),
    "b2pt_render_accumulate_device": (C.c_int, [_H, C.c_int32]),
    "b2pt_clear_accumulation": (C.c_int, [_H]),
    "b2pt_get_accumulation_device": (C.c_int, [_H, C.POINTER(C.c_void_p), c_uint32_p]),
    "b2pt_download_accumulation": (C.c_int, [_H, c_float_p, c_uint32_p]),
    "b2pt_display_rgb8": (C.c_int, [_H, C.c_int32, C.c_int32, C.POINTER(C.c_uint8)]),
    "b2pt_write_bmp": (C.c_int, [_H, C.c_char_p, C.c_int32]),
    "b2pt_write_png": (C.c_int, [_H, C.c_char_p, C.c_int32]),
    "b2pt_write_exr": (C.c_int, [_H, C.c_char_p]),
    "b2pt_upload_accumulation": (C.c_int, [_H, c_float_p, C.c_uint32]),
    "b2pt_set_kernel_timing": (C.c_int, [_H, C.c_int32]),
    "b2pt_get_stats": (C.c_int, [_H, C.POINTER(B2ptStats)]),
    "b2pt_reset_stats": (C.c_int, [_H]),
    "b2pt_trace_closest": (C.c_int, [_H, C.c_int64, c_float_p, c_float_p, C.c_float, C.c_float,
                                      c_int32_p, c_int32_p, c_float_p, c_float_p, c_float_p]),
    "b2pt_trace_any": (C.c_int, [_H, C.c_int64, c_float_p, c_float_p, C.c_float, c_float_p, c_uint8_p]),
    "b2pt_trace_any_noalloc": (C.c_int, [_H, C.c_int64, c_float_p, c_float_p, C.c_float, c_float_p, c_uint8_p]),
    "b2pt_bsdf_f": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int64, c_float_p, c_float_p,
                               c_float_p, c_float_p, c_uint32_p, c_float_p, c_uint32_p]),
    "b2pt_bsdf_sample": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int64, c_float_p, c_float_p,
                                    c_float_p, c_uint32_p, c_uint8_p, c_float_p, c_float_p, c_float_p,
                                    c_uint8_p, c_uint32_p]),
    "b2pt_bsdf_pdf": (C.c_int, [C.c_int32, C.c_int32, C.c_int64, c_float_p, c_float_p, c_float_p,
                                 C.c_int32, C.c_int32, c_float_p]),
    "b2pt_download_bvh": (C.c_int, [_H, C.c_void_p, c_int32_p]),
    # multi-GPU, one process (csrc/multi.cu)
    "b2pt_create_multi": (C.c_int, [c_int32_p, C.c_int32, C.POINTER(B2ptSceneDesc), C.POINTER(B2ptConfig), C.POINTER(_H)]),
    "b2pt_destroy_multi": (None, [_H]),
    "b2pt_multi_last_error": (C.c_char_p, [_H]),
    "b2pt_multi_device_count": (C.c_int32, [_H]),
    "b2pt_multi_renderer": (_H, [_H, C.c_int32]),
    "b2pt_multi_resize": (C.c_int, [_H, C.c_int32, C.c_int32]),
    "b2pt_multi_set_camera": (C.c_int, [_H, c_float_p, c_float_p, c_float_p]),
    "b2pt_multi_set_lights": (C.c_int, [_H, C.POINTER(B2ptPointLight), C.c_int32]),
    "b2pt_multi_set_max_bounces": (C.c_int, [_H, C.c_int32]),
    "b2pt_multi_set_bsdf_mode": (C.c_int, [_H, C.c_int32]),
    "b2pt_multi_set_rng": (C.c_int, [_H, C.c_int32, C.c_uint64]),
    "b2pt_multi_set_frame_id": (C.c_int, [_H, C.c_uint32]),
    "b2pt_multi_clear_accumulation": (C.c_int, [_H]),
    "b2pt_multi_render_accumulate": (C.c_int, [_H, C.c_int32, c_float_p]),
    # multi-GPU, one process per device (NCCL)
    "b2pt_dist_unique_id": (C.c_int, [c_uint8_p]),
    "b2pt_dist_init": (C.c_int, [_H, c_uint8_p, C.c_int32, C.c_int32]),
    "b2pt_dist_render_accumulate": (C.c_int, [_H, C.c_int32, c_float_p]),
    "b2pt_dist_shutdown": (C.c_int, [_H]),
}

_lib = None


class B2ptError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b2pt error {code}: {msg}")
        self.code = code


def load() -> C.CDLL:
    """Load libb2pt.so and bind every declared symbol; raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the path-tracing core)")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(handle, rc: int) -> None:
    if rc != B2PT_OK:
        msg = load().b2pt_last_error(handle)
        raise B2ptError(rc, msg.decode() if msg else "")
