"""CPU tier: the C-ABI library loads and exports every symbol include/b2pt.h
declares (no compute calls — there is no GPU here), argument checking that does
not need a device, and the C++ host shim compiles against the header."""
import ctypes as C
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    text = (ROOT / "include" / "b2pt.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2pt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from optixpathtracer_b200 import _lib
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 28
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert lib.b2pt_abi_version() == 4


def test_struct_sizes_match_the_header():
    from optixpathtracer_b200 import _lib
    src = r'''
    #include <stdio.h>
    #include "b2pt.h"
    int main(void) { printf("%zu %zu %zu %zu %zu %zu\n", sizeof(B2ptMesh), sizeof(B2ptTexture), sizeof(B2ptSceneDesc),
                            sizeof(B2ptPointLight), sizeof(B2ptConfig), sizeof(B2ptStats)); return 0; }'''
    exe = ROOT / "tests" / "hostsim" / "abi_sizes"
    subprocess.run(["gcc", "-x", "c", "-", "-I", str(ROOT / "include"), "-o", str(exe)], input=src.encode(), check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, check=True).stdout.split()]
    want = [C.sizeof(t) for t in (_lib.B2ptMesh, _lib.B2ptTexture, _lib.B2ptSceneDesc, _lib.B2ptPointLight,
                                  _lib.B2ptConfig, _lib.B2ptStats)]
    assert got == want


def test_argument_checks_without_a_device():
    from optixpathtracer_b200 import _lib
    lib = _lib.load()
    cfg = _lib.B2ptConfig()
    lib.b2pt_default_config(C.byref(cfg))
    assert (cfg.device, cfg.bsdf_mode, cfg.rng_mode, cfg.max_bounces) == (0, 0, 0, 8)
    h = C.c_void_p()
    assert lib.b2pt_create(None, None, C.byref(h)) == -1            # null scene
    assert lib.b2pt_resize(None, 4, 4) == -1 and lib.b2pt_render(None, None) == -1
    lib.b2pt_destroy(None)                                          # no-op
    # b2pt_create validates the config with the setters' ranges before it touches a device
    scene = _lib.B2ptSceneDesc()
    for field, bad in (("max_bounces", 100), ("max_bounces", -1), ("bsdf_mode", 5), ("bsdf_mode", -1), ("rng_mode", 3),
                       ("wave_paths_log2", 5), ("wave_paths_log2", 28), ("bvh_leaf_size", 9)):
        lib.b2pt_default_config(C.byref(cfg))
        setattr(cfg, field, bad)
        assert lib.b2pt_create(C.byref(scene), C.byref(cfg), C.byref(h)) == -1, (field, bad)
        assert not h.value and field.split("_")[0] in lib.b2pt_last_error(None).decode()


def test_cpp_shim_compiles():
    """include/OptixRenderer.hpp against stand-ins shaped like the reference's Model / Mesh / Texture / Camera."""
    src = r'''
    #include <memory>
    #include <vector>
    #include "OptixRenderer.hpp"
    struct vec2 { float x, y; }; struct vec3 { float x, y, z; }; struct ivec2 { int x, y; }; struct ivec3 { int x, y, z; };
    struct mat4 { float m[4][4]; float* operator[](int i) { return m[i]; } const float* operator[](int i) const { return m[i]; } };
    inline mat4 inverse(const mat4& a) { return a; }
    struct Mesh { std::vector<vec3> vertecies, normal; std::vector<vec2> texCoord; std::vector<ivec3> index;
                  vec3 albedo; float metallic, roughness; int albedoTex, normalTex, metalRoughTex;
                  mat4 GetModelMatrix() { return mat4(); } };
    struct Texture { uint32_t* pixel; ivec2 resolution; };
    struct Model { std::vector<std::unique_ptr<Mesh>> meshes; std::vector<std::unique_ptr<Texture>> textures; };
    struct Camera { vec3 position; mat4 GetViewMatrix() { return mat4(); } mat4 GetProjectionMatrix(float) { return mat4(); } };
    struct PointLight { vec3 position, color; };
    using OptixRenderer = b2pt_shim::OptixRendererT<Model>;
    void use(Model* m, Camera* c, std::vector<PointLight>* l, vec3* px) {
        OptixRenderer r("unused.ptx", m);
        ivec2 s{64, 36};
        r.Resize(s); r.SetCamera(c); r.SetLights(l); r.SetMaxBounces(2); r.Render(px);
    }'''
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-x", "c++", "-", "-I", str(ROOT / "include")],
                       input=src.encode(), capture_output=True)
    assert r.returncode == 0, r.stderr.decode()
