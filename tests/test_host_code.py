"""The rows either side of the hot path (SURVEY 8(f) N1 glTF ingestion, N3 image output, N4 progressive loop, and the
C++ drop-in shim of 8(b)) against the reference's own HOST code.

Oracle = oracle/_ref/libref_host.so: ModelLoading/ModelLoader.cpp + Mesh.cpp, Renderer/Images/WriteImage.cpp and
Renderer/Camera.cpp compiled in place from /root/reference (oracle/ref_host.cpp).  Its outputs on fixed inputs are
committed as tests/golden/host_code.npz + tests/golden/compat_scene/ (tools/make_golden_host.py), so the checks run
where the reference is not mounted; where the library exists it is re-run against the fixtures (pins the goldens).

  N1  load_gltf(compat=True) returns exactly what ModelLoader::LoadModel returns for the same file (every array
      bit-equal, model matrices included), and that scene renders the reference's frames
  N3  b2pt_write_exr / b2pt_write_bmp write, byte for byte, the files WriteImage::WriteEXR / WriteBMP write
  N4  the headless OptixView folds frames the way OptixView::AddNewFrameToBuffer + AddPathtracedFrame.frag do,
      checked on frames rendered by the REFERENCE's device code (not by the renderer under test)
  8b  include/OptixRenderer.hpp compiled against the reference's real Model / Camera / PointLight headers and
      linked to libb2pt.so (oracle/_ref/host_demo, the INTEGRATION.md recipe) renders the same frames as the
      Python host and hands them to the reference's own EXR writer
"""
import copy
import subprocess
from pathlib import Path

import numpy as np
import pytest

from optixpathtracer_b200 import gltf
from optixpathtracer_b200.camera import Camera
from tests import cases
from tests.util import RefHost, bits

f32 = np.float32
ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"
FIXTURE = GOLD / "compat_scene" / "scene.gltf"


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLD / "host_code.npz"))


def _assert_model_equals_gold(model, gold):
    assert len(model["meshes"]) == int(gold["n_meshes"]) and len(model["textures"]) == int(gold["n_textures"])
    for i, m in enumerate(model["meshes"]):
        for k in ("positions", "normals", "texcoords", "indices"):
            assert np.array_equal(np.asarray(m[k]), gold[f"mesh{i}_{k}"]), (i, k)
        assert np.array_equal(bits(np.asarray(m["model"], f32)), bits(gold[f"mesh{i}_model"])), (i, "model matrix")
        assert np.array_equal(np.asarray(m["albedo"], f32), gold[f"mesh{i}_albedo"])
        assert np.array_equal(np.array([m["metallic"], m["roughness"]], f32), gold[f"mesh{i}_mat"])
        assert [m["tex_albedo"], m["tex_normal"], m["tex_metal_rough"]] == gold[f"mesh{i}_tex"].tolist()
    for i, t in enumerate(model["textures"]):
        assert np.array_equal(t, gold[f"tex{i}"])


# ---- CPU tier -----------------------------------------------------------------------------------------
def test_compat_loader_returns_what_the_reference_loader_returns(gold):
    _assert_model_equals_gold(gltf.load_gltf(FIXTURE, compat=True), gold)


def test_compat_writer_round_trips_the_scene(tmp_path):
    """save_gltf(compat=True) -> load_gltf(compat=True) gives the scene back (TRS decomposition of the model matrices)."""
    sc = cases.frame_scenes()["textured"]
    back = gltf.load_gltf(gltf.save_gltf(sc, tmp_path / "s.gltf", compat=True), compat=True)
    assert len(back["meshes"]) == len(sc["meshes"])
    for a, b in zip(sc["meshes"], back["meshes"]):
        assert np.array_equal(a["positions"], b["positions"]) and np.array_equal(a["indices"], b["indices"])
        assert np.allclose(a["model"], b["model"], atol=1e-6)


def test_reference_host_code_reproduces_the_goldens(gold, tmp_path):
    if not RefHost.available():
        pytest.skip("oracle/_ref/libref_host.so not built")
    rh = RefHost()
    _assert_model_equals_gold(rh.load_model(FIXTURE), gold)
    rh.write_exr(gold["image_nan"].copy(), tmp_path / "a.exr")
    rh.write_bmp(gold["image"].copy(), tmp_path / "a.bmp")
    assert (tmp_path / "a.exr").read_bytes() == gold["exr_bytes"].tobytes()
    assert (tmp_path / "a.bmp").read_bytes() == gold["bmp_bytes"].tobytes()
    for row in gold["cameras"]:
        pos, view, proj = rh.camera(row[0:3], row[3:6], 40.0, float(row[6]))
        assert np.array_equal(np.concatenate([pos, view, proj]), row[7:])


def test_camera_matrices_match_camera_cpp(gold):
    """camera.py against Renderer/Camera.cpp + GlmHelperMethods.cpp: glm::lookAt / glm::perspective to the last ulps
    (numpy's float32 sin / cos against the C library's is the only difference)."""
    for row in gold["cameras"]:
        c = Camera()
        c.set_blender_position(row[0:3])
        c.set_blender_rotation(row[3:6])
        view = c.view_matrix().T.reshape(16)
        proj = c.projection_matrix(f32(row[6])).T.reshape(16)
        assert np.allclose(c.position, row[7:10], atol=0, rtol=0)
        assert np.allclose(view, row[10:26], atol=1e-6, rtol=1e-6)
        assert np.allclose(proj, row[26:42], atol=0, rtol=1e-6)


# ---- GPU tier -----------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_writers_write_the_reference_bytes(gold, tmp_path):
    from optixpathtracer_b200 import OptixRenderer
    img = gold["image"]
    h, w, _ = img.shape
    with OptixRenderer(None, cases.frame_scenes()["plain"]) as r:
        r.Resize((w, h))
        r.UploadAccumulation(gold["image_nan"], 1)
        r.WriteEXR(tmp_path / "a.exr")
        r.UploadAccumulation(img, 1)
        r.WriteBMP(tmp_path / "a.bmp")        # WriteBMP writes the linear pixels, clamped (WriteImage.cpp:9-33)
    assert (tmp_path / "a.exr").read_bytes() == gold["exr_bytes"].tobytes(), "EXR differs from WriteImage::WriteEXR's file"
    assert (tmp_path / "a.bmp").read_bytes() == gold["bmp_bytes"].tobytes(), "BMP differs from WriteImage::WriteBMP's file"


@pytest.mark.gpu
def test_scene_loaded_like_the_reference_renders_the_reference_frames():
    """fixture glTF -> load_gltf(compat=True) -> Render(): the frames the reference's device code produced for the
    scene the fixture was written from (tests/golden/frames.npz `textured`, AUTO mode)."""
    from optixpathtracer_b200 import OptixRenderer
    sc = gltf.load_gltf(FIXTURE, compat=True)
    src = cases.frame_scenes()["textured"]
    want = np.load(GOLD / "frames.npz")["textured_m0"]
    with OptixRenderer(None, sc, bsdf_mode="auto", rng_mode="lcg_host", max_bounces=cases.FRAME_BOUNCES) as r:
        r.Resize((cases.FRAME_W, cases.FRAME_H)); r.SetCamera(src["camera"]); r.SetLights(src["lights"])
        got = np.stack([r.Render() for _ in range(cases.FRAME_N)])
    same = (bits(got) == bits(want)).all(axis=-1)
    assert same.mean() >= 0.995, f"{(~same).sum()} pixels differ"
    assert (np.abs(got - want) <= 1e-6 + 1e-5 * np.abs(want)).all()


def _glsl_fold(frames, max_samples):
    """OptixView::AddNewFrameToBuffer (OptixView.cpp:212-255) + AddPathtracedFrame.frag:13-25 in fp32: running mean
    `mix(fb, new, 1 / samples)` when maxSamples < 0, else `fb + new * (1 / maxSamples)`."""
    fb = np.zeros_like(frames[0])
    for n, fr in enumerate(frames, 1):
        if max_samples < 0:
            w = f32(1.0) / f32(n)
            fb = fb * (f32(1.0) - w) + fr * w
        else:
            fb = fb + fr * (f32(1.0) / f32(max_samples))
    return fb


@pytest.mark.gpu
@pytest.mark.parametrize("max_samples", [-1024, 8])
def test_progressive_loop_folds_reference_frames_like_the_glsl_pass(gold, max_samples):
    from optixpathtracer_b200 import OptixView
    frames = gold["view_frames"]                       # rendered by the reference's device code (oracle/_ref)
    sc = cases.frame_scenes()["plain"]
    v = OptixView((cases.FRAME_W, cases.FRAME_H), sc, copy.deepcopy(sc["camera"]), max_samples=max_samples,
                  bsdf_mode="auto", rng_mode="lcg_host", max_bounces=cases.FRAME_BOUNCES)
    try:
        for k in range(1, len(frames) + 1):
            assert v.Step()
            want = _glsl_fold(frames[:k], max_samples)
            got = v.Image()
            # the device keeps a sum and divides once, the GLSL pass re-weights every frame: equal to rounding
            assert np.allclose(got, want, rtol=2e-6, atol=1e-7), f"after {k} frames"
            acc = frames[0].copy()
            for fr in frames[1:k]:
                acc = acc + fr
            denom = f32(k) if max_samples < 0 else f32(max_samples)
            assert np.array_equal(bits(got), bits(acc / denom)), "K9 adds the reference's frames in order"
        # a camera move clears the image (OptixView.cpp:141-151): to WHITE — what glClearColor(1,1,1,1) leaves, which
        # the fixed-count mode then adds to (fb + new / maxSamples); the running mean overwrites it (weight 1)
        v.OnKeyPressed("w")
        assert v.Step() and v.samples == 1
        one = v.Image()
        from optixpathtracer_b200 import OptixRenderer
        with OptixRenderer(None, sc, bsdf_mode="auto", rng_mode="lcg_host", max_bounces=cases.FRAME_BOUNCES) as r:
            r.Resize((cases.FRAME_W, cases.FRAME_H)); r.SetCamera(v.camera); r.SetLights(sc["lights"])
            r.SetFrameId(len(frames))
            fr = r.Render()
        want = fr if max_samples < 0 else f32(1.0) + fr / f32(max_samples)
        assert np.allclose(one, want, rtol=2e-6, atol=1e-7)
    finally:
        v.close()


@pytest.mark.gpu
def test_cpp_host_against_the_reference_headers_renders_the_same_frames(tmp_path):
    """oracle/_ref/host_demo = tests/integration/host_demo.cpp built against the reference's real headers with
    include/OptixRenderer.hpp swapped in, linked to libb2pt.so (oracle/Makefile)."""
    exe = ROOT / "oracle" / "_ref" / "host_demo"
    if not exe.exists():
        pytest.skip("oracle/_ref/host_demo not built")
    from optixpathtracer_b200 import OptixRenderer, scenes
    w, h, bounces, n = 96, 54, 6, 2
    raw, exr = tmp_path / "f.raw", tmp_path / "f.exr"
    p = subprocess.run([str(exe), str(FIXTURE.parent) + "/", FIXTURE.name, str(w), str(h), str(bounces), str(n), str(raw), str(exr)],
                       capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-1000:] + p.stderr[-2000:]
    got = np.fromfile(raw, f32).reshape(n, h, w, 3)
    sc = gltf.load_gltf(FIXTURE, compat=True)
    ref_scene = scenes.spheres_scene(segments=16, rings=8)     # Scene1's camera and lights (main.cpp:13-18, 83-84)
    with OptixRenderer(None, sc, bsdf_mode="auto", rng_mode="lcg_host", max_bounces=bounces) as r:
        r.Resize((w, h)); r.SetCamera(ref_scene["camera"]); r.SetLights(ref_scene["lights"])
        want = np.stack([r.Render() for _ in range(n)])
        r.UploadAccumulation(want[-1], 1)
        r.WriteEXR(tmp_path / "mine.exr")
    # the C++ host computes its camera with glm and the C library, the Python host with numpy: the matrices agree
    # to an ulp (test_camera_matrices_match_camera_cpp), so do the frames up to the few pixels an ulp of ray
    # direction moves across an edge
    same = (np.abs(got - want) <= 1e-5 + 1e-4 * np.abs(want)).all(axis=-1)
    assert same.mean() >= 0.99, f"{(~same).sum()} of {same.size} pixels differ"
    if np.array_equal(bits(got[-1]), bits(want[-1])):
        assert exr.read_bytes() == (tmp_path / "mine.exr").read_bytes()
