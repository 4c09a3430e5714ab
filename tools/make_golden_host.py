"""Generate the fixtures that pin the rows either side of the hot path to the reference's
own HOST code: tests/golden/compat_scene/ (a glTF the reference's loader can read) and tests/golden/host_code.npz.

Runs oracle/_ref/libref_host.so — ModelLoading/ModelLoader.cpp + Mesh.cpp, Renderer/Images/WriteImage.cpp,
Renderer/Camera.cpp compiled in place from /root/reference by oracle/Makefile (oracle/ref_host.cpp) — and, for the
progressive loop, oracle/_ref/liboracle_ref.so (frames of the reference's device code).

    python tools/make_golden_host.py
"""
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from optixpathtracer_b200 import gltf  # noqa: E402
from tests import cases  # noqa: E402
from tests.util import RefHost, RefOracle  # noqa: E402

f32 = np.float32


def host_image():
    """A fixed W x H x 3 buffer (row 0 = bottom) with values below 0, above 1 and — for the EXR — one NaN pixel."""
    rng = np.random.default_rng(2024)
    img = rng.uniform(-0.25, 1.5, (12, 20, 3)).astype(f32)
    img[0, 0] = (0.0, 1.0, 0.5)
    img[3, 4] = (1.0 / 255.0, 254.5 / 255.0, 0.999999)
    return img


def main():
    rh = RefHost()
    out = ROOT / "tests" / "golden"
    data = {}

    # ---- N1: what ModelLoader::LoadModel returns for a textured scene -----------------------------------
    sc = cases.frame_scenes()["textured"]
    d = out / "compat_scene"
    d.mkdir(exist_ok=True)
    p = gltf.save_gltf(sc, d / "scene.gltf", compat=True)
    model = rh.load_model(p)
    data["n_meshes"] = np.int32(len(model["meshes"]))
    for i, m in enumerate(model["meshes"]):
        for k in ("positions", "normals", "texcoords", "indices", "model", "albedo"):
            data[f"mesh{i}_{k}"] = np.asarray(m[k])
        data[f"mesh{i}_mat"] = np.array([m["metallic"], m["roughness"]], f32)
        data[f"mesh{i}_tex"] = np.array([m["tex_albedo"], m["tex_normal"], m["tex_metal_rough"]], np.int32)
    data["n_textures"] = np.int32(len(model["textures"]))
    for i, t in enumerate(model["textures"]):
        data[f"tex{i}"] = t

    # ---- N3: the bytes WriteImage::WriteEXR / WriteBMP write ---------------------------------------------
    img = host_image()
    with tempfile.TemporaryDirectory() as td:
        nan_img = img.copy()
        nan_img[5, 7, 1] = np.nan          # WriteEXR zeroes a pixel with a NaN component (WriteImage.cpp:52-55)
        rh.write_exr(nan_img.copy(), Path(td) / "a.exr")
        rh.write_bmp(img.copy(), Path(td) / "a.bmp")
        data["image"] = img
        data["image_nan"] = nan_img
        data["exr_bytes"] = np.frombuffer((Path(td) / "a.exr").read_bytes(), np.uint8)
        data["bmp_bytes"] = np.frombuffer((Path(td) / "a.bmp").read_bytes(), np.uint8)

    # ---- camera: main.cpp:83-84's set-up and two oblique ones --------------------------------------------
    cams = []
    for bp, br, asp in (((3.85382, 0.0, 1.0), (90.0, 0.0, 90.0), 1920 / 1080), ((2.0, -3.0, 1.5), (70.0, 0.0, 35.0), 1.0),
                        ((-1.0, 4.0, 0.3), (95.0, 0.0, 200.0), 4 / 3)):
        pos, view, proj = rh.camera(bp, br, 40.0, asp)
        cams.append(np.concatenate([np.array(bp, f32), np.array(br, f32), [f32(asp)], pos, view, proj]))
    data["cameras"] = np.array(cams, f32)

    # ---- N4: frames of the reference's device code for the progressive loop ------------------------------
    ro = RefOracle()
    plain = cases.frame_scenes()["plain"]
    _, fr, _ = ro.render(plain, cases.FRAME_W, cases.FRAME_H, bsdf_mode=0, max_bounces=cases.FRAME_BOUNCES, n_frames=6,
                         want_frames=True)
    data["view_frames"] = fr
    np.savez_compressed(out / "host_code.npz", **data)
    print("wrote", out / "host_code.npz", sorted(x.name for x in d.iterdir()))


if __name__ == "__main__":
    main()
