"""glTF ingestion: the robust reading and the reference-loader-compatible one
(ModelLoading/ModelLoader.cpp:91-245) on hand-built files."""
import base64
import json
import struct

import numpy as np
import pytest

from optixpathtracer_b200 import gltf
from tests import cases

f32 = np.float32


def _write(tmp_path, j, blob, name="t.gltf"):
    (tmp_path / "t.bin").write_bytes(blob)
    j["buffers"] = [{"byteLength": len(blob), "uri": "t.bin"}]
    p = tmp_path / name
    p.write_text(json.dumps(j))
    return p


def _hand_built(tmp_path, index_type):
    """Parent node (translate) -> child node (rotate 90 deg about y, scale 2) with one triangle pair;
    positions and normals INTERLEAVED (byteStride 24); uv as normalised u16; indices of `index_type`."""
    pos = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], f32)
    nrm = np.tile(np.array([0, 0, 1], f32), (4, 1))
    inter = np.concatenate([pos, nrm], axis=1).astype(f32).tobytes()
    uv = (np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float64) * 65535).astype(np.uint16).tobytes()
    ct, dt = {"u8": (5121, np.uint8), "u16": (5123, np.uint16), "u32": (5125, np.uint32)}[index_type]
    idx = np.array([0, 1, 2, 0, 2, 3], dt).tobytes()
    blob = inter + uv + idx + b"\0" * (-len(idx) % 4)
    s = np.sin(np.pi / 4)
    j = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0]}],
         "nodes": [{"name": "parent", "translation": [1, 2, 3], "children": [1]},
                   {"name": "child", "mesh": 0, "rotation": [0, s, 0, s], "scale": [2, 2, 2]}],
         "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1, "TEXCOORD_0": 2}, "indices": 3,
                                     "material": 0}]}],
         "materials": [{"pbrMetallicRoughness": {"baseColorFactor": [0.2, 0.4, 0.6, 1], "metallicFactor": 0.25,
                                                 "roughnessFactor": 0.75}}],
         "bufferViews": [{"buffer": 0, "byteOffset": 0, "byteLength": 96, "byteStride": 24},
                         {"buffer": 0, "byteOffset": 96, "byteLength": 16},
                         {"buffer": 0, "byteOffset": 112, "byteLength": len(idx)}],
         "accessors": [{"bufferView": 0, "byteOffset": 0, "componentType": 5126, "count": 4, "type": "VEC3"},
                       {"bufferView": 0, "byteOffset": 12, "componentType": 5126, "count": 4, "type": "VEC3"},
                       {"bufferView": 1, "componentType": 5123, "normalized": True, "count": 4, "type": "VEC2"},
                       {"bufferView": 2, "componentType": ct, "count": 6, "type": "SCALAR"}]}
    return _write(tmp_path, j, blob), pos, nrm


@pytest.mark.parametrize("index_type", ["u8", "u16", "u32"])
def test_hierarchy_stride_index_types_and_normalised_uv(tmp_path, index_type):
    p, pos, nrm = _hand_built(tmp_path, index_type)
    sc = gltf.load_gltf(p)
    assert len(sc["meshes"]) == 1
    m = sc["meshes"][0]
    assert np.array_equal(m["positions"], pos) and np.array_equal(m["normals"], nrm)
    assert np.allclose(m["texcoords"], [[0, 0], [1, 0], [1, 1], [0, 1]])
    assert np.array_equal(m["indices"], [[0, 1, 2], [0, 2, 3]]) and m["indices"].dtype == np.int32
    assert np.allclose(m["albedo"], [0.2, 0.4, 0.6]) and m["metallic"] == 0.25 and m["roughness"] == 0.75
    # world = T(1,2,3) * Ry(90) * S(2): x axis -> -z * 2
    want = np.array([[0, 0, 2, 1], [0, 2, 0, 2], [-2, 0, 0, 3], [0, 0, 0, 1]], f32)
    assert np.allclose(m["model"], want, atol=1e-6)


def test_compat_mode_is_the_reference_loader(tmp_path):
    """Root nodes only, TRS only, u16 indices whatever the accessor says, tightly packed floats."""
    p, pos, _ = _hand_built(tmp_path, "u16")
    with pytest.raises(gltf.GltfError):      # root node without a mesh: the reference dereferences it
        gltf.load_gltf(p, compat=True)
    j = json.loads(p.read_text())
    j["scenes"][0]["nodes"] = [1]            # make the mesh node a root: the parent's translation is lost
    p.write_text(json.dumps(j))
    with pytest.raises(gltf.GltfError, match="out of bounds"):   # u16 uvs read as packed floats
        gltf.load_gltf(p, compat=True)
    del j["meshes"][0]["primitives"][0]["attributes"]["TEXCOORD_0"]
    p.write_text(json.dumps(j))
    m = gltf.load_gltf(p, compat=True)["meshes"][0]
    assert np.allclose(m["model"][:3, 3], 0) and np.allclose(np.abs(m["model"][:3, :3]).sum(), 6.0, atol=1e-5)
    # tightly packed read of an interleaved view: vertex 1 is what sits 12 bytes after vertex 0 = normal 0
    assert np.array_equal(m["positions"][1], [0, 0, 1]) and np.array_equal(m["positions"][0], pos[0])
    full = gltf.load_gltf(p)["meshes"][0]
    assert np.array_equal(full["positions"], pos)


def test_missing_normals_indices_and_default_material(tmp_path):
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], f32)
    j = {"asset": {"version": "2.0"}, "scenes": [{"nodes": [0]}], "nodes": [{"mesh": 0}],
         "meshes": [{"primitives": [{"attributes": {"POSITION": 0}}, {"attributes": {"POSITION": 0}, "mode": 1}]}],
         "bufferViews": [{"buffer": 0, "byteLength": 36}],
         "accessors": [{"bufferView": 0, "componentType": 5126, "count": 3, "type": "VEC3"}]}
    sc = gltf.load_gltf(_write(tmp_path, j, pos.tobytes()))
    assert len(sc["meshes"]) == 1                       # the LINES primitive is not a surface
    m = sc["meshes"][0]
    assert m["normals"] is None and m["texcoords"] is None
    assert np.array_equal(m["indices"], [[0, 1, 2]])
    assert np.array_equal(m["albedo"], [1, 1, 1]) and m["tex_albedo"] == -1


def test_errors_are_reported(tmp_path):
    pos = np.zeros((3, 3), f32)
    j = {"asset": {"version": "2.0"}, "scenes": [{"nodes": [0]}], "nodes": [{"mesh": 0}],
         "meshes": [{"primitives": [{"attributes": {"POSITION": 0}, "indices": 1}]}],
         "bufferViews": [{"buffer": 0, "byteLength": 36}, {"buffer": 0, "byteOffset": 36, "byteLength": 6}],
         "accessors": [{"bufferView": 0, "componentType": 5126, "count": 3, "type": "VEC3"},
                       {"bufferView": 1, "componentType": 5123, "count": 3, "type": "SCALAR"}]}
    blob = pos.tobytes() + np.array([0, 1, 7], np.uint16).tobytes() + b"\0\0"
    with pytest.raises(gltf.GltfError, match="out of range"):
        gltf.load_gltf(_write(tmp_path, j, blob))
    j["accessors"][0]["count"] = 300
    with pytest.raises(gltf.GltfError, match="past its buffer"):
        gltf.load_gltf(_write(tmp_path, j, blob))
    j["asset"]["version"] = "1.0"
    with pytest.raises(gltf.GltfError):
        gltf.load_gltf(_write(tmp_path, j, blob))


@pytest.mark.parametrize("embed", [False, True])
def test_scene_round_trip_with_textures_and_lights(tmp_path, embed):
    sc = cases.frame_scenes()["textured"]
    p = gltf.save_gltf(sc, tmp_path / "scene.gltf", embed=embed)
    back = gltf.load_gltf(p)
    assert len(back["meshes"]) == len(sc["meshes"]) and len(back["textures"]) == len(sc["textures"])
    for a, b in zip(sc["meshes"], back["meshes"]):
        for k in ("positions", "normals", "texcoords", "indices"):
            if a[k] is None:
                assert b[k] is None
            else:
                assert np.array_equal(np.asarray(a[k]), b[k]), k
        assert np.array_equal(np.asarray(a["model"], f32).reshape(4, 4), b["model"])
        assert np.array_equal(np.asarray(a["albedo"], f32), b["albedo"])
        assert (f32(a["metallic"]), f32(a["roughness"])) == (f32(b["metallic"]), f32(b["roughness"]))
        assert (a["tex_albedo"], a["tex_normal"], a["tex_metal_rough"]) == (b["tex_albedo"], b["tex_normal"], b["tex_metal_rough"])
    for a, b in zip(sc["textures"], back["textures"]):
        assert np.array_equal(a, b)
    assert len(back["lights"]) == len(sc["lights"])
    for (pa, ca), (pb, cb) in zip(sc["lights"], back["lights"]):
        assert np.allclose(pa, pb) and np.allclose(ca, cb, rtol=1e-6)


def test_glb_container(tmp_path):
    p, pos, _ = _hand_built(tmp_path, "u16")
    j = json.loads(p.read_text())
    blob = (tmp_path / "t.bin").read_bytes()
    del j["buffers"][0]["uri"]
    js = json.dumps(j).encode()
    js += b" " * (-len(js) % 4)
    blob += b"\0" * (-len(blob) % 4)
    glb = struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + 8 + len(blob)) + struct.pack("<I4s", len(js), b"JSON") + js \
        + struct.pack("<I4s", len(blob), b"BIN\0") + blob
    (tmp_path / "t.glb").write_bytes(glb)
    m = gltf.load_gltf(tmp_path / "t.glb")["meshes"][0]
    assert np.array_equal(m["positions"], pos)
