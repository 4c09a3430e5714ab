"""glTF 2.0 scene ingestion for the B200 core.

Host-side mirror of `ModelLoading/ModelLoader.cpp` (tinygltf front-end) producing
the scene dict `OptixRenderer` takes (scenes.py): one mesh per glTF primitive,
material constants and texture ids, RGBA8 textures.  Flattening, the FLOAT12
pre-transform and the texture upload happen on the device (bvh_build.cu K1,
texture objects) — this module only parses.

`compat=True` reproduces the reference loader, limitations included
(ModelLoader.cpp:91-164): only the ROOT nodes of scene 0, each assumed to carry
a mesh; node TRS only (T * R * S, Mesh.cpp:6-23; `matrix` ignored); indices read
as u16 whatever the accessor says; positions/normals/uvs read as tightly packed
floats.  The default is the robust reading the survey asks for: the whole node
hierarchy with `matrix` or TRS, u8/u16/u32 or absent indices, byteStride,
normalised-integer texcoords, missing normals (None: the core shades with the
geometric normal), .glb containers, data: URIs, KHR_lights_punctual point
lights, and the first camera.

`save_gltf` writes a scene dict back out (one node per mesh, .gltf + .bin + PNG
textures): the README scene and C5 can be handed to any glTF consumer.
"""
import base64
import json
import struct
from pathlib import Path

import numpy as np

f32 = np.float32

_DTYPES = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_NCOMP = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


class GltfError(ValueError):
    pass


def _quat_matrix(q):
    x, y, z, w = [float(v) for v in q]
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w), 0],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w), 0],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y), 0],
                     [0, 0, 0, 1]], dtype=np.float64)


def _node_matrix(node, compat):
    if "matrix" in node and not compat:
        return np.array(node["matrix"], dtype=np.float64).reshape(4, 4).T   # glTF stores column-major
    t = np.eye(4); r = np.eye(4); s = np.eye(4)
    if node.get("translation"):
        t[:3, 3] = node["translation"]
    if node.get("rotation"):
        r = _quat_matrix(node["rotation"])          # glTF order x, y, z, w (ModelLoader.cpp:237-242)
    if node.get("scale"):
        s[0, 0], s[1, 1], s[2, 2] = node["scale"]
    return t @ r @ s                                 # Mesh.cpp:22


class _Doc:
    def __init__(self, path):
        self.path = Path(path)
        raw = self.path.read_bytes()
        self.glb_bin = None
        if raw[:4] == b"glTF":
            magic, version, length = struct.unpack_from("<4sII", raw, 0)
            if version != 2:
                raise GltfError(f"GLB version {version}")
            pos, chunks = 12, []
            while pos < length:
                n, typ = struct.unpack_from("<I4s", raw, pos)
                chunks.append((typ, raw[pos + 8:pos + 8 + n]))
                pos += 8 + n + (-n % 4)
            if not chunks or chunks[0][0] != b"JSON":
                raise GltfError("GLB without a JSON chunk")
            self.j = json.loads(chunks[0][1].decode("utf-8"))
            self.glb_bin = next((c[1] for c in chunks[1:] if c[0] == b"BIN\0"), None)
        else:
            self.j = json.loads(raw.decode("utf-8"))
        if str(self.j.get("asset", {}).get("version", "2.0")).split(".")[0] != "2":
            raise GltfError("not glTF 2.x")
        self._buffers = {}

    def uri_bytes(self, uri):
        if uri.startswith("data:"):
            return base64.b64decode(uri.split(",", 1)[1])
        return (self.path.parent / uri).read_bytes()

    def buffer(self, i):
        if i not in self._buffers:
            b = self.j["buffers"][i]
            self._buffers[i] = self.uri_bytes(b["uri"]) if "uri" in b else self.glb_bin
            if self._buffers[i] is None:
                raise GltfError(f"buffer {i} has no data")
        return self._buffers[i]

    def accessor(self, i, compat_as=None):
        """(count, ncomp) array of the accessor's component type.  compat_as = (dtype, ncomp):
        read tightly packed values of that type at the accessor's offset, as the reference does."""
        a = self.j["accessors"][i]
        if "bufferView" not in a:
            raise GltfError("accessor without a bufferView (sparse/zero accessors are not supported)")
        bv = self.j["bufferViews"][a["bufferView"]]
        buf = self.buffer(bv["buffer"])
        off = bv.get("byteOffset", 0) + a.get("byteOffset", 0)
        if compat_as is not None:
            dt, nc = np.dtype(compat_as[0]), compat_as[1]
            n = a["count"] * (1 if nc == 0 else nc)
            if off + n * dt.itemsize > len(buf):
                raise GltfError(f"compat: accessor {i} read as packed {dt.name} runs past its buffer "
                                "(the reference loader would read out of bounds here)")
            return np.frombuffer(buf, dt, count=n, offset=off).reshape(a["count"], max(nc, 1))
        dt = np.dtype(_DTYPES[a["componentType"]])
        nc = _NCOMP[a["type"]]
        stride = bv.get("byteStride", 0) or dt.itemsize * nc
        need = (a["count"] - 1) * stride + dt.itemsize * nc if a["count"] else 0
        if off + need > len(buf):
            raise GltfError(f"accessor {i} runs past its buffer")
        out = np.ndarray((a["count"], nc), dtype=dt, buffer=buf, offset=off, strides=(stride, dt.itemsize))
        out = np.array(out)
        if a.get("normalized") and dt.kind in "iu":
            m = float(np.iinfo(dt).max)
            out = np.maximum(out.astype(f32) / f32(m), f32(-1.0))
        return out


def _image_rgba(doc, img):
    from PIL import Image
    import io
    if "uri" in img:
        data = doc.uri_bytes(img["uri"])
    else:
        bv = doc.j["bufferViews"][img["bufferView"]]
        b = doc.buffer(bv["buffer"])
        data = b[bv.get("byteOffset", 0):bv.get("byteOffset", 0) + bv["byteLength"]]
    im = Image.open(io.BytesIO(data)).convert("RGBA")          # stbi_load(..., STBI_rgb_alpha), no flip (:60-77)
    return np.ascontiguousarray(np.asarray(im, dtype=np.uint8))


def load_gltf(path, compat: bool = False) -> dict:
    """Returns {"meshes": [...], "textures": [RGBA8 (H,W,4)], "lights": [(pos, colour)], "camera": Camera|None}."""
    doc = _Doc(path)
    j = doc.j
    meshes, lights, camera = [], [], None
    light_defs = j.get("extensions", {}).get("KHR_lights_punctual", {}).get("lights", [])

    def material(prim):
        m = {"albedo": np.array([1, 1, 1], dtype=f32), "metallic": 0.0, "roughness": 0.0,
             "tex_albedo": -1, "tex_normal": -1, "tex_metal_rough": -1}
        mats = j.get("materials", [])
        if not mats or prim.get("material", -1) < 0:               # ParseMaterial :167-171
            if not compat:
                m["metallic"], m["roughness"] = 1.0, 1.0            # glTF's default material
            return m
        mat = mats[prim["material"]]
        pbr = mat.get("pbrMetallicRoughness", {})
        m["tex_albedo"] = pbr.get("baseColorTexture", {}).get("index", -1)
        m["tex_metal_rough"] = pbr.get("metallicRoughnessTexture", {}).get("index", -1)
        m["tex_normal"] = mat.get("normalTexture", {}).get("index", -1)
        m["metallic"] = float(pbr.get("metallicFactor", 1.0))
        m["roughness"] = float(pbr.get("roughnessFactor", 1.0))
        m["albedo"] = np.array(pbr.get("baseColorFactor", [1, 1, 1, 1])[:3], dtype=f32)
        return m

    def add_primitives(node, world):
        mesh = j["meshes"][node["mesh"]]
        for k, prim in enumerate(mesh["primitives"]):
            if prim.get("mode", 4) != 4:
                if compat:
                    pass                                            # the reference never looks at `mode`
                else:
                    continue                                        # points / lines / strips: not surfaces
            att = prim["attributes"]
            if "POSITION" not in att:
                raise GltfError("primitive without POSITION")
            if compat:
                pos = doc.accessor(att["POSITION"], (np.float32, 3))
                nrm = doc.accessor(att["NORMAL"], (np.float32, 3))  # required by the reference (:118-127)
                uv = doc.accessor(att["TEXCOORD_0"], (np.float32, 2)) if "TEXCOORD_0" in att else None
                idx = doc.accessor(prim["indices"], (np.uint16, 0)).reshape(-1)
            else:
                pos = doc.accessor(att["POSITION"]).astype(f32)
                nrm = doc.accessor(att["NORMAL"]).astype(f32) if "NORMAL" in att else None
                uv = doc.accessor(att["TEXCOORD_0"]).astype(f32) if "TEXCOORD_0" in att else None
                idx = (doc.accessor(prim["indices"]).reshape(-1) if "indices" in prim
                       else np.arange(len(pos)))
            ntri = len(idx) // 3
            tri = np.ascontiguousarray(idx[:3 * ntri].astype(np.int64).reshape(ntri, 3))
            if not compat and ntri and (tri.max() >= len(pos) or tri.min() < 0):
                raise GltfError(f"mesh {node.get('name', '?')}: index {int(tri.max())} out of range ({len(pos)} vertices)")
            m = material(prim)
            name = node.get("name", mesh.get("name", "mesh")) + (f".{k}" if len(mesh["primitives"]) > 1 else "")
            meshes.append({"positions": np.ascontiguousarray(pos, dtype=f32),
                           "normals": None if nrm is None else np.ascontiguousarray(nrm, dtype=f32),
                           "texcoords": None if uv is None else np.ascontiguousarray(uv, dtype=f32),
                           "indices": tri.astype(np.int32), "model": world.astype(f32), "name": name, **m})

    def visit(node_id, parent, depth=0):
        if depth > 256:
            raise GltfError("node hierarchy too deep (cycle?)")
        nonlocal camera
        node = j["nodes"][node_id]
        world = parent @ _node_matrix(node, compat)
        if "mesh" in node:
            add_primitives(node, world)
        ext = node.get("extensions", {}).get("KHR_lights_punctual")
        if ext is not None and ext["light"] < len(light_defs):
            L = light_defs[ext["light"]]
            if L.get("type") == "point":
                col = np.array(L.get("color", [1, 1, 1]), dtype=f32) * f32(L.get("intensity", 1.0))
                lights.append((world[:3, 3].astype(f32), col))
        if "camera" in node and camera is None:
            from .camera import Camera
            cdef = j["cameras"][node["camera"]]
            if cdef.get("type") == "perspective":
                camera = Camera()
                camera.position = world[:3, 3].astype(f32)
                fwd = -(world[:3, 2] / np.linalg.norm(world[:3, 2]))      # glTF cameras look down -Z
                camera.rotation = np.array([np.degrees(-np.arcsin(np.clip(fwd[1], -1, 1))),
                                            np.degrees(np.arctan2(fwd[0], fwd[2])), 0.0], dtype=f32)
                camera.fov_radians = f32(cdef["perspective"]["yfov"])
        if not compat:
            for c in node.get("children", []):
                visit(c, world, depth + 1)

    scenes = j.get("scenes") or [{"nodes": list(range(len(j.get("nodes", []))))}]
    scene = scenes[0 if compat else j.get("scene", 0)]             # "Only the first scene ... will be loaded" (:13)
    for nid in scene.get("nodes", []):
        if compat and "mesh" not in j["nodes"][nid]:
            raise GltfError("compat: the reference loader dereferences node.mesh of every root node (:93)")
        visit(nid, np.eye(4))

    textures = []
    for t in j.get("textures", []):                                 # one entry per glTF *texture* (:48-85)
        textures.append(_image_rgba(doc, j["images"][t["source"]]))
    return {"meshes": meshes, "textures": textures, "lights": lights, "camera": camera}


def _decompose_trs(model):
    """translation, rotation (x, y, z, w), scale of an affine matrix without shear (T * R * S)."""
    m = np.asarray(model, dtype=np.float64).reshape(4, 4)
    t = m[:3, 3]
    s = np.linalg.norm(m[:3, :3], axis=0)
    if np.linalg.det(m[:3, :3]) < 0:
        s[0] = -s[0]
    r = m[:3, :3] / np.where(s == 0, 1.0, s)
    if not np.allclose(r.T @ r, np.eye(3), atol=1e-5):
        raise GltfError("compat: the model matrix has shear; the reference loader reads translation / rotation / scale only")
    w = np.sqrt(max(0.0, 1.0 + r[0, 0] + r[1, 1] + r[2, 2])) / 2.0
    if w > 1e-6:
        q = [(r[2, 1] - r[1, 2]) / (4 * w), (r[0, 2] - r[2, 0]) / (4 * w), (r[1, 0] - r[0, 1]) / (4 * w), w]
    else:   # rotation by pi: the axis is the eigenvector of eigenvalue 1
        ev, evec = np.linalg.eigh((r + r.T) / 2)
        ax = evec[:, np.argmax(ev)]
        q = [ax[0], ax[1], ax[2], 0.0]
    return t.tolist(), [float(v) for v in q], s.tolist()


def save_gltf(scene: dict, path, embed: bool = False, compat: bool = False):
    """Write a scene dict as glTF 2.0: `path`.gltf + .bin (+ PNG textures), or everything inline.
    compat=True writes what the reference's loader can read (ModelLoading/ModelLoader.cpp:91-245): node transforms
    as translation / rotation / scale (it ignores `matrix`), u16 indices, and no mesh-less root nodes (it
    dereferences node.mesh of every root: the lights stay in the file but out of the scene's roots)."""
    from PIL import Image
    import io
    path = Path(path)
    blob = bytearray()
    views, accessors, meshes, nodes, materials = [], [], [], [], []

    def add(arr, target=None, minmax=False):
        arr = np.ascontiguousarray(arr)
        while len(blob) % 4:
            blob.append(0)
        v = {"buffer": 0, "byteOffset": len(blob), "byteLength": arr.nbytes}
        if target:
            v["target"] = target
        views.append(v)
        blob.extend(arr.tobytes())
        ct = {np.dtype(np.float32): 5126, np.dtype(np.uint32): 5125, np.dtype(np.uint16): 5123}[arr.dtype]
        a = {"bufferView": len(views) - 1, "componentType": ct, "count": int(arr.shape[0]),
             "type": {1: "SCALAR", 2: "VEC2", 3: "VEC3"}[1 if arr.ndim == 1 else arr.shape[1]]}
        if minmax:
            a["min"], a["max"] = arr.min(axis=0).tolist(), arr.max(axis=0).tolist()
        accessors.append(a)
        return len(accessors) - 1

    for m in scene["meshes"]:
        att = {"POSITION": add(m["positions"].astype(f32), 34962, True)}
        if m.get("normals") is not None:
            att["NORMAL"] = add(m["normals"].astype(f32), 34962)
        if m.get("texcoords") is not None:
            att["TEXCOORD_0"] = add(m["texcoords"].astype(f32), 34962)
        idx = np.asarray(m["indices"]).reshape(-1)
        if compat and len(m["positions"]) > 65536:
            raise GltfError("compat: the reference loader reads 16-bit indices only (ModelLoader.cpp:145)")
        idx = idx.astype(np.uint16) if len(m["positions"]) <= 65536 else idx.astype(np.uint32)
        prim = {"attributes": att, "indices": add(idx, 34963), "material": len(materials)}
        pbr = {"baseColorFactor": [float(c) for c in m["albedo"]] + [1.0], "metallicFactor": float(m["metallic"]),
               "roughnessFactor": float(m["roughness"])}
        mat = {"pbrMetallicRoughness": pbr}
        if m.get("tex_albedo", -1) >= 0:
            pbr["baseColorTexture"] = {"index": int(m["tex_albedo"])}
            mat["alphaMode"] = "MASK"
        if m.get("tex_metal_rough", -1) >= 0:
            pbr["metallicRoughnessTexture"] = {"index": int(m["tex_metal_rough"])}
        if m.get("tex_normal", -1) >= 0:
            mat["normalTexture"] = {"index": int(m["tex_normal"])}
        materials.append(mat)
        meshes.append({"primitives": [prim], "name": m.get("name", "mesh")})
        model = np.asarray(m.get("model", np.eye(4)), dtype=np.float64).reshape(4, 4)
        if compat:
            t, q, sc = _decompose_trs(model)
            nodes.append({"mesh": len(meshes) - 1, "name": m.get("name", "mesh"), "translation": t, "rotation": q, "scale": sc})
        else:
            nodes.append({"mesh": len(meshes) - 1, "name": m.get("name", "mesh"), "matrix": model.T.reshape(16).tolist()})

    j = {"asset": {"version": "2.0", "generator": "optixpathtracer_b200.gltf"}, "scene": 0,
         "scenes": [{"nodes": list(range(len(nodes)))}], "nodes": nodes, "meshes": meshes, "materials": materials,
         "accessors": accessors, "bufferViews": views}
    images = []
    for i, t in enumerate(scene.get("textures", [])):
        bio = io.BytesIO()
        Image.fromarray(np.asarray(t, dtype=np.uint8), "RGBA").save(bio, format="PNG")
        if embed:
            images.append({"uri": "data:image/png;base64," + base64.b64encode(bio.getvalue()).decode()})
        else:
            name = f"{path.stem}_tex{i}.png"
            (path.parent / name).write_bytes(bio.getvalue())
            images.append({"uri": name})
    if images:
        j["images"] = images
        j["samplers"] = [{"magFilter": 9729, "minFilter": 9729, "wrapS": 10497, "wrapT": 10497}]
        j["textures"] = [{"source": i, "sampler": 0} for i in range(len(images))]
    if scene.get("lights"):
        j["extensionsUsed"] = ["KHR_lights_punctual"]
        defs = []
        for pos, col in scene["lights"]:
            col = np.asarray(col, dtype=np.float64)
            inten = float(col.max()) if col.max() > 0 else 1.0
            defs.append({"type": "point", "color": (col / inten).tolist(), "intensity": inten})
            j["nodes"].append({"name": f"light{len(defs) - 1}", "translation": [float(v) for v in pos],
                               "extensions": {"KHR_lights_punctual": {"light": len(defs) - 1}}})
            if not compat:
                j["scenes"][0]["nodes"].append(len(j["nodes"]) - 1)
        j["extensions"] = {"KHR_lights_punctual": {"lights": defs}}
    if embed:
        j["buffers"] = [{"byteLength": len(blob),
                         "uri": "data:application/octet-stream;base64," + base64.b64encode(bytes(blob)).decode()}]
    else:
        (path.parent / (path.stem + ".bin")).write_bytes(bytes(blob))
        j["buffers"] = [{"byteLength": len(blob), "uri": path.stem + ".bin"}]
    path.write_text(json.dumps(j))
    return path
