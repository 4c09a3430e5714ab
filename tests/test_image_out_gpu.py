"""K10 + file writers against a numpy restatement of the reference's display pass
(Renderer/OpenGL/Shader/Final.frag:12-36) and image writers (Renderer/Images/WriteImage.cpp:9-99).
The files are parsed back with nothing but struct/zlib."""
import struct
import zlib

import numpy as np
import pytest

from tests import cases

pytestmark = pytest.mark.gpu
W, H = 37, 21   # odd sizes: BMP row padding, non-multiple-of-block pixel counts


def linear_to_srgb8(c):   # Final.frag:12-17
    c = c.astype(np.float64)
    with np.errstate(invalid="ignore"):
        hi = (1.055 * np.power(c, np.float64(np.float32(1.0 / 2.4))).astype(np.float32) - np.float32(0.055))
    lo = np.float32(12.92) * c.astype(np.float32)
    return np.where(c < np.float32(0.0031308), lo, hi).astype(np.float32)


def to_u8(v):             # WriteImage.cpp:16-18: clamp, * 255, truncate; NaN clamps to 0 (fmaxf)
    v = np.where(np.isnan(v), np.float32(0), v)
    return (np.clip(v, 0.0, 1.0).astype(np.float32) * np.float32(255.0)).astype(np.uint8)


@pytest.fixture(scope="module")
def rend():
    from optixpathtracer_b200 import OptixRenderer
    sc = cases.frame_scenes()["plain"]
    r = OptixRenderer(None, sc)
    r.Resize((W, H))
    rng = np.random.default_rng(5)
    mean = rng.uniform(-0.1, 1.3, (H, W, 3)).astype(np.float32)
    mean[0, :8] = np.linspace(0.0, 0.01, 8, dtype=np.float32)[:, None]   # the linear toe of the OETF
    mean[3, 4, 1] = np.nan
    mean[5, 6] = np.inf
    spp = 7
    r.UploadAccumulation(mean * np.float32(spp), spp)
    got, n = r.DownloadAccumulation()
    assert n == spp
    yield r, (got / np.float32(spp)).astype(np.float32)   # what the kernels see: sum * (1/spp)
    r.close()


def _mean_as_kernel(r):
    s, n = r.DownloadAccumulation()
    return (s * np.float32(1.0 / n)).astype(np.float32)


def test_display_transform(rend):
    r, _ = rend
    mean = _mean_as_kernel(r)
    lin = r.Display(apply_oetf=False, top_row_first=False)
    assert np.array_equal(lin, to_u8(mean))                      # pure clamp/truncate: exact
    assert np.array_equal(r.Display(apply_oetf=False, top_row_first=True), lin[::-1])
    enc = r.Display(apply_oetf=True, top_row_first=False).astype(np.int32)
    ref = to_u8(linear_to_srgb8(mean)).astype(np.int32)
    d = np.abs(enc - ref)
    assert d.max() <= 1 and (d > 0).mean() < 0.005                # one count where pow lands on a boundary
    assert enc[3, 4, 1] == 0 and (enc[5, 6] == 255).all()         # NaN -> 0, +inf -> 255


def test_bmp_file(rend, tmp_path):
    r, _ = rend
    p = tmp_path / "out.bmp"
    r.WriteBMP(p)                                                 # WriteBMP writes the linear pixels
    raw = p.read_bytes()
    assert raw[:2] == b"BM"
    size, _, _, off = struct.unpack_from("<IHHI", raw, 2)
    hdr, w, h, planes, bpp, comp = struct.unpack_from("<IiiHHI", raw, 14)
    assert (size, off, hdr, w, h, planes, bpp, comp) == (len(raw), 54, 40, W, H, 1, 24, 0)
    stride = (W * 3 + 3) // 4 * 4
    rows = np.frombuffer(raw, np.uint8, offset=off).reshape(H, stride)[:, :W * 3].reshape(H, W, 3)
    assert np.array_equal(rows[:, :, ::-1], r.Display(apply_oetf=False, top_row_first=False))   # bottom-up, BGR


def test_png_file(rend, tmp_path):
    r, _ = rend
    p = tmp_path / "out.png"
    r.WritePNG(p)
    raw = p.read_bytes()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(raw):
        n, typ = struct.unpack_from(">I4s", raw, pos)
        data = raw[pos + 8:pos + 8 + n]
        assert struct.unpack_from(">I", raw, pos + 8 + n)[0] == zlib.crc32(typ + data)
        chunks.append((typ, data))
        pos += 12 + n
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    assert struct.unpack(">IIBBBBB", chunks[0][1]) == (W, H, 8, 2, 0, 0, 0)
    scan = np.frombuffer(zlib.decompress(chunks[1][1]), np.uint8).reshape(H, 1 + 3 * W)   # checks adler32 too
    assert (scan[:, 0] == 0).all()
    assert np.array_equal(scan[:, 1:].reshape(H, W, 3), r.Display(apply_oetf=True, top_row_first=True))


def test_exr_file(rend, tmp_path):
    r, _ = rend
    p = tmp_path / "out.exr"
    r.WriteEXR(p)
    raw = p.read_bytes()
    assert struct.unpack_from("<II", raw, 0) == (20000630, 2)
    pos, attrs = 8, {}
    while raw[pos] != 0:
        e = raw.index(b"\0", pos); name = raw[pos:e].decode(); pos = e + 1
        e = raw.index(b"\0", pos); typ = raw[pos:e].decode(); pos = e + 1
        n = struct.unpack_from("<I", raw, pos)[0]
        attrs[name] = (typ, raw[pos + 4:pos + 4 + n]); pos += 4 + n
    pos += 1
    assert attrs["compression"] == ("compression", b"\0") and attrs["lineOrder"][1] == b"\0"
    assert struct.unpack("<4i", attrs["dataWindow"][1]) == (0, 0, W - 1, H - 1)
    ch = attrs["channels"][1]
    names = [ch[i:i + 1] for i in (0, 18, 36)]
    assert names == [b"B", b"G", b"R"] and ch[-1] == 0            # WriteImage.cpp:76-78
    assert all(struct.unpack_from("<I", ch, i + 2)[0] == 2 for i in (0, 18, 36))   # FLOAT
    offs = struct.unpack_from(f"<{H}Q", raw, pos)
    img = np.empty((H, 3, W), np.float32)
    for y, o in enumerate(offs):
        yy, nbytes = struct.unpack_from("<iI", raw, o)
        assert (yy, nbytes) == (y, 12 * W)
        img[y] = np.frombuffer(raw, np.float32, count=3 * W, offset=o + 8).reshape(3, W)
    mean = _mean_as_kernel(r)
    want = mean[::-1].copy()                                     # top row first
    bad = np.isnan(want).any(axis=2)
    want[bad] = 0.0                                              # WriteImage.cpp:53-56
    assert np.array_equal(img[:, 0].view(np.uint32), want[:, :, 2].view(np.uint32))   # B plane
    assert np.array_equal(img[:, 1].view(np.uint32), want[:, :, 1].view(np.uint32))
    assert np.array_equal(img[:, 2].view(np.uint32), want[:, :, 0].view(np.uint32))


def test_image_output_argument_checks(rend, tmp_path):
    r, _ = rend
    from optixpathtracer_b200 import _lib
    lib = _lib.load()
    assert lib.b2pt_write_png(r._h, None, 1) < 0
    assert lib.b2pt_write_exr(r._h, str(tmp_path / "no" / "such" / "dir.exr").encode()) < 0
    assert lib.b2pt_display_rgb8(r._h, 1, 1, None) < 0
