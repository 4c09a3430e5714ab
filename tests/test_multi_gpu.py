"""GPU tier: the multi-GPU entry points (csrc/multi.cu) against a single-GPU render of the union of
the sample indices.

The path shards by samples: device g renders a disjoint range of absolute sample indices (the Philox key holds
the absolute index) and the fp32 sums are added once per call, so an N-device image must equal the 1-device
image of the same samples up to the re-association of that last sum (N partial sums instead of one running sum):
every pixel within a few ulps, and EXACTLY equal to the sum of the N single-device partial images added in
device order (the fused peer-memory reduce is deterministic).

  * one process, N devices (b2pt_create_multi): runs on one GPU too, with devices = [0, 0] (two renderers, one
    device — the reduce kernel then reads a same-device "peer"), and across real GPUs when the box has them;
  * one process per device (b2pt_dist_*, NCCL): world size 1 on any box; world size 2 as two sub-processes when
    the box has two GPUs.
"""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from tests import cases
from tests.util import bits

pytestmark = pytest.mark.gpu
f32 = np.float32
ROOT = Path(__file__).resolve().parent.parent
W, H, DEPTH, SPP = 96, 54, 6, 10


def _gpus():
    import torch
    return torch.cuda.device_count()


def _single(scene, first_frame, spp, mode="auto", device=0):
    from optixpathtracer_b200 import OptixRenderer
    with OptixRenderer(None, scene, device=device, bsdf_mode=mode, rng_mode="philox", max_bounces=DEPTH) as r:
        r.Resize((W, H)); r.SetCamera(scene["camera"]); r.SetLights(scene["lights"])
        r.SetFrameId(first_frame)
        r.RenderAccumulate(spp, download=False)
        s, n = r.DownloadAccumulation()
        assert n == spp
        return s


def _check_against_union(scene, mean, shares, mode="auto"):
    """mean = image of sum(shares) samples split in `shares`; compare with the 1-device render of all of them."""
    total = sum(shares)
    union = _single(scene, 0, total, mode) / f32(total)
    assert np.isfinite(mean).all()
    assert np.allclose(mean, union, rtol=2e-6, atol=1e-7), "N-device image differs from the 1-device image of the same samples"
    parts, at = [], 0
    for s in shares:
        parts.append(_single(scene, at, s, mode))
        at += s
    acc = parts[0].copy()
    for p in parts[1:]:
        acc = acc + p
    assert np.array_equal(bits(mean), bits(acc * (f32(1.0) / f32(total)))), "the reduce is not the device-order sum of the partial images"


@pytest.mark.parametrize("devices", [[0, 0], [0, 0, 0], "all"])
def test_one_process_n_devices_equals_the_union_of_samples(devices):
    from optixpathtracer_b200 import MultiRenderer
    if devices == "all":
        if _gpus() < 2:
            pytest.skip("needs two GPUs")
        devices = list(range(min(_gpus(), 8)))
    sc = cases.frame_scenes()["mixed"]
    n = len(devices)
    with MultiRenderer(sc, devices, bsdf_mode="auto", rng_mode="philox", max_bounces=DEPTH) as m:
        m.Resize((W, H)); m.SetCamera(sc["camera"]); m.SetLights(sc["lights"])
        mean = m.RenderAccumulate(SPP)
        shares = [SPP // n + (1 if g < SPP % n else 0) for g in range(n)]
        _check_against_union(sc, mean, shares)
        # a second call keeps accumulating: every device continues after the whole first call's samples
        mean2 = m.RenderAccumulate(n)
        st = [m.renderer(g).Stats() for g in range(n)]
        assert sum(s["samples"] for s in st) == (SPP + n) * W * H
    again = _single(sc, 0, SPP + n) / f32(SPP + n)
    assert np.allclose(mean2, again, rtol=3e-6, atol=1e-7)


def test_one_process_per_device_world_size_one():
    """The NCCL entry points with a communicator of one rank: same image as b2pt_render_accumulate."""
    from optixpathtracer_b200 import OptixRenderer
    sc = cases.frame_scenes()["mixed"]
    uid = OptixRenderer.DistUniqueId()
    with OptixRenderer(None, sc, bsdf_mode="auto", rng_mode="philox", max_bounces=DEPTH) as r:
        r.Resize((W, H)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
        r.DistInit(uid, 0, 1)
        out = np.empty((H, W, 3), f32)
        r.DistRenderAccumulate(SPP, out)
        r.DistShutdown()
    want = _single(sc, 0, SPP) * (f32(1.0) / f32(SPP))
    assert np.array_equal(bits(out), bits(want))


RANK_PROG = r'''
import sys, numpy as np
sys.path.insert(0, %r)
from optixpathtracer_b200 import OptixRenderer
from tests import cases
rank, world, idfile, out, spp = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4], int(sys.argv[5])
sc = cases.frame_scenes()["mixed"]
uid = np.fromfile(idfile, np.uint8)
with OptixRenderer(None, sc, device=rank, bsdf_mode="auto", rng_mode="philox", max_bounces=%d) as r:
    r.Resize((%d, %d)); r.SetCamera(sc["camera"]); r.SetLights(sc["lights"])
    r.DistInit(uid, rank, world)
    r.SetFrameId(rank * spp)                      # disjoint absolute sample indices
    img = np.empty((%d, %d, 3), np.float32)
    r.DistRenderAccumulate(spp, img)
    if rank == 0:
        np.save(out, img)
    r.DistShutdown()
'''


def test_one_process_per_device_world_size_two(tmp_path):
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    from optixpathtracer_b200 import OptixRenderer
    OptixRenderer.DistUniqueId().tofile(tmp_path / "id.bin")
    spp = SPP // 2
    prog = RANK_PROG % (str(ROOT), DEPTH, W, H, H, W)
    procs = [subprocess.Popen([sys.executable, "-c", prog, str(k), "2", str(tmp_path / "id.bin"), str(tmp_path / "out.npy"), str(spp)],
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for k in range(2)]
    for p in procs:
        o, e = p.communicate(timeout=600)
        assert p.returncode == 0, e[-2000:]
    mean = np.load(tmp_path / "out.npy")
    sc = cases.frame_scenes()["mixed"]
    union = _single(sc, 0, 2 * spp) / f32(2 * spp)
    assert np.allclose(mean, union, rtol=2e-6, atol=1e-7)
