"""Headless OptixView: the state machine of Renderer/OptixView.cpp:130-185, 212-255
over the device-side accumulation, and checkpoint / resume."""
import copy

import numpy as np
import pytest

from tests import cases
from tests.util import bits

pytestmark = pytest.mark.gpu
W, H = 40, 24


def _view(**kw):
    from optixpathtracer_b200 import OptixView
    sc = cases.frame_scenes()["mixed"]
    cam = copy.deepcopy(sc["camera"])
    return OptixView((W, H), sc, cam, rng_mode="lcg_host", max_bounces=4, **kw), sc


def _plain(sc, cam, first_frame, n):
    from optixpathtracer_b200 import OptixRenderer
    with OptixRenderer(None, sc, rng_mode="lcg_host", max_bounces=4) as r:
        r.Resize((W, H)); r.SetCamera(cam); r.SetLights(sc["lights"])
        r.SetFrameId(first_frame - 1)
        r.RenderAccumulate(n, download=False)
        return r.DownloadAccumulation()[0]


def test_running_mean_and_camera_reset():
    v, sc = _view()
    try:
        assert v.Run(3) == 3 and v.samples == 3
        assert np.array_equal(bits(v.Image()), bits(_plain(sc, v.camera, 1, 3) / np.float32(3)))
        v.OnKeyPressed("w")                      # camera moved: the next frame starts a new image (:141-151)
        assert v.clearFrameBuffer
        assert v.Step() and v.samples == 1 and not v.clearFrameBuffer
        # frame ids keep counting (OptixRenderer::Render pre-increments frame.id, never reset): frame 4
        assert v.optixRenderer.GetFrameId() == 4
        assert np.array_equal(bits(v.Image()), bits(_plain(sc, v.camera, 4, 1)))
        v.OnCursorMoved(1000.0, -5000.0)
        assert v.camera.rotation[0] == -80.0     # pitch clamp (:24)
    finally:
        v.close()


def test_fixed_sample_count_writes_the_exr_once(tmp_path):
    out = tmp_path / "final.exr"
    v, sc = _view(max_samples=4, finished_exr=out)
    try:
        assert v.Run(3) == 3
        assert np.array_equal(bits(v.Image()), bits(_plain(sc, v.camera, 1, 3) / np.float32(4)))   # fb + new/maxSamples
        assert v.Run(5) == 1 and v.samples == 4 and out.exists()          # 4th frame, then "Render is finished"
        stamp = out.stat().st_mtime_ns
        assert v.Run(3) == 0 and out.stat().st_mtime_ns == stamp
    finally:
        v.close()


def test_checkpoint_resume(tmp_path):
    a, sc = _view()
    b, _ = _view()
    try:
        a.Run(2)
        a.Save(tmp_path / "ck.npz")
        a.Run(3)
        b.Load(tmp_path / "ck.npz")
        assert b.samples == 2 and b.optixRenderer.GetFrameId() == 2
        b.Run(3)
        assert np.array_equal(bits(a.Image()), bits(b.Image()))
    finally:
        a.close(); b.close()
