"""A scene that went through glTF (save_gltf -> load_gltf) renders the same bits as the scene it came from."""
import numpy as np
import pytest

from optixpathtracer_b200 import gltf
from tests import cases
from tests.util import bits

pytestmark = pytest.mark.gpu


def _frames(scene, lights, camera, n=2):
    from optixpathtracer_b200 import OptixRenderer
    with OptixRenderer(None, scene, rng_mode="lcg_host", max_bounces=cases.FRAME_BOUNCES) as r:
        r.Resize((cases.FRAME_W, cases.FRAME_H)); r.SetCamera(camera); r.SetLights(lights)
        return [r.Render().copy() for _ in range(n)], r.Stats()


@pytest.mark.parametrize("name", ["mixed", "textured"])
def test_gltf_round_trip_renders_identically(tmp_path, name):
    sc = cases.frame_scenes()[name]
    back = gltf.load_gltf(gltf.save_gltf(sc, tmp_path / f"{name}.gltf"))
    a, sa = _frames(sc, sc["lights"], sc["camera"])
    b, sb = _frames(back, back["lights"], sc["camera"])
    for x, y in zip(a, b):
        assert np.array_equal(bits(x), bits(y))
    assert sa["radiance_rays"] == sb["radiance_rays"] and sa["n_triangles"] == sb["n_triangles"]
