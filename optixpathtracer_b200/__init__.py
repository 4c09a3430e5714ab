"""optixpathtracer_b200 — host-side mirror of the reference's renderer surface
over libb2pt.so, the B200-native path-tracing core.

Importing the package does not load the CUDA library; constructing an
`OptixRenderer` does, and fails loudly if it is not built.
"""
from .camera import Camera, camera_rays  # noqa: F401
from . import scenes  # noqa: F401


def __getattr__(name):
    if name == "OptixRenderer":
        from .api import OptixRenderer
        return OptixRenderer
    if name == "MultiRenderer":
        from .api import MultiRenderer
        return MultiRenderer
    if name == "OptixView":
        from .view import OptixView
        return OptixView
    raise AttributeError(name)
