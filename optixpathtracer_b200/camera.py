"""Host-side camera math: the matrices the reference hands to its device code.

Mirrors Renderer/Camera.cpp:37-70 (forward from Euler angles, glm::lookAt,
glm::perspective with the "horizontal" FOV used as fovy — SURVEY Q18) and
GlmHelperMethods.cpp:4-19 (Blender -> engine axes).  Scalar host code, fp32
like glm, outside the hot path: OptixRenderer::SetCamera (OptixRenderer.cpp:
662-668) inverts the two matrices on the host and uploads them.
"""
from __future__ import annotations

import numpy as np

f32 = np.float32


def blender_to_engine_position(p) -> np.ndarray:
    x, y, z = p
    return np.array([x, z, -y], dtype=f32)


def blender_to_engine_rotation(r) -> np.ndarray:
    x, y, z = r
    return np.array([90 - x, 180 + z, y], dtype=f32)


def _normalize(v):
    v = np.asarray(v, dtype=f32)
    return (v / f32(np.sqrt(np.dot(v, v)))).astype(f32)


class Camera:
    """Renderer/Camera.h:4-33."""

    def __init__(self):
        self.world_up = np.array([0, 1, 0], dtype=f32)
        self.position = np.zeros(3, dtype=f32)
        self.rotation = np.zeros(3, dtype=f32)   # degrees
        self.fov_radians = f32(np.radians(f32(40.0)))

    def set_blender_position(self, p):
        self.position = blender_to_engine_position(p)

    def set_blender_rotation(self, r):
        self.rotation = blender_to_engine_rotation(r)

    def forward(self) -> np.ndarray:
        rx, ry, _ = np.radians(self.rotation.astype(f32)).astype(f32)
        x = f32(np.sin(ry)) * f32(np.cos(rx))
        y = -f32(np.sin(rx))
        z = f32(np.cos(rx)) * f32(np.cos(ry))
        return _normalize([x, y, z])

    def view_matrix(self) -> np.ndarray:
        """glm::lookAt (right-handed); returned as a 4x4 with M[row, col]."""
        eye = self.position.astype(f32)
        f = _normalize(self.forward())
        s = _normalize(np.cross(f, self.world_up).astype(f32))
        u = np.cross(s, f).astype(f32)
        m = np.eye(4, dtype=f32)
        m[0, :3] = s
        m[1, :3] = u
        m[2, :3] = -f
        m[0, 3] = -np.dot(s, eye)
        m[1, 3] = -np.dot(u, eye)
        m[2, 3] = np.dot(f, eye)
        return m

    def projection_matrix(self, aspect: float, near: float = 0.1, far: float = 100.0) -> np.ndarray:
        """glm::perspective (RH, depth -1..1) with fovy = the 'horizontal' FOV."""
        t = f32(np.tan(self.fov_radians / f32(2)))
        m = np.zeros((4, 4), dtype=f32)
        m[0, 0] = f32(1) / (f32(aspect) * t)
        m[1, 1] = f32(1) / t
        m[2, 2] = -(f32(far) + f32(near)) / (f32(far) - f32(near))
        m[3, 2] = -f32(1)
        m[2, 3] = -(f32(2) * f32(far) * f32(near)) / (f32(far) - f32(near))
        return m

    def launch_matrices(self, width: int, height: int):
        """(position, inverse view, inverse projection) as column-major float32[16]
        — what LaunchParams.camera holds (LaunchParams.h:17-21)."""
        aspect = f32(width) / f32(height)
        inv_view = np.linalg.inv(self.view_matrix().astype(np.float64)).astype(f32)
        inv_proj = np.linalg.inv(self.projection_matrix(aspect).astype(np.float64)).astype(f32)
        return (self.position.astype(f32).copy(),
                np.ascontiguousarray(inv_view.T).reshape(16).copy(),
                np.ascontiguousarray(inv_proj.T).reshape(16).copy())


def camera_rays(position, inv_view_cm, inv_proj_cm, width: int, height: int):
    """fp32 numpy restatement of `Camera()` (devicePrograms.cu:601-623), used by
    tests to produce primary rays for the traversal parity hooks."""
    iv = np.asarray(inv_view_cm, dtype=f32).reshape(4, 4).T   # back to M[row, col]
    ip = np.asarray(inv_proj_cm, dtype=f32).reshape(4, 4).T
    xs = ((np.arange(width, dtype=f32) + f32(0.5)) / f32(width)) * f32(2) - f32(1)
    ys = ((np.arange(height, dtype=f32) + f32(0.5)) / f32(height)) * f32(2) - f32(1)
    X, Y = np.meshgrid(xs, ys)
    ndc = np.stack([X, Y, np.ones_like(X), np.ones_like(X)], axis=-1).astype(f32)
    pv = ndc @ ip.T
    pv[..., 3] = 0
    pw = (pv @ iv.T)[..., :4].astype(f32)
    n = np.sqrt((pw * pw).sum(-1, keepdims=True)).astype(f32)
    d = (pw / n)[..., :3].astype(f32)
    o = np.broadcast_to(np.asarray(position, dtype=f32), d.shape).copy()
    return o.reshape(-1, 3), d.reshape(-1, 3)
