"""Headless mirror of the reference's progressive render loop.

`Renderer/OptixView.{h,cpp}` drives the renderer from a GLFW window: every loop
iteration it sets the camera, renders one frame, clears the accumulated image if
the camera moved, folds the frame into the running image and — once
`maxSamples` frames are in — writes an EXR.  The window, the GL textures and the
two GLSL passes are gone (no graphics stack: "unavailable on this box"); what is
left is the loop's state machine over the B200 core's device-side accumulation
(K9) and the K10 writers, plus checkpoint / resume of (sum, samples, frame id),
which the reference has no counterpart for.
"""
from pathlib import Path

import numpy as np

from .api import OptixRenderer

f32 = np.float32


class OptixView:
    def __init__(self, view_size, model, camera, max_samples: int = -1, finished_exr=None, **renderer_kw):
        """OptixView::OptixView (OptixView.cpp:89-127).  max_samples < 0: running mean forever
        (main.cpp:85 uses -1024); > 0: stop after that many frames and write `finished_exr`."""
        self.maxSamples = int(max_samples)
        self.samples = 0
        self.clearFrameBuffer = False
        self.clearedToWhite = False   # a camera move clears the framebuffer with glClearColor(1, 1, 1, 1) (OptixView.cpp:144-146)
        self.camera = camera
        self.viewSize = tuple(int(v) for v in view_size)
        self.finished_exr = finished_exr
        self._finished_written = False
        self.optixRenderer = OptixRenderer(None, model, **renderer_kw)
        self.optixRenderer.Resize(self.viewSize)
        self.optixRenderer.SetLights(model["lights"])

    # -- what the GLFW callbacks do to the camera (OptixView.cpp:14-31, 48-86) ---------
    def OnCursorMoved(self, dx: float, dy: float):
        self.camera.rotation = (self.camera.rotation + np.array([dy, -dx, 0.0], dtype=f32) * f32(0.1)).astype(f32)
        self.camera.rotation[0] = np.clip(self.camera.rotation[0], -80.0, 80.0)
        self.clearFrameBuffer = True

    def OnKeyPressed(self, key: str):
        fwd = self.camera.forward()
        right = np.cross(fwd, self.camera.world_up).astype(f32)
        right /= np.linalg.norm(right)
        move = {"w": fwd, "s": -fwd, "d": right, "a": -right, "space": self.camera.world_up,
                "shift": -self.camera.world_up}.get(key.lower())
        if move is None or not np.any(move):
            return
        move = (move / np.linalg.norm(move)).astype(f32)
        self.clearFrameBuffer = True
        self.camera.position = (self.camera.position + move).astype(f32)

    def Resize(self, width, height=None):
        self.viewSize = tuple(width) if height is None else (int(width), int(height))
        self.optixRenderer.Resize(self.viewSize)   # a new size starts a new image
        self.clearedToWhite = False
        self.samples = 0

    # -- one iteration of OptixView::Run (OptixView.cpp:132-185) ------------------------
    def Step(self) -> bool:
        """Returns True when a frame was rendered and added."""
        r = self.optixRenderer
        if self.maxSamples < 0 or self.samples < self.maxSamples:
            r.SetCamera(self.camera)
            if self.clearFrameBuffer:          # the frame about to be rendered becomes sample 1 (:141-151)
                self.clearFrameBuffer = False
                r.ClearAccumulation()
                self.clearedToWhite = True
                self.samples = 0
            r.RenderAccumulate(1, download=False)   # DrawOptix + AddNewFrameToBuffer, on the device
            self.samples += 1
            return True
        if self.samples == self.maxSamples and not self._finished_written:
            self._finished_written = True      # "Render is finished" (:163-168)
            if self.finished_exr is not None:
                r.WriteEXR(self.finished_exr)
        return False

    def Run(self, iterations: int, snapshot_every: int = 0, snapshot_png=None) -> int:
        """`iterations` loop trips (the reference loops until the window closes); every
        `snapshot_every` added frames the displayed image goes to `snapshot_png` — the
        headless stand-in for the preview window."""
        added = 0
        for _ in range(int(iterations)):
            if self.Step():
                added += 1
                if snapshot_every and snapshot_png and self.samples % snapshot_every == 0:
                    self.optixRenderer.WritePNG(snapshot_png)
        return added

    def Image(self) -> np.ndarray:
        """The framebuffer the display pass would read: running mean (maxSamples < 0) or
        sum / maxSamples (AddPathtracedFrame.frag:18-24)."""
        s, n = self.optixRenderer.DownloadAccumulation()
        if self.maxSamples < 0:
            return s / f32(max(n, 1))       # mix(fb, new, 1 / samples): the first frame after a clear overwrites the white
        # fb + new / maxSamples adds to whatever the clear left: white after a camera move (the reference's quirk, kept)
        return s / f32(self.maxSamples) + (f32(1.0) if self.clearedToWhite else f32(0.0))

    # -- checkpoint / resume -------------------------------------------------------------
    def Save(self, path):
        s, n = self.optixRenderer.DownloadAccumulation()
        np.savez(Path(path), sum=s, samples=np.int64(self.samples), spp=np.int64(n),
                 frame_id=np.int64(self.optixRenderer.GetFrameId()), size=np.asarray(self.viewSize),
                 position=self.camera.position, rotation=self.camera.rotation, white=np.bool_(self.clearedToWhite))

    def Load(self, path):
        z = np.load(Path(path))
        assert tuple(z["size"]) == tuple(self.viewSize), "checkpoint was taken at another size"
        self.camera.position = z["position"].astype(f32)
        self.camera.rotation = z["rotation"].astype(f32)
        self.optixRenderer.UploadAccumulation(z["sum"], int(z["spp"]))
        self.optixRenderer.SetFrameId(int(z["frame_id"]))
        self.samples = int(z["samples"])
        self.clearedToWhite = bool(z["white"]) if "white" in z else False
        self.clearFrameBuffer = False

    def close(self):
        self.optixRenderer.close()
