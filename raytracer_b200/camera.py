"""Camera mirror of `src/raytracer/camera.odin` (fp32, column-major matrices).

`Camera_UBO` (camera.odin:15-20) is four mat4s: projection, view, inverse_view,
inverse_projection = 64 floats = 256 B, which is what `rt_trace_frame` takes.  Matrix
conventions follow Odin `core:math/linalg/glsl` (right-handed lookAt, GL perspective) with
`proj[1][1] *= -1` (camera.odin:80) so row 0 of the image is the top.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32

CAMERA_SPEED = F32(5.0)  # camera.odin:7
CAMERA_SENSIVITY = F32(0.001)  # camera.odin:8


def _normalize(v):
    v = np.asarray(v, F32)
    return (v / F32(np.sqrt(F32(np.dot(v, v))))).astype(F32)


def mat4_look_at(eye, centre, up) -> np.ndarray:
    eye = np.asarray(eye, F32)
    f = _normalize(np.asarray(centre, F32) - eye)
    s = _normalize(np.cross(f, np.asarray(up, F32)).astype(F32))
    u = np.cross(s, f).astype(F32)
    m = np.eye(4, dtype=F32)
    m[0, :3] = s
    m[1, :3] = u
    m[2, :3] = -f
    m[0, 3] = -np.dot(s, eye)
    m[1, 3] = -np.dot(u, eye)
    m[2, 3] = np.dot(f, eye)
    return m


def mat4_perspective(fovy, aspect, near, far) -> np.ndarray:
    t = F32(np.tan(F32(0.5) * F32(fovy)))
    m = np.zeros((4, 4), dtype=F32)
    m[0, 0] = F32(1) / (F32(aspect) * t)
    m[1, 1] = F32(1) / t
    m[2, 2] = -(F32(far) + F32(near)) / (F32(far) - F32(near))
    m[3, 2] = F32(-1)
    m[2, 3] = F32(-2) * F32(far) * F32(near) / (F32(far) - F32(near))
    return m


class Camera:
    """camera.odin:31-43; default pose is the application's (application.odin:52)."""

    def __init__(self, position=(0.0, 0.0, -3.0), aspect=1.0, target=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0),
                 fovy_degrees=45.0, near=0.1, far=1000.0):
        self.position = np.asarray(position, F32)
        self.up = np.asarray(up, F32)
        self.aspect = F32(aspect)
        self.speed = CAMERA_SPEED
        self.sensivity = CAMERA_SENSIVITY
        self.fovy = F32(np.radians(F32(fovy_degrees)))  # camera.odin:76 hard-codes 45 degrees
        self.near = F32(near)
        self.far = F32(far)
        self.dirty = False
        self.look_at(target, up)
        self.update_matrices()

    def look_at(self, target, up) -> None:  # camera.odin:62-65
        self.forward = _normalize(np.asarray(target, F32) - self.position)
        self.right = np.cross(self.forward, self.up).astype(F32)

    def on_resize(self, aspect) -> None:  # camera.odin:68-71
        self.aspect = F32(aspect)
        self.update_matrices()

    def update_matrices(self) -> None:  # camera.odin:74-85
        self.view = mat4_look_at(self.position, self.position + self.forward, self.up)
        self.proj = mat4_perspective(self.fovy, self.aspect, self.near, self.far)
        self.proj[1, 1] *= F32(-1)
        self.inverse_view = np.linalg.inv(self.view.astype(np.float64)).astype(F32)
        self.inverse_proj = np.linalg.inv(self.proj.astype(np.float64)).astype(F32)
        self.dirty = True

    def move(self, direction: str, delta_time: float) -> None:  # camera.odin:110-132
        step = self.speed * F32(delta_time)
        vec = {
            "Up": self.up, "Down": -self.up, "Forward": self.forward, "Backwards": -self.forward,
            "Right": self.right, "Left": -self.right,
        }[direction]
        self.position = (self.position + vec * step).astype(F32)
        self.update_matrices()

    def ubo(self) -> np.ndarray:
        """The 64 floats of Camera_UBO, each matrix column-major (camera.odin:15-20)."""
        out = np.empty((4, 16), dtype=F32)
        for i, m in enumerate((self.proj, self.view, self.inverse_view, self.inverse_proj)):
            out[i] = np.ascontiguousarray(m.T).reshape(16)
        return out.reshape(64)
