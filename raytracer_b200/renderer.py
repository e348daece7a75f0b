"""Host-side mirror of `Raytracing_Renderer` (src/raytracer/raytracing_renderer.odin).

Same names, same frame logic, the Vulkan calls replaced by the C ABI of librt.so:

    raytracing_renderer_init        :39-61    -> RaytracingRenderer(width, height)
    raytracing_renderer_set_scene   :98-106   -> set_scene        (gpu_scene_init   -> rt_build_scene)
    raytracing_renderer_on_resize   :108-126  -> on_resize        (new image        -> rt_resize)
    raytracing_renderer_begin_frame :128-187  -> begin_frame      (Scene_Change log -> rt_update_*)
    raytracing_renderer_render_scene:194-270  -> render_scene     (vkCmdTraceRaysKHR-> rt_trace_frame)
    raytracing_renderer_end_frame   :342-351  -> end_frame        (accumulation_frame += 1)
    the blit source image           :271-339  -> output_image / output_srgb8 (rt_readback*)

No window, swapchain or UI: the output is a headless device buffer read back on demand.
"""
from __future__ import annotations

from typing import Optional

import numpy as np

from . import librt
from .camera import Camera
from .scene import Scene, SceneChangeType

DEFAULT_BACKGROUND = (0.53, 0.81, 0.92)  # raytracing_renderer.odin:77


class RaytracingRenderer:
    def __init__(self, width: int, height: int, ctx=None, **ctx_kwargs):
        # `ctx`: an already created librt.Context (or any object with its methods: the CPU tests of the frame logic
        # pass a recorder); otherwise one is created here, which needs librt.so and a B200
        self.ctx = ctx if ctx is not None else librt.Context(**ctx_kwargs)
        self.scene: Optional[Scene] = None
        self.camera: Optional[Camera] = None
        self.accumulation_frame = 1  # raytracing_renderer.odin:45 (the first render resets it: camera.dirty)
        self.accumulation_limit: Optional[int] = None
        self.background_color = np.asarray(DEFAULT_BACKGROUND, np.float32)
        self.width, self.height = int(width), int(height)
        self.ctx.resize(self.width, self.height)
        self.framebuffer_resized = False

    def destroy(self):
        self.ctx.close()

    # ---- scene / camera -----------------------------------------------------------------------
    def set_scene(self, scene: Scene):
        assert scene is not None
        self.scene = scene
        self.scene_info = self.ctx.build_scene(scene)
        self.scene_info_meshes = len(scene.meshes)
        scene.changes.clear()

    def set_camera(self, camera: Camera):
        self.camera = camera

    def on_resize(self, width: int, height: int):
        self.width, self.height = int(width), int(height)
        self.ctx.resize(self.width, self.height)
        self.accumulation_frame = 0

    def set_accumulation_limit(self, new_limit: int):
        """The "Set Accumulation Limit" dialog (ui.odin:519-531): a positive limit at or below the current frame restarts
        the accumulation; zero or negative removes the limit."""
        if new_limit > 0:
            if self.accumulation_frame >= new_limit:
                self.accumulation_frame = 0
            self.accumulation_limit = int(new_limit)
        else:
            self.accumulation_limit = None

    def resume_from(self, image: np.ndarray, frames_done: int):
        """Checkpoint/resume of a long accumulation: `image` is the RGBA32F running mean over frames
        0 .. frames_done-1 (as `output_image` returned it); the next render_scene continues with frame `frames_done`."""
        self.ctx.upload_accum(image)
        self.accumulation_frame = int(frames_done)
        if self.camera is not None:
            self.camera.dirty = False

    # ---- frame --------------------------------------------------------------------------------
    def begin_frame(self):
        if self.framebuffer_resized:
            self.on_resize(self.width, self.height)
            self.framebuffer_resized = False
        sc = self.scene
        if sc is not None and sc.changes:
            self.accumulation_frame = 0
            n_meshes = self.scene_info_meshes
            for ch in sc.changes:
                t = ch.type
                if t == SceneChangeType.Material_Changed:
                    self.ctx.update_material(ch.index, sc.materials[ch.index])
                elif t == SceneChangeType.Object_Material_Changed:
                    self.ctx.update_object_material(ch.index, sc.objects[ch.index].material_index)
                elif t == SceneChangeType.Object_Transform_Changed:
                    self.ctx.update_object_transform(ch.index, sc.objects[ch.index].transform.model_matrix)
                elif t == SceneChangeType.Object_Mesh_Changed:
                    self.ctx.update_object_mesh(ch.index, sc.objects[ch.index].mesh_index)
                elif t == SceneChangeType.Material_Added:
                    self.ctx.add_material(sc.materials[ch.index])
                elif t == SceneChangeType.Material_Removed:
                    self.ctx.remove_material(ch.index)
                elif t == SceneChangeType.Object_Added:
                    o = sc.objects[ch.index]
                    if o.mesh_index >= n_meshes:
                        # the object instances a mesh added after the last build: meshes only travel with rt_build_scene
                        # (the reference has no change event for meshes at all), so the whole scene is built again
                        self.set_scene(sc)
                        return
                    self.ctx.add_object(o.mesh_index, o.material_index, o.transform.model_matrix)
                elif t == SceneChangeType.Object_Removed:
                    self.ctx.remove_object(ch.index)
            sc.changes.clear()

    def render_scene(self):
        cam = self.camera
        if cam.dirty:
            self.accumulation_frame = 0
            cam.dirty = False
        if self.accumulation_limit is None or self.accumulation_frame < self.accumulation_limit:
            self.ctx.trace_frame(cam.ubo(), self.background_color, self.accumulation_frame)

    def end_frame(self):
        if self.accumulation_limit is None or self.accumulation_frame < self.accumulation_limit:
            self.accumulation_frame += 1

    def render_frames(self, n: int):
        """n iterations of the application loop (application.odin:109-115) with nothing changing."""
        self.begin_frame()
        cam = self.camera
        if cam.dirty:
            self.accumulation_frame = 0
            cam.dirty = False
        if self.accumulation_limit is not None:
            n = max(0, min(n, self.accumulation_limit - self.accumulation_frame))
        if n:
            self.ctx.trace_frames(cam.ubo(), self.background_color, self.accumulation_frame, n)
            self.accumulation_frame += n

    # ---- output -------------------------------------------------------------------------------
    def output_image(self) -> np.ndarray:
        """The RGBA32F accumulation image the reference blits from."""
        return self.ctx.readback()

    def output_srgb8(self) -> np.ndarray:
        """What the blit to the B8G8R8A8_SRGB swapchain shows, as RGBA8."""
        return self.ctx.readback_srgb8()
