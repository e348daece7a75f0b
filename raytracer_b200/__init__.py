"""B200-native path-tracing core behind the reference's renderer seam.

Host-side mirror of the reference's scene/camera/renderer interfaces; the compute path is the
hand-written CUDA library `raytracer_b200/csrc/librt.so` reached through the C ABI declared in
`include/rt_api.h`.  There is no CPU fallback: importing works anywhere, rendering needs a B200.
"""
from .scene import (Material, Mesh, Object, Scene, SceneChange, SceneChangeType, Transform, create_plane,
                    create_sphere, make_mesh, model_matrix_from_trs, MATERIAL_DTYPE, VERTEX_DTYPE)
from .camera import Camera
from .scene_loader import SceneLoadError, load_scene, load_scene_from_file, load_scene_from_gltf

__all__ = [
    "Material", "Mesh", "Object", "Scene", "SceneChange", "SceneChangeType", "Transform", "create_plane",
    "create_sphere", "make_mesh", "model_matrix_from_trs", "MATERIAL_DTYPE", "VERTEX_DTYPE", "Camera",
    "SceneLoadError", "load_scene", "load_scene_from_file", "load_scene_from_gltf",
]
