"""Host-side scene model: the data the drop-in boundary consumes.

Mirrors the reference's CPU scene (`src/raytracer/scene.odin`): `Vertex` (:10-13),
`Object` (:41-46), `Transform` (:48-54), `Mesh` (:56-60), `Material` (:66-70), the change log
(:15-39, :95-163), the model matrix T*Rx*Ry*Rz*S (:213-224) and the two procedural meshes
`create_sphere` (:242-296) / `create_plane` (:298-320).  Everything here is plain numpy on the
host; nothing touches the GPU.  The arrays are laid out exactly like the reference's structs so
they can cross the C ABI (`include/rt_api.h`) without marshalling.
"""
from __future__ import annotations

import dataclasses
import enum
from typing import List, Optional, Sequence

import numpy as np

F32 = np.float32

# --- packed record layouts that cross the C ABI -------------------------------------------------
# Vertex: pos3 + normal3, 24 B (scene.odin:10-13, ray_common.glsl:48-51)
VERTEX_DTYPE = np.dtype([("pos", F32, 3), ("normal", F32, 3)])
# Material_Data: 44 B scalar layout (gpu_scene.odin:17-21, ray_common.glsl:3-11)
MATERIAL_DTYPE = np.dtype(
    [
        ("albedo", F32, 3),
        ("emission_color", F32, 3),
        ("emission_power", F32),
        ("roughness", F32),
        ("metallic", F32),
        ("transmission", F32),
        ("ior", F32),
    ]
)
assert VERTEX_DTYPE.itemsize == 24 and MATERIAL_DTYPE.itemsize == 44


class SceneChangeType(enum.Enum):
    """scene.odin:15-24."""

    Material_Changed = 0
    Material_Added = 1
    Material_Removed = 2
    Object_Material_Changed = 3
    Object_Added = 4
    Object_Removed = 5
    Object_Transform_Changed = 6
    Object_Mesh_Changed = 7


@dataclasses.dataclass
class SceneChange:
    type: SceneChangeType
    index: int = 0


@dataclasses.dataclass
class Material:
    name: str = ""
    albedo: Sequence[float] = (0.0, 0.0, 0.0)
    emission_color: Sequence[float] = (0.0, 0.0, 0.0)
    emission_power: float = 0.0
    roughness: float = 0.0
    metallic: float = 0.0
    transmission: float = 0.0
    ior: float = 0.0

    def packed(self) -> np.ndarray:
        rec = np.zeros((), dtype=MATERIAL_DTYPE)
        rec["albedo"] = np.asarray(self.albedo, dtype=F32)
        rec["emission_color"] = np.asarray(self.emission_color, dtype=F32)
        for k in ("emission_power", "roughness", "metallic", "transmission", "ior"):
            rec[k] = F32(getattr(self, k))
        return rec


@dataclasses.dataclass
class Mesh:
    name: str
    vertices: np.ndarray  # (V,) VERTEX_DTYPE
    indices: np.ndarray  # (I,) uint32

    @property
    def num_triangles(self) -> int:
        return len(self.indices) // 3


@dataclasses.dataclass
class Transform:
    position: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros(3, F32))
    rotation: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros(3, F32))
    scale: np.ndarray = dataclasses.field(default_factory=lambda: np.ones(3, F32))
    # column-major 4x4, stored as the 16 floats the GPU sees (model[c*4+r])
    model_matrix: np.ndarray = dataclasses.field(default_factory=lambda: np.eye(4, dtype=F32).T.reshape(16).copy())


@dataclasses.dataclass
class Object:
    name: str
    transform: Transform
    mesh_index: int = 0
    material_index: int = 0


# --- glm-style matrix helpers (Odin core:math/linalg/glsl semantics, fp32) ---------------------
def _mat4_translate(v) -> np.ndarray:
    m = np.eye(4, dtype=F32)
    m[:3, 3] = np.asarray(v, F32)
    return m


def _mat4_scale(v) -> np.ndarray:
    m = np.eye(4, dtype=F32)
    m[0, 0], m[1, 1], m[2, 2] = np.asarray(v, F32)
    return m


def _mat4_rotate(axis, radians) -> np.ndarray:
    c = F32(np.cos(F32(radians)))
    s = F32(np.sin(F32(radians)))
    a = np.asarray(axis, F32)
    a = a / F32(np.sqrt(np.dot(a, a)))
    t = a * (F32(1) - c)
    m = np.eye(4, dtype=F32)
    m[0, 0] = c + t[0] * a[0]
    m[1, 0] = t[0] * a[1] + s * a[2]
    m[2, 0] = t[0] * a[2] - s * a[1]
    m[0, 1] = t[1] * a[0] - s * a[2]
    m[1, 1] = c + t[1] * a[1]
    m[2, 1] = t[1] * a[2] + s * a[0]
    m[0, 2] = t[2] * a[0] + s * a[1]
    m[1, 2] = t[2] * a[1] - s * a[0]
    m[2, 2] = c + t[2] * a[2]
    return m


def model_matrix_from_trs(position, rotation_deg, scale) -> np.ndarray:
    """T * Rx * Ry * Rz * S with Euler angles in degrees (scene.odin:213-224).

    Returns the 16 column-major floats."""
    r = np.radians(np.asarray(rotation_deg, F32)).astype(F32)
    rot = _mat4_rotate((1, 0, 0), r[0]) @ _mat4_rotate((0, 1, 0), r[1]) @ _mat4_rotate((0, 0, 1), r[2])
    m = (_mat4_translate(position) @ rot.astype(F32) @ _mat4_scale(scale)).astype(F32)
    return np.ascontiguousarray(m.T).reshape(16)


class Scene:
    """scene.odin:31-39 plus the mutation API (:95-163) that feeds the change log."""

    def __init__(self) -> None:
        self.meshes: List[Mesh] = []
        self.objects: List[Object] = []
        self.materials: List[Material] = []
        self.changes: List[SceneChange] = []

    # -- construction -------------------------------------------------------------------------
    def add_material(self, material: Material) -> int:
        self.materials.append(material)
        self.changes.append(SceneChange(SceneChangeType.Material_Added, len(self.materials) - 1))
        return len(self.materials) - 1

    def add_mesh(self, mesh: Mesh) -> int:
        self.meshes.append(mesh)
        return len(self.meshes) - 1

    def add_object(self, name, mesh_index, material_index, position=(0, 0, 0), rotation=(0, 0, 0),
                   scale=(1, 1, 1)) -> int:
        if not (0 <= mesh_index < len(self.meshes)):
            raise IndexError("Invalid mesh index")  # scene.odin:178
        if not (0 <= material_index < len(self.materials)):
            raise IndexError("Invalid material index")  # scene.odin:179
        tr = Transform(np.asarray(position, F32), np.asarray(rotation, F32), np.asarray(scale, F32))
        tr.model_matrix = model_matrix_from_trs(tr.position, tr.rotation, tr.scale)
        self.objects.append(Object(name, tr, int(mesh_index), int(material_index)))
        self.changes.append(SceneChange(SceneChangeType.Object_Added, len(self.objects) - 1))
        return len(self.objects) - 1

    def delete_material(self, material_index: int) -> None:
        """scene_delete_material (scene.odin:104-116): unordered_remove (the last material takes the freed slot), objects that
        used the removed material fall back to material 0; objects that used the moved one follow it (defined here: the
        reference leaves their index dangling)."""
        last = len(self.materials) - 1
        self.materials[material_index] = self.materials[last]
        self.materials.pop()
        for i, o in enumerate(self.objects):
            if o.material_index == material_index:
                self.update_object_material(i, 0)
            elif o.material_index == last:
                o.material_index = material_index
        self.changes.append(SceneChange(SceneChangeType.Material_Removed, material_index))

    def delete_object(self, object_idx: int) -> None:
        """Object_Removed (scene.odin:21; the reference declares the change but ships no procedure that raises it): ordered
        removal, later objects move down one index."""
        del self.objects[object_idx]
        self.changes.append(SceneChange(SceneChangeType.Object_Removed, object_idx))

    # -- edits (each appends a Scene_Change, scene.odin:116-163) ------------------------------
    def update_material(self, material_idx: int, material: Material) -> None:
        self.materials[material_idx] = material
        self.changes.append(SceneChange(SceneChangeType.Material_Changed, material_idx))

    def update_object_material(self, object_idx: int, new_material_idx: int) -> None:
        self.objects[object_idx].material_index = int(new_material_idx)
        self.changes.append(SceneChange(SceneChangeType.Object_Material_Changed, object_idx))

    def _retransform(self, object_idx: int) -> None:
        tr = self.objects[object_idx].transform
        tr.model_matrix = model_matrix_from_trs(tr.position, tr.rotation, tr.scale)
        self.changes.append(SceneChange(SceneChangeType.Object_Transform_Changed, object_idx))

    def update_object_position(self, object_idx: int, p) -> None:
        self.objects[object_idx].transform.position = np.asarray(p, F32)
        self._retransform(object_idx)

    def update_object_rotation(self, object_idx: int, r) -> None:
        self.objects[object_idx].transform.rotation = np.asarray(r, F32)
        self._retransform(object_idx)

    def update_object_scale(self, object_idx: int, s) -> None:
        self.objects[object_idx].transform.scale = np.asarray(s, F32)
        self._retransform(object_idx)

    def update_object_mesh(self, object_idx: int, mesh_idx: int) -> None:
        self.objects[object_idx].mesh_index = int(mesh_idx)
        self.changes.append(SceneChange(SceneChangeType.Object_Mesh_Changed, object_idx))

    # -- views the boundary consumes -----------------------------------------------------------
    def packed_materials(self) -> np.ndarray:
        out = np.zeros(len(self.materials), dtype=MATERIAL_DTYPE)
        for i, m in enumerate(self.materials):
            out[i] = m.packed()
        return out

    def packed_objects(self):
        """(mesh_index u32[n], material_index u32[n], model f32[n,16])."""
        n = len(self.objects)
        mesh = np.array([o.mesh_index for o in self.objects], dtype=np.uint32).reshape(n)
        mat = np.array([o.material_index for o in self.objects], dtype=np.uint32).reshape(n)
        model = np.zeros((n, 16), dtype=F32)
        for i, o in enumerate(self.objects):
            model[i] = o.transform.model_matrix
        return mesh, mat, model

    def num_instanced_triangles(self) -> int:
        return sum(self.meshes[o.mesh_index].num_triangles for o in self.objects)

    def lights(self) -> List[int]:
        """Object indices that the reference treats as lights (gpu_scene.odin:366-379)."""
        return [i for i, o in enumerate(self.objects)
                if float(self.materials[o.material_index].emission_power) > 0.0]


def make_mesh(name: str, positions, normals, indices) -> Mesh:
    positions = np.asarray(positions, F32).reshape(-1, 3)
    v = np.zeros(len(positions), dtype=VERTEX_DTYPE)
    v["pos"] = positions
    if normals is not None:
        v["normal"] = np.asarray(normals, F32).reshape(-1, 3)
    return Mesh(name, v, np.ascontiguousarray(np.asarray(indices, dtype=np.uint32).reshape(-1)))


def create_plane() -> Mesh:
    """Unit quad in z=0 facing +z, two triangles (scene.odin:298-320)."""
    pos = [(-0.5, -0.5, 0), (0.5, -0.5, 0), (0.5, 0.5, 0), (-0.5, 0.5, 0)]
    nrm = [(0, 0, 1)] * 4
    return make_mesh("Plane", pos, nrm, [0, 1, 2, 0, 2, 3])


def create_sphere(stacks: int = 32, slices: int = 32) -> Mesh:
    """UV sphere, radius 1, normals = positions (scene.odin:242-296).

    994 vertices / 1,984 triangles at the default 32x32.  Index order follows the reference:
    the pole fans are interleaved (north, south per slice) and come before the quad strips."""
    pos = [(0.0, 1.0, 0.0)]
    pi = F32(np.pi)
    for i in range(stacks - 1):
        phi = pi * F32(i + 1) / F32(stacks)
        for j in range(slices):
            theta = F32(2.0) * pi * F32(j) / F32(slices)
            x = F32(np.sin(phi)) * F32(np.cos(theta))
            y = F32(np.cos(phi))
            z = F32(np.sin(phi)) * F32(np.sin(theta))
            pos.append((x, y, z))
    pos.append((0.0, -1.0, 0.0))
    nv = len(pos)
    idx: List[int] = []
    for i in range(slices):
        i0 = i + 1
        i1 = (i + 1) % slices + 1
        idx += [0, i1, i0]
        i0 = i + slices * (stacks - 2) + 1
        i1 = (i + 1) % slices + slices * (stacks - 2) + 1
        idx += [nv - 1, i0, i1]
    for j in range(stacks - 2):
        j0 = j * slices + 1
        j1 = (j + 1) * slices + 1
        for i in range(slices):
            i0 = j0 + i
            i1 = j0 + (i + 1) % slices
            i2 = j1 + (i + 1) % slices
            i3 = j1 + i
            idx += [i0, i1, i2, i0, i2, i3]
    p = np.asarray(pos, F32)
    return make_mesh("Sphere", p, p.copy(), idx)
