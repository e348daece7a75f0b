"""Compact scene snapshots (`.npz`): exactly the arrays that cross the C ABI.

Used for the committed parity fixtures (`tests/golden/scenes/*.npz`, made by
`tools/make_fixtures.py` from the reference's scene files through `scene_loader`), so GPU-box runs
do not need `/root/reference`.  A snapshot stores, per scene: the concatenated vertex array (24-B
records), the concatenated index array, per-mesh (vertex, index) offsets, per-object
(mesh, material, model[16]) and the 44-B material records.
"""
from __future__ import annotations

import numpy as np

from .scene import MATERIAL_DTYPE, VERTEX_DTYPE, Material, Mesh, Object, Scene, Transform

F32 = np.float32


def save_scene_npz(path: str, scene: Scene) -> None:
    nv = np.cumsum([0] + [len(m.vertices) for m in scene.meshes]).astype(np.int64)
    ni = np.cumsum([0] + [len(m.indices) for m in scene.meshes]).astype(np.int64)
    verts = np.concatenate([m.vertices for m in scene.meshes]) if scene.meshes else np.zeros(0, VERTEX_DTYPE)
    inds = np.concatenate([m.indices for m in scene.meshes]) if scene.meshes else np.zeros(0, np.uint32)
    om, omat, model = scene.packed_objects()
    np.savez_compressed(
        path, vertices=verts.view(F32).reshape(-1, 6), indices=inds.astype(np.uint32), mesh_vertex_offsets=nv,
        mesh_index_offsets=ni, object_mesh=om, object_material=omat, object_model=model,
        materials=scene.packed_materials().view(F32).reshape(-1, 11),
        mesh_names=np.array([m.name for m in scene.meshes]), object_names=np.array([o.name for o in scene.objects]),
        material_names=np.array([m.name for m in scene.materials]))


def load_scene_npz(path: str) -> Scene:
    z = np.load(path, allow_pickle=False)
    s = Scene()
    verts = np.ascontiguousarray(z["vertices"], F32).reshape(-1).view(VERTEX_DTYPE)
    inds = z["indices"].astype(np.uint32)
    nv, ni = z["mesh_vertex_offsets"], z["mesh_index_offsets"]
    for i in range(len(nv) - 1):
        s.meshes.append(Mesh(str(z["mesh_names"][i]), verts[nv[i]:nv[i + 1]].copy(), inds[ni[i]:ni[i + 1]].copy()))
    mats = np.ascontiguousarray(z["materials"], F32).reshape(-1).view(MATERIAL_DTYPE)
    for i, r in enumerate(mats):
        s.materials.append(Material(str(z["material_names"][i]), tuple(r["albedo"]), tuple(r["emission_color"]),
                                    F32(r["emission_power"]), F32(r["roughness"]), F32(r["metallic"]),
                                    F32(r["transmission"]), F32(r["ior"])))
    for i in range(len(z["object_mesh"])):
        tr = Transform(model_matrix=np.ascontiguousarray(z["object_model"][i], F32))
        s.objects.append(Object(str(z["object_names"][i]), tr, int(z["object_mesh"][i]), int(z["object_material"][i])))
    return s
