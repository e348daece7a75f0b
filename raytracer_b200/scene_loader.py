"""Fixture loaders with the reference's exact quirks (`src/raytracer/scene_loader.odin`).

* `load_scene_from_gltf` (:37-190): glTF/GLB subset.  Materials from factors only; one Mesh +
  one Object per primitive; the node's LOCAL transform only (parents ignored, :107-108); no
  textures; `emission_power` only from KHR_materials_emissive_strength (:91-96); cgltf defaults
  (baseColor 1, metallic 1, roughness 1) when `pbrMetallicRoughness` is present but a factor is not.
* `load_scene_from_file` (:193-253): the JSON scene format (dead code in the reference, used here
  for config 1).  Odin map iteration order is unspecified; this loader DEFINES it as file order.

cgltf itself ships with the Odin compiler and is not in the reference tree ("parity unpinned");
the subset below is restated from the glTF 2.0 specification.
"""
from __future__ import annotations

import base64
import json
import os
import struct
from typing import Optional, Tuple

import numpy as np

from .scene import (F32, Material, Mesh, Object, Scene, Transform, create_plane, create_sphere, make_mesh)


class SceneLoadError(Exception):
    """scene_loader.odin:30-34 (Invalid_File / Object_Material_Not_Found)."""


_COMPONENT = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_NCOMP = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


def _read_container(path: str) -> Tuple[dict, Optional[bytes]]:
    ext = os.path.splitext(path)[1]
    try:
        raw = open(path, "rb").read()
    except OSError as e:
        raise SceneLoadError(f"Invalid_File: {e}")
    if ext == ".gltf":
        return json.loads(raw.decode("utf-8")), None
    if ext == ".glb":
        magic, _ver, total = struct.unpack_from("<III", raw, 0)
        if magic != 0x46546C67:
            raise SceneLoadError("Invalid_File: bad GLB magic")
        off, doc, blob = 12, None, None
        while off < total:
            clen, ctype = struct.unpack_from("<II", raw, off)
            body = raw[off + 8: off + 8 + clen]
            if ctype == 0x4E4F534A:
                doc = json.loads(body.decode("utf-8"))
            elif ctype == 0x004E4942 and blob is None:
                blob = body
            off += 8 + clen
        if doc is None:
            raise SceneLoadError("Invalid_File: GLB without JSON chunk")
        return doc, blob
    raise SceneLoadError("Invalid_File: extension")  # scene_loader.odin:47-55


def _load_buffers(doc: dict, path: str, glb_blob: Optional[bytes]):
    out = []
    base = os.path.dirname(path)
    for i, b in enumerate(doc.get("buffers", [])):
        uri = b.get("uri")
        if uri is None:
            if glb_blob is None:
                raise SceneLoadError("Invalid_File: buffer without uri")
            out.append(glb_blob)
        elif uri.startswith("data:"):
            out.append(base64.b64decode(uri.split(",", 1)[1]))
        else:
            p = os.path.join(base, uri)
            if not os.path.exists(p):
                # several blobs are absent from the reference mount (.MISSING_LARGE_BLOBS)
                alt = os.environ.get("RT_ASSET_DIR")
                if alt and os.path.exists(os.path.join(alt, os.path.basename(uri))):
                    p = os.path.join(alt, os.path.basename(uri))
                else:
                    raise SceneLoadError(f"Invalid_File: missing buffer {uri}")
            out.append(open(p, "rb").read())
    return out


def _accessor(doc: dict, buffers, index: int) -> np.ndarray:
    acc = doc["accessors"][index]
    comp = np.dtype(_COMPONENT[acc["componentType"]])
    ncomp = _NCOMP[acc["type"]]
    count = acc["count"]
    if "bufferView" not in acc:
        return np.zeros((count, ncomp), dtype=comp)
    bv = doc["bufferViews"][acc["bufferView"]]
    start = bv.get("byteOffset", 0) + acc.get("byteOffset", 0)
    stride = bv.get("byteStride") or comp.itemsize * ncomp
    buf = buffers[bv["buffer"]]
    arr = np.ndarray((count, ncomp), dtype=comp, buffer=buf, offset=start, strides=(stride, comp.itemsize))
    arr = np.array(arr)
    if acc.get("normalized") and comp.kind in "iu":  # cgltf_accessor_read_float normalises
        arr = arr.astype(F32) / F32(np.iinfo(comp).max)
    return arr


def _node_local_matrix(n: dict) -> np.ndarray:
    """cgltf_node_transform_local: explicit matrix, else T*R*S in fp32. Returns 16 col-major floats."""
    if "matrix" in n:
        return np.asarray(n["matrix"], F32).reshape(16).copy()
    t = np.asarray(n.get("translation", (0, 0, 0)), F32)
    q = np.asarray(n.get("rotation", (0, 0, 0, 1)), F32)
    s = np.asarray(n.get("scale", (1, 1, 1)), F32)
    qx, qy, qz, qw = q
    one, two = F32(1), F32(2)
    lm = np.zeros(16, F32)
    lm[0] = (one - two * qy * qy - two * qz * qz) * s[0]
    lm[1] = (two * qx * qy + two * qz * qw) * s[0]
    lm[2] = (two * qx * qz - two * qy * qw) * s[0]
    lm[4] = (two * qx * qy - two * qz * qw) * s[1]
    lm[5] = (one - two * qx * qx - two * qz * qz) * s[1]
    lm[6] = (two * qy * qz + two * qx * qw) * s[1]
    lm[8] = (two * qx * qz + two * qy * qw) * s[2]
    lm[9] = (two * qy * qz - two * qx * qw) * s[2]
    lm[10] = (one - two * qx * qx - two * qy * qy) * s[2]
    lm[12:15] = t
    lm[15] = one
    return lm


def load_scene_from_gltf(path: str) -> Scene:
    doc, blob = _read_container(path)
    buffers = _load_buffers(doc, path, blob)
    scene = Scene()

    # ---------- materials (scene_loader.odin:80-99) ----------
    for m in doc.get("materials", []):
        mat = Material(name=m.get("name", ""))
        if "pbrMetallicRoughness" in m:
            pbr = m["pbrMetallicRoughness"]
            mat.albedo = tuple(pbr.get("baseColorFactor", (1.0, 1.0, 1.0, 1.0))[:3])
            mat.roughness = pbr.get("roughnessFactor", 1.0)
            mat.metallic = pbr.get("metallicFactor", 1.0)
        mat.emission_color = tuple(m.get("emissiveFactor", (0.0, 0.0, 0.0)))
        ext = m.get("extensions", {})
        if "KHR_materials_emissive_strength" in ext:
            mat.emission_power = ext["KHR_materials_emissive_strength"].get("emissiveStrength", 1.0)
        else:
            mat.emission_power = 0.0
        scene.materials.append(mat)

    # ---------- meshes + objects (scene_loader.odin:101-187) ----------
    for n in doc.get("nodes", []):
        if "mesh" not in n:
            continue
        world = _node_local_matrix(n)
        gm = doc["meshes"][n["mesh"]]
        for p in gm["primitives"]:
            attrs = p["attributes"]
            if "POSITION" not in attrs:
                raise SceneLoadError("primitive without POSITION")  # assert at :128
            pos = _accessor(doc, buffers, attrs["POSITION"]).astype(F32)
            nrm = _accessor(doc, buffers, attrs["NORMAL"]).astype(F32) if "NORMAL" in attrs else None
            if "indices" in p:
                inds = _accessor(doc, buffers, p["indices"]).reshape(-1).astype(np.uint32)
            else:
                inds = np.arange(len(pos), dtype=np.uint32)
            mesh = make_mesh(gm.get("name", ""), pos, nrm, inds)
            mesh_index = scene.add_mesh(mesh)
            tr = Transform(
                position=np.asarray(n.get("translation", (0, 0, 0)), F32),
                rotation=np.asarray(n.get("rotation", (0, 0, 0, 1))[:3], F32),
                scale=np.asarray(n.get("scale", (1, 1, 1)), F32),
                model_matrix=world.copy(),
            )
            obj = Object(n.get("name", ""), tr, mesh_index, int(p["material"]) if "material" in p else 0)
            scene.objects.append(obj)
    return scene


def load_scene_from_file(path: str) -> Scene:
    """JSON scenes (scene_loader.odin:193-253). Materials/objects in FILE order (defined here)."""
    try:
        doc = json.loads(open(path, "rb").read().decode("utf-8"))
    except (OSError, ValueError) as e:
        raise SceneLoadError(f"Invalid_File: {e}")
    scene = Scene()
    for name, m in doc.get("materials", {}).items():
        scene.add_material(Material(
            name=name,
            albedo=tuple(m.get("albedo", (0, 0, 0))),
            emission_color=tuple(m.get("emission_color", (0, 0, 0))),
            emission_power=m.get("emission_power", 0.0),
            roughness=m.get("roughness", 0.0),
            metallic=m.get("metallic", 0.0),
            transmission=m.get("transmission", 0.0),
            ior=m.get("ior", 0.0),
        ))
    meshes = {"Plane": scene.add_mesh(create_plane()), "Sphere": scene.add_mesh(create_sphere())}
    names = [m.name for m in scene.materials]
    for name, o in doc.get("objects", {}).items():
        if o.get("material") not in names:
            raise SceneLoadError(f"Object_Material_Not_Found: '{name}' uses '{o.get('material')}'")
        tr = o.get("transform", {})
        scene.add_object(name, meshes[o.get("mesh", "Plane")], names.index(o["material"]),
                         tr.get("position", (0, 0, 0)), tr.get("rotation", (0, 0, 0)), tr.get("scale", (0, 0, 0)))
    return scene


def load_scene(path: str) -> Scene:
    """Dispatch on extension. The application only calls the glTF loader (application.odin:41);
    `.json` goes to the JSON loader so config 1 is runnable."""
    if path.endswith(".json"):
        return load_scene_from_file(path)
    return load_scene_from_gltf(path)
