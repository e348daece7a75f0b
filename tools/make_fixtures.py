"""Regenerates tests/golden/ from the reference's scene files.  Runs in the build container only
(it reads /root/reference, which does not exist on the GPU box):

    python tools/make_fixtures.py

1. tests/golden/scenes/<name>.npz  — the reference's scenes as the reference loader sees them
   (raytracer_b200.scene_loader restates src/raytracer/scene_loader.odin), stored as the arrays that
   cross the C ABI.
2. tests/golden/oracle_<name>.npz  — oracle outputs on those scenes (frames, hit tables, counters),
   so that `-m "not gpu"` tests pin the oracle against drift and `-m gpu` tests can compare against
   committed numbers as well as against the live oracle.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import raytracer_b200 as rb  # noqa: E402
from oracle.binding import OracleScene  # noqa: E402
from raytracer_b200.scene_io import load_scene_npz, save_scene_npz  # noqa: E402

REF = "/root/reference"
SCENES = {
    "cornell_box": "scenes/cornell_box.json",
    "cornell_pbr": "scenes/cornell_pbr.glb",
    "conference": "models/conference.glb",
    "conference_mirror_chairs": "models/conference_mirror_chairs.glb",
    "conference_room_large_pbr": "models/conference_room_large_pbr.glb",
    "multi_dispersion": "scenes/multi-dispersion.gltf",
    "cornell_box_gltf": "models/cornell_box/scene_with_light.glb",
}
CLEAR = (0.53, 0.81, 0.92)


def sponza_layout():
    """Per-primitive triangle counts, material indices and the materials as the loader reads them, from the
    glTF JSON (present) of the Sponza model whose geometry blob is missing: raytracer_b200/data/sponza_layout.npz."""
    import json

    doc = json.load(open(os.path.join(REF, "models/sponza/sponza.gltf")))
    acc = doc["accessors"]
    counts, mats = [], []
    for m in doc["meshes"]:
        for p in m["primitives"]:
            counts.append(acc[p["indices"]]["count"] // 3)
            mats.append(int(p.get("material", 0)))
    table = []
    for m in doc["materials"]:  # scene_loader.odin:80-99 with cgltf defaults
        pbr = m.get("pbrMetallicRoughness")
        albedo = (pbr.get("baseColorFactor", (1, 1, 1, 1))[:3] if pbr is not None else (0, 0, 0))
        rough = pbr.get("roughnessFactor", 1.0) if pbr is not None else 0.0
        metal = pbr.get("metallicFactor", 1.0) if pbr is not None else 0.0
        emis = m.get("emissiveFactor", (0, 0, 0))
        ext = m.get("extensions", {})
        power = ext["KHR_materials_emissive_strength"].get("emissiveStrength", 1.0) if "KHR_materials_emissive_strength" in ext else 0.0
        table.append(list(albedo) + list(emis) + [power, rough, metal])
    out = os.path.join(ROOT, "raytracer_b200", "data")
    os.makedirs(out, exist_ok=True)
    np.savez_compressed(os.path.join(out, "sponza_layout.npz"), tri_counts=np.asarray(counts, np.int32),
                        material_index=np.asarray(mats, np.int32), materials=np.asarray(table, np.float32))
    print(f"sponza layout: {len(counts)} primitives, {sum(counts)} triangles, {len(table)} materials")


def main():
    sponza_layout()
    gold = os.path.join(ROOT, "tests", "golden")
    os.makedirs(os.path.join(gold, "scenes"), exist_ok=True)
    for name, rel in SCENES.items():
        scene = rb.load_scene(os.path.join(REF, rel))
        p = os.path.join(gold, "scenes", name + ".npz")
        save_scene_npz(p, scene)
        back = load_scene_npz(p)
        assert back.num_instanced_triangles() == scene.num_instanced_triangles()
        print(f"{name}: {len(scene.objects)} objects, {scene.num_instanced_triangles()} tris, "
              f"{len(scene.lights())} lights -> {os.path.getsize(p)} B")
    # oracle goldens on the two Cornell scenes (small frames, 3 accumulated frames)
    for name, (W, H) in {"cornell_box": (96, 96), "cornell_pbr": (96, 96)}.items():
        scene = load_scene_npz(os.path.join(gold, "scenes", name + ".npz"))
        orc = OracleScene(scene).build_bvh()
        cam = rb.Camera(aspect=W / H) if name == "cornell_box" else rb.Camera(position=(0.0, 1.0, 4.5), aspect=W / H,
                                                                               target=(0.0, 1.0, 0.0))
        img = None
        cnt = np.zeros(6, np.uint64)
        frames = []
        for f in range(3):
            img = orc.render(cam.ubo(), CLEAR, f, W, H, image=img, counters=cnt)
            frames.append(img.copy())
        nodes, tris, keys, ids = orc.bvh_arrays()
        np.savez_compressed(os.path.join(gold, f"oracle_{name}.npz"), frames=np.stack(frames), counters=cnt,
                            camera=cam.ubo(), n_nodes=len(nodes), bvh_depth=orc.bvh_depth,
                            nodes_crc=np.frombuffer(nodes.tobytes(), np.uint8).astype(np.uint64).sum(),
                            sorted_ids_head=ids[:256], morton_head=keys[:256])
        print(f"oracle_{name}: mean {frames[-1][..., :3].mean():.6f} counters {cnt}")


if __name__ == "__main__":
    main()
