#!/usr/bin/env python3
"""Literal constants and debug names of the reference's SHIPPED ray-tracing shader binaries.

The reference commits the compiled SPIR-V next to the GLSL (`shaders/rgen.spv`, `rchit.spv`, `rmiss.spv`,
`restir_rgen.spv`, `restir_rchit.spv`, `restir_rmiss.spv`); `src/raytracer/raytracing_renderer.odin:62-71` loads
those, not the sources.  No SPIR-V tool exists in this image, so this walks the binary itself (SPIR-V 1.x physical
layout: 5-word header, then `(word_count << 16) | opcode` instructions) and collects

  * every scalar OpConstant of a 32-bit float / int type (the numbers the driver would run with),
  * OpName strings (function and variable names glslang leaves in), and OpEntryPoint execution models,
  * the static call graph (OpFunction / OpFunctionCall): which of the GLSL functions the shipped `main` reaches,

into `tests/golden/spirv_constants.json`.  `tests/test_oracle.py::test_oracle_constants_are_the_shipped_ones`
then pins the CPU oracle's literals against that file: SURVEY.md §8(c) pin (2).  The reference mount is needed only
to REGENERATE the fixture (`python tools/spirv_constants.py`), never at test time.
"""
import json
import os
import struct
import sys

OP_NAME, OP_ENTRY_POINT, OP_TYPE_INT, OP_TYPE_FLOAT, OP_CONSTANT = 5, 15, 21, 22, 43
OP_FUNCTION, OP_FUNCTION_END, OP_FUNCTION_CALL = 54, 56, 57
EXEC_MODELS = {5313: "RayGenerationKHR", 5314: "IntersectionKHR", 5315: "AnyHitKHR", 5316: "ClosestHitKHR",
               5317: "MissKHR", 5318: "CallableKHR", 0: "Vertex", 4: "Fragment", 5: "GLCompute"}
MODULES = ["rgen.spv", "rchit.spv", "rmiss.spv", "shadow.spv", "shadow_rchit.spv",
           "restir_rgen.spv", "restir_rchit.spv", "restir_rmiss.spv"]


def _string(words):
    raw = b"".join(struct.pack("<I", w) for w in words)
    return raw.split(b"\0", 1)[0].decode("utf-8", "replace")


def parse(path):
    data = open(path, "rb").read()
    if len(data) < 20 or len(data) % 4:
        raise ValueError(f"{path}: not a SPIR-V module (size {len(data)})")
    w = struct.unpack(f"<{len(data) // 4}I", data)
    if w[0] != 0x07230203:
        raise ValueError(f"{path}: bad magic {w[0]:#x}")
    out = {"version": f"{(w[1] >> 16) & 0xff}.{(w[1] >> 8) & 0xff}", "bound": w[3], "entry_points": [], "names": [],
           "f32": [], "u32": [], "i32": []}
    types = {}
    id_name, calls, cur = {}, {}, None
    i = 5
    while i < len(w):
        n, op = w[i] >> 16, w[i] & 0xffff
        if n == 0:
            raise ValueError(f"{path}: zero-length instruction at word {i}")
        a = w[i + 1:i + n]
        if op == OP_TYPE_INT:
            types[a[0]] = ("i32" if a[2] else "u32") if a[1] == 32 else None
        elif op == OP_TYPE_FLOAT:
            types[a[0]] = "f32" if a[1] == 32 else None
        elif op == OP_CONSTANT:
            t = types.get(a[0])
            if t == "f32":
                out["f32"].append(struct.unpack("<f", struct.pack("<I", a[2]))[0])
            elif t == "u32":
                out["u32"].append(a[2])
            elif t == "i32":
                out["i32"].append(struct.unpack("<i", struct.pack("<I", a[2]))[0])
        elif op == OP_NAME:
            s = _string(a[1:])
            if s:
                out["names"].append(s)
                id_name[a[0]] = s
        elif op == OP_FUNCTION:
            cur = a[1]
            calls[cur] = set()
        elif op == OP_FUNCTION_END:
            cur = None
        elif op == OP_FUNCTION_CALL and cur is not None:
            calls[cur].add(a[2])
        elif op == OP_ENTRY_POINT:
            out["entry_points"].append([EXEC_MODELS.get(a[0], str(a[0])), _string(a[2:])])
        i += n
    for k in ("f32", "u32", "i32"):
        out[k] = sorted(set(out[k]))
    out["names"] = sorted(set(out["names"]))
    nm = lambda f: id_name.get(f, f"%{f}").split("(")[0]
    out["calls"] = {nm(f): sorted(nm(c) for c in cs) for f, cs in calls.items()}
    seen, todo = set(), ["main"]
    while todo:                       # functions reachable from the entry point
        f = todo.pop()
        if f not in seen:
            seen.add(f)
            todo += out["calls"].get(f, [])
    out["reachable"] = sorted(seen)
    return out


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "spirv_constants.json")
    res = {}
    for m in MODULES:
        p = os.path.join(ref, "shaders", m)
        if os.path.exists(p):
            res[m] = parse(p)
            # float32 values are stored as their exact bit patterns too, so the test compares bits, not decimals
            res[m]["f32_bits"] = [struct.unpack("<I", struct.pack("<f", v))[0] for v in res[m]["f32"]]
    with open(dst, "w") as f:
        json.dump(res, f, indent=1, sort_keys=True)
    for m, r in res.items():
        print(f"{m}: SPIR-V {r['version']}, {r['entry_points']}, {len(r['f32'])} f32, {len(r['u32'])} u32, "
              f"{len(r['i32'])} i32 constants, {len(r['names'])} names")


if __name__ == "__main__":
    main()
