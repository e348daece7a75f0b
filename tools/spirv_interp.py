#!/usr/bin/env python3
"""An interpreter for the reference's SHIPPED ray-tracing shader binaries — test infrastructure, like oracle/.

`src/raytracer/raytracing_renderer.odin:62-71` loads `shaders/rgen.spv`, `rchit.spv` and `rmiss.spv`; those binaries ARE the
reference's implementation of the hot path, and they are the one artefact of it that can be evaluated on this box (no graphics
stack, no SPIR-V tool, no Odin).  This module walks the SPIR-V (physical layout: 5-word header, `(word_count << 16) | opcode`
instructions) and EXECUTES it: every OpFMul / OpFAdd / OpFDiv / OpDot / OpExtInst / OpBranchConditional / OpFunctionCall in
the order the binary states, on IEEE-754 binary32 values, with the Vulkan environment replaced by plain Python objects:

  * descriptor sets (objects / materials / lights SSBOs, camera UBO, push constants, storage image) and the buffer
    references behind ObjectData.vertex_address / index_address are Python trees filled from a `raytracer_b200.Scene`
    exactly as `gpu_scene.odin:317-428` fills the reference's buffers;
  * `OpTraceRayKHR` and the inline ray query (`OpRayQuery*KHR`) — the Vulkan driver's part, which no file of the reference
    defines — are callbacks; tools/spirv_golden.py answers them with the oracle's closest-hit traversal and then runs
    `rchit.spv` / `rmiss.spv` on the shared payload, as the shader binding table would;
  * the GLSL.std.450 built-ins whose precision Vulkan leaves to the driver (normalize, length, dot, cross, reflect, mix,
    smoothstep, pow, sin, cos, inverse) take the definitions of the arithmetic contract (oracle/oracle_math.h):
    those are restated here in numpy float32, sin/cos through the oracle's exported Cody-Waite/Cephes routine.

What this pins, and what nothing else in the repo can: evaluation order, association, branch structure and the order of
the random draws of the shipped binary — independently of anybody's reading of the GLSL.  tools/spirv_golden.py turns it
into committed golden frames (tests/golden/spirv_frames.npz) that the oracle (CPU tests) and the CUDA path (-m gpu tests)
must reproduce bit for bit.  Nothing under raytracer_b200/ imports this file.
"""
from __future__ import annotations

import struct

import numpy as np

F32 = np.float32

# storage classes
SC_UNIFORM_CONSTANT, SC_INPUT, SC_UNIFORM, SC_PRIVATE, SC_FUNCTION, SC_PUSH, SC_STORAGE_BUFFER = 0, 1, 2, 6, 7, 9, 12
SC_RAY_PAYLOAD, SC_HIT_ATTRIBUTE, SC_INCOMING_PAYLOAD, SC_PHYSICAL = 5338, 5339, 5342, 5349
DEC_BUILTIN, DEC_BINDING, DEC_SET = 11, 33, 34
BUILTINS = {5319: "LaunchId", 5320: "LaunchSize", 5321: "WorldRayOrigin", 5322: "WorldRayDirection", 5325: "RayTmin",
            5326: "RayTmax", 5327: "InstanceCustomIndex", 5330: "ObjectToWorld", 5331: "WorldToObject", 5332: "HitT",
            7: "PrimitiveId", 6: "InstanceId"}


class Ptr:
    """A pointer: a root cell (a one-element list) and a path of indices into the nested lists below it."""
    __slots__ = ("root", "path")

    def __init__(self, root, path=()):
        self.root, self.path = root, path

    def load(self):
        v = self.root[0]
        for i in self.path:
            v = v[i]
        return _copy(v)

    def store(self, val):
        val = _copy(val)
        if not self.path:
            self.root[0] = val
            return
        v = self.root[0]
        for i in self.path[:-1]:
            v = v[i]
        v[self.path[-1]] = val

    def at(self, idx):
        return Ptr(self.root, self.path + tuple(idx))


def _copy(v):
    return [_copy(x) for x in v] if type(v) is list else v


def _string(words):
    raw = b"".join(struct.pack("<I", w) for w in words)
    return raw.split(b"\0", 1)[0].decode("utf-8", "replace")


def _f(bits):
    return np.frombuffer(struct.pack("<I", bits & 0xFFFFFFFF), dtype=np.float32)[0]


def _bits(f):
    return struct.unpack("<I", struct.pack("<f", float(f)))[0] if not isinstance(f, np.float32) else int(np.float32(f).view(np.uint32))


def _signed(x, width=32):
    x &= (1 << width) - 1
    return x - (1 << width) if x >> (width - 1) else x


# ---- the arithmetic contract's built-ins (oracle/oracle_math.h), restated on numpy float32 ----------------------------
def gmin(x, y):
    return y if y < x else x


def gmax(x, y):
    return y if x < y else x


def dot(a, b):
    acc = a[0] * b[0] + a[1] * b[1]
    for k in range(2, len(a)):
        acc = acc + a[k] * b[k]
    return acc


def cross(a, b):
    return [a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]]


def length(a):
    return np.sqrt(dot(a, a))


def normalize(a):
    inv = F32(1.0) / length(a)
    return [x * inv for x in a]


def inverse3(m):
    """inverse of a 3x3 (list of columns): the transpose of (a1 x a2, a2 x a0, a0 x a1) / det  (oracle: normal_matrix)."""
    a0, a1, a2 = m
    c0, c1, c2 = cross(a1, a2), cross(a2, a0), cross(a0, a1)
    inv = F32(1.0) / dot(a0, c0)
    c = [[x * inv for x in col] for col in (c0, c1, c2)]
    return [[c[r][col] for r in range(3)] for col in range(3)]


class CommittedInstances:
    """What rayQueryGetIntersectionInstanceCustomIndexEXT returns under the policy "all" (tools/spirv_golden.py): with
    gl_RayFlagsTerminateOnFirstHitEXT the hit a ray query commits is whichever the traversal meets first — implementation
    defined — so the value stands for EVERY instance the ray meets in range, and compares equal to x only when x is the
    only one.  `instIdx == lightInstanceCustomIndex` (simple.rchit:381) then reads "no other instance occludes", the
    deterministic semantics both the oracle and the CUDA path define."""

    def __init__(self, first, others):
        self.first, self.others = int(first), bool(others)

    def __and__(self, mask):
        return self

    def __eq__(self, x):
        return (not self.others) and self.first == (int(x) & 0xFFFFFFFF)

    def __ne__(self, x):
        return not self.__eq__(x)

    def __int__(self):
        return self.first


class Module:
    def __init__(self, path):
        data = open(path, "rb").read()
        w = struct.unpack(f"<{len(data) // 4}I", data)
        if w[0] != 0x07230203:
            raise ValueError(f"{path}: not SPIR-V")
        self.path = path
        self.types, self.consts, self.names, self.member_names = {}, {}, {}, {}
        self.decor, self.globals_decl, self.functions, self.entry = {}, [], {}, None
        self.ext_sets = {}
        cur = None
        i = 5
        while i < len(w):
            n, op = w[i] >> 16, w[i] & 0xFFFF
            a = w[i + 1:i + n]
            i += n
            if cur is not None:
                if op == 56:  # FunctionEnd
                    cur = None
                elif op == 55:
                    cur["params"].append(a[1])
                elif op == 248:
                    cur["blocks"][a[0]] = []
                    cur["order"].append(a[0])
                    cur["_b"] = cur["blocks"][a[0]]
                else:
                    cur["_b"].append((op, a))
                continue
            if op == 5:
                self.names[a[0]] = _string(a[1:])
            elif op == 6:
                self.member_names[(a[0], a[1])] = _string(a[2:])
            elif op == 11:
                self.ext_sets[a[0]] = _string(a[1:])
            elif op == 15:
                self.entry = a[1]
            elif op == 71:
                self.decor.setdefault(a[0], {})[a[1]] = a[2] if len(a) > 2 else True
            elif op == 19:
                self.types[a[0]] = ("void",)
            elif op == 20:
                self.types[a[0]] = ("bool",)
            elif op == 21:
                self.types[a[0]] = ("int", a[1], a[2])
            elif op == 22:
                self.types[a[0]] = ("float", a[1])
            elif op == 23:
                self.types[a[0]] = ("vec", a[1], a[2])
            elif op == 24:
                self.types[a[0]] = ("mat", a[1], a[2])
            elif op == 25:
                self.types[a[0]] = ("image",)
            elif op == 28:
                self.types[a[0]] = ("array", a[1], a[2])
            elif op == 29:
                self.types[a[0]] = ("rtarray", a[1])
            elif op == 30:
                self.types[a[0]] = ("struct", list(a[1:]))
            elif op == 32:
                self.types[a[0]] = ("ptr", a[1], a[2])
            elif op == 33:
                self.types[a[0]] = ("func",)
            elif op == 39:
                self.types.setdefault(a[0], ("ptr", a[1], None))
            elif op == 5341:
                self.types[a[0]] = ("accel",)
            elif op == 4472:
                self.types[a[0]] = ("rayquery",)
            elif op == 41:
                self.consts[a[1]] = True
            elif op == 42:
                self.consts[a[1]] = False
            elif op == 43:
                t = self.types[a[0]]
                if t[0] == "float":
                    self.consts[a[1]] = _f(a[2])
                else:
                    self.consts[a[1]] = a[2] | ((a[3] << 32) if t[1] == 64 else 0)
            elif op == 44:
                self.consts[a[1]] = [self.consts[c] for c in a[2:]]
            elif op in (46, 1):
                self.consts[a[1]] = self.default(a[0])
            elif op == 59:
                self.globals_decl.append((a[1], a[0], a[2]))
            elif op == 54:
                cur = {"id": a[1], "params": [], "blocks": {}, "order": [], "type": a[0]}
                self.functions[a[1]] = cur
        self.by_name = {v: k for k, v in self.names.items()}
        self.stats = {"instructions": 0}

    # ---- types --------------------------------------------------------------------------------------------------------
    def default(self, tid):
        t = self.types[tid]
        k = t[0]
        if k == "bool":
            return False
        if k == "int":
            return 0
        if k == "float":
            return F32(0.0)
        if k == "vec":
            return [self.default(t[1]) for _ in range(t[2])]
        if k == "mat":
            return [self.default(t[1]) for _ in range(t[2])]
        if k == "array":
            return [self.default(t[1]) for _ in range(self.consts[t[2]])]
        if k == "rtarray":
            return []
        if k == "struct":
            return [self.default(m) for m in t[1]]
        return None  # pointers, acceleration structures, images, ray queries

    def function_named(self, prefix):
        """id of the function whose OpName is `prefix` or starts with `prefix(` (glslang mangles the signature in)."""
        for i, n in self.names.items():
            if i in self.functions and (n == prefix or n.startswith(prefix + "(")):
                return i
        raise KeyError(prefix)


class Machine:
    """One module + its global variables + the environment callbacks."""

    def __init__(self, module: Module, sincos=None):
        self.m = module
        self.g = dict(module.consts)  # id -> value: constants, then global pointers
        self.vars = {}                # name -> Ptr of every global variable
        self.builtin = {}
        self.addresses = {}           # device address -> root cell of a buffer_reference block
        self.trace_ray = None         # callback(machine, origin, tmin, direction, tmax, payload_ptr)
        self.ray_query = None         # callback(origin, tmin, direction, tmax) -> None | instance custom index
        self.images = {}              # variable name -> numpy (H, W, 4) float32 (storage images; texel conversion is the caller's)
        self.sincos = sincos          # callable(float32) -> (sin, cos) of the arithmetic contract
        for vid, tid, storage in module.globals_decl:
            pointee = module.types[tid][2]
            name = module.names.get(vid, f"%{vid}")
            # an image variable "holds" its own name: OpLoad hands it to OpImageRead / OpImageWrite, which look the array up
            p = Ptr([name if module.types[pointee][0] == "image" else module.default(pointee)])
            self.g[vid] = p
            self.vars[name] = p
            b = module.decor.get(vid, {}).get(DEC_BUILTIN)
            if b is not None:
                self.builtin[BUILTINS.get(b, b)] = p

    # ---- execution ----------------------------------------------------------------------------------------------------
    def run(self, fn=None, args=()):
        with np.errstate(all="ignore"):
            return self._call(self.m.entry if fn is None else fn, list(args))

    def _call(self, fid, args):
        fn = self.m.functions[fid]
        v = {}
        for pid, a in zip(fn["params"], args):
            v[pid] = a
        g = self.g

        def val(i):
            r = v.get(i)
            return r if r is not None or i in v else g[i]

        label, prev = fn["order"][0], None
        types = self.m.types
        n_inst = 0
        while True:
            nxt = None
            for op, a in fn["blocks"][label]:
                n_inst += 1
                if op == 61:      # Load
                    v[a[1]] = val(a[2]).load()
                elif op == 62:    # Store
                    val(a[0]).store(val(a[1]))
                elif op == 65:    # AccessChain
                    v[a[1]] = val(a[2]).at([val(i) for i in a[3:]])
                elif op == 59:    # Variable (Function storage)
                    v[a[1]] = Ptr([_copy(val(a[3])) if len(a) > 3 else self.m.default(types[a[0]][2])])
                elif op == 133:   # FMul
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), _mul)
                elif op == 129:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), _add)
                elif op == 131:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), _sub)
                elif op == 136:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), _div)
                elif op == 127:   # FNegate
                    v[a[1]] = _ew1(val(a[2]), lambda x: -x)
                elif op == 142:   # VectorTimesScalar
                    s = val(a[3])
                    v[a[1]] = [x * s for x in val(a[2])]
                elif op == 145:   # MatrixTimesVector: ((c0*x + c1*y) + c2*z) + c3*w
                    mtx, vec = val(a[2]), val(a[3])
                    out = []
                    for r in range(len(mtx[0])):
                        acc = mtx[0][r] * vec[0] + mtx[1][r] * vec[1]
                        for c in range(2, len(mtx)):
                            acc = acc + mtx[c][r] * vec[c]
                        out.append(acc)
                    v[a[1]] = out
                elif op == 146:   # MatrixTimesMatrix: column c of the result = M1 x (column c of M2), summed left to right
                    m1, m2 = val(a[2]), val(a[3])
                    res = []
                    for col in m2:
                        out = []
                        for r in range(len(m1[0])):
                            acc = m1[0][r] * col[0] + m1[1][r] * col[1]
                            for k in range(2, len(m1)):
                                acc = acc + m1[k][r] * col[k]
                            out.append(acc)
                        res.append(out)
                    v[a[1]] = res
                elif op == 148:
                    v[a[1]] = dot(val(a[2]), val(a[3]))
                elif op == 81:    # CompositeExtract
                    x = val(a[2])
                    for i in a[3:]:
                        x = x[i]
                    v[a[1]] = _copy(x)
                elif op == 80:    # CompositeConstruct
                    t = types[a[0]]
                    parts = [val(i) for i in a[2:]]
                    if t[0] == "vec":
                        flat = []
                        for p in parts:
                            flat.extend(p) if type(p) is list else flat.append(p)
                        v[a[1]] = flat
                    else:
                        v[a[1]] = [_copy(p) for p in parts]
                elif op == 79:    # VectorShuffle
                    both = list(val(a[2])) + list(val(a[3]))
                    v[a[1]] = [both[i] for i in a[4:]]
                elif op == 84:    # Transpose
                    mtx = val(a[2])
                    v[a[1]] = [[mtx[c][r] for c in range(len(mtx))] for r in range(len(mtx[0]))]
                elif op == 12:    # ExtInst
                    v[a[1]] = self._ext(a[3], [val(i) for i in a[4:]])
                elif op == 57:    # FunctionCall
                    v[a[1]] = self._call(a[2], [val(i) for i in a[3:]])
                elif op == 400:   # CopyLogical
                    v[a[1]] = _copy(val(a[2]))
                elif op == 68:    # ArrayLength
                    v[a[1]] = len(val(a[2]).load()[a[3]])
                # ---- integers / conversions ----
                elif op == 128:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: (x + y) & 0xFFFFFFFF)
                elif op == 130:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: (x - y) & 0xFFFFFFFF)
                elif op == 132:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: (x * y) & 0xFFFFFFFF)
                elif op == 112:   # ConvertUToF
                    v[a[1]] = _ew1(val(a[2]), lambda x: F32(x))
                elif op == 111:   # ConvertSToF
                    v[a[1]] = _ew1(val(a[2]), lambda x: F32(_signed(x)))
                elif op == 109:   # ConvertFToU (round toward zero)
                    v[a[1]] = _ew1(val(a[2]), lambda x: int(x) & 0xFFFFFFFF)
                elif op == 110:
                    v[a[1]] = _ew1(val(a[2]), lambda x: int(x) & 0xFFFFFFFF)
                elif op == 113:   # UConvert
                    v[a[1]] = val(a[2])
                elif op == 120:   # ConvertUToPtr: a device address becomes a pointer to the block registered there
                    v[a[1]] = Ptr(self.addresses[val(a[2])])
                elif op == 124:   # Bitcast
                    v[a[1]] = self._bitcast(a[0], val(a[2]))
                elif op == 194:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: (x & 0xFFFFFFFF) >> y)
                elif op == 196:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: (x << y) & 0xFFFFFFFF)
                elif op == 197:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x | y)
                elif op == 198:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x ^ y)
                elif op == 199:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x & y)
                # ---- comparisons / logic ----
                elif op == 184:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: bool(x < y))
                elif op == 186:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: bool(x > y))
                elif op == 188:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: bool(x <= y))
                elif op == 190:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: bool(x >= y))
                elif op == 180:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: bool(x == y))
                elif op in (182, 183):
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: bool(x != y))
                elif op == 170:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: (x & 0xFFFFFFFF) == (y & 0xFFFFFFFF))
                elif op == 171:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: (x & 0xFFFFFFFF) != (y & 0xFFFFFFFF))
                elif op == 176:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x < y)
                elif op == 174:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x >= y)
                elif op == 172:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x > y)
                elif op == 178:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x <= y)
                elif op == 177:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: _signed(x) < _signed(y))
                elif op == 179:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: _signed(x) <= _signed(y))
                elif op == 173:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: _signed(x) > _signed(y))
                elif op == 175:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: _signed(x) >= _signed(y))
                elif op == 166:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x or y)
                elif op == 167:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x and y)
                elif op == 164:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x == y)
                elif op == 165:
                    v[a[1]] = _ew2(val(a[2]), val(a[3]), lambda x, y: x != y)
                elif op == 168:
                    v[a[1]] = _ew1(val(a[2]), lambda x: not x)
                elif op == 169:   # Select
                    c, x, y = val(a[2]), val(a[3]), val(a[4])
                    v[a[1]] = [xx if cc else yy for cc, xx, yy in zip(c, x, y)] if type(c) is list else (x if c else y)
                elif op == 154:
                    v[a[1]] = any(val(a[2]))
                elif op == 155:
                    v[a[1]] = all(val(a[2]))
                elif op == 156:
                    v[a[1]] = _ew1(val(a[2]), lambda x: bool(np.isnan(x)))
                elif op == 157:
                    v[a[1]] = _ew1(val(a[2]), lambda x: bool(np.isinf(x)))
                # ---- control flow ----
                elif op == 245:   # Phi
                    for k in range(2, len(a), 2):
                        if a[k + 1] == prev:
                            v[a[1]] = val(a[k])
                            break
                    else:
                        raise RuntimeError("phi without a matching predecessor")
                elif op in (246, 247):
                    pass
                elif op == 249:
                    nxt = a[0]
                elif op == 250:
                    nxt = a[1] if val(a[0]) else a[2]
                elif op == 253:
                    self.m.stats["instructions"] += n_inst
                    return None
                elif op == 254:
                    self.m.stats["instructions"] += n_inst
                    return _copy(val(a[0]))
                # ---- environment ----
                elif op == 4445:  # TraceRayKHR
                    self.trace_ray(self, val(a[6]), val(a[7]), val(a[8]), val(a[9]), val(a[10]))
                elif op == 4473:  # RayQueryInitialize
                    val(a[0]).store(["rq", val(a[4]), val(a[5]), val(a[6]), val(a[7])])
                elif op == 4477:  # RayQueryProceed: opaque triangles only -> traversal completes, nothing to confirm
                    p = val(a[2])
                    q = p.load()
                    p.store(["done", self.ray_query(q[1], q[2], q[3], q[4])])
                    v[a[1]] = False
                elif op == 4479:  # committed intersection type: 0 none, 1 triangle
                    v[a[1]] = 0 if val(a[2]).load()[1] is None else 1
                elif op == 6019:  # committed instance custom index
                    c = val(a[2]).load()[1]
                    v[a[1]] = c if isinstance(c, CommittedInstances) else int(c)
                elif op == 98:    # ImageRead
                    c = val(a[3])
                    v[a[1]] = [F32(x) for x in self.images[val(a[2])][_signed(c[1]), _signed(c[0])]]
                elif op == 99:    # ImageWrite
                    c = val(a[1])
                    self.images[val(a[0])][_signed(c[1]), _signed(c[0])] = [F32(x) for x in val(a[2])]
                else:
                    raise NotImplementedError(f"{self.m.path}: opcode {op} {a}")
                if nxt is not None:
                    break
            if nxt is None:
                raise RuntimeError("block fell through without a terminator")
            prev, label = label, nxt

    def _bitcast(self, tid, x):
        t = self.m.types[tid]
        if t[0] == "float":
            return _f(x)
        if t[0] == "int":
            return _bits(x) if isinstance(x, np.floating) else (x & 0xFFFFFFFF)
        if t[0] == "vec":
            return [self._bitcast(t[1], c) for c in x]
        raise NotImplementedError(f"bitcast to {t}")

    def _ext(self, inst, o):
        if inst == 69:
            return normalize(o[0])
        if inst == 66:
            return length(o[0])
        if inst == 68:
            return cross(o[0], o[1])
        if inst == 40:
            return _ew2(o[0], o[1], gmax)
        if inst == 37:
            return _ew2(o[0], o[1], gmin)
        if inst == 43:  # clamp = min(max(x, lo), hi)
            return _ew2(_ew2(o[0], o[1], gmax), o[2], gmin)
        if inst == 46:  # mix = x * (1 - a) + y * a
            one = F32(1.0)
            f = lambda x, y, t: x * (one - t) + y * t  # noqa: E731
            return [f(x, y, t) for x, y, t in zip(*o)] if type(o[0]) is list else f(*o)
        if inst == 49:  # smoothstep
            e0, e1, x = o
            t = gmin(gmax((x - e0) / (e1 - e0), F32(0.0)), F32(1.0))
            return (t * t) * (F32(3.0) - F32(2.0) * t)
        if inst == 71:  # reflect(I, N) = I - N * (2 * dot(N, I))
            k = F32(2.0) * dot(o[1], o[0])
            return [i - n * k for i, n in zip(o[0], o[1])]
        if inst == 31:
            return _ew1(o[0], np.sqrt)
        if inst == 4:
            return _ew1(o[0], np.abs)
        if inst == 6:
            return _ew1(o[0], lambda x: F32(float(int(x > 0) - int(x < 0))))
        if inst == 26:  # pow: the shaders only raise to the fifth power (simple.rchit:97): ((x^2)^2) x
            if float(o[1]) != 5.0:
                raise NotImplementedError(f"pow(x, {o[1]})")
            x2 = o[0] * o[0]
            return (x2 * x2) * o[0]
        if inst == 13:
            return self.sincos(o[0])[0]
        if inst == 14:
            return self.sincos(o[0])[1]
        if inst == 38:
            return min(o[0], o[1])
        if inst == 39:
            return min(_signed(o[0]), _signed(o[1])) & 0xFFFFFFFF
        if inst == 34:
            return inverse3(o[0])
        raise NotImplementedError(f"GLSL.std.450 instruction {inst}")


def _mul(x, y):
    return x * y


def _add(x, y):
    return x + y


def _sub(x, y):
    return x - y


def _div(x, y):
    return x / y


def _ew1(x, f):
    return [f(c) for c in x] if type(x) is list else f(x)


def _ew2(x, y, f):
    return [f(a, b) for a, b in zip(x, y)] if type(x) is list else f(x, y)
