"""CPU suite, part 3: the C-ABI library loads, exports every symbol include/rt_api.h declares, and refuses to
run without a B200 (no CPU fallback).  No compute call is made here."""
import ctypes as C
import os
import re
import subprocess

import pytest

from conftest import ROOT

from raytracer_b200 import librt


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(librt.LIB_PATH):
        import __graft_entry__ as g

        g.build()
    return librt.load()


def _declared():
    hdr = open(os.path.join(ROOT, "include", "rt_api.h")).read()
    return sorted(set(re.findall(r"RT_API\s+[\w\s\*]+?\b(rt_\w+)\s*\(", hdr)))


def test_header_and_binding_agree():
    assert _declared() == sorted(librt.EXPORTS)


def test_every_declared_symbol_is_exported(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", librt.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\sT\s+(rt_\w+)", out))
    assert set(_declared()) <= exported
    for n in _declared():
        assert hasattr(lib, n)
    # nothing but the C ABI leaks out of the library (-fvisibility=hidden)
    leaked = [s for s in re.findall(r"\sT\s+(\S+)", out) if not s.startswith("rt_")]
    assert leaked == []


def test_library_is_sm100a_only(lib):
    out = subprocess.run(["cuobjdump", "-lelf", librt.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_struct_sizes_match_the_header(lib, tmp_path):
    """The ctypes mirrors against the header itself: a C program that includes include/rt_api.h prints the sizes (and the
    offsets of the fields added for the multi-GPU fan-out) as gcc lays them out."""
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "rt_api.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(rt_config),sizeof(rt_object),sizeof(rt_mesh),sizeof(rt_counters),sizeof(rt_scene_info),sizeof(rt_restir_params),"
                   "sizeof(rt_material),sizeof(rt_vertex),sizeof(rt_reservoir),offsetof(rt_config,n_devices),offsetof(rt_config,strip_balance));return 0;}\n")
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(librt.RtConfig), C.sizeof(librt.RtObject), C.sizeof(librt.RtMesh), C.sizeof(librt.RtCounters),
            C.sizeof(librt.RtSceneInfo), C.sizeof(librt.RtRestirParams), 44, 24, librt.RESERVOIR_DTYPE.itemsize,
            librt.RtConfig.n_devices.offset, librt.RtConfig.strip_balance.offset]
    assert got == want
    assert C.sizeof(librt.RtConfig) == 96 and C.sizeof(librt.RtObject) == 72 and C.sizeof(librt.RtMesh) == 32
    assert C.sizeof(librt.RtCounters) == 80 and C.sizeof(librt.RtSceneInfo) == 32 and C.sizeof(librt.RtRestirParams) == 32


def test_no_cpu_fallback(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the failure path is for GPU-less hosts")
    cfg = librt.RtConfig()
    lib.rt_default_config(C.byref(cfg))
    assert (cfg.max_depth, cfg.rr_start_depth, cfg.radiance_clamp) == (3, 3, 5.0)  # simple.rgen:23,56,121
    h = C.c_void_p()
    rc = lib.rt_create(C.byref(cfg), C.byref(h))
    assert rc == -2 and not h.value  # RT_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.rt_last_error(None)
    with pytest.raises(librt.RtError):
        librt.Context()


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "raytracer_b200")
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|#include\s+\"[^\"]*oracle|liboracle", re.M)
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                assert not pat.search(open(os.path.join(dirpath, f), errors="ignore").read()), f


def _build_c_host(tmp_path):
    exe = str(tmp_path / "host_three_calls")
    csrc = os.path.join(ROOT, "raytracer_b200", "csrc")
    subprocess.run(["gcc", "-std=c99", "-O2", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "examples", "host_three_calls.c"), "-L", csrc, "-lrt", f"-Wl,-rpath,{csrc}", "-o", exe], check=True)
    return exe


def test_a_plain_c_host_compiles_and_links_against_the_abi(lib, tmp_path):
    """examples/host_three_calls.c: the reference renderer's three calls from C, GPUs listed in rt_config — no Python, no torch.
    Here it must compile with -Wall -Werror against include/rt_api.h and link against librt.so; tests/test_gpu_multi.py runs it."""
    exe = _build_c_host(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr
