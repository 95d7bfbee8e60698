"""The C-ABI boundary without a GPU: libtcif.so / libtcif_host.so load, export every symbol the
headers declare, fail loudly without a device, and the product never touches oracle/."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from thermocapillaryinterfoam_b200 import abi, mesh as hm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"TCIF_API[^;(]*?\b(tcif_\w+)\s*\(", src)))


def test_headers_declare_something():
    assert len(_declared("tcif.h")) >= 35
    assert len(_declared("tcif_mesh.h")) >= 8


def test_libtcif_exports_every_declared_symbol():
    L = abi.lib()
    missing = [s for s in _declared("tcif.h") if not hasattr(L, s)]
    assert not missing, missing


def test_libtcif_host_exports_every_declared_symbol():
    L = hm.lib()
    missing = [s for s in _declared("tcif_mesh.h") if not hasattr(L, s)]
    assert not missing, missing


def test_only_declared_symbols_are_exported():
    out = subprocess.check_output(["nm", "-D", "--defined-only", abi.LIB_PATH], text=True)
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    exported = {s for s in exported if not s.startswith("_")}
    assert exported == set(_declared("tcif.h")), exported ^ set(_declared("tcif.h"))


def test_library_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "--list-elf", abi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_version_and_timer_names():
    assert "sm_100a" in abi.version()
    n = abi.lib().tcif_timer_count()
    names = [abi.lib().tcif_timer_name(i).decode() for i in range(n)]
    assert len(set(names)) == n and "mules_iter" in names and "div_tensor" in names


def test_no_device_is_a_loud_error_not_a_fallback():
    if abi.device_count() > 0:
        pytest.skip("a GPU is visible")
    m = hm.box_block((4, 4, 4))
    with pytest.raises(abi.TcifError, match="no CUDA device"):
        abi.Context(m, 2)


def test_null_arguments_are_errors():
    L = abi.lib()
    h = C.c_void_p()
    assert L.tcif_create(None, None, C.byref(h)) != 0
    assert b"null" in L.tcif_last_error(None)
    assert L.tcif_stage(None, C.c_double(1.0)) != 0
    assert L.tcif_launch_count(None) == 0


def test_product_never_references_the_oracle():
    pkg = os.path.join(ROOT, "thermocapillaryinterfoam_b200")
    bad = []
    for d, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(d, fn), errors="ignore").read()
                if re.search(r"^\s*(from|import)\s+oracle\b|#include[^\n]*oracle|libtcif_oracle|\borc_\w+\(", txt, flags=re.M):
                    bad.append(os.path.join(d, fn))
    assert not bad, bad
    for h in ("tcif.h", "tcif_mesh.h"):
        assert "orc_" not in open(os.path.join(ROOT, "include", h)).read()
    # and the shared objects do not link it
    for so in (abi.LIB_PATH, hm.LIB_PATH):
        ldd = subprocess.run(["ldd", so], capture_output=True, text=True).stdout
        assert "oracle" not in ldd


def test_host_mesh_desc_is_abi_compatible():
    """tcif_mesh_get_desc fills exactly the struct tcif_create consumes."""
    m = hm.box_block((3, 2, 2))
    d = m.desc
    assert (d.nCells, d.nInternalFaces, d.nPatches, d.nBoundaryFaces) == (12, 20, 6, 32)
    assert np.array_equal(m.owner < m.neighbour, np.ones(20, bool))
    assert C.sizeof(abi.MeshDesc) == 4 * 4 + 14 * C.sizeof(C.c_void_p)
