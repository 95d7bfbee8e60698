"""The compiled stand-in for foam_shim/ (tests/harness/shim_harness.cpp): the C ABI driven from C++ with
OpenFOAM-shaped containers (AoS vectors, per-patch boundary fields flattened with `empty` patches skipped, stamped
transfers), one time step of the reference's call sequence, checked against the oracle.

CPU: the harness must compile warning-free against include/*.h (C++17 -Wall -Wextra -Werror), the headers must be
plain C99, and without a GPU the program must refuse to run (exit 77) instead of falling back.
GPU: the replay must pass for the 2-D case with `empty` patches (config 1) and the 3-D 3-phase case (config 5)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "harness", "shim_harness.cpp")
EXE = os.path.join(ROOT, "tests", "harness", "_build", "shim_harness")
PKG = os.path.join(ROOT, "thermocapillaryinterfoam_b200")
ORC = os.path.join(ROOT, "oracle")


def _build():
    from thermocapillaryinterfoam_b200 import abi, mesh as hm
    from oracle import oracle as orc
    abi.lib(); hm.lib(); orc.lib()                      # the three libraries the harness links (built by __graft_entry__.build())
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    deps = [SRC, os.path.join(ROOT, "include", "tcif.h"), os.path.join(ROOT, "include", "tcif_mesh.h"),
            os.path.join(ORC, "tcif_oracle.h")]
    if os.path.exists(EXE) and all(os.path.getmtime(d) <= os.path.getmtime(EXE) for d in deps):
        return
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           "-I" + ORC, SRC, "-o", EXE, "-L" + PKG, "-L" + ORC, "-ltcif", "-ltcif_host", "-ltcif_oracle",
                           "-Wl,-rpath," + PKG, "-Wl,-rpath," + ORC])


def test_headers_are_plain_c99():
    for h in ("tcif.h", "tcif_mesh.h"):
        subprocess.check_call(["gcc", "-std=c99", "-fsyntax-only", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                               "-x", "c", os.path.join(ROOT, "include", h)])


def test_harness_compiles_and_refuses_to_run_without_a_gpu():
    _build()
    from thermocapillaryinterfoam_b200 import abi
    if abi.device_count() > 0:
        pytest.skip("a GPU is visible: the replay itself is the -m gpu test")
    r = subprocess.run([EXE, "1", "8"], capture_output=True, text=True)
    assert r.returncode == 77 and "no CUDA device" in r.stdout, (r.returncode, r.stdout, r.stderr)


@pytest.mark.gpu
@pytest.mark.parametrize("config,n", [(1, 32), (5, 16)])
def test_harness_replays_a_time_step(config, n):
    _build()
    r = subprocess.run([EXE, str(config), str(n)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "harness: OK" in r.stdout, (r.returncode, r.stdout[-3000:], r.stderr[-3000:])
