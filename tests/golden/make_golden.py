#!/usr/bin/env python
"""Regenerates tests/golden/*.npz from the CPU oracle (run from the repo root: python tests/golden/make_golden.py).

The reference ships no golden vectors for this path and cannot be run here (OpenFOAM v2106 is absent), so these
fixtures pin the ORACLE's current output, not the reference's: they guard against silent drift of the oracle
and give the GPU tests a target that does not depend on the oracle being rebuilt identically."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import oracle as orc          # noqa: E402
from thermocapillaryinterfoam_b200 import cases  # noqa: E402
from helpers import global_mesh           # noqa: E402

CASES = [("static_droplet_2d_n12", 1, 12), ("thermocapillary_2d_n10", 2, 10), ("marangoni_3d_n6", 3, 6),
         ("three_phase_3d_n6", 5, 6)]
RHO = np.array([1000.0, 800.0, 1.2])


def build(config, n):
    spec = cases.case(config, n)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    P = spec.nPhases
    st = orc.stage(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                   f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    so = orc.solve(m, orc.mules_cfg(3), orc.alpha_ctl(2, 1, "vanLeer", "linear", True, 1e-13, 300), f["alpha_i"],
                   f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], RHO[:P])
    out = dict(stage_F_i=st["F_i"], stage_F_b=st["F_b"], solve_rhoPhi_i=so["rhoPhi_i"])
    for p in range(P):
        out[f"stage_alpha_i{p}"] = st["alpha_i"][p]
        out[f"stage_alphaPhi_i{p}"] = st["alphaPhi_i"][p]
        out[f"solve_alpha_i{p}"] = so["alpha_i"][p]
        out[f"solve_alphaPhi_i{p}"] = so["alphaPhi_i"][p]
    return out


if __name__ == "__main__":
    for name, config, n in CASES:
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), **build(config, n))
        print("wrote", name)
