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


# dmdt (Lee::Kexp), alphaTransfer Su/Sp, solve() with calculateSuSp() in the correctors
MASS_CASE = ("mass_transfer_3d_n8", 5, 8)
MASS_RHO = [7000.0, 1.2, 6500.0]
MASS_PAIRS = [(0, 1), (0, 2), (1, 2)]
MASS_MODELS = [{"from": 0, "to": 2, "C": 0.5, "Tactivate": 300.05, "alphaMin": 0.0},
               {"from": 2, "to": 0, "C": -0.25, "Tactivate": 299.98, "alphaMin": 0.01},
               {"from": 0, "to": 1, "C": 3.0, "Tactivate": 300.02, "alphaMin": 0.0}]


def mass_inputs(config, n):
    spec = cases.case(config, n)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    # deterministic non-solenoidal perturbation of phi so that div(phi) != 0
    k = np.arange(m.nIF, dtype=np.float64)
    f["phi_i"] = f["phi_i"] + 1e-3 * np.abs(f["phi_i"]).max() * np.sin(0.37 * k + 0.11 * k * k / max(m.nIF, 1))
    return spec, m, f


def build_mass(config, n):
    spec, m, f = mass_inputs(config, n)
    mass = orc.MassHandle(MASS_MODELS, MASS_PAIRS, MASS_RHO)
    dm = orc.correct_mass_sources(m, mass, f["alpha_i"], f["T_i"])
    Su, Sp, mx = orc.alpha_transfer(m, mass, f["alpha_i"], dm, f["phi_i"], f["phi_b"])
    so = orc.solve_mt(m, orc.mules_cfg(3), orc.alpha_ctl(2, 1, "vanLeer", "linear", True, 1e-13, 300), f["alpha_i"],
                      f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], MASS_RHO, mass, dm)
    out = dict(ddtAlphaMax=np.array([mx]))
    for i in range(len(MASS_MODELS)):
        out[f"dmdt{i}"] = dm[i]
    for p in range(3):
        out[f"Su{p}"] = Su[p]; out[f"Sp{p}"] = Sp[p]; out[f"solve_alpha_i{p}"] = so["alpha_i"][p]
    return out


# laserDTRM::calculate() - reflecting cells, nHat, Q of a Gaussian beam onto the first blob
LASER_CASE = ("laser_dtrm_3d_n10", 5, 10)
LASER_MODELS = [{"kind": "FresnelLaser", "from": 2, "to": 0, "epsilon": 0.25},
                {"kind": "Fresnel", "from": 2, "to": 1, "nk1": (1.0, 0.0), "nk2": (3.6, 5.0)}]
LASER_ABS = dict(a=(30.0, 12.0, 0.2), e=(0.3, 0.1, 0.0), E=(0.0, 0.0, 0.0))


def laser_inputs(config, n):
    from thermocapillaryinterfoam_b200 import mesh as hm
    spec = cases.case(config, n)
    m = global_mesh(spec)
    v = tuple(hm.vertices(spec.n[d], spec.lengths[d]) for d in range(3))
    geom = cases.box_polymesh(m, *v)
    f = cases.synthetic_fields(spec, m)
    d = np.array([0.2, -1.0, 0.1]); d /= np.linalg.norm(d)
    r = np.cross(d, [0.0, 0.0, 1.0]); r /= np.linalg.norm(r)
    p0, p1, Ip, dA = cases.laser_seeds((0.45, 0.93, 0.5), d, r, 6, 12, 0.2, geom["boundsMag"], mode="Gaussian", power=150.0,
                                       sigma=0.09)
    seeds = dict(p0=p0, p1=p1, I=Ip, dA=dA, cell=cases.find_cells_box(spec.n, *v, p0))
    return spec, m, geom, f, seeds


def build_laser(config, n):
    spec, m, geom, f, s = laser_inputs(config, n)
    al = f["alpha_i"]
    a, e, E = (orc.local_density(al, LASER_ABS[k]) for k in ("a", "e", "E"))
    r = orc.laser_calculate(m, geom, LASER_MODELS, al, f["alpha_b"], f["alphaBc"], a, e, E, f["T_i"], s["p0"], s["p1"], s["I"],
                            s["dA"], s["cell"])
    return dict(Q=r["Q"], reflectingCells=r["reflectingCells"], nHat=r["nHat"], nHatPoints=r["nHatPoints"],
                nParticles=np.array([r["nParticles"]]))


if __name__ == "__main__":
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", LASER_CASE[0] + ".npz"), **build_laser(*LASER_CASE[1:]))
    print("wrote", LASER_CASE[0])
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", MASS_CASE[0] + ".npz"), **build_mass(*MASS_CASE[1:]))
    print("wrote", MASS_CASE[0])
    for name, config, n in CASES:
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), **build(config, n))
        print("wrote", name)
