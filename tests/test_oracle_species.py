"""Known-answer tests of the ORACLE's solveYi restatement (MultiComponentPhaseModel.C:182-388).  CPU only."""
import numpy as np
import pytest

from oracle import oracle as orc
from thermocapillaryinterfoam_b200 import cases
from helpers import global_mesh


def _case(n=10, config=5):
    spec = cases.case(config, n)
    m = global_mesh(spec)
    return spec, m, cases.synthetic_fields(spec, m)


def _uniform_alphas(m, f):
    """phase fractions without an interface: grad(alpha) = 0, so nHatf = 0 and phir = 0 (no compression flux)"""
    vals = (0.3, 0.3, 0.4)
    return [np.full(m.nCells, v) for v in vals], [np.full(m.nBF, v) for v in vals]


def test_uniform_species_stay_uniform():
    # div-free phi, closed walls, no sources, no interface: a uniform mass fraction is a fixed point.  (WITH an
    # interface it is not: the compression term sum_q flux(-flux(-phir, alpha_q), Yi) = Yi*phir*(1 - alpha1)_f has a
    # divergence -- the reference compresses the species with the phase, :256-276.)
    spec, m, f = _case()
    f["alpha_i"], f["alpha_b"] = _uniform_alphas(m, f)
    Y = [np.full(m.nCells, 0.3), np.full(m.nCells, 0.45), np.full(m.nCells, 0.25)]
    Yb = [np.full(m.nBF, 0.3), np.full(m.nBF, 0.45), np.full(m.nBF, 0.25)]
    yi, yb = orc.solve_yi(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], 0, Y, Yb,
                          np.zeros((3, m.nPatches), np.int32), 2, f["phi_i"], f["phi_b"], f["deltaT"])
    for i in range(3):
        assert np.abs(yi[i] - Y[i]).max() < 1e-12


@pytest.mark.parametrize("sphi,sphir", [("vanLeer", "linear"), ("upwind", "upwind"), ("linear", "vanLeer")])
def test_bounds_closure_and_conservation(sphi, sphir):
    spec, m, f = _case(12)
    X = np.asarray(m.C)
    y0 = 0.5 * (1 + np.tanh((X[:, 0] - 0.5) / 0.08)) * 0.6          # a front, 0 .. 0.6
    y1 = 0.3 * (1 - y0 / 0.6)
    Y = [y0, y1, 1 - y0 - y1]
    fc = np.asarray(m.faceCells)
    Yb = [y[fc].copy() for y in Y]
    yi, yb = orc.solve_yi(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], 1, Y, Yb,
                          np.zeros((3, m.nPatches), np.int32), 2, f["phi_i"], f["phi_b"], f["deltaT"],
                          schemePhi=sphi, schemePhir=sphir)
    for i in range(3):
        assert yi[i].min() >= -1e-12 and yi[i].max() <= 1 + 1e-12
    assert np.abs(yi[0] + yi[1] + yi[2] - 1).max() < 1e-12           # inert species closes the sum
    V = np.asarray(m.V)
    for i in range(2):                                                # phi_b = 0, Su = Sp = 0: mass of the solved species
        assert abs((V * yi[i]).sum() - (V * Y[i]).sum()) < 1e-13 * V.sum()
    assert np.abs(yi[0] - Y[0]).max() > 1e-5


def test_sources_act_with_the_right_sign():
    spec, m, f = _case(8)
    f["alpha_i"], f["alpha_b"] = _uniform_alphas(m, f)
    Y = [np.full(m.nCells, 0.2), np.full(m.nCells, 0.8)]
    Yb = [np.full(m.nBF, 0.2), np.full(m.nBF, 0.8)]
    Su = [np.full(m.nCells, 0.3), None]
    yi, _ = orc.solve_yi(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], 0, Y, Yb,
                         np.zeros((2, m.nPatches), np.int32), 1, f["phi_i"], f["phi_b"], f["deltaT"], Su=Su,
                         schemePhi="upwind", schemePhir="upwind")
    # dY/dt = Su (uniform field, div-free flow): Y_new = Y + dt*Su
    assert np.abs(yi[0] - (0.2 + 0.3 * f["deltaT"])).max() < 1e-12
    assert np.abs(yi[1] - (1 - yi[0])).max() < 1e-15


@pytest.mark.parametrize("nSub", [2, 4])
def test_sub_cycles_keep_bounds_closure_and_mass(nSub):
    """nYiSubCycles > 1 (MultiComponentPhaseModel.C:345-364): n explicit sub-steps of deltaT/n with the same flux."""
    spec, m, f = _case(12)
    X = np.asarray(m.C)
    y0 = 0.5 * (1 + np.tanh((X[:, 0] - 0.5) / 0.08)) * 0.6
    y1 = 0.3 * (1 - y0 / 0.6)
    Y = [y0, y1, 1 - y0 - y1]
    fc = np.asarray(m.faceCells)
    Yb = [y[fc].copy() for y in Y]
    args = (m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], 1, Y, Yb, np.zeros((3, m.nPatches), np.int32), 2,
            f["phi_i"], f["phi_b"], f["deltaT"])
    yi, _ = orc.solve_yi(*args, nYiSubCycles=nSub)
    y1c, _ = orc.solve_yi(*args)
    for i in range(3):
        assert yi[i].min() >= -1e-12 and yi[i].max() <= 1 + 1e-12
    assert np.abs(yi[0] + yi[1] + yi[2] - 1).max() < 1e-12
    assert max(np.abs(yi[i] - y1c[i]).max() for i in range(2)) > 1e-9     # not the single-step result
    V = np.asarray(m.V)
    for i in range(2):                                                    # closed box, no sources: mass is kept
        assert abs((V * yi[i]).sum() - (V * Y[i]).sum()) < 1e-13 * V.sum()


def test_sub_cycles_with_uniform_source_integrate_it_over_the_whole_step():
    """Uniform Y, no interface, Su = const, Sp = 0: the first sub-cycle starts from the real old-time field (subCycle's
    constructor protects it), every later one from the current field, each adds Su*deltaT/n: Y_end = Y + Su*deltaT,
    the single-step answer."""
    spec, m, f = _case(8)
    f["alpha_i"], f["alpha_b"] = _uniform_alphas(m, f)
    Y = [np.full(m.nCells, 0.2), np.full(m.nCells, 0.8)]
    Yb = [np.full(m.nBF, 0.2), np.full(m.nBF, 0.8)]
    su = 0.3
    Su = [np.full(m.nCells, su), None]
    dt = f["deltaT"]
    args = (m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], 0, Y, Yb, np.zeros((2, m.nPatches), np.int32), 1,
            0.0 * f["phi_i"], 0.0 * f["phi_b"], dt)          # fluid at rest: only the source acts
    y1, _ = orc.solve_yi(*args, Su=Su)
    assert np.allclose(y1[0], 0.2 + su * dt, rtol=0, atol=1e-13)
    y3, _ = orc.solve_yi(*args, Su=Su, nYiSubCycles=3)
    assert np.allclose(y3[0], 0.2 + su * dt, rtol=0, atol=1e-13)
