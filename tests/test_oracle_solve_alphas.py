"""Known answers for the oracle's restatement of the rest of multiphaseSystem::solveAlphas()
(flux assembly, nHatf, implicit upwind predictor, rhoPhi / renormalisation, sub-cycling)."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spl

from oracle import oracle as orc, mesh_oracle as mo
from thermocapillaryinterfoam_b200 import cases
from helpers import global_mesh

ZG, FV = orc.BC_ZERO_GRADIENT, orc.BC_FIXED_VALUE


def _flux(m, umag=0.1):
    def u(X):
        v = np.zeros_like(X)
        v[:, 0] = umag * np.sin(np.pi * X[:, 0]) * np.cos(np.pi * X[:, 1])
        v[:, 1] = -umag * np.cos(np.pi * X[:, 0]) * np.sin(np.pi * X[:, 1])
        return v
    return np.einsum("ij,ij->i", m.Sf, u(m.Cf)), np.zeros(m.nBF)


# ---- fvc::flux with upwind / linear / vanLeer ----------------------------------------------------------
def test_conv_flux_upwind_and_linear():
    m = mo.box_mesh(7, 6, 5, (1.0, 1.0, 1.0), (0,) * 6, (2.0, None, 0.5))
    rng = np.random.default_rng(1)
    a = rng.random(m.nCells); ab = a[m.faceCells]
    phii, phib = _flux(m)
    bc = np.zeros(m.nPatches, np.int32)
    ui, ub = orc.conv_flux(m, "upwind", phii, phib, a, ab, bc)
    ri, rb = orc.upwind_flux(m, phii, phib, a, ab)
    assert np.array_equal(ui, ri) and np.array_equal(ub, rb)
    li, lb = orc.conv_flux(m, "linear", phii, phib, a, ab, bc)
    si, sb = orc.interpolate(m, a, ab)
    assert np.array_equal(li, phii * si) and np.array_equal(lb, phib * sb)


def test_conv_flux_vanleer_limits():
    """Linear field: r == 1 -> limiter 1 -> central (linear) weights.  Step field: face values stay between
    the two cell values (TVD) and reduce to upwind next to the local extremum."""
    m = mo.box_mesh(16, 4, 3)
    bc = np.zeros(m.nPatches, np.int32)
    phii = np.einsum("ij,j->i", m.Sf, np.array([0.3, 0.1, -0.2])); phib = np.zeros(m.nBF)
    a = 0.1 + 0.5 * m.C[:, 0] + 0.2 * m.C[:, 1]
    ab = 0.1 + 0.5 * m.Cfb[:, 0] + 0.2 * m.Cfb[:, 1]
    bcf = np.full(m.nPatches, FV, np.int32)
    vi, _ = orc.conv_flux(m, "vanLeer", phii, phib, a, ab, bcf)
    li, _ = orc.conv_flux(m, "linear", phii, phib, a, ab, bcf)
    assert np.allclose(vi, li, rtol=0, atol=1e-15)
    a = (m.C[:, 0] < 0.5).astype(float); ab = a[m.faceCells]
    vi, _ = orc.conv_flux(m, "vanLeer", phii, phib, a, ab, bc)
    ui, _ = orc.conv_flux(m, "upwind", phii, phib, a, ab, bc)
    nz = phii != 0
    af = vi[nz] / phii[nz]
    lo = np.minimum(a[m.owner], a[m.neighbour])[nz]; hi = np.maximum(a[m.owner], a[m.neighbour])[nz]
    assert np.all(af >= lo - 1e-15) and np.all(af <= hi + 1e-15)
    assert np.array_equal(vi, ui)        # a pure step has zero upwind-side gradient ratio everywhere it matters


# ---- nHatf ------------------------------------------------------------------------------------------------
def test_nhatf_planar_interface():
    m = mo.box_mesh(6, 20, 4)
    a1 = 0.5 * (1.0 - np.tanh((m.C[:, 1] - 0.5) / 0.1)); a2 = 1.0 - a1
    bc = np.zeros(m.nPatches, np.int32)
    ni, nb = orc.nHatf(m, a1, a1[m.faceCells], bc, a2, a2[m.faceCells], bc)
    # interface normal = grad(a1)/|grad a1| = -e_y: nHatf = -Sf_y (to within deltaN/|g|) on interior faces
    inner = np.ones(m.nCells, bool); inner[m.faceCells] = False
    f = inner[m.owner] & inner[m.neighbour]
    g, _ = orc.grad(m, a1, a1[m.faceCells], bc)
    gmin = np.abs(g[inner, 1]).min()
    assert np.allclose(ni[f], -m.Sf[f, 1], rtol=0, atol=2 * orc.deltaN(m) / gmin * np.abs(m.Sf).max() + 1e-15)


# ---- flux assembly ----------------------------------------------------------------------------------------
def test_flux_assembly_two_phase_antisymmetry_and_walls():
    spec = cases.case(3, 12)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    oi, ob = orc.flux_assembly(m, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], "linear", "linear")
    # two phases with a1 + a2 = 1 and linear schemes: the two HO fluxes sum to phi, the compression parts cancel
    assert np.allclose(oi[0] + oi[1], f["phi_i"], rtol=0, atol=1e-15 * np.abs(f["phi_i"]).max() * 10)
    assert np.array_equal(ob[0], f["phi_b"] * f["alpha_b"][0])          # walls: phi_b * alpha_b (here 0)
    # compression sharpens: its flux is phic * nHatf * a1f * a2f
    base = f["phi_i"] * orc.interpolate(m, f["alpha_i"][0], f["alpha_b"][0])[0]
    comp = oi[0] - base
    nh, _ = orc.nHatf(m, f["alpha_i"][0], f["alpha_b"][0], f["alphaBc"][0], f["alpha_i"][1], f["alpha_b"][1], f["alphaBc"][1])
    a1f = orc.interpolate(m, f["alpha_i"][0], f["alpha_b"][0])[0]; a2f = orc.interpolate(m, f["alpha_i"][1], f["alpha_b"][1])[0]
    phic = np.abs(f["phi_i"] / np.linalg.norm(m.Sf, axis=1))
    assert np.allclose(comp, phic * nh * a2f * a1f, rtol=1e-12, atol=1e-18)


def test_calpha_zero_removes_compression():
    spec = cases.case(5, 8)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    oi, _ = orc.flux_assembly(m, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], "upwind", "linear",
                              cAlpha=np.zeros((3, 3)))
    for p in range(3):
        ui, _ = orc.upwind_flux(m, f["phi_i"], f["phi_b"], f["alpha_i"][p], f["alpha_b"][p])
        assert np.allclose(oi[p], ui, rtol=0, atol=1e-300)


# ---- implicit upwind predictor: solves the matrix the reference assembles ---------------------------------
@pytest.mark.parametrize("fixed", [False, True])
def test_upwind_predictor_solves_the_linear_system(fixed):
    m = mo.box_mesh(9, 8, 1, (1.0, 1.0, 0.1), (0, 0, 0, 0, 1, 1))
    rng = np.random.default_rng(2)
    psi0 = rng.random(m.nCells)
    # through-flow so that boundary coefficients matter: inflow at xMin, outflow at xMax
    u = np.array([0.2, 0.05, 0.0])
    phii = m.Sf @ u; phib = m.Sfb @ u
    bc = np.zeros(m.nPatches, np.int32)
    psib = psi0[m.faceCells].copy()
    if fixed:
        bc[0] = FV; psib[: m.patchSize[0]] = 0.7
    dt = 0.4 * (1.0 / 9) / 0.2
    Su = 0.1 * rng.random(m.nCells); Sp = -0.2 * rng.random(m.nCells)
    qi, qb, pi, pb, n = orc.upwind_predictor(m, dt, psi0, psib, psi0, bc, phii, phib, np.zeros(m.nIF), np.zeros(m.nBF),
                                              Su=Su, Sp=Sp, tol=1e-14, maxIter=500)
    assert 1 < n < 500
    # independent assembly (scipy): ddt + upwind fvmDiv == Su + fvm::Sp
    o, nb = m.owner.astype(int), m.neighbour.astype(int)
    A = sp.lil_matrix((m.nCells, m.nCells)); b = m.V * psi0 / dt + m.V * Su
    A.setdiag(m.V / dt - m.V * Sp)
    for f_ in range(m.nIF):
        ph = phii[f_]
        A[o[f_], o[f_]] += max(ph, 0.0); A[o[f_], nb[f_]] += min(ph, 0.0)
        A[nb[f_], nb[f_]] += max(-ph, 0.0); A[nb[f_], o[f_]] += -max(ph, 0.0)
    for p in range(m.nPatches):
        for i in range(m.patchSize[p]):
            bf = m.patchStart[p] + i; c = m.faceCells[bf]
            if bc[p] == FV:
                b[c] -= phib[bf] * psib[bf]
            else:
                A[c, c] += phib[bf]
    x = spl.spsolve(A.tocsr(), b)
    assert np.allclose(qi, x, rtol=0, atol=1e-12)
    # alpha1Eqn.flux() = phi * psi_upwind(new)
    ui, ub = orc.upwind_flux(m, phii, phib, qi, qb)
    assert np.array_equal(pi, ui) and np.array_equal(pb, ub)
    if not fixed:
        assert np.array_equal(qb, qi[m.faceCells])


# ---- solve(): conservation, renormalisation, rhoPhi, sub-cycles ------------------------------------------
@pytest.mark.parametrize("implicit", [True, False])
def test_solve_invariants(implicit):
    spec = cases.case(5, 12)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    rho = np.array([1000.0, 800.0, 1.2])
    ctl = orc.alpha_ctl(nAlphaCorr=2, schemePhi="vanLeer", schemePhir="linear", implicitPredictor=implicit)
    r = orc.solve(m, orc.mules_cfg(3), ctl, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"],
                  f["deltaT"], rho)
    s = sum(r["alpha_i"])
    assert np.abs(s - 1.0).max() < 1e-6 and np.abs(s - 1.0).max() > 0         # one renormalisation step, not exact
    for p in range(3):
        assert r["alpha_i"][p].min() > -1e-9 and r["alpha_i"][p].max() < 1 + 1e-9
        assert np.abs(r["alpha_i"][p] - f["alpha_i"][p]).max() > 1e-6
    # rhoPhi = sum_p rho_p * alphaPhi_p for uniform densities
    assert np.allclose(r["rhoPhi_i"], sum(rho[p] * r["alphaPhi_i"][p] for p in range(3)), rtol=1e-14, atol=0)
    # closed box: total volume of every phase changes only through the renormalisation (tiny)
    for p in range(3):
        v0, v1 = (m.V * f["alpha_i"][p]).sum(), (m.V * r["alpha_i"][p]).sum()
        assert abs(v1 - v0) < 1e-6 * max(v0, 1e-3)


def test_solve_subcycles_accumulate_rhophi():
    spec = cases.case(3, 10)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    rho = np.array([1000.0, 1.0])
    cfg = orc.mules_cfg(3)
    one = orc.solve(m, cfg, orc.alpha_ctl(1, 1), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"],
                    0.5 * f["deltaT"], rho)
    two_manual = orc.solve(m, cfg, orc.alpha_ctl(1, 1), one["alpha_i"], one["alpha_b"], f["alphaBc"], f["phi_i"],
                           f["phi_b"], 0.5 * f["deltaT"], rho)
    sub = orc.solve(m, cfg, orc.alpha_ctl(1, 2), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"],
                    f["deltaT"], rho)
    for p in range(2):
        assert np.array_equal(sub["alpha_i"][p], two_manual["alpha_i"][p])
    assert np.allclose(sub["rhoPhi_i"], 0.5 * one["rhoPhi_i"] + 0.5 * two_manual["rhoPhi_i"], rtol=1e-14, atol=1e-300)
