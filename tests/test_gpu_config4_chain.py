"""BASELINE.json configs[3] (3-D laser melt pool, 3 phases, Lee melting, laserDTRM heat input): the device pieces
chained as the solver's time step chains them, against the same chain in the oracle -- each piece has its own
parity test, this one checks that they work on each other's OUTPUT (device-resident alpha, Su/Sp, force, Q).

    fluid.correctMassSources(T)                      thermocapillaryInterFoam.C:104   MassTransferPhaseSystem.C:447-497
    fluid.solve()  with calculateSuSp() in the       :105                             multiphaseSystem.C:128-429
                   correctors
    fluid.divCapillaryStress()   (UEqn.H:29)         on the NEW alpha                 phaseSystem.C:1038-1105
    radiation->correct() -> laserDTRM::calculate()   (TEqn.H) on the NEW alpha        laserDTRM.C:542-735

Tolerances: alpha after solve() 1e-11 (the Jacobi predictor may stop one sweep apart), the force of that alpha 1e-9 of
max|F| (a gradient of a 1e-11 difference over a cell), Q 1e-9 of max|Q| with identical particle counts and reflecting
cells.
"""
import numpy as np
import pytest

from oracle import oracle as orc, mesh_oracle as mo
from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm
from helpers import assert_field_close

pytestmark = pytest.mark.gpu

# synthetic 3-phase box (cases.synthetic_fields): phases 0 and 1 are the two blobs (metal, melt), phase 2 the gas around
RHO = [7000.0, 6500.0, 1.2]
PAIRS = [(0, 1), (0, 2), (1, 2)]
# (C small enough that the blobs survive the two steps: with C = 0.5 the first solve() melts phase 0 away and no ray is
# reflected any more)
LEE = [dict(**{"from": 0, "to": 2}, C=0.002, Tactivate=300.05, alphaMin=0.0),
       dict(**{"from": 2, "to": 0}, C=-0.001, Tactivate=299.98, alphaMin=0.01)]
REFL = [dict(kind="FresnelLaser", **{"from": 2, "to": 0}, epsilon=0.25),          # gas -> metal
        dict(kind="Fresnel", **{"from": 2, "to": 1}, nk1=(1.0, 0.0), nk2=(3.6, 5.0))]     # gas -> melt
ABS = ((60.0, 25.0, 0.3), (0.2, 0.1, 0.0), (0.0, 0.0, 0.0))


@pytest.mark.parametrize("n,nSteps", [(16, 2)])
def test_melt_pool_time_steps_chain(n, nSteps):
    spec = cases.case(4, n)
    grading = (1.3, None, 0.8)
    m = mo.box_mesh(*spec.n, spec.lengths, spec.patch_types, grading)
    xs, ys, zs = (hm.vertices(spec.n[d], spec.lengths[d], grading[d]) for d in range(3))
    geom = cases.box_polymesh(m, xs, ys, zs)
    f = cases.synthetic_fields(spec, m)
    rng = np.random.default_rng(4)
    f["phi_i"] = f["phi_i"] + 1e-3 * np.abs(f["phi_i"]).max() * rng.standard_normal(m.nIF)      # div(phi) != 0: the divU terms act
    P = spec.nPhases
    ctx = abi.Context(m, P)
    cases.load_case(ctx, spec, f)
    fc = np.asarray(m.faceCells, dtype=np.int64)
    ctx.set_cell_centres(m.C, np.asarray(m.Cfb) - np.asarray(m.C)[fc])
    for p in range(P):
        ctx.set_rho(p, RHO[p])
    for md in LEE:
        ctx.add_mass_transfer_lee(md["from"], md["to"], md["C"], md["Tactivate"], md["alphaMin"], abi.VAR_T, True)
    ctx.set_phase_pairs(PAIRS)
    ctx.set_mass_transfer_in_solve(True)
    ctx.laser_set_polymesh(m.Cf, m.Cfb, geom)
    ctx.laser_set_reflection(REFL, alphaCut=0.5)
    ctx.laser_set_local_density_absorption([0, 1, 2], *ABS)
    d = np.array([0.25, -1.0, 0.1]); d /= np.linalg.norm(d)
    r = np.cross(d, [0.0, 0.0, 1.0]); r /= np.linalg.norm(r)
    p0, p1, Ip, dA = cases.laser_seeds((0.5, 0.93, 0.5), d, r, 8, 16, 0.18, geom["boundsMag"], mode="Gaussian", power=200.0, sigma=0.08)
    cell = cases.find_cells_box(spec.n, xs, ys, zs, p0)
    assert (cell >= 0).all()
    mass = orc.MassHandle(LEE, PAIRS, RHO, None)
    ctl = abi.alpha_controls(nAlphaCorr=2, nAlphaSubCycles=1, schemePhi="vanLeer", schemePhir="linear", predictorTol=1e-13,
                             predictorMaxIter=400)
    octl = orc.alpha_ctl(nAlphaCorr=2, nAlphaSubCycles=1, predictorTol=1e-13, predictorMaxIter=400)
    ai, ab = [a.copy() for a in f["alpha_i"]], [a.copy() for a in f["alpha_b"]]
    T = f["T_i"].copy()
    for step in range(nSteps):
        ctx.set_phi(f["phi_i"], f["phi_b"])
        ctx.set_temperature(T, f["T_b"])
        # device chain, alpha stays resident between the calls
        ctx.correct_mass_sources()
        ctx.solve_alphas(ctl, f["deltaT"])
        ctx.div_capillary_stress()
        Fi, Fb = ctx.get_force()
        Q, nPart = ctx.laser_calculate(p0, p1, Ip, dA, cell)
        marks = ctx.laser_get()["reflectingCells"]
        # oracle chain
        dm = orc.correct_mass_sources(m, mass, ai, T)
        assert max(np.abs(x).max() for x in dm) > 0, "no melting in this case"
        ref = orc.solve_mt(m, orc.mules_cfg(3), octl, ai, ab, f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], RHO, mass, dm)
        ai, ab = ref["alpha_i"], ref["alpha_b"]
        Fr, Fbr = orc.div_capillary_stress(m, ai, ab, f["alphaBc"], f["sigma"], T, f["T_b"])
        lz = orc.laser_calculate(m, geom, REFL, ai, ab, f["alphaBc"], orc.local_density(ai, ABS[0]), orc.local_density(ai, ABS[1]),
                                 orc.local_density(ai, ABS[2]), T, p0, p1, Ip, dA, cell)
        for p in range(P):
            ga, gb = ctx.get_alpha(p)
            assert_field_close(ga, ai[p], f"alpha{p}, step {step}", rtol=1e-11)
            assert_field_close(gb, ab[p], f"alpha{p} boundary, step {step}", rtol=1e-11)
        assert_field_close(ctx.get_rho_phi()[0], ref["rhoPhi_i"], f"rhoPhi, step {step}", rtol=1e-11)
        assert np.abs(Fr).max() > 0
        assert_field_close(Fi, Fr, f"force, step {step}", rtol=1e-9)
        assert_field_close(Fb, Fbr, f"force boundary, step {step}", rtol=1e-9)
        assert np.array_equal(marks, lz["reflectingCells"]), f"reflecting cells, step {step}"
        assert nPart == lz["nParticles"] > len(Ip), (nPart, lz["nParticles"])
        assert np.abs(Q - lz["Q"]).max() <= 1e-9 * np.abs(lz["Q"]).max(), f"Q, step {step}"
        # the heat input warms the pool for the next step (a stand-in for TEqn; the same update on both sides)
        T = T + 1e-3 * lz["Q"] / max(np.abs(lz["Q"]).max(), 1e-300)
    ctx.close()
