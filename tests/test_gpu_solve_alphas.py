"""GPU parity of the rest of multiphaseSystem::solveAlphas() (flux assembly, implicit upwind predictor,
solve() with correctors / sub-cycles, rhoPhi, renormalisation) against the CPU oracle.

Tolerance: 1e-12 relative per entry + 1e-12*||field||_inf floor (helpers.assert_field_close).  The flux
assembly consumes Gauss gradients that the GPU computes with x*(1/V) and fma (tcif_device.cuh), so it is
not bit-identical; everything stays far inside the tolerance."""
import threading

import numpy as np
import pytest

from oracle import oracle as orc
from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm
from helpers import assert_field_close, global_mesh

pytestmark = pytest.mark.gpu

RHO = np.array([1000.0, 800.0, 1.2, 50.0])


def _delta_b(m):
    fc = np.asarray(m.faceCells, dtype=np.int64)
    db = np.asarray(m.Cfb) - np.asarray(m.C)[fc]
    nbrC = getattr(m, "nbrC", None)
    for p in range(len(m.patchType)):
        if m.patchType[p] == 3:
            sl = slice(int(m.patchStart[p]), int(m.patchStart[p] + m.patchSize[p]))
            db[sl] = np.asarray(nbrC)[sl] - np.asarray(m.C)[fc[sl]]
    return db


def _ctx(m, spec, f, **kw):
    ctx = abi.Context(m, spec.nPhases, **kw)
    cases.load_case(ctx, spec, f)
    ctx.set_cell_centres(m.C, _delta_b(m))
    for p in range(spec.nPhases):
        ctx.set_rho(p, RHO[p])
    return ctx


@pytest.mark.parametrize("config,n,sphi,sphir", [(5, 14, "vanLeer", "linear"), (5, 10, "upwind", "vanLeer"),
                                                 (3, 14, "linear", "linear"), (1, 24, "vanLeer", "vanLeer")])
def test_flux_assembly_matches_oracle(config, n, sphi, sphir):
    spec = cases.case(config, n)
    m = global_mesh(spec, grading=(2.0, None, None))
    f = cases.synthetic_fields(spec, m)
    ctx = _ctx(m, spec, f)
    ctx.flux_assembly(sphi, sphir)
    oi, ob = orc.flux_assembly(m, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], sphi, sphir)
    for p in range(spec.nPhases):
        gi, gb = ctx.get_flux_assembly(p)
        assert np.abs(oi[p]).max() > 0
        assert_field_close(gi, oi[p], f"phiAlphaCorr{p}")
        assert_field_close(gb, ob[p], f"phiAlphaCorr{p} boundary")
    ctx.close()


def test_flux_assembly_calpha_table():
    spec = cases.case(5, 10)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    cA = np.array([[1.0, 0.5, 2.0], [0.5, 1.0, 0.25], [2.0, 0.25, 1.0]])
    ctx = _ctx(m, spec, f)
    ctx.set_calpha(cA)
    ctx.flux_assembly("vanLeer", "linear")
    oi, _ = orc.flux_assembly(m, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], "vanLeer", "linear", cA)
    for p in range(3):
        assert_field_close(ctx.get_flux_assembly(p)[0], oi[p], f"phiAlphaCorr{p}")
    ctx.close()


def test_vanleer_without_geometry_is_an_error():
    spec = cases.case(3, 8)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    ctx = abi.Context(m, 2)
    cases.load_case(ctx, spec, f)
    with pytest.raises(abi.TcifError, match="cell centres"):
        ctx.flux_assembly("vanLeer", "linear")
    ctx.flux_assembly("upwind", "linear")
    ctx.close()


@pytest.mark.parametrize("fixed", [False, True])
def test_upwind_predictor_matches_oracle(fixed):
    spec = cases.case(3, 12)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    u = np.array([0.1, 0.03, -0.02])                     # through-flow: boundary coefficients matter
    f["phi_i"] = m.Sf @ u; f["phi_b"] = m.Sfb @ u
    bc = f["alphaBc"].copy()
    if fixed:
        bc[0, 0] = abi.BC_FIXED_VALUE
        f["alpha_b"][0][: m.patchSize[0]] = 0.6
    f["alphaBc"] = bc
    rng = np.random.default_rng(4)
    Su, Sp = 0.05 * rng.random(m.nCells), -0.1 * rng.random(m.nCells)
    ctx = _ctx(m, spec, f)
    ctx.set_su_sp(0, Su, Sp)
    for p in range(2):                                   # start the flux from zero
        ctx.set_phi_alpha(p, np.zeros(m.nIF), np.zeros(m.nBF))
    dt = f["deltaT"]
    n = ctx.upwind_predictor(0, dt, 1e-13, 300)
    qi, qb, pi, pb, nref = orc.upwind_predictor(m, dt, f["alpha_i"][0], f["alpha_b"][0], f["alpha_i"][0], bc[0],
                                                f["phi_i"], f["phi_b"], np.zeros(m.nIF), np.zeros(m.nBF), Su=Su, Sp=Sp,
                                                tol=1e-13, maxIter=300)
    assert n == nref and 1 < n < 300
    ai, ab = ctx.get_alpha(0)
    gi, gb = ctx.get_phi_alpha(0)
    assert np.array_equal(ai, qi) and np.array_equal(ab, qb)          # same sweeps, same order: bit-identical
    assert np.array_equal(gi, pi) and np.array_equal(gb, pb)
    ctx.close()


@pytest.mark.parametrize("config,n,implicit,ncorr,nsub", [(5, 12, True, 1, 1), (5, 12, False, 2, 1), (3, 12, True, 2, 2),
                                                          (1, 24, True, 1, 3)])
def test_solve_alphas_matches_oracle(config, n, implicit, ncorr, nsub):
    spec = cases.case(config, n)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    P = spec.nPhases
    ctx = _ctx(m, spec, f)
    ctl = abi.alpha_controls(ncorr, nsub, "vanLeer", "linear", implicit, 1e-13, 300)
    ctx.solve_alphas(ctl, f["deltaT"])
    octl = orc.alpha_ctl(ncorr, nsub, "vanLeer", "linear", implicit, 1e-13, 300)
    ref = orc.solve(m, orc.mules_cfg(3), octl, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"],
                    f["deltaT"], RHO[:P])
    for p in range(P):
        ai, ab = ctx.get_alpha(p)
        pi, pb = ctx.get_phi_alpha(p)
        assert np.abs(ai - f["alpha_i"][p]).max() > 0
        assert_field_close(ai, ref["alpha_i"][p], f"alpha{p}")
        assert_field_close(ab, ref["alpha_b"][p], f"alpha{p} boundary")
        assert_field_close(pi, ref["alphaPhi_i"][p], f"alphaPhi{p}")
        assert_field_close(pb, ref["alphaPhi_b"][p], f"alphaPhi{p} boundary")
    ri, rb = ctx.get_rho_phi()
    assert_field_close(ri, ref["rhoPhi_i"], "rhoPhi")
    assert_field_close(rb, ref["rhoPhi_b"], "rhoPhi boundary")
    ctx.close()


def test_solve_alphas_decomposed_matches_serial():
    """2x2x2 processor blocks (in-process group) == serial oracle, incl. the predictor's halo per sweep."""
    spec = cases.case(5, 12)
    g = global_mesh(spec)
    f = cases.synthetic_fields(spec, g)
    octl = orc.alpha_ctl(2, 1, "vanLeer", "linear", True, 1e-13, 300)
    ref = orc.solve(g, orc.mules_cfg(3), octl, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"],
                    f["deltaT"], RHO[:3])
    procs, nd = (2, 2, 2), 8
    meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, r) for r in range(nd)]
    fs = [cases.synthetic_fields(spec, m) for m in meshes]
    grp = abi.Group(nd)
    ctxs = [abi.Context(m, 3) for m in meshes]
    outs, errs = [None] * nd, []
    ctl = abi.alpha_controls(2, 1, "vanLeer", "linear", True, 1e-13, 300)

    def work(r):
        try:
            c = ctxs[r]
            c.comm_init_local(r, grp)
            cases.load_case(c, spec, fs[r])
            c.set_cell_centres(meshes[r].C, _delta_b(meshes[r]))
            for p in range(3):
                c.set_rho(p, RHO[p])
            c.solve_alphas(ctl, fs[r]["deltaT"])
            outs[r] = dict(alpha=[c.get_alpha(p)[0] for p in range(3)], aphi=[c.get_phi_alpha(p)[0] for p in range(3)],
                           rhoPhi=c.get_rho_phi()[0])
        except Exception as e:          # noqa: BLE001
            errs.append((r, e))
    ts = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(nd)]
    [t.start() for t in ts]
    [t.join(timeout=300) for t in ts]
    assert not errs, errs
    for r in range(nd):
        gid = np.asarray(meshes[r].cellGlobal); fa = np.asarray(meshes[r].faceProcAddressing) - 1
        for p in range(3):
            # the predictor converges to 1e-13 on both sides but may stop one sweep apart
            assert_field_close(outs[r]["alpha"][p], ref["alpha_i"][p][gid], f"alpha{p}", rtol=1e-11)
            assert_field_close(outs[r]["aphi"][p], ref["alphaPhi_i"][p][fa], f"alphaPhi{p}", rtol=1e-11)
        assert_field_close(outs[r]["rhoPhi"], ref["rhoPhi_i"][fa], "rhoPhi", rtol=1e-11)
    for c in ctxs:
        c.close()
    grp.close()


def test_time_step_host_matches_oracle():
    spec = cases.case(5, 12)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    ctx = _ctx(m, spec, f)
    ctl = abi.alpha_controls(1, 1, "vanLeer", "linear", True, 1e-13, 300)
    ai = [x.copy() for x in f["alpha_i"]]; ab = [x.copy() for x in f["alpha_b"]]
    ri, rb, Fi, Fb = np.zeros(m.nIF), np.zeros(m.nBF), np.zeros((m.nCells, 3)), np.zeros((m.nBF, 3))
    pi = [np.zeros(m.nIF) for _ in range(3)]; pb = [np.zeros(m.nBF) for _ in range(3)]
    ctx.time_step_host(ctl, f["deltaT"], ai, ab, f["T_i"], f["T_b"], f["phi_i"], f["phi_b"], ri, rb, Fi, Fb, pi, pb)
    ref = orc.solve(m, orc.mules_cfg(3), orc.alpha_ctl(1, 1, "vanLeer", "linear", True, 1e-13, 300), f["alpha_i"],
                    f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], RHO[:3])
    Fr, Fbr = orc.div_capillary_stress(m, ref["alpha_i"], ref["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"])
    for p in range(3):
        assert_field_close(ai[p], ref["alpha_i"][p], f"alpha{p}")
        assert_field_close(ab[p], ref["alpha_b"][p], f"alpha{p} boundary")
        assert_field_close(pi[p], ref["alphaPhi_i"][p], f"alphaPhi{p}")
    assert_field_close(ri, ref["rhoPhi_i"], "rhoPhi")
    assert_field_close(Fi, Fr, "force on the new alpha")
    assert_field_close(Fb, Fbr, "force boundary")
    ctx.close()
