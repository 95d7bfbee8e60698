"""GPU parity of the species MULES (MultiComponentPhaseModel::solveYi) against the CPU oracle.
Tolerance 1e-11 relative + 1e-11*||.||inf (gradients use x*(1/V)+fma on the device, the predictor may stop one
sweep apart); the phase-fraction state must come back untouched."""
import threading

import numpy as np
import pytest

from oracle import oracle as orc
from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm
from helpers import assert_field_close, global_mesh

pytestmark = pytest.mark.gpu


def _species(m, nS, seed=3):
    X = np.asarray(m.C)
    rng = np.random.default_rng(seed)
    raw = [0.2 + 0.15 * np.sin(5.0 * X[:, 0] + k) * np.cos(4.0 * X[:, 1] - k) + 0.02 * rng.standard_normal(m.nCells)
           for k in range(nS)]
    tot = sum(raw) + 0.3
    Y = [r / tot for r in raw]
    fc = np.asarray(m.faceCells, dtype=np.int64)
    Yb = [y[fc].copy() for y in Y]
    return Y, Yb


def _delta_b(m):
    fc = np.asarray(m.faceCells, dtype=np.int64)
    return np.asarray(m.Cfb) - np.asarray(m.C)[fc]


@pytest.mark.parametrize("config,n,ph,sphi,sphir,cYi", [(5, 12, 0, "vanLeer", "linear", 1.0), (5, 10, 2, "upwind", "vanLeer", 0.5),
                                                      (3, 12, 1, "linear", "linear", 1.0), (2, 24, 0, "vanLeer", "vanLeer", 1.0)])
def test_solve_yi_matches_oracle(config, n, ph, sphi, sphir, cYi):
    spec = cases.case(config, n)
    m = global_mesh(spec, grading=(1.5, None, None))
    f = cases.synthetic_fields(spec, m)
    nS = 3
    Y, Yb = _species(m, nS)
    ybc = np.zeros((nS, m.nPatches), np.int32)
    ybc[0, 0] = abi.BC_FIXED_VALUE if hasattr(abi, "BC_FIXED_VALUE") else 1
    Yb[0][int(m.patchStart[0]):int(m.patchStart[0] + m.patchSize[0])] = 0.25
    rng = np.random.default_rng(9)
    Su = [1e-2 * rng.uniform(0, 1, m.nCells), None, None]
    Sp = [None, -1e-2 * rng.uniform(0, 1, m.nCells), None]
    ctx = abi.Context(m, spec.nPhases)
    cases.load_case(ctx, spec, f)
    ctx.set_cell_centres(m.C, _delta_b(m))
    a_before = [ctx.get_alpha(p) for p in range(spec.nPhases)]
    yi, yb = ctx.solve_yi(ph, Y, Yb, ybc, 2, f["deltaT"], cYi=cYi, schemePhi=sphi, schemePhir=sphir, Su=Su, Sp=Sp)
    ri, rb = orc.solve_yi(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], ph, Y, Yb, ybc, 2,
                          f["phi_i"], f["phi_b"], f["deltaT"], cYi=cYi, schemePhi=sphi, schemePhir=sphir, Su=Su, Sp=Sp)
    for i in range(nS):
        assert np.abs(ri[i] - Y[i]).max() > 1e-6                  # something happened
        assert_field_close(yi[i], ri[i], f"Y{i}", rtol=1e-11)
        assert_field_close(yb[i], rb[i], f"Y{i} boundary", rtol=1e-11)
    for p in range(spec.nPhases):                                 # phase fractions untouched
        ai, ab = ctx.get_alpha(p)
        assert np.array_equal(ai, a_before[p][0]) and np.array_equal(ab, a_before[p][1])
    # and the alpha path still works afterwards
    ctx.stage(f["deltaT"])
    ref = orc.stage(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                    f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    for p in range(spec.nPhases):
        assert_field_close(ctx.get_alpha(p)[0], ref["alpha_i"][p], f"alpha{p} after solve_yi")
    ctx.close()


def test_solve_yi_errors():
    spec = cases.case(5, 8)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    Y, Yb = _species(m, 2)
    ctx = abi.Context(m, 3)
    cases.load_case(ctx, spec, f)
    with pytest.raises(RuntimeError):      # vanLeer without cell centres
        ctx.solve_yi(0, Y, Yb, np.zeros((2, m.nPatches), np.int32), 1, f["deltaT"])
    with pytest.raises(RuntimeError):      # deltaT must be positive
        ctx.solve_yi(0, Y, Yb, np.zeros((2, m.nPatches), np.int32), 1, -1.0, schemePhi="upwind", schemePhir="upwind")
    ctx.close()


@pytest.mark.parametrize("procs", [(2, 1, 1), (2, 2, 2)])
def test_decomposed_solve_yi_matches_serial(procs):
    """Processor patches: halo of Yi, grad(Yi), grad(alpha) and the coupled form of the face interpolation."""
    spec = cases.case(5, 12)
    g = global_mesh(spec)
    fg = cases.synthetic_fields(spec, g)
    nS = 3
    Y, Yb = _species(g, nS)
    ybc = np.zeros((nS, g.nPatches), np.int32)
    ri, rb = orc.solve_yi(g, orc.mules_cfg(3), fg["alpha_i"], fg["alpha_b"], fg["alphaBc"], 0, Y, Yb, ybc, 2,
                          fg["phi_i"], fg["phi_b"], fg["deltaT"])
    nd = procs[0] * procs[1] * procs[2]
    meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, r) for r in range(nd)]
    fs = [cases.synthetic_fields(spec, m) for m in meshes]
    grp = abi.Group(nd)
    ctxs = [abi.Context(m, spec.nPhases) for m in meshes]
    outs, errs = [None] * nd, []

    def work(r):
        try:
            m, f, ctx = meshes[r], fs[r], ctxs[r]
            ctx.comm_init_local(r, grp)
            cases.load_case(ctx, spec, f)
            ctx.set_cell_centres(f["C"], f["deltab"])
            gid = np.asarray(m.cellGlobal)
            fc = np.asarray(m.faceCells, dtype=np.int64)
            yl = [y[gid] for y in Y]
            ybl = [y[fc].copy() for y in yl]                      # walls zeroGradient; processor values come from the halo
            bc = np.zeros((nS, m.nPatches), np.int32)
            yi, yb = ctx.solve_yi(0, yl, ybl, bc, 2, f["deltaT"])
            outs[r] = (gid, yi)
        except Exception as e:          # noqa: BLE001
            errs.append((r, e))
    ts = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(nd)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(120)
    assert not errs, errs
    for r in range(nd):
        gid, yi = outs[r]
        for i in range(nS):
            assert_field_close(yi[i], ri[i][gid], f"Y{i} rank {r}", rtol=1e-11)
    for c in ctxs:
        c.close()
    grp.close()


@pytest.mark.parametrize("nSub", [2, 3])
def test_solve_yi_sub_cycles_match_oracle(nSub):
    """nYiSubCycles > 1 (MultiComponentPhaseModel.C:345-364): n explicitSolve calls with deltaT/n, Yi.oldTime() re-stored
    from the current Yi at every sub-cycle, the flux left in place by the previous limit()."""
    spec = cases.case(5, 12)
    m = global_mesh(spec, grading=(1.5, None, None))
    f = cases.synthetic_fields(spec, m)
    nS = 3
    Y, Yb = _species(m, nS)
    ybc = np.zeros((nS, m.nPatches), np.int32)
    rng = np.random.default_rng(2)
    Su = [1e-2 * rng.uniform(0, 1, m.nCells), None, None]
    ctx = abi.Context(m, spec.nPhases)
    cases.load_case(ctx, spec, f)
    ctx.set_cell_centres(m.C, _delta_b(m))
    yi, yb = ctx.solve_yi(0, Y, Yb, ybc, 2, f["deltaT"], Su=Su, nYiSubCycles=nSub)
    ri, rb = orc.solve_yi(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], 0, Y, Yb, ybc, 2,
                          f["phi_i"], f["phi_b"], f["deltaT"], Su=Su, nYiSubCycles=nSub)
    r1, _ = orc.solve_yi(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], 0, Y, Yb, ybc, 2,
                         f["phi_i"], f["phi_b"], f["deltaT"], Su=Su)
    assert max(np.abs(ri[i] - r1[i]).max() for i in range(nS)) > 1e-9      # sub-cycling changes the result
    for i in range(nS):
        assert_field_close(yi[i], ri[i], f"Y{i}", rtol=1e-11)
        assert_field_close(yb[i], rb[i], f"Y{i} boundary", rtol=1e-11)
    ctx.close()
