"""GPU parity of SURVEY.md 8(f)-3 against the CPU oracle: correctMassSources (Lee::Kexp), alphaTransfer Su/Sp,
solve() with calculateSuSp() inside the correctors, addInterfacePorosity.  The kernels keep the reference's
operation order and the library is built with -fmad=false, so Su/Sp/dmdt are expected bit-identical (asserted);
alpha after a solve within 1e-12 (helpers.assert_field_close)."""
import numpy as np
import pytest

from oracle import oracle as orc
from thermocapillaryinterfoam_b200 import abi, cases
from helpers import assert_field_close, global_mesh

pytestmark = pytest.mark.gpu

RHO = [7000.0, 1.2, 6500.0]
PAIRS = [(0, 1), (0, 2), (1, 2)]
MODELS = [dict(**{"from": 0, "to": 2}, C=0.5, Tactivate=300.05, alphaMin=0.0),
          dict(**{"from": 2, "to": 0}, C=-0.25, Tactivate=299.98, alphaMin=0.01),
          dict(**{"from": 0, "to": 1}, C=3.0, Tactivate=300.02, alphaMin=0.0)]


def _setup(n=12, config=5, models=MODELS, pairs=PAIRS, rho_field=False, seed=11):
    spec = cases.case(config, n)
    m = global_mesh(spec, grading=(1.5, None, None))
    f = cases.synthetic_fields(spec, m)
    rng = np.random.default_rng(seed)
    # a flux with a non-zero divergence so that the div(phi) term is exercised
    f["phi_i"] = f["phi_i"] + 1e-3 * np.abs(f["phi_i"]).max() * rng.standard_normal(m.nIF)
    ctx = abi.Context(m, spec.nPhases)
    cases.load_case(ctx, spec, f)
    rf = None
    P = spec.nPhases
    if rho_field:
        rf = [RHO[p] * (1 + 0.05 * rng.standard_normal(m.nCells)) for p in range(P)]
        for p in range(P):
            ctx.set_rho(p, RHO[p], rf[p], np.full(m.nBF, RHO[p]))
    else:
        for p in range(P):
            ctx.set_rho(p, RHO[p])
    ctx.set_phi(f["phi_i"], f["phi_b"])
    for md in models:
        ctx.add_mass_transfer_lee(md["from"], md["to"], md["C"], md["Tactivate"], md.get("alphaMin", 0.0),
                                  md.get("modelVariable", abi.VAR_T), md.get("includeDivU", True))
    ctx.set_phase_pairs(pairs)
    mass = orc.MassHandle(models, pairs, RHO[:P], rf)
    return spec, m, f, ctx, mass


@pytest.mark.parametrize("rho_field", [False, True])
def test_correct_mass_sources_and_alpha_transfer_bitwise(rho_field):
    spec, m, f, ctx, mass = _setup(rho_field=rho_field)
    ctx.correct_mass_sources()
    dm = orc.correct_mass_sources(m, mass, f["alpha_i"], f["T_i"])
    assert max(d.max() for d in dm) > 0
    for md, d in zip(MODELS, dm):
        assert np.array_equal(ctx.get_dmdt(md["from"], md["to"]), d)
    mx = ctx.alpha_transfer()
    Su, Sp, mxo = orc.alpha_transfer(m, mass, f["alpha_i"], dm, f["phi_i"], f["phi_b"])
    assert any(np.abs(s).max() > 0 for s in Sp)
    for p in range(spec.nPhases):
        su, sp = ctx.get_su_sp(p)
        assert np.array_equal(su, Su[p]), f"Su{p}: max diff {np.abs(su - Su[p]).max():.3e}"
        assert np.array_equal(sp, Sp[p]), f"Sp{p}"
    assert mx == mxo
    ctx.close()


def test_alpha_transfer_host_dmdt_alpha_variable_and_no_divU():
    models = [dict(**{"from": 1, "to": 0}, C=2.0, Tactivate=0.5, alphaMin=0.1, modelVariable=abi.VAR_ALPHA, includeDivU=False),
              dict(**{"from": 2, "to": 1}, C=1.0, Tactivate=300.0)]
    spec, m, f, ctx, mass = _setup(models=models, pairs=[(1, 2), (0, 1), (0, 2)])
    rng = np.random.default_rng(5)
    dm = [np.full(m.nCells, 9.0), rng.uniform(-1, 1, m.nCells)]
    ctx.set_dmdt(1, 0, dm[0]); ctx.set_dmdt(2, 1, dm[1])
    mx = ctx.alpha_transfer()
    Su, Sp, mxo = orc.alpha_transfer(m, mass, f["alpha_i"], dm, f["phi_i"], f["phi_b"])
    for p in range(3):
        su, sp = ctx.get_su_sp(p)
        assert np.array_equal(su, Su[p]) and np.array_equal(sp, Sp[p])
    assert mx == mxo
    ctx.close()


def test_two_phase_2d_case():
    models = [dict(**{"from": 0, "to": 1}, C=1.5, Tactivate=300.01)]
    spec, m, f, ctx, mass = _setup(n=32, config=2, models=models, pairs=[(0, 1)])
    ctx.correct_mass_sources()
    ctx.alpha_transfer()
    dm = orc.correct_mass_sources(m, mass, f["alpha_i"], f["T_i"])
    Su, Sp, _ = orc.alpha_transfer(m, mass, f["alpha_i"], dm, f["phi_i"], f["phi_b"])
    for p in range(2):
        su, sp = ctx.get_su_sp(p)
        assert np.array_equal(su, Su[p]) and np.array_equal(sp, Sp[p])
    ctx.close()


@pytest.mark.parametrize("nAlphaCorr,nSub", [(1, 1), (2, 2)])
def test_solve_with_mass_transfer_in_the_correctors(nAlphaCorr, nSub):
    spec, m, f, ctx, mass = _setup(n=12)
    fc = np.asarray(m.faceCells, dtype=np.int64)
    ctx.set_cell_centres(m.C, np.asarray(m.Cfb) - np.asarray(m.C)[fc])
    ctx.correct_mass_sources()
    ctx.set_mass_transfer_in_solve(True)
    ctl = abi.alpha_controls(nAlphaCorr=nAlphaCorr, nAlphaSubCycles=nSub, schemePhi="vanLeer", schemePhir="linear")
    ctx.solve_alphas(ctl, f["deltaT"])
    dm = orc.correct_mass_sources(m, mass, f["alpha_i"], f["T_i"])
    ref = orc.solve_mt(m, orc.mules_cfg(3), orc.alpha_ctl(nAlphaCorr=nAlphaCorr, nAlphaSubCycles=nSub), f["alpha_i"], f["alpha_b"],
                       f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], RHO, mass, dm)
    plain = orc.solve(m, orc.mules_cfg(3), orc.alpha_ctl(nAlphaCorr=nAlphaCorr, nAlphaSubCycles=nSub), f["alpha_i"], f["alpha_b"],
                      f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], RHO)
    assert max(np.abs(plain["alpha_i"][p] - ref["alpha_i"][p]).max() for p in range(3)) > 1e-9
    for p in range(3):
        ai, ab = ctx.get_alpha(p)
        assert_field_close(ai, ref["alpha_i"][p], f"alpha{p}", rtol=1e-11)
    ri, rb = ctx.get_rho_phi()
    assert_field_close(ri, ref["rhoPhi_i"], "rhoPhi", rtol=1e-11)
    ctx.close()


def test_interface_porosity_bitwise():
    spec, m, f, ctx, mass = _setup(models=[], n=10)
    d0 = np.linspace(1.0, 2.0, m.nCells)
    got = ctx.interface_porosity([0, 2], [1e6, 5.0], d0)
    ref = d0 + orc.interface_porosity(m, [0, 2], [1e6, 5.0], f["alpha_i"])
    # the oracle accumulates into a zero field and the sum d0 + (t0 + t2) is re-associated: 2 ulp
    assert np.allclose(got, ref, rtol=4e-16, atol=0)
    ctx.close()


def test_errors():
    spec, m, f, ctx, mass = _setup(models=[], n=8)
    with pytest.raises(RuntimeError):
        ctx.add_mass_transfer_lee(0, 0, 1.0, 300.0)
    ctx.add_mass_transfer_lee(0, 1, 1.0, 300.0)
    with pytest.raises(RuntimeError):
        ctx.add_mass_transfer_lee(0, 1, 1.0, 300.0)          # same key twice
    with pytest.raises(RuntimeError):
        ctx.set_dmdt(1, 2, np.zeros(m.nCells))               # no such model
    ctx.close()


@pytest.mark.parametrize("procs", [(2, 1, 1), (2, 2, 2)])
def test_decomposed_alpha_transfer_matches_serial(procs):
    """Processor patches: div(phi) of a boundary-adjacent cell includes the processor-face fluxes, ddtAlphaMax is a
    gMax over ranks.  Su/Sp are cell-local after that, so the decomposed result must be bit-identical to the serial
    oracle up to the face-sum order of div(phi) (1e-12)."""
    import threading
    from thermocapillaryinterfoam_b200 import mesh as hm
    spec = cases.case(5, 12)
    g = global_mesh(spec)
    fg = cases.synthetic_fields(spec, g)
    rng = np.random.default_rng(11)
    phig = fg["phi_i"] + 1e-3 * np.abs(fg["phi_i"]).max() * rng.standard_normal(g.nIF)
    mass = orc.MassHandle(MODELS, PAIRS, RHO)
    dm = orc.correct_mass_sources(g, mass, fg["alpha_i"], fg["T_i"])
    Su, Sp, mxo = orc.alpha_transfer(g, mass, fg["alpha_i"], dm, phig, fg["phi_b"])
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
            # the perturbed global flux restricted to this rank (processor faces: sign by orientation)
            fa = np.asarray(m.faceProcAddressing) - 1
            phi_i = phig[fa]
            phi_b = np.array(f["phi_b"])
            for k in range(m.nPatches):
                if m.patchType[k] == 3:
                    sl = slice(int(m.patchStart[k]), int(m.patchStart[k] + m.patchSize[k]))
                    a = np.asarray(m.bfaceProcAddressing)[sl]
                    phi_b[sl] = np.sign(a) * phig[np.abs(a) - 1]
            ctx.set_phi(phi_i, phi_b)
            for p in range(3):
                ctx.set_rho(p, RHO[p])
            for md in MODELS:
                ctx.add_mass_transfer_lee(md["from"], md["to"], md["C"], md["Tactivate"], md.get("alphaMin", 0.0))
            ctx.set_phase_pairs(PAIRS)
            ctx.correct_mass_sources()
            mx = ctx.alpha_transfer()
            outs[r] = dict(mx=mx, su=[ctx.get_su_sp(p) for p in range(3)], gid=np.asarray(m.cellGlobal))
        except Exception as e:          # noqa: BLE001
            errs.append((r, e))
    ts = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(nd)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(120)
    assert not errs, errs
    for r in range(nd):
        gid = outs[r]["gid"]
        assert outs[r]["mx"] == mxo
        for p in range(3):
            assert_field_close(outs[r]["su"][p][0], Su[p][gid], f"Su{p} rank {r}")
            assert np.array_equal(outs[r]["su"][p][1], Sp[p][gid]), f"Sp{p} rank {r}"
    for c in ctxs:
        c.close()
    grp.close()
