"""Out-of-bounds writes next to a device array (the stand-in for a compute-sanitizer memcheck pass, which the GPU pool
does not allow): contexts created under TCIF_GUARD=1 keep a 512-byte guard zone filled with a pattern on either side
of every device allocation; after a 2x2x1-rank decomposed stage + solve() and after one laserDTRM calculate() -- the
two paths the round-1 review asked to see under memcheck -- every zone must still hold the pattern, and the results
must equal the ones of unguarded contexts bit for bit (a guard zone that shifted an array would show up there)."""
import threading

import numpy as np
import pytest

from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm
import test_gpu_laser as tl

pytestmark = pytest.mark.gpu


def _decomposed_stage_and_solve(spec, procs):
    nd = procs[0] * procs[1] * procs[2]
    meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, r) for r in range(nd)]
    fs = [cases.synthetic_fields(spec, m) for m in meshes]
    grp = abi.Group(nd)
    ctxs = [abi.Context(m, spec.nPhases) for m in meshes]
    outs, errs = [None] * nd, []
    ctl = abi.alpha_controls(nAlphaCorr=1, nAlphaSubCycles=2, schemePhi="vanLeer", schemePhir="linear", predictorTol=1e-12,
                             predictorMaxIter=200)

    def work(r):
        try:
            c, f, m = ctxs[r], fs[r], meshes[r]
            c.comm_init_local(r, grp)
            cases.load_case(c, spec, f)
            c.stage(f["deltaT"])
            F = c.get_force()[0].copy()
            fc = np.asarray(m.faceCells, dtype=np.int64)
            c.set_cell_centres(np.asarray(m.C), np.asarray(m.Cfb) - np.asarray(m.C)[fc])
            for p in range(spec.nPhases):
                c.set_rho(p, (1000.0, 800.0, 1.2, 50.0)[p])
            c.solve_alphas(ctl, f["deltaT"])
            outs[r] = (F, [c.get_alpha(p)[0].copy() for p in range(spec.nPhases)], c.get_rho_phi()[0].copy())
        except Exception as e:          # noqa: BLE001
            errs.append((r, repr(e)))
    ts = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(nd)]
    [t.start() for t in ts]
    [t.join(timeout=300) for t in ts]
    assert not errs, errs
    return ctxs, grp, outs


def test_no_write_outside_any_array_in_a_decomposed_stage_and_solve(monkeypatch):
    spec = cases.case(5, 12)
    procs = (2, 2, 1)
    ctxs, grp, plain = _decomposed_stage_and_solve(spec, procs)
    for c in ctxs:
        with pytest.raises(abi.TcifError, match="TCIF_GUARD"):
            c.debug_guard_check()
        c.close()
    grp.close()
    monkeypatch.setenv("TCIF_GUARD", "1")
    ctxs, grp, guarded = _decomposed_stage_and_solve(spec, procs)
    for r, c in enumerate(ctxs):
        zones, corrupt = c.debug_guard_check()
        assert zones > 100 and corrupt == 0, (r, zones, corrupt)
        assert np.array_equal(guarded[r][0], plain[r][0]) and np.array_equal(guarded[r][2], plain[r][2])
        for a, b in zip(guarded[r][1], plain[r][1]):
            assert np.array_equal(a, b)
        c.close()
    grp.close()


def test_no_write_outside_any_array_in_laser_calculate(monkeypatch):
    monkeypatch.setenv("TCIF_GUARD", "1")
    spec, m, v, geom, f, ctx = tl._setup()
    ctx.laser_set_reflection(tl.MODELS, alphaCut=0.5)
    ctx.laser_set_local_density_absorption([0, 1, 2], (60.0, 25.0, 0.3), (0.2, 0.1, 0.0), (0.0, 0.0, 0.0))
    p0, p1, Ip, dA, cell = tl._seeds(spec, v, geom, (0.5, 0.93, 0.5), (0.25, -1.0, 0.1))
    Q, nPart = ctx.laser_calculate(p0, p1, Ip, dA, cell)
    assert nPart > len(Ip) and np.abs(Q).max() > 0
    cells, vals = ctx.laser_get_sparse(capacity=16)
    zones, corrupt = ctx.debug_guard_check()
    assert zones > 100 and corrupt == 0, (zones, corrupt)
    ctx.close()
