"""Oracle parity at the BASELINE.json sizes, on the code path bench.py times.

configs[0..3] at their stated sizes (64^2, 512^2, 128^3, 200^3) against the SERIAL oracle stage
(`orc_stage`, restating phaseSystem.C:1038-1105 and multiphaseSystem.C:296-375): MULES results (alpha, alphaPhi)
bit-identical, force within 1e-12 (helpers.RTOL: relative per cell + 1e-12*||F||_inf floor).
configs[4] at 400^3 = 64 M cells against the THREADED oracle (16 sub-domains, `orc_stage_threads`), 1e-12 --
the decomposed oracle sums the faces of a cell in another order than the serial one, so bitwise is not expected.

The CUDA side runs with the library defaults (TCIF_TILE_ORDER / TCIF_PF_AHEAD / TCIF_INLINE_BND unset), i.e. the
strip-wise tile order and the L2 prefetch are ON where bench.py has them on; the 400^3 case asserts that.
Everything goes through the C ABI.
"""
import gc
import os

import numpy as np
import pytest

from oracle import oracle as orc
from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm
from helpers import assert_field_close

pytestmark = pytest.mark.gpu

_ENV = ("TCIF_TILE_ORDER", "TCIF_PF_AHEAD", "TCIF_STRIP_CELLS", "TCIF_INLINE_BND", "TCIF_TSUM")


def _free_hbm():
    import torch
    return torch.cuda.mem_get_info(0)[0]


def _host_ram_gb():
    with open("/proc/meminfo") as fh:
        for ln in fh:
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) / 1e6
    return 0.0


def _cuda_stage(spec, m, f):
    """One tcif_stage on the whole mesh; returns F, alpha[], alphaPhi[] (internal + boundary) and the tile-order flag."""
    for k in _ENV:
        assert k not in os.environ, f"{k} is set: this test must run the library defaults"
    ctx = abi.Context(m, spec.nPhases, device=0)
    cases.load_case(ctx, spec, f)
    ctx.stage(f["deltaT"])
    out = dict(F=ctx.get_force(), alpha=[ctx.get_alpha(p) for p in range(spec.nPhases)],
               alphaPhi=[ctx.get_phi_alpha(p) for p in range(spec.nPhases)], strips=ctx.tile_strips)
    ctx.close()
    return out


@pytest.mark.parametrize("config,need_gb", [(1, 1), (2, 1), (3, 4), (4, 14)])
def test_stage_matches_serial_oracle_at_baseline_size(config, need_gb):
    if _free_hbm() < need_gb * 1e9:
        pytest.skip(f"needs {need_gb} GB of free HBM")
    spec = cases.case(config)                         # the BASELINE.json size
    m = hm.box_block(spec.n, spec.lengths, spec.patch_types, (1, 1, 1), 0)
    f = cases.synthetic_fields(spec, m)
    got = _cuda_stage(spec, m, f)
    # (the strip-wise tile order switches on where one mesh plane of all streams exceeds the L2 share: 400^3, not 200^3)
    ref = orc.stage(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                    f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    assert np.abs(ref["F_i"]).max() > 0
    assert_field_close(got["F"][0], ref["F_i"], "force")
    assert_field_close(got["F"][1], ref["F_b"], "force boundary")
    ndiff = 0
    for p in range(spec.nPhases):
        assert np.abs(ref["alpha_i"][p] - f["alpha_i"][p]).max() > 0, "alpha did not move"
        ndiff += int((got["alpha"][p][0] != ref["alpha_i"][p]).sum()) + int((got["alpha"][p][1] != ref["alpha_b"][p]).sum())
        ndiff += int((got["alphaPhi"][p][0] != ref["alphaPhi_i"][p]).sum()) + int((got["alphaPhi"][p][1] != ref["alphaPhi_b"][p]).sum())
    assert ndiff == 0, f"{ndiff} MULES values differ from the serial oracle in the last bit(s)"
    m.close()


def test_stage_matches_threaded_oracle_at_400_cubed():
    """configs[4], 64 M cells: one B200 against the oracle on 16 sub-domains (one thread each)."""
    if _free_hbm() < 90e9:
        pytest.skip("needs 90 GB of free HBM")
    if _host_ram_gb() < 110:
        pytest.skip("needs 110 GB of host memory for the 64 M-cell oracle run")
    spec = cases.case(5)
    P = spec.nPhases
    m = hm.box_block(spec.n, spec.lengths, spec.patch_types, (1, 1, 1), 0)
    f = cases.synthetic_fields(spec, m)
    got = _cuda_stage(spec, m, f)
    assert got["strips"] == 1, "strip-wise tile order expected at 400^3 (the path bench.py times)"
    dt = f["deltaT"]
    del f
    m.close()
    del m
    gc.collect()
    procs = (4, 2, 2)
    nd = 16
    meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, r) for r in range(nd)]
    fs = [cases.synthetic_fields(spec, mm) for mm in meshes]
    _, res = orc.stage_threads(meshes, orc.mules_cfg(3), [x["alpha_i"] for x in fs], [x["alpha_b"] for x in fs],
                               [x["alphaBc"] for x in fs], spec.sigma, [x["T_i"] for x in fs], [x["T_b"] for x in fs],
                               [x["phi_i"] for x in fs], [x["phi_b"] for x in fs], [x["phiAlpha_i"] for x in fs],
                               [x["phiAlpha_b"] for x in fs], dt, nStages=1)
    del fs
    Fg, ag, pg = got["F"][0], [a[0] for a in got["alpha"]], [a[0] for a in got["alphaPhi"]]
    scaleF = np.abs(Fg).max()
    scaleP = [np.abs(x).max() for x in pg]
    assert scaleF > 0
    rt = 1e-12
    for d in range(nd):
        cg = np.asarray(meshes[d].cellGlobal)
        fg = np.asarray(meshes[d].faceProcAddressing) - 1          # global internal face of every local internal face
        r = res[d]
        a, b = Fg[cg], r["F_i"]
        assert (np.abs(a - b) <= rt * np.maximum(np.abs(a), np.abs(b)) + rt * scaleF).all(), f"force, sub-domain {d}"
        for p in range(P):
            a, b = ag[p][cg], r["alpha_i"][p]
            assert (np.abs(a - b) <= rt * np.maximum(np.abs(a), np.abs(b)) + rt).all(), f"alpha{p}, sub-domain {d}"
            a, b = pg[p][fg], r["alphaPhi_i"][p]
            assert (np.abs(a - b) <= rt * np.maximum(np.abs(a), np.abs(b)) + rt * scaleP[p]).all(), f"alphaPhi{p}, sub-domain {d}"
    for mm in meshes:
        mm.close()
