"""GPU parity: the CUDA path through the C ABI against the CPU oracle on identical inputs.

Tolerance (north_star): 1e-12 relative per cell with the absolute floor of helpers.RTOL*||field||_inf;
built with -fmad=false the serial CUDA path is in fact expected to be bit-identical to the oracle,
which `test_serial_bitwise` asserts.
"""
import numpy as np
import pytest

from oracle import oracle as orc
from thermocapillaryinterfoam_b200 import abi, cases
from helpers import assert_field_close, global_mesh

pytestmark = pytest.mark.gpu


def _ctx(mesh, spec, f, **kw):
    ctx = abi.Context(mesh, spec.nPhases, **kw)
    cases.load_case(ctx, spec, f)
    return ctx


@pytest.mark.parametrize("config,n", [(1, 32), (2, 48), (3, 20), (5, 24)])
def test_force_matches_oracle(config, n):
    spec = cases.case(config, n)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    ctx = _ctx(m, spec, f)
    ctx.div_capillary_stress()
    Fi, Fb = ctx.get_force()
    Ri, Rb = orc.div_capillary_stress(m, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"])
    assert np.abs(Ri).max() > 0
    assert_field_close(Fi, Ri, "force internal")
    assert_field_close(Fb, Rb, "force boundary")
    ctx.close()


@pytest.mark.parametrize("config,n", [(2, 40), (5, 16)])
def test_grad_matches_oracle(config, n):
    spec = cases.case(config, n)
    m = global_mesh(spec, grading=(3.0, 0.5, None))
    f = cases.synthetic_fields(spec, m)
    bc = f["alphaBc"].copy()
    bc[0, 0] = abi.BC_FIXED_VALUE        # a fixedValue wall exercises the snGrad correction
    f["alphaBc"] = bc
    f["alpha_b"][0][: m.patchSize[0]] = 0.25
    ctx = _ctx(m, spec, f)
    # A Gauss gradient is a cancelling sum of terms Sf*alpha_f/V: where alpha is nearly uniform the result
    # is tiny while each term is O(|alpha| |Sf|/V), so rounding (the GPU path uses x*(1/V) and fma) is
    # bounded by a few ulp of the TERMS, not of the result: 64 eps * max|alpha| * max(sum|Sf|/V).
    floor = 64 * np.finfo(float).eps * 6 * np.abs(m.Sf).max() / m.V.min()
    for p in range(spec.nPhases):
        gi, gb = ctx.grad_alpha(p)
        ri, rb = orc.grad(m, f["alpha_i"][p], f["alpha_b"][p], bc[p])
        assert_field_close(gi, ri, f"grad alpha{p}", floor=floor)
        assert_field_close(gb, rb, f"grad alpha{p} boundary", floor=floor)
    ctx.close()


@pytest.mark.parametrize("config,n,L", [(1, 32, 3), (3, 16, 3), (5, 20, 3), (5, 12, 1), (5, 12, 0), (3, 12, 5)])
def test_stage_matches_oracle(config, n, L):
    spec = cases.case(config, n)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    ctx = _ctx(m, spec, f, nLimiterIter=L)
    ctx.stage(f["deltaT"])
    ref = orc.stage(m, orc.mules_cfg(L), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                    f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    Fi, Fb = ctx.get_force()
    assert_field_close(Fi, ref["F_i"], "force")
    assert_field_close(Fb, ref["F_b"], "force boundary")
    for p in range(spec.nPhases):
        ai, ab = ctx.get_alpha(p)
        pi, pb = ctx.get_phi_alpha(p)
        assert np.abs(ai - f["alpha_i"][p]).max() > 0, "alpha did not move"
        assert_field_close(ai, ref["alpha_i"][p], f"alpha{p}")
        assert_field_close(ab, ref["alpha_b"][p], f"alpha{p} boundary")
        assert_field_close(pi, ref["alphaPhi_i"][p], f"alphaPhi{p}")
        assert_field_close(pb, ref["alphaPhi_b"][p], f"alphaPhi{p} boundary")
    ctx.close()


def test_serial_bitwise():
    """-fmad=false + reference accumulation order => MULES results bit-identical to the serial oracle."""
    spec = cases.case(5, 16)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    ctx = _ctx(m, spec, f)
    ctx.stage(f["deltaT"])
    ref = orc.stage(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                    f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    ndiff = 0
    for p in range(spec.nPhases):
        ai, _ = ctx.get_alpha(p)
        pi, _ = ctx.get_phi_alpha(p)
        ndiff += int((ai != ref["alpha_i"][p]).sum()) + int((pi != ref["alphaPhi_i"][p]).sum())
    assert ndiff == 0, f"{ndiff} MULES values differ in the last bit(s)"
    # the force kernels trade the reference's per-component divisions for one reciprocal per cell and
    # contract multiply-adds (TCIF_FORCE_EXACT=0, tcif_device.cuh): a few ulp, not bit-identical
    Fi, _ = ctx.get_force()
    assert_field_close(Fi, ref["F_i"], "force", rtol=1e-14)
    if "exact" in abi.version():
        assert int((Fi != ref["F_i"]).sum()) == 0
    ctx.close()


def test_fine_grained_mules_calls():
    """tcif_mules_limit / limit_sum / add_upwind_flux / explicit_solve == the fused stage."""
    spec = cases.case(5, 14)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    a = _ctx(m, spec, f)
    b = _ctx(m, spec, f)
    dt = f["deltaT"]
    a.stage(dt)
    b.store_old_alphas()
    for p in range(3):
        b.mules_limit(p, 1.0 / dt, True)
    b.mules_limit_sum()
    for p in range(3):
        b.add_upwind_flux(p)
        b.mules_explicit_solve(p, dt)
    for p in range(3):
        assert np.array_equal(a.get_alpha(p)[0], b.get_alpha(p)[0])
        assert np.array_equal(a.get_phi_alpha(p)[0], b.get_phi_alpha(p)[0])
    a.close(); b.close()


def test_stage_host_roundtrip():
    spec = cases.case(3, 16)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    ctx = _ctx(m, spec, f)
    ai = [x.copy() for x in f["alpha_i"]]; ab = [x.copy() for x in f["alpha_b"]]
    pi = [np.zeros(m.nIF) for _ in range(2)]; pb = [np.zeros(m.nBF) for _ in range(2)]
    Fi, Fb = np.zeros((m.nCells, 3)), np.zeros((m.nBF, 3))
    ctx.stage_host(f["deltaT"], ai, ab, f["T_i"], f["T_b"], f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"],
                   pi, pb, Fi, Fb)
    ctx.stage_host(f["deltaT"], [x.copy() for x in f["alpha_i"]], [x.copy() for x in f["alpha_b"]], f["T_i"], f["T_b"],
                   f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], pi, pb, Fi, Fb)   # repeatable: inputs untouched
    ref = orc.stage(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                    f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    assert_field_close(Fi, ref["F_i"], "force")
    for p in range(2):
        assert_field_close(ai[p], ref["alpha_i"][p], "alpha")
        assert_field_close(pi[p], ref["alphaPhi_i"][p], "alphaPhi")
    ctx.close()


def test_ldu_csr_bit_exact():
    from oracle import mesh_oracle as mo
    spec = cases.case(5, 9)
    m = global_mesh(spec)
    ctx = abi.Context(m, 3)
    got = ctx.ldu_csr()
    ref = mo.ldu_csr(m.owner, m.neighbour, m.nCells)
    for g, r in zip(got, ref):
        assert np.array_equal(g, r)
    ctx.close()


def test_mass_conservation_and_bounds():
    spec = cases.case(5, 24)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    ctx = _ctx(m, spec, f)
    before = [ctx.alpha_stats(p) for p in range(3)]
    ctx.stage(f["deltaT"])
    after = [ctx.alpha_stats(p) for p in range(3)]
    for p in range(3):
        # phi_b = 0 on every wall: sum(V*alpha) can only change by round-off
        assert abs(after[p][0] - before[p][0]) <= 1e-13 * max(1.0, abs(before[p][0]))
        assert after[p][1] >= -1e-12 and after[p][2] <= 1.0 + 1e-12
    ctx.close()


def test_errors_are_loud():
    spec = cases.case(1, 8)
    m = global_mesh(spec)
    ctx = abi.Context(m, 2)
    with pytest.raises(abi.TcifError):
        ctx.div_capillary_stress()           # no surfaceTension entry (phaseSystem.C:1135)
    with pytest.raises(abi.TcifError):
        ctx.set_sigma(0, 1, ("fancy", [(1, 2)]))
    ctx.set_sigma(0, 1, ("polynomial", [(1.0, 0.0), (-1e-4, 1.0)]))
    with pytest.raises(abi.TcifError):
        ctx.div_capillary_stress()           # temperatureDependent without T
    with pytest.raises(abi.TcifError):
        abi.Context(m, 7)
    ctx.close()


def test_flux_of_force_and_cache():
    """fvc::flux(divCapillaryStress()) as pEqn.H:35 consumes it; F is recomputed only when alpha/T/sigma change."""
    spec = cases.case(5, 14)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    ctx = _ctx(m, spec, f)
    Ri, Rb = orc.div_capillary_stress(m, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"])
    ri, rb = orc.flux_of_vector(m, Ri, Rb)
    gi, gb = ctx.flux_of_force()
    assert_field_close(gi, ri, "flux(F)")
    assert_field_close(gb, rb, "flux(F) boundary")
    assert ctx.force_counters() == (1, 0)
    ctx.div_capillary_stress(); ctx.flux_of_force()
    assert ctx.force_counters() == (1, 2)                     # same alpha, T: served from the cache
    ctx.set_temperature(f["T_i"] + 1.0, f["T_b"] + 1.0)       # sigma(T) changed
    ctx.div_capillary_stress()
    assert ctx.force_counters() == (2, 2)
    ctx.stage(f["deltaT"])                                    # its force is a hit; then MULES moves alpha
    ctx.div_capillary_stress()
    assert ctx.force_counters() == (3, 3)
    ctx.set_force_cache(False)
    ctx.div_capillary_stress()
    assert ctx.force_counters() == (4, 3)
    ctx.close()
