"""Golden fixtures (tests/golden/*.npz, written by tests/golden/make_golden.py from the oracle):
the oracle must keep reproducing them (CPU; 1e-13 -- the synthetic inputs go through numpy's tanh/sin, whose last
bit may differ between host CPUs), and the CUDA path must match them within the 1e-12 contract (GPU)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden as mg  # noqa: E402
from thermocapillaryinterfoam_b200 import abi, cases  # noqa: E402
from helpers import assert_field_close, global_mesh   # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name,config,n", mg.CASES)
def test_oracle_reproduces_golden(name, config, n):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    got = mg.build(config, n)
    assert set(z.files) == set(got)
    for k in z.files:
        assert_field_close(got[k], z[k], k, rtol=1e-13)


@pytest.mark.gpu
@pytest.mark.parametrize("name,config,n", mg.CASES)
def test_cuda_path_matches_golden(name, config, n):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    spec = cases.case(config, n)
    m = global_mesh(spec)
    f = cases.synthetic_fields(spec, m)
    P = spec.nPhases
    ctx = abi.Context(m, P)
    cases.load_case(ctx, spec, f)
    ctx.stage(f["deltaT"])
    assert_field_close(ctx.get_force()[0], z["stage_F_i"], "force")
    for p in range(P):
        assert_field_close(ctx.get_alpha(p)[0], z[f"stage_alpha_i{p}"], f"alpha{p}")
        assert_field_close(ctx.get_phi_alpha(p)[0], z[f"stage_alphaPhi_i{p}"], f"alphaPhi{p}")
    ctx.close()
    ctx = abi.Context(m, P)
    cases.load_case(ctx, spec, f)
    ctx.set_cell_centres(f["C"], f["deltab"])
    for p in range(P):
        ctx.set_rho(p, mg.RHO[p])
    ctx.solve_alphas(abi.alpha_controls(2, 1, "vanLeer", "linear", True, 1e-13, 300), f["deltaT"])
    for p in range(P):
        assert_field_close(ctx.get_alpha(p)[0], z[f"solve_alpha_i{p}"], f"solve alpha{p}")
        assert_field_close(ctx.get_phi_alpha(p)[0], z[f"solve_alphaPhi_i{p}"], f"solve alphaPhi{p}")
    assert_field_close(ctx.get_rho_phi()[0], z["solve_rhoPhi_i"], "rhoPhi")
    ctx.close()


def test_oracle_reproduces_mass_transfer_golden():
    z = np.load(os.path.join(GOLD, mg.MASS_CASE[0] + ".npz"))
    got = mg.build_mass(*mg.MASS_CASE[1:])
    assert set(z.files) == set(got)
    for k in z.files:
        assert_field_close(got[k], z[k], k, rtol=1e-13)


@pytest.mark.gpu
def test_cuda_path_matches_mass_transfer_golden():
    z = np.load(os.path.join(GOLD, mg.MASS_CASE[0] + ".npz"))
    spec, m, f = mg.mass_inputs(*mg.MASS_CASE[1:])
    ctx = abi.Context(m, 3)
    cases.load_case(ctx, spec, f)
    ctx.set_cell_centres(f["C"], f["deltab"])
    for p in range(3):
        ctx.set_rho(p, mg.MASS_RHO[p])
    for md in mg.MASS_MODELS:
        ctx.add_mass_transfer_lee(md["from"], md["to"], md["C"], md["Tactivate"], md["alphaMin"])
    ctx.set_phase_pairs(mg.MASS_PAIRS)
    ctx.correct_mass_sources()
    for i, md in enumerate(mg.MASS_MODELS):
        assert_field_close(ctx.get_dmdt(md["from"], md["to"]), z[f"dmdt{i}"], f"dmdt{i}")
    mx = ctx.alpha_transfer()
    assert abs(mx - float(z["ddtAlphaMax"][0])) <= 1e-12 * abs(mx)
    for p in range(3):
        su, sp = ctx.get_su_sp(p)
        assert_field_close(su, z[f"Su{p}"], f"Su{p}")
        assert_field_close(sp, z[f"Sp{p}"], f"Sp{p}")
    ctx.set_mass_transfer_in_solve(True)
    ctx.solve_alphas(abi.alpha_controls(2, 1, "vanLeer", "linear", True, 1e-13, 300), f["deltaT"])
    for p in range(3):
        assert_field_close(ctx.get_alpha(p)[0], z[f"solve_alpha_i{p}"], f"solve alpha{p}", rtol=1e-11)
    ctx.close()


def test_oracle_reproduces_laser_golden():
    z = np.load(os.path.join(GOLD, mg.LASER_CASE[0] + ".npz"))
    got = mg.build_laser(*mg.LASER_CASE[1:])
    assert set(z.files) == set(got)
    assert int(z["nParticles"][0]) > 72                     # the fixture exercises reflection
    for k in z.files:
        if k in ("reflectingCells", "nParticles"):
            assert np.array_equal(got[k], z[k]), k
        else:
            assert_field_close(got[k], z[k], k, rtol=1e-13)


@pytest.mark.gpu
def test_cuda_path_matches_laser_golden():
    z = np.load(os.path.join(GOLD, mg.LASER_CASE[0] + ".npz"))
    spec, m, geom, f, s = mg.laser_inputs(*mg.LASER_CASE[1:])
    ctx = abi.Context(m, 3)
    cases.load_case(ctx, spec, f)
    ctx.set_cell_centres(f["C"], f["deltab"])
    ctx.laser_set_polymesh(m.Cf, m.Cfb, geom)
    ctx.laser_set_reflection(mg.LASER_MODELS)
    ctx.laser_set_local_density_absorption([0, 1, 2], mg.LASER_ABS["a"], mg.LASER_ABS["e"], mg.LASER_ABS["E"])
    Q, n = ctx.laser_calculate(s["p0"], s["p1"], s["I"], s["dA"], s["cell"])
    got = ctx.laser_get()
    assert n == int(z["nParticles"][0])
    assert np.array_equal(got["reflectingCells"], z["reflectingCells"])
    assert_field_close(got["nHat"], z["nHat"], "nHat")
    assert_field_close(got["nHatPoints"], z["nHatPoints"], "nHatPoints")
    assert np.abs(Q - z["Q"]).max() <= 1e-11 * np.abs(z["Q"]).max()
    ctx.close()
