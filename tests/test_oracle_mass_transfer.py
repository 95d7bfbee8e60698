"""Known-answer tests of the ORACLE for SURVEY.md 8(f)-3: alphaTransfer Su/Sp (MassTransferPhaseSystem.C:503-704),
Lee::Kexp/KSu (Lee.C:52-167), fvc::div(phi), addInterfacePorosity + VollerPrakash::S().  CPU only."""
import numpy as np

from oracle import oracle as orc, mesh_oracle as mo
from thermocapillaryinterfoam_b200 import cases
from helpers import global_mesh


def _mesh(n=6):
    spec = cases.case(5, n)
    m = global_mesh(spec)
    return spec, m, cases.synthetic_fields(spec, m)


def _flux_of(m, u):
    """phi = S_f . u(face centre) on internal and boundary faces."""
    return np.einsum("ij,ij->i", m.Sf, u(m.Cf)), np.einsum("ij,ij->i", m.Sfb, u(m.Cfb))


def test_div_phi_of_linear_velocity_is_exact():
    # K10: u = (x, 2y, 3z) -> div u = 6 in every cell, boundary cells included (Gauss theorem is exact for
    # a linear field with face-centre quadrature)
    spec, m, f = _mesh(7)
    phii, phib = _flux_of(m, lambda X: X * np.array([1.0, 2.0, 3.0]))
    d = orc.div_phi(m, phii, phib)
    assert np.abs(d - 6.0).max() < 1e-11


def test_lee_kexp_hand_values():
    # Kexp = C*from*rho*pos(from-alphaMin)*(T-Tact)/Tact * pos(+-(T-Tact))   (Lee.C:52-85)
    spec, m, f = _mesh(4)
    nC = m.nCells
    a0 = np.linspace(-0.1, 1.1, nC)
    T = np.linspace(250.0, 450.0, nC)[::-1].copy()
    alphas = [a0, 1 - a0, np.zeros(nC)]
    for Cc, Tact, amin in [(10.0, 330.0, 0.0), (-5.0, 320.0, 0.2)]:
        mass = orc.MassHandle([{"from": 0, "to": 1, "C": Cc, "Tactivate": Tact, "alphaMin": amin}], [(0, 1)], [7000.0, 1.0, 2.0])
        got = orc.correct_mass_sources(m, mass, alphas, T)[0]
        fr = np.clip(a0, 0, 1)
        coeff = Cc * fr * 7000.0 * (fr - amin >= 0) * (T - Tact) / Tact
        ref = coeff * ((T - Tact >= 0) if Cc > 0 else (Tact - T >= 0))
        assert np.allclose(got, ref, rtol=1e-15, atol=0)
        assert (got >= 0).all()          # evaporation/melting rates are non-negative for either sign of C


def test_alpha_transfer_branches_two_phases():
    # phi = 0 -> div(phi) = 0; one model 0 -> 1 with a stored dmdt of both signs; every branch of :613-697
    spec, m, f = _mesh(4)
    nC = m.nCells
    rng = np.random.default_rng(7)
    a0 = rng.uniform(-0.05, 1.05, nC)
    alphas = [a0, 1 - a0, np.zeros(nC)]
    dm = rng.uniform(-3.0, 3.0, nC)
    dm[:5] = 0.0
    phii, phib = np.zeros(m.nIF), np.zeros(m.nBF)
    for rho in ([1000.0, 10.0, 1.0], [10.0, 1000.0, 1.0]):
        mass = orc.MassHandle([{"from": 0, "to": 1, "C": 1.0, "Tactivate": 300.0}], [(0, 1)], rho)
        Su, Sp, mx = orc.alpha_transfer(m, mass, alphas, [dm], phii, phib)
        c1, c2 = 1 / rho[0], 1 / rho[1]
        net = -dm                                  # dmdt21 - dmdt12
        c12 = c1 - c2
        aL = np.clip(a0, 0, 1)
        imp1 = ((net > 0) & (c12 > 0)) | ((net < 0) & (c12 < 0))
        su1 = c1 * net - np.where(~imp1 & (net != 0), net * c12 * aL, 0.0)
        sp1 = -np.where(imp1, net * c12, 0.0)
        assert np.allclose(Su[0], su1, rtol=1e-14, atol=1e-18) and np.allclose(Sp[0], sp1, rtol=1e-14, atol=1e-18)
        assert (Sp[0] <= 0).all() and (Sp[1] <= 0).all()      # implicit parts only ever stabilise
        # volume consistency: sum_p (Su_p + Sp_p alpha_p) = net*(c1-c2)*(1 - a1L - a2L) -> 0 where a1+a2 = 1 in [0,1]
        inside = (a0 >= 0) & (a0 <= 1)
        tot = Su[0] + Sp[0] * aL + Su[1] + Sp[1] * np.clip(1 - a0, 0, 1)
        assert np.abs(tot[inside]).max() < 1e-14 * max(1.0, np.abs(net).max() * abs(c12))
        assert np.all(Su[2] == 0) and np.all(Sp[2] == 0)
        assert mx == max((0.0 * c1), (dm * c2).max())            # :700-702, dmdt21 = 0


def test_alpha_transfer_without_divU_conserves_mass():
    # includeDivU false: Su_p = +-dmdtNet/rho_p only -> rho1 Su1 + rho2 Su2 = 0, Sp = 0
    spec, m, f = _mesh(4)
    nC = m.nCells
    rng = np.random.default_rng(3)
    a0 = rng.uniform(0, 1, nC)
    dm = rng.uniform(0, 2.0, nC)
    mass = orc.MassHandle([{"from": 1, "to": 0, "C": 1.0, "Tactivate": 300.0, "includeDivU": False}], [(0, 1)], [500.0, 2.0, 1.0])
    Su, Sp, _ = orc.alpha_transfer(m, mass, [a0, 1 - a0, np.zeros(nC)], [dm], f["phi_i"] + 1e-3, f["phi_b"])
    assert np.abs(500.0 * Su[0] + 2.0 * Su[1]).max() < 1e-13 * dm.max()
    assert all(np.all(s == 0) for s in Sp)


def test_alpha_transfer_div_term_added_once_per_pair():
    # no models, includeDivU stays true: every pair adds div(phi)*clamp(alpha) to both of its phases (:599-606),
    # so with 3 pairs a phase that sits in two pairs receives the term twice -- parity means reproducing that
    spec, m, f = _mesh(5)
    phii, phib = _flux_of(m, lambda X: X * np.array([1.0, 2.0, 3.0]))
    mass = orc.MassHandle([], [(0, 1), (0, 2), (1, 2)], [1.0, 1.0, 1.0])
    Su, Sp, _ = orc.alpha_transfer(m, mass, f["alpha_i"], None, phii, phib)
    d = orc.div_phi(m, phii, phib)
    for p in range(3):
        assert np.allclose(Su[p], 2 * d * np.clip(f["alpha_i"][p], 0, 1), rtol=1e-14, atol=1e-15)


def test_ksu_alpha_variable_replaces_stored_dmdt():
    # modelVariable == alpha: dmdt12 = KSu(alpha, phase1) = -C*from*rho*pos(from-alphaMin)*pos(alpha1 - Tactivate) (:541-552)
    spec, m, f = _mesh(4)
    nC = m.nCells
    a0 = np.linspace(0, 1, nC)
    model = {"from": 0, "to": 1, "C": 2.0, "Tactivate": 0.5, "alphaMin": 0.1, "modelVariable": orc.VAR_ALPHA, "includeDivU": False}
    mass = orc.MassHandle([model], [(0, 1)], [4.0, 8.0, 1.0])
    Su, Sp, _ = orc.alpha_transfer(m, mass, [a0, 1 - a0, np.zeros(nC)], [np.full(nC, 123.0)], np.zeros(m.nIF), np.zeros(m.nBF))
    d12 = -(2.0 * a0 * 4.0 * (a0 - 0.1 >= 0)) * (a0 - 0.5 >= 0)
    assert np.allclose(Su[0], (1 / 4.0) * (0 - d12), rtol=1e-15) and np.allclose(Su[1], (1 / 8.0) * d12, rtol=1e-15)


def test_interface_porosity_hand_value():
    spec, m, f = _mesh(4)
    a = f["alpha_i"]
    got = orc.interface_porosity(m, [0, 2], [1e6, 5.0], a)
    ref = m.V * 1e6 * a[0] ** 2 / ((1 - a[0]) ** 3 + 1e-3) + m.V * 5.0 * a[2] ** 2 / ((1 - a[2]) ** 3 + 1e-3)
    assert np.allclose(got, ref, rtol=1e-14)


def test_solve_mt_equals_solve_with_precomputed_sources_for_one_corrector():
    # nAlphaCorr = 1, no sub-cycles: calculateSuSp() sees the alpha the step starts from
    spec, m, f = _mesh(8)
    rho = [1000.0, 800.0, 1.2]
    mass = orc.MassHandle([{"from": 0, "to": 2, "C": 0.5, "Tactivate": 300.05}], [(0, 1), (0, 2), (1, 2)], rho)
    dm = orc.correct_mass_sources(m, mass, f["alpha_i"], f["T_i"])
    assert dm[0].max() > 0
    Su, Sp, _ = orc.alpha_transfer(m, mass, f["alpha_i"], dm, f["phi_i"], f["phi_b"])
    cfg, ctl = orc.mules_cfg(3), orc.alpha_ctl(schemePhi="vanLeer", schemePhir="linear")
    a = orc.solve(m, cfg, ctl, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], rho, Su=Su, Sp=Sp)
    b = orc.solve_mt(m, cfg, ctl, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], rho, mass, dm)
    for p in range(3):
        assert np.array_equal(a["alpha_i"][p], b["alpha_i"][p])
    ref = orc.solve(m, cfg, ctl, f["alpha_i"], f["alpha_b"], f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], rho)
    assert max(np.abs(ref["alpha_i"][p] - b["alpha_i"][p]).max() for p in range(3)) > 0      # the sources act
