"""Known-answer tests that pin the CPU oracle.

The reference ships no test or golden vector for this path and OpenFOAM v2106 is absent
("parity unpinned"), so the oracle is pinned by closed-form answers of the operators it restates.
"""
import numpy as np
import pytest

from oracle import oracle as orc, mesh_oracle as mo

ZG, FV = orc.BC_ZERO_GRADIENT, orc.BC_FIXED_VALUE


def _patch_slices(m):
    return [slice(int(m.patchStart[p]), int(m.patchStart[p] + m.patchSize[p])) for p in range(m.nPatches)]


def _interior(m):
    mask = np.ones(m.nCells, bool)
    mask[m.faceCells] = False
    return mask


# ---- K1: Gauss-linear gradient of a linear field ------------------------------------------------
@pytest.mark.parametrize("n", [(6, 5, 4), (8, 8, 1)])
def test_k1_gradient_of_linear_field(n):
    pt = (0, 0, 0, 0, 1, 1) if n[2] == 1 else (0,) * 6
    m = mo.box_mesh(*n, (1.0, 0.8, 0.6), pt)
    a = np.array([0.3, -0.7, 0.0 if n[2] == 1 else 1.1])
    ai = m.C @ a + 0.2
    # fixedValue on every wall with the exact face value: gradient is exact in every cell
    ab = m.Cfb @ a + 0.2
    bc = np.full(m.nPatches, FV, np.int32)
    gi, gb = orc.grad(m, ai, ab, bc)
    assert np.allclose(gi, a, rtol=0, atol=1e-12)
    # boundary gradient: g_b = g_c + n (snGrad - n.g_c); snGrad = dc*(a_b - a_c) = n.a exactly
    assert np.allclose(gb, a, rtol=0, atol=1e-11)
    # zeroGradient walls: interior exact, wall-normal component reduced in wall cells, boundary
    # normal gradient removed (snGrad = 0)
    ab0 = ai[m.faceCells]
    gi0, gb0 = orc.grad(m, ai, ab0, np.full(m.nPatches, ZG, np.int32))
    inner = _interior(m)
    assert np.allclose(gi0[inner], a, rtol=0, atol=1e-12)
    nb = m.Sfb / m.magSfb[:, None]
    assert np.allclose(np.einsum("ij,ij->i", nb, gb0), 0.0, atol=1e-12)


def test_k1_graded_mesh_constant_field():
    m = mo.box_mesh(5, 4, 3, (1.0, 1.0, 1.0), (0,) * 6, (4.0, 0.3, None))
    ai = np.full(m.nCells, 0.37)
    gi, gb = orc.grad(m, ai, ai[m.faceCells], np.zeros(m.nPatches, np.int32))
    scale = 0.37 * np.abs(m.Sfb).max() / m.V.min()
    assert np.abs(gi).max() <= 1e-13 * scale        # closed surface: sum Sf = 0 to round-off
    assert np.abs(gb).max() <= 1e-13 * scale


# ---- A.2 linear interpolation -------------------------------------------------------------------
def test_interpolate_linear_exact_on_graded_mesh():
    m = mo.box_mesh(6, 3, 2, (1.0, 1.0, 1.0), (0,) * 6, (5.0, None, 0.5))
    a = np.array([1.0, 2.0, -3.0])
    vi, vb = m.C @ a, m.Cfb @ a
    si, sb = orc.interpolate(m, vi, vb)
    assert np.allclose(si, m.Cf @ a, atol=1e-13)     # linear weights reproduce linear fields
    assert np.array_equal(sb, vb)


# ---- A.5 sigma Function1 ------------------------------------------------------------------------
def test_sigma_function1():
    assert orc.sigma_eval(("constant", 0.07), 123.0) == 0.07
    poly = ("polynomial", [(0.1, 0.0), (-1e-4, 1.0), (2e-9, 2.0)])
    T = 350.0
    assert orc.sigma_eval(poly, T) == pytest.approx(0.1 - 1e-4 * T + 2e-9 * T * T, rel=1e-15)
    tab = ("table", [(300.0, 0.07), (400.0, 0.06), (600.0, 0.01)])
    assert orc.sigma_eval(tab, 250.0) == 0.07        # clamped
    assert orc.sigma_eval(tab, 700.0) == 0.01
    assert orc.sigma_eval(tab, 350.0) == pytest.approx(0.065, rel=1e-14)
    assert orc.sigma_eval(tab, 500.0) == pytest.approx(0.035, rel=1e-14)


def test_deltaN():
    m = mo.box_mesh(4, 4, 4, (2.0, 2.0, 2.0))
    assert orc.deltaN(m) == pytest.approx(1e-8 / 0.5, rel=1e-14)


# ---- K2/K3/K4/K5: capillary stress force ---------------------------------------------------------
def _two_phase(m, a1i, a1b):
    return [a1i, 1.0 - a1i], [a1b, 1.0 - a1b], np.zeros((2, m.nPatches), np.int32)


def test_k2_planar_interface_constant_sigma_has_no_force_inside():
    """alpha = smooth function of y only, sigma constant: T depends on y only and T_yy == T_xy == 0
    (n = e_y), so div T vanishes away from the walls."""
    n = (6, 24, 5)
    m = mo.box_mesh(*n)
    prof = lambda X: 0.5 * (1.0 - np.tanh((X[:, 1] - 0.5) / 0.08))
    a1i = prof(m.C)
    ai, ab, bc = _two_phase(m, a1i, a1i[m.faceCells])
    Ti, Tb = np.full(m.nCells, 300.0), np.full(m.nBF, 300.0)
    Fi, Fb = orc.div_capillary_stress(m, ai, ab, bc, [("constant", 0.07)], Ti, Tb)
    inner = _interior(m)
    g, _ = orc.grad(m, ai[0], ab[0], bc[0])
    scale = 2 * 0.07 * np.abs(g).max() / (1.0 / 24)
    assert np.abs(Fi[inner][:, [0, 2]]).max() <= 1e-12 * scale
    # normal component: T_yy = 2 sigma |g| (1 - (|g|/(|g|+deltaN))^2) ~ 4 sigma deltaN, i.e. only the
    # deltaN regularisation of nHat survives (phaseSystem.C:1287-1294); its divergence is bounded by it
    assert np.abs(Fi[inner][:, 1]).max() <= 4 * 0.07 * orc.deltaN(m) / (1.0 / 24)


def test_k3_telescoping_sum():
    """sum_c V_c F_c == sum over boundary faces of S_b . T_b: internal faces cancel pairwise.  With
    alpha uniform near the walls T_b == 0, so the volume integral of the force vanishes to round-off."""
    m = mo.box_mesh(14, 12, 10, (1.0, 1.0, 1.0), (0,) * 6, (2.0, None, 0.7))
    r = np.linalg.norm(m.C - np.array([0.45, 0.5, 0.55]), axis=1)
    a1i = 0.5 * (1.0 - np.tanh((r - 0.15) / 0.04))
    a1i[a1i < 1e-5] = 0.0                                    # exactly uniform far field
    ai, ab, bc = _two_phase(m, a1i, a1i[m.faceCells])
    assert np.all(ab[0] == 0.0)
    Ti = 300.0 + 100.0 * m.C[:, 0]
    Tb = 300.0 + 100.0 * m.Cfb[:, 0]
    sig = [("polynomial", [(0.1, 0.0), (-1e-4, 1.0)])]
    Fi, _ = orc.div_capillary_stress(m, ai, ab, bc, sig, Ti, Tb)
    tot = (m.V[:, None] * Fi).sum(axis=0)
    ref = np.abs(m.V[:, None] * Fi).sum(axis=0).max()
    assert ref > 0
    assert np.abs(tot).max() <= 1e-12 * ref


def test_k4_marangoni_flat_interface():
    """Flat interface normal to y, sigma = s0 - gam (T - T0), T linear in x:  T_xx = 2 sigma |grad a|,
    so integral F_x dy = 2 dsigma/dx = -2 gam dT/dx (the coded factor 2); F_y integrates to 0."""
    nx, ny = 8, 64
    m = mo.box_mesh(nx, ny, 1, (1.0, 1.0, 1.0 / nx), (0, 0, 0, 0, 1, 1))
    a1i = 0.5 * (1.0 - np.tanh((m.C[:, 1] - 0.5) / 0.05))
    ai, ab, bc = _two_phase(m, a1i, a1i[m.faceCells])
    s0, gam, G = 0.07, 1e-4, 100.0
    Ti, Tb = 300.0 + G * m.C[:, 0], 300.0 + G * m.Cfb[:, 0]
    sig = [("polynomial", [(s0 + gam * 300.0, 0.0), (-gam, 1.0)])]
    Fi, _ = orc.div_capillary_stress(m, ai, ab, bc, sig, Ti, Tb)
    F = Fi.reshape(ny, nx, 3)
    dy = 1.0 / ny
    cols = slice(1, nx - 1)                                  # away from the x walls
    Ix = F[:, cols, 0].sum(axis=0) * dy
    Iy = F[:, cols, 1].sum(axis=0) * dy
    # integral of 2|grad a| dy across the interface = 2 (discrete sum of central differences telescopes)
    assert np.allclose(Ix, -2.0 * gam * G, rtol=1e-6)
    assert np.abs(Iy).max() <= 1e-10 * 2.0 * gam * G
    assert np.abs(F[..., 2]).max() == 0.0


def test_k5_laplace_jump_circle():
    """2-D circle radius R, constant sigma: -int F.n dn across the interface -> 2 sigma / R (coded factor 2)."""
    n = 160
    m = mo.box_mesh(n, n, 1, (1.0, 1.0, 1.0 / n), (0, 0, 0, 0, 1, 1))
    R, s0, wdt = 0.25, 0.07, 3.0 / n
    d = m.C[:, :2] - 0.5
    r = np.linalg.norm(d, axis=1)
    a1i = 0.5 * (1.0 - np.tanh((r - R) / wdt))
    ai, ab, bc = _two_phase(m, a1i, a1i[m.faceCells])
    Fi, _ = orc.div_capillary_stress(m, ai, ab, bc, [("constant", s0)], np.zeros(m.nCells), np.zeros(m.nBF))
    # area integral of F.e_r / (2 pi r) over the annulus = radial line integral averaged over angle
    er = d / np.maximum(r, 1e-300)[:, None]
    Fr = np.einsum("ij,ij->i", Fi[:, :2], er)
    band = np.abs(r - R) < 8 * wdt
    jump = -(Fr[band] / (2 * np.pi * r[band])).sum() * (1.0 / n) ** 2
    assert jump == pytest.approx(2.0 * s0 / R, rel=0.03)


def test_pair_order_three_phases_symmetric_sum():
    """F is a sum over unordered pairs; permuting phases (and sigma entries accordingly) changes
    only rounding (T is even in g)."""
    m = mo.box_mesh(10, 9, 8)
    X = m.C
    b1 = 0.5 * (1 - np.tanh((np.linalg.norm(X - [0.35, 0.5, 0.5], axis=1) - 0.2) / 0.06))
    b2 = 0.5 * (1 - np.tanh((np.linalg.norm(X - [0.65, 0.5, 0.5], axis=1) - 0.2) / 0.06))
    a3 = np.maximum(1 - b1 - b2, 0)
    s = b1 + b2 + a3
    ai = [b1 / s, b2 / s, a3 / s]
    ab = [a[m.faceCells] for a in ai]
    bc = np.zeros((3, m.nPatches), np.int32)
    T = np.full(m.nCells, 300.0), np.full(m.nBF, 300.0)
    sg = [("constant", 0.07), ("constant", 0.05), ("constant", 0.03)]      # (01),(02),(12)
    F1, _ = orc.div_capillary_stress(m, ai, ab, bc, sg, *T)
    perm = [2, 0, 1]                                                       # new phase k = old perm[k]
    sg2 = [("constant", 0.05), ("constant", 0.03), ("constant", 0.07)]     # (20),(21),(01)
    F2, _ = orc.div_capillary_stress(m, [ai[k] for k in perm], [ab[k] for k in perm], bc, sg2, *T)
    assert np.allclose(F1, F2, rtol=0, atol=1e-12 * np.abs(F1).max())


# ---- K6: MULES ------------------------------------------------------------------------------------
def _rot_flux(m, umag=0.1):
    """phi = Sf . u(Cf) with the solenoidal u = (sin pi x cos pi y, -cos pi x sin pi y, 0) umag (zero on walls)."""
    def u(X):
        v = np.zeros_like(X)
        v[:, 0] = umag * np.sin(np.pi * X[:, 0]) * np.cos(np.pi * X[:, 1])
        v[:, 1] = -umag * np.cos(np.pi * X[:, 0]) * np.sin(np.pi * X[:, 1])
        return v
    return np.einsum("ij,ij->i", m.Sf, u(m.Cf)), np.zeros(m.nBF)


def _ho_flux(m, phii, psi):
    o, n = m.owner, m.neighbour
    return phii * (m.w * (psi[o] - psi[n]) + psi[n])


def test_k6_limiter_lambda_in_unit_interval_and_bounded_update():
    m = mo.box_mesh(24, 24, 1, (1.0, 1.0, 1.0 / 24), (0, 0, 0, 0, 1, 1))
    r = np.linalg.norm(m.C[:, :2] - [0.5, 0.7], axis=1)
    psi = (r < 0.18).astype(float)                            # sharp disc: unlimited HO flux over/undershoots
    psib = psi[m.faceCells]
    bc = np.zeros(m.nPatches, np.int32)
    phii, phib = _rot_flux(m)
    dt = 0.4 * (1.0 / 24) / 0.1
    cfg = orc.mules_cfg(3)
    z = np.zeros(m.nCells)
    ho = _ho_flux(m, phii, psi)
    bdi, bdb = orc.upwind_flux(m, phii, phib, psi, psib)
    lam_i, lam_b = orc.mules_limiter(m, cfg, 1.0 / dt, psi, psib, psi, bc, bdi, bdb, ho - bdi, np.zeros(m.nBF), z, z)
    assert lam_i.min() >= 0.0 and lam_i.max() <= 1.0
    assert (lam_i < 1.0).any() and (lam_i == 1.0).any()
    # unlimited HO update violates the bounds, limited one does not; mass is conserved to round-off
    def update(flux):
        d = np.zeros(m.nCells)
        np.add.at(d, m.owner, flux); np.subtract.at(d, m.neighbour, flux)
        return psi - dt * d / m.V
    raw = update(ho)
    assert raw.max() > 1.0 + 1e-3 or raw.min() < -1e-3
    qi, qb, pi, pb = orc.mules_explicit_solve(m, cfg, dt, psi, psib, psi, bc, phii, phib, ho, np.zeros(m.nBF), z, z)
    assert qi.min() >= -1e-12 and qi.max() <= 1.0 + 1e-12
    assert np.array_equal(qb, qi[m.faceCells])                 # zeroGradient re-evaluated
    assert abs((m.V * qi).sum() - (m.V * psi).sum()) <= 1e-13 * (m.V * psi).sum()
    assert np.array_equal(qi, update(pi)) or np.allclose(qi, update(pi), rtol=0, atol=1e-15)


def test_k6_uniform_field_stays_uniform_and_smooth_field_unlimited():
    m = mo.box_mesh(12, 10, 6)
    phii, phib = _rot_flux(m)
    bc = np.zeros(m.nPatches, np.int32)
    z = np.zeros(m.nCells)
    dt = 0.2 * (1.0 / 12) / 0.1
    cfg = orc.mules_cfg(3)
    psi = np.full(m.nCells, 0.4)
    # div-free phi discretely: sum of face fluxes per cell
    d = np.zeros(m.nCells)
    np.add.at(d, m.owner, phii); np.subtract.at(d, m.neighbour, phii)
    ho = _ho_flux(m, phii, psi)
    qi, _, _, _ = orc.mules_explicit_solve(m, cfg, dt, psi, psi[m.faceCells], psi, bc, phii, phib, ho, np.zeros(m.nBF), z, z)
    assert np.allclose(qi, 0.4 - dt * 0.4 * d / m.V, rtol=0, atol=1e-14)
    # a monotone field well inside (0,1) with a small time step: lambda == 1 everywhere
    psi = 0.2 + 0.3 * m.C[:, 0] + 0.25 * m.C[:, 1]            # monotone: no interior local extremum
    ho = _ho_flux(m, phii, psi)
    bdi, bdb = orc.upwind_flux(m, phii, phib, psi, psi[m.faceCells])
    lam_i, _ = orc.mules_limiter(m, cfg, 1.0 / (0.02 * dt), psi, psi[m.faceCells], psi, bc, bdi, bdb, ho - bdi,
                                 np.zeros(m.nBF), z, z)
    assert np.array_equal(lam_i, np.ones(m.nIF))


def test_k6_limit_returncorr_consistency():
    """limit(returnCorr=false) == upwind flux + limit(returnCorr=true); boundary faces of non-coupled
    patches keep their flux (phiBD_b := phiPsi_b)."""
    m = mo.box_mesh(10, 9, 4)
    rng = np.random.default_rng(5)
    psi = rng.random(m.nCells)
    psib = psi[m.faceCells]
    phii, phib = _rot_flux(m)
    bc = np.zeros(m.nPatches, np.int32)
    z = np.zeros(m.nCells)
    cfg = orc.mules_cfg(3)
    ho, hob = _ho_flux(m, phii, psi), 0.01 * rng.standard_normal(m.nBF)
    ci, cb = orc.mules_limit(m, cfg, 50.0, psi, psib, psi, bc, phii, phib, ho, hob, z, z, True)
    fi, fb = orc.mules_limit(m, cfg, 50.0, psi, psib, psi, bc, phii, phib, ho, hob, z, z, False)
    bdi, _ = orc.upwind_flux(m, phii, phib, psi, psib)
    assert np.allclose(fi, bdi + ci, rtol=0, atol=1e-16)
    assert np.array_equal(cb, np.zeros(m.nBF)) and np.array_equal(fb, hob)


def test_k6_one_dimensional_step_is_monotone():
    """1-D advection of a step with CFL 0.5: the limited solution stays monotone and in [0,1]."""
    n = 40
    m = mo.box_mesh(n, 1, 1, (1.0, 1.0 / n, 1.0 / n), (0, 0, 1, 1, 1, 1))
    psi = (m.C[:, 0] < 0.3).astype(float)
    A = (1.0 / n) ** 2
    phii = np.full(m.nIF, 1.0 * A)
    phib = np.array([-A, A])                                  # inflow at xMin, outflow at xMax
    bc = np.array([FV, ZG, ZG, ZG, ZG, ZG], np.int32)
    dt = 0.5 / n
    cfg = orc.mules_cfg(3)
    z = np.zeros(m.nCells)
    for _ in range(20):
        psib = np.array([1.0, psi[-1]])
        ho = _ho_flux(m, phii, psi)
        hob = phib * psib
        psi, psib, _, _ = orc.mules_explicit_solve(m, cfg, dt, psi, psib, psi, bc, phii, phib, ho, hob, z, z)
        assert psi.min() >= -1e-14 and psi.max() <= 1 + 1e-14
        assert np.all(np.diff(psi) <= 1e-14)


def test_k6_boundary_extrema_switch():
    """Appendix B-2: non-fixed-value wall patches contribute global bounds (default) or the patch value."""
    m = mo.box_mesh(6, 6, 1, (1.0, 1.0, 1.0 / 6), (0, 0, 0, 0, 1, 1))
    psi = np.full(m.nCells, 0.5)
    psib = psi[m.faceCells]
    phii, phib = _rot_flux(m)
    rng = np.random.default_rng(0)
    corr = 1e-3 * rng.standard_normal(m.nIF)
    bdi, bdb = orc.upwind_flux(m, phii, phib, psi, psib)
    z = np.zeros(m.nCells)
    bc = np.zeros(m.nPatches, np.int32)
    lg, _ = orc.mules_limiter(m, orc.mules_cfg(3, boundaryExtrema=orc.BEXT_GLOBAL), 1.0, psi, psib, psi, bc, bdi, bdb, corr, np.zeros(m.nBF), z, z)
    lp, _ = orc.mules_limiter(m, orc.mules_cfg(3, boundaryExtrema=orc.BEXT_PATCHVALUE), 1.0, psi, psib, psi, bc, bdi, bdb, corr, np.zeros(m.nBF), z, z)
    wall = np.zeros(m.nCells, bool); wall[m.faceCells] = True
    inner_faces = ~wall[m.owner] & ~wall[m.neighbour]
    assert (lp <= lg + 1e-16).all()                           # patch-value extrema are the tighter ones
    assert (lp[~inner_faces] < lg[~inner_faces]).any()


# ---- K7: limitSum hand vectors ------------------------------------------------------------------
def test_k7_limit_sum_hand_vectors():
    m = mo.box_mesh(4, 1, 1, (1.0, 0.25, 0.25), (0, 0, 1, 1, 1, 1))       # 3 internal faces
    zb = [np.zeros(m.nBF)] * 3
    c = [np.array([1.0, 2.0, 1e-310]), np.array([2.0, -1.0, 0.0]), np.array([3.0, -0.5, -1e-320])]
    out, _ = orc.mules_limit_sum(m, c, zb)
    # face 0: all positive, sumNeg = 0 -> positives scaled by 0
    assert [o[0] for o in out] == [0.0, 0.0, 0.0]
    # face 1: sumPos 2, sumNeg -1.5, sum > 0 -> positives scaled by 0.75
    assert [o[1] for o in out] == [1.5, -1.0, -0.5]
    # face 2: magnitudes below VSMALL are left alone
    assert [o[2] for o in out] == [1e-310, 0.0, -1e-320]
    c = [np.array([1.0, 0.0, 0.0]), np.array([-3.0, 0.0, 0.0]), np.array([-1.0, 0.0, 0.0])]
    out, _ = orc.mules_limit_sum(m, c, zb)
    assert [o[0] for o in out] == [1.0, -0.75, -0.25]          # sum < 0: negatives scaled by -sumPos/sumNeg
    assert sum(o[0] for o in out) == 0.0


def test_flux_of_vector_uniform_and_linear():
    m = mo.box_mesh(5, 4, 3, (1.0, 1.0, 1.0), (0,) * 6, (2.0, None, 0.5))
    F0 = np.array([0.3, -1.2, 0.7])
    fi, fb = orc.flux_of_vector(m, np.tile(F0, (m.nCells, 1)), np.tile(F0, (m.nBF, 1)))
    assert np.allclose(fi, m.Sf @ F0, rtol=1e-15, atol=1e-18) and np.allclose(fb, m.Sfb @ F0, rtol=1e-15, atol=1e-18)
    A = np.array([[1.0, 2.0, 0.0], [0.0, -1.0, 3.0], [0.5, 0.0, 1.0]])
    fi, _ = orc.flux_of_vector(m, m.C @ A.T, m.Cfb @ A.T)
    assert np.allclose(fi, np.einsum("ij,ij->i", m.Cf @ A.T, m.Sf), rtol=1e-13, atol=1e-16)   # linear weights: exact


# ---- hand-derived known answers of MULES::limiter on a row of three unit cells ----------------------------------------
# [UPSTREAM] OpenFOAM-v2106 src/finiteVolume/fvMatrices/solvers/MULES/MULESTemplates.C, MULES::limiter(): the branches
# SURVEY.md 8(a) M4 / A.6 flags -- (i) the local extrema are taken over the NEIGHBOURS (the cell's own value is not
# among them), (ii) on non-coupled patches that do not fix the value the extrema take the global bounds psiMax/psiMin
# (`boundaryExtrema` = global; the patch-value alternative is the run-time switch), (iii) the denominators carry
# ROOTVSMALL.  The source is not on this machine (parity unpinned); what IS pinned here is that the oracle does
# exactly what that restatement says, by numbers worked out by hand below -- so a reader with the v2106 source can check
# the three branches against nine literals.
def _row3():
    m = mo.box_mesh(3, 1, 1, (3.0, 1.0, 1.0), (0, 0, 1, 1, 1, 1))      # x walls, y/z empty: V = 1, |Sf| = 1
    assert m.nIF == 2 and m.nBF == 2 and np.allclose(m.V, 1.0)
    return m


def test_k6_limiter_hand_derived_row_global_and_patch_value_branches():
    """psi = (0.9, 0.5, 0.1), phi = +0.5 on both faces, HO flux (1.0, 0.6), dt = 1, Su = Sp = 0, walls zeroGradient.
         phiBD   = phi*psi_upwind          = (0.45, 0.25)       phiCorr = (0.55, 0.35)   (both > 0)
         sumPhiBD (own +, nei -)           = (0.45, -0.20, -0.25)
         sumPhip  (outflow corrections)    = (0.55, 0.35, 0)    mSumPhim (inflow) = (0, 0.55, 0.35)
       global bounds on the walls: psiMaxn = 1, psiMinn = 0 in every cell (cell 1: neighbours 0.9 / 0.1 clipped by nothing)
         Q+ = psiMaxn - psi0 + sumPhiBD    = (0.55, 0.30, 0.65)  Q- = psi0 - psiMinn - sumPhiBD = (0.45, 0.60, 0.35)   [cell 1: 0.5-0.1+0.2]
       iteration 1 (lambda = 1):  lambdap_0 = (mSumlPhim_0 + Q-_0)/sumPhip_0 = 0.45/0.55 = 9/11, every other lambda+- >= 1
         face 0: min(lambdap_0, lambdam_1) = 9/11;  face 1: 1.   Further iterations leave them (lambdap_0 stays 9/11).
       explicitSolve: phiPsi = phiBD + lambda phiCorr = (0.45 + 0.45, 0.6);  psi_new = psi0 - div = (0, 0.8, 0.7):
       cell 0 lands EXACTLY on psiMin -- the limiter's definition of "as much anti-diffusion as boundedness allows"."""
    m = _row3()
    psi = np.array([0.9, 0.5, 0.1]); psib = psi[m.faceCells]
    phii, phib = np.array([0.5, 0.5]), np.zeros(2)
    ho = np.array([1.0, 0.6])
    bc = np.zeros(m.nPatches, np.int32)
    z = np.zeros(3)
    bdi, bdb = orc.upwind_flux(m, phii, phib, psi, psib)
    assert np.allclose(bdi, [0.45, 0.25], rtol=0, atol=1e-16)
    lam, _ = orc.mules_limiter(m, orc.mules_cfg(3), 1.0, psi, psib, psi, bc, bdi, bdb, ho - bdi, np.zeros(2), z, z)
    assert np.allclose(lam, [9.0 / 11.0, 1.0], rtol=0, atol=2e-16), lam
    qi, _, pi, _ = orc.mules_explicit_solve(m, orc.mules_cfg(3), 1.0, psi, psib, psi, bc, phii, phib, ho - bdi + bdi, np.zeros(2), z, z)
    assert np.allclose(pi, [0.9, 0.6], rtol=0, atol=2e-16) and np.allclose(qi, [0.0, 0.8, 0.7], rtol=0, atol=3e-16), (pi, qi)
    # patch-value branch: the wall of cell 0 contributes psi_b = 0.9 instead of the global bounds -> psiMinn_0 = min(0.5, 0.9)
    #   Q-_0 = 0.9 - 0.5 - 0.45 = -0.05 < 0  ->  lambdap_0 = clamp(-0.05/0.55) = 0: face 0 gets NO anti-diffusion
    #   psi_new_0 = 0.9 - 0.45 = 0.45;  cell 1: 0.5 + 0.45 - 0.6 = 0.35   [face 1: psiMaxn_2 = max(0.5, psi_b = 0.1) = 0.5,
    #   Q+_2 = 0.5 - 0.1 - 0.25 = 0.15, lambdam_2 = 0.15/0.35 = 3/7 -> phiPsi_1 = 0.25 + 0.15 = 0.4 -> psi_new = (0.45, 0.55, 0.5)]
    cfgp = orc.mules_cfg(3, boundaryExtrema=orc.BEXT_PATCHVALUE)
    lam2, _ = orc.mules_limiter(m, cfgp, 1.0, psi, psib, psi, bc, bdi, bdb, ho - bdi, np.zeros(2), z, z)
    assert np.allclose(lam2, [0.0, 3.0 / 7.0], rtol=0, atol=2e-16), lam2
    q2, _, p2, _ = orc.mules_explicit_solve(m, cfgp, 1.0, psi, psib, psi, bc, phii, phib, ho, np.zeros(2), z, z)
    assert np.allclose(p2, [0.45, 0.4], rtol=0, atol=2e-16) and np.allclose(q2, [0.45, 0.55, 0.5], rtol=0, atol=3e-16), (p2, q2)


def test_k6_limiter_hand_derived_own_value_is_not_a_local_extremum():
    """psi = (0.2, 0.9, 0.3): cell 1 is a local MAXIMUM.  phi = +0.5, HO flux (0.3, 0.75), dt = 1.
         phiBD = (0.10, 0.45), phiCorr = (0.20, 0.30);  cell 1: sumPhiBD = -0.10 + 0.45 = 0.35,
         inflow correction mSumPhim_1 = 0.20 (face 0), outflow sumPhip_1 = 0.30 (face 1).
       Neighbours only: psiMaxn_1 = max(0.2, 0.3) = 0.3  ->  Q+_1 = 0.3 - 0.9 + 0.35 = -0.25
         iteration 1: lambdam_1 = clamp((sumlPhip_1 + Q+_1)/mSumPhim_1) = (0.30 - 0.25)/0.20 = 0.25  -> face 0: 0.25
       (had the own value 0.9 been included: Q+_1 = +0.35, lambdam_1 = 1, face 0 unlimited).
         face 1: lambdap_1 = (mSumlPhim_1 + Q-_1)/sumPhip_1 with psiMinn_1 = 0.2: Q-_1 = 0.9 - 0.2 - 0.35 = 0.35 -> (0.2+0.35)/0.3 > 1;
                 lambdam_2: psiMaxn_2 = 1 (wall, global), Q+_2 = 1 - 0.3 - 0.45 = 0.25 -> 0.25/0.30 = 5/6  -> face 1: 5/6
       iteration 2 (lambda = (0.25, 5/6)): sumlPhip_1 = 5/6*0.3 = 0.25 -> lambdam_1 = (0.25 - 0.25)/0.2 = 0 -> face 0: 0;
                 mSumlPhim_1 = 0.25*0.2 = 0.05 -> lambdap_1 = (0.05 + 0.35)/0.3 > 1; lambdam_2 unchanged -> face 1: 5/6
       iteration 3 changes nothing: lambda = (0, 5/6)."""
    m = _row3()
    psi = np.array([0.2, 0.9, 0.3]); psib = psi[m.faceCells]
    phii, phib = np.array([0.5, 0.5]), np.zeros(2)
    ho = np.array([0.3, 0.75])
    bc = np.zeros(m.nPatches, np.int32)
    z = np.zeros(3)
    bdi, bdb = orc.upwind_flux(m, phii, phib, psi, psib)
    for L, expect in ((1, [0.25, 5.0 / 6.0]), (2, [0.0, 5.0 / 6.0]), (3, [0.0, 5.0 / 6.0])):
        lam, _ = orc.mules_limiter(m, orc.mules_cfg(L), 1.0, psi, psib, psi, bc, bdi, bdb, ho - bdi, np.zeros(2), z, z)
        assert np.allclose(lam, expect, rtol=0, atol=2e-15), (L, lam)      # a few roundings in Q and the quotient


def test_k6_limiter_rootvsmall_keeps_cells_without_corrections_unlimited():
    """A cell with no outflow (inflow) correction has sumPhip (mSumPhim) = 0: the quotient is Q/ROOTVSMALL = +-1e150*Q,
    clamped to 1 (Q > 0) or 0 (Q < 0) -- never NaN.  psi = (0.3, 0.5, 0.7), phiBD = (0.15, 0.25), one small correction
    (+0.05) on face 0: Q-_0 = 0.3 - 0.15 = 0.15 -> lambdap_0 = 0.15/0.05 -> 1; Q+_1 = 0.7 - 0.5 + 0.1 = 0.3 -> lambdam_1 -> 1;
    face 1 (phiCorr = 0, not > 0) takes min(lambdam_1, lambdap_2) with lambdap_2 = (0 + 0.95)/(0 + ROOTVSMALL) -> 1."""
    m = _row3()
    psi = np.array([0.3, 0.5, 0.7]); psib = psi[m.faceCells]
    bc = np.zeros(m.nPatches, np.int32)
    z = np.zeros(3)
    bdi, bdb = np.array([0.15, 0.25]), np.zeros(2)
    lam, _ = orc.mules_limiter(m, orc.mules_cfg(3), 1.0, psi, psib, psi, bc, bdi, bdb, np.array([0.05, 0.0]), np.zeros(2), z, z)
    assert np.isfinite(lam).all() and np.array_equal(lam, [1.0, 1.0]), lam
