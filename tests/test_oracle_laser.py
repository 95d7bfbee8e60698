"""CPU known-answer tests of the laserDTRM restatement in the oracle: seeding (laserDTRM.C:169-331),
the cell walk with Beer-Lambert absorption (DTRMParticle.C:94-266), reflecting-cell marking (laserDTRM.C:596-652) and
Fresnel reflection.  The expected values are computed here with plain numpy/Python on cases that are one-dimensional
or analytic, independently of the C walk."""
import numpy as np
import pytest

from oracle import oracle as orc, mesh_oracle as mo
from thermocapillaryinterfoam_b200 import cases, mesh as hm


def _box(n, L=(1.0, 1.0, 1.0), grading=None):
    g = grading or (None, None, None)
    m = mo.box_mesh(*n, L, (0,) * 6, grading)
    xs, ys, zs = (hm.vertices(n[d], L[d], g[d]) for d in range(3))
    return m, (xs, ys, zs), cases.box_polymesh(m, xs, ys, zs)


def test_polymesh_faces_match_the_mesh_geometry():
    m, v, geom = _box((5, 4, 3), grading=(1.7, None, 0.6))
    P, fp = geom["points"], geom["fpPts"].reshape(-1, 4)
    Cf = np.concatenate([m.Cf, m.Cfb]); Sf = np.concatenate([m.Sf, m.Sfb])
    assert np.allclose(P[fp].mean(axis=1), Cf, atol=1e-14)
    S = 0.5 * np.cross(P[fp[:, 2]] - P[fp[:, 0]], P[fp[:, 3]] - P[fp[:, 1]])
    assert np.allclose(S, Sf, atol=1e-15)        # right-handed about the area vector (owner -> neighbour / outward)


@pytest.mark.parametrize("mode", ["Gaussian", "uniform", "GaussianPeak"])
def test_seeds(mode):
    kw = dict(mode=mode, power=250.0, sigma=0.04, I0=3.0e4)
    d = np.array([0.2, -1.0, 0.1]); d /= np.linalg.norm(d)
    r = np.cross(d, [0, 0, 1.0]); r /= np.linalg.norm(r)
    a = cases.laser_seeds((0.5, 0.9, 0.5), d, r, 7, 16, 0.1, 1.7, **kw)
    b = orc.laser_seed((0.5, 0.9, 0.5), d, r, 7, 16, 0.1, 1.7, **kw)
    for x, y in zip(a, b):
        assert np.allclose(x, y, rtol=1e-13, atol=1e-15)
    p0, p1, Ip, dA = b
    assert np.allclose((p0 - [0.5, 0.9, 0.5]) @ d, 0, atol=1e-15)            # seeds lie on the disk normal to lDir
    assert np.allclose(p1 - p0, 0.5 * 1.7 * d, atol=1e-14)
    assert np.isclose(dA.sum(), np.pi * 0.1 ** 2, rtol=1e-12)
    if mode == "uniform":
        assert np.isclose((Ip * dA).sum(), 250.0, rtol=1e-12)


def test_find_cell_matches_the_structured_lookup():
    m, (xs, ys, zs), geom = _box((6, 5, 4), grading=(1.5, None, None))
    rng = np.random.default_rng(3)
    pts = rng.uniform(-0.1, 1.1, (60, 3))
    assert np.array_equal(orc.find_cells(m, geom, pts), cases.find_cells_box((6, 5, 4), xs, ys, zs, pts))


def test_beer_lambert_along_a_column():
    """One vertical ray through a uniform absorbing, emitting medium: per cell crossed I1 = (I + ds*(e*sigma*T^4/pi + E))
    /(1 + ds*a) and Q*V += (I - max(I1, 0))*dA until I <= 1% (DTRMParticle.C:238-262)."""
    n = (3, 10, 3)
    m, (xs, ys, zs), geom = _box(n, grading=(None, 1.4, None))
    nC = m.nCells
    a, e, E, T = np.full(nC, 2.5), np.full(nC, 0.3), np.full(nC, 0.2), np.full(nC, 900.0)
    al = [np.ones(nC), np.zeros(nC)]; ab = [np.ones(m.nBF), np.zeros(m.nBF)]
    p0 = np.array([[0.5, 0.97, 0.5]]); p1 = p0 + [[0, -geom["boundsMag"], 0]]
    cell = cases.find_cells_box(n, xs, ys, zs, p0)
    I0, dA = 5.0e3, 1e-4
    r = orc.laser_calculate(m, geom, [], al, ab, np.zeros((2, 6), np.int32), a, e, E, T, p0, p1, [I0], [dA], cell)
    assert r["nParticles"] == 1
    emit = 0.3 * orc.SIGMA_SB * 900.0 ** 4 / np.pi + 0.2
    I, y, QV = I0, 0.97, np.zeros(nC)
    for j in range(n[1] - 1, -1, -1):
        ds = y - ys[j]; y = ys[j]
        I1 = (I + ds * emit) / (1 + ds * 2.5)
        QV[1 + 3 * (j + 10 * 1)] += (I - max(I1, 0.0)) * dA
        I = I1
        if I <= 0.01 * I0:
            break
    assert np.allclose(r["Q"] * m.V, QV, rtol=1e-12, atol=1e-18 * I0)
    assert (r["Q"] != 0).sum() == (QV != 0).sum()


def test_oblique_ray_path_lengths_sum_to_the_chord():
    """a = 0 with a uniform source E: Q*V in a cell = -(I1 - I)*dA = -ds*E*dA, so sum(Q*V)/(-E*dA) is the chord
    length of the ray inside the box (independent of the cells it crosses)."""
    n = (7, 6, 5)
    m, (xs, ys, zs), geom = _box(n, grading=(1.3, 0.8, None))
    nC = m.nCells
    al = [np.ones(nC)]; ab = [np.ones(m.nBF)]
    p0 = np.array([[0.13, 0.91, 0.22]]); d = np.array([0.55, -0.7, 0.45]); d /= np.linalg.norm(d)
    p1 = p0 + geom["boundsMag"] * d
    cell = cases.find_cells_box(n, xs, ys, zs, p0)
    r = orc.laser_calculate(m, geom, [], al, ab, np.zeros((1, 6), np.int32), np.zeros(nC), np.zeros(nC), np.full(nC, 2.0),
                            np.full(nC, 300.0), p0, p1, [1.0], [1.0], cell)
    t_exit = min(((1.0 if d[k] > 0 else 0.0) - p0[0, k]) / d[k] for k in range(3))
    assert np.isclose(-(r["Q"] * m.V).sum() / 2.0, t_exit, rtol=1e-12)
    # every cell on the segment got its share and no other
    ts = np.linspace(0, t_exit, 4001)[1:-1]
    on = np.unique(cases.find_cells_box(n, xs, ys, zs, p0 + ts[:, None] * d))
    assert set(np.nonzero(r["Q"])[0]) <= set(on) | {int(cell[0])}
    assert len(np.nonzero(r["Q"])[0]) >= len(on) - 2


def _layered(n=(4, 16, 4), y0=0.5, ny_w=1.5):
    m, v, geom = _box(n)
    w = ny_w / n[1]
    metal = lambda Y: 0.5 * (1.0 - np.tanh((Y - y0) / w))
    al = [metal(m.C[:, 1]), 1.0 - metal(m.C[:, 1])]
    ab = [metal(m.Cfb[:, 1]), 1.0 - metal(m.Cfb[:, 1])]
    return m, v, geom, al, ab


def test_reflecting_cells_of_a_flat_interface():
    """laserDTRM.C:596-652: marked where alpha_to >= alphaCut, |nHatPhase| > 0.99 and grad(from).grad(to) < 0;
    nHat there is the unit normal pointing from the metal (to) into the gas (from)."""
    m, v, geom, al, ab = _layered()
    nC = m.nCells
    models = [dict(kind="FresnelLaser", **{"from": 1, "to": 0}, epsilon=0.3)]
    z = np.zeros(nC)
    r = orc.laser_calculate(m, geom, models, al, ab, np.ones((2, 6), np.int32), z, z, z, z + 300, np.zeros((0, 3)), np.zeros((0, 3)),
                            [], [], [])
    mk = r["reflectingCells"]
    assert set(np.unique(mk)) == {-1, 0}
    assert (al[0][mk == 0] >= 0.5).all()
    nh = r["nHat"][mk == 0]
    assert np.all(nh[:, [0, 2]] == 0) and np.all(nh[:, 1] > 0.99) and np.all(nh[:, 1] <= 1.0)
    assert np.all(r["nHat"][mk < 0] == 0)
    # the gas side of alphaCut is never marked; the metal side is, as far as grad(alpha) stays above deltaN
    assert (mk[al[0] < 0.5] == -1).all()
    # points on the boundary patches carry the (zero) patch value, interior points the weighted cell average
    P = geom["points"]
    onb = (np.abs(P - 0.5).max(axis=1) > 0.5 - 1e-12)
    assert np.all(r["nHatPoints"][onb] == 0)
    assert np.abs(r["nHatPoints"][~onb]).max() > 0.4


def test_normal_incidence_reflection_known_answer():
    """Vertical ray onto a flat pool: column walk in Python following DTRMParticle.C:94-266 with the normal (0,1,0).
    The reflected particle starts 0.1*cbrt(V) above the face where the first reflecting cell was left, goes straight
    up with I*rho, and (transmissiveId == -1) treats its first cell as reflecting again with cosTheta < 0."""
    n = (4, 16, 4)
    m, (xs, ys, zs), geom, al, ab = _layered(n)
    nC = m.nCells
    eps = 0.3
    models = [dict(kind="FresnelLaser", **{"from": 1, "to": 0}, epsilon=eps)]
    aC = [40.0, 0.5]
    a = orc.local_density(al, aC); z = np.zeros(nC)
    x0, z0 = 0.6, 0.4          # inside cell column (2, :, 1), away from the cell edges
    p0 = np.array([[x0, 0.99, z0]]); p1 = p0 + [[0, -geom["boundsMag"], 0]]
    cell = cases.find_cells_box(n, xs, ys, zs, p0)
    I0, dA = 1.0e4, 1e-5
    r = orc.laser_calculate(m, geom, models, al, ab, np.ones((2, 6), np.int32), a, z, z, z + 300, p0, p1, [I0], [dA], cell)
    mk = r["reflectingCells"]
    col = lambda j: 2 + 4 * (j + 16 * 1)
    QV = np.zeros(nC)
    rho1 = 0.5 * ((1 + (1 - eps) ** 2) / (1 + (1 + eps) ** 2) + (eps ** 2 - 2 * eps + 2) / (eps ** 2 + 2 * eps + 2))
    rho1 = min(max(rho1, 0.0), 0.98)
    I, y, tid, refl = I0, 0.99, -1, None
    for j in range(15, -1, -1):
        c = col(j); ds = y - ys[j]; y = ys[j]
        zone = mk[c]
        rho = 0.0
        if zone > -1 and tid != zone:
            if I > 0.01 * I0:
                rho = rho1
                refl = (j, I * rho, y + 0.1 * np.cbrt(m.V[c]))
            tid = zone
        Itr = I * (1 - rho)
        I1 = Itr / (1 + ds * a[c])
        QV[c] += (Itr - max(I1, 0.0)) * dA
        I = I1
        if I <= 0.01 * I0:
            break
    assert refl is not None
    j, I, y = refl
    Ir0 = I
    for jj in range(j, 16):
        c = col(jj); ds = ys[jj + 1] - y; y = ys[jj + 1]
        I1 = I / (1 + ds * a[c])          # first cell: reflecting zone again, cosTheta < 0 -> rho = 0
        QV[c] += (I - max(I1, 0.0)) * dA
        I = I1
        if I <= 0.01 * Ir0:
            break
    assert r["nParticles"] == 2
    assert np.allclose(r["Q"] * m.V, QV, rtol=1e-9, atol=1e-16 * I0)


def test_linearity_and_mirror_symmetry_with_reflection():
    """Q is linear in the seed intensities, and a mirrored pair of oblique rays over mirror-symmetric fields
    deposits mirror images."""
    n = (8, 12, 8)
    m, (xs, ys, zs), geom, al, ab = _layered(n, ny_w=1.2)
    nC = m.nCells
    models = [dict(kind="Fresnel", **{"from": 1, "to": 0}, nk1=(1.0, 0.0), nk2=(3.2, 4.1))]
    a = orc.local_density(al, [15.0, 0.2]); z = np.zeros(nC)
    d1 = np.array([0.5, -1.0, 0.0]); d1 /= np.linalg.norm(d1)
    d2 = d1 * [-1, 1, 1]
    p0 = np.array([[0.3, 0.97, 0.45], [0.7, 0.97, 0.45]])
    p1 = p0 + geom["boundsMag"] * np.array([d1, d2])
    cell = cases.find_cells_box(n, xs, ys, zs, p0)
    run = lambda I: orc.laser_calculate(m, geom, models, al, ab, np.ones((2, 6), np.int32), a, z, z, z + 300, p0, p1, I,
                                        [1e-5, 1e-5], cell)
    r1, r2 = run([1.0e4, 1.0e4]), run([2.5e4, 2.5e4])
    assert r1["nParticles"] >= 4                      # both rays reflected at least once
    assert np.allclose(r2["Q"], 2.5 * r1["Q"], rtol=1e-12, atol=1e-18)
    Q = r1["Q"].reshape(n[2], n[1], n[0])
    assert np.allclose(Q, Q[:, :, ::-1], rtol=1e-9, atol=1e-12 * np.abs(Q).max())
    assert (r1["Q"] >= 0).all() and r1["Q"].max() > 0
    assert (r1["Q"] * m.V).sum() <= 2 * 1.0e4 * 1e-5 * (1 + 1e-12)


def test_cloud_overflow_is_reported():
    n = (4, 16, 4)
    m, (xs, ys, zs), geom, al, ab = _layered(n)
    models = [dict(kind="FresnelLaser", **{"from": 1, "to": 0}, epsilon=0.3)]
    z = np.zeros(m.nCells)
    p0 = np.array([[0.6, 0.99, 0.4]]); p1 = p0 + [[0, -2.0, 0]]
    cell = cases.find_cells_box(n, xs, ys, zs, p0)
    with pytest.raises(RuntimeError):
        orc.laser_calculate(m, geom, models, al, ab, np.ones((2, 6), np.int32), z + 1.0, z, z, z + 300, p0, p1, [1.0], [1.0],
                            cell, maxParticles=1)


@pytest.mark.parametrize("procs", [(2, 1, 1), (2, 2, 1), (2, 2, 2)])
def test_block_polymesh_and_shared_points_of_processor_sub_meshes(procs):
    """Host side of the decomposed laser walk: every sub-mesh's points/faces reproduce its own geometry, the points of
    processor patches carry global labels that at least two ranks hold, and all other points are private."""
    spec = cases.case(5, 8)
    xs, ys, zs = (hm.vertices(spec.n[d], spec.lengths[d], g) for d, g in enumerate((1.3, None, None)))
    gm = mo.box_mesh(*spec.n, spec.lengths, spec.patch_types, (1.3, None, None))
    geomG = cases.box_polymesh(gm, xs, ys, zs)
    nd = procs[0] * procs[1] * procs[2]
    holders = {}
    for rk in range(nd):
        m = hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, rk, grading=(1.3, None, None))
        geom = cases.box_block_polymesh(spec.n, xs, ys, zs, procs, rk, m)
        assert np.allclose(geom["points"], geomG["points"][geom["pointGlobal"]], atol=0)
        P, fp = geom["points"], geom["fpPts"].reshape(-1, 4)
        Cf = np.concatenate([m.Cf, m.Cfb]); Sf = np.concatenate([m.Sf, m.Sfb])
        assert np.allclose(P[fp].mean(axis=1), Cf, atol=1e-14)
        assert np.allclose(0.5 * np.cross(P[fp[:, 2]] - P[fp[:, 0]], P[fp[:, 3]] - P[fp[:, 1]]), Sf, atol=1e-15)
        assert geom["boundsMag"] == geomG["boundsMag"]
        assert np.array_equal(geom["pointGlobal"][geom["sharedLocal"]], geom["sharedGlobal"])
        for gpt in geom["pointGlobal"]:
            holders.setdefault(int(gpt), set()).add(rk)
        geom_shared = set(int(x) for x in geom["sharedGlobal"])
        for gpt in geom["pointGlobal"]:
            holders.setdefault(("shared", int(gpt)), set())
            if int(gpt) in geom_shared:
                holders[("shared", int(gpt))].add(rk)
    assert len([k for k in holders if not isinstance(k, tuple)]) == len(geomG["points"])
    for k, ranks in holders.items():
        if isinstance(k, tuple):
            continue
        # a point held by several ranks is declared shared by each of them, a private point by none
        assert holders[("shared", k)] == (ranks if len(ranks) > 1 else set()), (k, ranks)


@pytest.mark.parametrize("procs", [(2, 1, 1), (1, 2, 1), (2, 2, 2)])
def test_decomposed_walk_on_cpu_matches_the_serial_walk(procs):
    """N > 1 on the CPU: all ranks in lock-step rounds (orc_laser_calculate_ranks) - hand-over at processor patches,
    point values summed over the ranks sharing a point, reflected particles located across ranks - against the serial
    walk on the undecomposed mesh.  The same algorithm the device runs over NCCL (tests/test_gpu_laser.py)."""
    n = 16
    spec = cases.case(5, n)
    xs, ys, zs = (hm.vertices(spec.n[d], spec.lengths[d]) for d in range(3))
    g = mo.box_mesh(*spec.n, spec.lengths, spec.patch_types)
    geomG = cases.box_polymesh(g, xs, ys, zs)
    fg = cases.synthetic_fields(spec, g)
    models = [dict(kind="FresnelLaser", **{"from": 2, "to": 0}, epsilon=0.25),
              dict(kind="Fresnel", **{"from": 2, "to": 1}, nk1=(1.0, 0.0), nk2=(3.6, 5.0))]
    d = np.array([0.5, -1.0, 0.2]); d /= np.linalg.norm(d)
    r = np.cross(d, [0.0, 0.0, 1.0]); r /= np.linalg.norm(r)
    p0, p1, Ip, dA = cases.laser_seeds((0.3, 0.93, 0.45), d, r, 8, 16, 0.15, geomG["boundsMag"], sigma=0.07, power=200.0)
    cellG = cases.find_cells_box(spec.n, xs, ys, zs, p0)
    al = fg["alpha_i"]
    a, e, E = (orc.local_density(al, c) for c in ((25.0, 10.0, 0.3), (0.1, 0.0, 0.0), (0.0, 0.0, 0.0)))
    ref = orc.laser_calculate(g, geomG, models, al, fg["alpha_b"], fg["alphaBc"], a, e, E, fg["T_i"], p0, p1, Ip, dA, cellG)
    nd = procs[0] * procs[1] * procs[2]
    meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, rk) for rk in range(nd)]
    geoms = [cases.box_block_polymesh(spec.n, xs, ys, zs, procs, rk, m) for rk, m in enumerate(meshes)]
    gids = [np.asarray(m.cellGlobal) for m in meshes]
    cells = []
    for gid in gids:
        g2l = np.full(g.nCells, -1, np.int32); g2l[gid] = np.arange(len(gid), dtype=np.int32)
        cells.append(g2l[cellG])
    out = orc.laser_calculate_ranks(meshes, geoms, models, [ref["reflectingCells"][i] for i in gids], [ref["nHat"][i] for i in gids],
                                    [gm["pointGlobal"] for gm in geoms], [a[i] for i in gids], [e[i] for i in gids],
                                    [E[i] for i in gids], [fg["T_i"][i] for i in gids], p0, p1, Ip, dA, cells)
    assert int(out["created"].sum()) == ref["nParticles"] > len(Ip)
    assert sum(1 for q in out["Q"] if q.any()) >= 2           # the beam crosses the processor planes
    scale = np.abs(ref["Q"]).max()
    for rk in range(nd):
        assert np.abs(out["nHatPoints"][rk] - ref["nHatPoints"][geoms[rk]["pointGlobal"]]).max() <= 1e-12
        assert np.abs(out["Q"][rk] - ref["Q"][gids[rk]]).max() <= 1e-10 * scale, rk
    assert sum((q != 0).sum() for q in out["Q"]) == (ref["Q"] != 0).sum()
