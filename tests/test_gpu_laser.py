"""GPU parity of laserDTRM through the C ABI against the CPU oracle: reflecting cells and nHat
(bit-identical: same operations, fmad off), the point interpolation, the cell walk with absorption and Fresnel
reflection.  Q is summed with fp64 atomics on the device, in seed order on the CPU: tolerance 1e-11 of max|Q|;
sin/cos/tan/cbrt of the Fresnel model and of the insertion offset differ by ulps between libm and CUDA."""
import numpy as np
import pytest

from oracle import oracle as orc, mesh_oracle as mo
from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm

pytestmark = pytest.mark.gpu

MODELS = [dict(kind="FresnelLaser", **{"from": 2, "to": 0}, epsilon=0.25),
          dict(kind="Fresnel", **{"from": 2, "to": 1}, nk1=(1.0, 0.0), nk2=(3.6, 5.0))]


def _setup(n=16, grading=(1.4, None, 0.8), config=5):
    spec = cases.case(config, n)
    g = grading or (None, None, None)
    m = mo.box_mesh(*spec.n, spec.lengths, spec.patch_types, grading)
    xs, ys, zs = (hm.vertices(spec.n[d], spec.lengths[d], g[d]) for d in range(3))
    geom = cases.box_polymesh(m, xs, ys, zs)
    f = cases.synthetic_fields(spec, m)
    ctx = abi.Context(m, spec.nPhases)
    cases.load_case(ctx, spec, f)
    fc = np.asarray(m.faceCells, dtype=np.int64)
    ctx.set_cell_centres(m.C, np.asarray(m.Cfb) - np.asarray(m.C)[fc])
    ctx.laser_set_polymesh(m.Cf, m.Cfb, geom)
    return spec, m, (xs, ys, zs), geom, f, ctx


def _seeds(spec, v, geom, pos, d, nr=8, nTheta=16, R=0.18):
    d = np.asarray(d, float); d /= np.linalg.norm(d)
    r = np.cross(d, [0.0, 0.0, 1.0]); r /= np.linalg.norm(r)
    p0, p1, Ip, dA = cases.laser_seeds(pos, d, r, nr, nTheta, R, geom["boundsMag"], mode="Gaussian", power=200.0, sigma=0.08)
    return p0, p1, Ip, dA, cases.find_cells_box(spec.n, *v, p0)


def _close(a, b, what, rtol=1e-11):
    scale = max(np.abs(a).max(), np.abs(b).max(), 1e-300)
    err = np.abs(a - b).max()
    assert err <= rtol * scale, f"{what}: max err {err:.3e} (scale {scale:.3e})"


@pytest.mark.parametrize("aC", [(60.0, 25.0, 0.3), (4.0, 2.0, 0.05)])
def test_calculate_matches_oracle(aC):
    spec, m, v, geom, f, ctx = _setup()
    eC, EC = (0.2, 0.1, 0.0), (0.0, 0.0, 0.0)
    ctx.laser_set_reflection(MODELS, alphaCut=0.5)
    ctx.laser_set_local_density_absorption([0, 1, 2], aC, eC, EC)
    p0, p1, Ip, dA, cell = _seeds(spec, v, geom, (0.5, 0.93, 0.5), (0.25, -1.0, 0.1))
    assert (cell >= 0).all()
    Q, nPart = ctx.laser_calculate(p0, p1, Ip, dA, cell)
    got = ctx.laser_get()
    al = f["alpha_i"]
    ref = orc.laser_calculate(m, geom, MODELS, al, f["alpha_b"], f["alphaBc"], orc.local_density(al, aC), orc.local_density(al, eC),
                              orc.local_density(al, EC), f["T_i"], p0, p1, Ip, dA, cell)
    assert (ref["reflectingCells"] >= 0).sum() > 20 and len(np.unique(ref["reflectingCells"])) == 3
    assert np.array_equal(got["reflectingCells"], ref["reflectingCells"])
    _close(got["nHat"], ref["nHat"], "nHat", 1e-12)
    _close(got["nHatPoints"], ref["nHatPoints"], "nHatPoints", 1e-12)
    assert ref["nParticles"] > len(Ip), "the case must exercise reflection"
    assert nPart == ref["nParticles"]
    assert ctx.laser_generations >= 2
    assert np.array_equal(Q, got["Q"])
    cells, vals = ctx.laser_get_sparse(capacity=16)          # grows to what is needed
    assert len(cells) == (Q != 0).sum() > 16 and len(np.unique(cells)) == len(cells)
    Qs = np.zeros(m.nCells); Qs[cells] = vals
    assert np.array_equal(Qs, Q)
    _close(Q, ref["Q"], "Q")
    assert np.array_equal(Q != 0, ref["Q"] != 0)
    ctx.close()


def test_no_reflection_and_host_absorption_fields():
    spec, m, v, geom, f, ctx = _setup(n=12, grading=None)
    rng = np.random.default_rng(4)
    a, e, E = rng.uniform(0.5, 8.0, m.nCells), rng.uniform(0, 0.5, m.nCells), rng.uniform(0, 3.0, m.nCells)
    ctx.laser_set_reflection([])
    ctx.laser_set_absorption_fields(a, e, E)
    p0, p1, Ip, dA, cell = _seeds(spec, v, geom, (0.4, 0.5, 1.2), (0.3, 0.2, -1.0), nr=6, nTheta=12, R=0.3)
    assert (cell < 0).all()            # seeded outside the box: findCell misses, nothing is tracked
    Q, nPart = ctx.laser_calculate(p0, p1, Ip, dA, cell)
    assert nPart == 0 and not Q.any()
    p0, p1, Ip, dA, cell = _seeds(spec, v, geom, (0.4, 0.5, 0.97), (0.3, 0.2, -1.0), nr=6, nTheta=12, R=0.3)
    assert (cell >= 0).any()
    Q, nPart = ctx.laser_calculate(p0, p1, Ip, dA, cell)
    al = f["alpha_i"]
    ref = orc.laser_calculate(m, geom, [], al, f["alpha_b"], f["alphaBc"], a, e, E, f["T_i"], p0, p1, Ip, dA, cell)
    assert nPart == ref["nParticles"] == (cell >= 0).sum()
    _close(Q, ref["Q"], "Q")
    assert (ctx.laser_get()["reflectingCells"] == -1).all()
    ctx.close()


def test_noReflection_model_marks_but_does_not_spawn():
    spec, m, v, geom, f, ctx = _setup(n=12, grading=None)
    models = [dict(kind="noReflection", **{"from": 2, "to": 0})]
    ctx.laser_set_reflection(models, alphaCut=0.4)
    ctx.laser_set_local_density_absorption([0, 2], (30.0, 0.1), (0.0, 0.0), (0.0, 0.0))
    p0, p1, Ip, dA, cell = _seeds(spec, v, geom, (0.35, 0.93, 0.5), (0.0, -1.0, 0.0), nr=5, nTheta=8, R=0.1)
    Q, nPart = ctx.laser_calculate(p0, p1, Ip, dA, cell)
    al = f["alpha_i"]
    a = orc.local_density([al[0], al[2]], (30.0, 0.1)); z = np.zeros(m.nCells)
    ref = orc.laser_calculate(m, geom, models, al, f["alpha_b"], f["alphaBc"], a, z, z, f["T_i"], p0, p1, Ip, dA, cell, alphaCut=0.4)
    assert nPart == ref["nParticles"] == len(Ip)
    assert (ref["reflectingCells"] == 0).any()
    assert np.array_equal(ctx.laser_get()["reflectingCells"], ref["reflectingCells"])
    _close(Q, ref["Q"], "Q")
    ctx.close()


def test_energy_is_linear_in_the_seed_intensity_on_the_device():
    spec, m, v, geom, f, ctx = _setup(n=12, grading=None)
    ctx.laser_set_reflection(MODELS)
    ctx.laser_set_local_density_absorption([0, 1, 2], (20.0, 10.0, 0.2), (0, 0, 0), (0, 0, 0))
    p0, p1, Ip, dA, cell = _seeds(spec, v, geom, (0.5, 0.93, 0.5), (0.1, -1.0, 0.0))
    Q1, n1 = ctx.laser_calculate(p0, p1, Ip, dA, cell)
    Q2, n2 = ctx.laser_calculate(p0, p1, 4.0 * Ip, dA, cell)
    assert n1 == n2
    _close(Q2, 4.0 * Q1, "linearity")
    assert (Q1 >= 0).all() and (Q1 * m.V).sum() <= (Ip * dA).sum() * (1 + 1e-12)
    ctx.close()


def test_errors():
    spec = cases.case(5, 8)
    m = mo.box_mesh(*spec.n, spec.lengths, spec.patch_types)
    xs = ys = zs = hm.vertices(8, 1.0)
    geom = cases.box_polymesh(m, xs, ys, zs)
    f = cases.synthetic_fields(spec, m)
    ctx = abi.Context(m, spec.nPhases)
    cases.load_case(ctx, spec, f)
    one = (np.array([[0.5, 0.9, 0.5]]), np.array([[0.5, -1.0, 0.5]]), [1.0], [1.0], [0])
    with pytest.raises(RuntimeError):
        ctx.laser_calculate(*one)                           # no polymesh
    with pytest.raises(RuntimeError):
        ctx.laser_set_polymesh(m.Cf, m.Cfb, geom)           # no cell centres
    fc = np.asarray(m.faceCells, dtype=np.int64)
    ctx.set_cell_centres(m.C, np.asarray(m.Cfb) - np.asarray(m.C)[fc])
    ctx.laser_set_polymesh(m.Cf, m.Cfb, geom)
    with pytest.raises(RuntimeError):
        ctx.laser_calculate(*one)                           # no absorption model
    with pytest.raises(RuntimeError):
        ctx.laser_set_reflection([dict(kind="Fresnel", **{"from": 1, "to": 1})])
    ctx.laser_set_local_density_absorption([0, 1, 2], (5.0, 5.0, 0.5), (0, 0, 0), (0, 0, 0))
    ctx.laser_set_reflection(MODELS)
    p0, p1, Ip, dA, cell = _seeds(spec, (xs, ys, zs), geom, (0.5, 0.93, 0.5), (0.25, -1.0, 0.1))
    with pytest.raises(RuntimeError):
        ctx.laser_calculate(p0, p1, Ip, dA, cell, maxParticles=len(Ip))     # reflected particles do not fit
    Q, n = ctx.laser_calculate(p0, p1, Ip, dA, cell)        # the context is usable afterwards
    assert n > len(Ip) and Q.max() > 0
    ctx.close()


@pytest.mark.parametrize("procs", [(2, 1, 1), (1, 2, 1), (2, 2, 2)])
def test_decomposed_matches_serial_oracle(procs):
    """Processor patches: grad(alpha) with halos, point interpolation summed over the ranks sharing a point,
    particles handed to the neighbour rank when they reach a processor patch (in-process group, one thread per
    rank).  Against the SERIAL oracle on the undecomposed mesh: the point weights are re-associated (1e-10)."""
    import threading
    n = 16
    spec = cases.case(5, n)
    g = mo.box_mesh(*spec.n, spec.lengths, spec.patch_types)
    xs, ys, zs = (hm.vertices(spec.n[d], spec.lengths[d]) for d in range(3))
    geomG = cases.box_polymesh(g, xs, ys, zs)
    fg = cases.synthetic_fields(spec, g)
    aC, eC, EC = (25.0, 10.0, 0.3), (0.1, 0.0, 0.0), (0.0, 0.0, 0.0)
    d = np.array([0.5, -1.0, 0.2]); d /= np.linalg.norm(d)
    r = np.cross(d, [0.0, 0.0, 1.0]); r /= np.linalg.norm(r)
    p0, p1, Ip, dA = cases.laser_seeds((0.3, 0.93, 0.45), d, r, 8, 16, 0.15, geomG["boundsMag"], sigma=0.07, power=200.0)
    cellG = cases.find_cells_box(spec.n, xs, ys, zs, p0)
    assert (cellG >= 0).all()
    al = fg["alpha_i"]
    ref = orc.laser_calculate(g, geomG, MODELS, al, fg["alpha_b"], fg["alphaBc"], orc.local_density(al, aC),
                              orc.local_density(al, eC), orc.local_density(al, EC), fg["T_i"], p0, p1, Ip, dA, cellG)
    assert ref["nParticles"] > len(Ip)
    nd = procs[0] * procs[1] * procs[2]
    meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, rk) for rk in range(nd)]
    fs = [cases.synthetic_fields(spec, m) for m in meshes]
    grp = abi.Group(nd)
    ctxs = [abi.Context(m, spec.nPhases) for m in meshes]
    outs, errs = [None] * nd, []

    def work(rk):
        try:
            m, f, ctx = meshes[rk], fs[rk], ctxs[rk]
            ctx.comm_init_local(rk, grp)
            cases.load_case(ctx, spec, f)
            ctx.set_cell_centres(f["C"], f["deltab"])
            geom = cases.box_block_polymesh(spec.n, xs, ys, zs, procs, rk, m)
            ctx.laser_set_polymesh(m.Cf, m.Cfb, geom)
            ctx.laser_set_shared_points(geom["sharedLocal"], geom["sharedGlobal"])
            ctx.laser_set_reflection(MODELS)
            ctx.laser_set_local_density_absorption([0, 1, 2], aC, eC, EC)
            gid = np.asarray(m.cellGlobal)
            g2l = np.full(g.nCells, -1, np.int32); g2l[gid] = np.arange(len(gid), dtype=np.int32)
            Q, created = ctx.laser_calculate(p0, p1, Ip, dA, g2l[cellG])
            got = ctx.laser_get()
            outs[rk] = dict(Q=Q, created=created, gid=gid, refl=got["reflectingCells"], nHat=got["nHat"],
                            nPts=got["nHatPoints"], pgid=geom["pointGlobal"])
        except Exception as e:          # noqa: BLE001
            errs.append((rk, e))
    ts = [threading.Thread(target=work, args=(rk,), daemon=True) for rk in range(nd)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(180)
    assert not errs, errs
    assert all(o is not None for o in outs), "a rank did not finish"
    assert sum(o["created"] for o in outs) == ref["nParticles"]
    scale = np.abs(ref["Q"]).max()
    for rk, o in enumerate(outs):
        assert np.array_equal(o["refl"], ref["reflectingCells"][o["gid"]]), f"reflectingCells rank {rk}"
        assert np.abs(o["nHat"] - ref["nHat"][o["gid"]]).max() <= 1e-12, f"nHat rank {rk}"
        assert np.abs(o["nPts"] - ref["nHatPoints"][o["pgid"]]).max() <= 1e-12, f"nHatPoints rank {rk}"
        err = np.abs(o["Q"] - ref["Q"][o["gid"]]).max()
        assert err <= 1e-10 * scale, f"Q rank {rk}: {err:.3e} of {scale:.3e}"
    assert sum((o["Q"] != 0).sum() for o in outs) == (ref["Q"] != 0).sum()
    for c in ctxs:
        c.close()
    grp.close()


def test_local_find_cell_reproduces_the_decomposed_reference_walk():
    """tcif_laser_set_local_find_cell(1): a reflected particle whose insertion point lies behind a processor patch is
    dropped, as the reference's rank-local mesh().findCell does (DTRMParticle.C:189-206).  The decomposed device run then
    equals the decomposed walk of the oracle with the same rule (counts exact, Q to 1e-10 of max), which is what a
    decomposed run of the reference computes -- and both lose particles against the serial walk here."""
    import threading
    n, procs = 16, (2, 2, 2)
    spec = cases.case(5, n)
    g = mo.box_mesh(*spec.n, spec.lengths, spec.patch_types)
    xs, ys, zs = (hm.vertices(spec.n[d], spec.lengths[d]) for d in range(3))
    geomG = cases.box_polymesh(g, xs, ys, zs)
    fg = cases.synthetic_fields(spec, g)
    aC, eC, EC = (25.0, 10.0, 0.3), (0.1, 0.0, 0.0), (0.0, 0.0, 0.0)
    d = np.array([0.5, -1.0, 0.2]); d /= np.linalg.norm(d)
    r = np.cross(d, [0.0, 0.0, 1.0]); r /= np.linalg.norm(r)
    p0, p1, Ip, dA = cases.laser_seeds((0.3, 0.93, 0.45), d, r, 8, 16, 0.15, geomG["boundsMag"], sigma=0.07, power=200.0)
    cellG = cases.find_cells_box(spec.n, xs, ys, zs, p0)
    al = fg["alpha_i"]
    a, e, E = (orc.local_density(al, c) for c in (aC, eC, EC))
    serial = orc.laser_calculate(g, geomG, MODELS, al, fg["alpha_b"], fg["alphaBc"], a, e, E, fg["T_i"], p0, p1, Ip, dA, cellG)
    nd = 8
    meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, rk) for rk in range(nd)]
    geoms = [cases.box_block_polymesh(spec.n, xs, ys, zs, procs, rk, m) for rk, m in enumerate(meshes)]
    gids = [np.asarray(m.cellGlobal) for m in meshes]
    cells = []
    for gid in gids:
        g2l = np.full(g.nCells, -1, np.int32); g2l[gid] = np.arange(len(gid), dtype=np.int32)
        cells.append(g2l[cellG])
    ref = {}
    for local in (False, True):
        ref[local] = orc.laser_calculate_ranks(meshes, geoms, MODELS, [serial["reflectingCells"][i] for i in gids],
                                               [serial["nHat"][i] for i in gids], [gm["pointGlobal"] for gm in geoms],
                                               [a[i] for i in gids], [e[i] for i in gids], [E[i] for i in gids],
                                               [fg["T_i"][i] for i in gids], p0, p1, Ip, dA, cells, localFindCell=local)
    assert int(ref[False]["created"].sum()) == serial["nParticles"]
    assert int(ref[True]["created"].sum()) < serial["nParticles"], "no reflected particle is inserted behind a processor patch in this case"
    fs = [cases.synthetic_fields(spec, m) for m in meshes]
    grp = abi.Group(nd)
    ctxs = [abi.Context(m, spec.nPhases) for m in meshes]
    outs, errs = [None] * nd, []

    def work(rk):
        try:
            m, f, ctx = meshes[rk], fs[rk], ctxs[rk]
            ctx.comm_init_local(rk, grp)
            cases.load_case(ctx, spec, f)
            ctx.set_cell_centres(f["C"], f["deltab"])
            ctx.laser_set_polymesh(m.Cf, m.Cfb, geoms[rk])
            ctx.laser_set_shared_points(geoms[rk]["sharedLocal"], geoms[rk]["sharedGlobal"])
            ctx.laser_set_shared_points(geoms[rk]["sharedLocal"], geoms[rk]["sharedGlobal"])     # idempotent (collective)
            ctx.laser_set_reflection(MODELS)
            ctx.laser_set_local_density_absorption([0, 1, 2], aC, eC, EC)
            ctx.laser_set_local_find_cell(True)
            own = cells[rk] >= 0                        # each rank hands over ONLY its own seeds, as the solver does
            Q, created = ctx.laser_calculate(p0[own], p1[own], Ip[own], dA[own], cells[rk][own])
            outs[rk] = (Q, created)
        except Exception as ex:          # noqa: BLE001
            errs.append((rk, ex))
    ts = [threading.Thread(target=work, args=(rk,), daemon=True) for rk in range(nd)]
    [t.start() for t in ts]
    [t.join(180) for t in ts]
    assert not errs, errs
    assert all(o is not None for o in outs), "a rank did not finish"
    assert sum(o[1] for o in outs) == int(ref[True]["created"].sum())
    scale = max(np.abs(q).max() for q in ref[True]["Q"])
    for rk in range(nd):
        assert np.abs(outs[rk][0] - ref[True]["Q"][rk]).max() <= 1e-10 * scale, rk
    for c in ctxs:
        c.close()
    grp.close()


# ---- NCCL, one process per GPU: hand-over by ncclSend/ncclRecv, point sums and counts by ncclAllGather ---------------
_NCCL_ABS = ((25.0, 10.0, 0.3), (0.1, 0.0, 0.0), (0.0, 0.0, 0.0))


def _nccl_case(n):
    spec = cases.case(5, n)
    xs, ys, zs = (hm.vertices(spec.n[d], spec.lengths[d]) for d in range(3))
    d = np.array([0.5, -1.0, 0.2]); d /= np.linalg.norm(d)
    r = np.cross(d, [0.0, 0.0, 1.0]); r /= np.linalg.norm(r)
    p0, p1, Ip, dA = cases.laser_seeds((0.3, 0.93, 0.45), d, r, 8, 16, 0.15, float(np.sqrt(3.0)), sigma=0.07, power=200.0)
    return spec, (xs, ys, zs), (p0, p1, Ip, dA), cases.find_cells_box(spec.n, xs, ys, zs, p0)


def _nccl_laser_worker(rank, world, n, out):
    import os, time
    spec, v, (p0, p1, Ip, dA), cellG = _nccl_case(n)
    procs = hm.procs_for(world)
    m = hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, rank)
    f = cases.synthetic_fields(spec, m)
    idfile = os.path.join(out, "nccl_id")
    if rank == 0:
        with open(idfile + ".tmp", "wb") as fh:
            fh.write(abi.nccl_unique_id())
        os.rename(idfile + ".tmp", idfile)
    t0 = time.time()
    while not os.path.exists(idfile):
        if time.time() - t0 > 120:
            raise RuntimeError("no NCCL id")
        time.sleep(0.05)
    ctx = abi.Context(m, spec.nPhases, device=rank)
    ctx.comm_init_nccl(rank, world, open(idfile, "rb").read())
    cases.load_case(ctx, spec, f)
    ctx.set_cell_centres(f["C"], f["deltab"])
    geom = cases.box_block_polymesh(spec.n, *v, procs, rank, m)
    ctx.laser_set_polymesh(m.Cf, m.Cfb, geom)
    ctx.laser_set_shared_points(geom["sharedLocal"], geom["sharedGlobal"])
    ctx.laser_set_reflection(MODELS)
    ctx.laser_set_local_density_absorption([0, 1, 2], *_NCCL_ABS)
    gid = np.asarray(m.cellGlobal)
    g2l = np.full(spec.nCells, -1, np.int32); g2l[gid] = np.arange(len(gid), dtype=np.int32)
    # each rank hands over ONLY the seeds whose cell it found, as laserDTRM::initialise() seeds its local cloud: a rank
    # with few or no seeds of its own still receives its neighbours' particles (capacities come from the global count)
    own = g2l[cellG] >= 0
    Q, created = ctx.laser_calculate(p0[own], p1[own], Ip[own], dA[own], g2l[cellG][own])
    np.savez(os.path.join(out, f"rank{rank}.npz"), Q=Q, created=np.array([created]), gid=gid)
    ctx.close()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("world", [2, 4, 8])
def test_nccl_ranks_match_serial_oracle(world):
    if abi.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import multiprocessing as mp
    import os
    import tempfile
    n = 16
    spec, v, (p0, p1, Ip, dA), cellG = _nccl_case(n)
    g = mo.box_mesh(*spec.n, spec.lengths, spec.patch_types)
    geomG = cases.box_polymesh(g, *v)
    fg = cases.synthetic_fields(spec, g)
    al = fg["alpha_i"]
    ref = orc.laser_calculate(g, geomG, MODELS, al, fg["alpha_b"], fg["alphaBc"], *(orc.local_density(al, c) for c in _NCCL_ABS),
                              fg["T_i"], p0, p1, Ip, dA, cellG)
    ctxm = mp.get_context("spawn")
    with tempfile.TemporaryDirectory() as out:
        ps = [ctxm.Process(target=_nccl_laser_worker, args=(r, world, n, out)) for r in range(world)]
        [p.start() for p in ps]
        [p.join(timeout=500) for p in ps]
        assert all(p.exitcode == 0 for p in ps), [p.exitcode for p in ps]
        zs = [np.load(os.path.join(out, f"rank{r}.npz")) for r in range(world)]
    assert sum(int(z["created"][0]) for z in zs) == ref["nParticles"]
    scale = np.abs(ref["Q"]).max()
    for r, z in enumerate(zs):
        err = np.abs(z["Q"] - ref["Q"][z["gid"]]).max()
        assert err <= 1e-10 * scale, f"Q rank {r}: {err:.3e} of {scale:.3e}"


def test_walk_table_and_ell_gathers_agree(monkeypatch):
    """The per-cell face records (default) and the gathers through the ELL tables (TCIF_LASER_TABLE=0, also the path
    of meshes with more than 8 faces per cell) walk the same faces in the same order."""
    res = []
    for flag in ("1", "0"):
        monkeypatch.setenv("TCIF_LASER_TABLE", flag)
        spec, m, v, geom, f, ctx = _setup()
        ctx.laser_set_reflection(MODELS)
        ctx.laser_set_local_density_absorption([0, 1, 2], (8.0, 4.0, 0.1), (0, 0, 0), (0, 0, 0))
        p0, p1, Ip, dA, cell = _seeds(spec, v, geom, (0.5, 0.93, 0.5), (0.25, -1.0, 0.1))
        res.append(ctx.laser_calculate(p0, p1, Ip, dA, cell))
        ctx.close()
    assert res[0][1] == res[1][1] > 128
    _close(res[0][0], res[1][0], "Q table vs ELL", 1e-13)
    assert np.array_equal(res[0][0] != 0, res[1][0] != 0)
