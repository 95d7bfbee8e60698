"""Processor-patch path on the GPU: decomposed == serial oracle.

* in-process group: N contexts on ONE GPU, one host thread per rank, halo by device copies;
* NCCL: one process per GPU (skipped unless >= 2 devices are visible), halo by ncclSend/ncclRecv.
"""
import os
import tempfile
import threading

import numpy as np
import pytest

from oracle import oracle as orc
from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm
from helpers import assert_field_close, global_mesh

pytestmark = pytest.mark.gpu


def _serial(spec, L=3):
    g = global_mesh(spec)
    f = cases.synthetic_fields(spec, g)
    ref = orc.stage(g, orc.mules_cfg(L), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                    f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    ref["V"] = g.V
    return g, ref


def _compare(spec, ref, m, out):
    gid = np.asarray(m.cellGlobal)
    assert_field_close(out["F_i"], ref["F_i"][gid], "force")
    fa = np.asarray(m.faceProcAddressing) - 1
    for p in range(spec.nPhases):
        assert_field_close(out["alpha_i"][p], ref["alpha_i"][p][gid], f"alpha{p}")
        assert_field_close(out["alphaPhi_i"][p], ref["alphaPhi_i"][p][fa], f"alphaPhi{p}")
        for k in range(m.nPatches):
            sl = slice(int(m.patchStart[k]), int(m.patchStart[k] + m.patchSize[k]))
            a = np.asarray(m.bfaceProcAddressing)[sl]
            if m.patchType[k] == 3:
                # processor patch: alpha_b = neighbour-rank cell value; F_b likewise; flux = +-serial flux
                nb = np.asarray(m.nbrGlobal)[sl]
                assert_field_close(out["alpha_b"][p][sl], ref["alpha_i"][p][nb], f"alpha{p} halo")
                assert_field_close(out["alphaPhi_b"][p][sl], np.sign(a) * ref["alphaPhi_i"][p][np.abs(a) - 1],
                                   f"alphaPhi{p} processor patch")
                if p == 0:
                    assert_field_close(out["F_b"][sl], ref["F_i"][nb], "force halo")
            else:
                assert_field_close(out["alpha_b"][p][sl], ref["alpha_b"][p][a - 1], f"alpha{p} wall")


def _run_rank(ctx, spec, f):
    cases.load_case(ctx, spec, f)
    ctx.stage(f["deltaT"])
    Fi, Fb = ctx.get_force()
    al = [ctx.get_alpha(p) for p in range(spec.nPhases)]
    ph = [ctx.get_phi_alpha(p) for p in range(spec.nPhases)]
    st = [ctx.alpha_stats(p) for p in range(spec.nPhases)]
    return dict(F_i=Fi, F_b=Fb, alpha_i=[a[0] for a in al], alpha_b=[a[1] for a in al],
                alphaPhi_i=[a[0] for a in ph], alphaPhi_b=[a[1] for a in ph], stats=st)


@pytest.mark.parametrize("config,n,procs", [(5, 12, (2, 1, 1)), (5, 12, (2, 2, 2)), (1, 16, (2, 2, 1)),
                                            (3, 12, (3, 1, 2))])
def test_local_group_matches_serial(config, n, procs):
    spec = cases.case(config, n)
    g, ref = _serial(spec)
    nd = procs[0] * procs[1] * procs[2]
    meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, r) for r in range(nd)]
    fs = [cases.synthetic_fields(spec, m) for m in meshes]
    grp = abi.Group(nd)
    ctxs = [abi.Context(m, spec.nPhases) for m in meshes]
    outs, errs = [None] * nd, []

    def work(r):
        try:
            ctxs[r].comm_init_local(r, grp)
            outs[r] = _run_rank(ctxs[r], spec, fs[r])
        except Exception as e:          # noqa: BLE001
            errs.append((r, e))
    ts = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(nd)]
    [t.start() for t in ts]
    [t.join(timeout=300) for t in ts]
    assert not errs, errs
    for r in range(nd):
        _compare(spec, ref, meshes[r], outs[r])
    # global reductions: every rank reports the same sum(V alpha), equal to the serial one
    for p in range(spec.nPhases):
        tot = float((ref["V"] * ref["alpha_i"][p]).sum())
        for r in range(nd):
            assert outs[r]["stats"][p] == outs[0]["stats"][p]
            assert abs(outs[r]["stats"][p][0] - tot) <= 1e-13 * abs(tot)
    for c in ctxs:
        c.close()
    grp.close()


def test_irregular_decomposition_local_group():
    """A scotch-like irregular cell->rank map (multiple neighbours, ragged patches)."""
    spec = cases.case(5, 10)
    g, ref = _serial(spec)
    gh = hm.box_block(spec.n, spec.lengths, spec.patch_types)
    rng = np.random.default_rng(11)
    cp = ((gh.C[:, 0] + 0.4 * gh.C[:, 1] + 0.03 * rng.standard_normal(gh.nCells)) * 3).astype(np.int32).clip(0, 2)
    meshes = [hm.decompose(gh, cp, 3, r) for r in range(3)]
    fs = [cases.synthetic_fields(spec, m) for m in meshes]
    grp = abi.Group(3)
    ctxs = [abi.Context(m, 3) for m in meshes]
    outs, errs = [None] * 3, []

    def work(r):
        try:
            ctxs[r].comm_init_local(r, grp)
            outs[r] = _run_rank(ctxs[r], spec, fs[r])
        except Exception as e:          # noqa: BLE001
            errs.append((r, e))
    ts = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(3)]
    [t.start() for t in ts]
    [t.join(timeout=300) for t in ts]
    assert not errs, errs
    for r in range(3):
        _compare(spec, ref, meshes[r], outs[r])
    for c in ctxs:
        c.close()
    grp.close()


def test_processor_patches_without_communicator_is_an_error():
    spec = cases.case(5, 8)
    m = hm.box_block(spec.n, spec.lengths, spec.patch_types, (2, 1, 1), 0)
    f = cases.synthetic_fields(spec, m)
    ctx = abi.Context(m, 3)
    cases.load_case(ctx, spec, f)
    with pytest.raises(abi.TcifError, match="communicator"):
        ctx.stage(f["deltaT"])
    ctx.close()


# ---- NCCL, one process per GPU ----------------------------------------------------------------------
def _nccl_worker(rank, world, config, n, out, transport="peer"):
    os.environ["TCIF_PEER_HALO"] = "1" if transport == "peer" else "0"
    spec = cases.case(config, n)
    m = hm.box_block(spec.n, spec.lengths, spec.patch_types, hm.procs_for(world), rank)
    f = cases.synthetic_fields(spec, m)
    idfile = os.path.join(out, "nccl_id")
    if rank == 0:
        with open(idfile + ".tmp", "wb") as fh:
            fh.write(abi.nccl_unique_id())
        os.rename(idfile + ".tmp", idfile)
    import time
    t0 = time.time()
    while not os.path.exists(idfile):
        if time.time() - t0 > 120:
            raise RuntimeError("no NCCL id")
        time.sleep(0.05)
    uid = open(idfile, "rb").read()
    ctx = abi.Context(m, spec.nPhases, device=rank)
    ctx.comm_init_nccl(rank, world, uid)
    assert ctx.comm_transport == transport, (ctx.comm_transport, transport)
    res = _run_rank(ctx, spec, f)
    np.savez(os.path.join(out, f"rank{rank}.npz"), F_i=res["F_i"], F_b=res["F_b"], stats=np.array(res["stats"]),
             **{f"{k}{p}": res[k][p] for k in ("alpha_i", "alpha_b", "alphaPhi_i", "alphaPhi_b")
                for p in range(spec.nPhases)})
    ctx.close()


@pytest.mark.timeout(900)
@pytest.mark.parametrize("world", [2, 4, 8])
def test_nccl_ranks_match_serial(world):
    """One process per GPU: processor-patch values over peer memory (default) and over ncclSend/ncclRecv; both
    against the serial oracle, and bit-identical to each other."""
    if abi.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import multiprocessing as mp
    config, n = 5, 16
    spec = cases.case(config, n)
    g, ref = _serial(spec)
    ctxm = mp.get_context("spawn")
    results = {}
    for transport in ("peer", "nccl"):
        with tempfile.TemporaryDirectory() as out:
            ps = [ctxm.Process(target=_nccl_worker, args=(r, world, config, n, out, transport)) for r in range(world)]
            [p.start() for p in ps]
            [p.join(timeout=400) for p in ps]
            assert all(p.exitcode == 0 for p in ps), (transport, [p.exitcode for p in ps])
            stats0 = None
            for r in range(world):
                m = hm.box_block(spec.n, spec.lengths, spec.patch_types, hm.procs_for(world), r)
                z = np.load(os.path.join(out, f"rank{r}.npz"))
                res = dict(F_i=z["F_i"], F_b=z["F_b"],
                           **{k: [z[f"{k}{p}"] for p in range(3)] for k in ("alpha_i", "alpha_b", "alphaPhi_i", "alphaPhi_b")})
                _compare(spec, ref, m, res)
                stats0 = z["stats"] if stats0 is None else stats0
                assert np.array_equal(z["stats"], stats0)
                results[(transport, r)] = res
    for r in range(world):
        a, b = results[("peer", r)], results[("nccl", r)]
        assert np.array_equal(a["F_i"], b["F_i"]) and np.array_equal(a["F_b"], b["F_b"]), f"force differs between transports, rank {r}"
        for k in ("alpha_i", "alpha_b", "alphaPhi_i", "alphaPhi_b"):
            for p in range(3):
                assert np.array_equal(a[k][p], b[k][p]), (k, p, r)
