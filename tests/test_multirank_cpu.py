"""N>1 host logic on CPU: decomposed == serial for the oracle, once with the
oracle's own shared-memory thread exchange and once over torch.distributed `gloo` (world_size 2, one
process per rank -- the same rendezvous / one-rank-per-sub-domain plumbing bench.py uses with NCCL)."""
import os
import tempfile

import numpy as np
import pytest

from oracle import oracle as orc
from thermocapillaryinterfoam_b200 import cases, mesh as hm
from helpers import assert_field_close, global_mesh


def _serial(spec, L=3):
    g = global_mesh(spec)
    f = cases.synthetic_fields(spec, g)
    ref = orc.stage(g, orc.mules_cfg(L), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                    f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    return g, f, ref


def _check_against_serial(spec, ref, meshes, results):
    P = spec.nPhases
    for m, r in zip(meshes, results):
        gid = np.asarray(m.cellGlobal)
        assert_field_close(r["F_i"], ref["F_i"][gid], "force")
        for p in range(P):
            assert_field_close(r["alpha_i"][p], ref["alpha_i"][p][gid], f"alpha{p}")
            fa = np.asarray(m.faceProcAddressing) - 1             # internal faces keep orientation
            assert_field_close(r["alphaPhi_i"][p], ref["alphaPhi_i"][p][fa], f"alphaPhi{p}")
            # processor faces: +-(global internal face + 1); flipped faces carry the opposite flux
            for k in range(m.nPatches):
                if m.patchType[k] != 3:
                    continue
                sl = slice(int(m.patchStart[k]), int(m.patchStart[k] + m.patchSize[k]))
                a = np.asarray(m.bfaceProcAddressing)[sl]
                assert_field_close(r["alphaPhi_b"][p][sl], np.sign(a) * ref["alphaPhi_i"][p][np.abs(a) - 1],
                                   f"alphaPhi{p} processor patch")
    # mass: sum over sub-domains == serial to round-off
    for p in range(P):
        tot = sum(float((np.asarray(m.V) * r["alpha_i"][p]).sum()) for m, r in zip(meshes, results))
        g = float((ref["V"] * ref["alpha_i"][p]).sum())
        assert abs(tot - g) <= 1e-13 * abs(g)


@pytest.mark.parametrize("config,n,procs", [(5, 12, (2, 1, 1)), (5, 12, (2, 2, 2)), (3, 12, (2, 2, 1)),
                                            (1, 16, (2, 2, 1))])
def test_decomposed_threads_match_serial(config, n, procs):
    spec = cases.case(config, n)
    g, f, ref = _serial(spec)
    ref["V"] = g.V
    nd = procs[0] * procs[1] * procs[2]
    meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, r) for r in range(nd)]
    fs = [cases.synthetic_fields(spec, m) for m in meshes]
    _, res = orc.stage_threads(meshes, orc.mules_cfg(3), [x["alpha_i"] for x in fs], [x["alpha_b"] for x in fs],
                               [x["alphaBc"] for x in fs], spec.sigma, [x["T_i"] for x in fs], [x["T_b"] for x in fs],
                               [x["phi_i"] for x in fs], [x["phi_b"] for x in fs], [x["phiAlpha_i"] for x in fs],
                               [x["phiAlpha_b"] for x in fs], spec.deltaT)
    _check_against_serial(spec, ref, meshes, res)


# ---- gloo, world_size 2 --------------------------------------------------------------------------
def _gloo_worker(rank, world, initfile, config, n, out):
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"file://{initfile}", rank=rank, world_size=world)
    try:
        spec = cases.case(config, n)
        procs = hm.procs_for(world)
        m = hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, rank)
        f = cases.synthetic_fields(spec, m)
        patches = [(int(m.patchNbrProc[k]), int(m.patchSize[k])) for k in range(m.nPatches) if m.patchType[k] == 3]

        def halo(ncomp, send):
            recv = np.empty_like(send)
            off, reqs, bufs = 0, [], []
            for peer, sz in patches:            # one message per processor patch, like processorFvPatchField
                s = torch.from_numpy(np.ascontiguousarray(send[off:off + sz]))
                r = torch.empty_like(s)
                reqs += [dist.isend(s, peer), dist.irecv(r, peer)]
                bufs.append((off, sz, r, s))
                off += sz
            for q in reqs:
                q.wait()
            for off, sz, r, _ in bufs:
                recv[off:off + sz] = r.numpy()
            return recv

        def reduce(op, vals):
            t = torch.from_numpy(vals.copy())
            dist.all_reduce(t, op={0: dist.ReduceOp.SUM, 1: dist.ReduceOp.MIN, 2: dist.ReduceOp.MAX}[op])
            return t.numpy()

        comm, keep = orc.make_comm(halo, reduce)
        res = orc.stage(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                        f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"], comm=comm)
        np.savez(os.path.join(out, f"rank{rank}.npz"), F_i=res["F_i"],
                 **{f"alpha{p}": res["alpha_i"][p] for p in range(spec.nPhases)},
                 **{f"alphaPhi{p}": res["alphaPhi_i"][p] for p in range(spec.nPhases)},
                 **{f"alphaPhib{p}": res["alphaPhi_b"][p] for p in range(spec.nPhases)})
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_gloo_world_size_2_matches_serial():
    import torch.multiprocessing as mp
    config, n, world = 5, 10, 2
    spec = cases.case(config, n)
    g, f, ref = _serial(spec)
    ref["V"] = g.V
    with tempfile.TemporaryDirectory() as out:
        initfile = os.path.join(out, "rendezvous")
        mp.spawn(_gloo_worker, args=(world, initfile, config, n, out), nprocs=world, join=True)
        meshes = [hm.box_block(spec.n, spec.lengths, spec.patch_types, hm.procs_for(world), r) for r in range(world)]
        res = []
        for r in range(world):
            z = np.load(os.path.join(out, f"rank{r}.npz"))
            res.append(dict(F_i=z["F_i"], alpha_i=[z[f"alpha{p}"] for p in range(3)],
                            alphaPhi_i=[z[f"alphaPhi{p}"] for p in range(3)],
                            alphaPhi_b=[z[f"alphaPhib{p}"] for p in range(3)]))
    _check_against_serial(spec, ref, meshes, res)
