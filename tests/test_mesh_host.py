"""Product mesh generator / decomposer (C++, libtcif_host.so) against the numpy mesh oracle: bit-exact."""
import numpy as np
import pytest

from oracle import mesh_oracle as mo
from thermocapillaryinterfoam_b200 import mesh as hm

FIELDS = ["owner", "neighbour", "Sf", "w", "V", "C", "Cf", "patchType", "patchStart", "patchSize", "patchNbrProc",
          "faceCells", "Sfb", "magSfb", "deltaCoeffsb", "wb", "Cfb"]


def same(a, b, what):
    for k in FIELDS:
        x, y = np.asarray(getattr(a, k)), np.asarray(getattr(b, k))
        assert x.shape == y.shape, (what, k, x.shape, y.shape)
        assert np.array_equal(x, y), (what, k)


@pytest.mark.parametrize("n,pt,grading", [((5, 4, 3), (0,) * 6, None), ((7, 6, 1), (0, 0, 0, 0, 1, 1), None),
                                          ((4, 5, 6), (0, 0, 2, 0, 0, 0), (2.0, 0.5, 3.0))])
def test_box_matches_oracle(n, pt, grading):
    L = (1.0, 0.7, 1.3)
    ref = mo.box_mesh(*n, L, pt, grading)
    got = hm.box_block(n, L, pt, grading=grading)
    same(got, ref, "box")
    assert np.array_equal(got.cellGlobal, np.arange(ref.nCells))
    assert np.array_equal(got.faceProcAddressing, np.arange(ref.nIF) + 1)


@pytest.mark.parametrize("n,procs", [((6, 4, 4), (2, 1, 1)), ((6, 6, 4), (2, 2, 1)), ((6, 5, 4), (2, 2, 2)),
                                     ((7, 5, 3), (3, 2, 1)), ((8, 8, 1), (2, 2, 1))])
def test_blocks_match_oracle_decomposition(n, procs):
    pt = (0, 0, 0, 0, 1, 1) if n[2] == 1 else (0,) * 6
    L = (1.0, 1.0, 0.5)
    grading = (1.5, None, None)
    g = mo.box_mesh(*n, L, pt, grading)
    nP = procs[0] * procs[1] * procs[2]
    cp = mo.simple_decomposition(*n, *procs)
    assert np.array_equal(cp, hm.simple_map(n, procs))
    subs = mo.decompose(g, cp, nP)
    gh = hm.box_block(n, L, pt, grading=grading)
    for r in range(nP):
        direct = hm.box_block(n, L, pt, procs, r, grading)
        general = hm.decompose(gh, cp, nP, r)
        same(direct, subs[r], f"direct block {r}")
        same(general, subs[r], f"general decompose {r}")
        for m in (direct, general):
            assert np.array_equal(m.cellGlobal, subs[r].cellProcAddressing)
            assert np.array_equal(m.faceProcAddressing, subs[r].faceProcAddressing)
            assert np.array_equal(m.bfaceProcAddressing, subs[r].bfaceProcAddressing)
            assert np.array_equal(m.boundaryProcAddressing, subs[r].boundaryProcAddressing)
        # neighbour info used by the case harness
        for p in range(direct.nPatches):
            if direct.patchType[p] != 3:
                continue
            sl = slice(direct.patchStart[p], direct.patchStart[p] + direct.patchSize[p])
            a = direct.bfaceProcAddressing[sl]
            f = np.abs(a) - 1
            nb = np.where(a < 0, g.owner[f], g.neighbour[f])
            assert np.array_equal(direct.nbrGlobal[sl], nb)
            assert np.array_equal(direct.nbrC[sl], g.C[nb])
            assert np.array_equal(general.nbrGlobal[sl], nb)


def test_scotch_like_irregular_map():
    """Any cell->rank labelList (what decomposePar -method scotch writes to cellDecomposition) is accepted."""
    n = (6, 5, 4)
    g = mo.box_mesh(*n)
    rng = np.random.default_rng(3)
    cp = ((g.C[:, 0] + 0.3 * g.C[:, 1] + 0.05 * rng.standard_normal(g.nCells)) * 3).astype(np.int32).clip(0, 2)
    subs = mo.decompose(g, cp, 3)
    gh = hm.box_block(n)
    for r in range(3):
        same(hm.decompose(gh, cp, 3, r), subs[r], f"irregular {r}")


def test_processor_faces_pair_up():
    n, procs = (6, 6, 6), (2, 2, 2)
    ms = [hm.box_block(n, procs=procs, rank=r) for r in range(8)]
    for r, m in enumerate(ms):
        for p in range(m.nPatches):
            if m.patchType[p] != 3:
                continue
            e = int(m.patchNbrProc[p])
            q = [k for k in range(ms[e].nPatches) if ms[e].patchNbrProc[k] == r][0]
            a = m.bfaceProcAddressing[m.patchStart[p]:m.patchStart[p] + m.patchSize[p]]
            b = ms[e].bfaceProcAddressing[ms[e].patchStart[q]:ms[e].patchStart[q] + ms[e].patchSize[q]]
            assert np.array_equal(a, -b)          # same faces, same order, opposite orientation
            Sa = m.Sfb[m.patchStart[p]:m.patchStart[p] + m.patchSize[p]]
            Sb = ms[e].Sfb[ms[e].patchStart[q]:ms[e].patchStart[q] + ms[e].patchSize[q]]
            assert np.array_equal(Sa, -Sb)
