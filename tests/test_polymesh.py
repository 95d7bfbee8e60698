"""Non-hex meshes and the polyMesh / cellDecomposition input path.

CPU: the C++ reader (libtcif_host.so, csrc/polymesh_host.cpp) against the numpy restatement
oracle/polymesh_oracle.py on tet / prism / warped-hex / randomly renumbered meshes written as ASCII polyMesh
directories -- addressing bit-exact, geometry bit-exact; the hex case against the block generator; a
cellDecomposition labelList read back and used to decompose (C++ vs numpy sub-meshes bit-exact).
GPU (-m gpu): force, stage and solve() parity with the oracle on those meshes, i.e. the K = 4 instantiation of the
fixed-K sweeps (tets renumbered so that a cell owns all four faces) and the generic-K kernels (Ko up to 6), serial and
decomposed by an irregular map (in-process group).
"""
import os

import numpy as np
import pytest

from oracle import mesh_oracle as mo, oracle as orc, polymesh_oracle as po
from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm
from helpers import assert_field_close, add_neighbour_info

MESHES = {
    "tet": dict(kind="tet"),
    "tet_renumbered": dict(kind="tet", renumber=3),            # Ko = Kn = 4: the K = 4 sweeps
    "prism": dict(kind="prism"),
    "prism_renumbered": dict(kind="prism", renumber=11),       # Ko = Kn = 5: generic-K kernels
    "hex_renumbered": dict(kind="hex", renumber=7),            # Ko = Kn = 6: generic-K kernels
    "hex_warped": dict(kind="hex", jitter=0.2),                # non-planar quads, non-orthogonal
}
ARRAYS = ("owner", "neighbour", "Sf", "w", "V", "C", "Cf", "patchType", "patchStart", "patchSize", "faceCells", "Sfb",
          "magSfb", "deltaCoeffsb", "wb", "Cfb")


def _poly(name, n=(5, 4, 3), lengths=(1.0, 0.8, 0.6)):
    return po.split_box(*n, lengths, **MESHES[name])


@pytest.mark.parametrize("name", sorted(MESHES))
def test_cpp_reader_matches_numpy_oracle_bit_exactly(name, tmp_path):
    pm = _poly(name)
    po.write_polymesh(str(tmp_path), *pm)
    ref = po.from_polymesh(*po.read_polymesh(str(tmp_path)))
    got = hm.read_polymesh(str(tmp_path))
    assert got.nCells == ref.nCells
    for a in ARRAYS:
        g = np.asarray(getattr(got, a if a != "deltaCoeffsb" else "deltaCoeffsb"))
        r = np.asarray(getattr(ref, a))
        assert g.shape == r.shape and np.array_equal(g, r), f"{name}: {a} differs (max {np.abs(g - r).max() if g.size else 0})"
    # what the reader must guarantee for the device layout
    assert np.all(ref.owner < ref.neighbour) and np.all(np.diff(ref.owner) >= 0)
    acc = np.zeros((ref.nCells, 3))
    np.add.at(acc, ref.owner, ref.Sf); np.add.at(acc, ref.neighbour, -ref.Sf); np.add.at(acc, ref.faceCells, ref.Sfb)
    assert np.abs(acc).max() < 1e-15, "cells are not closed"
    assert abs(ref.V.sum() - 0.48) < 1e-14
    got.close()


def test_unsplit_box_equals_the_block_generator():
    a = po.split_box_mesh(6, 5, 4, (1.0, 0.8, 0.6), "hex")
    b = mo.box_mesh(6, 5, 4, (1.0, 0.8, 0.6))
    for k in ("owner", "neighbour", "faceCells", "patchStart", "patchSize"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k
    for k in ("Sf", "V", "C", "Cf", "w", "Sfb", "Cfb", "deltaCoeffsb"):
        assert np.allclose(getattr(a, k), getattr(b, k), rtol=1e-14, atol=1e-15), k   # general formulas vs the analytic hex ones: rounding only


def test_slot_counts_of_the_test_meshes():
    for name, (ko, kn) in {"tet": (3, 3), "tet_renumbered": (4, 4), "prism_renumbered": (5, 5), "hex_renumbered": (6, 6)}.items():
        m = po.from_polymesh(*_poly(name, (6, 5, 4)))
        assert (np.bincount(m.owner).max(), np.bincount(m.neighbour).max()) == (ko, kn), name


def test_cell_decomposition_file_roundtrip_and_decompose(tmp_path):
    pm = _poly("prism_renumbered")
    po.write_polymesh(str(tmp_path / "polyMesh"), *pm)
    g = po.from_polymesh(*pm)
    rng = np.random.default_rng(5)
    cellProc = ((g.C[:, 0] * 2).astype(np.int32) + 2 * (g.C[:, 1] > 0.3 + 0.2 * g.C[:, 2]).astype(np.int32)) % 3   # irregular, 3 ranks
    cellProc[rng.integers(0, g.nCells, 6)] = 2
    path = tmp_path / "cellDecomposition"
    with open(path, "w") as fh:
        fh.write("FoamFile\n{\n    version 2.0;\n    format ascii;\n    class labelList;\n    object cellDecomposition;\n}\n// scotch\n"
                 f"\n{g.nCells}\n(\n" + "\n".join(str(int(x)) for x in cellProc) + "\n)\n")
    got = hm.read_label_list(str(path))
    assert np.array_equal(got, cellProc) and np.array_equal(po.read_label_list(str(path)), cellProc)
    gh = hm.read_polymesh(str(tmp_path / "polyMesh"))
    subs = mo.decompose(g, cellProc, 3)
    for r in range(3):
        s = hm.decompose(gh, got, 3, r)
        for a in ARRAYS + ("patchNbrProc",):
            assert np.array_equal(np.asarray(getattr(s, a)), np.asarray(getattr(subs[r], a))), (r, a)
        assert np.array_equal(np.asarray(s.cellGlobal), subs[r].cellProcAddressing)
        s.close()
    gh.close()


def test_reader_errors_are_loud(tmp_path):
    with pytest.raises(hm.TcifError, match="cannot open"):
        hm.read_polymesh(str(tmp_path / "nowhere"))
    pm = list(_poly("tet", (2, 2, 2)))
    pm[4] = [(n, "cyclic", s, k) for (n, t, s, k) in pm[4]]
    po.write_polymesh(str(tmp_path), *pm)
    with pytest.raises(hm.TcifError, match="not supported"):
        hm.read_polymesh(str(tmp_path))


# ---- GPU parity on non-hex meshes -----------------------------------------------------------------------------------
def _spec_fields(m, nPhases=3):
    spec = cases.CaseSpec("polymesh_box", (10, 10, 10), nPhases, [cases.sigma_linear(s) for s in cases.SIGMA0[:nPhases * (nPhases - 1) // 2]],
                          False, 5)
    return spec, cases.synthetic_fields(spec, m)


@pytest.mark.gpu
@pytest.mark.parametrize("name,fixedK", [("tet_renumbered", 4), ("prism", 3), ("prism_renumbered", 0), ("hex_renumbered", 0), ("hex_warped", 3)])
def test_stage_and_solve_match_oracle_on_nonhex_meshes(name, fixedK):
    m = po.from_polymesh(*_poly(name, (8, 8, 8), (1.0, 1.0, 1.0)))
    spec, f = _spec_fields(m)
    ctx = abi.Context(m, 3, device=0)
    assert ctx.mesh_info["fixedK"] == fixedK, ctx.mesh_info
    cases.load_case(ctx, spec, f)
    ctx.stage(f["deltaT"])
    ref = orc.stage(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                    f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    Fi, Fb = ctx.get_force()
    assert np.abs(ref["F_i"]).max() > 0
    assert_field_close(Fi, ref["F_i"], "force")
    assert_field_close(Fb, ref["F_b"], "force boundary")
    for p in range(3):
        ai, ab = ctx.get_alpha(p)
        pi, pb = ctx.get_phi_alpha(p)
        assert np.array_equal(ai, ref["alpha_i"][p]) and np.array_equal(pi, ref["alphaPhi_i"][p]), f"MULES not bit-identical, phase {p}"
        assert np.array_equal(ab, ref["alpha_b"][p]) and np.array_equal(pb, ref["alphaPhi_b"][p])
    # solve(): flux assembly (vanLeer + compression), predictor, MULES, rhoPhi
    ctx2 = abi.Context(m, 3, device=0)
    cases.load_case(ctx2, spec, f)
    ctx2.set_cell_centres(np.asarray(f["C"]), np.asarray(f["deltab"]))
    rho = (1000.0, 800.0, 1.2)
    for p in range(3):
        ctx2.set_rho(p, rho[p])
    ctx2.solve_alphas(abi.alpha_controls(1, 1, "vanLeer", "linear", True, 1e-13, 400), f["deltaT"])
    r2 = orc.solve(m, orc.mules_cfg(3), orc.alpha_ctl(1, 1, "vanLeer", "linear", True, 1e-13, 400), f["alpha_i"], f["alpha_b"],
                   f["alphaBc"], f["phi_i"], f["phi_b"], f["deltaT"], list(rho))
    for p in range(3):
        assert_field_close(ctx2.get_alpha(p)[0], r2["alpha_i"][p], f"solve alpha{p}", rtol=1e-11)
        assert_field_close(ctx2.get_phi_alpha(p)[0], r2["alphaPhi_i"][p], f"solve alphaPhi{p}", rtol=1e-11)
    assert_field_close(ctx2.get_rho_phi()[0], r2["rhoPhi_i"], "rhoPhi", rtol=1e-11)
    ctx.close(); ctx2.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["tet_renumbered", "hex_renumbered"])
def test_irregular_decomposition_of_a_nonhex_mesh_matches_serial_oracle(name):
    import threading
    g = po.from_polymesh(*_poly(name, (8, 8, 8), (1.0, 1.0, 1.0)))
    nR = 3
    cellProc = ((g.C[:, 0] * 2).astype(np.int32) + 2 * (g.C[:, 1] > 0.3 + 0.3 * g.C[:, 2]).astype(np.int32)) % nR
    subs = add_neighbour_info(g, mo.decompose(g, cellProc, nR))
    spec, fg = _spec_fields(g)
    ref = orc.stage(g, orc.mules_cfg(3), fg["alpha_i"], fg["alpha_b"], fg["alphaBc"], fg["sigma"], fg["T_i"], fg["T_b"],
                    fg["phi_i"], fg["phi_b"], fg["phiAlpha_i"], fg["phiAlpha_b"], fg["deltaT"])
    grp = abi.Group(nR)
    out, errs = [None] * nR, []

    def rank(r):
        try:
            s = subs[r]
            f = cases.synthetic_fields(spec, s)
            # the HO flux of a sub-mesh must be the global one (same values, flipped on flipped processor faces)
            ctx = abi.Context(s, 3, device=0)
            ctx.comm_init_local(r, grp)
            cases.load_case(ctx, spec, f)
            ctx.stage(f["deltaT"])
            out[r] = (ctx.get_force()[0], [ctx.get_alpha(p)[0] for p in range(3)])
            ctx.close()
        except Exception as e:      # noqa: BLE001
            errs.append((r, repr(e)))
    th = [threading.Thread(target=rank, args=(r,)) for r in range(nR)]
    [t.start() for t in th]; [t.join() for t in th]
    grp.close()
    assert not errs, errs
    scaleF = np.abs(ref["F_i"]).max()
    for r in range(nR):
        cg = subs[r].cellProcAddressing
        assert np.abs(out[r][0] - ref["F_i"][cg]).max() <= 1e-12 * scaleF, f"force, rank {r}"
        for p in range(3):
            assert np.abs(out[r][1][p] - ref["alpha_i"][p][cg]).max() <= 1e-12, f"alpha{p}, rank {r}"
