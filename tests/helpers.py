"""Shared test helpers: meshes from the numpy mesh ORACLE, synthetic fields, comparisons."""
from __future__ import annotations

import numpy as np

from oracle import mesh_oracle as mo
from thermocapillaryinterfoam_b200 import cases

# |a-b| <= RTOL*max(|a|,|b|) + RTOL*||field||_inf  (floor stated explicitly)
RTOL = 1e-12


def assert_field_close(a, b, what="", rtol=RTOL, floor=0.0):
    """`floor`: extra absolute tolerance for cancellation-dominated fields (stated by the caller)."""
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if a.size == 0:
        return
    scale = max(np.max(np.abs(a)), np.max(np.abs(b)))
    err = np.abs(a - b)
    tol = rtol * np.maximum(np.abs(a), np.abs(b)) + rtol * scale + floor
    bad = err > tol
    assert not bad.any(), f"{what}: {bad.sum()} of {a.size} entries differ, max err {err.max():.3e} (scale {scale:.3e})"


def global_mesh(spec: cases.CaseSpec, grading=None):
    nx, ny, nz = spec.n
    return mo.box_mesh(nx, ny, nz, spec.lengths, spec.patch_types, grading)


def add_neighbour_info(g: mo.Mesh, subs):
    """cellGlobal / nbrC / nbrGlobal on processor sub-meshes (needed by cases.synthetic_fields)."""
    for s in subs:
        s.cellGlobal = s.cellProcAddressing.astype(np.int64)
        nBF = s.nBF
        s.nbrC = np.zeros((nBF, 3))
        s.nbrGlobal = np.zeros(nBF, np.int64)
        for p in range(s.nPatches):
            if s.patchType[p] != mo.PATCH_PROCESSOR:
                continue
            sl = slice(int(s.patchStart[p]), int(s.patchStart[p] + s.patchSize[p]))
            a = s.bfaceProcAddressing[sl]
            f = np.abs(a) - 1
            nb = np.where(a < 0, g.owner[f], g.neighbour[f])     # flipped: local cell is the global neighbour
            s.nbrGlobal[sl] = nb
            s.nbrC[sl] = g.C[nb]
    return subs


def decomposed(spec, procs, grading=None):
    g = global_mesh(spec, grading)
    px, py, pz = procs
    cp = mo.simple_decomposition(*spec.n, px, py, pz)
    subs = mo.decompose(g, cp, px * py * pz)
    return g, add_neighbour_info(g, subs), cp
