"""numpy ORACLE for general polyhedral meshes (test infrastructure only): OpenFOAM polyMesh -> the arrays of
mesh_oracle.Mesh, generators of non-hex test meshes, and an ASCII polyMesh / labelList reader and writer.

Why: the reference runs on arbitrary polyhedral meshes (blockMesh / snappyHexMesh output decomposed by scotch);
hex boxes only ever exercise the K = 2/3 instantiations of the fixed-K sweeps.  Here:

  * `split_box(n, kind)`  each hex of an nx*ny*nz box split into 2 prisms ("prism", cells with 5 faces) or into the
                          6 Kuhn tetrahedra ("tet", 4 faces: the K = 4 sweeps); `renumber=seed` permutes the cell
                          labels at random, which makes some cell own all of its faces (Ko up to 6: the generic-K
                          kernels) -- a legal OpenFOAM mesh (what renumberMesh with a bad ordering would write);
  * `from_polymesh(...)`  geometry by the [UPSTREAM] OpenFOAM formulas, restated from
                          src/OpenFOAM/meshes/primitiveMesh/primitiveMeshFaceCentresAndAreas.C (triangle fan about the
                          vertex average: Sf = 1/2 sum n_i, Cf = 1/3 sum(a_i c_i)/sum a_i, triangles directly),
                          primitiveMeshCellCentresAndVols.C (pyramids about the face-centre average) and
                          surfaceInterpolation::makeWeights (mesh_oracle.linear_weights); boundary deltaCoeffs =
                          1/|Cf - C|, boundary weights 1;
  * `write_polymesh` / `read_polymesh` / `read_label_list`  the ASCII files constant/polyMesh/{points,faces,owner,
                          neighbour,boundary} and a `cellDecomposition` labelList as decomposePar -cellDist writes it.

PARITY UNPINNED: no OpenFOAM here to pin these conventions; the C++ reader in
thermocapillaryinterfoam_b200/csrc/mesh_host.cpp is checked bit-exactly against this file.
"""
from __future__ import annotations

import os
import re

import numpy as np

from .mesh_oracle import Mesh, linear_weights, dot3, PATCH_EMPTY, PATCH_GENERIC, PATCH_WEDGE, PATCH_PROCESSOR

VSMALL = 1e-300
PATCH_TYPE_NAMES = {"patch": PATCH_GENERIC, "wall": PATCH_GENERIC, "symmetryPlane": PATCH_GENERIC, "symmetry": PATCH_GENERIC,
                    "empty": PATCH_EMPTY, "wedge": PATCH_WEDGE, "processor": PATCH_PROCESSOR}


def _cross(a, b):
    return np.stack([a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1], a[:, 2] * b[:, 0] - a[:, 0] * b[:, 2],
                     a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]], axis=1)


def _mag(a):
    return np.sqrt(a[:, 0] * a[:, 0] + a[:, 1] * a[:, 1] + a[:, 2] * a[:, 2])


def face_geometry(points, faces):
    """Sf, Cf of every face (primitiveMesh::makeFaceCentresAndAreas): faces = list of vertex-label sequences."""
    points = np.asarray(points, np.float64)
    nF = len(faces)
    Sf, Cf = np.zeros((nF, 3)), np.zeros((nF, 3))
    nv = np.array([len(f) for f in faces])
    for n in np.unique(nv):
        sel = np.nonzero(nv == n)[0]
        P = points[np.array([faces[i] for i in sel], np.int64)]          # [m, n, 3]
        if n == 3:
            Cf[sel] = (1.0 / 3.0) * (P[:, 0] + P[:, 1] + P[:, 2])
            Sf[sel] = 0.5 * _cross(P[:, 1] - P[:, 0], P[:, 2] - P[:, 0])
            continue
        fC = P[:, 0].copy()
        for k in range(1, n):
            fC = fC + P[:, k]
        fC = fC / n
        sumN, sumA, sumAc = np.zeros((len(sel), 3)), np.zeros(len(sel)), np.zeros((len(sel), 3))
        for k in range(n):                       # triangle (p_k, p_k+1, fC), ascending k: the accumulation order
            p, pn = P[:, k], P[:, (k + 1) % n]
            c = p + pn + fC
            nn = _cross(pn - p, fC - p)
            a = _mag(nn)
            sumN = sumN + nn
            sumA = sumA + a
            sumAc = sumAc + a[:, None] * c
        ok = sumA > VSMALL
        Cf[sel] = np.where(ok[:, None], (1.0 / 3.0) * sumAc / np.where(ok, sumA, 1.0)[:, None], fC)
        Sf[sel] = 0.5 * sumN
    return Sf, Cf


def cell_geometry(nCells, owner, neighbour, Sf, Cf):
    """C, V (primitiveMesh::makeCellCentresAndVols): owner [nFaces], neighbour [nIF]; faces visited in ascending label."""
    nF, nIF = len(owner), len(neighbour)
    cEst, nCellFaces = np.zeros((nCells, 3)), np.zeros(nCells)
    # np.add.at adds in index order = ascending face label: owner contributions first, then neighbour (as OpenFOAM's two loops)
    np.add.at(cEst, owner, Cf)
    np.add.at(nCellFaces, owner, 1.0)
    np.add.at(cEst, neighbour, Cf[:nIF])
    np.add.at(nCellFaces, neighbour, 1.0)
    cEst = cEst / nCellFaces[:, None]
    C, V = np.zeros((nCells, 3)), np.zeros(nCells)
    pyr = dot3(Sf, Cf - cEst[owner])
    pc = 0.75 * Cf + 0.25 * cEst[owner]
    np.add.at(C, owner, pyr[:, None] * pc)
    np.add.at(V, owner, pyr)
    pyrN = dot3(Sf[:nIF], cEst[neighbour] - Cf[:nIF])
    pcN = 0.75 * Cf[:nIF] + 0.25 * cEst[neighbour]
    np.add.at(C, neighbour, pyrN[:, None] * pcN)
    np.add.at(V, neighbour, pyrN)
    C = C / V[:, None]
    V = V * (1.0 / 3.0)
    return C, V


def from_polymesh(points, faces, owner, neighbour, patches) -> Mesh:
    """patches: list of (name, type-name-or-id, startFace, nFaces, neighbProcNo=-1) in polyMesh face numbering."""
    owner, neighbour = np.asarray(owner, np.int32), np.asarray(neighbour, np.int32)
    nIF = len(neighbour)
    nCells = int(max(owner.max(), neighbour.max() if nIF else 0)) + 1
    Sf, Cf = face_geometry(points, faces)
    C, V = cell_geometry(nCells, owner.astype(np.int64), neighbour.astype(np.int64), Sf, Cf)
    w = linear_weights(Sf[:nIF], Cf[:nIF], C[owner[:nIF]], C[neighbour])
    ptype, pstart, psize, pnbr, names, sel = [], [], [], [], [], []
    pos = 0
    for pt in patches:
        name, typ, start, n = pt[:4]
        t = PATCH_TYPE_NAMES[typ] if isinstance(typ, str) else int(typ)
        names.append(name); ptype.append(t); pstart.append(pos); pnbr.append(int(pt[4]) if len(pt) > 4 else -1)
        if t == PATCH_EMPTY:
            psize.append(0)
            continue
        psize.append(int(n)); pos += int(n)
        sel.append(np.arange(start, start + n))
    sel = np.concatenate(sel) if sel else np.zeros(0, np.int64)
    Sfb, Cfb, fc = Sf[sel], Cf[sel], owner[sel]
    d = Cfb - C[fc]
    return Mesh(nCells=nCells, owner=owner[:nIF].copy(), neighbour=neighbour, Sf=np.ascontiguousarray(Sf[:nIF]), w=w, V=V, C=C,
                Cf=np.ascontiguousarray(Cf[:nIF]), patchType=np.array(ptype, np.int32), patchStart=np.array(pstart, np.int32),
                patchSize=np.array(psize, np.int32), patchNbrProc=np.array(pnbr, np.int32), faceCells=fc.astype(np.int32),
                Sfb=np.ascontiguousarray(Sfb), magSfb=_mag(Sfb), deltaCoeffsb=1.0 / _mag(d), wb=np.ones(len(fc)),
                Cfb=np.ascontiguousarray(Cfb), patchNames=names)


# ---- generators ---------------------------------------------------------------------------------------------------
_KUHN = [(0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)]


def split_box(nx, ny, nz, lengths=(1.0, 1.0, 1.0), kind="tet", patch_types=(0,) * 6, renumber=None, jitter=0.0, seed=1):
    """polyMesh (points, faces, owner, neighbour, patches) of a box whose hexes are split ("tet", "prism") or kept
    ("hex").  renumber=seed: random cell relabelling.  jitter: interior vertices moved by jitter*h (non-orthogonal,
    warped quads)."""
    vid = lambda i, j, k: i + (nx + 1) * (j + (ny + 1) * k)
    K, J, I = np.meshgrid(np.arange(nz + 1), np.arange(ny + 1), np.arange(nx + 1), indexing="ij")
    pts = np.stack([I * (lengths[0] / nx), J * (lengths[1] / ny), K * (lengths[2] / nz)], axis=-1).reshape(-1, 3).astype(np.float64)
    if jitter:
        rng = np.random.default_rng(seed)
        inner = ((I > 0) & (I < nx) & (J > 0) & (J < ny) & (K > 0) & (K < nz)).reshape(-1)
        h = min(lengths[0] / nx, lengths[1] / ny, lengths[2] / nz)
        pts[inner] += jitter * h * rng.uniform(-1, 1, (int(inner.sum()), 3))
    cells = []                                   # each: list of faces (vertex tuples, any orientation)
    for k in range(nz):
        for j in range(ny):
            for i in range(nx):
                v = lambda a, b, c: vid(i + a, j + b, k + c)
                if kind == "hex":
                    cells.append([(v(0, 0, 0), v(0, 1, 0), v(0, 1, 1), v(0, 0, 1)), (v(1, 0, 0), v(1, 1, 0), v(1, 1, 1), v(1, 0, 1)),
                                  (v(0, 0, 0), v(1, 0, 0), v(1, 0, 1), v(0, 0, 1)), (v(0, 1, 0), v(1, 1, 0), v(1, 1, 1), v(0, 1, 1)),
                                  (v(0, 0, 0), v(1, 0, 0), v(1, 1, 0), v(0, 1, 0)), (v(0, 0, 1), v(1, 0, 1), v(1, 1, 1), v(0, 1, 1))])
                elif kind == "prism":
                    for tri in (((0, 0), (1, 0), (1, 1)), ((0, 0), (1, 1), (0, 1))):
                        b = [v(a, c, 0) for a, c in tri]; t = [v(a, c, 1) for a, c in tri]
                        cells.append([tuple(b), tuple(t)] + [(b[e], b[(e + 1) % 3], t[(e + 1) % 3], t[e]) for e in range(3)])
                elif kind == "tet":
                    for perm in _KUHN:            # path (0,0,0) -> +e_perm[0] -> +e_perm[1] -> (1,1,1)
                        p = [0, 0, 0]
                        vs = [v(0, 0, 0)]
                        for ax in perm:
                            p[ax] = 1
                            vs.append(v(*p))
                        cells.append([tuple(vs[m] for m in range(4) if m != o) for o in range(4)])
                else:
                    raise ValueError(kind)
    nC = len(cells)
    label = np.arange(nC)
    if renumber is not None:
        label = np.random.default_rng(renumber).permutation(nC)
    # faces: key = sorted vertex tuple -> [(cell, vertex order)]
    table = {}
    for c, fl in enumerate(cells):
        for f in fl:
            table.setdefault(tuple(sorted(f)), []).append((int(label[c]), f))
    internal, boundary = [], []
    for key, uses in table.items():
        if len(uses) == 2:
            (c0, f0), (c1, _) = sorted(uses, key=lambda u: u[0])
            internal.append((c0, c1, f0))
        else:
            boundary.append((uses[0][0], uses[0][1]))
    internal.sort(key=lambda t: (t[0], t[1]))      # upper-triangular order
    # cell centroids (vertex averages are enough to orient the faces)
    cc = np.zeros((nC, 3))
    for c, fl in enumerate(cells):
        vs = sorted({x for f in fl for x in f})
        cc[label[c]] = pts[vs].mean(axis=0)

    def oriented(f, c_from, toward):
        S, Cf = face_geometry(pts, [f])
        return tuple(f) if float(S[0] @ (toward - Cf[0])) > 0 else tuple(reversed(f))
    faces, owner, neighbour = [], [], []
    for c0, c1, f in internal:
        faces.append(oriented(f, c0, cc[c1])); owner.append(c0); neighbour.append(c1)
    L = np.array(lengths, np.float64)
    groups = [[] for _ in range(6)]
    for c, f in boundary:
        S, Cf = face_geometry(pts, [f])
        f = tuple(f) if float(S[0] @ (Cf[0] - cc[c])) > 0 else tuple(reversed(f))
        ctr = pts[list(f)].mean(axis=0)
        d = int(np.argmax(np.maximum(np.abs(ctr) < 1e-12 * L.max(), np.abs(ctr - L) < 1e-12 * L.max())))
        side = 2 * d + (1 if abs(ctr[d] - L[d]) < 1e-12 * L.max() else 0)
        groups[side].append((c, f))
    patches = []
    names = ["xMin", "xMax", "yMin", "yMax", "zMin", "zMax"]
    for p in range(6):
        groups[p].sort(key=lambda t: (t[0], t[1]))
        patches.append((names[p], int(patch_types[p]), len(faces), len(groups[p])))
        for c, f in groups[p]:
            faces.append(f); owner.append(c)
    return pts, faces, np.array(owner, np.int32), np.array(neighbour, np.int32), patches


def split_box_mesh(*a, **kw) -> Mesh:
    return from_polymesh(*split_box(*a, **kw))


# ---- ASCII polyMesh files ---------------------------------------------------------------------------------------
_HEADER = """FoamFile
{{
    version     2.0;
    format      ascii;
    class       {cls};
    location    "constant/polyMesh";
    object      {obj};
}}
"""
_TYPE_OUT = {PATCH_GENERIC: "patch", PATCH_EMPTY: "empty", PATCH_WEDGE: "wedge", PATCH_PROCESSOR: "processor"}


def write_polymesh(dirname, points, faces, owner, neighbour, patches):
    os.makedirs(dirname, exist_ok=True)
    with open(os.path.join(dirname, "points"), "w") as fh:
        fh.write(_HEADER.format(cls="vectorField", obj="points") + f"\n{len(points)}\n(\n")
        for p in points:
            fh.write(f"({float(p[0])!r} {float(p[1])!r} {float(p[2])!r})\n")
        fh.write(")\n")
    with open(os.path.join(dirname, "faces"), "w") as fh:
        fh.write(_HEADER.format(cls="faceList", obj="faces") + f"\n{len(faces)}\n(\n")
        for f in faces:
            fh.write(f"{len(f)}(" + " ".join(str(int(x)) for x in f) + ")\n")
        fh.write(")\n")
    for name, arr in (("owner", owner), ("neighbour", neighbour)):
        with open(os.path.join(dirname, name), "w") as fh:
            fh.write(_HEADER.format(cls="labelList", obj=name) + f"\n{len(arr)}\n(\n" + "\n".join(str(int(x)) for x in arr) + "\n)\n")
    with open(os.path.join(dirname, "boundary"), "w") as fh:
        fh.write(_HEADER.format(cls="polyBoundaryMesh", obj="boundary") + f"\n{len(patches)}\n(\n")
        for pt in patches:
            name, typ, start, n = pt[:4]
            tn = typ if isinstance(typ, str) else _TYPE_OUT[int(typ)]
            fh.write(f"    {name}\n    {{\n        type            {tn};\n        nFaces          {int(n)};\n"
                     f"        startFace       {int(start)};\n")
            if len(pt) > 4 and int(pt[4]) >= 0:
                fh.write(f"        myProcNo        0;\n        neighbProcNo    {int(pt[4])};\n")
            fh.write("    }\n")
        fh.write(")\n")


def _body(path):
    """file content after the FoamFile header, comments stripped"""
    txt = open(path).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"//[^\n]*", "", txt)
    m = re.search(r"FoamFile\s*\{.*?\}", txt, flags=re.S)
    return txt[m.end():] if m else txt


def read_label_list(path) -> np.ndarray:
    """labelList file (owner, neighbour, cellDecomposition, ...): `N ( l0 l1 ... )` or the uniform form `N{l}`."""
    b = _body(path)
    m = re.search(r"(\d+)\s*\{\s*(-?\d+)\s*\}", b)
    if m and "(" not in b[:m.start()]:
        return np.full(int(m.group(1)), int(m.group(2)), np.int32)
    m = re.search(r"(\d+)\s*\(", b)
    n = int(m.group(1))
    vals = np.array(b[m.end():b.rindex(")")].split(), dtype=np.int64)
    assert len(vals) == n, (path, n, len(vals))
    return vals.astype(np.int32)


def read_polymesh(dirname):
    b = _body(os.path.join(dirname, "points"))
    m = re.search(r"(\d+)\s*\(", b)
    nums = np.array(re.findall(r"[-+0-9.eE]+", b[m.end():b.rindex(")")]), dtype=np.float64)
    points = nums.reshape(int(m.group(1)), 3)
    b = _body(os.path.join(dirname, "faces"))
    m = re.search(r"(\d+)\s*\(", b)
    faces = [tuple(int(x) for x in g.split()) for g in re.findall(r"\d+\s*\(([^()]*)\)", b[m.end():])]
    assert len(faces) == int(m.group(1))
    owner = read_label_list(os.path.join(dirname, "owner"))
    neighbour = read_label_list(os.path.join(dirname, "neighbour"))
    b = _body(os.path.join(dirname, "boundary"))
    patches = []
    for name, blk in re.findall(r"(\w+)\s*\{([^{}]*)\}", b):
        d = dict(re.findall(r"(\w+)\s+([^;]+);", blk))
        patches.append((name, d["type"].strip(), int(d["startFace"]), int(d["nFaces"]), int(d.get("neighbProcNo", -1))))
    return points, faces, owner, neighbour, patches
