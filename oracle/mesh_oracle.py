"""numpy ORACLE for mesh addressing and decomposition (test infrastructure only).

Independent restatement (vectorised numpy, no code shared with the C++ generator in
thermocapillaryinterfoam_b200/csrc/mesh_host.cpp) of the OpenFOAM conventions the hot path
relies on:

* single-block hex mesh, cell id = i + nx*(j + ny*k); internal faces in upper-triangular
  order (per cell ascending: +x, +y, +z face), owner < neighbour, Sf owner->neighbour;
  boundary patches xMin,xMax,yMin,yMax,zMin,zMax, faces in ascending face-cell order;
  `empty` patches carry no field faces (size 0);
* linear weights w = |Sf.(C_N-Cf)| / (|Sf.(Cf-C_P)| + |Sf.(C_N-Cf)|);
* decomposition into processor sub-meshes: local cells/faces keep global relative order,
  original patches first, then one processor patch per neighbour rank (ascending), faces
  in global face order on both sides, flipped where the local cell is the global
  neighbour; processor weights from unit-normal projections (processorFvPatch::makeWeights).

PARITY UNPINNED: no blockMesh / decomposePar exists in this environment to pin these
conventions against; the product generator is checked bit-exactly against this file.
"""
from __future__ import annotations

from dataclasses import dataclass, field
import numpy as np

PATCH_GENERIC, PATCH_EMPTY, PATCH_WEDGE, PATCH_PROCESSOR = 0, 1, 2, 3
VSMALL = 1e-300


@dataclass
class Mesh:
    nCells: int
    owner: np.ndarray        # int32 [nIF]
    neighbour: np.ndarray    # int32 [nIF]
    Sf: np.ndarray           # [nIF,3]
    w: np.ndarray            # [nIF]
    V: np.ndarray            # [nCells]
    C: np.ndarray            # [nCells,3]
    Cf: np.ndarray           # [nIF,3]
    patchType: np.ndarray    # int32 [nPatches]
    patchStart: np.ndarray   # int32 [nPatches] offset into boundary arrays
    patchSize: np.ndarray    # int32 [nPatches]
    patchNbrProc: np.ndarray  # int32 [nPatches] (-1 unless processor)
    faceCells: np.ndarray    # int32 [nBF]
    Sfb: np.ndarray          # [nBF,3]
    magSfb: np.ndarray       # [nBF]
    deltaCoeffsb: np.ndarray  # [nBF]
    wb: np.ndarray           # [nBF]
    Cfb: np.ndarray          # [nBF,3]
    patchNames: list = field(default_factory=list)
    # decomposition addressing (sub-meshes only)
    cellProcAddressing: np.ndarray | None = None   # local cell -> global cell
    faceProcAddressing: np.ndarray | None = None   # local internal face -> +-(global face+1)
    bfaceProcAddressing: np.ndarray | None = None  # local boundary face -> +-(global internal face+1) for proc faces, global bf index+1 else
    boundaryProcAddressing: np.ndarray | None = None  # local patch -> global patch (-1 for processor)

    @property
    def nIF(self):
        return int(self.owner.shape[0])

    @property
    def nBF(self):
        return int(self.faceCells.shape[0])

    @property
    def nPatches(self):
        return int(self.patchType.shape[0])


def dot3(a, b):
    """row-wise a.b summed left to right (x, y, z): the order of the C++ side; einsum may contract differently"""
    return a[:, 0] * b[:, 0] + a[:, 1] * b[:, 1] + a[:, 2] * b[:, 2]


def linear_weights(Sf, Cf, CP, CN):
    a = np.abs(dot3(Sf, CN - Cf))
    b = np.abs(dot3(Sf, Cf - CP))
    return a / (b + a)


def hex_mesh(xs, ys, zs, patch_types=(0, 0, 0, 0, 0, 0)) -> Mesh:
    """Rectilinear hex block from vertex coordinate arrays (len n+1 each)."""
    xs, ys, zs = (np.asarray(a, dtype=np.float64) for a in (xs, ys, zs))
    nx, ny, nz = len(xs) - 1, len(ys) - 1, len(zs) - 1
    nC = nx * ny * nz
    dx, dy, dz = np.diff(xs), np.diff(ys), np.diff(zs)
    xc, yc, zc = 0.5 * (xs[:-1] + xs[1:]), 0.5 * (ys[:-1] + ys[1:]), 0.5 * (zs[:-1] + zs[1:])
    K, J, I = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    cid = (I + nx * (J + ny * K)).astype(np.int64)
    C = np.stack([xc[I], yc[J], zc[K]], axis=-1).reshape(nC, 3)
    V = (dx[I] * dy[J] * dz[K]).reshape(nC)
    # internal faces: array [nz,ny,nx,3] with validity mask, flattened cell-major dir-minor
    valid = np.stack([I < nx - 1, J < ny - 1, K < nz - 1], axis=-1)
    off = np.array([1, nx, nx * ny], dtype=np.int64)
    own3 = np.broadcast_to(cid[..., None], valid.shape)
    nei3 = cid[..., None] + off
    owner = own3[valid].astype(np.int32)
    neighbour = nei3[valid].astype(np.int32)
    zero = np.zeros_like(V).reshape(nz, ny, nx)
    area = np.stack([dy[J] * dz[K], dx[I] * dz[K], dx[I] * dy[J]], axis=-1)
    Sf3 = np.zeros(valid.shape + (3,))
    for d in range(3):
        Sf3[..., d, d] = area[..., d]
    Sf = Sf3[valid]
    # face centres
    Cf3 = np.empty(valid.shape + (3,))
    xf, yf, zf = xs[1:], ys[1:], zs[1:]
    Cf3[..., 0, :] = np.stack([xf[I], yc[J], zc[K]], axis=-1)
    Cf3[..., 1, :] = np.stack([xc[I], yf[J], zc[K]], axis=-1)
    Cf3[..., 2, :] = np.stack([xc[I], yc[J], zf[K]], axis=-1)
    Cf = Cf3[valid]
    w = linear_weights(Sf, Cf, C[owner], C[neighbour])
    del zero
    # boundary patches
    sel = [I == 0, I == nx - 1, J == 0, J == ny - 1, K == 0, K == nz - 1]
    names = ["xMin", "xMax", "yMin", "yMax", "zMin", "zMax"]
    pstart, psize, fcs, Sfbs, Cfbs = [], [], [], [], []
    pos = 0
    for p in range(6):
        d, hi = p // 2, p % 2
        if patch_types[p] == PATCH_EMPTY:
            pstart.append(pos); psize.append(0)
            continue
        m = sel[p].reshape(-1)
        fc = np.nonzero(m)[0]           # ascending cell order
        ii, jj, kk = fc % nx, (fc // nx) % ny, fc // (nx * ny)
        a = [dy[jj] * dz[kk], dx[ii] * dz[kk], dx[ii] * dy[jj]][d]
        S = np.zeros((len(fc), 3))
        S[:, d] = a if hi else -a
        cf = np.stack([xc[ii], yc[jj], zc[kk]], axis=-1)
        cf[:, d] = [xs, ys, zs][d][-1 if hi else 0]
        pstart.append(pos); psize.append(len(fc)); pos += len(fc)
        fcs.append(fc); Sfbs.append(S); Cfbs.append(cf)
    faceCells = np.concatenate(fcs).astype(np.int32) if fcs else np.zeros(0, np.int32)
    Sfb = np.concatenate(Sfbs) if Sfbs else np.zeros((0, 3))
    Cfb = np.concatenate(Cfbs) if Cfbs else np.zeros((0, 3))
    magSfb = np.sqrt(Sfb[:, 0] * Sfb[:, 0] + Sfb[:, 1] * Sfb[:, 1] + Sfb[:, 2] * Sfb[:, 2])
    dlt = Cfb - C[faceCells]
    deltaCoeffsb = 1.0 / np.sqrt(dlt[:, 0] * dlt[:, 0] + dlt[:, 1] * dlt[:, 1] + dlt[:, 2] * dlt[:, 2])
    return Mesh(
        nCells=nC, owner=owner, neighbour=neighbour, Sf=np.ascontiguousarray(Sf), w=w, V=V, C=C,
        Cf=np.ascontiguousarray(Cf), patchType=np.array(patch_types, np.int32),
        patchStart=np.array(pstart, np.int32), patchSize=np.array(psize, np.int32),
        patchNbrProc=np.full(6, -1, np.int32), faceCells=faceCells, Sfb=Sfb, magSfb=magSfb,
        deltaCoeffsb=deltaCoeffsb, wb=np.ones(len(faceCells)), Cfb=Cfb, patchNames=names)


def box_mesh(nx, ny, nz, lengths=(1.0, 1.0, 1.0), patch_types=(0, 0, 0, 0, 0, 0), grading=None) -> Mesh:
    """Uniform (or geometrically graded) box; vertices at L*i/n."""
    def verts(n, L, r):
        t = np.arange(n + 1, dtype=np.float64) / n
        if r is not None and r != 1.0:
            t = (np.power(r, t) - 1.0) / (r - 1.0)
        return L * t
    g = grading or (None, None, None)
    return hex_mesh(verts(nx, lengths[0], g[0]), verts(ny, lengths[1], g[1]), verts(nz, lengths[2], g[2]), patch_types)


def simple_decomposition(nx, ny, nz, px, py, pz) -> np.ndarray:
    """OpenFOAM `simple (px py pz)` on a uniform box: contiguous blocks, proc = bi + px*(bj + py*bk)."""
    i = np.arange(nx) * px // nx
    j = np.arange(ny) * py // ny
    k = np.arange(nz) * pz // nz
    K, J, I = np.meshgrid(k, j, i, indexing="ij")
    return (I + px * (J + py * K)).reshape(-1).astype(np.int32)


def decompose(g: Mesh, cellProc: np.ndarray, nProcs: int) -> list[Mesh]:
    """Processor sub-meshes of `g` for the cell->rank map `cellProc` (decomposePar conventions)."""
    cellProc = np.asarray(cellProc, np.int32)
    po, pn = cellProc[g.owner], cellProc[g.neighbour]
    subs = []
    for r in range(nProcs):
        cells = np.nonzero(cellProc == r)[0]
        g2l = np.full(g.nCells, -1, np.int64)
        g2l[cells] = np.arange(len(cells))
        fint = np.nonzero((po == r) & (pn == r))[0]
        owner = g2l[g.owner[fint]].astype(np.int32)
        neighbour = g2l[g.neighbour[fint]].astype(np.int32)
        ptype, pstart, psize, pnbr, bpa = [], [], [], [], []
        fcs, Sfb, magSfb, dcb, wb, Cfb, bfa = [], [], [], [], [], [], []
        pos = 0
        for p in range(g.nPatches):
            s, n = int(g.patchStart[p]), int(g.patchSize[p])
            idx = s + np.nonzero(cellProc[g.faceCells[s:s + n]] == r)[0]
            ptype.append(int(g.patchType[p])); pstart.append(pos); psize.append(len(idx)); pnbr.append(-1); bpa.append(p)
            pos += len(idx)
            fcs.append(g2l[g.faceCells[idx]]); Sfb.append(g.Sfb[idx]); magSfb.append(g.magSfb[idx])
            dcb.append(g.deltaCoeffsb[idx]); wb.append(g.wb[idx]); Cfb.append(g.Cfb[idx]); bfa.append(idx + 1)
        for e in range(nProcs):
            if e == r:
                continue
            f = np.nonzero(((po == r) & (pn == e)) | ((po == e) & (pn == r)))[0]   # global face order
            if len(f) == 0:
                continue
            flip = pn[f] == r                     # local cell is the global neighbour
            lc_g = np.where(flip, g.neighbour[f], g.owner[f])
            nc_g = np.where(flip, g.owner[f], g.neighbour[f])
            S = np.where(flip[:, None], -g.Sf[f], g.Sf[f])
            mag = np.sqrt(S[:, 0] * S[:, 0] + S[:, 1] * S[:, 1] + S[:, 2] * S[:, 2])
            nf = S / mag[:, None]
            dP = dot3(nf, g.Cf[f] - g.C[lc_g])
            SN = -S
            nN = SN / (mag[:, None] + VSMALL)
            dN = dot3(nN, g.Cf[f] - g.C[nc_g])
            wgt = dN / (dP + dN)
            dl = g.C[nc_g] - g.C[lc_g]
            ptype.append(PATCH_PROCESSOR); pstart.append(pos); psize.append(len(f)); pnbr.append(e); bpa.append(-1)
            pos += len(f)
            fcs.append(g2l[lc_g]); Sfb.append(S); magSfb.append(mag)
            dcb.append(1.0 / np.sqrt(dl[:, 0] * dl[:, 0] + dl[:, 1] * dl[:, 1] + dl[:, 2] * dl[:, 2]))
            wb.append(wgt); Cfb.append(g.Cf[f]); bfa.append(np.where(flip, -(f + 1), f + 1))
        cat = lambda xs, shape=(0,): (np.concatenate(xs) if len(xs) else np.zeros(shape))
        subs.append(Mesh(
            nCells=len(cells), owner=owner, neighbour=neighbour, Sf=np.ascontiguousarray(g.Sf[fint]),
            w=g.w[fint].copy(), V=g.V[cells].copy(), C=g.C[cells].copy(), Cf=g.Cf[fint].copy(),
            patchType=np.array(ptype, np.int32), patchStart=np.array(pstart, np.int32),
            patchSize=np.array(psize, np.int32), patchNbrProc=np.array(pnbr, np.int32),
            faceCells=cat(fcs).astype(np.int32), Sfb=np.ascontiguousarray(cat(Sfb, (0, 3))).reshape(-1, 3),
            magSfb=cat(magSfb), deltaCoeffsb=cat(dcb), wb=cat(wb), Cfb=np.ascontiguousarray(cat(Cfb, (0, 3))).reshape(-1, 3),
            patchNames=list(g.patchNames) + [f"procBoundary{r}to{e}" for e in pnbr if e >= 0],
            cellProcAddressing=cells.astype(np.int32), faceProcAddressing=(fint + 1).astype(np.int32),
            bfaceProcAddressing=cat(bfa).astype(np.int64), boundaryProcAddressing=np.array(bpa, np.int32)))
    return subs


def ldu_csr(owner, neighbour, nCells):
    """lduAddressing::ownerStartAddr / losortAddr / losortStartAddr."""
    ownerStart = np.zeros(nCells + 1, np.int32)
    np.add.at(ownerStart, owner.astype(np.int64) + 1, 1)
    ownerStart = np.cumsum(ownerStart).astype(np.int32)
    losort = np.argsort(neighbour, kind="stable").astype(np.int32)
    losortStart = np.zeros(nCells + 1, np.int32)
    np.add.at(losortStart, neighbour.astype(np.int64) + 1, 1)
    losortStart = np.cumsum(losortStart).astype(np.int32)
    return ownerStart, losort, losortStart
