"""Synthetic inputs of the BASELINE.json configs (SURVEY.md 8(d) "Synthetic inputs").

Harness code (numpy): builds the frozen fields the hot path is timed and checked on.  Every
field is an analytic function of position plus hash noise keyed on the GLOBAL cell id, so a
processor sub-mesh can build its own part without the global arrays and decomposed runs see
exactly the serial values.
"""
from __future__ import annotations

from dataclasses import dataclass
import numpy as np

from .abi import PATCH_PROCESSOR, BC_ZERO_GRADIENT

T0, TGRAD, TNOISE = 300.0, 100.0, 1e-3
SIGMA0 = (0.07, 0.05, 0.03, 0.04, 0.06, 0.02)
GAMMA = 1e-4
UMAG = 0.1
SEED = 20261104


def _hash_uniform(gid: np.ndarray, seed: int) -> np.ndarray:
    """splitmix64 of (global id, seed) -> U(-1,1)."""
    z = (gid.astype(np.uint64) + np.uint64(seed)) * np.uint64(0x9E3779B97F4A7C15)
    z ^= z >> np.uint64(30)
    z *= np.uint64(0xBF58476D1CE4E5B9)
    z ^= z >> np.uint64(27)
    z *= np.uint64(0x94D049BB133111EB)
    z ^= z >> np.uint64(31)
    return (z >> np.uint64(11)).astype(np.float64) * (2.0 / 9007199254740992.0) - 1.0


@dataclass
class CaseSpec:
    name: str
    n: tuple            # (nx, ny, nz)
    nPhases: int
    sigma: list         # per pair spec for set_sigma
    two_d: bool
    config_id: int

    @property
    def nCells(self):
        return self.n[0] * self.n[1] * self.n[2]

    @property
    def dx(self):
        return 1.0 / self.n[0]

    @property
    def lengths(self):
        return (1.0, 1.0 * self.n[1] / self.n[0], self.dx if self.two_d else 1.0 * self.n[2] / self.n[0])

    @property
    def patch_types(self):
        from .abi import PATCH_EMPTY, PATCH_GENERIC
        return (0, 0, 0, 0, PATCH_EMPTY, PATCH_EMPTY) if self.two_d else (PATCH_GENERIC,) * 6

    @property
    def deltaT(self):
        return 0.25 * self.dx / UMAG       # CFL 0.25


def sigma_linear(s0):
    """sigma = s0 - GAMMA*(T - T0) as Function1 polynomial ((a e)...)."""
    return ("polynomial", [(s0 + GAMMA * T0, 0.0), (-GAMMA, 1.0)])


def case(config: int, n: int | None = None) -> CaseSpec:
    """BASELINE.json configs[config-1]; `n` overrides the resolution (parity tests use small n)."""
    if config == 1:
        n = n or 64
        return CaseSpec("static_droplet_2d", (n, n, 1), 2, [("constant", SIGMA0[0])], True, 1)
    if config == 2:
        n = n or 512
        return CaseSpec("thermocapillary_migration_2d", (n, n, 1), 2, [sigma_linear(SIGMA0[0])], True, 2)
    if config == 3:
        n = n or 128
        return CaseSpec("marangoni_cavity_3d", (n, n, n), 2, [sigma_linear(SIGMA0[0])], False, 3)
    if config == 4:
        n = n or 200
        return CaseSpec("melt_pool_3d_3phase", (n, n, n), 3, [sigma_linear(s) for s in SIGMA0[:3]], False, 4)
    if config == 5:
        n = n or 400
        return CaseSpec("synthetic_3phase_box", (n, n, n), 3, [sigma_linear(s) for s in SIGMA0[:3]], False, 5)
    raise ValueError(f"unknown config {config}")


def alphas_at(spec: CaseSpec, X: np.ndarray) -> list[np.ndarray]:
    """Phase fractions at points X[n,3]: smoothed circles/spheres, width 1.5 dx."""
    wdt = 1.5 * spec.dx
    zc = 0.5 * spec.lengths[2]

    def blob(cx, R):
        d = X - np.array([cx, 0.5, zc])
        if spec.two_d:
            d = d[:, :2]
        r = np.sqrt(np.einsum("ij,ij->i", d, d))
        return 0.5 * (1.0 - np.tanh((r - R) / wdt))
    if spec.nPhases == 2:
        a1 = blob(0.5, 0.25)
        return [a1, 1.0 - a1]
    a1, a2 = blob(0.35, 0.2), blob(0.65, 0.2)
    a3 = np.maximum(1.0 - a1 - a2, 0.0)
    s = a1 + a2 + a3
    out = [a1 / s, a2 / s, a3 / s]
    if spec.nPhases == 4:       # split the carrier phase by height
        f = 0.5 * (1.0 + np.tanh((X[:, 1] - 0.5) / wdt))
        out = [out[0], out[1], out[2] * f, out[2] * (1.0 - f)]
    return out


def temperature_at(spec: CaseSpec, X: np.ndarray, gid: np.ndarray | None) -> np.ndarray:
    T = T0 + TGRAD * X[:, 0]
    if gid is not None:
        T = T + TNOISE * _hash_uniform(gid, SEED + spec.config_id)
    return T


def velocity_at(X: np.ndarray) -> np.ndarray:
    u = np.zeros_like(X)
    u[:, 0] = UMAG * np.sin(np.pi * X[:, 0]) * np.cos(np.pi * X[:, 1])
    u[:, 1] = -UMAG * np.cos(np.pi * X[:, 0]) * np.sin(np.pi * X[:, 1])
    return u


def synthetic_fields(spec: CaseSpec, mesh) -> dict:
    """Frozen inputs of one stage on `mesh` (global mesh or processor sub-mesh).

    mesh needs: C, Cf, Cfb, Sf, Sfb, owner, neighbour, w, wb, faceCells, patchType, patchStart, patchSize and,
    for sub-meshes, cellGlobal [nCells], nbrC [nBF,3], nbrGlobal [nBF] (valid on processor faces).
    Wall patches: alpha zeroGradient, T_b evaluated on the wall (fixedValue), phi_b = 0.
    """
    P = spec.nPhases
    nC, nBF = mesh.nCells, len(mesh.faceCells)
    gid = getattr(mesh, "cellGlobal", None)
    if gid is None:
        gid = np.arange(nC, dtype=np.int64)
    ai = alphas_at(spec, mesh.C)
    Ti = temperature_at(spec, mesh.C, gid)
    coupled = np.zeros(nBF, bool)
    for p in range(len(mesh.patchType)):
        if mesh.patchType[p] == PATCH_PROCESSOR:
            coupled[mesh.patchStart[p]:mesh.patchStart[p] + mesh.patchSize[p]] = True
    fc = np.asarray(mesh.faceCells, dtype=np.int64)
    ab = [a[fc].copy() for a in ai]                      # zeroGradient
    Tb = temperature_at(spec, mesh.Cfb, None) if nBF else np.zeros(0)
    phib = np.zeros(nBF)
    if coupled.any():
        k = np.nonzero(coupled)[0]
        an = alphas_at(spec, mesh.nbrC[k])
        for p in range(P):
            ab[p][k] = an[p]
        Tb[k] = temperature_at(spec, mesh.nbrC[k], np.asarray(mesh.nbrGlobal)[k])
        phib[k] = np.einsum("ij,ij->i", mesh.Sfb[k], velocity_at(mesh.Cfb[k]))
    phii = np.einsum("ij,ij->i", mesh.Sf, velocity_at(mesh.Cf))
    o, n = np.asarray(mesh.owner, np.int64), np.asarray(mesh.neighbour, np.int64)
    pai, pab = [], []
    for p in range(P):                                   # HO flux = phi * linear(alpha)
        pai.append(phii * (mesh.w * (ai[p][o] - ai[p][n]) + ai[p][n]))
        b = phib * ab[p]
        if coupled.any():
            k = np.nonzero(coupled)[0]
            b[k] = phib[k] * (mesh.wb[k] * ai[p][fc[k]] + (1.0 - mesh.wb[k]) * ab[p][k])
        pab.append(b)
    bc = np.full((P, len(mesh.patchType)), BC_ZERO_GRADIENT, np.int32)
    deltab = (np.asarray(mesh.Cfb) - np.asarray(mesh.C)[fc]) if nBF else np.zeros((0, 3))      # fvPatch::delta()
    if coupled.any():
        k = np.nonzero(coupled)[0]
        deltab[k] = np.asarray(mesh.nbrC)[k] - np.asarray(mesh.C)[fc[k]]
    return dict(C=np.array(mesh.C), deltab=deltab, alpha_i=ai, alpha_b=ab, T_i=Ti, T_b=Tb, phi_i=phii, phi_b=phib, phiAlpha_i=pai, phiAlpha_b=pab,
                alphaBc=bc, sigma=spec.sigma, deltaT=spec.deltaT)


def load_case(ctx, spec: CaseSpec, f: dict):
    """Upload the frozen inputs into a Context."""
    P = spec.nPhases
    q = 0
    for p1 in range(P):
        for p2 in range(p1 + 1, P):
            ctx.set_sigma(p1, p2, f["sigma"][q])
            q += 1
    for p in range(P):
        ctx.set_alpha_bc(p, f["alphaBc"][p])
        ctx.set_alpha(p, f["alpha_i"][p], f["alpha_b"][p])
        ctx.set_phi_alpha(p, f["phiAlpha_i"][p], f["phiAlpha_b"][p])
    ctx.set_temperature(f["T_i"], f["T_b"])
    ctx.set_phi(f["phi_i"], f["phi_b"])
    ctx.store_old_alphas()


# ---- laserDTRM inputs -----------------------------------------------------------------------
def box_polymesh(mesh, xs, ys, zs) -> dict:
    """mesh.points(), mesh.faces() and tetBasePtIs of a rectilinear hex box whose vertex coordinates are xs, ys, zs
    (what blockMesh would write; the shim reads them off the polyMesh).  Faces are ordered internal, then boundary;
    the vertices of a face turn right-handed about its area vector.  Vertex id = i + (nx+1)*(j + (ny+1)*k)."""
    xs, ys, zs = (np.asarray(v, np.float64) for v in (xs, ys, zs))
    co = (xs, ys, zs)
    nvx, nvy = len(xs), len(ys)
    Cf = np.concatenate([np.asarray(mesh.Cf).reshape(-1, 3), np.asarray(mesh.Cfb).reshape(-1, 3)])
    Sf = np.concatenate([np.asarray(mesh.Sf).reshape(-1, 3), np.asarray(mesh.Sfb).reshape(-1, 3)])
    nF = len(Cf)
    ax = np.argmax(np.abs(Sf), axis=1)
    idx = np.zeros((nF, 3), np.int64)
    for d in range(3):
        mid = 0.5 * (co[d][1:] + co[d][:-1])
        plane = np.searchsorted(mid, Cf[:, d])                       # nearest vertex plane
        cellidx = np.clip(np.searchsorted(co[d], Cf[:, d]) - 1, 0, len(co[d]) - 2)
        idx[:, d] = np.where(ax == d, plane, cellidx)
    # the two in-plane axes (e1, e2) with e1 x e2 = +normal
    e1 = (ax + 1) % 3
    e2 = (ax + 2) % 3
    corners = np.zeros((nF, 4, 3), np.int64)
    for k, (s1, s2) in enumerate(((0, 0), (1, 0), (1, 1), (0, 1))):
        c = idx.copy()
        c[np.arange(nF), e1] += s1
        c[np.arange(nF), e2] += s2
        corners[:, k] = c
    neg = Sf[np.arange(nF), ax] < 0
    corners[neg] = corners[neg][:, [0, 3, 2, 1]]
    vid = corners[..., 0] + nvx * (corners[..., 1] + nvy * corners[..., 2])
    Z, Y, X = np.meshgrid(zs, ys, xs, indexing="ij")
    points = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)
    ext = np.array([xs[-1] - xs[0], ys[-1] - ys[0], zs[-1] - zs[0]])
    return dict(points=points, fpStart=(4 * np.arange(nF + 1)).astype(np.int32), fpPts=vid.reshape(-1).astype(np.int32),
                tetBasePt=np.zeros(nF, np.int32), boundsMag=float(np.sqrt(np.dot(ext, ext))))


def find_cells_box(n, xs, ys, zs, pts) -> np.ndarray:
    """mesh.findCell on the rectilinear box (cell id = i + nx*(j + ny*k)); -1 outside."""
    pts = np.asarray(pts, np.float64).reshape(-1, 3)
    ijk, ok = [], np.ones(len(pts), bool)
    for d, v in enumerate((xs, ys, zs)):
        i = np.searchsorted(np.asarray(v), pts[:, d], side="right") - 1
        ok &= (i >= 0) & (i < n[d])
        ijk.append(np.clip(i, 0, n[d] - 1))
    return np.where(ok, ijk[0] + n[0] * (ijk[1] + n[1] * ijk[2]), -1).astype(np.int32)


def laser_seeds(lPosition, lDir, rArea, nr, nTheta, focalLaserRadius, maxTrackLength, mode="Gaussian", power=1.0,
                sigma=1.0, I0=0.0):
    """The polar disk of DTRM seeds, laserDTRM::initialise() (laserDTRM.C:199-283): p0, p1, Ip, dA [nr*nTheta].
    In the drop-in the solver's own initialise() produces these (it also owns findCell and the Function1s)."""
    SMALL = 1e-15
    lPosition, lDir, rArea = (np.asarray(v, np.float64) for v in (lPosition, lDir, rArea))
    dr, dTheta = focalLaserRadius / nr, 2.0 * np.pi / nTheta
    r1 = SMALL + dr * np.arange(nr)
    r2 = r1 + dr
    rP = (r1 + r2) / 2
    th1 = SMALL + dTheta * np.arange(nTheta)
    th2 = th1 + dTheta
    thP = (th1 + th2) / 2.0
    # quaternion(lDir, thetaP).R() & (rP*rArea): Rodrigues rotation about the unit axis lDir
    c, s = np.cos(thP)[None, :, None], np.sin(thP)[None, :, None]
    v = (rP[:, None, None] * rArea[None, None, :])
    rot = v * c + np.cross(lDir, v) * s + lDir * np.einsum("k,ijk->ij", lDir, v)[..., None] * (1 - c)
    p0 = (lPosition + rot).reshape(-1, 3)
    p1 = p0 + 0.5 * maxTrackLength * lDir
    if mode == "GaussianPeak":
        Ip = I0 * np.exp(-2.0 * rP ** 2 / sigma ** 2)
    elif mode == "Gaussian":
        Ip = power / (2.0 * np.pi * sigma ** 2) * np.exp(-rP ** 2 / 2.0 / sigma ** 2)
    elif mode == "uniform":
        Ip = np.full(nr, power / (np.pi * focalLaserRadius ** 2))
    else:
        raise ValueError(mode)
    dA = ((r2 ** 2 - r1 ** 2)[:, None] * (th2 - th1)[None, :] / 2.0)
    return p0, p1, np.repeat(Ip, nTheta), dA.reshape(-1)


def box_block_polymesh(n, xs, ys, zs, procs, rank, mesh) -> dict:
    """box_polymesh of processor sub-mesh `rank` of the `simple procs` decomposition (mesh: that sub-mesh), plus
    what tcif_laser_set_shared_points needs: the local points on processor patches and their global labels."""
    co = (np.asarray(xs, np.float64), np.asarray(ys, np.float64), np.asarray(zs, np.float64))
    b = (rank % procs[0], (rank // procs[0]) % procs[1], rank // (procs[0] * procs[1]))
    lo, sub = [], []
    for d in range(3):
        own = np.nonzero(np.arange(n[d]) * procs[d] // n[d] == b[d])[0]
        lo.append(int(own[0]))
        sub.append(co[d][own[0]:own[-1] + 2])
    geom = box_polymesh(mesh, *sub)
    nvx, nvy = len(sub[0]), len(sub[1])
    pid = np.arange(len(geom["points"]))
    i, j, k = pid % nvx, (pid // nvx) % nvy, pid // (nvx * nvy)
    gid = (i + lo[0]) + (n[0] + 1) * ((j + lo[1]) + (n[1] + 1) * (k + lo[2]))
    nIF = len(np.asarray(mesh.owner))
    fp = geom["fpPts"].reshape(-1, 4)
    shared = []
    for p in range(len(mesh.patchType)):
        if mesh.patchType[p] == 3:
            s, e = int(mesh.patchStart[p]), int(mesh.patchStart[p] + mesh.patchSize[p])
            shared.append(fp[nIF + s:nIF + e].reshape(-1))
    shared = np.unique(np.concatenate(shared)) if shared else np.zeros(0, np.int64)
    geom["sharedLocal"] = shared.astype(np.int32)
    geom["sharedGlobal"] = gid[shared].astype(np.int64)
    geom["pointGlobal"] = gid.astype(np.int64)
    ext = np.array([co[d][-1] - co[d][0] for d in range(3)])
    geom["boundsMag"] = float(np.sqrt(np.dot(ext, ext)))          # mesh.bounds() is global (reduced over ranks)
    return geom
