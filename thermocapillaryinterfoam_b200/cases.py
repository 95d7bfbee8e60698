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
