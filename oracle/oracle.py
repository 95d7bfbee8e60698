"""ctypes binding of the CPU ORACLE (oracle/tcif_oracle.c).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs; never by the product package.  PARITY UNPINNED (see tcif_oracle.h)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)

BC_ZERO_GRADIENT, BC_FIXED_VALUE = 0, 1
SIGMA_CONSTANT, SIGMA_POLYNOMIAL, SIGMA_TABLE = 0, 1, 2
BEXT_GLOBAL, BEXT_PATCHVALUE = 0, 1


class OrcMesh(C.Structure):
    _fields_ = [("nCells", C.c_int32), ("nIF", C.c_int32), ("nPatches", C.c_int32), ("nBF", C.c_int32),
                ("owner", c_ip), ("neighbour", c_ip), ("Sf", c_dp), ("w", c_dp), ("V", c_dp),
                ("patchType", c_ip), ("patchStart", c_ip), ("patchSize", c_ip), ("patchNbrProc", c_ip),
                ("faceCells", c_ip), ("Sfb", c_dp), ("magSfb", c_dp), ("deltaCoeffsb", c_dp), ("wb", c_dp),
                ("C", c_dp), ("deltab", c_dp)]


class OrcSigma(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n", C.c_int32), ("c", C.c_double * 64)]


class OrcAlphaCtl(C.Structure):
    _fields_ = [("nAlphaCorr", C.c_int32), ("nAlphaSubCycles", C.c_int32), ("schemePhi", C.c_int32),
                ("schemePhir", C.c_int32), ("implicitPredictor", C.c_int32), ("predictorTol", C.c_double),
                ("predictorMaxIter", C.c_int32)]


SCHEMES = {"upwind": 0, "linear": 1, "vanLeer": 2}


def alpha_ctl(nAlphaCorr=1, nAlphaSubCycles=1, schemePhi="vanLeer", schemePhir="linear", implicitPredictor=True,
              predictorTol=1e-12, predictorMaxIter=200):
    return OrcAlphaCtl(nAlphaCorr, nAlphaSubCycles, SCHEMES[schemePhi], SCHEMES[schemePhir], int(implicitPredictor),
                       predictorTol, predictorMaxIter)


class OrcMulesCfg(C.Structure):
    _fields_ = [("nLimiterIter", C.c_int32), ("smoothLimiter", C.c_double), ("extremaCoeff", C.c_double),
                ("boundaryExtrema", C.c_int32)]


HALO_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, c_dp, c_dp)
REDUCE_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, c_dp)


class OrcComm(C.Structure):
    _fields_ = [("user", C.c_void_p), ("halo", HALO_FN), ("reduce", REDUCE_FN)]


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libtcif_oracle.so")
    src = [os.path.join(_HERE, f) for f in ("tcif_oracle.c", "tcif_oracle.h", "tcif_oracle_alpha.inc", "tcif_oracle_mass.inc", "tcif_oracle_species.inc", "tcif_oracle_laser.inc")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src if os.path.exists(s)):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_deltaN.restype = C.c_double
        _LIB.orc_sigma_eval.restype = C.c_double
        _LIB.orc_stage_threads.restype = C.c_double
        _LIB.orc_upwind_predictor.restype = C.c_int
        _LIB.orc_laser_calculate.restype = C.c_int
        _LIB.orc_find_cell.restype = C.c_int
    return _LIB


def _d(a):
    return a.ctypes.data_as(c_dp) if a is not None else None


def _i(a):
    return a.ctypes.data_as(c_ip) if a is not None else None


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class MeshHandle:
    """Keeps contiguous copies alive and exposes the orc_mesh struct."""

    def __init__(self, m):
        self.m = m
        self.keep = dict(
            owner=i32(m.owner), neighbour=i32(m.neighbour), Sf=f64(m.Sf), w=f64(m.w), V=f64(m.V),
            patchType=i32(m.patchType), patchStart=i32(m.patchStart), patchSize=i32(m.patchSize),
            patchNbrProc=i32(m.patchNbrProc), faceCells=i32(m.faceCells), Sfb=f64(m.Sfb), magSfb=f64(m.magSfb),
            deltaCoeffsb=f64(m.deltaCoeffsb), wb=f64(m.wb))
        Cc = getattr(m, "C", None)
        if Cc is not None:
            # fvPatch::delta(): own centre -> neighbour centre on processor faces (nbrC), -> face centre elsewhere
            fc = np.asarray(m.faceCells, dtype=np.int64)
            db = f64(m.Cfb) - f64(Cc)[fc] if len(fc) else np.zeros((0, 3))
            nbrC = getattr(m, "nbrC", None)
            if nbrC is not None and len(fc):
                for p in range(len(m.patchType)):
                    if m.patchType[p] == 3:
                        sl = slice(int(m.patchStart[p]), int(m.patchStart[p] + m.patchSize[p]))
                        db[sl] = f64(nbrC)[sl] - f64(Cc)[fc[sl]]
            self.keep["C"] = f64(Cc); self.keep["deltab"] = f64(db)
        k = self.keep
        self.s = OrcMesh(m.nCells, len(k["owner"]), len(k["patchType"]), len(k["faceCells"]),
                         _i(k["owner"]), _i(k["neighbour"]), _d(k["Sf"]), _d(k["w"]), _d(k["V"]),
                         _i(k["patchType"]), _i(k["patchStart"]), _i(k["patchSize"]), _i(k["patchNbrProc"]),
                         _i(k["faceCells"]), _d(k["Sfb"]), _d(k["magSfb"]), _d(k["deltaCoeffsb"]), _d(k["wb"]),
                         _d(k.get("C")), _d(k.get("deltab")))

    @property
    def ref(self):
        return C.byref(self.s)


def make_sigma(spec) -> OrcSigma:
    """spec: ("constant", s0) | ("polynomial", [(a,e),...]) | ("table", [(T,s),...])."""
    s = OrcSigma()
    kind, val = spec
    if kind == "constant":
        s.kind, s.n = SIGMA_CONSTANT, 1
        s.c[0] = float(val)
    else:
        s.kind = SIGMA_POLYNOMIAL if kind == "polynomial" else SIGMA_TABLE
        s.n = len(val)
        for i, (a, b) in enumerate(val):
            s.c[2 * i], s.c[2 * i + 1] = float(a), float(b)
    return s


def sigma_array(specs):
    arr = (OrcSigma * max(1, len(specs)))()
    for i, sp in enumerate(specs):
        arr[i] = make_sigma(sp)
    return arr


def mules_cfg(nLimiterIter=3, smoothLimiter=0.0, extremaCoeff=0.0, boundaryExtrema=BEXT_GLOBAL):
    return OrcMulesCfg(nLimiterIter, smoothLimiter, extremaCoeff, boundaryExtrema)


def _pp(arrs):
    """list of float64 arrays -> double** """
    p = (c_dp * len(arrs))()
    for k, a in enumerate(arrs):
        p[k] = _d(a)
    return p


def make_comm(halo=None, reduce=None):
    """halo(ncomp, send[nc,ncomp]) -> recv array; reduce(op, vals) -> vals. Returns (OrcComm, keepalive)."""
    def _h(user, ncomp, nc, send, recv):
        s = np.ctypeslib.as_array(send, shape=(nc, ncomp))
        r = np.ctypeslib.as_array(recv, shape=(nc, ncomp))
        r[...] = halo(ncomp, s.copy())

    def _r(user, op, n, vals):
        v = np.ctypeslib.as_array(vals, shape=(n,))
        v[...] = reduce(op, v.copy())
    hf, rf = HALO_FN(_h), REDUCE_FN(_r)
    return OrcComm(None, hf, rf), (hf, rf)


def grad(m, ai, ab, bc, comm=None):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    gi, gb = np.zeros((h.s.nCells, 3)), np.zeros((h.s.nBF, 3))
    ai, ab, bc = f64(ai), f64(ab), i32(bc)
    lib().orc_grad(h.ref, C.byref(comm) if comm else None, _d(ai), _d(ab), _i(bc), _d(gi), _d(gb))
    return gi, gb


def interpolate(m, vi, vb):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    vi, vb = f64(vi), f64(vb)
    ncomp = 1 if vi.ndim == 1 else vi.shape[1]
    si = np.zeros((h.s.nIF,) + vi.shape[1:]); sb = np.zeros((h.s.nBF,) + vi.shape[1:])
    lib().orc_interpolate(h.ref, ncomp, _d(vi), _d(vb), _d(si), _d(sb))
    return si, sb


def div_tensor(m, Ti, Tb, comm=None):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    Ti, Tb = f64(Ti), f64(Tb)
    di, db = np.zeros((h.s.nCells, 3)), np.zeros((h.s.nBF, 3))
    lib().orc_div_tensor(h.ref, C.byref(comm) if comm else None, _d(Ti), _d(Tb), _d(di), _d(db))
    return di, db


def sigma_eval(spec, T):
    s = make_sigma(spec)
    return lib().orc_sigma_eval(C.byref(s), C.c_double(T))


def deltaN(m, comm=None):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    return lib().orc_deltaN(h.ref, C.byref(comm) if comm else None)


def div_capillary_stress(m, alphas_i, alphas_b, alphaBc, sigma_specs, Ti, Tb, comm=None):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    P = len(alphas_i)
    ai = [f64(a) for a in alphas_i]; ab = [f64(a) for a in alphas_b]
    bc = i32(alphaBc).reshape(P, -1)
    Ti, Tb = f64(Ti), f64(Tb)
    Fi, Fb = np.zeros((h.s.nCells, 3)), np.zeros((h.s.nBF, 3))
    sg = sigma_array(sigma_specs)
    lib().orc_div_capillary_stress(h.ref, C.byref(comm) if comm else None, P, _pp(ai), _pp(ab), _i(bc), sg,
                                   _d(Ti), _d(Tb), _d(Fi), _d(Fb))
    return Fi, Fb


def mules_limiter(m, cfg, rDeltaT, psii, psib, psi0, psiBc, phiBDi, phiBDb, phiCorri, phiCorrb, Sp, Su,
                  psiMax=1.0, psiMin=0.0, comm=None):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    a = [f64(x) for x in (psii, psib, psi0, phiBDi, phiBDb, phiCorri, phiCorrb, Sp, Su)]
    bc = i32(psiBc)
    li, lb = np.zeros(h.s.nIF), np.zeros(h.s.nBF)
    lib().orc_mules_limiter(h.ref, C.byref(comm) if comm else None, C.byref(cfg), C.c_double(rDeltaT),
                            _d(a[0]), _d(a[1]), _d(a[2]), _i(bc), _d(a[3]), _d(a[4]), _d(a[5]), _d(a[6]),
                            _d(a[7]), _d(a[8]), C.c_double(psiMax), C.c_double(psiMin), _d(li), _d(lb))
    return li, lb


def mules_limit(m, cfg, rDeltaT, psii, psib, psi0, psiBc, phii, phib, phiPsii, phiPsib, Sp, Su,
                returnCorr, psiMax=1.0, psiMin=0.0, comm=None):
    """Returns new (phiPsii, phiPsib)."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    a = [f64(x) for x in (psii, psib, psi0, phii, phib, Sp, Su)]
    pi, pb = f64(phiPsii).copy(), f64(phiPsib).copy()
    bc = i32(psiBc)
    lib().orc_mules_limit(h.ref, C.byref(comm) if comm else None, C.byref(cfg), C.c_double(rDeltaT),
                          _d(a[0]), _d(a[1]), _d(a[2]), _i(bc), _d(a[3]), _d(a[4]), _d(pi), _d(pb),
                          _d(a[5]), _d(a[6]), C.c_double(psiMax), C.c_double(psiMin), int(bool(returnCorr)))
    return pi, pb


def mules_limit_sum(m, phiCorri, phiCorrb):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    ci = [f64(x).copy() for x in phiCorri]; cb = [f64(x).copy() for x in phiCorrb]
    lib().orc_mules_limit_sum(h.ref, len(ci), _pp(ci), _pp(cb))
    return ci, cb


def mules_explicit_solve(m, cfg, deltaT, psii, psib, psi0, psiBc, phii, phib, phiPsii, phiPsib, Sp, Su,
                         psiMax=1.0, psiMin=0.0, comm=None):
    """Returns (psii, psib, phiPsii, phiPsib) new."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    a = [f64(x) for x in (psi0, phii, phib, Sp, Su)]
    qi, qb = f64(psii).copy(), f64(psib).copy()
    pi, pb = f64(phiPsii).copy(), f64(phiPsib).copy()
    bc = i32(psiBc)
    lib().orc_mules_explicit_solve(h.ref, C.byref(comm) if comm else None, C.byref(cfg), C.c_double(deltaT),
                                   _d(qi), _d(qb), _d(a[0]), _i(bc), _d(a[1]), _d(a[2]), _d(pi), _d(pb),
                                   _d(a[3]), _d(a[4]), C.c_double(psiMax), C.c_double(psiMin))
    return qi, qb, pi, pb


def upwind_flux(m, phii, phib, psii, psib):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    a = [f64(x) for x in (phii, phib, psii, psib)]
    oi, ob = np.zeros(h.s.nIF), np.zeros(h.s.nBF)
    lib().orc_upwind_flux(h.ref, _d(a[0]), _d(a[1]), _d(a[2]), _d(a[3]), _d(oi), _d(ob), 0)
    return oi, ob


def stage(m, cfg, alphas_i, alphas_b, alphaBc, sigma_specs, Ti, Tb, phii, phib, phiAlpha_i, phiAlpha_b,
          deltaT, comm=None):
    """One metric stage.  Returns dict(F_i,F_b, alpha_i[], alpha_b[], alphaPhi_i[], alphaPhi_b[])."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    P = len(alphas_i)
    ai = [f64(a).copy() for a in alphas_i]; ab = [f64(a).copy() for a in alphas_b]
    pi = [f64(a).copy() for a in phiAlpha_i]; pb = [f64(a).copy() for a in phiAlpha_b]
    bc = i32(alphaBc).reshape(P, -1)
    Ti, Tb, phii, phib = f64(Ti), f64(Tb), f64(phii), f64(phib)
    Fi, Fb = np.zeros((h.s.nCells, 3)), np.zeros((h.s.nBF, 3))
    sg = sigma_array(sigma_specs)
    lib().orc_stage(h.ref, C.byref(comm) if comm else None, C.byref(cfg), P, _pp(ai), _pp(ab), _i(bc), sg,
                    _d(Ti), _d(Tb), _d(phii), _d(phib), _pp(pi), _pp(pb), None, None, C.c_double(deltaT),
                    _d(Fi), _d(Fb))
    return dict(F_i=Fi, F_b=Fb, alpha_i=ai, alpha_b=ab, alphaPhi_i=pi, alphaPhi_b=pb)


def stage_threads(meshes, cfg, alphas_i, alphas_b, alphaBc, sigma_specs, Ti, Tb, phii, phib, phiAlpha_i,
                  phiAlpha_b, deltaT, nStages=1):
    """Domain-decomposed stage, one thread per sub-domain.  Per-domain lists.  Returns (seconds, results[])."""
    nD = len(meshes)
    hs = [MeshHandle(m) for m in meshes]
    marr = (OrcMesh * nD)(*[h.s for h in hs])
    P = len(alphas_i[0])
    ai = [[f64(a).copy() for a in alphas_i[d]] for d in range(nD)]
    ab = [[f64(a).copy() for a in alphas_b[d]] for d in range(nD)]
    pi = [[f64(a).copy() for a in phiAlpha_i[d]] for d in range(nD)]
    pb = [[f64(a).copy() for a in phiAlpha_b[d]] for d in range(nD)]
    bcs = [i32(alphaBc[d]).reshape(P, -1) for d in range(nD)]
    Tis = [f64(x) for x in Ti]; Tbs = [f64(x) for x in Tb]
    phis = [f64(x) for x in phii]; phibs = [f64(x) for x in phib]
    Fis = [np.zeros((h.s.nCells, 3)) for h in hs]; Fbs = [np.zeros((h.s.nBF, 3)) for h in hs]
    ppp = C.POINTER(c_dp)

    def ppl(lists):
        keep = [_pp(l) for l in lists]
        arr = (ppp * nD)(*[C.cast(k, ppp) for k in keep])
        return arr, keep
    aiA, k1 = ppl(ai); abA, k2 = ppl(ab); piA, k3 = ppl(pi); pbA, k4 = ppl(pb)
    bcA = (c_ip * nD)(*[_i(b) for b in bcs])
    sg = sigma_array(sigma_specs)
    secs = lib().orc_stage_threads(nD, marr, C.byref(cfg), P, aiA, abA, bcA, sg, _pp(Tis), _pp(Tbs), _pp(phis),
                                   _pp(phibs), piA, pbA, C.c_double(deltaT), _pp(Fis), _pp(Fbs), int(nStages))
    res = [dict(F_i=Fis[d], F_b=Fbs[d], alpha_i=ai[d], alpha_b=ab[d], alphaPhi_i=pi[d], alphaPhi_b=pb[d])
           for d in range(nD)]
    return secs, res


# ---- rest of solveAlphas ---------------------------------------------------------
def conv_flux(m, scheme, fluxi, fluxb, ai, ab, bc, comm=None):
    """fvc::flux(flux, alpha, scheme)."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    a = [f64(x) for x in (fluxi, fluxb, ai, ab)]
    bc = i32(bc)
    oi, ob = np.zeros(h.s.nIF), np.zeros(h.s.nBF)
    lib().orc_conv_flux(h.ref, C.byref(comm) if comm else None, SCHEMES[scheme], _d(a[0]), _d(a[1]), _d(a[2]), _d(a[3]),
                        _i(bc), _d(oi), _d(ob))
    return oi, ob


def nHatf(m, a1i, a1b, bc1, a2i, a2b, bc2, comm=None):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    a = [f64(x) for x in (a1i, a1b, a2i, a2b)]
    b1, b2 = i32(bc1), i32(bc2)
    oi, ob = np.zeros(h.s.nIF), np.zeros(h.s.nBF)
    lib().orc_nHatf(h.ref, C.byref(comm) if comm else None, _d(a[0]), _d(a[1]), _i(b1), _d(a[2]), _d(a[3]), _i(b2),
                    _d(oi), _d(ob))
    return oi, ob


def flux_assembly(m, alphas_i, alphas_b, alphaBc, phii, phib, schemePhi="vanLeer", schemePhir="linear", cAlpha=None,
                  comm=None):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    P = len(alphas_i)
    ai = [f64(a) for a in alphas_i]; ab = [f64(a) for a in alphas_b]
    bc = i32(alphaBc).reshape(P, -1)
    phii, phib = f64(phii), f64(phib)
    cA = f64(np.ones((P, P)) if cAlpha is None else cAlpha)
    oi = [np.zeros(h.s.nIF) for _ in range(P)]; ob = [np.zeros(h.s.nBF) for _ in range(P)]
    lib().orc_flux_assembly(h.ref, C.byref(comm) if comm else None, P, _pp(ai), _pp(ab), _i(bc), _d(phii), _d(phib),
                            SCHEMES[schemePhi], SCHEMES[schemePhir], _d(cA), _pp(oi), _pp(ob))
    return oi, ob


def upwind_predictor(m, deltaT, psii, psib, psi0, psiBc, phii, phib, phiAlpha_i, phiAlpha_b, Su=None, Sp=None,
                     tol=1e-12, maxIter=200, comm=None):
    """Returns (psii, psib, phiAlpha_i, phiAlpha_b, sweeps)."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    qi, qb = f64(psii).copy(), f64(psib).copy()
    pi, pb = f64(phiAlpha_i).copy(), f64(phiAlpha_b).copy()
    a = [f64(x) for x in (psi0, phii, phib)]
    su = f64(Su) if Su is not None else None
    sp = f64(Sp) if Sp is not None else None
    bc = i32(psiBc)
    n = lib().orc_upwind_predictor(h.ref, C.byref(comm) if comm else None, C.c_double(deltaT), _d(qi), _d(qb), _d(a[0]),
                                   _i(bc), _d(a[1]), _d(a[2]), _d(su), _d(sp), C.c_double(tol), int(maxIter), _d(pi), _d(pb))
    return qi, qb, pi, pb, int(n)


def solve(m, cfg, ctl, alphas_i, alphas_b, alphaBc, phii, phib, deltaT, rho, cAlpha=None, Su=None, Sp=None, comm=None):
    """multiphaseSystem::solve().  rho: per-phase uniform densities.  Returns dict(alpha_i, alpha_b, alphaPhi_i,
    alphaPhi_b, rhoPhi_i, rhoPhi_b)."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    P = len(alphas_i)
    ai = [f64(a).copy() for a in alphas_i]; ab = [f64(a).copy() for a in alphas_b]
    bc = i32(alphaBc).reshape(P, -1)
    phii, phib = f64(phii), f64(phib)
    cA = f64(np.ones((P, P)) if cAlpha is None else cAlpha)
    rho = f64(rho)
    su = [f64(x) for x in Su] if Su is not None else None
    sp = [f64(x) for x in Sp] if Sp is not None else None
    pi = [np.zeros(h.s.nIF) for _ in range(P)]; pb = [np.zeros(h.s.nBF) for _ in range(P)]
    ri, rb = np.zeros(h.s.nIF), np.zeros(h.s.nBF)
    lib().orc_solve(h.ref, C.byref(comm) if comm else None, C.byref(cfg), C.byref(ctl), P, _pp(ai), _pp(ab), _i(bc),
                    _d(phii), _d(phib), _d(cA), _d(rho), None, None, _pp(su) if su else None, _pp(sp) if sp else None,
                    C.c_double(deltaT), _pp(pi), _pp(pb), _d(ri), _d(rb))
    return dict(alpha_i=ai, alpha_b=ab, alphaPhi_i=pi, alphaPhi_b=pb, rhoPhi_i=ri, rhoPhi_b=rb)


def flux_of_vector(m, Fi, Fb):
    """fvc::flux(F) = interpolate(F) & Sf."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    Fi, Fb = f64(Fi), f64(Fb)
    oi, ob = np.zeros(h.s.nIF), np.zeros(h.s.nBF)
    lib().orc_flux_of_vector(h.ref, _d(Fi), _d(Fb), _d(oi), _d(ob))
    return oi, ob


# ---- Su/Sp of the phase-fraction equations, porous interface drag --------------------
VAR_T, VAR_ALPHA = 0, 1


class OrcLee(C.Structure):
    _fields_ = [("from_", C.c_int32), ("to", C.c_int32), ("C", C.c_double), ("Tactivate", C.c_double),
                ("alphaMin", C.c_double), ("modelVariable", C.c_int32), ("includeDivU", C.c_int32)]


class OrcMass(C.Structure):
    _fields_ = [("nModels", C.c_int32), ("models", C.POINTER(OrcLee)), ("nPairs", C.c_int32), ("pairs", c_ip),
                ("rhoUniform", c_dp), ("rhoi", C.POINTER(c_dp))]


class MassHandle:
    """orc_mass built from python values; keeps the arrays alive.  models: list of dicts (from, to, C, Tactivate,
    alphaMin, modelVariable, includeDivU); pairs: [(p1, p2), ...]; rho: uniform per-phase densities; rho_fields:
    optional per-phase arrays (None entries fall back to rho)."""

    def __init__(self, models, pairs, rho, rho_fields=None):
        self.models = (OrcLee * max(len(models), 1))()
        for i, md in enumerate(models):
            self.models[i] = OrcLee(md["from"], md["to"], md["C"], md["Tactivate"], md.get("alphaMin", 0.0),
                                    md.get("modelVariable", VAR_T), int(md.get("includeDivU", True)))
        self.pairs = i32(np.asarray(pairs, dtype=np.int32).reshape(-1))
        self.rho = f64(rho)
        self.rf = [None if r is None else f64(r) for r in rho_fields] if rho_fields is not None else None
        self.rfp = None
        if self.rf is not None:
            self.rfp = (c_dp * len(self.rf))(*[(_d(r) if r is not None else None) for r in self.rf])
        self.s = OrcMass(len(models), self.models, len(pairs), _i(self.pairs), _d(self.rho),
                         C.cast(self.rfp, C.POINTER(c_dp)) if self.rfp is not None else None)

    @property
    def ref(self):
        return C.byref(self.s)


def div_phi(m, phii, phib):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    phii, phib = f64(phii), f64(phib)
    out = np.zeros(h.s.nCells)
    lib().orc_div_phi(h.ref, _d(phii), _d(phib), _d(out))
    return out


def correct_mass_sources(m, mass, alphas_i, T):
    """dmdt[i] = Kexp(T) of model i (MassTransferPhaseSystem::correctMassSources)."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    ai = [f64(a) for a in alphas_i]
    T = f64(T)
    out = [np.zeros(h.s.nCells) for _ in range(mass.s.nModels)]
    lib().orc_correct_mass_sources(h.ref, mass.ref, _pp(ai), _d(T), _pp(out))
    return out


def alpha_transfer(m, mass, alphas_i, dmdt, phii, phib, comm=None):
    """alphaTransfer on zeroed Su/Sp.  Returns (Su[P], Sp[P], ddtAlphaMax)."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    P = len(alphas_i)
    ai = [f64(a) for a in alphas_i]
    dm = [f64(d) for d in dmdt] if dmdt else None
    phii, phib = f64(phii), f64(phib)
    Su = [np.zeros(h.s.nCells) for _ in range(P)]; Sp = [np.zeros(h.s.nCells) for _ in range(P)]
    L = lib()
    L.orc_alpha_transfer.restype = C.c_double
    mx = L.orc_alpha_transfer(h.ref, C.byref(comm) if comm else None, mass.ref, P, _pp(ai), _pp(dm) if dm else None,
                              _d(phii), _d(phib), _pp(Su), _pp(Sp))
    return Su, Sp, float(mx)


def interface_porosity(m, solidPhase, Cu, alphas_i):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    ai = [f64(a) for a in alphas_i]
    sp, cu = i32(solidPhase), f64(Cu)
    out = np.zeros(h.s.nCells)
    lib().orc_interface_porosity(h.ref, len(sp), _i(sp), _d(cu), _pp(ai), _d(out))
    return out


def solve_mt(m, cfg, ctl, alphas_i, alphas_b, alphaBc, phii, phib, deltaT, rho, mass, dmdt, cAlpha=None, comm=None):
    """multiphaseSystem::solve() with calculateSuSp() inside every corrector."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    P = len(alphas_i)
    ai = [f64(a).copy() for a in alphas_i]; ab = [f64(a).copy() for a in alphas_b]
    bc = i32(alphaBc).reshape(P, -1)
    phii, phib = f64(phii), f64(phib)
    cA = f64(np.ones((P, P)) if cAlpha is None else cAlpha)
    rho = f64(rho)
    dm = [f64(d) for d in dmdt] if dmdt else None
    pi = [np.zeros(h.s.nIF) for _ in range(P)]; pb = [np.zeros(h.s.nBF) for _ in range(P)]
    ri, rb = np.zeros(h.s.nIF), np.zeros(h.s.nBF)
    lib().orc_solve_mt(h.ref, C.byref(comm) if comm else None, C.byref(cfg), C.byref(ctl), P, _pp(ai), _pp(ab), _i(bc),
                       _d(phii), _d(phib), _d(cA), _d(rho), None, None, None, None,
                       C.c_double(deltaT), _pp(pi), _pp(pb), _d(ri), _d(rb), mass.ref, _pp(dm) if dm else None)
    return dict(alpha_i=ai, alpha_b=ab, alphaPhi_i=pi, alphaPhi_b=pb, rhoPhi_i=ri, rhoPhi_b=rb)


def solve_yi(m, cfg, alphas_i, alphas_b, alphaBc, ph, Y_i, Y_b, YBc, inertIndex, phii, phib, deltaT, cYi=1.0,
             schemePhi="vanLeer", schemePhir="linear", Su=None, Sp=None, tol=1e-12, maxIter=200, comm=None, nYiSubCycles=1):
    """MultiComponentPhaseModel::solveYi for the species of phase `ph`.  Returns (Y_i, Y_b) lists."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    P, S = len(alphas_i), len(Y_i)
    ai = [f64(a) for a in alphas_i]; ab = [f64(a) for a in alphas_b]
    bc = i32(alphaBc).reshape(P, -1)
    yi = [f64(a).copy() for a in Y_i]; yb = [f64(a).copy() for a in Y_b]
    ybc = i32(YBc).reshape(S, -1)
    phii, phib = f64(phii), f64(phib)
    z = np.zeros(h.s.nCells)
    su = [f64(x) if x is not None else z for x in Su] if Su is not None else None
    sp = [f64(x) if x is not None else z for x in Sp] if Sp is not None else None
    lib().orc_solve_yi(h.ref, C.byref(comm) if comm else None, C.byref(cfg), P, _pp(ai), _pp(ab), _i(bc), int(ph), S,
                       int(inertIndex), _pp(yi), _pp(yb), _i(ybc), _d(phii), _d(phib), C.c_double(cYi),
                       SCHEMES[schemePhi], SCHEMES[schemePhir], _pp(su) if su else None, _pp(sp) if sp else None,
                       C.c_double(deltaT), C.c_double(tol), int(maxIter), int(nYiSubCycles))
    return yi, yb


# ---- laserDTRM -----------------------------------------------------------------------------
REFL_KINDS = {"noReflection": 0, "Fresnel": 1, "FresnelLaser": 2}
PD_MODES = {"Gaussian": 0, "uniform": 2, "GaussianPeak": 3}
SIGMA_SB = 5.670374419e-8          # physicoChemical::sigma


class OrcLaserGeom(C.Structure):
    _fields_ = [("Cf", c_dp), ("Cfb", c_dp), ("nPoints", C.c_int32), ("points", c_dp), ("fpStart", c_ip),
                ("fpPts", c_ip), ("tetBasePt", c_ip), ("boundsMag", C.c_double)]


class OrcReflection(C.Structure):
    _fields_ = [("kind", C.c_int32), ("from_", C.c_int32), ("to", C.c_int32), ("nk1", C.c_double * 2),
                ("nk2", C.c_double * 2), ("epsilon", C.c_double)]


class OrcLaserCfg(C.Structure):
    _fields_ = [("nModels", C.c_int32), ("models", C.POINTER(OrcReflection)), ("alphaCut", C.c_double),
                ("sigmaSB", C.c_double), ("maxParticles", C.c_int32), ("localFindCell", C.c_int32)]


class LaserGeomHandle:
    """geom: dict(points [nP,3], fpStart [nIF+nBF+1], fpPts, tetBasePt [nIF+nBF], boundsMag) + the mesh's Cf/Cfb."""

    def __init__(self, m, geom):
        self.keep = dict(Cf=f64(m.Cf), Cfb=f64(m.Cfb), points=f64(geom["points"]), fpStart=i32(geom["fpStart"]),
                         fpPts=i32(geom["fpPts"]), tetBasePt=i32(geom["tetBasePt"]))
        k = self.keep
        self.s = OrcLaserGeom(_d(k["Cf"]), _d(k["Cfb"]), len(k["points"]), _d(k["points"]), _i(k["fpStart"]),
                              _i(k["fpPts"]), _i(k["tetBasePt"]), float(geom["boundsMag"]))

    @property
    def ref(self):
        return C.byref(self.s)


def reflection_array(models):
    """models: list of dict(kind, from, to, nk1, nk2, epsilon) in the reflectionModel table's iteration order."""
    arr = (OrcReflection * max(len(models), 1))()
    for i, md in enumerate(models):
        arr[i].kind = REFL_KINDS[md["kind"]]; arr[i].from_ = md["from"]; arr[i].to = md["to"]
        arr[i].nk1[:] = md.get("nk1", (1.0, 0.0)); arr[i].nk2[:] = md.get("nk2", (1.0, 0.0))
        arr[i].epsilon = md.get("epsilon", 0.0)
    return arr


def laser_seed(lPosition, lDir, rArea, nr, nTheta, focalLaserRadius, maxTrackLength, mode="Gaussian", power=1.0,
               sigma=1.0, I0=0.0):
    n = nr * nTheta
    p0, p1, Ip, dA = np.zeros((n, 3)), np.zeros((n, 3)), np.zeros(n), np.zeros(n)
    v3 = lambda v: (C.c_double * 3)(*[float(x) for x in v])
    lib().orc_laser_seed(v3(lPosition), v3(lDir), v3(rArea), int(nr), int(nTheta), C.c_double(focalLaserRadius),
                         C.c_double(maxTrackLength), PD_MODES[mode], C.c_double(power), C.c_double(sigma), C.c_double(I0),
                         _d(p0), _d(p1), _d(Ip), _d(dA))
    return p0, p1, Ip, dA


def find_cells(m, geom, pts):
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    g = geom if isinstance(geom, LaserGeomHandle) else LaserGeomHandle(h.m, geom)
    pts = f64(pts).reshape(-1, 3)
    return np.array([lib().orc_find_cell(h.ref, g.ref, (C.c_double * 3)(*p)) for p in pts], np.int32)


def local_density(alphas_i, coeff):
    ai = [f64(a) for a in alphas_i]
    out = np.zeros(len(ai[0]))
    lib().orc_local_density(len(out), len(ai), _pp(ai), _d(f64(coeff)), _d(out))
    return out


def laser_calculate(m, geom, models, alphas_i, alphas_b, alphaBc, a, e, E, T, p0, p1, I, dA, cell, alphaCut=0.5,
                    sigmaSB=SIGMA_SB, maxParticles=0, comm=None):
    """laserDTRM::calculate().  Returns dict(Q, reflectingCells, nHat, nHatPoints, nParticles)."""
    h = m if isinstance(m, MeshHandle) else MeshHandle(m)
    g = geom if isinstance(geom, LaserGeomHandle) else LaserGeomHandle(h.m, geom)
    P = len(alphas_i)
    ai = [f64(x) for x in alphas_i]; ab = [f64(x) for x in alphas_b]
    bc = i32(alphaBc).reshape(P, -1)
    arr = reflection_array(models)
    n = len(I)
    cfg = OrcLaserCfg(len(models), arr, float(alphaCut), float(sigmaSB), int(maxParticles or 64 * max(n, 1)), 0)
    nC = h.s.nCells
    Q = np.zeros(nC); refl = np.zeros(nC, np.int32); nHat = np.zeros((nC, 3)); nPts = np.zeros((g.s.nPoints, 3))
    nFinal = C.c_int32(0)
    p0, p1, I, dA, cell = f64(p0), f64(p1), f64(I), f64(dA), i32(cell)
    a, e, E, T = f64(a), f64(e), f64(E), f64(T)
    rc = lib().orc_laser_calculate(h.ref, C.byref(comm) if comm else None, g.ref, C.byref(cfg), P, _pp(ai), _pp(ab), _i(bc),
                                   _d(a), _d(e), _d(E), _d(T), n, _d(p0), _d(p1), _d(I), _d(dA), _i(cell), _d(Q), _i(refl),
                                   _d(nHat), _d(nPts), C.byref(nFinal))
    if rc:
        raise RuntimeError(f"orc_laser_calculate failed: rc={rc}")
    return dict(Q=Q, reflectingCells=refl, nHat=nHat, nHatPoints=nPts, nParticles=nFinal.value)


def laser_calculate_ranks(meshes, geoms, models, refl, nHat, pointGlobal, a, e, E, T, p0, p1, I, dA, cells, alphaCut=0.5,
                          sigmaSB=SIGMA_SB, maxParticles=0, localFindCell=False):
    """The decomposed laser walk with all ranks on one CPU (orc_laser_calculate_ranks).  Per-rank lists: meshes, geoms,
    refl/nHat (the marking restricted to the rank's cells), pointGlobal, a/e/E/T, cells (local cell of each seed or -1).
    Returns dict(Q=[...], created=[...], nHatPoints=[...])."""
    R = len(meshes)
    hs = [m if isinstance(m, MeshHandle) else MeshHandle(m) for m in meshes]
    gs = [g if isinstance(g, LaserGeomHandle) else LaserGeomHandle(h.m, g) for h, g in zip(hs, geoms)]
    arr = reflection_array(models)
    n = len(I)
    cfg = OrcLaserCfg(len(models), arr, float(alphaCut), float(sigmaSB), int(maxParticles or 64 * max(n, 1)), int(bool(localFindCell)))
    mp = (C.POINTER(OrcMesh) * R)(*[C.pointer(h.s) for h in hs])
    gp = (C.POINTER(OrcLaserGeom) * R)(*[C.pointer(g.s) for g in gs])
    refl = [i32(x) for x in refl]; nHat = [f64(x) for x in nHat]
    pg = [np.ascontiguousarray(x, dtype=np.int64) for x in pointGlobal]
    a, e, E, T = ([f64(x) for x in v] for v in (a, e, E, T))
    cells = [i32(x) for x in cells]
    Q = [np.zeros(h.s.nCells) for h in hs]
    nPts = [np.zeros((g.s.nPoints, 3)) for g in gs]
    created = np.zeros(R, np.int32)
    ipp = lambda xs: (c_ip * R)(*[_i(x) for x in xs])
    lpp = (C.POINTER(C.c_int64) * R)(*[x.ctypes.data_as(C.POINTER(C.c_int64)) for x in pg])
    p0, p1, I, dA = f64(p0), f64(p1), f64(I), f64(dA)
    fn = lib().orc_laser_calculate_ranks
    fn.restype = C.c_int
    rc = fn(R, mp, gp, C.byref(cfg), ipp(refl), _pp(nHat), lpp, _pp(a), _pp(e), _pp(E), _pp(T), n, _d(p0), _d(p1), _d(I), _d(dA),
            ipp(cells), _pp(Q), _i(created), _pp(nPts))
    if rc:
        raise RuntimeError(f"orc_laser_calculate_ranks failed: rc={rc}")
    return dict(Q=Q, created=created, nHatPoints=nPts)
