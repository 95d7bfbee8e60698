"""ctypes binding of the C ABI in include/tcif.h (libtcif.so, hand-written CUDA for sm_100a).

This is the *only* compute path of the package: there is no CPU fallback.  Loading fails
loudly when the CUDA library has not been built (``python -c 'import __graft_entry__ as g;
g.build()'``) and ``Context`` creation fails when no B200 is visible.

The binding is deliberately thin: one Python method per exported entry point, numpy arrays in
OpenFOAM layout (vectors AoS [n,3], vol field = internal[nCells] + boundary[nBoundaryFaces]).
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# TCIF_LIB: alternative build of the same library (kernel-tuning experiments only)
LIB_PATH = os.environ.get("TCIF_LIB") or os.path.join(_HERE, "libtcif.so")

PATCH_GENERIC, PATCH_EMPTY, PATCH_WEDGE, PATCH_PROCESSOR = 0, 1, 2, 3
BC_ZERO_GRADIENT, BC_FIXED_VALUE = 0, 1
SIGMA_CONSTANT, SIGMA_POLYNOMIAL, SIGMA_TABLE = 0, 1, 2
BEXT_GLOBAL, BEXT_PATCHVALUE = 0, 1

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)


class TcifError(RuntimeError):
    """Non-zero return of a tcif_* entry point (the shim turns these into FatalError)."""


class MeshDesc(C.Structure):
    _fields_ = [("nCells", C.c_int32), ("nInternalFaces", C.c_int32), ("nPatches", C.c_int32),
                ("nBoundaryFaces", C.c_int32),
                ("owner", c_ip), ("neighbour", c_ip), ("Sf", c_dp), ("weights", c_dp), ("V", c_dp),
                ("patchType", c_ip), ("patchStart", c_ip), ("patchSize", c_ip), ("patchNeighbProcNo", c_ip),
                ("faceCells", c_ip), ("SfBoundary", c_dp), ("magSfBoundary", c_dp),
                ("deltaCoeffsBoundary", c_dp), ("weightsBoundary", c_dp)]


VAR_T, VAR_ALPHA = 0, 1


class LeeModel(C.Structure):
    """tcif_lee_model (include/tcif.h)"""
    _fields_ = [("from_", C.c_int32), ("to", C.c_int32), ("C", C.c_double), ("Tactivate", C.c_double),
                ("alphaMin", C.c_double), ("modelVariable", C.c_int32), ("includeDivU", C.c_int32)]


class ReflectionModel(C.Structure):
    """tcif_reflection_model (include/tcif.h)"""
    _fields_ = [("kind", C.c_int32), ("from_", C.c_int32), ("to", C.c_int32), ("nk1", C.c_double * 2),
                ("nk2", C.c_double * 2), ("epsilon", C.c_double)]


REFL_KINDS = {"noReflection": 0, "Fresnel": 1, "FresnelLaser": 2}
SIGMA_SB = 5.670374419e-8       # physicoChemical::sigma


class AlphaControls(C.Structure):
    """tcif_alpha_controls: fvSolution "alpha" controls of solve()/solveAlphas()."""
    _fields_ = [("nAlphaCorr", C.c_int32), ("nAlphaSubCycles", C.c_int32), ("schemePhi", C.c_int32),
                ("schemePhir", C.c_int32), ("implicitPredictor", C.c_int32), ("predictorTol", C.c_double),
                ("predictorMaxIter", C.c_int32)]


SCHEMES = {"upwind": 0, "linear": 1, "vanLeer": 2}


def alpha_controls(nAlphaCorr=1, nAlphaSubCycles=1, schemePhi="vanLeer", schemePhir="linear", implicitPredictor=True,
                   predictorTol=1e-12, predictorMaxIter=200):
    return AlphaControls(nAlphaCorr, nAlphaSubCycles, SCHEMES[schemePhi], SCHEMES[schemePhir], int(implicitPredictor),
                         predictorTol, predictorMaxIter)


class Config(C.Structure):
    _fields_ = [("nPhases", C.c_int32), ("nLimiterIter", C.c_int32), ("smoothLimiter", C.c_double),
                ("extremaCoeff", C.c_double), ("boundaryExtrema", C.c_int32), ("device", C.c_int32),
                ("flags", C.c_int32)]


_lib = None


def lib():
    """Load libtcif.so.  Raises if it has not been built -- never falls back to anything."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise TcifError(f"{LIB_PATH} is missing: build it with __graft_entry__.build() "
                        "(nvcc -gencode arch=compute_100a,code=sm_100a); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    L.tcif_last_error.restype = C.c_char_p
    L.tcif_last_error.argtypes = [C.c_void_p]
    L.tcif_version.restype = C.c_char_p
    L.tcif_timer_name.restype = C.c_char_p
    L.tcif_host_alloc.restype = C.c_void_p
    L.tcif_host_alloc.argtypes = [C.c_uint64]
    L.tcif_host_free.argtypes = [C.c_void_p]
    L.tcif_launch_count.restype = C.c_int64
    L.tcif_launch_count.argtypes = [C.c_void_p]
    L.tcif_device_bytes.restype = C.c_int64
    L.tcif_device_bytes.argtypes = [C.c_void_p]
    L.tcif_mesh_info.argtypes = [C.c_void_p, C.POINTER(C.c_int32)]
    L.tcif_comm_transport.argtypes = [C.c_void_p]
    L.tcif_set_alpha_stamped.argtypes = [C.c_void_p, C.c_int, C.c_uint64, c_dp, c_dp, C.POINTER(C.c_int32)]
    L.tcif_set_temperature_stamped.argtypes = [C.c_void_p, C.c_uint64, c_dp, c_dp, C.POINTER(C.c_int32)]
    L.tcif_mark_alpha.argtypes = [C.c_void_p, C.c_int, C.c_uint64]
    L.tcif_transfer_counters.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
    L.tcif_debug_guard_check.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.tcif_host_alpha_stamps.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.tcif_host_sparse_returns.argtypes = [C.c_void_p, C.c_int]
    L.tcif_host_bytes_moved.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.tcif_predictor_sweeps.argtypes = [C.c_void_p, C.c_int]
    L.tcif_destroy.argtypes = [C.c_void_p]
    L.tcif_destroy.restype = None
    L.tcif_group_destroy.argtypes = [C.c_void_p]
    L.tcif_group_destroy.restype = None
    _lib = L
    return L


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _d(a):
    return a.ctypes.data_as(c_dp) if a is not None else None


def _i(a):
    return a.ctypes.data_as(c_ip) if a is not None else None


def _pp(arrs):
    p = (c_dp * max(1, len(arrs)))()
    for k, a in enumerate(arrs):
        p[k] = _d(a)
    return p


def device_count() -> int:
    return int(lib().tcif_device_count())


def version() -> str:
    return lib().tcif_version().decode()


def pinned_empty(shape, dtype=np.float64):
    """numpy array on cudaHostAlloc'd memory (tcif_host_alloc); freed when the array dies."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = lib().tcif_host_alloc(C.c_uint64(max(n, 1)))
    if not p:
        raise TcifError("tcif_host_alloc failed")
    buf = (C.c_char * max(n, 1)).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[id(buf)] = (buf, p)
    import weakref
    weakref.finalize(arr, _free_pinned, id(buf))
    return arr


_PINNED: dict = {}


def _free_pinned(key):
    buf, p = _PINNED.pop(key, (None, None))
    if p:
        lib().tcif_host_free(C.c_void_p(p))


class Group:
    """In-process rank group: several contexts on one GPU, one host thread per rank (tests)."""

    def __init__(self, nRanks: int):
        self.h = C.c_void_p()
        if lib().tcif_group_create(int(nRanks), C.byref(self.h)):
            raise TcifError("tcif_group_create failed")
        self.n = nRanks

    def close(self):
        if self.h:
            lib().tcif_group_destroy(self.h)
            self.h = C.c_void_p()


def nccl_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    if lib().tcif_comm_unique_id(buf):
        raise TcifError(lib().tcif_last_error(None).decode())
    return buf.raw


class Context:
    """One rank's device-resident mesh + fields (tcif_ctx)."""

    def __init__(self, mesh, nPhases, nLimiterIter=3, smoothLimiter=0.0, extremaCoeff=0.0,
                 boundaryExtrema=BEXT_GLOBAL, device=-1):
        """`mesh` is any object with the attributes of tcif_mesh_desc's arrays (owner, neighbour, Sf,
        w, V, patchType, patchStart, patchSize, patchNbrProc, faceCells, Sfb, magSfb, deltaCoeffsb, wb)."""
        L = lib()
        self.nCells = int(mesh.nCells)
        k = dict(owner=_i32(mesh.owner), neighbour=_i32(mesh.neighbour), Sf=_f64(mesh.Sf), w=_f64(mesh.w),
                 V=_f64(mesh.V), patchType=_i32(mesh.patchType), patchStart=_i32(mesh.patchStart),
                 patchSize=_i32(mesh.patchSize), patchNbrProc=_i32(mesh.patchNbrProc),
                 faceCells=_i32(mesh.faceCells), Sfb=_f64(mesh.Sfb), magSfb=_f64(mesh.magSfb),
                 dcb=_f64(mesh.deltaCoeffsb), wb=_f64(mesh.wb))
        self.nIF = int(k["owner"].shape[0])
        self.nBF = int(k["faceCells"].shape[0])
        self.nPatches = int(k["patchType"].shape[0])
        self.P = int(nPhases)
        desc = MeshDesc(self.nCells, self.nIF, self.nPatches, self.nBF, _i(k["owner"]), _i(k["neighbour"]),
                        _d(k["Sf"]), _d(k["w"]), _d(k["V"]), _i(k["patchType"]), _i(k["patchStart"]),
                        _i(k["patchSize"]), _i(k["patchNbrProc"]), _i(k["faceCells"]), _d(k["Sfb"]),
                        _d(k["magSfb"]), _d(k["dcb"]), _d(k["wb"]))
        cfg = Config(self.P, int(nLimiterIter), float(smoothLimiter), float(extremaCoeff), int(boundaryExtrema),
                     int(device), 0)
        self.h = C.c_void_p()
        if L.tcif_create(C.byref(desc), C.byref(cfg), C.byref(self.h)):
            raise TcifError(L.tcif_last_error(None).decode())
        self._group = None

    # -- plumbing
    def _ck(self, rc):
        if rc:
            raise TcifError(lib().tcif_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            lib().tcif_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def synchronize(self):
        self._ck(lib().tcif_synchronize(self.h))

    def set_stream(self, cuda_stream_ptr):
        self._ck(lib().tcif_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def comm_init_local(self, rank, group: Group):
        self._group = group
        self._ck(lib().tcif_comm_init_local(self.h, int(rank), group.h))

    def comm_init_nccl(self, rank, nRanks, id128: bytes):
        self._ck(lib().tcif_comm_init_nccl(self.h, int(rank), int(nRanks), C.c_char_p(id128)))

    def ldu_csr(self):
        os_, lo, ls = (np.zeros(self.nCells + 1, np.int32), np.zeros(self.nIF, np.int32),
                       np.zeros(self.nCells + 1, np.int32))
        self._ck(lib().tcif_get_ldu_csr(self.h, _i(os_), _i(lo), _i(ls)))
        return os_, lo, ls

    # -- fields
    def set_alpha_bc(self, phase, kinds):
        k = _i32(kinds)
        assert k.shape[0] == self.nPatches
        self._ck(lib().tcif_set_alpha_bc(self.h, int(phase), _i(k)))

    def set_alpha(self, phase, internal, boundary):
        a, b = _f64(internal), _f64(boundary)
        assert a.shape[0] == self.nCells and b.shape[0] == self.nBF
        self._ck(lib().tcif_set_alpha(self.h, int(phase), _d(a), _d(b)))

    def get_alpha(self, phase):
        a, b = np.empty(self.nCells), np.empty(self.nBF)
        self._ck(lib().tcif_get_alpha(self.h, int(phase), _d(a), _d(b)))
        return a, b

    def set_alpha_stamped(self, phase, stamp, internal, boundary):
        """tcif_set_alpha_stamped: uploads only when `stamp` differs from the device copy's; returns whether bytes moved."""
        a, b = _f64(internal), _f64(boundary)
        assert a.shape[0] == self.nCells and b.shape[0] == self.nBF
        up = C.c_int32(0)
        self._ck(lib().tcif_set_alpha_stamped(self.h, int(phase), C.c_uint64(int(stamp)), _d(a), _d(b), C.byref(up)))
        return bool(up.value)

    def mark_alpha(self, phase, stamp):
        self._ck(lib().tcif_mark_alpha(self.h, int(phase), C.c_uint64(int(stamp))))

    def set_temperature_stamped(self, stamp, internal, boundary):
        a, b = _f64(internal), _f64(boundary)
        up = C.c_int32(0)
        self._ck(lib().tcif_set_temperature_stamped(self.h, C.c_uint64(int(stamp)), _d(a), _d(b), C.byref(up)))
        return bool(up.value)

    @property
    def transfer_counters(self):
        out = (C.c_int64 * 4)()
        self._ck(lib().tcif_transfer_counters(self.h, out))
        return dict(alphaUploads=int(out[0]), alphaSkipped=int(out[1]), TUploads=int(out[2]), TSkipped=int(out[3]))

    def store_old_alphas(self):
        self._ck(lib().tcif_store_old_alphas(self.h))

    def set_alpha_old(self, phase, internal):
        a = _f64(internal)
        self._ck(lib().tcif_set_alpha_old(self.h, int(phase), _d(a)))

    def set_temperature(self, internal, boundary):
        a, b = _f64(internal), _f64(boundary)
        self._ck(lib().tcif_set_temperature(self.h, _d(a), _d(b)))

    def set_sigma(self, phase1, phase2, spec):
        """spec: ("constant", s0) | ("polynomial", [(a,e),..]) | ("table", [(T,s),..])."""
        kind, val = spec
        if kind == "constant":
            co, n, k = _f64([val]), 1, SIGMA_CONSTANT
        else:
            co = _f64(np.asarray(val, dtype=np.float64).reshape(-1))
            n = len(val)
            k = SIGMA_POLYNOMIAL if kind == "polynomial" else SIGMA_TABLE if kind == "table" else 99
        self._ck(lib().tcif_set_sigma(self.h, int(phase1), int(phase2), k, n, _d(co)))

    def set_phi(self, internal, boundary):
        a, b = _f64(internal), _f64(boundary)
        assert a.shape[0] == self.nIF and b.shape[0] == self.nBF
        self._ck(lib().tcif_set_phi(self.h, _d(a), _d(b)))

    def set_phi_alpha(self, phase, internal, boundary):
        a, b = _f64(internal), _f64(boundary)
        assert a.shape[0] == self.nIF and b.shape[0] == self.nBF
        self._ck(lib().tcif_set_phi_alpha(self.h, int(phase), _d(a), _d(b)))

    def get_phi_alpha(self, phase):
        a, b = np.empty(self.nIF), np.empty(self.nBF)
        self._ck(lib().tcif_get_phi_alpha(self.h, int(phase), _d(a), _d(b)))
        return a, b

    def set_su_sp(self, phase, Su=None, Sp=None):
        su = _f64(Su) if Su is not None else None
        sp = _f64(Sp) if Sp is not None else None
        self._ck(lib().tcif_set_su_sp(self.h, int(phase), _d(su), _d(sp)))

    # -- Su/Sp from the mass-transfer models (MassTransferPhaseSystem::alphaTransfer)
    def mass_transfer_clear(self):
        self._ck(lib().tcif_mass_transfer_clear(self.h))

    def add_mass_transfer_lee(self, from_, to, C_, Tactivate, alphaMin=0.0, modelVariable=VAR_T, includeDivU=True):
        mdl = LeeModel(int(from_), int(to), float(C_), float(Tactivate), float(alphaMin), int(modelVariable), int(includeDivU))
        self._ck(lib().tcif_add_mass_transfer_lee(self.h, C.byref(mdl)))

    def set_phase_pairs(self, pairs):
        a = np.ascontiguousarray(np.asarray(pairs, dtype=np.int32).reshape(-1))
        self._ck(lib().tcif_set_phase_pairs(self.h, len(a) // 2, a.ctypes.data_as(C.POINTER(C.c_int32))))

    def set_dmdt(self, from_, to, internal):
        a = _f64(internal) if internal is not None else None
        self._ck(lib().tcif_set_dmdt(self.h, int(from_), int(to), _d(a)))

    def get_dmdt(self, from_, to):
        a = np.empty(self.nCells)
        self._ck(lib().tcif_get_dmdt(self.h, int(from_), int(to), _d(a)))
        return a

    def correct_mass_sources(self):
        self._ck(lib().tcif_correct_mass_sources(self.h))

    def alpha_transfer(self):
        mx = C.c_double(0.0)
        self._ck(lib().tcif_alpha_transfer(self.h, C.byref(mx)))
        return mx.value

    def set_mass_transfer_in_solve(self, on=True):
        self._ck(lib().tcif_set_mass_transfer_in_solve(self.h, int(on)))

    def get_su_sp(self, phase):
        su, sp = np.empty(self.nCells), np.empty(self.nCells)
        self._ck(lib().tcif_get_su_sp(self.h, int(phase), _d(su), _d(sp)))
        return su, sp

    def interface_porosity(self, solidPhase, Cu, Udiag):
        sp = np.ascontiguousarray(solidPhase, dtype=np.int32)
        cu = _f64(Cu)
        d = _f64(Udiag).copy()
        self._ck(lib().tcif_interface_porosity(self.h, len(sp), sp.ctypes.data_as(C.POINTER(C.c_int32)), _d(cu), _d(d)))
        return d

    # ---- laserDTRM ----
    def laser_set_polymesh(self, Cf, Cfb, geom):
        """geom: dict(points, fpStart, fpPts, tetBasePt, boundsMag) as cases.box_polymesh builds it."""
        cf, cfb, pts = _f64(Cf), _f64(Cfb), _f64(geom["points"])
        fs, fp, tb = _i32(geom["fpStart"]), _i32(geom["fpPts"]), _i32(geom["tetBasePt"])
        self._nPoints = len(pts.reshape(-1, 3))
        self._ck(lib().tcif_laser_set_polymesh(self.h, _d(cf), _d(cfb), self._nPoints, _d(pts), _i(fs), _i(fp), _i(tb),
                                               C.c_double(geom["boundsMag"])))

    def laser_set_shared_points(self, localPoint, globalPoint):
        """Collective: the local points on processor patches and their global labels."""
        lp = _i32(localPoint)
        gp = np.ascontiguousarray(globalPoint, dtype=np.int64)
        self._ck(lib().tcif_laser_set_shared_points(self.h, len(lp), _i(lp), gp.ctypes.data_as(C.POINTER(C.c_int64))))

    def laser_set_local_find_cell(self, on=True):
        self._ck(lib().tcif_laser_set_local_find_cell(self.h, int(bool(on))))

    def laser_set_reflection(self, models, alphaCut=0.5):
        arr = (ReflectionModel * max(len(models), 1))()
        for i, md in enumerate(models):
            arr[i].kind = REFL_KINDS[md["kind"]]; arr[i].from_ = md["from"]; arr[i].to = md["to"]
            arr[i].nk1[:] = md.get("nk1", (1.0, 0.0)); arr[i].nk2[:] = md.get("nk2", (1.0, 0.0))
            arr[i].epsilon = md.get("epsilon", 0.0)
        self._ck(lib().tcif_laser_set_reflection(self.h, len(models), arr, C.c_double(alphaCut)))

    def laser_set_local_density_absorption(self, phases, aCoeff, eCoeff, ECoeff):
        ph = _i32(phases)
        a, e, E = _f64(aCoeff), _f64(eCoeff), _f64(ECoeff)
        self._ck(lib().tcif_laser_set_local_density_absorption(self.h, len(ph), _i(ph), _d(a), _d(e), _d(E)))

    def laser_set_absorption_fields(self, a, e, E):
        a, e, E = _f64(a), _f64(e), _f64(E)
        self._ck(lib().tcif_laser_set_absorption_fields(self.h, _d(a), _d(e), _d(E)))

    def laser_calculate(self, p0, p1, I, dA, cell, sigmaSB=SIGMA_SB, maxParticles=0, dense=True):
        """laserDTRM::calculate(); returns (Q [nCells], particles created on this rank).  dense=False leaves Q on the
        device (read it with laser_get_sparse / laser_get) and returns (None, particles)."""
        p0, p1, I, dA, cell = _f64(p0), _f64(p1), _f64(I), _f64(dA), _i32(cell)
        Q = np.zeros(self.nCells) if dense else None
        n = C.c_int32(0)
        self._ck(lib().tcif_laser_calculate(self.h, len(I), _d(p0), _d(p1), _d(I), _d(dA), _i(cell), C.c_double(sigmaSB),
                                            int(maxParticles), _d(Q), C.byref(n)))
        return Q, n.value

    def laser_get_sparse(self, capacity=1 << 16):
        """(cells, values) of the cells that hold energy after the last laser_calculate, unordered."""
        while True:
            cells, vals, n = np.empty(capacity, np.int32), np.empty(capacity), C.c_int32(0)
            rc = lib().tcif_laser_get_sparse(self.h, int(capacity), _i(cells), _d(vals), C.byref(n))
            if rc and n.value > capacity:
                capacity = int(n.value)
                continue
            self._ck(rc)
            return cells[:n.value].copy(), vals[:n.value].copy()

    def laser_get(self):
        Q = np.zeros(self.nCells); refl = np.zeros(self.nCells, np.int32)
        nHat = np.zeros((self.nCells, 3)); nPts = np.zeros((self._nPoints, 3))
        self._ck(lib().tcif_laser_get(self.h, _d(Q), _i(refl), _d(nHat), _d(nPts)))
        return dict(Q=Q, reflectingCells=refl, nHat=nHat, nHatPoints=nPts)

    @property
    def laser_generations(self):
        return lib().tcif_laser_generations(self.h)

    def solve_yi(self, phase, Y_i, Y_b, Ybc, inertIndex, deltaT, cYi=1.0, schemePhi="vanLeer", schemePhir="linear",
                 Su=None, Sp=None, nYiSubCycles=1, tol=1e-12, maxIter=200):
        """MultiComponentPhaseModel::solveYi for the species of `phase`; returns new (Y_i, Y_b) lists."""
        S = len(Y_i)
        yi = [_f64(a).copy() for a in Y_i]
        yb = [_f64(a).copy() for a in Y_b]
        bc = np.ascontiguousarray(np.asarray(Ybc, dtype=np.int32).reshape(S, -1))
        dpp = C.POINTER(C.c_double) * S

        def pp(arrs):
            if arrs is None:
                return None
            keep = [(_f64(a) if a is not None else None) for a in arrs]
            return dpp(*[(_d(a) if a is not None else None) for a in keep]), keep
        su, sp = pp(Su), pp(Sp)
        self._ck(lib().tcif_solve_yi(self.h, int(phase), S, int(inertIndex), dpp(*[_d(a) for a in yi]),
                                     dpp(*[_d(a) for a in yb]), bc.ctypes.data_as(C.POINTER(C.c_int32)),
                                     su[0] if su else None, sp[0] if sp else None, C.c_double(cYi),
                                     SCHEMES[schemePhi], SCHEMES[schemePhir], int(nYiSubCycles), C.c_double(deltaT),
                                     C.c_double(tol), int(maxIter)))
        return yi, yb

    # -- hot path
    def div_capillary_stress(self):
        self._ck(lib().tcif_div_capillary_stress(self.h))

    def get_force(self):
        Fi, Fb = np.empty((self.nCells, 3)), np.empty((self.nBF, 3))
        self._ck(lib().tcif_get_force(self.h, _d(Fi), _d(Fb)))
        return Fi, Fb

    def div_capillary_stress_host(self, alphas_i, alphas_b, Ti=None, Tb=None, Fi=None, Fb=None):
        ai = [_f64(a) for a in alphas_i]
        ab = [_f64(a) for a in alphas_b]
        Ti = _f64(Ti) if Ti is not None else None
        Tb = _f64(Tb) if Tb is not None else None
        Fi = np.empty((self.nCells, 3)) if Fi is None else Fi
        Fb = np.empty((self.nBF, 3)) if Fb is None else Fb
        self._ck(lib().tcif_div_capillary_stress_host(self.h, _pp(ai), _pp(ab), _d(Ti), _d(Tb), _d(Fi), _d(Fb)))
        return Fi, Fb

    def flux_of_force(self):
        a, b = np.empty(self.nIF), np.empty(self.nBF)
        self._ck(lib().tcif_flux_of_force(self.h, _d(a), _d(b)))
        return a, b

    def set_force_cache(self, on=True):
        self._ck(lib().tcif_set_force_cache(self.h, int(bool(on))))

    def force_counters(self):
        e, h = C.c_int64(), C.c_int64()
        self._ck(lib().tcif_force_counters(self.h, C.byref(e), C.byref(h)))
        return int(e.value), int(h.value)

    def grad_alpha(self, phase):
        gi, gb = np.empty((self.nCells, 3)), np.empty((self.nBF, 3))
        self._ck(lib().tcif_grad_alpha(self.h, int(phase), _d(gi), _d(gb)))
        return gi, gb

    def mules_limit(self, phase, rDeltaT, returnCorr):
        self._ck(lib().tcif_mules_limit(self.h, int(phase), C.c_double(rDeltaT), int(bool(returnCorr))))

    def mules_limit_sum(self):
        self._ck(lib().tcif_mules_limit_sum(self.h))

    def add_upwind_flux(self, phase):
        self._ck(lib().tcif_add_upwind_flux(self.h, int(phase)))

    def mules_explicit_solve(self, phase, deltaT):
        self._ck(lib().tcif_mules_explicit_solve(self.h, int(phase), C.c_double(deltaT)))

    def stage(self, deltaT):
        self._ck(lib().tcif_stage(self.h, C.c_double(deltaT)))

    def stage_host(self, deltaT, alphas_i, alphas_b, Ti, Tb, phii, phib, phiAlpha_i, phiAlpha_b, alphaPhi_i, alphaPhi_b,
                   Fi, Fb):
        """alphas in/out, phiAlpha (HO flux) read-only, alphaPhi/F outputs; float64 C-contiguous arrays."""
        for a in list(alphas_i) + list(alphas_b) + list(phiAlpha_i) + list(phiAlpha_b) + list(alphaPhi_i) + \
                list(alphaPhi_b) + [Fi, Fb]:
            assert a.dtype == np.float64 and a.flags.c_contiguous
        self._ck(lib().tcif_stage_host(self.h, C.c_double(deltaT), _pp(alphas_i), _pp(alphas_b), _d(Ti), _d(Tb),
                                       _d(phii), _d(phib), _pp(phiAlpha_i), _pp(phiAlpha_b), _pp(alphaPhi_i),
                                       _pp(alphaPhi_b), _d(Fi), _d(Fb)))

    # -- rest of solveAlphas
    def set_cell_centres(self, Ccells, deltaBoundary):
        a, b = _f64(Ccells), _f64(deltaBoundary)
        assert a.shape == (self.nCells, 3) and b.shape == (self.nBF, 3)
        self._ck(lib().tcif_set_cell_centres(self.h, _d(a), _d(b)))

    def set_calpha(self, table):
        t = _f64(table)
        assert t.shape == (self.P, self.P)
        self._ck(lib().tcif_set_calpha(self.h, _d(t)))

    def set_rho(self, phase, uniform, internal=None, boundary=None):
        i = _f64(internal) if internal is not None else None
        b = _f64(boundary) if boundary is not None else None
        self._ck(lib().tcif_set_rho(self.h, int(phase), C.c_double(uniform), _d(i), _d(b)))

    def flux_assembly(self, schemePhi="vanLeer", schemePhir="linear"):
        self._ck(lib().tcif_flux_assembly(self.h, SCHEMES[schemePhi], SCHEMES[schemePhir]))

    def get_flux_assembly(self, phase):
        a, b = np.empty(self.nIF), np.empty(self.nBF)
        self._ck(lib().tcif_get_flux_assembly(self.h, int(phase), _d(a), _d(b)))
        return a, b

    def upwind_predictor(self, phase, deltaT, tol=1e-12, maxIter=200):
        self._ck(lib().tcif_upwind_predictor(self.h, int(phase), C.c_double(deltaT), C.c_double(tol), int(maxIter)))
        return int(lib().tcif_predictor_sweeps(self.h, int(phase)))

    def solve_alphas(self, ctl: AlphaControls, deltaT):
        self._ck(lib().tcif_solve_alphas(self.h, C.byref(ctl), C.c_double(deltaT)))

    def time_step_host(self, ctl, deltaT, alphas_i, alphas_b, Ti, Tb, phii, phib, rhoPhi_i, rhoPhi_b, Fi, Fb,
                       alphaPhi_i=None, alphaPhi_b=None):
        """solve() + divCapillaryStress() with host buffers (alphas in/out; rhoPhi, F, optional alphaPhi out)."""
        self._ck(lib().tcif_time_step_host(self.h, C.byref(ctl), C.c_double(deltaT), _pp(alphas_i), _pp(alphas_b), _d(Ti),
                                           _d(Tb), _d(phii), _d(phib), _pp(alphaPhi_i) if alphaPhi_i else None,
                                           _pp(alphaPhi_b) if alphaPhi_b else None, _d(rhoPhi_i), _d(rhoPhi_b), _d(Fi),
                                           _d(Fb)))

    def host_alpha_stamps(self, stamps_in=None, stamps_out=None):
        """one-shot stamps of the alpha arrays of the NEXT stage_host / time_step_host call (tcif_host_alpha_stamps)"""
        P = self.P
        a = (C.c_uint64 * P)(*[int(x) for x in (stamps_in or [0] * P)])
        b = (C.c_uint64 * P)(*[int(x) for x in (stamps_out or [0] * P)])
        self._ck(lib().tcif_host_alpha_stamps(self.h, a, b))

    def host_sparse_returns(self, on=True):
        self._ck(lib().tcif_host_sparse_returns(self.h, int(bool(on))))

    @property
    def host_bytes_moved(self):
        a, b = C.c_int64(), C.c_int64()
        self._ck(lib().tcif_host_bytes_moved(self.h, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def get_rho_phi(self):
        a, b = np.empty(self.nIF), np.empty(self.nBF)
        self._ck(lib().tcif_get_rho_phi(self.h, _d(a), _d(b)))
        return a, b

    def alpha_stats(self, phase):
        out = (C.c_double * 3)()
        self._ck(lib().tcif_alpha_stats(self.h, int(phase), out))
        return float(out[0]), float(out[1]), float(out[2])

    def alpha_stats_static(self, phase):
        """tcif_alpha_stats_static: rank-local (sum V alpha, min, max) with the scratch in a static device array."""
        out = (C.c_double * 3)()
        self._ck(lib().tcif_alpha_stats_static(self.h, int(phase), out))
        return float(out[0]), float(out[1]), float(out[2])

    # -- measurement
    def timers_enable(self, on=True):
        self._ck(lib().tcif_timers_enable(self.h, int(bool(on))))

    def timers_reset(self):
        self._ck(lib().tcif_timers_reset(self.h))

    def timers_read(self):
        n = lib().tcif_timer_count()
        ms, ln, by = (C.c_double * n)(), (C.c_int64 * n)(), (C.c_double * n)()
        self._ck(lib().tcif_timers_read(self.h, ms, ln, by))
        return {lib().tcif_timer_name(i).decode(): dict(ms=ms[i], launches=int(ln[i]), bytes=by[i]) for i in range(n)}

    @property
    def launch_count(self):
        return int(lib().tcif_launch_count(self.h))

    @property
    def device_bytes(self):
        return int(lib().tcif_device_bytes(self.h))

    def debug_guard_check(self):
        """(zones, corrupted zones) of a context created under TCIF_GUARD=1."""
        nz, nc = C.c_int64(0), C.c_int64(0)
        self._ck(lib().tcif_debug_guard_check(self.h, C.byref(nz), C.byref(nc)))
        return int(nz.value), int(nc.value)

    @property
    def mesh_info(self):
        out = (C.c_int32 * 8)()
        self._ck(lib().tcif_mesh_info(self.h, out))
        return dict(Ko=out[0], Kn=out[1], fixedK=out[2], tileStrips=out[3], prefetchTiles=out[4], boundaryCells=out[5],
                    deferredCells=out[6], processorFaces=out[7])

    @property
    def comm_transport(self):
        return {0: "none", 1: "local", 2: "nccl", 3: "peer"}[int(lib().tcif_comm_transport(self.h))]

    @property
    def tile_strips(self):
        return self.mesh_info["tileStrips"]
