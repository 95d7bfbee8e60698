"""ctypes binding of include/tcif_mesh.h (libtcif_host.so): hex block meshes in blockMesh ordering and
processor sub-meshes in decomposePar conventions, as numpy views on the C++ object's arrays."""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

from .abi import MeshDesc, TcifError, c_dp, c_ip

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtcif_host.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise TcifError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        L = C.CDLL(LIB_PATH)
        L.tcif_mesh_last_error.restype = C.c_char_p
        L.tcif_mesh_free.argtypes = [C.c_void_p]
        L.tcif_mesh_free.restype = None
        _lib = L
    return _lib


def _view(ptr, n, dtype, shape=None):
    if n == 0:
        return np.zeros(shape if shape else (0,), dtype)
    a = np.ctypeslib.as_array(ptr, shape=(n,))
    a = a.view(dtype) if a.dtype != dtype else a
    return a.reshape(shape) if shape else a


class HostMesh:
    """Owns a tcif_mesh; attributes mirror tcif_mesh_desc plus centres and decomposition addressing."""

    def __init__(self, handle):
        L = lib()
        self.h = handle
        d = MeshDesc()
        if L.tcif_mesh_get_desc(self.h, C.byref(d)):
            raise TcifError(L.tcif_mesh_last_error().decode())
        self.desc = d
        self.nCells, nIF, nP, nBF = d.nCells, d.nInternalFaces, d.nPatches, d.nBoundaryFaces
        self.owner = _view(d.owner, nIF, np.int32)
        self.neighbour = _view(d.neighbour, nIF, np.int32)
        self.Sf = _view(d.Sf, 3 * nIF, np.float64, (nIF, 3))
        self.w = _view(d.weights, nIF, np.float64)
        self.V = _view(d.V, self.nCells, np.float64)
        self.patchType = _view(d.patchType, nP, np.int32)
        self.patchStart = _view(d.patchStart, nP, np.int32)
        self.patchSize = _view(d.patchSize, nP, np.int32)
        self.patchNbrProc = _view(d.patchNeighbProcNo, nP, np.int32)
        self.faceCells = _view(d.faceCells, nBF, np.int32)
        self.Sfb = _view(d.SfBoundary, 3 * nBF, np.float64, (nBF, 3))
        self.magSfb = _view(d.magSfBoundary, nBF, np.float64)
        self.deltaCoeffsb = _view(d.deltaCoeffsBoundary, nBF, np.float64)
        self.wb = _view(d.weightsBoundary, nBF, np.float64)
        self.C = self._f64("C", (self.nCells, 3))
        self.Cf = self._f64("Cf", (nIF, 3))
        self.Cfb = self._f64("Cfb", (nBF, 3))
        self.nbrC = self._f64("nbrC", (nBF, 3))
        self.cellGlobal = self._i64("cellGlobal")
        self.nbrGlobal = self._i64("nbrGlobal")
        self.faceProcAddressing = self._i64("faceProcAddressing")
        self.bfaceProcAddressing = self._i64("bfaceProcAddressing")
        self.boundaryProcAddressing = self._i64("boundaryProcAddressing")

    nIF = property(lambda self: int(self.owner.shape[0]))
    nBF = property(lambda self: int(self.faceCells.shape[0]))
    nPatches = property(lambda self: int(self.patchType.shape[0]))

    def _f64(self, name, shape):
        p, n = c_dp(), C.c_int64()
        if lib().tcif_mesh_array_f64(self.h, name.encode(), C.byref(p), C.byref(n)):
            raise TcifError(lib().tcif_mesh_last_error().decode())
        return _view(p, n.value, np.float64, shape)

    def _i64(self, name):
        p, n = C.POINTER(C.c_int64)(), C.c_int64()
        if lib().tcif_mesh_array_i64(self.h, name.encode(), C.byref(p), C.byref(n)):
            raise TcifError(lib().tcif_mesh_last_error().decode())
        return _view(p, n.value, np.int64)

    def close(self):
        if self.h:
            lib().tcif_mesh_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def vertices(n, L, grading=None):
    t = np.arange(n + 1, dtype=np.float64) / n
    if grading is not None and grading != 1.0:
        t = (np.power(grading, t) - 1.0) / (grading - 1.0)
    return L * t


def box_block(n, lengths=(1.0, 1.0, 1.0), patch_types=(0,) * 6, procs=(1, 1, 1), rank=0, grading=None) -> HostMesh:
    """Sub-mesh `rank` of the `simple procs` decomposition of an nx*ny*nz box (procs=(1,1,1): whole mesh)."""
    g = grading or (None, None, None)
    xs, ys, zs = (np.ascontiguousarray(vertices(n[d], lengths[d], g[d])) for d in range(3))
    pt = np.ascontiguousarray(patch_types, dtype=np.int32)
    h = C.c_void_p()
    rc = lib().tcif_mesh_box_block(int(n[0]), int(n[1]), int(n[2]), xs.ctypes.data_as(c_dp), ys.ctypes.data_as(c_dp),
                                   zs.ctypes.data_as(c_dp), pt.ctypes.data_as(c_ip), int(procs[0]), int(procs[1]),
                                   int(procs[2]), int(rank), C.byref(h))
    if rc:
        raise TcifError(lib().tcif_mesh_last_error().decode())
    return HostMesh(h)


def simple_map(n, procs) -> np.ndarray:
    out = np.zeros(n[0] * n[1] * n[2], np.int32)
    if lib().tcif_mesh_simple_map(int(n[0]), int(n[1]), int(n[2]), int(procs[0]), int(procs[1]), int(procs[2]),
                                  out.ctypes.data_as(c_ip)):
        raise TcifError(lib().tcif_mesh_last_error().decode())
    return out


def decompose(g: HostMesh, cellProc, nProcs, rank) -> HostMesh:
    cp = np.ascontiguousarray(cellProc, dtype=np.int32)
    assert cp.shape[0] == g.nCells
    h = C.c_void_p()
    if lib().tcif_mesh_decompose(g.h, cp.ctypes.data_as(c_ip), int(nProcs), int(rank), C.byref(h)):
        raise TcifError(lib().tcif_mesh_last_error().decode())
    return HostMesh(h)


def read_polymesh(polyMeshDir) -> HostMesh:
    """An OpenFOAM ASCII polyMesh directory (any cell shape) -> host mesh (tcif_mesh_read_polymesh)."""
    h = C.c_void_p()
    if lib().tcif_mesh_read_polymesh(str(polyMeshDir).encode(), C.byref(h)):
        raise TcifError(lib().tcif_mesh_last_error().decode())
    return HostMesh(h)


def read_label_list(path) -> np.ndarray:
    """A labelList file such as decomposePar's cellDecomposition (tcif_mesh_read_label_list)."""
    n = C.c_int32(0)
    lib().tcif_mesh_read_label_list(str(path).encode(), 0, None, C.byref(n))
    if n.value == 0:
        raise TcifError(lib().tcif_mesh_last_error().decode())
    out = np.zeros(n.value, np.int32)
    if lib().tcif_mesh_read_label_list(str(path).encode(), n.value, out.ctypes.data_as(c_ip), C.byref(n)):
        raise TcifError(lib().tcif_mesh_last_error().decode())
    return out


def procs_for(nRanks: int):
    """Block layout used for 1/2/4/8 GPUs."""
    return {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}[nRanks]
