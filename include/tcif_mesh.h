/*
 * tcif_mesh.h -- host-side mesh model (C ABI of libtcif_host.so, plain C++, no CUDA).
 *
 * Stands in for what the OpenFOAM shim reads off an fvMesh and for the two utilities the
 * reference workflow relies on but does not contain (blockMesh for single-block hex cases,
 * decomposePar for processor sub-meshes; reference README.md:47-49 pins OpenFOAM v2106, the
 * WENOEXT parallel test decomposes with `method simple; n (2 2 2)`,
 * WENOEXT/tests/Cases/globalFvMeshTestCase/system/decomposeParDict:17-25).
 * Conventions: cell id = i + nx*(j + ny*k); internal faces in
 * upper-triangular order; patches xMin,xMax,yMin,yMax,zMin,zMax; `empty` patches carry no
 * field faces; a sub-mesh keeps global relative cell/face order, original patches first, then
 * one processor patch per neighbour rank (ascending), faces in global face order, flipped where
 * the local cell is the global neighbour.
 */
#ifndef TCIF_MESH_H
#define TCIF_MESH_H

#include <stdint.h>
#include "tcif.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tcif_mesh tcif_mesh;

/* Processor sub-mesh `rank` of the `simple (px py pz)` decomposition of a rectilinear hex block
 * with vertex coordinates xs[nx+1], ys[ny+1], zs[nz+1]; px=py=pz=1, rank=0 is the undecomposed
 * mesh.  Built directly (the global mesh is never materialised). */
TCIF_API int tcif_mesh_box_block(int32_t nx, int32_t ny, int32_t nz, const double *xs, const double *ys,
                                 const double *zs, const int32_t patchTypes[6], int32_t px, int32_t py,
                                 int32_t pz, int32_t rank, tcif_mesh **out);
/* Processor sub-mesh `rank` of ANY cell->rank map (e.g. a scotch cellDecomposition labelList). */
TCIF_API int tcif_mesh_decompose(const tcif_mesh *global, const int32_t *cellProc, int32_t nProcs, int32_t rank,
                                 tcif_mesh **out);
/* `simple (px py pz)` cell->rank map of a uniform box. */
TCIF_API int tcif_mesh_simple_map(int32_t nx, int32_t ny, int32_t nz, int32_t px, int32_t py, int32_t pz,
                                  int32_t *cellProc);
/* An OpenFOAM ASCII polyMesh directory (<case>/constant/polyMesh or <case>/processorN/constant/polyMesh: points,
 * faces, owner, neighbour, boundary) of ANY cell shape; geometry by OpenFOAM's primitiveMesh formulas, linear weights.
 * Patch types: patch | wall | symmetry(Plane) (generic), empty, wedge, processor (neighbProcNo is read). */
TCIF_API int tcif_mesh_read_polymesh(const char *polyMeshDir, tcif_mesh **out);
/* A labelList file, e.g. the `cellDecomposition` that decomposePar -cellDist writes for scotch: *count labels into
 * labels[capacity] (call with capacity 0 to learn the count; returns 1 when capacity is too small). */
TCIF_API int tcif_mesh_read_label_list(const char *path, int32_t capacity, int32_t *labels, int32_t *count);
TCIF_API void tcif_mesh_free(tcif_mesh *m);
/* Borrowed view for tcif_create (valid while the mesh lives). */
TCIF_API int tcif_mesh_get_desc(const tcif_mesh *m, tcif_mesh_desc *out);
/* Named extra arrays.  double: "C" [nCells*3], "Cf" [nIF*3], "Cfb" [nBF*3], "nbrC" [nBF*3]
 * (neighbour-rank cell centre on processor faces).  int64: "cellGlobal" (cellProcAddressing),
 * "nbrGlobal" [nBF], "faceProcAddressing" [nIF] (global face + 1), "bfaceProcAddressing" [nBF]
 * (+-(global internal face + 1) on processor faces, global boundary face + 1 elsewhere),
 * "boundaryProcAddressing" [nPatches].  Returns 0 and sets ptr/count, or 1 if unknown. */
TCIF_API int tcif_mesh_array_f64(const tcif_mesh *m, const char *name, const double **ptr, int64_t *count);
TCIF_API int tcif_mesh_array_i64(const tcif_mesh *m, const char *name, const int64_t **ptr, int64_t *count);
TCIF_API const char *tcif_mesh_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
