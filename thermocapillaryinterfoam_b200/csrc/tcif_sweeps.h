// tcif_sweeps.h -- host-side launch table of the fixed-K sweeps (tcif_volume.cuh).  The kernels are instantiated
// per slot count K in their own translation units (tcif_sweeps_k2.cu, _k3.cu, _k4.cu: compiled in parallel),
// tcif.cu picks the table of the mesh at hand.  def = the O(surface) launch over DMesh::bndCells.
#pragma once
#include "tcif_device.cuh"
#include "tcif_alpha.cuh"

namespace tcif {

struct SweepFns {
    // grid sizes are computed inside from DMesh (nTiles for the volume launch, nBndCells for the deferred one)
    void (*grad_tensor)(int P, bool def, cudaStream_t st, const DMesh &m, const ForceArgs &a);
    void (*div_tensor)(int P, bool def, cudaStream_t st, const DMesh &m, const ForceArgs &a);
    void (*setup)(int nPh, bool def, cudaStream_t st, const DMesh &m, const MulesBatch &b, const MulesParams &prm);
    void (*iter)(int nPh, bool first, bool def, cudaStream_t st, const DMesh &m, const MulesBatch &b);
    void (*fin)(int nPh, int mode, bool first, bool lsum, bool def, cudaStream_t st, const DMesh &m, const MulesBatch &b,
                const MulesParams &prm, int nIter);
    // surfaceIntegrate + Euler update from the stored fluxes (second half of the split explicitSolve)
    void (*update)(int nPh, bool def, cudaStream_t st, const DMesh &m, const MulesBatch &b, const MulesParams &prm);
    void (*pred_jacobi)(int nPh, bool def, cudaStream_t st, const DMesh &m, const PredBatch &b);
};

const SweepFns *sweep_fns_k2();
const SweepFns *sweep_fns_k3();
const SweepFns *sweep_fns_k4();

}  // namespace tcif
