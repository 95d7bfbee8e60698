// tcif_species.cuh -- flux assembly of the species MULES:
//   MultiComponentPhaseModel::solveYi   phaseModel/MultiComponentPhaseModel/MultiComponentPhaseModel.C:182-297
//     phic  = min(cYi*|phi|/|Sf|, max(phic))                                             (:199-201)
//     phir  = sum_{alpha2 != alpha1} phic * nHatf(alpha1, alpha2); 0 on non-coupled patches   (:203-228)
//     phiYiCorr = fvc::flux(phi, Yi, div(phi,Yi))
//               + sum_{alpha2 != alpha1} fvc::flux(-fvc::flux(-phir, alpha2, S), Yi, S)   (:236-276)
//     inflow patches keep phi_b*Yi_b                                                     (:279-297)
// One face-local sweep (the reference: ~6 surface fields per species and phase).  The limiter, the implicit
// upwind predictor and explicitSolve that follow are the MULES kernels of the phase fractions, run on Yi.
#pragma once
#include "tcif_alpha.cuh"

namespace tcif {

struct YiFluxArgs {
    int P, ph, schemePhi, schemePhir;
    const double *alpha[4], *alphab[4];
    const double *GA, *GAb;          // grad(alpha_p): [(p*3+d)*stride + c], [(p*3+d)*nBF + bf] (processor: neighbour cell)
    const double *Y, *Yb;
    const double *GY, *GYb;          // grad(Yi): [d*stride + c], [d*nBF + bf]
    const double *Cx, *Cy, *Cz, *dbx, *dby, *dbz;
    const double *phi, *phib;
    double deltaN, cYi, maxPhic;
    double *out, *outb;
};

// flux of ONE face from the two cell states; COUPLED selects OpenFOAM's w*P + (1-w)*N form of processor patches
template <int P, bool COUPLED>
__device__ __forceinline__ double yi_face_flux(const YiFluxArgs &a, double phi, double w, double sx, double sy, double sz,
                                               const double *aP, const double *aN, const double (*gP)[3], const double (*gN)[3],
                                               double yP, double yN, const double *gyP, const double *gyN,
                                               const double *d, double phic0) {
    auto mix = [&](double W, double vP, double vN) { return COUPLED ? W * vP + (1.0 - W) * vN : W * (vP - vN) + vN; };
    double af[P], gfa[P][3];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        af[p] = mix(w, aP[p], aN[p]);
#pragma unroll
        for (int k = 0; k < 3; ++k) gfa[p][k] = mix(w, gP[p][k], gN[p][k]);
    }
    const double phic = dmin(a.cYi * phic0, a.maxPhic);
    double phir = 0.0 * phi;
#pragma unroll
    for (int q = 0; q < P; ++q) {
        if (q == a.ph) continue;
        // nHatf(alpha1, alpha_q); evaluated for the ordered pair (min, max) and negated, as face_fluxes does
        const int p1 = a.ph < q ? a.ph : q, p2 = a.ph < q ? q : a.ph;
        double f1 = 0, f2 = 0, g1[3] = {0, 0, 0}, g2[3] = {0, 0, 0};
#pragma unroll
        for (int p = 0; p < P; ++p) {
            if (p == p1) { f1 = af[p]; g1[0] = gfa[p][0]; g1[1] = gfa[p][1]; g1[2] = gfa[p][2]; }
            if (p == p2) { f2 = af[p]; g2[0] = gfa[p][0]; g2[1] = gfa[p][1]; g2[2] = gfa[p][2]; }
        }
        const double gx = f2 * g1[0] - f1 * g2[0], gy = f2 * g1[1] - f1 * g2[1], gz = f2 * g1[2] - f1 * g2[2];
        const double rden = f_rcp(sqrt(gx * gx + gy * gy + gz * gz) + a.deltaN);
        double s = f_div(gx, rden) * sx;
        s += f_div(gy, rden) * sy;
        s += f_div(gz, rden) * sz;
        phir += phic * (a.ph < q ? s : -s);
    }
    double W = scheme_weight(a.schemePhi, w, phi, yP, yN, gyP, gyN, d);
    double f = phi * mix(W, yP, yN);
#pragma unroll
    for (int q = 0; q < P; ++q) {
        if (q == a.ph) continue;
        const double t1 = -phir;
        const double W2 = scheme_weight(a.schemePhir, w, t1, aP[q], aN[q], gP[q], gN[q], d);
        const double inner = -(t1 * mix(W2, aP[q], aN[q]));
        const double W1 = scheme_weight(a.schemePhir, w, inner, yP, yN, gyP, gyN, d);
        f += inner * mix(W1, yP, yN);
    }
    return f;
}

// internal faces: thread c handles the faces it owns
template <int P>
__global__ void __launch_bounds__(128, 3)
k_flux_yi(const __grid_constant__ DMesh m, const __grid_constant__ YiFluxArgs a) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= m.nC) return;
    double aP[P], gP[P][3], gyP[3];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        aP[p] = a.alpha[p][c];
#pragma unroll
        for (int k = 0; k < 3; ++k) gP[p][k] = a.GA[(size_t)(p * 3 + k) * m.stride + c];
    }
    const double yP = a.Y[c];
#pragma unroll
    for (int k = 0; k < 3; ++k) gyP[k] = a.GY[(size_t)k * m.stride + c];
    const bool geo = a.Cx != nullptr;
    double cP[3] = {0, 0, 0};
    if (geo) { cP[0] = a.Cx[c]; cP[1] = a.Cy[c]; cP[2] = a.Cz[c]; }
    for (int k = 0; k < m.Ko; ++k) {
        const int fo = k * m.stride + c;
        const int n = m.nbrOwn[fo];
        if (n < 0) continue;
        double aN[P], gN[P][3], gyN[3], d[3] = {0, 0, 0};
#pragma unroll
        for (int p = 0; p < P; ++p) {
            aN[p] = a.alpha[p][n];
#pragma unroll
            for (int kk = 0; kk < 3; ++kk) gN[p][kk] = a.GA[(size_t)(p * 3 + kk) * m.stride + n];
        }
        const double yN = a.Y[n];
#pragma unroll
        for (int kk = 0; kk < 3; ++kk) gyN[kk] = a.GY[(size_t)kk * m.stride + n];
        if (geo) { d[0] = a.Cx[n] - cP[0]; d[1] = a.Cy[n] - cP[1]; d[2] = a.Cz[n] - cP[2]; }
        const double sx = m.Sfx[fo], sy = m.Sfy[fo], sz = m.Sfz[fo], phi = a.phi[fo];
        const double phic0 = fabs(phi / sqrt(sx * sx + sy * sy + sz * sz));
        a.out[fo] = yi_face_flux<P, false>(a, phi, m.w[fo], sx, sy, sz, aP, aN, gP, gN, yP, yN, gyP, gyN, d, phic0);
    }
}

// boundary faces: processor faces use the halo values; elsewhere phi_b*Yi_b (phir_b == 0 there, and the inflow
// fix of :279-297 writes the same product)
template <int P>
__global__ void __launch_bounds__(TCIF_TPB)
k_flux_yi_bnd(const __grid_constant__ DMesh m, const __grid_constant__ YiFluxArgs a) {
    const int bf = blockIdx.x * blockDim.x + threadIdx.x;
    if (bf >= m.nBF) return;
    const double phi = a.phib[bf];
    if (m.bfType[bf] != PATCH_PROCESSOR) { a.outb[bf] = phi * a.Yb[bf]; return; }
    const int c = m.faceCells[bf];
    double aP[P], aN[P], gP[P][3], gN[P][3], gyP[3], gyN[3], d[3] = {0, 0, 0};
#pragma unroll
    for (int p = 0; p < P; ++p) {
        aP[p] = a.alpha[p][c]; aN[p] = a.alphab[p][bf];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            gP[p][k] = a.GA[(size_t)(p * 3 + k) * m.stride + c];
            gN[p][k] = a.GAb[(size_t)(p * 3 + k) * m.nBF + bf];
        }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) { gyP[k] = a.GY[(size_t)k * m.stride + c]; gyN[k] = a.GYb[(size_t)k * m.nBF + bf]; }
    if (a.dbx) { d[0] = a.dbx[bf]; d[1] = a.dby[bf]; d[2] = a.dbz[bf]; }
    const double phic0 = fabs(phi / m.magSfb[bf]);
    a.outb[bf] = yi_face_flux<P, true>(a, phi, m.wb[bf], m.Sfbx[bf], m.Sfby[bf], m.Sfbz[bf], aP, aN, gP, gN,
                                       a.Y[c], a.Yb[bf], gyP, gyN, d, phic0);
}

}  // namespace tcif
