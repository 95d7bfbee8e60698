// tcif_mass.cuh -- Su/Sp of the phase-fraction equations and the porous interface drag on the device
// (SURVEY.md 8(f)-3).  Replaces, for the MULES path,
//   MassTransferPhaseSystem::alphaTransfer(Su, Sp)    MassTransferPhaseSystem.C:503-704
//   MassTransferPhaseSystem::correctMassSources(T)    MassTransferPhaseSystem.C:447-497
//   meltingEvaporationModels::Lee::Kexp / KSu         massTransferModels/Lee/Lee.C:52-85, 130-167
//   phaseSystem::coeffs = 1/rho                       phaseSystem.C:1139-1145
//   phaseSystem::addInterfacePorosity + VollerPrakash::S()   phaseSystem.C:1148-1183, VollerPrakash.C:66-72
// One cell-centric sweep does what the reference does in ~12 full-field sweeps per pair: fvc::div(phi) is
// gathered over the cell's faces once (reference accumulation order), every pair is visited in
// totalPhasePairs order with Su/Sp of all phases in registers, and each output array is written once
// (solveAlphas zeroes Su/Sp immediately before, multiphaseSystem.C:282-286, so the result is stored, not
// accumulated).  Compulsory HBM bytes per cell: slot indices 4(Ko+Kn), phi 8Ko, V 8, alpha 8P, stored dmdt
// 8 per model, Su/Sp 16P.
#pragma once
#include "tcif_device.cuh"

namespace tcif {

enum { MASS_VAR_T = 0, MASS_VAR_ALPHA = 1 };
#define TCIF_MAX_MASS_MODELS 12
#define TCIF_MAX_PAIRS 6

struct MassArgs {
    int P, nPairs, nModels;
    int pairs[2 * TCIF_MAX_PAIRS];             // (phase1, phase2) in totalPhasePairs iteration order
    int m12[TCIF_MAX_PAIRS], m21[TCIF_MAX_PAIRS];   // model index phase1->phase2 / phase2->phase1 (-1: none)
    int divU[TCIF_MAX_PAIRS];                  // value of includeDivU when the pair is visited (:512, 555, 583)
    int anyDivU;
    int mFrom[TCIF_MAX_MASS_MODELS], mVar[TCIF_MAX_MASS_MODELS];
    double mC[TCIF_MAX_MASS_MODELS], mTact[TCIF_MAX_MASS_MODELS], mAmin[TCIF_MAX_MASS_MODELS];
    const double *dmdt[TCIF_MAX_MASS_MODELS];  // stored dmdt_[key] (null = zero field)
    const double *alpha[4];
    double rhoU[4];
    const double *rho[4];                      // thermo().rho() as a field (null: uniform rhoU)
    const double *phi, *phib;
    double *Su[4], *Sp[4];
    long long *ddtMax;                         // [2] order-preserving integer image of max(dmdt21*coeffs1), max(dmdt12*coeffs2)
};

__device__ __forceinline__ double pos0(double x) { return x >= 0 ? 1.0 : 0.0; }          // Foam::pos
__device__ __forceinline__ double clamp01(double a) { return dmin(dmax(a, 0.0), 1.0); }  // min(max(a,0),1)

// doubles <-> integers with the same ordering, so that atomicMax(long long) is a max over doubles
__device__ __host__ __forceinline__ long long dbl_ord(double x) {
    long long b;
    memcpy(&b, &x, 8);
    return b >= 0 ? b : (b ^ 0x7fffffffffffffffLL);
}
__device__ __host__ __forceinline__ double ord_dbl(long long b) {
    b = b >= 0 ? b : (b ^ 0x7fffffffffffffffLL);
    double x;
    memcpy(&x, &b, 8);
    return x;
}

// Lee::KSu(alpha, refValue) (Lee.C:130-167), only called when the model variable is alpha
__device__ __forceinline__ double lee_ksu(const MassArgs &a, int i, double alphaFrom, double rhoFrom, double ref) {
    double from = clamp01(alphaFrom);
    double coeff = a.mC[i] * from * rhoFrom * pos0(from - a.mAmin[i]);
    return a.mC[i] > 0 ? -coeff * pos0(ref - a.mTact[i]) : -coeff * pos0(a.mTact[i] - ref);
}

// correctMassSources: dmdt_[key] = Kexp(T) (Lee.C:52-85)
__global__ void __launch_bounds__(TCIF_TPB)
k_lee_kexp(int nC, double C, double Tact, double aMin, const double *alphaFrom, const double *rhoField, double rhoU,
           const double *T, double *out) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nC) return;
    double from = clamp01(alphaFrom[c]);
    double rho = rhoField ? rhoField[c] : rhoU;
    double Tc = T[c];
    double coeff = C * from * rho * pos0(from - aMin) * (Tc - Tact) / Tact;
    out[c] = C > 0 ? coeff * pos0(Tc - Tact) : coeff * pos0(Tact - Tc);
}

// alphaTransfer for every pair; BND=false: interior cells (internal faces only), BND=true: the boundary-adjacent
// cells including their patch faces.  LAST pair's maxima feed ddtAlphaMax (:700-702 overwrites it per pair).
template <bool BND>
__global__ void __launch_bounds__(TCIF_TPB)
k_alpha_transfer(const __grid_constant__ DMesh m, const __grid_constant__ MassArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int c = -1;
    if (BND) { if (i < m.nBndCells) c = m.bndCells[i]; }
    else     { if (i < m.nC && !is_bnd_cell(m, i)) c = i; }
    double mx21 = -1e300, mx12 = -1e300;
    if (c >= 0) {
        // fvc::div(phi) = surfaceIntegrate: neighbour-side faces (-), owned faces (+), patch faces (+), then / V
        double divPhi = 0.0;
        if (a.anyDivU) {
            for (int k = 0; k < m.Kn; ++k) {
                int fd = m.neiFace[k * m.stride + c];
                if (fd >= 0) divPhi -= a.phi[fd];
            }
            for (int k = 0; k < m.Ko; ++k) {
                size_t fd = (size_t)k * m.stride + c;
                if (m.nbrOwn[fd] >= 0) divPhi += a.phi[fd];
            }
            if (BND)
                for (int j = m.bndCellStart[i]; j < m.bndCellStart[i + 1]; ++j) divPhi += a.phib[m.bndFaces[j]];
            divPhi /= m.V[c];
        }
        double al[4], rh[4], Su[4], Sp[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            if (p < a.P) { al[p] = a.alpha[p][c]; rh[p] = a.rho[p] ? a.rho[p][c] : a.rhoU[p]; }
            else { al[p] = 0.0; rh[p] = 1.0; }
            Su[p] = 0.0; Sp[p] = 0.0;
        }
        for (int q = 0; q < a.nPairs; ++q) {
            const int p1 = a.pairs[2 * q], p2 = a.pairs[2 * q + 1];
            double alpha1 = 0, alpha2 = 0, rho1 = 1, rho2 = 1;
#pragma unroll
            for (int p = 0; p < 4; ++p) { if (p == p1) { alpha1 = al[p]; rho1 = rh[p]; } if (p == p2) { alpha2 = al[p]; rho2 = rh[p]; } }
            const double coeffs1 = 1.0 / rho1, coeffs2 = 1.0 / rho2;
            double d12 = 0.0, d21 = 0.0;
            const int i12 = a.m12[q], i21 = a.m21[q];
            if (i12 >= 0) {
                d12 = a.dmdt[i12] ? a.dmdt[i12][c] : 0.0;
                if (a.mVar[i12] == MASS_VAR_ALPHA) d12 = lee_ksu(a, i12, alpha1, rho1, alpha1);     // from == phase1
            }
            if (i21 >= 0) {
                d21 = a.dmdt[i21] ? a.dmdt[i21][c] : 0.0;
                if (a.mVar[i21] == MASS_VAR_ALPHA) d21 = lee_ksu(a, i21, alpha2, rho2, alpha2);     // from == phase2
            }
            const double dNet = d21 - d12, c12 = coeffs1 - coeffs2;
            const bool divU = a.divU[q] != 0;
            double s1 = 0, s2 = 0, q1 = 0, q2 = 0;       // Su/Sp of phase1, phase2 (registers, selected below)
#pragma unroll
            for (int p = 0; p < 4; ++p) { if (p == p1) { s1 = Su[p]; q1 = Sp[p]; } if (p == p2) { s2 = Su[p]; q2 = Sp[p]; } }
            if (divU) { s1 += divPhi * clamp01(alpha1); s2 += divPhi * clamp01(alpha2); }
            {
                const double dm21 = dNet, c12c = c12, a1L = dmax(dmin(alpha1, 1.0), 0.0);
                s1 += coeffs1 * dm21;
                if (divU) {
                    if (dm21 > 0) {
                        if (c12c > 0) q1 -= dm21 * c12c;
                        else if (c12c < 0) s1 -= dm21 * c12c * a1L;
                    } else if (dm21 < 0) {
                        if (c12c > 0) s1 -= dm21 * c12c * a1L;
                        else if (c12c < 0) q1 -= dm21 * c12c;
                    }
                }
            }
            {
                const double dm12 = -dNet, c21c = -c12, a2L = dmax(dmin(alpha2, 1.0), 0.0);
                s2 += coeffs2 * dm12;
                if (divU) {
                    if (dm12 > 0) {
                        if (c21c > 0) q2 -= dm12 * c21c;
                        else if (c21c < 0) s2 -= dm12 * c21c * a2L;
                    } else if (dm12 < 0) {
                        if (c21c > 0) s2 -= c21c * dm12 * a2L;
                        else if (c21c < 0) q2 -= dm12 * c21c;
                    }
                }
            }
#pragma unroll
            for (int p = 0; p < 4; ++p) { if (p == p1) { Su[p] = s1; Sp[p] = q1; } if (p == p2) { Su[p] = s2; Sp[p] = q2; } }
            if (q == a.nPairs - 1) { mx21 = d21 * coeffs1; mx12 = d12 * coeffs2; }
        }
#pragma unroll
        for (int p = 0; p < 4; ++p)
            if (p < a.P) { a.Su[p][c] = Su[p]; a.Sp[p][c] = Sp[p]; }
    }
    // block maxima -> two atomics per CTA
    __shared__ double s21[TCIF_TPB / 32], s12[TCIF_TPB / 32];
    for (int o = 16; o > 0; o >>= 1) {
        mx21 = dmax(mx21, __shfl_xor_sync(0xffffffffu, mx21, o));
        mx12 = dmax(mx12, __shfl_xor_sync(0xffffffffu, mx12, o));
    }
    if ((threadIdx.x & 31) == 0) { s21[threadIdx.x >> 5] = mx21; s12[threadIdx.x >> 5] = mx12; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < TCIF_TPB / 32; ++w) { mx21 = dmax(mx21, s21[w]); mx12 = dmax(mx12, s12[w]); }
        atomicMax(a.ddtMax, dbl_ord(mx21));
        atomicMax(a.ddtMax + 1, dbl_ord(mx12));
    }
}

// addInterfacePorosity: Udiag += V * Cu*sqr(a_s)/(pow3(1 - a_s) + 1e-3), one term per model in table order
struct PorosityArgs { int n; const double *alpha[8]; double Cu[8]; };
__global__ void __launch_bounds__(TCIF_TPB)
k_interface_porosity(int nC, const double *V, const __grid_constant__ PorosityArgs a, double *Udiag) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nC) return;
    double d = Udiag[c];
    const double v = V[c];
    for (int i = 0; i < a.n; ++i) {
        double al = a.alpha[i][c], om = 1.0 - al;
        d += v * (a.Cu[i] * (al * al) / (om * om * om + 1e-3));
    }
    Udiag[c] = d;
}

}  // namespace tcif
