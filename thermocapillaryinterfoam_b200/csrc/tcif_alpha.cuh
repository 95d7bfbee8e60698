// tcif_alpha.cuh -- device kernels for the rest of multiphaseSystem::solveAlphas():
//   k_flux_assembly / k_flux_assembly_bnd   multiphaseSystem.C:181-274 + phaseSystem.C:1251-1305 (nHatfv/nHatf)
//   k_pred_setup / k_pred_jacobi (+ _bnd)    implicit upwind predictor, multiphaseSystem.C:348-363
//   k_rho_phi, k_renormalise                 multiphaseSystem.C:397-406, 414-426
// Same layout and cell-centric, atomic-free structure as tcif_device.cuh.
#pragma once
#include "tcif_device.cuh"

namespace tcif {

enum { SCHEME_UPWIND = 0, SCHEME_LINEAR = 1, SCHEME_VANLEER = 2 };

__device__ __forceinline__ double fsign(double x) { return x >= 0 ? 1.0 : -1.0; }   // Foam::sign

// weight W of the owner value, face = W*(P - N) + N: upwind | linear | vanLeer (limitedScheme<NVDTVD>)
__device__ __forceinline__ double scheme_weight(int scheme, double cdw, double flux, double phiP, double phiN,
                                                const double *gP, const double *gN, const double *d) {
    if (scheme == SCHEME_LINEAR) return cdw;
    double up = flux >= 0 ? 1.0 : 0.0;
    if (scheme == SCHEME_UPWIND) return up;
    double gradf = phiN - phiP, gradcf;
    if (flux > 0) gradcf = d[0] * gP[0] + d[1] * gP[1] + d[2] * gP[2];
    else          gradcf = d[0] * gN[0] + d[1] * gN[1] + d[2] * gN[2];
    double r;
    if (fabs(gradcf) >= 1000 * fabs(gradf)) r = 2 * 1000 * fsign(gradcf) * fsign(gradf) - 1;
    else r = 2 * (gradcf / gradf) - 1;
    double lim = (r + fabs(r)) / (1 + fabs(r));
    return lim * cdw + (1.0 - lim) * up;
}

struct FluxArgs {
    int P, schemePhi, schemePhir;
    const double *alpha[4], *alphab[4];
    const double *G, *Gb;            // [(p*3+d)*stride + c], [(p*3+d)*nBF + bf]  (Gb: neighbour-rank values on processor faces)
    const double *Cx, *Cy, *Cz;      // cell centres (vanLeer only)
    const double *dbx, *dby, *dbz;   // fvPatch::delta() on boundary faces (vanLeer only)
    const double *phi, *phib;
    double deltaN;
    double pa[12], pb[12];           // phic_k = min(pa[k]*phic0, pb[k]) for the k-th (phase1, phase2) visit
    double *out[4], *outb[4];        // phiAlphaCorr[p]
};

// all P fluxes of ONE face from the two cell states (P = owner side, N = neighbour side)
template <int P>
__device__ __forceinline__ void face_fluxes(const FluxArgs &a, double phi, double w, double sx, double sy, double sz,
                                            const double *aP, const double *aN, const double (*gP)[3], const double (*gN)[3],
                                            const double *d, double phic0, double *out) {
    // nHatf for the unordered pairs: gf(p2,p1) = -gf(p1,p2) exactly, so nHatf is antisymmetric bit for bit
    double nh[P][P];
    double af[P], gfa[P][3];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        af[p] = w * (aP[p] - aN[p]) + aN[p];
#pragma unroll
        for (int k = 0; k < 3; ++k) gfa[p][k] = w * (gP[p][k] - gN[p][k]) + gN[p][k];
    }
#pragma unroll
    for (int p1 = 0; p1 < P; ++p1)
#pragma unroll
        for (int p2 = p1 + 1; p2 < P; ++p2) {
            double gx = af[p2] * gfa[p1][0] - af[p1] * gfa[p2][0];
            double gy = af[p2] * gfa[p1][1] - af[p1] * gfa[p2][1];
            double gz = af[p2] * gfa[p1][2] - af[p1] * gfa[p2][2];
            // nHatfv = gradAlphaf/(mag(gradAlphaf) + deltaN): one reciprocal instead of three divisions (like the
            // force kernels, TCIF_FORCE_EXACT=0: a few ulp, far inside the 1e-12 contract)
            double rden = f_rcp(sqrt(gx * gx + gy * gy + gz * gz) + a.deltaN);
            double s = f_div(gx, rden) * sx;
            s += f_div(gy, rden) * sy;
            s += f_div(gz, rden) * sz;
            nh[p1][p2] = s; nh[p2][p1] = -s;
        }
    int step = 0;
#pragma unroll
    for (int p1 = 0; p1 < P; ++p1) {
        double W = scheme_weight(a.schemePhi, w, phi, aP[p1], aN[p1], gP[p1], gN[p1], d);
        double f = phi * (W * (aP[p1] - aN[p1]) + aN[p1]);
#pragma unroll
        for (int p2 = 0; p2 < P; ++p2) {
            if (p2 == p1) continue;
            double phic = dmin(a.pa[step] * phic0, a.pb[step]);
            ++step;
            double t1 = -(phic * nh[p1][p2]);
            double W2 = scheme_weight(a.schemePhir, w, t1, aP[p2], aN[p2], gP[p2], gN[p2], d);
            double inner = -(t1 * (W2 * (aP[p2] - aN[p2]) + aN[p2]));
            double W1 = scheme_weight(a.schemePhir, w, inner, aP[p1], aN[p1], gP[p1], gN[p1], d);
            f += inner * (W1 * (aP[p1] - aN[p1]) + aN[p1]);
        }
        out[p1] = f;
    }
}

// internal faces: thread c handles the faces it owns (face-local work: no boundary twin needed)
#define TCIF_TPB_FLUX 128      // ~140 live registers per thread: 3 CTAs of 128 instead of 1 CTA of 256 per SM
template <int P>
__global__ void __launch_bounds__(TCIF_TPB_FLUX, 3)
k_flux_assembly(const __grid_constant__ DMesh m, const __grid_constant__ FluxArgs a) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= m.nC) return;
    double aP[P], gP[P][3];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        aP[p] = a.alpha[p][c];
#pragma unroll
        for (int k = 0; k < 3; ++k) gP[p][k] = a.G[(size_t)(p * 3 + k) * m.stride + c];
    }
    const bool geo = a.Cx != nullptr;
    double cP[3] = {0, 0, 0};
    if (geo) { cP[0] = a.Cx[c]; cP[1] = a.Cy[c]; cP[2] = a.Cz[c]; }
    for (int k = 0; k < m.Ko; ++k) {
        const int fo = k * m.stride + c;
        const int n = m.nbrOwn[fo];
        if (n < 0) continue;
        double aN[P], gN[P][3], d[3] = {0, 0, 0}, out[P];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            aN[p] = a.alpha[p][n];
#pragma unroll
            for (int kk = 0; kk < 3; ++kk) gN[p][kk] = a.G[(size_t)(p * 3 + kk) * m.stride + n];
        }
        if (geo) { d[0] = a.Cx[n] - cP[0]; d[1] = a.Cy[n] - cP[1]; d[2] = a.Cz[n] - cP[2]; }
        const double sx = m.Sfx[fo], sy = m.Sfy[fo], sz = m.Sfz[fo], phi = a.phi[fo];
        const double phic0 = fabs(phi / sqrt(sx * sx + sy * sy + sz * sz));
        face_fluxes<P>(a, phi, m.w[fo], sx, sy, sz, aP, aN, gP, gN, d, phic0, out);
#pragma unroll
        for (int p = 0; p < P; ++p) a.out[p][fo] = out[p];
    }
}

// boundary faces: processor faces use the halo values; elsewhere phi_b*alpha_b (phic_b == 0; the inflow
// fix of :254-274 writes the same product)
template <int P>
__global__ void __launch_bounds__(TCIF_TPB)
k_flux_assembly_bnd(const __grid_constant__ DMesh m, const __grid_constant__ FluxArgs a) {
    const int bf = blockIdx.x * blockDim.x + threadIdx.x;
    if (bf >= m.nBF) return;
    const double phi = a.phib[bf];
    if (m.bfType[bf] != PATCH_PROCESSOR) {
#pragma unroll
        for (int p = 0; p < P; ++p) a.outb[p][bf] = phi * a.alphab[p][bf];
        return;
    }
    const int c = m.faceCells[bf];
    double aP[P], aN[P], gP[P][3], gN[P][3], d[3] = {0, 0, 0}, out[P];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        aP[p] = a.alpha[p][c]; aN[p] = a.alphab[p][bf];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            gP[p][k] = a.G[(size_t)(p * 3 + k) * m.stride + c];
            gN[p][k] = a.Gb[(size_t)(p * 3 + k) * m.nBF + bf];
        }
    }
    if (a.dbx) { d[0] = a.dbx[bf]; d[1] = a.dby[bf]; d[2] = a.dbz[bf]; }
    const double sx = m.Sfbx[bf], sy = m.Sfby[bf], sz = m.Sfbz[bf];
    const double phic0 = fabs(phi / m.magSfb[bf]);
    // coupled interpolation is w*P + (1-w)*N in OpenFOAM; face_fluxes uses w*(P-N)+N -- the oracle's
    // orc_interpolate does the former on coupled patches, so do it here too
    double W, af[P], gfa[P][3], nh[P][P];
    const double w = m.wb[bf];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        af[p] = w * aP[p] + (1.0 - w) * aN[p];
#pragma unroll
        for (int k = 0; k < 3; ++k) gfa[p][k] = w * gP[p][k] + (1.0 - w) * gN[p][k];
    }
#pragma unroll
    for (int p1 = 0; p1 < P; ++p1)
#pragma unroll
        for (int p2 = p1 + 1; p2 < P; ++p2) {
            double gx = af[p2] * gfa[p1][0] - af[p1] * gfa[p2][0];
            double gy = af[p2] * gfa[p1][1] - af[p1] * gfa[p2][1];
            double gz = af[p2] * gfa[p1][2] - af[p1] * gfa[p2][2];
            double den = sqrt(gx * gx + gy * gy + gz * gz) + a.deltaN;
            double s = (gx / den) * sx;
            s += (gy / den) * sy;
            s += (gz / den) * sz;
            nh[p1][p2] = s; nh[p2][p1] = -s;
        }
    int step = 0;
#pragma unroll
    for (int p1 = 0; p1 < P; ++p1) {
        W = scheme_weight(a.schemePhi, w, phi, aP[p1], aN[p1], gP[p1], gN[p1], d);
        double f = phi * (W * aP[p1] + (1.0 - W) * aN[p1]);
#pragma unroll
        for (int p2 = 0; p2 < P; ++p2) {
            if (p2 == p1) continue;
            double phic = dmin(a.pa[step] * phic0, a.pb[step]);
            ++step;
            double t1 = -(phic * nh[p1][p2]);
            double W2 = scheme_weight(a.schemePhir, w, t1, aP[p2], aN[p2], gP[p2], gN[p2], d);
            double inner = -(t1 * (W2 * aP[p2] + (1.0 - W2) * aN[p2]));
            double W1 = scheme_weight(a.schemePhir, w, inner, aP[p1], aN[p1], gP[p1], gN[p1], d);
            f += inner * (W1 * aP[p1] + (1.0 - W1) * aN[p1]);
        }
        out[p1] = f;
    }
#pragma unroll
    for (int p = 0; p < P; ++p) a.outb[p][bf] = out[p];
}

// ---- implicit upwind predictor (multiphaseSystem.C:348-363) -------------------------------------------
//   (V/dt - V Sp + sum_out phi) psi_c + sum_in phi psi_nb = V psi0/dt + V Su     (EulerDdt + upwind fvmDiv == Su + Sp)
// Jacobi sweeps, phases batched like the MULES kernels.  active[p] == 0 freezes a converged phase.
struct PredPh {
    const double *psi0, *Su, *Sp;
    const int *bc;
    const double *psib;          // boundary values (fixedValue patches: the value; processor: halo of the iterate)
    double *diag, *rhs;          // [stride]
    const double *x;             // iterate in
    double *xn;                  // iterate out
    int active;
};
struct PredBatch { int nPh; PredPh ph[4]; const double *phi, *phib; double rDeltaT; double *dmax; /* [4] */ };

__device__ __forceinline__ void atomic_max_nonneg(double *addr, double v) {   // v >= 0: bit pattern is monotone
    atomicMax((unsigned long long *)addr, (unsigned long long)__double_as_longlong(v));
}

template <bool BND>
__global__ void __launch_bounds__(TCIF_TPB)
k_pred_setup(const __grid_constant__ DMesh m, const __grid_constant__ PredBatch b) {
    const int p = blockIdx.x % b.nPh;
    const int i = (blockIdx.x / b.nPh) * blockDim.x + threadIdx.x;
    const PredPh &ph = b.ph[p];
    int c;
    if (BND) { if (i >= m.nBndCells) return; c = m.bndCells[i]; }
    else     { if (i >= m.nC) return; c = i; if (is_bnd_cell(m, c)) return; }
    const double V = m.V[c], sp = ph.Sp ? ph.Sp[c] : 0.0, su = ph.Su ? ph.Su[c] : 0.0;
    double dg = V * b.rDeltaT - V * sp, r = (V * b.rDeltaT) * ph.psi0[c] + V * su;
    for (int k = 0; k < m.Kn; ++k) {
        int fd = m.neiFace[k * m.stride + c];
        if (fd >= 0) dg += dmax(-b.phi[fd], 0.0);
    }
    for (int k = 0; k < m.Ko; ++k) {
        int fo = k * m.stride + c;
        if (m.nbrOwn[fo] >= 0) dg += dmax(b.phi[fo], 0.0);
    }
    if (BND) {
        for (int j = m.bndCellStart[i]; j < m.bndCellStart[i + 1]; ++j) {
            int bf = m.bndFaces[j];
            double phb = b.phib[bf];
            if (m.bfType[bf] == PATCH_PROCESSOR) dg += dmax(phb, 0.0);
            else if (ph.bc[m.bfPatch[bf]] == BC_FIXED_VALUE) r -= phb * ph.psib[bf];
            else dg += phb;
        }
    }
    ph.diag[c] = dg; ph.rhs[c] = r;
}

template <bool BND>
__global__ void __launch_bounds__(TCIF_TPB)
k_pred_jacobi(const __grid_constant__ DMesh m, const __grid_constant__ PredBatch b) {
    const int p = blockIdx.x % b.nPh;
    const int i = (blockIdx.x / b.nPh) * blockDim.x + threadIdx.x;
    const PredPh &ph = b.ph[p];
    if (!ph.active) return;
    int c = -1;
    double dl = 0.0;
    bool on;
    if (BND) { on = i < m.nBndCells; if (on) c = m.bndCells[i]; }
    else     { on = i < m.nC; c = i; if (on && is_bnd_cell(m, c)) on = false; }
    if (on) {
        double s = ph.rhs[c];
        for (int k = 0; k < m.Kn; ++k) {
            int fd = m.neiFace[k * m.stride + c];
            if (fd >= 0) s += dmax(b.phi[fd], 0.0) * ph.x[owner_of(m, fd)];
        }
        for (int k = 0; k < m.Ko; ++k) {
            int fo = k * m.stride + c, n = m.nbrOwn[fo];
            if (n >= 0) s -= dmin(b.phi[fo], 0.0) * ph.x[n];
        }
        if (BND) {
            for (int j = m.bndCellStart[i]; j < m.bndCellStart[i + 1]; ++j) {
                int bf = m.bndFaces[j];
                if (m.bfType[bf] == PATCH_PROCESSOR) s -= dmin(b.phib[bf], 0.0) * ph.psib[bf];
            }
        }
        double xn = s / ph.diag[c];
        ph.xn[c] = xn;
        dl = fabs(xn - ph.x[c]);
    }
    // block maximum of |dpsi| -> one atomic per warp
    for (int o = 16; o > 0; o >>= 1) dl = dmax(dl, __shfl_xor_sync(0xffffffffu, dl, o));
    if ((threadIdx.x & 31) == 0 && dl > 0.0) atomic_max_nonneg(b.dmax + p, dl);
}

// ---- rhoPhi = sum_p interpolate(rho_p) * alphaPhi_p  (:397-406); uniform rho_p when rho[p] == nullptr ----
struct RhoPhiArgs {
    int P;
    double rhoU[4];
    const double *rho[4], *rhob[4];
    const double *aPhi[4], *aPhib[4];
    double *out, *outb;
    double wgt; int accumulate;      // sub-cycling: out += wgt * rhoPhi   (multiphaseSystem.C:158)
};
static __global__ void __launch_bounds__(TCIF_TPB)
k_rho_phi(const __grid_constant__ DMesh m, const __grid_constant__ RhoPhiArgs a) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= m.nC) return;
    for (int k = 0; k < m.Ko; ++k) {
        const int fo = k * m.stride + c, n = m.nbrOwn[fo];
        if (n < 0) continue;
        const double w = m.w[fo];
        double s = 0.0;
        for (int p = 0; p < a.P; ++p) {
            double rP = a.rho[p] ? a.rho[p][c] : a.rhoU[p], rN = a.rho[p] ? a.rho[p][n] : a.rhoU[p];
            s += (w * (rP - rN) + rN) * a.aPhi[p][fo];
        }
        a.out[fo] = a.accumulate ? a.out[fo] + a.wgt * s : a.wgt * s;      // wgt == 1 without sub-cycles
    }
}
static __global__ void __launch_bounds__(TCIF_TPB)
k_rho_phi_bnd(const __grid_constant__ DMesh m, const __grid_constant__ RhoPhiArgs a) {
    const int bf = blockIdx.x * blockDim.x + threadIdx.x;
    if (bf >= m.nBF) return;
    const bool coupled = m.bfType[bf] == PATCH_PROCESSOR;
    const double w = m.wb[bf];
    double s = 0.0;
    for (int p = 0; p < a.P; ++p) {
        double rf;
        if (!a.rho[p]) rf = coupled ? w * a.rhoU[p] + (1.0 - w) * a.rhoU[p] : a.rhoU[p];
        else rf = coupled ? w * a.rho[p][m.faceCells[bf]] + (1.0 - w) * a.rhob[p][bf] : a.rhob[p][bf];
        s += rf * a.aPhib[p][bf];
    }
    a.outb[bf] = a.accumulate ? a.outb[bf] + a.wgt * s : a.wgt * s;
}

// ---- alpha += alpha*(1 - sum alpha) on cells and patch values (:414-426; no BC re-evaluation) --
struct RenormArgs { int P; double *a[4]; };
static __global__ void __launch_bounds__(TCIF_TPB)
k_renormalise(int n, const __grid_constant__ RenormArgs r) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v[4], s = 0.0;
    for (int p = 0; p < r.P; ++p) { v[p] = r.a[p][i]; s += v[p]; }
    const double corr = 1.0 - s;
    for (int p = 0; p < r.P; ++p) r.a[p][i] = v[p] + v[p] * corr;
}

// max over all face slots / boundary faces of phic0 = |phi|/|Sf| (multiphaseSystem.C:181, 237)
static __global__ void __launch_bounds__(TCIF_TPB)
k_phic_max(const __grid_constant__ DMesh m, const double *phi, const double *phib, double *out, int allPatches = 0) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double v = 0.0;
    if (i < m.nFd) {
        double sx = m.Sfx[i], sy = m.Sfy[i], sz = m.Sfz[i];
        double mg = sqrt(sx * sx + sy * sy + sz * sz);
        if (mg > 0.0) v = fabs(phi[i] / mg);
    } else if (i - m.nFd < m.nBF) {
        int bf = i - m.nFd;
        if (allPatches || m.bfType[bf] == PATCH_PROCESSOR) v = fabs(phib[bf] / m.magSfb[bf]);
    }
    for (int o = 16; o > 0; o >>= 1) v = dmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((threadIdx.x & 31) == 0 && v > 0.0) atomic_max_nonneg(out, v);
}

// ---- fvc::flux(F) = linear(F) & Sf  (pEqn.H:31-38) ------------------------------------------------------------
static __global__ void __launch_bounds__(TCIF_TPB)
k_flux_of_force(const __grid_constant__ DMesh m, const double *F, double *out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= m.nC) return;
    const double fx = F[c], fy = F[(size_t)m.stride + c], fz = F[(size_t)2 * m.stride + c];
    for (int k = 0; k < m.Ko; ++k) {
        const int fo = k * m.stride + c, n = m.nbrOwn[fo];
        if (n < 0) continue;
        const double w = m.w[fo];
        const double gx = F[n], gy = F[(size_t)m.stride + n], gz = F[(size_t)2 * m.stride + n];
        const double ix = w * (fx - gx) + gx, iy = w * (fy - gy) + gy, iz = w * (fz - gz) + gz;
        out[fo] = ix * m.Sfx[fo] + iy * m.Sfy[fo] + iz * m.Sfz[fo];
    }
}
static __global__ void __launch_bounds__(TCIF_TPB)
k_flux_of_force_bnd(const __grid_constant__ DMesh m, const double *F, const double *Fb, double *outb) {
    const int bf = blockIdx.x * blockDim.x + threadIdx.x;
    if (bf >= m.nBF) return;
    double ix = Fb[bf], iy = Fb[(size_t)m.nBF + bf], iz = Fb[(size_t)2 * m.nBF + bf];
    if (m.bfType[bf] == PATCH_PROCESSOR) {
        const int c = m.faceCells[bf];
        const double w = m.wb[bf];
        ix = w * F[c] + (1.0 - w) * ix; iy = w * F[(size_t)m.stride + c] + (1.0 - w) * iy; iz = w * F[(size_t)2 * m.stride + c] + (1.0 - w) * iz;
    }
    outb[bf] = ix * m.Sfbx[bf] + iy * m.Sfby[bf] + iz * m.Sfbz[bf];
}

}  // namespace tcif
