// tcif_device.cuh -- device-side data model and sm_100a kernels of the interface hot path.
//
// Layout (DESIGN.md "Data layout in HBM"): all fields FP64 SoA.  Internal faces live in an
// owner-slot-major padded ELL layout: device face index fd = k*stride + owner, k = rank of the
// face among its owner's faces in ascending OpenFOAM face index, stride = nCells rounded up
// to 32.  Owner-side accesses of a cell-centric kernel are therefore implicit and perfectly
// coalesced; the neighbour side needs ONE int32 per half-face (neiFace) that yields both the
// face slot and the owner cell (fd mod stride).  Per-cell accumulation order is the
// reference's (ascending face index: neighbour-side faces, then owner-side faces, then
// boundary faces in patch order), so results are deterministic without atomics and, built
// with -fmad=false, bit-identical to the serial oracle.
//
// Each cell-centric kernel exists in two instantiations: BND=false sweeps every cell over
// its internal faces only (boundary-adjacent cells get provisional values); BND=true re-does
// the boundary-adjacent cells including their patch faces and overwrites them.  The second
// launch is O(surface) and is where halo values of processor patches are consumed, so the
// O(volume) launch overlaps the exchange.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define TCIF_TPB 256
#define TCIF_SMALL 1e-15
#define TCIF_VSMALL 1e-300
#define TCIF_ROOTVSMALL 1e-150

// Force kernels (grad/tensor/div): TCIF_FORCE_EXACT=1 keeps the reference's operation sequence (true
// divisions, separate multiply/add) and is bit-identical to the serial oracle; the default (0) divides
// once per cell / pair (x * (1/V) instead of x / V) and contracts multiply-add chains with explicit
// fma(): the first ncu capture showed these two kernels instruction-issue bound (FP64 divisions are
// ~35 instructions each), and the change is <= a few ulp per operation, far inside the 1e-12 contract.
// The MULES kernels always keep the exact sequence (-fmad=false).
#ifndef TCIF_FORCE_EXACT
#define TCIF_FORCE_EXACT 0
#endif

namespace tcif {

#if TCIF_FORCE_EXACT
__device__ __forceinline__ double f_rcp(double x) { return x; }                 // "reciprocal" kept as divisor
__device__ __forceinline__ double f_div(double a, double r) { return a / r; }
__device__ __forceinline__ double f_mad(double a, double b, double c) { return a * b + c; }
#else
__device__ __forceinline__ double f_rcp(double x) { return 1.0 / x; }
__device__ __forceinline__ double f_div(double a, double r) { return a * r; }
__device__ __forceinline__ double f_mad(double a, double b, double c) { return fma(a, b, c); }
#endif

enum { PATCH_GENERIC = 0, PATCH_EMPTY = 1, PATCH_WEDGE = 2, PATCH_PROCESSOR = 3 };
enum { BC_ZERO_GRADIENT = 0, BC_FIXED_VALUE = 1 };
enum { SIGMA_CONSTANT = 0, SIGMA_POLYNOMIAL = 1, SIGMA_TABLE = 2 };

struct SigmaModel { int kind, n; double c[64]; };

struct DMesh {
    int nC, stride, Ko, Kn, nFd;
    int nBF, nBndCells, nPatches, nCoupled;
    int pfTiles;            // L2 prefetch distance of the volume sweeps in 256-cell tiles (0 = off)
    int nTiles;             // ceil(nC / 256)
    const int *tileOrder;   // [nTiles] traversal order of the 256-cell tiles (null = ascending), see build_tile_order()
    const int *nbrOwn;      // [Ko*stride] neighbour cell of the k-th owner face (-1: none)
    const int *neiFace;     // [Kn*stride] device face index of the k-th neighbour-side face (-1: none)
    const double *Sfx, *Sfy, *Sfz, *w;   // [Ko*stride]
    const double *V;        // [nC]
    const unsigned *bndMask;// [stride/32] bit c: cell c owns >=1 boundary face
    const int *bndWordRank; // [stride/32] boundary-adjacent cells below cell 32*w: rank of a mask bit = index in bndCells
    double *part;           // [partSlots*nBndCells] partial records of the boundary-adjacent cells (tcif_volume.cuh), SoA
    const int *bndCells;    // [nBndCells] ascending cell ids owning >=1 boundary face
    const int *bndCellStart;// [nBndCells+1]
    const int *bndFaces;    // boundary face ids grouped by cell, ascending
    const int4 *bndRec;     // [2*nBndCells] packed record of a boundary-adjacent cell: {n, bf0, bf1, bf2}, {tp0, tp1, tp2, 0};
                            //   tp = patch type | patch id << 4; unused slots repeat face 0; n < 0: more than 3 faces (use the lists)
    const int *faceCells;   // [nBF]
    const int *bfPatch;     // [nBF]
    const int *bfType;      // [nBF] patch type of the face
    const double *Sfbx, *Sfby, *Sfbz, *magSfb, *dcb, *wb;  // [nBF]
    const int *coupledFaces;// [nCoupled] boundary face ids of processor faces, ascending
    const int *coupledOff;  // [nCoupled] offset (in coupled-face units) of the face's processor patch
    const int *coupledSize; // [nCoupled] size of the face's processor patch
    const int *coupledPatchIdx; // [nCoupled] index of the face's processor patch among this rank's processor patches
};

// hint: pull the sector holding *p from HBM into L2 (no register destination, never faults on a valid mapping)
__device__ __forceinline__ void pf_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// own-slot load of an array that neighbouring cells gather later in the same sweep: ask L2 to keep the line
#ifndef TCIF_EVL
#define TCIF_EVL 0
#endif
__device__ __forceinline__ double ld_keep(const double *p) {
#if TCIF_EVL
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    double v;
    asm volatile("ld.global.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
    return v;
#else
    return *p;
#endif
}
__device__ __forceinline__ void pf_l2_keep(const void *p) {
#if TCIF_EVL
    asm volatile("prefetch.global.L2::evict_last [%0];" ::"l"(p));
#else
    pf_l2(p);
#endif
}

__device__ __forceinline__ bool is_bnd_cell(const DMesh &m, int c) { return (m.bndMask[c >> 5] >> (c & 31)) & 1u; }

__device__ __forceinline__ int owner_of(const DMesh &m, int fd) {
    int o = fd;
    while (o >= m.stride) o -= m.stride;
    return o;
}

__device__ __forceinline__ double dmin(double a, double b) { return a < b ? a : b; }   // Foam::min
__device__ __forceinline__ double dmax(double a, double b) { return a > b ? a : b; }   // Foam::max

// Function1<scalar> "sigma" (interfaceProperties temperatureDependent), SURVEY.md A.5
__device__ __forceinline__ double sigma_eval(const SigmaModel &s, double T) {
    if (s.kind == SIGMA_CONSTANT) return s.c[0];
    if (s.kind == SIGMA_POLYNOMIAL) {
        double y = 0.0;
        for (int i = 0; i < s.n; ++i) {
            double e = s.c[2 * i + 1], pw;
            if (e == 0.0) pw = 1.0; else if (e == 1.0) pw = T; else if (e == 2.0) pw = T * T; else pw = pow(T, e);
            y += s.c[2 * i] * pw;
        }
        return y;
    }
    int n = s.n;
    if (T <= s.c[0]) return s.c[1];
    if (T >= s.c[2 * (n - 1)]) return s.c[2 * (n - 1) + 1];
    int i = 0;
    while (i < n - 2 && T >= s.c[2 * (i + 1)]) ++i;
    double t0 = s.c[2 * i], t1 = s.c[2 * (i + 1)];
    double w0 = (t1 - T) / (t1 - t0), w1 = 1.0 - w0;
    return w0 * s.c[2 * i + 1] + w1 * s.c[2 * (i + 1) + 1];
}

// ------------------------------------------------------------------------------------
// Force: K1 grad(alpha_p) for all phases + per-pair capillary stress tensor (6 unique
// components), K2 Gauss divergence of the tensors, accumulated pair by pair.
// ------------------------------------------------------------------------------------
struct ForceArgs {
    const double *alpha[4];    // [nC]
    const double *alphab[4];   // [nBF]
    const int *bc[4];          // [nPatches]
    const double *T, *Tb;      // temperature
    const SigmaModel *sig;     // [Q]
    double deltaN;
    double *Tq;                // [(q*6+k)*stride + c]
    double *gBnd;              // [(p*3+d)*nBndCells + i]
    double *TqHalo;            // [(q*6+k)*nBF + bf]   neighbour-rank tensors on coupled faces
    double *G;                 // optional [(p*3+d)*stride + c] (tcif_grad_alpha / flux assembly)
    double *F;                 // [d*stride + c]
    int tsum;                  // 1: Tq holds ONE tensor, the sum over pairs (div is linear), 0: one per pair
};

// T = (2|g|) * (sigma * (I - n n)),  g = a2*g1 - a1*g2,  n = g/(|g| + deltaN)
// (phaseSystem.C:1074-1096, 1281-1294); symmetric -> xx,xy,xz,yy,yz,zz.
__device__ __forceinline__ void pair_tensor(double a1, double a2, const double *g1, const double *g2,
                                            double sigma, double deltaN, double *T6) {
    double gx = a2 * g1[0] - a1 * g2[0];
    double gy = a2 * g1[1] - a1 * g2[1];
    double gz = a2 * g1[2] - a1 * g2[2];
    double mg = sqrt(gx * gx + gy * gy + gz * gz);
    double rden = f_rcp(mg + deltaN);
    double nx = f_div(gx, rden), ny = f_div(gy, rden), nz = f_div(gz, rden);
    double tm = 2.0 * mg;
    T6[0] = tm * (sigma * (1.0 - nx * nx));
    T6[1] = tm * (sigma * (0.0 - nx * ny));
    T6[2] = tm * (sigma * (0.0 - nx * nz));
    T6[3] = tm * (sigma * (1.0 - ny * ny));
    T6[4] = tm * (sigma * (0.0 - ny * nz));
    T6[5] = tm * (sigma * (1.0 - nz * nz));
}

template <int P, int KO, int KN, bool BND, bool WRITE_G>
__global__ void __launch_bounds__(TCIF_TPB)
k_grad_tensor(const __grid_constant__ DMesh m, const __grid_constant__ ForceArgs a) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int c;
    if (BND) { if (i >= m.nBndCells) return; c = m.bndCells[i]; }
    else     { if (i >= m.nC) return; c = i; if (is_bnd_cell(m, c)) return; }
    double aS[P], g[P][3];
#pragma unroll
    for (int p = 0; p < P; ++p) { aS[p] = a.alpha[p][c]; g[p][0] = g[p][1] = g[p][2] = 0.0; }
    const int kn = KN ? KN : m.Kn, ko = KO ? KO : m.Ko;
#pragma unroll
    for (int k = 0; k < kn; ++k) {           // faces whose neighbour is c (ascending face index)
        int fd = m.neiFace[k * m.stride + c];
        if (fd >= 0) {
            int o = owner_of(m, fd);
            double w = m.w[fd], sx = m.Sfx[fd], sy = m.Sfy[fd], sz = m.Sfz[fd];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                double aP = a.alpha[p][o], aN = aS[p];
                double af = f_mad(w, aP - aN, aN);
                g[p][0] = f_mad(-sx, af, g[p][0]); g[p][1] = f_mad(-sy, af, g[p][1]); g[p][2] = f_mad(-sz, af, g[p][2]);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < ko; ++k) {           // faces owned by c
        int n = m.nbrOwn[k * m.stride + c];
        if (n >= 0) {
            int fd = k * m.stride + c;
            double w = m.w[fd], sx = m.Sfx[fd], sy = m.Sfy[fd], sz = m.Sfz[fd];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                double aP = aS[p], aN = a.alpha[p][n];
                double af = f_mad(w, aP - aN, aN);
                g[p][0] = f_mad(sx, af, g[p][0]); g[p][1] = f_mad(sy, af, g[p][1]); g[p][2] = f_mad(sz, af, g[p][2]);
            }
        }
    }
    if (BND) {
        for (int j = m.bndCellStart[i]; j < m.bndCellStart[i + 1]; ++j) {
            int bf = m.bndFaces[j];
            bool coupled = m.bfType[bf] == PATCH_PROCESSOR;
            double wb = m.wb[bf], sx = m.Sfbx[bf], sy = m.Sfby[bf], sz = m.Sfbz[bf];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                double ab = a.alphab[p][bf];
                double af = coupled ? wb * aS[p] + (1.0 - wb) * ab : ab;
                g[p][0] += sx * af; g[p][1] += sy * af; g[p][2] += sz * af;
            }
        }
    }
    double rV = f_rcp(m.V[c]);
#pragma unroll
    for (int p = 0; p < P; ++p) { g[p][0] = f_div(g[p][0], rV); g[p][1] = f_div(g[p][1], rV); g[p][2] = f_div(g[p][2], rV); }
    if (BND) {
#pragma unroll
        for (int p = 0; p < P; ++p)
#pragma unroll
            for (int d = 0; d < 3; ++d) a.gBnd[(size_t)(p * 3 + d) * m.nBndCells + i] = g[p][d];
    }
    if (WRITE_G) {
#pragma unroll
        for (int p = 0; p < P; ++p)
#pragma unroll
            for (int d = 0; d < 3; ++d) a.G[(size_t)(p * 3 + d) * m.stride + c] = g[p][d];
    }
    if (a.Tq) {
        double Tc = a.T ? a.T[c] : 0.0;
        double Ts6[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        int q = 0;
#pragma unroll
        for (int p1 = 0; p1 < P; ++p1)
#pragma unroll
            for (int p2 = p1 + 1; p2 < P; ++p2, ++q) {
                double T6[6];
                double sg = sigma_eval(a.sig[q], Tc);
                pair_tensor(aS[p1], aS[p2], g[p1], g[p2], sg, a.deltaN, T6);
                if (a.tsum) {
#pragma unroll
                    for (int k = 0; k < 6; ++k) Ts6[k] += T6[k];
                } else {
#pragma unroll
                    for (int k = 0; k < 6; ++k) a.Tq[(size_t)(q * 6 + k) * m.stride + c] = T6[k];
                }
            }
        if (a.tsum) {
#pragma unroll
            for (int k = 0; k < 6; ++k) a.Tq[(size_t)k * m.stride + c] = Ts6[k];
        }
    }
}

// ssf = S & (w*(TP - TN) + TN) with symmetric T6; sign = +1 (c is owner) / -1 (c is neighbour)
__device__ __forceinline__ void face_div_acc(double w, double sx, double sy, double sz,
                                             const double *TP, const double *TN, double sgn, double *acc) {
    double t[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) t[k] = f_mad(w, TP[k] - TN[k], TN[k]);
    double fx = f_mad(sz, t[2], f_mad(sy, t[1], sx * t[0]));
    double fy = f_mad(sz, t[4], f_mad(sy, t[3], sx * t[1]));
    double fz = f_mad(sz, t[5], f_mad(sy, t[4], sx * t[2]));
    if (sgn > 0) { acc[0] += fx; acc[1] += fy; acc[2] += fz; }
    else         { acc[0] -= fx; acc[1] -= fy; acc[2] -= fz; }
}

template <int P, int KO, int KN, bool BND>
__global__ void __launch_bounds__(TCIF_TPB)
k_div_tensor(const __grid_constant__ DMesh m, const __grid_constant__ ForceArgs a) {
    constexpr int Q = P * (P - 1) / 2;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int c;
    if (BND) { if (i >= m.nBndCells) return; c = m.bndCells[i]; }
    else     { if (i >= m.nC) return; c = i; if (is_bnd_cell(m, c)) return; }
    const int kn = KN ? KN : m.Kn, ko = KO ? KO : m.Ko;
    double rV = f_rcp(m.V[c]);
    double F[3] = {0.0, 0.0, 0.0};
    const int nQ = a.tsum ? 1 : Q;
#pragma unroll 1
    for (int q = 0; q < nQ; ++q) {
        const double *Tq = a.Tq + (size_t)q * 6 * m.stride;
        double Ts[6], acc[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int k = 0; k < 6; ++k) Ts[k] = Tq[(size_t)k * m.stride + c];
#pragma unroll
        for (int k = 0; k < kn; ++k) {
            int fd = m.neiFace[k * m.stride + c];
            if (fd >= 0) {
                int o = owner_of(m, fd);
                double To[6];
#pragma unroll
                for (int kk = 0; kk < 6; ++kk) To[kk] = Tq[(size_t)kk * m.stride + o];
                face_div_acc(m.w[fd], m.Sfx[fd], m.Sfy[fd], m.Sfz[fd], To, Ts, -1.0, acc);
            }
        }
#pragma unroll
        for (int k = 0; k < ko; ++k) {
            int n = m.nbrOwn[k * m.stride + c];
            if (n >= 0) {
                int fd = k * m.stride + c;
                double Tn[6];
#pragma unroll
                for (int kk = 0; kk < 6; ++kk) Tn[kk] = Tq[(size_t)kk * m.stride + n];
                face_div_acc(m.w[fd], m.Sfx[fd], m.Sfy[fd], m.Sfz[fd], Ts, Tn, 1.0, acc);
            }
        }
        if (BND) {
            for (int j = m.bndCellStart[i]; j < m.bndCellStart[i + 1]; ++j) {
                int bf = m.bndFaces[j];
                double sx = m.Sfbx[bf], sy = m.Sfby[bf], sz = m.Sfbz[bf];
                double t[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
                if (m.bfType[bf] == PATCH_PROCESSOR) {
                    double wb = m.wb[bf];
#pragma unroll
                    for (int kk = 0; kk < 6; ++kk)
                        t[kk] = wb * Ts[kk] + (1.0 - wb) * a.TqHalo[(size_t)(q * 6 + kk) * m.nBF + bf];
                } else {
                    // patch values: alpha_b, grad_b = g_c + n (snGrad - n.g_c), sigma(T_b); pair q only, or
                    // (tsum) all pairs summed in pair order
                    double mag = m.magSfb[bf];
                    double nx = sx / mag, ny = sy / mag, nz = sz / mag;
                    int qq = 0;
                    for (int p1 = 0; p1 < P; ++p1)
                        for (int p2 = p1 + 1; p2 < P; ++p2, ++qq) {
                            if (!a.tsum && qq != q) continue;
                            double gb[2][3], ab[2];
                            int pp[2] = {p1, p2};
#pragma unroll
                            for (int s = 0; s < 2; ++s) {
                                int p = pp[s];
                                double gc0 = a.gBnd[(size_t)(p * 3 + 0) * m.nBndCells + i];
                                double gc1 = a.gBnd[(size_t)(p * 3 + 1) * m.nBndCells + i];
                                double gc2 = a.gBnd[(size_t)(p * 3 + 2) * m.nBndCells + i];
                                ab[s] = a.alphab[p][bf];
                                double sn = 0.0;
                                if (a.bc[p][m.bfPatch[bf]] == BC_FIXED_VALUE) sn = m.dcb[bf] * (ab[s] - a.alpha[p][c]);
                                double ng = nx * gc0 + ny * gc1 + nz * gc2;
                                double sc = sn - ng;
                                gb[s][0] = gc0 + nx * sc; gb[s][1] = gc1 + ny * sc; gb[s][2] = gc2 + nz * sc;
                            }
                            double sg = sigma_eval(a.sig[qq], a.Tb ? a.Tb[bf] : 0.0);
                            double tq[6];
                            pair_tensor(ab[0], ab[1], gb[0], gb[1], sg, a.deltaN, tq);
                            if (a.tsum) {
#pragma unroll
                                for (int kk = 0; kk < 6; ++kk) t[kk] += tq[kk];
                            } else {
#pragma unroll
                                for (int kk = 0; kk < 6; ++kk) t[kk] = tq[kk];
                            }
                        }
                }
                acc[0] += sx * t[0] + sy * t[1] + sz * t[2];
                acc[1] += sx * t[1] + sy * t[3] + sz * t[4];
                acc[2] += sx * t[2] + sy * t[4] + sz * t[5];
            }
        }
        F[0] += f_div(acc[0], rV); F[1] += f_div(acc[1], rV); F[2] += f_div(acc[2], rV);
    }
    a.F[c] = F[0]; a.F[(size_t)m.stride + c] = F[1]; a.F[(size_t)2 * m.stride + c] = F[2];
}

// ------------------------------------------------------------------------------------
// MULES fused into cell-centric sweeps:
//   k_mules_setup : upwind flux phiBD, phiCorr, local extrema, sums, Q+/Q-, and the cell pass
//                   of limiter iteration 1 (lambda == 1 there)
//   k_mules_iter  : face pass of iteration j fused with the cell pass of iteration j+1
//   k_mules_final : face pass of the last iteration + apply (returnCorr) or
//                   apply + surfaceIntegrate + explicit Euler update (explicitSolve)
// ------------------------------------------------------------------------------------
struct MulesPh {
    const double *psi, *psi0, *psib;    // alpha, alpha.oldTime(), patch values (halo on processor)
    double *psi0w;                      // = psi0, written by the set-up sweep when MulesParams::storeOld
    const double *Su, *Sp;              // may be null (= 0)
    const int *bc;                      // [nPatches]
    const double *phiIn, *phiInb;       // HO flux read by the set-up sweep (may alias phiPsi)
    double *phiPsi, *phiPsib;           // out: limited flux / correction
    double *phiCorr, *phiCorrb, *phiBDb;
    double *Qp, *Qm, *sP, *sM;
    double *lpA, *lmA, *lpB, *lmB;      // lambda+/- cell values, ping-pong
    double *lamA, *lamB, *lamb;         // face lambda ping-pong (internal), boundary
    double *lpHalo, *lmHalo;            // [nBF] neighbour-rank lambda+/- on coupled faces
    double *psiNew;                     // explicit update target
    const double *lpIn, *lmIn, *lamIn;  // ping-pong resolved on the host for the current sweep (kvp_* kernels)
    double *lpOut, *lmOut, *lamOut;
};
struct MulesBatch { int nPh; MulesPh ph[4]; };
struct MulesParams {
    double rDeltaT, psiMax, psiMin, extremaCoeff, smoothLimiter;
    int boundaryExtrema, addUpwind;
    int storeOld;                       // set-up sweep: psi.oldTime() := psi (fixed-K kernels; the generic path copies on the host side)
    const double *phi, *phib;
};

// upwind<scalar>::flux: phi * (w*(psiP - psiN) + psiN), w = pos0(phi)
__device__ __forceinline__ double upwind_flux(double ph, double psiP, double psiN) {
    double w = ph >= 0 ? 1.0 : 0.0;
    return ph * (w * (psiP - psiN) + psiN);
}

template <int KO, int KN, bool BND>
__global__ void __launch_bounds__(TCIF_TPB)
k_mules_setup(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b,
              const __grid_constant__ MulesParams prm) {
    const int p = blockIdx.x % b.nPh;
    const int i = (blockIdx.x / b.nPh) * blockDim.x + threadIdx.x;
    const MulesPh &ph = b.ph[p];
    int c;
    if (BND) { if (i >= m.nBndCells) return; c = m.bndCells[i]; }
    else     { if (i >= m.nC) return; c = i; if (is_bnd_cell(m, c)) return; }
    const int kn = KN ? KN : m.Kn, ko = KO ? KO : m.Ko;
    const double psiC = ph.psi[c];
    double psiMaxn = prm.psiMin, psiMinn = prm.psiMax, sBD = 0.0, sP = 0.0, sM = 0.0;
#pragma unroll
    for (int k = 0; k < kn; ++k) {
        int fd = m.neiFace[k * m.stride + c];
        if (fd >= 0) {
            int o = owner_of(m, fd);
            double psiO = ph.psi[o];
            double bd = upwind_flux(prm.phi[fd], psiO, psiC);
            double pp = ph.phiIn[fd];
            if (prm.addUpwind) pp = pp + bd;
            double pc = pp - bd;
            psiMaxn = dmax(psiMaxn, psiO); psiMinn = dmin(psiMinn, psiO);
            sBD -= bd;
            if (pc > 0) sM += pc; else sP -= pc;
        }
    }
#pragma unroll
    for (int k = 0; k < ko; ++k) {
        int n = m.nbrOwn[k * m.stride + c];
        if (n >= 0) {
            int fd = k * m.stride + c;
            double psiN = ph.psi[n];
            double bd = upwind_flux(prm.phi[fd], psiC, psiN);
            double pp = ph.phiIn[fd];
            if (prm.addUpwind) pp = pp + bd;
            double pc = pp - bd;
            ph.phiCorr[fd] = pc;
            psiMaxn = dmax(psiMaxn, psiN); psiMinn = dmin(psiMinn, psiN);
            sBD += bd;
            if (pc > 0) sP += pc; else sM -= pc;
        }
    }
    if (BND) {
        for (int j = m.bndCellStart[i]; j < m.bndCellStart[i + 1]; ++j) {
            int bf = m.bndFaces[j];
            bool coupled = m.bfType[bf] == PATCH_PROCESSOR;
            double psb = ph.psib[bf], phb = prm.phib[bf];
            double pp = ph.phiInb[bf], bd;
            if (coupled) {
                double w = phb >= 0 ? 1.0 : 0.0;
                double up = phb * (w * psiC + (1.0 - w) * psb);
                if (prm.addUpwind) pp = pp + up;
                bd = up;
            } else {
                if (prm.addUpwind) pp = pp + phb * psb;
                bd = pp;                       // phiBD_b := phiPsi_b on non-coupled patches
            }
            double pc = pp - bd;
            ph.phiBDb[bf] = bd; ph.phiCorrb[bf] = pc; ph.lamb[bf] = 1.0;
            if (coupled || ph.bc[m.bfPatch[bf]] == BC_FIXED_VALUE || prm.boundaryExtrema == 1) {
                psiMaxn = dmax(psiMaxn, psb); psiMinn = dmin(psiMinn, psb);
            } else {
                psiMaxn = dmax(psiMaxn, prm.psiMax); psiMinn = dmin(psiMinn, prm.psiMin);
            }
            sBD += bd;
            if (pc > 0) sP += pc; else sM -= pc;
        }
    }
    psiMaxn = dmin(psiMaxn + prm.extremaCoeff * (prm.psiMax - prm.psiMin), prm.psiMax);
    psiMinn = dmax(psiMinn - prm.extremaCoeff * (prm.psiMax - prm.psiMin), prm.psiMin);
    if (prm.smoothLimiter > TCIF_SMALL) {
        double s = prm.smoothLimiter;
        psiMaxn = dmin(s * psiC + (1.0 - s) * psiMaxn, prm.psiMax);
        psiMinn = dmax(s * psiC + (1.0 - s) * psiMinn, prm.psiMin);
    }
    const double V = m.V[c], psi0 = ph.psi0[c];
    const double Su = ph.Su ? ph.Su[c] : 0.0, Sp = ph.Sp ? ph.Sp[c] : 0.0;
    double Qp = V * ((prm.rDeltaT - Sp) * psiMaxn - Su) - (V * prm.rDeltaT) * psi0 + sBD;
    double Qm = V * (Su - (prm.rDeltaT - Sp) * psiMinn) + (V * prm.rDeltaT) * psi0 - sBD;
    ph.Qp[c] = Qp; ph.Qm[c] = Qm; ph.sP[c] = sP; ph.sM[c] = sM;
    // limiter iteration 1, cell pass: lambda == 1 so sumlPhip == sumPhip, mSumlPhim == mSumPhim
    ph.lmA[c] = dmax(dmin((sP + Qp) / (sM + TCIF_ROOTVSMALL), 1.0), 0.0);
    ph.lpA[c] = dmax(dmin((sM + Qm) / (sP + TCIF_ROOTVSMALL), 1.0), 0.0);
}

// face rule of the limiter (internal face, owner o / neighbour n)
__device__ __forceinline__ double lambda_face(double lam, double pc, double lpO, double lmO, double lpN, double lmN) {
    // = pc > 0 ? min(lam, min(lpO, lmN)) : min(lam, min(lmO, lpN)), selected first (same values, fewer instructions)
    const bool pos = pc > 0;
    return dmin(lam, dmin(pos ? lpO : lmO, pos ? lmN : lpN));
}

// boundary-face rule (wedge / coupled two-sided / outflow-only), SURVEY.md A.6
__device__ __forceinline__ double lambda_bface(int type, double lam, double pc, double bd, double lpC, double lmC,
                                               double lpH, double lmH) {
    if (type == PATCH_WEDGE) return 0.0;
    if (type == PATCH_PROCESSOR) {
        // own side: min(lam, pc>0 ? lp[c] : lm[c]); neighbour side sees -pc; syncFaceList(min)
        double own = pc > 0 ? lpC : lmC;
        double oth = (-pc) > 0 ? lpH : lmH;
        return dmin(dmin(lam, own), oth);
    }
    if ((bd + pc) > TCIF_SMALL * TCIF_SMALL) return dmin(lam, pc > 0 ? lpC : lmC);
    return lam;
}

// first != 0: lambda_in == 1 everywhere (iteration 1), lamIn is not read.
template <int KO, int KN, bool BND>
__global__ void __launch_bounds__(TCIF_TPB)
k_mules_iter(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b, int first, int flip) {
    const int p = blockIdx.x % b.nPh;
    const int i = (blockIdx.x / b.nPh) * blockDim.x + threadIdx.x;
    const MulesPh &ph = b.ph[p];
    int c;
    if (BND) { if (i >= m.nBndCells) return; c = m.bndCells[i]; }
    else     { if (i >= m.nC) return; c = i; if (is_bnd_cell(m, c)) return; }
    const double *lpIn = flip ? ph.lpB : ph.lpA, *lmIn = flip ? ph.lmB : ph.lmA;
    double *lpOut = flip ? ph.lpA : ph.lpB, *lmOut = flip ? ph.lmA : ph.lmB;
    const double *lamIn = flip ? ph.lamB : ph.lamA;
    double *lamOut = flip ? ph.lamA : ph.lamB;
    const int kn = KN ? KN : m.Kn, ko = KO ? KO : m.Ko;
    const double lpC = lpIn[c], lmC = lmIn[c];
    double slp = 0.0, slm = 0.0;
#pragma unroll
    for (int k = 0; k < kn; ++k) {
        int fd = m.neiFace[k * m.stride + c];
        if (fd >= 0) {
            int o = owner_of(m, fd);
            double pc = ph.phiCorr[fd];
            double lam = first ? 1.0 : lamIn[fd];
            lam = lambda_face(lam, pc, lpIn[o], lmIn[o], lpC, lmC);
            double l = lam * pc;
            if (l > 0) slm += l; else slp -= l;
        }
    }
#pragma unroll
    for (int k = 0; k < ko; ++k) {
        int n = m.nbrOwn[k * m.stride + c];
        if (n >= 0) {
            int fd = k * m.stride + c;
            double pc = ph.phiCorr[fd];
            double lam = first ? 1.0 : lamIn[fd];
            lam = lambda_face(lam, pc, lpC, lmC, lpIn[n], lmIn[n]);
            lamOut[fd] = lam;
            double l = lam * pc;
            if (l > 0) slp += l; else slm -= l;
        }
    }
    if (BND) {
        for (int j = m.bndCellStart[i]; j < m.bndCellStart[i + 1]; ++j) {
            int bf = m.bndFaces[j];
            double pc = ph.phiCorrb[bf];
            double lam = lambda_bface(m.bfType[bf], ph.lamb[bf], pc, ph.phiBDb[bf], lpC, lmC,
                                      ph.lpHalo[bf], ph.lmHalo[bf]);
            ph.lamb[bf] = lam;
            double l = lam * pc;
            if (l > 0) slp += l; else slm -= l;
        }
    }
    lmOut[c] = dmax(dmin((slp + ph.Qp[c]) / (ph.sM[c] + TCIF_ROOTVSMALL), 1.0), 0.0);
    lpOut[c] = dmax(dmin((slm + ph.Qm[c]) / (ph.sP[c] + TCIF_ROOTVSMALL), 1.0), 0.0);
}

// MODE 0: returnCorr  -> phiPsi = lambda*phiCorr           (owner-side faces only)
// MODE 1: !returnCorr -> phiPsi = phiBD + lambda*phiCorr   (owner-side faces only)
// MODE 2: explicitSolve: MODE 1 + surfaceIntegrate + Euler update into psiNew
template <int KO, int KN, bool BND, int MODE>
__global__ void __launch_bounds__(TCIF_TPB)
k_mules_final(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b,
              const __grid_constant__ MulesParams prm, int first, int flip, int nIter) {
    const int p = blockIdx.x % b.nPh;
    const int i = (blockIdx.x / b.nPh) * blockDim.x + threadIdx.x;
    const MulesPh &ph = b.ph[p];
    int c;
    if (BND) { if (i >= m.nBndCells) return; c = m.bndCells[i]; }
    else     { if (i >= m.nC) return; c = i; if (is_bnd_cell(m, c)) return; }
    const double *lpIn = flip ? ph.lpB : ph.lpA, *lmIn = flip ? ph.lmB : ph.lmA;
    const double *lamIn = flip ? ph.lamB : ph.lamA;
    const int kn = KN ? KN : m.Kn, ko = KO ? KO : m.Ko;
    const double lpC = lpIn[c], lmC = lmIn[c];
    const double psiC = (MODE >= 1) ? ph.psi[c] : 0.0;
    double sum = 0.0;
    if (MODE == 2) {
#pragma unroll
        for (int k = 0; k < kn; ++k) {
            int fd = m.neiFace[k * m.stride + c];
            if (fd >= 0) {
                int o = owner_of(m, fd);
                double pc = ph.phiCorr[fd];
                double lam = first ? 1.0 : lamIn[fd];
                if (nIter > 0) lam = lambda_face(lam, pc, lpIn[o], lmIn[o], lpC, lmC);
                double bd = upwind_flux(prm.phi[fd], ph.psi[o], psiC);
                sum -= bd + lam * pc;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < ko; ++k) {
        int n = m.nbrOwn[k * m.stride + c];
        if (n >= 0) {
            int fd = k * m.stride + c;
            double pc = ph.phiCorr[fd];
            double lam = first ? 1.0 : lamIn[fd];
            if (nIter > 0) lam = lambda_face(lam, pc, lpC, lmC, lpIn[n], lmIn[n]);
            double fl;
            if (MODE == 0) fl = pc * lam;
            else           fl = upwind_flux(prm.phi[fd], psiC, ph.psi[n]) + lam * pc;
            ph.phiPsi[fd] = fl;
            sum += fl;
        }
    }
    if (BND) {
        for (int j = m.bndCellStart[i]; j < m.bndCellStart[i + 1]; ++j) {
            int bf = m.bndFaces[j];
            double pc = ph.phiCorrb[bf], bd = ph.phiBDb[bf];
            double lam = ph.lamb[bf];
            if (nIter > 0) lam = lambda_bface(m.bfType[bf], lam, pc, bd, lpC, lmC, ph.lpHalo[bf], ph.lmHalo[bf]);
            ph.lamb[bf] = lam;
            double fl = (MODE == 0) ? pc * lam : bd + lam * pc;
            ph.phiPsib[bf] = fl;
            sum += fl;
        }
    }
    if (MODE == 2) {
        const double V = m.V[c];
        const double Su = ph.Su ? ph.Su[c] : 0.0, Sp = ph.Sp ? ph.Sp[c] : 0.0;
        double d = sum / V;
        ph.psiNew[c] = (ph.psi0[c] * prm.rDeltaT + Su - d) / (prm.rDeltaT - Sp);
    }
}

// MULES::limitSum on the internal faces (device face slots; padded slots hold 0) .
struct LimitSumArgs { int nPh; double *f[4]; };
static __global__ void __launch_bounds__(TCIF_TPB)
k_limit_sum(int n, const __grid_constant__ LimitSumArgs a) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n) return;
    double v[4], sumPos = 0.0, sumNeg = 0.0;
    for (int p = 0; p < a.nPh; ++p) {
        v[p] = a.f[p][f];
        if (v[p] > 0) sumPos += v[p]; else sumNeg += v[p];
    }
    double sum = sumPos + sumNeg;
    if (sum > 0 && sumPos > TCIF_VSMALL) {
        double lambda = -sumNeg / sumPos;
        for (int p = 0; p < a.nPh; ++p) if (v[p] > 0) a.f[p][f] = v[p] * lambda;
    } else if (sum < 0 && sumNeg < -TCIF_VSMALL) {
        double lambda = -sumPos / sumNeg;
        for (int p = 0; p < a.nPh; ++p) if (v[p] < 0) a.f[p][f] = v[p] * lambda;
    }
}
// same on the coupled boundary faces
static __global__ void k_limit_sum_coupled(int n, const int *coupledFaces, const __grid_constant__ LimitSumArgs a) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int f = coupledFaces[k];
    double v[4], sumPos = 0.0, sumNeg = 0.0;
    for (int p = 0; p < a.nPh; ++p) {
        v[p] = a.f[p][f];
        if (v[p] > 0) sumPos += v[p]; else sumNeg += v[p];
    }
    double sum = sumPos + sumNeg;
    if (sum > 0 && sumPos > TCIF_VSMALL) {
        double lambda = -sumNeg / sumPos;
        for (int p = 0; p < a.nPh; ++p) if (v[p] > 0) a.f[p][f] = v[p] * lambda;
    } else if (sum < 0 && sumNeg < -TCIF_VSMALL) {
        double lambda = -sumPos / sumNeg;
        for (int p = 0; p < a.nPh; ++p) if (v[p] < 0) a.f[p][f] = v[p] * lambda;
    }
}

// phiAlpha += upwind flux (stand-alone form of the fused addUpwind path)
template <int KO>
__global__ void __launch_bounds__(TCIF_TPB)
k_add_upwind(const __grid_constant__ DMesh m, const double *phi, const double *psi, double *phiPsi) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= m.nC) return;
    const int ko = KO ? KO : m.Ko;
    double psiC = psi[c];
#pragma unroll
    for (int k = 0; k < ko; ++k) {
        int n = m.nbrOwn[k * m.stride + c];
        if (n >= 0) {
            int fd = k * m.stride + c;
            phiPsi[fd] = phiPsi[fd] + upwind_flux(phi[fd], psiC, psi[n]);
        }
    }
}
static __global__ void k_add_upwind_bnd(const __grid_constant__ DMesh m, const double *phib, const double *psi,
                                 const double *psib, double *phiPsib) {
    int bf = blockIdx.x * blockDim.x + threadIdx.x;
    if (bf >= m.nBF) return;
    double phb = phib[bf], v;
    if (m.bfType[bf] == PATCH_PROCESSOR) {
        double w = phb >= 0 ? 1.0 : 0.0;
        v = phb * (w * psi[m.faceCells[bf]] + (1.0 - w) * psib[bf]);
    } else v = phb * psib[bf];
    phiPsib[bf] = phiPsib[bf] + v;
}

// ------------------------------------------------------------------------------------
// boundary evaluation, halo pack/unpack, layout conversion, reductions
// ------------------------------------------------------------------------------------
// vol-field correctBoundaryConditions on non-coupled patches: zeroGradient -> adjacent cell; all phases of a batch in one launch.
struct BcArgs { int n; const double *vi[4]; double *vb[4]; const int *bc[4]; };
static __global__ void k_correct_bc_scalar(const __grid_constant__ DMesh m, const __grid_constant__ BcArgs a) {
    int bf = blockIdx.x * blockDim.x + threadIdx.x;
    if (bf >= m.nBF) return;
    if (m.bfType[bf] == PATCH_PROCESSOR) return;
    const int fc = m.faceCells[bf], pa = m.bfPatch[bf];
    for (int k = 0; k < a.n; ++k)
        if (a.bc[k][pa] == BC_ZERO_GRADIENT) a.vb[k][bf] = a.vi[k][fc];
}
// extrapolatedCalculated patches of a result field (ncomp SoA comps, cell stride `cs`, bnd stride nBF)
static __global__ void k_extrapolate_bnd(const __grid_constant__ DMesh m, int ncomp, size_t cs, const double *vi, double *vb) {
    int bf = blockIdx.x * blockDim.x + threadIdx.x;
    if (bf >= m.nBF) return;
    if (m.bfType[bf] == PATCH_PROCESSOR) return;
    int fc = m.faceCells[bf];
    for (int d = 0; d < ncomp; ++d) vb[(size_t)d * m.nBF + bf] = vi[d * cs + fc];
}
// gaussGrad::correctBoundaryConditions on non-coupled patches of one phase gradient
static __global__ void k_grad_bnd_correct(const __grid_constant__ DMesh m, const double *alpha, const double *alphab,
                                   const int *bc, double *gb) {
    int bf = blockIdx.x * blockDim.x + threadIdx.x;
    if (bf >= m.nBF) return;
    if (m.bfType[bf] == PATCH_PROCESSOR) return;
    double mag = m.magSfb[bf];
    double nx = m.Sfbx[bf] / mag, ny = m.Sfby[bf] / mag, nz = m.Sfbz[bf] / mag;
    double sn = 0.0;
    if (bc[m.bfPatch[bf]] == BC_FIXED_VALUE) sn = m.dcb[bf] * (alphab[bf] - alpha[m.faceCells[bf]]);
    double g0 = gb[bf], g1 = gb[(size_t)m.nBF + bf], g2 = gb[(size_t)2 * m.nBF + bf];
    double ng = nx * g0 + ny * g1 + nz * g2;
    double s = sn - ng;
    gb[bf] = g0 + nx * s; gb[(size_t)m.nBF + bf] = g1 + ny * s; gb[(size_t)2 * m.nBF + bf] = g2 + nz * s;
}

struct HaloArgs { int ncomp; const double *src[36]; double *dst[36]; };
// Packed buffers are patch-major so that one processor patch is ONE contiguous message:
// buf[off_p*ncomp + d*size_p + i], i = rank of the face inside its patch (same order on both ranks).
__device__ __forceinline__ size_t halo_slot(const DMesh &m, int k, int d, int ncomp) {
    int off = m.coupledOff[k], sz = m.coupledSize[k];
    return (size_t)off * ncomp + (size_t)d * sz + (k - off);
}
// patchInternalField of cell arrays
static __global__ void k_halo_pack(const __grid_constant__ DMesh m, const __grid_constant__ HaloArgs h, double *send) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m.nCoupled) return;
    int fc = m.faceCells[m.coupledFaces[k]];
    for (int d = 0; d < h.ncomp; ++d) send[halo_slot(m, k, d, h.ncomp)] = h.src[d][fc];
}
static __global__ void k_halo_unpack(const __grid_constant__ DMesh m, const __grid_constant__ HaloArgs h, const double *recv) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m.nCoupled) return;
    int bf = m.coupledFaces[k];
    for (int d = 0; d < h.ncomp; ++d) h.dst[d][bf] = recv[halo_slot(m, k, d, h.ncomp)];
}

// ---- processor-patch exchange over peer memory (NVLink): no pack buffer, no NCCL kernel ----------------------------
// Every rank owns a receive buffer (two parities) and one flag word per source rank; neighbours map it with
// cudaIpcOpenMemHandle.  k_halo_push stores patchInternalField straight into the neighbour's buffer, in the layout
// k_halo_unpack reads (patch-major, component-major inside a patch); the last CTA to finish releases the flags
// (value = exchange number) at system scope.  k_halo_wait (one warp) acquires this rank's flags, k_halo_pull then
// reads the buffer with cache-volatile loads.  Parity double-buffering is safe because a rank pushes exchange n+1 only
// after it has pulled exchange n, i.e. after every neighbour has pushed n -- which each did after pulling n-1.
#define TCIF_MAX_PEER_PATCHES 64
struct HaloPeer {
    int nPatches;
    double *base[TCIF_MAX_PEER_PATCHES];              // neighbour's receive buffer (mapped), start of the parity-0 half
    int peerOff[TCIF_MAX_PEER_PATCHES];               // offset (coupled-face units) of the matching patch in the NEIGHBOUR's layout
    unsigned *remoteFlag[TCIF_MAX_PEER_PATCHES];      // neighbour's flag word for this rank
    const unsigned *localFlag[TCIF_MAX_PEER_PATCHES]; // this rank's flag word for that neighbour
    size_t stride[TCIF_MAX_PEER_PATCHES];             // doubles per parity half of the NEIGHBOUR's buffer
};
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
static __global__ void __launch_bounds__(TCIF_TPB)
k_halo_push(const __grid_constant__ DMesh m, const __grid_constant__ HaloArgs h, const __grid_constant__ HaloPeer hp,
            unsigned parity, unsigned seq, int *counter) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < m.nCoupled) {
        const int pi = m.coupledPatchIdx[k], off = m.coupledOff[k], sz = m.coupledSize[k];
        const int fc = m.faceCells[m.coupledFaces[k]];
        double *dst = hp.base[pi] + parity * hp.stride[pi] + (size_t)hp.peerOff[pi] * h.ncomp + (k - off);
        for (int d = 0; d < h.ncomp; ++d) dst[(size_t)d * sz] = h.src[d][fc];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int done = atomicAdd(counter, 1);
        if (done == (int)gridDim.x - 1) {            // every CTA's stores are ordered before this point (fence + atomic)
            *counter = 0;
            __threadfence_system();
            for (int p = 0; p < hp.nPatches; ++p) st_release_sys(hp.remoteFlag[p], seq);
        }
    }
}
// one warp: wait until every neighbour has released exchange `seq`; gives up after ~2 s and raises *status
static __global__ void k_halo_wait(const __grid_constant__ HaloPeer hp, unsigned seq, int *status) {
    const long long t0 = clock64();
    for (int p = threadIdx.x; p < hp.nPatches; p += blockDim.x) {
        while ((int)(ld_acquire_sys(hp.localFlag[p]) - seq) < 0) {
            if (clock64() - t0 > 4000000000LL) { atomicExch(status, 1); break; }
            __nanosleep(200);
        }
    }
}
static __global__ void __launch_bounds__(TCIF_TPB)
k_halo_pull(const __grid_constant__ DMesh m, const __grid_constant__ HaloArgs h, const double *recv) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m.nCoupled) return;
    const int bf = m.coupledFaces[k];
    for (int d = 0; d < h.ncomp; ++d) h.dst[d][bf] = __ldcv(recv + halo_slot(m, k, d, h.ncomp));
}

// host order <-> device slots
static __global__ void k_face_scatter(int nIF, const int *faceMap, const double *src, double *dst) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < nIF) dst[faceMap[f]] = src[f];
}
static __global__ void k_face_gather(int nIF, const int *faceMap, const double *src, double *dst) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < nIF) dst[f] = src[faceMap[f]];
}
static __global__ void k_face_scatter_vec(int nIF, const int *faceMap, const double *src, double *x, double *y, double *z) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < nIF) { int fd = faceMap[f]; x[fd] = src[3 * (size_t)f]; y[fd] = src[3 * (size_t)f + 1]; z[fd] = src[3 * (size_t)f + 2]; }
}
static __global__ void k_soa_to_aos3(int n, size_t stride, const double *src, double *dst) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { dst[3 * (size_t)i] = src[i]; dst[3 * (size_t)i + 1] = src[stride + i]; dst[3 * (size_t)i + 2] = src[2 * stride + i]; }
}
static __global__ void k_aos3_to_soa(int n, size_t stride, const double *src, double *dst) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { dst[i] = src[3 * (size_t)i]; dst[stride + i] = src[3 * (size_t)i + 1]; dst[2 * stride + i] = src[3 * (size_t)i + 2]; }
}

// ---- delta read-back (tcif_host_sparse_returns): only the entries that differ from what the host already holds ---
// `shadow` is the device's record of the host array.  Entries whose VALUE differs are appended as (index, value) with
// one atomic per warp; the order of the pairs is arbitrary, the host array after the scatter is not.  count may
// exceed cap (the caller then falls back to the dense copy).
static __global__ void __launch_bounds__(TCIF_TPB)
k_delta_compact(int n, const double *cur, const double *shadow, int cap, int *idx, double *val, int *counter) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double v = 0.0;
    bool ch = false;
    if (i < n) { v = cur[i]; ch = v != shadow[i]; }          // by VALUE: -0.0 == +0.0 is not a change (NaN always is)
    const unsigned mask = __ballot_sync(0xffffffffu, ch);
    if (!mask) return;
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == __ffs(mask) - 1) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(0xffffffffu, base, __ffs(mask) - 1);
    const int pos = base + __popc(mask & ((1u << lane) - 1u));
    if (ch && pos < cap) { idx[pos] = i; val[pos] = v; }
}
// 3-component SoA field (stride between components) -> (index, x, y, z)
static __global__ void __launch_bounds__(TCIF_TPB)
k_delta_compact3(int n, size_t stride, const double *cur, const double *shadow, int cap, int *idx, double *val, int *counter) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double x = 0.0, y = 0.0, z = 0.0;
    bool ch = false;
    if (i < n) {
        x = cur[i]; y = cur[stride + i]; z = cur[2 * stride + i];
        ch = x != shadow[i] || y != shadow[stride + i] || z != shadow[2 * stride + i];
    }
    const unsigned mask = __ballot_sync(0xffffffffu, ch);
    if (!mask) return;
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == __ffs(mask) - 1) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(0xffffffffu, base, __ffs(mask) - 1);
    const int pos = base + __popc(mask & ((1u << lane) - 1u));
    if (ch && pos < cap) { idx[pos] = i; val[3 * (size_t)pos] = x; val[3 * (size_t)pos + 1] = y; val[3 * (size_t)pos + 2] = z; }
}
// shadow[idx[k]] := the values just sent (ncomp components, SoA with `stride`)
static __global__ void __launch_bounds__(TCIF_TPB)
k_delta_apply(int n, int ncomp, size_t stride, const int *idx, const double *val, double *shadow) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int i = idx[k];
    for (int d = 0; d < ncomp; ++d) shadow[d * stride + i] = val[(size_t)ncomp * k + d];
}

// sum(V*a), min(a), max(a): deterministic two-level tree
static __global__ void __launch_bounds__(TCIF_TPB)
k_stats_partial(int n, const double *V, const double *a, double *part) {
    __shared__ double s0[TCIF_TPB], s1[TCIF_TPB], s2[TCIF_TPB];
    double sum = 0.0, mn = 1e300, mx = -1e300;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double v = a[i];
        sum += V[i] * v; mn = dmin(mn, v); mx = dmax(mx, v);
    }
    s0[threadIdx.x] = sum; s1[threadIdx.x] = mn; s2[threadIdx.x] = mx;
    __syncthreads();
    for (int s = blockDim.x / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            s0[threadIdx.x] += s0[threadIdx.x + s];
            s1[threadIdx.x] = dmin(s1[threadIdx.x], s1[threadIdx.x + s]);
            s2[threadIdx.x] = dmax(s2[threadIdx.x], s2[threadIdx.x + s]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) { part[blockIdx.x * 3] = s0[0]; part[blockIdx.x * 3 + 1] = s1[0]; part[blockIdx.x * 3 + 2] = s2[0]; }
}
static __global__ void k_stats_final(int nPart, const double *part, double *out3) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double sum = 0.0, mn = 1e300, mx = -1e300;
        for (int i = 0; i < nPart; ++i) { sum += part[i * 3]; mn = dmin(mn, part[i * 3 + 1]); mx = dmax(mx, part[i * 3 + 2]); }
        out3[0] = sum; out3[1] = mn; out3[2] = mx;
    }
}

}  // namespace tcif
