// tcif_laser.cuh -- laserDTRM on the device: the discrete-ray laser heat source.
//   laserDTRM::calculate                 laserDTRM/laserDTRM.C:542-735   (reflecting cells :596-652, Q/V :671)
//   laserDTRM::nHatfv                    laserDTRM/laserDTRM.C:115-135   (deltaN = 1e-7/cbrt(average(V)))
//   DTRMParticle::move                   laserDTRM/DTRMParticle/DTRMParticle.C:94-266
//   Fresnel / FresnelLaser / noReflection  laserDTRM/reflectionModel/*/*.C
//   localDensityAbsorptionEmission       laserDTRM/localDensityAbsorptionEmission/localDensityAbsorptionEmission.C:86-186
// One thread walks one ray cell to cell over the planes of the cell's faces (what particle::trackToAndHitFace
// does in convex cells with planar faces), Beer-Lambert absorption per cell crossed, Fresnel reflection in the
// marked interface cells.  Reflected rays are appended to the cloud and walked by the next launch.  The energy
// goes to Q with fp64 atomics (order of the additions is the only difference to the serial reference).
// The cell's faces come from the ELL tables of DMesh: owner slots, neighbour-side slots, boundary faces - the
// same order as the CPU oracle, so that ties at edges resolve alike.
#pragma once
#include "tcif_device.cuh"

namespace tcif {

#define TCIF_SMALL 1e-15
#define TCIF_ROOTSMALL 3e-8
#define TCIF_VGREAT 1e300

enum { REFL_NONE = 0, REFL_FRESNEL = 1, REFL_FRESNEL_LASER = 2 };
enum { LASER_ERR_PROCESSOR = 2, LASER_ERR_OVERFLOW = 3, LASER_ERR_STUCK = 4 };

struct LaserRefl { int kind, from, to; double nk1[2], nk2[2], epsilon; };

struct LaserPart {
    double pos[3], p0[3], p1[3], I0, I, dA, frac;
    int cell, tid, face, keep;          // face: last face hit (device id: slot fd, nFd + bf; -1 none)
};

struct LaserGeo {
    const int *bndIdx;                  // [nC] index into bndCells / bndCellStart, -1 for interior cells
    const double *Cfx, *Cfy, *Cfz;      // [nFd] face centres in the slot layout of Sf
    const double *Cfbx, *Cfby, *Cfbz;   // [nBF]
    const double *Cx, *Cy, *Cz;         // [stride] cell centres
    const int *fdToFace;                // [nFd] slot -> mesh face (-1: padding)
    int nIF, nPoints;
    const double *points;               // [nPoints*3]
    const int *fpStart, *fpPts, *tetBasePt;   // mesh.faces() as CSR (internal, then boundary faces), tetBasePtIs
    const int *pcStart, *pc;            // point -> cells (ascending) and their normalised inverse-distance weights
    const double *pw;
    const unsigned char *isPatchPoint;
    double boundsMag;
    // walk table (W > 0): per cell W face records in the walk's order - outward area vector and face centre
    // [(c*W + j)*6 doubles], device face id and the cell behind it [(c*W + j)*2 ints, -1: padding / boundary]:
    // one contiguous read per cell crossed instead of gathers through the ELL tables
    int W;
    const double *wd;
    const int *wi;
};

struct LaserArgs {
    int nModels, P, nAbs, cap;
    LaserRefl models[8];
    double alphaCut, sigmaSB, deltaN;
    const double *alpha[4];
    const double *G;                    // grad(alpha_p): [(p*3+d)*stride + c]
    int *refl;                          // [nC] reflection model id of the cell, -1: none
    double *nHat;                       // [3*stride] SoA
    double *nHatPts;                    // [nPoints*3]
    int absPhase[4];
    double aCoeff[4], eCoeff[4], ECoeff[4];
    const double *aF, *eF, *EF;         // a/e/E fields of another absorptionEmission model (null: local density)
    const double *T;
    double *Q;
    LaserPart *cloud;
    int *counters;                      // [0] particles in the cloud, [1] error code, [4 + i] outbox fill of coupled patch i
    // processor patches: a particle reaching one is handed to the neighbour rank (hitProcessorPatch, DTRMParticle.C:268-275)
    LaserPart *outbox;                  // [nCoupledPatches * mailCap], null on a single domain
    int mailCap;
    const int *patchInfo;               // [nPatches*2]: coupled-patch slot (-1: none), first boundary face of the patch
    int localFindCell;                  // 1: reference behaviour, a reflected particle inserted behind a processor patch is dropped
    int maxSteps;                       // bound on the cells one particle may cross in one launch (LASER_ERR_STUCK beyond)
};

__device__ __forceinline__ double l_dot(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
__device__ __forceinline__ double l_mag(const double *v) { return sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

// ---- reflecting cells and their normal (laserDTRM.C:596-652) --------------------------------------------------
__global__ void __launch_bounds__(TCIF_TPB)
k_laser_mark(const __grid_constant__ DMesh m, const __grid_constant__ LaserArgs a) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= m.nC) return;
    int refl = -1;
    double nh[3] = {0.0, 0.0, 0.0};
    for (int id = 0; id < a.nModels; ++id) {
        const int from = a.models[id].from, to = a.models[id].to;
        const double aF = a.alpha[from][c], aT = a.alpha[to][c];
        double gF[3], gT[3], gA[3], nP[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            gF[d] = a.G[(size_t)(from * 3 + d) * m.stride + c];
            gT[d] = a.G[(size_t)(to * 3 + d) * m.stride + c];
            gA[d] = aT * gF[d] - aF * gT[d];
        }
        const double den = l_mag(gA) + a.deltaN;
#pragma unroll
        for (int d = 0; d < 3; ++d) nP[d] = gA[d] / den;
        const double gradAlphaf = l_dot(gF, gT);
        const double nearI = (aT - a.alphaCut >= 0) ? 1.0 : 0.0;
        const double mask = nearI * gradAlphaf;
        if (nearI != 0.0 && l_mag(nP) > 0.99 && mask < 0) {
            refl = id;
            if (l_mag(nh) == 0.0) { nh[0] += nP[0]; nh[1] += nP[1]; nh[2] += nP[2]; }
        }
    }
    a.refl[c] = refl;
#pragma unroll
    for (int d = 0; d < 3; ++d) a.nHat[(size_t)d * m.stride + c] = nh[d];
}

// ---- volPointInterpolation of nHat (patch values are zero) ------------------------------------------------------
__global__ void __launch_bounds__(TCIF_TPB)
k_laser_points(const __grid_constant__ DMesh m, const __grid_constant__ LaserGeo g, const double *nHat, double *out) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= g.nPoints) return;
    double s[3] = {0.0, 0.0, 0.0};
    if (!g.isPatchPoint[p])
        for (int i = g.pcStart[p]; i < g.pcStart[p + 1]; ++i) {
            const double w = g.pw[i];
            const int c = g.pc[i];
#pragma unroll
            for (int d = 0; d < 3; ++d) s[d] += w * nHat[(size_t)d * m.stride + c];
        }
    out[(size_t)p * 3] = s[0]; out[(size_t)p * 3 + 1] = s[1]; out[(size_t)p * 3 + 2] = s[2];
}

// ---- face access by device id F: slot fd < nFd, boundary face nFd + bf -----------------------------------------
__device__ __forceinline__ void l_face_geom(const DMesh &m, const LaserGeo &g, int F, double S[3], double Cf[3]) {
    if (F < m.nFd) {
        S[0] = m.Sfx[F]; S[1] = m.Sfy[F]; S[2] = m.Sfz[F];
        Cf[0] = g.Cfx[F]; Cf[1] = g.Cfy[F]; Cf[2] = g.Cfz[F];
    } else {
        const int b = F - m.nFd;
        S[0] = m.Sfbx[b]; S[1] = m.Sfby[b]; S[2] = m.Sfbz[b];
        Cf[0] = g.Cfbx[b]; Cf[1] = g.Cfby[b]; Cf[2] = g.Cfbz[b];
    }
}
__device__ __forceinline__ int l_mesh_face(const DMesh &m, const LaserGeo &g, int F) {
    return F < m.nFd ? g.fdToFace[F] : g.nIF + (F - m.nFd);
}

// visits the faces of cell c in the oracle's order; fn(F, sign, otherCell, Sout, Cf) returns true to stop.
// Sout = sign*Sf points out of the cell (the sign is exact, so dot(x, Sout) == sign*dot(x, Sf) bit for bit).
template <class Fn> __device__ __forceinline__ void l_for_faces_ell(const DMesh &m, const LaserGeo &g, int c, Fn &&fn) {
    double S[3], Cf[3];
    for (int k = 0; k < m.Ko; ++k) {
        const int fd = k * m.stride + c, n = m.nbrOwn[fd];
        if (n < 0) continue;
        l_face_geom(m, g, fd, S, Cf);
        if (fn(fd, 1.0, n, S, Cf)) return;
    }
    for (int k = 0; k < m.Kn; ++k) {
        const int fd = m.neiFace[k * m.stride + c];
        if (fd < 0) continue;
        l_face_geom(m, g, fd, S, Cf);
        S[0] = -S[0]; S[1] = -S[1]; S[2] = -S[2];
        if (fn(fd, -1.0, owner_of(m, fd), S, Cf)) return;
    }
    const int bi = g.bndIdx[c];
    if (bi >= 0)
        for (int j = m.bndCellStart[bi]; j < m.bndCellStart[bi + 1]; ++j) {
            const int F = m.nFd + m.bndFaces[j];
            l_face_geom(m, g, F, S, Cf);
            if (fn(F, 1.0, -1, S, Cf)) return;
        }
}

template <int WW, class Fn> __device__ __forceinline__ void l_for_faces_table(const LaserGeo &g, int c, int W, Fn &&fn) {
    const double2 *rd = reinterpret_cast<const double2 *>(g.wd + (size_t)c * W * 6);
    const int2 *ri = reinterpret_cast<const int2 *>(g.wi + (size_t)c * W * 2);
    if (WW > 0) {
        double2 r[WW > 0 ? WW : 1][3];
        int2 id[WW > 0 ? WW : 1];
#pragma unroll
        for (int j = 0; j < WW; ++j) { id[j] = ri[j]; r[j][0] = rd[3 * j]; r[j][1] = rd[3 * j + 1]; r[j][2] = rd[3 * j + 2]; }
#pragma unroll
        for (int j = 0; j < WW; ++j) {
            if (id[j].x < 0) return;
            const double S[3] = {r[j][0].x, r[j][0].y, r[j][1].x}, Cf[3] = {r[j][1].y, r[j][2].x, r[j][2].y};
            const bool nei = (id[j].y & 0x40000000) != 0 && id[j].y >= 0;
            if (fn(id[j].x, nei ? -1.0 : 1.0, nei ? (id[j].y & 0x3fffffff) : id[j].y, S, Cf)) return;
        }
    } else {
        for (int j = 0; j < W; ++j) {
            const int2 idj = ri[j];
            if (idj.x < 0) return;
            const double2 a = rd[3 * j], b = rd[3 * j + 1], cc = rd[3 * j + 2];
            const double S[3] = {a.x, a.y, b.x}, Cf[3] = {b.y, cc.x, cc.y};
            const bool nei = (idj.y & 0x40000000) != 0 && idj.y >= 0;
            if (fn(idj.x, nei ? -1.0 : 1.0, nei ? (idj.y & 0x3fffffff) : idj.y, S, Cf)) return;
        }
    }
}

template <class Fn> __device__ __forceinline__ void l_for_faces(const DMesh &m, const LaserGeo &g, int c, Fn &&fn) {
    if (g.W == 6) l_for_faces_table<6>(g, c, 6, fn);
    else if (g.W > 0) l_for_faces_table<0>(g, c, g.W, fn);
    else l_for_faces_ell(m, g, c, fn);
}

// fills the walk table of cell c (slot j: face j of the ELL enumeration; bit 30 of `other` marks a face seen from
// its neighbour side, which interpolation in the cell's tets needs for the orientation of the face triangle)
__global__ void __launch_bounds__(TCIF_TPB)
k_laser_build_walk(const __grid_constant__ DMesh m, const __grid_constant__ LaserGeo g, int W, double *wd, int *wi) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= m.nC) return;
    int j = 0;
    l_for_faces_ell(m, g, c, [&](int F, double sg, int oc, const double *S, const double *Cf) {
        if (j < W) {
            double *r = wd + ((size_t)c * W + j) * 6;
            r[0] = S[0]; r[1] = S[1]; r[2] = S[2]; r[3] = Cf[0]; r[4] = Cf[1]; r[5] = Cf[2];
            wi[((size_t)c * W + j) * 2] = F;
            wi[((size_t)c * W + j) * 2 + 1] = (sg < 0 && oc >= 0) ? (oc | 0x40000000) : oc;
        }
        ++j;
        return false;
    });
    for (; j < W; ++j) {
        double *r = wd + ((size_t)c * W + j) * 6;
        for (int k = 0; k < 6; ++k) r[k] = 0.0;
        wi[((size_t)c * W + j) * 2] = -1; wi[((size_t)c * W + j) * 2 + 1] = -1;
    }
}

// particle::trackToAndHitFace as a walk over the face planes; returns an error code or 0
__device__ int l_track(const DMesh &m, const LaserGeo &g, LaserPart &p, const double disp[3], double f) {
    if (p.keep != 1) return 0;
    double lam = TCIF_VGREAT;
    int hit = -1, other = -1;
    l_for_faces(m, g, p.cell, [&](int F, double, int oc, const double *S, const double *Cf) {
        const double dn = l_dot(disp, S);
        if (dn > 0) {
            const double d[3] = {Cf[0] - p.pos[0], Cf[1] - p.pos[1], Cf[2] - p.pos[2]};
            double l = l_dot(d, S) / dn;
            if (l < 0) l = 0;
            if (l < lam) { lam = l; hit = F; other = oc; }
        }
        return false;
    });
    if (hit < 0 || lam >= 1.0) {
        for (int d = 0; d < 3; ++d) p.pos[d] += disp[d];
        p.frac += f;
        if (p.frac > 1.0 - 1e-14) p.frac = 1.0;
        p.face = -1;
        return 0;
    }
    for (int d = 0; d < 3; ++d) p.pos[d] += lam * disp[d];
    p.frac += lam * f;
    p.face = hit;
    if (hit < m.nFd) p.cell = other;
    else {
        // hitProcessorPatch: switchProcessor (DTRMParticle.C:268-275); hitWallPatch (:278-285) and the other
        // non-coupled patches remove the particle
        p.keep = m.bfType[hit - m.nFd] == PATCH_PROCESSOR ? 2 : 0;
    }
    return 0;
}

// mesh.findCell for the reflected particle: walk from cell c through the first violated face plane.
// -1: outside the mesh; *procFace = boundary face when the walk stopped at a processor patch
__device__ int l_walk_find(const DMesh &m, const LaserGeo &g, int c, const double pt[3], int *procFace) {
    *procFace = -1;
    for (int step = 0; step < 1000; ++step) {
        int next = -2, via = -1;
        l_for_faces(m, g, c, [&](int F, double, int oc, const double *S, const double *Cf) {
            const double d[3] = {pt[0] - Cf[0], pt[1] - Cf[1], pt[2] - Cf[2]};
            if (l_dot(d, S) > 0) { next = oc; via = F; return true; }
            return false;
        });
        if (next == -2) return c;
        if (next < 0) {
            if (m.bfType[via - m.nFd] == PATCH_PROCESSOR) *procFace = via - m.nFd;
            return -1;
        }
        c = next;
    }
    return -1;
}

// ---- cellPointWeight / interpolationCellPoint -----------------------------------------------------------------
// triangle::pointToBarycentric (Ericson, Real-time collision detection, p.47-48)
__device__ double l_tri_bary(const double *a, const double *b, const double *c, const double *pt, double w[3]) {
    double v0[3], v1[3], v2[3];
    for (int d = 0; d < 3; ++d) { v0[d] = b[d] - a[d]; v1[d] = c[d] - a[d]; v2[d] = pt[d] - a[d]; }
    const double d00 = l_dot(v0, v0), d01 = l_dot(v0, v1), d11 = l_dot(v1, v1), d20 = l_dot(v2, v0), d21 = l_dot(v2, v1);
    const double denom = d00 * d11 - d01 * d01;
    if (fabs(denom) < TCIF_SMALL) { w[0] = w[1] = w[2] = 1.0 / 3.0; return denom; }
    w[1] = (d11 * d20 - d01 * d21) / denom;
    w[2] = (d00 * d21 - d01 * d20) / denom;
    w[0] = 1.0 - w[1] - w[2];
    return denom;
}

// vertex labels of tet `tetPt` (1..n-2) of mesh face f seen from its owner (tetIndices::faceTriIs)
__device__ __forceinline__ void l_face_tri(const LaserGeo &g, int f, int tetPt, int v[3]) {
    const int s = g.fpStart[f], n = g.fpStart[f + 1] - s;
    int base = g.tetBasePt[f];
    if (base < 0) base = 0;
    const int ia = (base + tetPt) % n, ib = (ia + 1) % n;
    v[0] = g.fpPts[s + base]; v[1] = g.fpPts[s + ia]; v[2] = g.fpPts[s + ib];
}

// cellPointWeight::findTriangle + interpolate on mesh face f with area vector S
__device__ void l_interp_on_face(const LaserGeo &g, int f, const double *S, const double *pos, const double *pv, double out[3]) {
    const double tol = TCIF_SMALL;
    const int n = g.fpStart[f + 1] - g.fpStart[f];
    const double faceAreaSqr = l_dot(S, S);
    double minUVClose = TCIF_VGREAT, w[3];
    int close = 1, v[3];
    for (int tp = 1; tp < n - 1; ++tp) {
        l_face_tri(g, f, tp, v);
        const double det = l_tri_bary(g.points + (size_t)v[0] * 3, g.points + (size_t)v[1] * 3, g.points + (size_t)v[2] * 3, pos, w);
        if (0.25 * fabs(det) / faceAreaSqr > tol) {
            const double u = w[0], vv = w[1];
            if ((u + tol > 0) && (vv + tol > 0) && (u + vv < 1 + tol)) { close = -tp; break; }
            double minU = fabs(u), minV = fabs(vv);
            if (minU > 1.0) minU -= 1.0;
            if (minV > 1.0) minV -= 1.0;
            const double minUV = fabs(minU + minV);
            if (minUV < minUVClose) { minUVClose = minUV; close = tp; }
        }
    }
    if (close > 0) {
        l_face_tri(g, f, close, v);
        l_tri_bary(g.points + (size_t)v[0] * 3, g.points + (size_t)v[1] * 3, g.points + (size_t)v[2] * 3, pos, w);
    }
    for (int d = 0; d < 3; ++d) {
        double t = pv[(size_t)v[0] * 3 + d] * w[0];
        t += pv[(size_t)v[1] * 3 + d] * w[1];
        t += pv[(size_t)v[2] * 3 + d] * w[2];
        out[d] = t;
    }
}

// tetrahedron::pointToBarycentric: a = cell centre, (b, c, d) the face triangle
__device__ double l_tet_bary(const double *a, const double *b, const double *c, const double *d, const double *pt, double w[4]) {
    double e1[3], e2[3], e3[3], r[3];
    for (int k = 0; k < 3; ++k) { e1[k] = b[k] - a[k]; e2[k] = c[k] - a[k]; e3[k] = d[k] - a[k]; r[k] = pt[k] - a[k]; }
    const double c23[3] = {e2[1] * e3[2] - e2[2] * e3[1], e2[2] * e3[0] - e2[0] * e3[2], e2[0] * e3[1] - e2[1] * e3[0]};
    const double c31[3] = {e3[1] * e1[2] - e3[2] * e1[1], e3[2] * e1[0] - e3[0] * e1[2], e3[0] * e1[1] - e3[1] * e1[0]};
    const double c12[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
    const double det = l_dot(e1, c23);
    if (fabs(det) < TCIF_SMALL * TCIF_SMALL) { w[0] = w[1] = w[2] = w[3] = 0.25; return det; }
    w[1] = l_dot(r, c23) / det; w[2] = l_dot(r, c31) / det; w[3] = l_dot(r, c12) / det;
    w[0] = 1.0 - w[1] - w[2] - w[3];
    return det;
}

// cellPointWeight::findTetrahedron + interpolate inside cell c (the track ended inside the cell)
__device__ void l_interp_in_cell(const DMesh &m, const LaserGeo &g, int c, const double *pos, const double *nHat,
                                 const double *pv, double out[3]) {
    const double tol = TCIF_SMALL, V = m.V[c];
    const double C[3] = {g.Cx[c], g.Cy[c], g.Cz[c]};
    double best = TCIF_VGREAT, bw[4] = {1, 0, 0, 0};
    int bv[3] = {0, 0, 0};
    l_for_faces(m, g, c, [&](int F, double sg, int, const double *, const double *) {
        const int f = l_mesh_face(m, g, F), n = g.fpStart[f + 1] - g.fpStart[f];
        for (int tp = 1; tp < n - 1; ++tp) {
            int v[3];
            double w[4];
            l_face_tri(g, f, tp, v);
            if (sg < 0) { int s = v[1]; v[1] = v[2]; v[2] = s; }
            const double det = l_tet_bary(C, g.points + (size_t)v[0] * 3, g.points + (size_t)v[1] * 3, g.points + (size_t)v[2] * 3, pos, w);
            if (fabs(det / V) > tol) {
                const double u = w[0], vv = w[1], ww = w[2];
                if ((u + tol > 0) && (vv + tol > 0) && (ww + tol > 0) && (u + vv + ww < 1 + tol)) {
                    for (int k = 0; k < 4; ++k) bw[k] = w[k];
                    for (int k = 0; k < 3; ++k) bv[k] = v[k];
                    return true;
                }
                double mu = fabs(u), mv = fabs(vv), mw = fabs(ww);
                if (mu > 1.0) mu -= 1.0;
                if (mv > 1.0) mv -= 1.0;
                if (mw > 1.0) mw -= 1.0;
                const double muvw = fabs(mu + mv + mw);
                if (muvw < best) {
                    best = muvw;
                    for (int k = 0; k < 4; ++k) bw[k] = w[k];
                    for (int k = 0; k < 3; ++k) bv[k] = v[k];
                }
            }
        }
        return false;
    });
    for (int d = 0; d < 3; ++d) {
        double s = nHat[(size_t)d * m.stride + c] * bw[0];
        s += pv[(size_t)bv[0] * 3 + d] * bw[1];
        s += pv[(size_t)bv[1] * 3 + d] * bw[2];
        s += pv[(size_t)bv[2] * 3 + d] * bw[3];
        out[d] = s;
    }
}

// ---- reflection models ------------------------------------------------------------------------------------------
// Fresnel::rho (Fresnel.C:66-117), called with cosTheta as DTRMParticle::move does (DTRMParticle.C:170-183)
__device__ double l_fresnel_rho(const LaserRefl &r, double incidentAngle) {
    const double n1 = r.nk1[0] * r.nk1[0];
    const double n2 = r.nk2[0] * r.nk2[0], k2 = r.nk2[1] * r.nk2[1];
    const double sinTheta1 = sin(incidentAngle);
    const double b = n2 - k2 - n1 * (sinTheta1 * sinTheta1);
    const double sqrP = 0.5 * (sqrt(b * b + 4 * n2 * k2) + b);
    const double sqrQ = 0.5 * (sqrt(b * b + 4 * n2 * k2) - b);
    const double cosTheta1 = cos(incidentAngle), tanTheta1 = tan(incidentAngle);
    const double t1 = sqrt(n1) * cosTheta1 - sqrt(sqrP), t2 = sqrt(n1) * cosTheta1 + sqrt(sqrP);
    const double rhoP = (t1 * t1 + sqrQ) / (t2 * t2 + sqrQ);
    const double t3 = sqrt(sqrP) - sqrt(n1) * sinTheta1 * tanTheta1, t4 = sqrt(sqrP) + sqrt(n1) * sinTheta1 * tanTheta1;
    const double rhoN = ((t3 * t3 + sqrQ) / (t4 * t4 + sqrQ)) * rhoP;
    return 0.5 * (rhoP + rhoN);
}
// FresnelLaser::rho (FresnelLaser.C:66-83)
__device__ double l_fresnel_laser_rho(const LaserRefl &r, double cosTheta) {
    const double eps = r.epsilon;
    const double a = 1 - eps * cosTheta, b = 1 + eps * cosTheta;
    return 0.5 * ((1 + a * a) / (1 + b * b)
                  + (eps * eps - 2 * eps * cosTheta + 2 * (cosTheta * cosTheta)) / (eps * eps + 2 * eps * cosTheta + 2 * (cosTheta * cosTheta)));
}

// ---- DTRMParticle::move (DTRMParticle.C:94-266): particles [first, last) of the cloud ---------------------------
__global__ void __launch_bounds__(128)
k_laser_move(const __grid_constant__ DMesh m, const __grid_constant__ LaserGeo g, const __grid_constant__ LaserArgs a,
             int first, int last) {
    const int ip = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (ip >= last) return;
    const double pi = 3.14159265358979323846;
    LaserPart p = a.cloud[ip];
    if (p.keep != 1) return;            // a received particle whose insertion point was not found on this rank
    int err = 0;
    int steps = 0;
    do {
        // l_track clips a negative crossing parameter to 0: on warped cells the plane walk could cycle without advancing
        // p.frac; a hung kernel needs a device reset, so the walk is bounded and reports instead
        if (++steps > a.maxSteps) { err = LASER_ERR_STUCK; break; }
        const double pos0[3] = {p.pos[0], p.pos[1], p.pos[2]};
        const int cell1 = p.cell;
        const double f = 1 - p.frac;
        double disp[3];
        for (int d = 0; d < 3; ++d) disp[d] = f * (p.p1[d] - p.p0[d]);                 // :110-111
        // cell data first: these loads overlap the face records of the track
        const int zone = a.refl[cell1];
        double ac, ec, Ec;                                                            // interpolationCell of a, e, E, T
        if (a.aF) { ac = a.aF[cell1]; ec = a.eF[cell1]; Ec = a.EF[cell1]; }
        else {
            ac = 0.0; ec = 0.0; Ec = 0.0;
            for (int i = 0; i < a.nAbs; ++i) {
                const double al = dmax(a.alpha[a.absPhase[i]][cell1], 0.0);
                ac += al * a.aCoeff[i]; ec += al * a.eCoeff[i]; Ec += al * a.ECoeff[i];
            }
        }
        const double Tc = a.T[cell1];
        const double emit = ec * a.sigmaSB * ((Tc * Tc) * (Tc * Tc)) / pi + Ec;
        if ((err = l_track(m, g, p, disp, f))) break;
        const double p1[3] = {p.pos[0], p.pos[1], p.pos[2]};
        double dsv[3] = {p1[0] - pos0[0], p1[1] - pos0[1], p1[2] - pos0[2]};
        double ds = l_mag(dsv);
        if (ds == 0) {                                                                // :127-132
            if ((err = l_track(m, g, p, disp, f))) break;
            for (int d = 0; d < 3; ++d) dsv[d] = p1[d] - p.pos[d];
            ds = l_mag(dsv);
        }
        if (zone > -1 && (p.tid == -1 || p.tid != zone)) {                            // :139-146
            double rho = 0;
            if (p.I > 0.01 * p.I0 && ds > 0) {                                        // :151
                const double pDir[3] = {dsv[0] / ds, dsv[1] / ds, dsv[2] / ds};
                double nh[3];
                if (p.face >= 0) {
                    double S[3], Cf[3];
                    l_face_geom(m, g, p.face, S, Cf);
                    l_interp_on_face(g, l_mesh_face(m, g, p.face), S, p.pos, a.nHatPts, nh);
                } else l_interp_in_cell(m, g, cell1, p.pos, a.nHat, a.nHatPts, nh);
                const double mg = l_mag(nh) + TCIF_ROOTSMALL;
                for (int d = 0; d < 3; ++d) nh[d] /= mg;
                const double cosTheta = -l_dot(pDir, nh);
                if (cosTheta > TCIF_SMALL) {                                          // :162
                    const LaserRefl &r = a.models[zone];
                    const double rr = r.kind == REFL_FRESNEL ? l_fresnel_rho(r, cosTheta)
                                    : r.kind == REFL_FRESNEL_LASER ? l_fresnel_laser_rho(r, cosTheta) : 0.0;
                    rho = dmin(dmax(rr, 0.0), 0.98);                                  // :171-183
                    const double delaM = cbrt(m.V[cell1]);
                    double insertP[3];
                    for (int d = 0; d < 3; ++d) insertP[d] = p.pos[d] - pDir[d] * 0.1 * delaM;
                    int procFace;
                    const int cellI = l_walk_find(m, g, cell1, insertP, &procFace);
                    // noReflection: R() = 0 and rho = 0, the new particle would deposit exactly nothing
                    if ((cellI > -1 || (procFace >= 0 && a.outbox && !a.localFindCell)) && r.kind != REFL_NONE) {
                        const double s2 = 2.0 * (-l_dot(pDir, nh));                   // R(i, n) = i + 2(-i & n) n
                        LaserPart q;
                        for (int d = 0; d < 3; ++d) {
                            const double nd = pDir[d] + s2 * nh[d];
                            q.pos[d] = q.p0[d] = insertP[d];
                            q.p1[d] = insertP[d] + nd * g.boundsMag;                  // :196-198
                        }
                        q.I0 = q.I = p.I * rho; q.dA = p.dA; q.frac = 0; q.cell = cellI; q.tid = -1; q.face = -1; q.keep = 1;
                        if (cellI > -1) {
                            const int slot = atomicAdd(a.counters, 1);
                            if (slot >= a.cap) { err = LASER_ERR_OVERFLOW; break; }
                            a.cloud[slot] = q;
                        } else {
                            // the insertion point lies behind a processor patch: the neighbour rank locates it
                            // (the reference's local findCell would drop the particle on a decomposed mesh only)
                            const int pa = m.bfPatch[procFace], slot = a.patchInfo[2 * pa];
                            const int k = atomicAdd(a.counters + 4 + slot, 1);
                            if (k >= a.mailCap) { err = LASER_ERR_OVERFLOW; break; }
                            q.keep = 3;
                            q.face = procFace - a.patchInfo[2 * pa + 1];
                            a.outbox[(size_t)slot * a.mailCap + k] = q;
                        }
                    }
                }
            }
            p.tid = zone;                                                             // :212
            const double Itran = p.I * (1.0 - rho);
            const double I1 = (Itran + ds * emit) / (1 + ds * ac);                    // :221-226
            atomicAdd(a.Q + cell1, (Itran - dmax(I1, 0.0)) * p.dA);
            p.I = I1;
            if (p.I <= 0.01 * p.I0) { p.frac = 1.0; break; }
        } else {
            const double I1 = (p.I + ds * emit) / (1 + ds * ac);                      // :245-249
            atomicAdd(a.Q + cell1, (p.I - dmax(I1, 0.0)) * p.dA);
            p.I = I1;
            if (p.I <= 0.01 * p.I0) { p.frac = 1.0; break; }
        }
    } while (p.keep == 1 && p.frac < 1);
    if (!err && p.keep == 2 && p.frac < 1) {
        // hand the live particle to the rank behind the patch; `face` carries its index within the patch
        const int bf = p.face - m.nFd, pa = m.bfPatch[bf];
        const int slot = a.outbox ? a.patchInfo[2 * pa] : -1;
        if (slot < 0) err = LASER_ERR_PROCESSOR;
        else {
            const int k = atomicAdd(a.counters + 4 + slot, 1);
            if (k >= a.mailCap) err = LASER_ERR_OVERFLOW;
            else {
                LaserPart q = p;
                q.face = bf - a.patchInfo[2 * pa + 1];
                a.outbox[(size_t)slot * a.mailCap + k] = q;
            }
        }
    }
    if (err) atomicMax(a.counters + 1, err);
    a.cloud[ip] = p;
}

// particles received through a coupled patch: continue in the cell behind the same face of this rank; a reflected
// particle still to be located (keep == 3) walks from that cell to its insertion point
__global__ void k_laser_receive(const __grid_constant__ DMesh m, const __grid_constant__ LaserGeo g, int n, const LaserPart *inbox,
                                int patchStart, LaserPart *dst, int *counters) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    LaserPart q = inbox[i];
    const int bf = patchStart + q.face;
    if (q.keep == 3) {
        int procFace;
        q.cell = l_walk_find(m, g, m.faceCells[bf], q.pos, &procFace);
        q.face = -1;
        q.keep = q.cell >= 0 ? 1 : 0;
        if (q.keep) atomicAdd(counters + 2, 1);
        else q.frac = 1.0;
    } else {
        q.cell = m.faceCells[bf];
        q.face = m.nFd + bf;
        q.keep = 1;
    }
    dst[i] = q;
}

__global__ void k_laser_points_gather(int n, const int *pts, const double *pv, double *out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const int p = pts[i]; out[3 * i] = pv[3 * (size_t)p]; out[3 * i + 1] = pv[3 * (size_t)p + 1]; out[3 * i + 2] = pv[3 * (size_t)p + 2]; }
}
__global__ void k_laser_points_scatter(int n, const int *pts, const double *in, double *pv) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const int p = pts[i]; pv[3 * (size_t)p] = in[3 * i]; pv[3 * (size_t)p + 1] = in[3 * i + 1]; pv[3 * (size_t)p + 2] = in[3 * i + 2]; }
}

__global__ void k_laser_finish(int n, const double *V, double *Q) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) Q[c] /= V[c];                                                          // laserDTRM.C:671
}

// the cells the beam deposited in, unordered: Q is zero almost everywhere, the host zeroes its field anyway
// (laserDTRM.C:580) and scatters these
__global__ void k_laser_compact(int n, const double *Q, int cap, int *cells, double *vals, int *counter) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    const double q = Q[c];
    if (q != 0.0) {
        const int k = atomicAdd(counter, 1);
        if (k < cap) { cells[k] = c; vals[k] = q; }
    }
}

__global__ void k_laser_bnd_index(int nBndCells, const int *bndCells, int *bndIdx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nBndCells) bndIdx[bndCells[i]] = i;
}

__global__ void k_laser_fd_to_face(int nIF, const int *faceMap, int *fdToFace) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < nIF) fdToFace[faceMap[f]] = f;
}

}  // namespace tcif
