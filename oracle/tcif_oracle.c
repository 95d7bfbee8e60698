/*
 * tcif_oracle.c -- CPU ORACLE (test infrastructure only; see tcif_oracle.h header).
 * PARITY UNPINNED (no OpenFOAM v2106 source/binary and no reference golden vectors
 * exist in this environment); restates the reference composition
 *   /root/reference/phasesSystem/phaseSystem/phaseSystem.C:1038-1105, 1132-1136, 1274-1295
 *   /root/reference/phasesSystem/multiphaseSystem/multiphaseSystem.C:296-375
 * over the published OpenFOAM-v2106 operators, one sweep per
 * operator, reference accumulation order.  Build: gcc -O3 -ffp-contract=off (no FMA
 * contraction, no -march=native), see oracle/Makefile.
 */
#define _GNU_SOURCE
#include "tcif_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define ORC_SMALL 1e-15
#define ORC_VSMALL 1e-300
#define ORC_ROOTVSMALL 1e-150

static void *xmalloc(size_t n) {
    void *p = malloc(n ? n : 1);
    if (!p) { fprintf(stderr, "tcif_oracle: out of memory (%zu bytes)\n", n); abort(); }
    return p;
}
static double *dnew(size_t n) { return (double *)xmalloc(n * sizeof(double)); }
static double *dzero(size_t n) { double *p = dnew(n); memset(p, 0, n * sizeof(double)); return p; }
static inline double dmin(double a, double b) { return a < b ? a : b; }   /* Foam::min */
static inline double dmax(double a, double b) { return a > b ? a : b; }   /* Foam::max */
static inline int is_coupled(const orc_mesh *m, int p) { return m->patchType[p] == ORC_PATCH_PROCESSOR; }

static int n_coupled_faces(const orc_mesh *m) {
    int n = 0;
    for (int p = 0; p < m->nPatches; ++p) if (is_coupled(m, p)) n += m->patchSize[p];
    return n;
}

/* processorFvPatchField::evaluate for all coupled patches at once: boundary values of
 * coupled faces := neighbour-rank patchInternalField. */
static void halo_swap(const orc_mesh *m, const orc_comm *c, int ncomp, const double *vi, double *vb) {
    int nc = n_coupled_faces(m);
    if (nc == 0) return;
    if (!c || !c->halo) { fprintf(stderr, "tcif_oracle: processor patch without halo hook\n"); abort(); }
    double *send = dnew((size_t)nc * ncomp), *recv = dnew((size_t)nc * ncomp);
    int k = 0;
    for (int p = 0; p < m->nPatches; ++p) {
        if (!is_coupled(m, p)) continue;
        for (int i = 0; i < m->patchSize[p]; ++i, ++k) {
            int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
            for (int d = 0; d < ncomp; ++d) send[(size_t)k * ncomp + d] = vi[(size_t)fc * ncomp + d];
        }
    }
    c->halo(c->user, ncomp, nc, send, recv);
    k = 0;
    for (int p = 0; p < m->nPatches; ++p) {
        if (!is_coupled(m, p)) continue;
        for (int i = 0; i < m->patchSize[p]; ++i, ++k) {
            int bf = m->patchStart[p] + i;
            for (int d = 0; d < ncomp; ++d) vb[(size_t)bf * ncomp + d] = recv[(size_t)k * ncomp + d];
        }
    }
    free(send); free(recv);
}

/* syncTools::syncFaceList(minEqOp) on a surface field's coupled boundary values. */
static void halo_min_faces(const orc_mesh *m, const orc_comm *c, double *lamb) {
    int nc = n_coupled_faces(m);
    if (nc == 0) return;
    double *send = dnew(nc), *recv = dnew(nc);
    int k = 0;
    for (int p = 0; p < m->nPatches; ++p) {
        if (!is_coupled(m, p)) continue;
        for (int i = 0; i < m->patchSize[p]; ++i, ++k) send[k] = lamb[m->patchStart[p] + i];
    }
    c->halo(c->user, 1, nc, send, recv);
    k = 0;
    for (int p = 0; p < m->nPatches; ++p) {
        if (!is_coupled(m, p)) continue;
        for (int i = 0; i < m->patchSize[p]; ++i, ++k) {
            int bf = m->patchStart[p] + i;
            lamb[bf] = dmin(lamb[bf], recv[k]);
        }
    }
    free(send); free(recv);
}

static void reduce(const orc_comm *c, int op, int n, double *v) {
    if (c && c->reduce) c->reduce(c->user, op, n, v);
}

/* ------------------------------------------------------------------ A.2 */
void orc_interpolate(const orc_mesh *m, int ncomp, const double *vi, const double *vb,
                     double *sfi, double *sfb) {
    for (int f = 0; f < m->nIF; ++f) {
        const double *vP = vi + (size_t)m->owner[f] * ncomp, *vN = vi + (size_t)m->neighbour[f] * ncomp;
        double w = m->w[f];
        for (int d = 0; d < ncomp; ++d) sfi[(size_t)f * ncomp + d] = w * (vP[d] - vN[d]) + vN[d];
    }
    for (int p = 0; p < m->nPatches; ++p) {
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i;
            if (is_coupled(m, p)) {
                const double *vP = vi + (size_t)m->faceCells[bf] * ncomp;
                double w = m->wb[bf];
                for (int d = 0; d < ncomp; ++d)
                    sfb[(size_t)bf * ncomp + d] = w * vP[d] + (1.0 - w) * vb[(size_t)bf * ncomp + d];
            } else {
                for (int d = 0; d < ncomp; ++d) sfb[(size_t)bf * ncomp + d] = vb[(size_t)bf * ncomp + d];
            }
        }
    }
}

/* ------------------------------------------------------------------ A.3 */
void orc_grad(const orc_mesh *m, const orc_comm *c, const double *ai, const double *ab,
              const int32_t *alphaBc, double *gi, double *gb) {
    double *sfi = dnew(m->nIF), *sfb = dnew(m->nBF);
    orc_interpolate(m, 1, ai, ab, sfi, sfb);
    memset(gi, 0, sizeof(double) * 3 * (size_t)m->nCells);
    for (int f = 0; f < m->nIF; ++f) {
        int o = m->owner[f], n = m->neighbour[f];
        for (int d = 0; d < 3; ++d) {
            double t = m->Sf[(size_t)f * 3 + d] * sfi[f];
            gi[(size_t)o * 3 + d] += t;
            gi[(size_t)n * 3 + d] -= t;
        }
    }
    for (int p = 0; p < m->nPatches; ++p)
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
            for (int d = 0; d < 3; ++d) gi[(size_t)fc * 3 + d] += m->Sfb[(size_t)bf * 3 + d] * sfb[bf];
        }
    for (int cI = 0; cI < m->nCells; ++cI)
        for (int d = 0; d < 3; ++d) gi[(size_t)cI * 3 + d] /= m->V[cI];
    /* gGrad.correctBoundaryConditions(): extrapolatedCalculated / processor */
    for (int p = 0; p < m->nPatches; ++p) {
        if (is_coupled(m, p)) continue;
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
            for (int d = 0; d < 3; ++d) gb[(size_t)bf * 3 + d] = gi[(size_t)fc * 3 + d];
        }
    }
    halo_swap(m, c, 3, gi, gb);
    /* gaussGrad::correctBoundaryConditions: non-coupled patches only */
    for (int p = 0; p < m->nPatches; ++p) {
        if (is_coupled(m, p)) continue;
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
            double n[3];
            for (int d = 0; d < 3; ++d) n[d] = m->Sfb[(size_t)bf * 3 + d] / m->magSfb[bf];
            double sn = 0.0;
            if (alphaBc[p] == ORC_BC_FIXED_VALUE) sn = m->deltaCoeffsb[bf] * (ab[bf] - ai[fc]);
            double *g = gb + (size_t)bf * 3;
            double ng = n[0] * g[0] + n[1] * g[1] + n[2] * g[2];
            double s = sn - ng;
            for (int d = 0; d < 3; ++d) g[d] += n[d] * s;
        }
    }
    free(sfi); free(sfb);
}

/* ------------------------------------------------------------------ A.4 */
void orc_div_tensor(const orc_mesh *m, const orc_comm *c, const double *Ti, const double *Tb,
                    double *di, double *db) {
    double *ssfi = dnew((size_t)m->nIF * 3), *ssfb = dnew((size_t)m->nBF * 3);
    /* dotInterpolate(Sf, T) */
    for (int f = 0; f < m->nIF; ++f) {
        const double *TP = Ti + (size_t)m->owner[f] * 9, *TN = Ti + (size_t)m->neighbour[f] * 9;
        const double *S = m->Sf + (size_t)f * 3;
        double w = m->w[f], t[9];
        for (int k = 0; k < 9; ++k) t[k] = w * (TP[k] - TN[k]) + TN[k];
        for (int j = 0; j < 3; ++j) ssfi[(size_t)f * 3 + j] = S[0] * t[j] + S[1] * t[3 + j] + S[2] * t[6 + j];
    }
    for (int p = 0; p < m->nPatches; ++p)
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i;
            const double *S = m->Sfb + (size_t)bf * 3;
            double t[9];
            if (is_coupled(m, p)) {
                const double *TP = Ti + (size_t)m->faceCells[bf] * 9;
                double w = m->wb[bf];
                for (int k = 0; k < 9; ++k) t[k] = w * TP[k] + (1.0 - w) * Tb[(size_t)bf * 9 + k];
            } else {
                for (int k = 0; k < 9; ++k) t[k] = Tb[(size_t)bf * 9 + k];
            }
            for (int j = 0; j < 3; ++j) ssfb[(size_t)bf * 3 + j] = S[0] * t[j] + S[1] * t[3 + j] + S[2] * t[6 + j];
        }
    /* fvc::surfaceIntegrate */
    memset(di, 0, sizeof(double) * 3 * (size_t)m->nCells);
    for (int f = 0; f < m->nIF; ++f) {
        int o = m->owner[f], n = m->neighbour[f];
        for (int d = 0; d < 3; ++d) {
            di[(size_t)o * 3 + d] += ssfi[(size_t)f * 3 + d];
            di[(size_t)n * 3 + d] -= ssfi[(size_t)f * 3 + d];
        }
    }
    for (int p = 0; p < m->nPatches; ++p)
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
            for (int d = 0; d < 3; ++d) di[(size_t)fc * 3 + d] += ssfb[(size_t)bf * 3 + d];
        }
    for (int cI = 0; cI < m->nCells; ++cI)
        for (int d = 0; d < 3; ++d) di[(size_t)cI * 3 + d] /= m->V[cI];
    /* correctBoundaryConditions */
    for (int p = 0; p < m->nPatches; ++p) {
        if (is_coupled(m, p)) continue;
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
            for (int d = 0; d < 3; ++d) db[(size_t)bf * 3 + d] = di[(size_t)fc * 3 + d];
        }
    }
    halo_swap(m, c, 3, di, db);
    free(ssfi); free(ssfb);
}

/* ------------------------------------------------------------------ A.5 */
double orc_sigma_eval(const orc_sigma *s, double T) {
    if (s->kind == ORC_SIGMA_CONSTANT) return s->c[0];
    if (s->kind == ORC_SIGMA_POLYNOMIAL) {
        double y = 0.0;
        for (int i = 0; i < s->n; ++i) y += s->c[2 * i] * pow(T, s->c[2 * i + 1]);
        return y;
    }
    /* table, linear interpolation weights, outOfBounds clamp */
    int n = s->n;
    if (T <= s->c[0]) return s->c[1];
    if (T >= s->c[2 * (n - 1)]) return s->c[2 * (n - 1) + 1];
    int i = 0;
    while (i < n - 2 && T >= s->c[2 * (i + 1)]) ++i;
    double t0 = s->c[2 * i], t1 = s->c[2 * (i + 1)];
    double w0 = (t1 - T) / (t1 - t0), w1 = 1.0 - w0;
    return w0 * s->c[2 * i + 1] + w1 * s->c[2 * (i + 1) + 1];
}

void orc_sigma_field(const orc_mesh *m, const orc_sigma *s, const double *Ti, const double *Tb,
                     double *si, double *sb) {
    for (int cI = 0; cI < m->nCells; ++cI) si[cI] = orc_sigma_eval(s, s->kind == ORC_SIGMA_CONSTANT ? 0.0 : Ti[cI]);
    for (int bf = 0; bf < m->nBF; ++bf) sb[bf] = orc_sigma_eval(s, s->kind == ORC_SIGMA_CONSTANT ? 0.0 : Tb[bf]);
}

double orc_deltaN(const orc_mesh *m, const orc_comm *c) {
    double v[2] = {0.0, (double)m->nCells};
    for (int cI = 0; cI < m->nCells; ++cI) v[0] += m->V[cI];
    reduce(c, ORC_RED_SUM, 2, v);
    return 1e-8 / cbrt(v[0] / v[1]);
}

/* gradAlpha = alpha2*grad(alpha1) - alpha1*grad(alpha2), internal + boundary
 * (phaseSystem.C:1074-1078, 1281-1285). */
static void pair_grad(const orc_mesh *m, const orc_comm *c,
                      const double *a1i, const double *a1b, const int32_t *bc1,
                      const double *a2i, const double *a2b, const int32_t *bc2,
                      double *gi, double *gb) {
    size_t nC = m->nCells, nB = m->nBF;
    double *g1i = dnew(nC * 3), *g1b = dnew(nB * 3), *g2i = dnew(nC * 3), *g2b = dnew(nB * 3);
    orc_grad(m, c, a1i, a1b, bc1, g1i, g1b);
    orc_grad(m, c, a2i, a2b, bc2, g2i, g2b);
    for (size_t k = 0; k < nC; ++k)
        for (int d = 0; d < 3; ++d) gi[k * 3 + d] = a2i[k] * g1i[k * 3 + d] - a1i[k] * g2i[k * 3 + d];
    for (size_t k = 0; k < nB; ++k)
        for (int d = 0; d < 3; ++d) gb[k * 3 + d] = a2b[k] * g1b[k * 3 + d] - a1b[k] * g2b[k * 3 + d];
    free(g1i); free(g1b); free(g2i); free(g2b);
}

static inline double vmag(const double *v) { return sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

void orc_nHat(const orc_mesh *m, const orc_comm *c,
              const double *a1i, const double *a1b, const int32_t *bc1,
              const double *a2i, const double *a2b, const int32_t *bc2,
              double *ni, double *nb) {
    size_t nC = m->nCells, nB = m->nBF;
    double *gi = dnew(nC * 3), *gb = dnew(nB * 3);
    pair_grad(m, c, a1i, a1b, bc1, a2i, a2b, bc2, gi, gb);
    double deltaN = orc_deltaN(m, c);
    for (size_t k = 0; k < nC; ++k) {
        double den = vmag(gi + k * 3) + deltaN;
        for (int d = 0; d < 3; ++d) ni[k * 3 + d] = gi[k * 3 + d] / den;
    }
    for (size_t k = 0; k < nB; ++k) {
        double den = vmag(gb + k * 3) + deltaN;
        for (int d = 0; d < 3; ++d) nb[k * 3 + d] = gb[k * 3 + d] / den;
    }
    free(gi); free(gb);
}

/* T = (2*mag(g)) * (sigma * (I - n1*n2))   (phaseSystem.C:1083-1096, note N1) */
static void stress_tensor(size_t n, const double *g, const double *sig, const double *n1, const double *n2, double *T) {
    for (size_t k = 0; k < n; ++k) {
        double tm = 2.0 * vmag(g + k * 3);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) {
                double I = (i == j) ? 1.0 : 0.0;
                double nn = n1[k * 3 + i] * n2[k * 3 + j];
                T[k * 9 + i * 3 + j] = tm * (sig[k] * (I - nn));
            }
    }
}

/* ------------------------------------------------- phaseSystem.C:1038-1105 */
void orc_div_capillary_stress(const orc_mesh *m, const orc_comm *c, int nPhases,
                              const double *const *alphai, const double *const *alphab,
                              const int32_t *alphaBc, const orc_sigma *sigma,
                              const double *Ti, const double *Tb,
                              double *Fi, double *Fb) {
    size_t nC = m->nCells, nB = m->nBF;
    memset(Fi, 0, sizeof(double) * 3 * nC);
    memset(Fb, 0, sizeof(double) * 3 * nB);
    double *gi = dnew(nC * 3), *gb = dnew(nB * 3);
    double *n1i = dnew(nC * 3), *n1b = dnew(nB * 3), *n2i = dnew(nC * 3), *n2b = dnew(nB * 3);
    double *si = dnew(nC), *sb = dnew(nB);
    double *TTi = dnew(nC * 9), *TTb = dnew(nB * 9);
    double *di = dnew(nC * 3), *db = dnew(nB * 3);
    int q = 0;
    for (int p1 = 0; p1 < nPhases; ++p1)
        for (int p2 = p1 + 1; p2 < nPhases; ++p2, ++q) {
            const int32_t *bc1 = alphaBc + (size_t)p1 * m->nPatches, *bc2 = alphaBc + (size_t)p2 * m->nPatches;
            pair_grad(m, c, alphai[p1], alphab[p1], bc1, alphai[p2], alphab[p2], bc2, gi, gb);
            /* nHat(alpha1,alpha2) is evaluated twice by the reference (:1092) */
            orc_nHat(m, c, alphai[p1], alphab[p1], bc1, alphai[p2], alphab[p2], bc2, n1i, n1b);
            orc_nHat(m, c, alphai[p1], alphab[p1], bc1, alphai[p2], alphab[p2], bc2, n2i, n2b);
            orc_sigma_field(m, &sigma[q], Ti, Tb, si, sb);
            stress_tensor(nC, gi, si, n1i, n2i, TTi);
            stress_tensor(nB, gb, sb, n1b, n2b, TTb);
            orc_div_tensor(m, c, TTi, TTb, di, db);
            for (size_t k = 0; k < nC * 3; ++k) Fi[k] += di[k];
            for (size_t k = 0; k < nB * 3; ++k) Fb[k] += db[k];
        }
    free(gi); free(gb); free(n1i); free(n1b); free(n2i); free(n2b); free(si); free(sb);
    free(TTi); free(TTb); free(di); free(db);
}

/* ------------------------------------------------------------------ A.6 */
void orc_mules_limiter(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg,
                       double rDeltaT, const double *psii, const double *psib,
                       const double *psi0, const int32_t *psiBc,
                       const double *phiBDi, const double *phiBDb,
                       const double *phiCorri, const double *phiCorrb,
                       const double *Sp, const double *Su, double psiMax, double psiMin,
                       double *lami, double *lamb) {
    int nC = m->nCells, nIF = m->nIF;
    double *psiMaxn = dnew(nC), *psiMinn = dnew(nC);
    double *sumPhiBD = dzero(nC), *sumPhip = dzero(nC), *mSumPhim = dzero(nC);
    for (int k = 0; k < nC; ++k) { psiMaxn[k] = psiMin; psiMinn[k] = psiMax; }
    for (int f = 0; f < nIF; ++f) {
        int own = m->owner[f], nei = m->neighbour[f];
        psiMaxn[own] = dmax(psiMaxn[own], psii[nei]);
        psiMinn[own] = dmin(psiMinn[own], psii[nei]);
        psiMaxn[nei] = dmax(psiMaxn[nei], psii[own]);
        psiMinn[nei] = dmin(psiMinn[nei], psii[own]);
        sumPhiBD[own] += phiBDi[f];
        sumPhiBD[nei] -= phiBDi[f];
        double pc = phiCorri[f];
        if (pc > 0) { sumPhip[own] += pc; mSumPhim[nei] += pc; }
        else        { mSumPhim[own] -= pc; sumPhip[nei] -= pc; }
    }
    for (int p = 0; p < m->nPatches; ++p) {
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
            if (is_coupled(m, p) || psiBc[p] == ORC_BC_FIXED_VALUE || cfg->boundaryExtrema == ORC_BEXT_PATCHVALUE) {
                /* coupled: patchNeighbourField (= halo value held in psib); fixesValue: patch value */
                psiMaxn[fc] = dmax(psiMaxn[fc], psib[bf]);
                psiMinn[fc] = dmin(psiMinn[fc], psib[bf]);
            } else {
                psiMaxn[fc] = dmax(psiMaxn[fc], psiMax);
                psiMinn[fc] = dmin(psiMinn[fc], psiMin);
            }
        }
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
            sumPhiBD[fc] += phiBDb[bf];
            double pc = phiCorrb[bf];
            if (pc > 0) sumPhip[fc] += pc; else mSumPhim[fc] -= pc;
        }
    }
    for (int k = 0; k < nC; ++k) {
        psiMaxn[k] = dmin(psiMaxn[k] + cfg->extremaCoeff * (psiMax - psiMin), psiMax);
        psiMinn[k] = dmax(psiMinn[k] - cfg->extremaCoeff * (psiMax - psiMin), psiMin);
    }
    if (cfg->smoothLimiter > ORC_SMALL) {
        double s = cfg->smoothLimiter;
        for (int k = 0; k < nC; ++k) {
            psiMaxn[k] = dmin(s * psii[k] + (1.0 - s) * psiMaxn[k], psiMax);
            psiMinn[k] = dmax(s * psii[k] + (1.0 - s) * psiMinn[k], psiMin);
        }
    }
    for (int k = 0; k < nC; ++k) {       /* rho = rho.oldTime() = 1 (geometricOneField) */
        double V = m->V[k];
        double qp = V * ((rDeltaT - Sp[k]) * psiMaxn[k] - Su[k]) - (V * rDeltaT) * psi0[k] + sumPhiBD[k];
        double qm = V * (Su[k] - (rDeltaT - Sp[k]) * psiMinn[k]) + (V * rDeltaT) * psi0[k] - sumPhiBD[k];
        psiMaxn[k] = qp; psiMinn[k] = qm;
    }
    for (int f = 0; f < nIF; ++f) lami[f] = 1.0;
    for (int b = 0; b < m->nBF; ++b) lamb[b] = 1.0;
    double *sumlPhip = dnew(nC), *mSumlPhim = dnew(nC);
    for (int j = 0; j < cfg->nLimiterIter; ++j) {
        memset(sumlPhip, 0, sizeof(double) * nC);
        memset(mSumlPhim, 0, sizeof(double) * nC);
        for (int f = 0; f < nIF; ++f) {
            int own = m->owner[f], nei = m->neighbour[f];
            double lp = lami[f] * phiCorri[f];
            if (lp > 0) { sumlPhip[own] += lp; mSumlPhim[nei] += lp; }
            else        { mSumlPhim[own] -= lp; sumlPhip[nei] -= lp; }
        }
        for (int p = 0; p < m->nPatches; ++p)
            for (int i = 0; i < m->patchSize[p]; ++i) {
                int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
                double lp = lamb[bf] * phiCorrb[bf];
                if (lp > 0) sumlPhip[fc] += lp; else mSumlPhim[fc] -= lp;
            }
        for (int k = 0; k < nC; ++k) {
            sumlPhip[k]  = dmax(dmin((sumlPhip[k]  + psiMaxn[k]) / (mSumPhim[k] + ORC_ROOTVSMALL), 1.0), 0.0);
            mSumlPhim[k] = dmax(dmin((mSumlPhim[k] + psiMinn[k]) / (sumPhip[k]  + ORC_ROOTVSMALL), 1.0), 0.0);
        }
        const double *lambdam = sumlPhip, *lambdap = mSumlPhim;
        for (int f = 0; f < nIF; ++f) {
            int own = m->owner[f], nei = m->neighbour[f];
            if (phiCorri[f] > 0) lami[f] = dmin(lami[f], dmin(lambdap[own], lambdam[nei]));
            else                 lami[f] = dmin(lami[f], dmin(lambdam[own], lambdap[nei]));
        }
        for (int p = 0; p < m->nPatches; ++p) {
            for (int i = 0; i < m->patchSize[p]; ++i) {
                int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
                if (m->patchType[p] == ORC_PATCH_WEDGE) {
                    lamb[bf] = 0.0;
                } else if (is_coupled(m, p)) {
                    if (phiCorrb[bf] > 0) lamb[bf] = dmin(lamb[bf], lambdap[fc]);
                    else                  lamb[bf] = dmin(lamb[bf], lambdam[fc]);
                } else if ((phiBDb[bf] + phiCorrb[bf]) > ORC_SMALL * ORC_SMALL) {   /* outflow faces only */
                    if (phiCorrb[bf] > 0) lamb[bf] = dmin(lamb[bf], lambdap[fc]);
                    else                  lamb[bf] = dmin(lamb[bf], lambdam[fc]);
                }
            }
        }
        halo_min_faces(m, c, lamb);
    }
    free(psiMaxn); free(psiMinn); free(sumPhiBD); free(sumPhip); free(mSumPhim);
    free(sumlPhip); free(mSumlPhim);
}

/* upwind<scalar>::flux(psi) = phi * interpolate(psi, weights = pos0(phi)) */
void orc_upwind_flux(const orc_mesh *m, const double *phii, const double *phib,
                     const double *psii, const double *psib, double *outi, double *outb, int add) {
    for (int f = 0; f < m->nIF; ++f) {
        double w = phii[f] >= 0 ? 1.0 : 0.0;
        double pP = psii[m->owner[f]], pN = psii[m->neighbour[f]];
        double v = phii[f] * (w * (pP - pN) + pN);
        outi[f] = add ? outi[f] + v : v;
    }
    for (int p = 0; p < m->nPatches; ++p)
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i;
            double v;
            if (is_coupled(m, p)) {
                double w = phib[bf] >= 0 ? 1.0 : 0.0;
                v = phib[bf] * (w * psii[m->faceCells[bf]] + (1.0 - w) * psib[bf]);
            } else {
                v = phib[bf] * psib[bf];
            }
            outb[bf] = add ? outb[bf] + v : v;
        }
}

void orc_mules_limit(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg,
                     double rDeltaT, const double *psii, const double *psib,
                     const double *psi0, const int32_t *psiBc,
                     const double *phii, const double *phib,
                     double *phiPsii, double *phiPsib,
                     const double *Sp, const double *Su, double psiMax, double psiMin,
                     int returnCorr) {
    int nIF = m->nIF, nBF = m->nBF;
    double *phiBDi = dnew(nIF), *phiBDb = dnew(nBF), *lami = dnew(nIF), *lamb = dnew(nBF);
    orc_upwind_flux(m, phii, phib, psii, psib, phiBDi, phiBDb, 0);
    for (int p = 0; p < m->nPatches; ++p) {
        if (is_coupled(m, p)) continue;
        for (int i = 0; i < m->patchSize[p]; ++i) { int bf = m->patchStart[p] + i; phiBDb[bf] = phiPsib[bf]; }
    }
    for (int f = 0; f < nIF; ++f) phiPsii[f] -= phiBDi[f];       /* phiCorr = phiPsi - phiBD (aliases phiPsi) */
    for (int b = 0; b < nBF; ++b) phiPsib[b] -= phiBDb[b];
    orc_mules_limiter(m, c, cfg, rDeltaT, psii, psib, psi0, psiBc, phiBDi, phiBDb, phiPsii, phiPsib,
                      Sp, Su, psiMax, psiMin, lami, lamb);
    if (returnCorr) {
        for (int f = 0; f < nIF; ++f) phiPsii[f] *= lami[f];
        for (int b = 0; b < nBF; ++b) phiPsib[b] *= lamb[b];
    } else {
        for (int f = 0; f < nIF; ++f) phiPsii[f] = phiBDi[f] + lami[f] * phiPsii[f];
        for (int b = 0; b < nBF; ++b) phiPsib[b] = phiBDb[b] + lamb[b] * phiPsib[b];
    }
    free(phiBDi); free(phiBDb); free(lami); free(lamb);
}

static void limit_sum_list(int n, int nPhases, double *const *fl) {
    for (int f = 0; f < n; ++f) {
        double sumPos = 0, sumNeg = 0;
        for (int p = 0; p < nPhases; ++p) {
            if (fl[p][f] > 0) sumPos += fl[p][f]; else sumNeg += fl[p][f];
        }
        double sum = sumPos + sumNeg;
        if (sum > 0 && sumPos > ORC_VSMALL) {
            double lambda = -sumNeg / sumPos;
            for (int p = 0; p < nPhases; ++p) if (fl[p][f] > 0) fl[p][f] *= lambda;
        } else if (sum < 0 && sumNeg < -ORC_VSMALL) {
            double lambda = -sumPos / sumNeg;
            for (int p = 0; p < nPhases; ++p) if (fl[p][f] < 0) fl[p][f] *= lambda;
        }
    }
}

void orc_mules_limit_sum(const orc_mesh *m, int nPhases, double *const *phiCorri, double *const *phiCorrb) {
    limit_sum_list(m->nIF, nPhases, phiCorri);
    double **pl = (double **)xmalloc(sizeof(double *) * nPhases);
    for (int p = 0; p < m->nPatches; ++p) {
        if (!is_coupled(m, p)) continue;
        for (int k = 0; k < nPhases; ++k) pl[k] = phiCorrb[k] + m->patchStart[p];
        limit_sum_list(m->patchSize[p], nPhases, pl);
    }
    free(pl);
}

void orc_correct_bcs(const orc_mesh *m, const orc_comm *c, int ncomp, const double *vi, double *vb, const int32_t *bc) {
    for (int p = 0; p < m->nPatches; ++p) {
        if (is_coupled(m, p) || bc[p] != ORC_BC_ZERO_GRADIENT) continue;
        for (int i = 0; i < m->patchSize[p]; ++i) {
            int bf = m->patchStart[p] + i, fc = m->faceCells[bf];
            for (int d = 0; d < ncomp; ++d) vb[(size_t)bf * ncomp + d] = vi[(size_t)fc * ncomp + d];
        }
    }
    halo_swap(m, c, ncomp, vi, vb);
}

void orc_mules_explicit_solve(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg,
                              double deltaT, double *psii, double *psib,
                              const double *psi0, const int32_t *psiBc,
                              const double *phii, const double *phib,
                              double *phiPsii, double *phiPsib,
                              const double *Sp, const double *Su, double psiMax, double psiMin) {
    int nC = m->nCells;
    orc_correct_bcs(m, c, 1, psii, psib, psiBc);
    double rDeltaT = 1.0 / deltaT;
    orc_mules_limit(m, c, cfg, rDeltaT, psii, psib, psi0, psiBc, phii, phib, phiPsii, phiPsib,
                    Sp, Su, psiMax, psiMin, 0);
    /* psiIf = 0; fvc::surfaceIntegrate(psiIf, phiPsi) */
    double *d = dzero(nC);
    for (int f = 0; f < m->nIF; ++f) { d[m->owner[f]] += phiPsii[f]; d[m->neighbour[f]] -= phiPsii[f]; }
    for (int p = 0; p < m->nPatches; ++p)
        for (int i = 0; i < m->patchSize[p]; ++i) { int bf = m->patchStart[p] + i; d[m->faceCells[bf]] += phiPsib[bf]; }
    for (int k = 0; k < nC; ++k) d[k] /= m->V[k];
    for (int k = 0; k < nC; ++k) psii[k] = (psi0[k] * rDeltaT + Su[k] - d[k]) / (rDeltaT - Sp[k]);
    orc_correct_bcs(m, c, 1, psii, psib, psiBc);
    free(d);
}

/* --------------------------------------------------------------- stage */
void orc_stage(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg, int nPhases,
               double *const *alphai, double *const *alphab, const int32_t *alphaBc,
               const orc_sigma *sigma, const double *Ti, const double *Tb,
               const double *phii, const double *phib,
               double *const *phiAlphai, double *const *phiAlphab,
               const double *const *Su, const double *const *Sp, double deltaT,
               double *Fi, double *Fb) {
    int nC = m->nCells;
    orc_div_capillary_stress(m, c, nPhases, (const double *const *)alphai, (const double *const *)alphab,
                             alphaBc, sigma, Ti, Tb, Fi, Fb);
    /* old-time fields: psi.oldTime() */
    double **a0 = (double **)xmalloc(sizeof(double *) * nPhases);
    double *zero = dzero(nC);
    for (int p = 0; p < nPhases; ++p) { a0[p] = dnew(nC); memcpy(a0[p], alphai[p], sizeof(double) * nC); }
    double rDeltaT = 1.0 / deltaT;
    for (int p = 0; p < nPhases; ++p)      /* multiphaseSystem.C:296-319 */
        orc_mules_limit(m, c, cfg, rDeltaT, alphai[p], alphab[p], a0[p], alphaBc + (size_t)p * m->nPatches,
                        phii, phib, phiAlphai[p], phiAlphab[p], Sp ? Sp[p] : zero, Su ? Su[p] : zero, 1.0, 0.0, 1);
    orc_mules_limit_sum(m, nPhases, phiAlphai, phiAlphab);   /* :321 */
    for (int p = 0; p < nPhases; ++p) {    /* :336-380 with the bounded upwind flux of :346 standing in for the predictor */
        orc_upwind_flux(m, phii, phib, alphai[p], alphab[p], phiAlphai[p], phiAlphab[p], 1);
        orc_mules_explicit_solve(m, c, cfg, deltaT, alphai[p], alphab[p], a0[p], alphaBc + (size_t)p * m->nPatches,
                                 phii, phib, phiAlphai[p], phiAlphab[p], Sp ? Sp[p] : zero, Su ? Su[p] : zero, 1.0, 0.0);
    }
    for (int p = 0; p < nPhases; ++p) free(a0[p]);
    free(a0); free(zero);
}

/* ---------------------------------------- threaded multi-domain driver */
typedef struct thr_shared {
    int nDom;
    const orc_mesh *meshes;
    pthread_barrier_t bar;
    const double **sendPtr;     /* per domain: current packed send buffer */
    double *redbuf;             /* nDom * 8 */
} thr_shared;

typedef struct thr_ctx { thr_shared *sh; int dom; } thr_ctx;

static void thr_halo(void *user, int ncomp, int nc, const double *send, double *recv) {
    thr_ctx *t = (thr_ctx *)user;
    thr_shared *sh = t->sh;
    (void)nc;
    sh->sendPtr[t->dom] = send;
    pthread_barrier_wait(&sh->bar);
    const orc_mesh *m = &sh->meshes[t->dom];
    int k = 0;
    for (int p = 0; p < m->nPatches; ++p) {
        if (m->patchType[p] != ORC_PATCH_PROCESSOR) continue;
        int e = m->patchNbrProc[p];
        const orc_mesh *me = &sh->meshes[e];
        /* offset of the matching patch (nbrProc == dom) in e's packed buffer */
        int off = 0, found = 0;
        for (int q = 0; q < me->nPatches; ++q) {
            if (me->patchType[q] != ORC_PATCH_PROCESSOR) continue;
            if (me->patchNbrProc[q] == t->dom) { found = 1; break; }
            off += me->patchSize[q];
        }
        if (!found) { fprintf(stderr, "tcif_oracle: unmatched processor patch\n"); abort(); }
        memcpy(recv + (size_t)k * ncomp, sh->sendPtr[e] + (size_t)off * ncomp, sizeof(double) * ncomp * m->patchSize[p]);
        k += m->patchSize[p];
    }
    pthread_barrier_wait(&sh->bar);
}

static void thr_reduce(void *user, int op, int n, double *v) {
    thr_ctx *t = (thr_ctx *)user;
    thr_shared *sh = t->sh;
    if (n > 8) abort();
    memcpy(sh->redbuf + (size_t)t->dom * 8, v, sizeof(double) * n);
    pthread_barrier_wait(&sh->bar);
    for (int i = 0; i < n; ++i) {
        double r = sh->redbuf[i];
        for (int d = 1; d < sh->nDom; ++d) {
            double x = sh->redbuf[(size_t)d * 8 + i];
            r = op == ORC_RED_SUM ? r + x : op == ORC_RED_MIN ? dmin(r, x) : dmax(r, x);
        }
        v[i] = r;
    }
    pthread_barrier_wait(&sh->bar);
}

typedef struct thr_args {
    thr_ctx tc; const orc_mules_cfg *cfg; int nPhases;
    double *const *alphai; double *const *alphab; const int32_t *alphaBc; const orc_sigma *sigma;
    const double *Ti, *Tb, *phii, *phib; double *const *phiAlphai; double *const *phiAlphab;
    double deltaT; double *Fi, *Fb; int nStages; double seconds;
} thr_args;

static double now_s(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }

/* first touch: a worker copies what it is going to stream (its sub-mesh and its fields) into pages it allocates and
 * writes itself, so that on a multi-socket host they live on the worker's own NUMA node -- what an MPI rank of the
 * reference gets for free by reading its processor directory itself.  The callers' arrays were written by ONE thread. */
static double *dclone(const double *src, size_t n) { double *d = dnew(n ? n : 1); if (n) memcpy(d, src, sizeof(double) * n); return d; }
static int32_t *iclone(const int32_t *src, size_t n) { int32_t *d = (int32_t *)xmalloc(sizeof(int32_t) * (n ? n : 1)); if (n) memcpy(d, src, sizeof(int32_t) * n); return d; }

static void *thr_main(void *p) {
    thr_args *a = (thr_args *)p;
    const orc_mesh *m0 = &a->tc.sh->meshes[a->tc.dom];
    orc_comm comm = { &a->tc, thr_halo, thr_reduce };
    int nP = a->nPhases;
    const size_t nC = m0->nCells, nIF = m0->nIF, nBF = m0->nBF;
    /* ORC_FIRST_TOUCH=0: stream the caller's arrays as they are (A/B switch for the NUMA effect) */
    const char *e = getenv("ORC_FIRST_TOUCH");
    const int ft = !(e && e[0] == '0');
    /* local sub-mesh (patch tables stay shared: the halo hook reads the neighbours' through sh->meshes) */
    orc_mesh lm = *m0;
    int32_t *own = NULL, *nei = NULL, *fc = NULL;
    double *Sf = NULL, *w = NULL, *V = NULL, *Sfb = NULL, *magSfb = NULL, *dcb = NULL, *wb = NULL;
    if (ft) {
        own = iclone(m0->owner, nIF); nei = iclone(m0->neighbour, nIF); fc = iclone(m0->faceCells, nBF);
        Sf = dclone(m0->Sf, 3 * nIF); w = dclone(m0->w, nIF); V = dclone(m0->V, nC); Sfb = dclone(m0->Sfb, 3 * nBF);
        magSfb = dclone(m0->magSfb, nBF); dcb = dclone(m0->deltaCoeffsb, nBF); wb = dclone(m0->wb, nBF);
        lm.owner = own; lm.neighbour = nei; lm.faceCells = fc; lm.Sf = Sf; lm.w = w; lm.V = V; lm.Sfb = Sfb;
        lm.magSfb = magSfb; lm.deltaCoeffsb = dcb; lm.wb = wb;
    }
    const orc_mesh *m = &lm;
    double **ai = (double **)xmalloc(sizeof(double *) * nP), **ab = (double **)xmalloc(sizeof(double *) * nP);
    double **pai = (double **)xmalloc(sizeof(double *) * nP), **pab = (double **)xmalloc(sizeof(double *) * nP);
    /* keep the HO flux so every stage starts from the same phiAlpha input */
    double **pa0i = (double **)xmalloc(sizeof(double *) * nP), **pa0b = (double **)xmalloc(sizeof(double *) * nP);
    for (int k = 0; k < nP; ++k) {
        ai[k] = ft ? dclone(a->alphai[k], nC) : a->alphai[k]; ab[k] = ft ? dclone(a->alphab[k], nBF) : a->alphab[k];
        pai[k] = ft ? dclone(a->phiAlphai[k], nIF) : a->phiAlphai[k]; pab[k] = ft ? dclone(a->phiAlphab[k], nBF) : a->phiAlphab[k];
        pa0i[k] = dclone(a->phiAlphai[k], nIF); pa0b[k] = dclone(a->phiAlphab[k], nBF);
    }
    const double *Ti = a->Ti, *Tb = a->Tb, *phii = a->phii, *phib = a->phib;
    double *TiL = NULL, *TbL = NULL, *phiiL = NULL, *phibL = NULL, *Fi = a->Fi, *Fb = a->Fb;
    if (ft) {
        if (a->Ti) Ti = TiL = dclone(a->Ti, nC);
        if (a->Tb) Tb = TbL = dclone(a->Tb, nBF);
        phii = phiiL = dclone(a->phii, nIF); phib = phibL = dclone(a->phib, nBF);
        Fi = dnew(3 * nC + 1); Fb = dnew(3 * nBF + 1);
    }
    pthread_barrier_wait(&a->tc.sh->bar);
    double t0 = now_s();
    for (int s = 0; s < a->nStages; ++s) {
        if (s > 0)
            for (int k = 0; k < nP; ++k) {
                memcpy(pai[k], pa0i[k], sizeof(double) * nIF);
                memcpy(pab[k], pa0b[k], sizeof(double) * nBF);
            }
        orc_stage(m, &comm, a->cfg, nP, ai, ab, a->alphaBc, a->sigma, Ti, Tb, phii, phib, pai, pab, NULL, NULL, a->deltaT, Fi, Fb);
    }
    pthread_barrier_wait(&a->tc.sh->bar);
    a->seconds = now_s() - t0;
    for (int k = 0; k < nP; ++k) {
        if (ft) {      /* results back into the caller's arrays */
            memcpy(a->alphai[k], ai[k], sizeof(double) * nC); memcpy(a->alphab[k], ab[k], sizeof(double) * nBF);
            memcpy(a->phiAlphai[k], pai[k], sizeof(double) * nIF); memcpy(a->phiAlphab[k], pab[k], sizeof(double) * nBF);
            free(ai[k]); free(ab[k]); free(pai[k]); free(pab[k]);
        }
        free(pa0i[k]); free(pa0b[k]);
    }
    if (ft) {
        memcpy(a->Fi, Fi, sizeof(double) * 3 * nC); memcpy(a->Fb, Fb, sizeof(double) * 3 * nBF);
        free(Fi); free(Fb); free(TiL); free(TbL); free(phiiL); free(phibL);
        free(own); free(nei); free(fc); free(Sf); free(w); free(V); free(Sfb); free(magSfb); free(dcb); free(wb);
    }
    free(ai); free(ab); free(pai); free(pab); free(pa0i); free(pa0b);
    return NULL;
}

double orc_stage_threads(int nDom, const orc_mesh *meshes, const orc_mules_cfg *cfg, int nPhases,
                         double *const *const *alphai, double *const *const *alphab,
                         const int32_t *const *alphaBc, const orc_sigma *sigma,
                         const double *const *Ti, const double *const *Tb,
                         const double *const *phii, const double *const *phib,
                         double *const *const *phiAlphai, double *const *const *phiAlphab,
                         double deltaT, double *const *Fi, double *const *Fb, int nStages) {
    thr_shared sh;
    sh.nDom = nDom; sh.meshes = meshes;
    sh.sendPtr = (const double **)xmalloc(sizeof(double *) * nDom);
    sh.redbuf = dnew((size_t)nDom * 8);
    pthread_barrier_init(&sh.bar, NULL, nDom);
    thr_args *args = (thr_args *)xmalloc(sizeof(thr_args) * nDom);
    pthread_t *th = (pthread_t *)xmalloc(sizeof(pthread_t) * nDom);
    for (int d = 0; d < nDom; ++d) {
        thr_args *a = &args[d];
        a->tc.sh = &sh; a->tc.dom = d; a->cfg = cfg; a->nPhases = nPhases;
        a->alphai = alphai[d]; a->alphab = alphab[d]; a->alphaBc = alphaBc[d]; a->sigma = sigma;
        a->Ti = Ti[d]; a->Tb = Tb[d]; a->phii = phii[d]; a->phib = phib[d];
        a->phiAlphai = phiAlphai[d]; a->phiAlphab = phiAlphab[d];
        a->deltaT = deltaT; a->Fi = Fi[d]; a->Fb = Fb[d]; a->nStages = nStages; a->seconds = 0;
        pthread_create(&th[d], NULL, thr_main, a);
    }
    double tmax = 0;
    for (int d = 0; d < nDom; ++d) { pthread_join(th[d], NULL); if (args[d].seconds > tmax) tmax = args[d].seconds; }
    pthread_barrier_destroy(&sh.bar);
    free(sh.sendPtr); free(sh.redbuf); free(args); free(th);
    return tmax;
}
#include "tcif_oracle_mass.inc"
#include "tcif_oracle_alpha.inc"
#include "tcif_oracle_species.inc"
#include "tcif_oracle_laser.inc"
