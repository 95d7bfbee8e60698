/*
 * tcif_oracle.h -- CPU ORACLE for the thermocapillaryInterFoam interface hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product path
 * (thermocapillaryinterfoam_b200/, include/tcif.h) never includes, links or calls it.
 *
 * PARITY UNPINNED: the arithmetic of this path lives in OpenFOAM v2106 (OpenCFD/ESI),
 * a third-party dependency that is neither vendored under /root/reference nor
 * installed in this image (README.md:5,11,47-49 and .github/workflows/build.yml:23 of
 * the reference pin it by prose / CI image only), and the reference ships no test,
 * tutorial or golden vector for this path.  The oracle therefore
 * restates (i) the reference's own composition of operators
 *     phasesSystem/phaseSystem/phaseSystem.C:1038-1105   divCapillaryStress()
 *     phasesSystem/phaseSystem/phaseSystem.C:1274-1295   nHat()
 *     phasesSystem/phaseSystem/phaseSystem.C:1132-1136   surfaceTensionCoeff()
 *     phasesSystem/multiphaseSystem/multiphaseSystem.C:296-375  limit/limitSum/explicitSolve calls
 * and (ii) the published OpenFOAM-v2106 algorithms those calls resolve to
 * (gaussGrad, gaussDivScheme, linear interpolation, MULES limiter / limit / limitSum /
 * explicitSolve, surfaceTensionModels constant + temperatureDependent), loop for loop,
 * one memory sweep per OpenFOAM operator with full-field temporaries, in the
 * reference's accumulation order.  It is pinned by known-answer tests only
 * (tests/test_oracle_known_answers.py).
 *
 * Layout conventions = OpenFOAM's: scalar = double, label = int32, vectors AoS (x,y,z),
 * tensors AoS row-major (xx,xy,xz,yx,yy,yz,zx,zy,zz); a vol field is internal[nCells]
 * plus boundary[nBF] (patch arrays concatenated in patch order; empty patches have
 * size 0); a surface field is internal[nIF] plus boundary[nBF].  On coupled
 * (processor) patches the boundary value of a vol field is the neighbour-rank cell
 * value (= patchNeighbourField()).
 */
#ifndef TCIF_ORACLE_H
#define TCIF_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_PATCH_GENERIC = 0, ORC_PATCH_EMPTY = 1, ORC_PATCH_WEDGE = 2, ORC_PATCH_PROCESSOR = 3 };
enum { ORC_BC_ZERO_GRADIENT = 0, ORC_BC_FIXED_VALUE = 1 };
enum { ORC_SIGMA_CONSTANT = 0, ORC_SIGMA_POLYNOMIAL = 1, ORC_SIGMA_TABLE = 2 };
enum { ORC_RED_SUM = 0, ORC_RED_MIN = 1, ORC_RED_MAX = 2 };
enum { ORC_BEXT_GLOBAL = 0, ORC_BEXT_PATCHVALUE = 1 };

typedef struct orc_mesh {
    int32_t nCells, nIF, nPatches, nBF;
    const int32_t *owner, *neighbour;          /* [nIF]  lduAddr lower/upper            */
    const double *Sf;                          /* [nIF*3]                               */
    const double *w;                           /* [nIF]  mesh.weights() (linear)        */
    const double *V;                           /* [nCells]                              */
    const int32_t *patchType, *patchStart, *patchSize, *patchNbrProc;  /* [nPatches]    */
    const int32_t *faceCells;                  /* [nBF]                                 */
    const double *Sfb;                         /* [nBF*3]                               */
    const double *magSfb;                      /* [nBF]                                 */
    const double *deltaCoeffsb;                /* [nBF]                                 */
    const double *wb;                          /* [nBF]  1 on non-coupled patches       */
    /* only needed by the NVD/TVD convection schemes (vanLeer), may be NULL otherwise */
    const double *C;                           /* [nCells*3] cell centres mesh.C()      */
    const double *deltab;                      /* [nBF*3] fvPatch::delta(): on processor
                                                  faces own centre -> neighbour centre   */
} orc_mesh;

/* Function1<scalar> "sigma": constant | polynomial ((a e)...) | table ((T s)...) clamp */
typedef struct orc_sigma {
    int32_t kind, n;
    double c[64];      /* constant: c[0]; polynomial: (a_i,e_i) pairs; table: (T_i,s_i) pairs */
} orc_sigma;

/* Per-rank communication hooks (Pstream stand-in).  halo(): `send` holds, for every
 * coupled boundary face in ascending boundary index, ncomp interleaved doubles of the
 * local patchInternalField; `recv` returns the neighbour rank's values in the same
 * layout.  reduce(): in-place all-reduce of n doubles.  NULL hooks = serial run. */
typedef void (*orc_halo_fn)(void *user, int ncomp, int nCoupledFaces, const double *send, double *recv);
typedef void (*orc_reduce_fn)(void *user, int op, int n, double *vals);
typedef struct orc_comm { void *user; orc_halo_fn halo; orc_reduce_fn reduce; } orc_comm;

typedef struct orc_mules_cfg {
    int32_t nLimiterIter;        /* 3   */
    double smoothLimiter;        /* 0   */
    double extremaCoeff;         /* 0   */
    int32_t boundaryExtrema;     /* ORC_BEXT_GLOBAL */
} orc_mules_cfg;

/* ---- operators ---- */

/* A.2 fvc::interpolate(vf) with `linear`, ncomp components interleaved. */
void orc_interpolate(const orc_mesh *m, int ncomp, const double *vi, const double *vb,
                     double *sfi, double *sfb);

/* A.3 fvc::grad(alpha) with Gauss linear + gaussGrad::correctBoundaryConditions.
 * alphaBc[nPatches] gives the BC kind of alpha on non-coupled patches. */
void orc_grad(const orc_mesh *m, const orc_comm *c, const double *ai, const double *ab,
              const int32_t *alphaBc, double *gi, double *gb);

/* A.4 fvc::div(T), T a volTensorField (9 comps). */
void orc_div_tensor(const orc_mesh *m, const orc_comm *c, const double *Ti, const double *Tb,
                    double *di, double *db);

/* A.5 sigma field from temperature. */
void orc_sigma_field(const orc_mesh *m, const orc_sigma *s, const double *Ti, const double *Tb,
                     double *si, double *sb);
double orc_sigma_eval(const orc_sigma *s, double T);

/* deltaN = 1e-8/cbrt(average(V))  (phaseSystem.C:1287-1291). */
double orc_deltaN(const orc_mesh *m, const orc_comm *c);

/* phaseSystem::nHat (phaseSystem.C:1274-1295). */
void orc_nHat(const orc_mesh *m, const orc_comm *c,
              const double *a1i, const double *a1b, const int32_t *bc1,
              const double *a2i, const double *a2b, const int32_t *bc2,
              double *ni, double *nb);

/* phaseSystem::divCapillaryStress (phaseSystem.C:1038-1105).  alphas in table order;
 * sigma[q] for pairs q enumerated (0,1),(0,2),..,(1,2),..; alphaBc[p*nPatches+patch]. */
void orc_div_capillary_stress(const orc_mesh *m, const orc_comm *c, int nPhases,
                              const double *const *alphai, const double *const *alphab,
                              const int32_t *alphaBc, const orc_sigma *sigma,
                              const double *Ti, const double *Tb,
                              double *Fi, double *Fb);

/* A.6 MULES::limiter: lambda (internal [nIF] + boundary [nBF]) out. */
void orc_mules_limiter(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg,
                       double rDeltaT, const double *psii, const double *psib,
                       const double *psi0, const int32_t *psiBc,
                       const double *phiBDi, const double *phiBDb,
                       const double *phiCorri, const double *phiCorrb,
                       const double *Sp, const double *Su, double psiMax, double psiMin,
                       double *lami, double *lamb);

/* A.6 MULES::limit: phiPsi in/out. */
void orc_mules_limit(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg,
                     double rDeltaT, const double *psii, const double *psib,
                     const double *psi0, const int32_t *psiBc,
                     const double *phii, const double *phib,
                     double *phiPsii, double *phiPsib,
                     const double *Sp, const double *Su, double psiMax, double psiMin,
                     int returnCorr);

/* A.6 MULES::limitSum over nPhases correction fluxes (internal + coupled patches). */
void orc_mules_limit_sum(const orc_mesh *m, int nPhases, double *const *phiCorri,
                         double *const *phiCorrb);

/* A.6 MULES::explicitSolve: psi (internal+boundary) and phiPsi in/out. */
void orc_mules_explicit_solve(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg,
                              double deltaT, double *psii, double *psib,
                              const double *psi0, const int32_t *psiBc,
                              const double *phii, const double *phib,
                              double *phiPsii, double *phiPsib,
                              const double *Sp, const double *Su, double psiMax, double psiMin);

/* upwind flux phi*psi_upwind (A.7), added into out (out += flux) when add != 0. */
void orc_upwind_flux(const orc_mesh *m, const double *phii, const double *phib,
                     const double *psii, const double *psib,
                     double *outi, double *outb, int add);

/* vol-field boundary evaluation: zeroGradient -> adjacent cell, fixedValue -> kept,
 * processor -> halo. */
void orc_correct_bcs(const orc_mesh *m, const orc_comm *c, int ncomp, const double *vi,
                     double *vb, const int32_t *bc);

/* One metric "stage" (SURVEY.md 8(d)): divCapillaryStress over all pairs + one MULES
 * corrector over all phases (limit(returnCorr) x P, limitSum, then per phase
 * phiAlpha = corr + upwind flux; explicitSolve).  alpha (i,b) are updated in place;
 * phiAlpha[p] holds the HO flux on entry and alphaPhi on exit. */
void orc_stage(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg, int nPhases,
               double *const *alphai, double *const *alphab, const int32_t *alphaBc,
               const orc_sigma *sigma, const double *Ti, const double *Tb,
               const double *phii, const double *phib,
               double *const *phiAlphai, double *const *phiAlphab,
               const double *const *Su, const double *const *Sp, double deltaT,
               double *Fi, double *Fb);

/* Run `orc_stage` on nDom sub-domains at once, one worker thread per sub-domain with
 * processor-patch halo exchange through shared memory (the CPU baseline of SURVEY.md
 * 8(d): "domain-decomposed on all host cores").  All per-domain pointers are arrays of
 * length nDom.  nStages repeats the stage (alpha carried over, phiAlpha re-read from
 * phiAlpha0).  Returns wall seconds of the stage loop. */
double orc_stage_threads(int nDom, const orc_mesh *meshes, const orc_mules_cfg *cfg, int nPhases,
                         double *const *const *alphai, double *const *const *alphab,
                         const int32_t *const *alphaBc, const orc_sigma *sigma,
                         const double *const *Ti, const double *const *Tb,
                         const double *const *phii, const double *const *phib,
                         double *const *const *phiAlphai, double *const *const *phiAlphab,
                         double deltaT, double *const *Fi, double *const *Fb, int nStages);

/* ---- rest of multiphaseSystem::solveAlphas() ---------------- */
enum { ORC_SCHEME_UPWIND = 0, ORC_SCHEME_LINEAR = 1, ORC_SCHEME_VANLEER = 2 };

/* fvc::flux(flux, alpha, "div(..)") = flux_f * alpha_f(scheme; upwind direction from flux)
 * (multiphaseSystem.C:209-214, 246-251).  gradient for vanLeer: fvc::grad, Gauss linear. */
void orc_conv_flux(const orc_mesh *m, const orc_comm *c, int scheme,
                   const double *fluxi, const double *fluxb,
                   const double *ai, const double *ab, const int32_t *alphaBc,
                   double *outi, double *outb);

/* phaseSystem::nHatf (phaseSystem.C:1251-1272, 1297-1305): nHatfv & Sf on all faces. */
void orc_nHatf(const orc_mesh *m, const orc_comm *c,
               const double *a1i, const double *a1b, const int32_t *bc1,
               const double *a2i, const double *a2b, const int32_t *bc2,
               double *outi, double *outb);

/* Flux assembly for every phase (multiphaseSystem.C:181-274).  cAlpha[p1*nPhases+p2].
 * out: phiAlphaCorr[p] (internal, boundary). */
void orc_flux_assembly(const orc_mesh *m, const orc_comm *c, int nPhases,
                       const double *const *alphai, const double *const *alphab, const int32_t *alphaBc,
                       const double *phii, const double *phib, int schemePhi, int schemePhir,
                       const double *cAlpha, double *const *outi, double *const *outb);

/* Implicit upwind predictor (multiphaseSystem.C:348-363): solves
 *   (V/dt - V Sp + sum_out phi) psi_c + sum_in phi psi_nb = V psi0/dt + V Su
 * by Jacobi sweeps until max|dpsi| <= tol (globally) or maxIter, re-evaluates the boundary values and
 * ADDS alpha1Eqn.flux() = phi*psi_upwind(new) into phiAlpha.  Returns the sweep count. */
int orc_upwind_predictor(const orc_mesh *m, const orc_comm *c, double deltaT,
                         double *psii, double *psib, const double *psi0, const int32_t *psiBc,
                         const double *phii, const double *phib,
                         const double *Su, const double *Sp, double tol, int maxIter,
                         double *phiAlphai, double *phiAlphab);

typedef struct orc_alpha_ctl {
    int32_t nAlphaCorr, nAlphaSubCycles;
    int32_t schemePhi, schemePhir;       /* div(phi,alpha), div(phir,alpha,alpha) */
    int32_t implicitPredictor;           /* 1: :348-363 as coded; 0: the commented-out bounded upwind flux of :346 */
    double predictorTol;
    int32_t predictorMaxIter;
} orc_alpha_ctl;

/* multiphaseSystem::solve() (multiphaseSystem.C:128-168) = nAlphaSubCycles x solveAlphas() (:170-429)
 * incl. renormalisation and rhoPhi.  alpha (i,b) in/out; alpha0 is taken at the start of every
 * sub-cycle; rho[p]: uniform phase density when rhoi[p] == NULL.  Su/Sp NULL = no mass transfer.
 * out: alphaPhi[p], rhoPhi (the Delta-t weighted sum over sub-cycles). */
void orc_solve(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg, const orc_alpha_ctl *ctl,
               int nPhases, double *const *alphai, double *const *alphab, const int32_t *alphaBc,
               const double *phii, const double *phib, const double *cAlpha,
               const double *rhoUniform, const double *const *rhoi, const double *const *rhob,
               const double *const *Su, const double *const *Sp, double deltaT,
               double *const *alphaPhii, double *const *alphaPhib, double *rhoPhii, double *rhoPhib);

/* fvc::flux(volVectorField) with linear interpolation (pEqn.H:31-38 consumes divCapillaryStress() this way). */
void orc_flux_of_vector(const orc_mesh *m, const double *Fi, const double *Fb, double *outi, double *outb);

/* ---- Su/Sp of the phase-fraction equations and porous interface drag -------- */
enum { ORC_VAR_T = 0, ORC_VAR_ALPHA = 1 };      /* interfaceCompositionModel::modelVariable */
/* meltingEvaporationModels::Lee (Lee.C:35-45): one entry of massTransferModel (from to) */
typedef struct orc_lee {
    int32_t from, to;
    double C, Tactivate, alphaMin;
    int32_t modelVariable;     /* ORC_VAR_* */
    int32_t includeDivU;
} orc_lee;
typedef struct orc_mass {
    int32_t nModels; const orc_lee *models;
    int32_t nPairs; const int32_t *pairs;       /* totalPhasePairs in iteration order: (phase1, phase2) */
    const double *rhoUniform;                   /* [nPhases] used when rhoi == NULL or rhoi[p] == NULL */
    const double *const *rhoi;                  /* thermo().rho() per phase, may be NULL */
} orc_mass;
void orc_div_phi(const orc_mesh *m, const double *phii, const double *phib, double *out);
void orc_lee_kexp(int nC, const orc_lee *mdl, const double *alphaFrom, const double *rhoFrom, const double *T, double *out);
int orc_lee_ksu(int nC, const orc_lee *mdl, int variable, const double *alphaFrom, const double *rhoFrom,
                const double *refValue, double *out);
/* MassTransferPhaseSystem::correctMassSources(T) (:447-497): dmdt[i] = Kexp(T) of model i */
void orc_correct_mass_sources(const orc_mesh *m, const orc_mass *mt, const double *const *alphai, const double *T,
                              double *const *dmdt);
/* MassTransferPhaseSystem::alphaTransfer(Su, Sp) (:503-704); Su/Sp are added to; returns ddtAlphaMax */
double orc_alpha_transfer(const orc_mesh *m, const orc_comm *c, const orc_mass *mt, int nPhases,
                          const double *const *alphai, const double *const *dmdt,
                          const double *phii, const double *phib, double *const *Su, double *const *Sp);
/* phaseSystem::addInterfacePorosity (phaseSystem.C:1148-1183) with VollerPrakash::S() (VollerPrakash.C:66-72) */
void orc_interface_porosity(const orc_mesh *m, int nModels, const int32_t *solidPhase, const double *Cu,
                            const double *const *alphai, double *Udiag);
/* orc_solve with calculateSuSp() inside every corrector (multiphaseSystem.C:280-292): Su/Sp zeroed, then
 * alphaTransfer from the current alpha and the stored dmdt.  mt == NULL: identical to orc_solve. */
void orc_solve_mt(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg, const orc_alpha_ctl *ctl,
                  int nPhases, double *const *alphai, double *const *alphab, const int32_t *alphaBc,
                  const double *phii, const double *phib, const double *cAlpha,
                  const double *rhoUniform, const double *const *rhoi, const double *const *rhob,
                  const double *const *Su, const double *const *Sp, double deltaT,
                  double *const *alphaPhii, double *const *alphaPhib, double *rhoPhii, double *rhoPhib,
                  const orc_mass *mt, const double *const *dmdt);

/* MultiComponentPhaseModel::solveYi (MultiComponentPhaseModel.C:182-388): species MULES of phase `ph`.
 * Y (internal, boundary) in/out [nSpecies]; YBc [nSpecies*nPatches]; Su/Sp per species may be NULL.
 * cYi, nYiSubCycles, the two div schemes and the Yi linear-solver tolerance come from the case. */
void orc_solve_yi(const orc_mesh *m, const orc_comm *c, const orc_mules_cfg *cfg, int nPhases,
                  const double *const *alphai, const double *const *alphab, const int32_t *alphaBc, int ph,
                  int nSpecies, int inertIndex, double *const *Yi, double *const *Yb, const int32_t *YBc,
                  const double *phii, const double *phib, double cYi, int schemePhi, int schemePhir,
                  const double *const *Su, const double *const *Sp, double deltaT, double tol, int maxIter,
                  int nYiSubCycles);

/* ---- laserDTRM: discrete-ray laser heat source ---------------------------------------- */
/* polyMesh data the cell walk and interpolationCellPoint need on top of orc_mesh */
typedef struct orc_laser_geom {
    const double *Cf;          /* [nIF*3] mesh.faceCentres(), internal faces  */
    const double *Cfb;         /* [nBF*3] boundary faces                      */
    int32_t nPoints;
    const double *points;      /* [nPoints*3] mesh.points()                   */
    const int32_t *fpStart;    /* [nIF+nBF+1] CSR of mesh.faces(): internal faces, then boundary faces */
    const int32_t *fpPts;
    const int32_t *tetBasePt;  /* [nIF+nBF] mesh.tetBasePtIs()                */
    double boundsMag;          /* mesh.bounds().mag()                         */
} orc_laser_geom;
enum { ORC_REFL_NONE = 0, ORC_REFL_FRESNEL = 1, ORC_REFL_FRESNEL_LASER = 2 };
/* one entry of the reflectionModel table, in the table's iteration order (laserDTRM.C:137-166) */
typedef struct orc_reflection {
    int32_t kind, from, to;
    double nk1[2], nk2[2];     /* Fresnel (Fresnel.C:57-58)   */
    double epsilon;            /* FresnelLaser (FresnelLaser.C:57) */
} orc_reflection;
typedef struct orc_laser_cfg {
    int32_t nModels; const orc_reflection *models;
    double alphaCut;           /* laserDTRM.C:379 (default 0.5)  */
    double sigmaSB;            /* physicoChemical::sigma          */
    int32_t maxParticles;      /* capacity of the cloud incl. reflected particles */
    int32_t localFindCell;     /* 1: a reflected particle whose insertion point lies behind a processor patch is dropped, as the
                                  reference's rank-local mesh().findCell does (DTRMParticle.C:189-206); 0: the neighbour rank
                                  locates it (decomposition-independent result) */
} orc_laser_cfg;
enum { ORC_PD_GAUSSIAN = 0, ORC_PD_UNIFORM = 2, ORC_PD_GAUSSIAN_PEAK = 3 };
/* laserDTRM::initialise() (laserDTRM.C:169-331) without findCell: the polar disk of seeds.  rArea is the unit
 * vector normal to lDir the reference derives from Random(1234) (:186-197); out arrays [nr*nTheta]. */
void orc_laser_seed(const double lPosition[3], const double lDir[3], const double rArea[3], int nr, int nTheta,
                    double focalLaserRadius, double maxTrackLength, int mode, double power, double sigma, double I0,
                    double *p0, double *p1, double *Ip, double *dA);
/* mesh.findCell(p) for convex cells with planar faces (brute force); -1 if outside */
int orc_find_cell(const orc_mesh *m, const orc_laser_geom *g, const double p[3]);
/* localDensityAbsorptionEmission::aCont/eCont/ECont (localDensityAbsorptionEmission.C:86-186):
 * out = sum_i max(alpha_i, 0)*coeff_i over the listed phases */
void orc_local_density(int nC, int nNames, const double *const *alphai, const double *coeff, double *out);
/* laserDTRM::calculate() (laserDTRM.C:542-735) with DTRMParticle::move (DTRMParticle.C:94-266): marks the
 * reflecting cells, walks the particles cell to cell, deposits Q and divides by V.  Single domain (a particle
 * reaching a processor patch returns 2; cloud overflow returns 3).  Optional outputs may be NULL. */
int orc_laser_calculate(const orc_mesh *m, const orc_comm *c, const orc_laser_geom *g, const orc_laser_cfg *cfg,
                        int nPhases, const double *const *alphai, const double *const *alphab, const int32_t *alphaBc,
                        const double *a, const double *e, const double *E, const double *T,
                        int nParticles, const double *p0, const double *p1, const double *I, const double *dA,
                        const int32_t *cell, double *Q, int32_t *reflectingCells, double *nHat, double *nHatPoints,
                        int32_t *nParticlesFinal);

/* The decomposed run restated on one CPU: all ranks in lock-step rounds, particles handed over at processor
 * patches, point values summed over the ranks sharing a point (pointGlobal: global point labels per rank).
 * refl / nHat per rank are the marking of laserDTRM.C:596-652 for the rank's cells; created [nRanks] out. */
int orc_laser_calculate_ranks(int nRanks, const orc_mesh *const *ms, const orc_laser_geom *const *gs, const orc_laser_cfg *cfg,
                              const int32_t *const *refl, const double *const *nHat, const int64_t *const *pointGlobal,
                              const double *const *a, const double *const *e, const double *const *E, const double *const *T,
                              int nParticles, const double *p0, const double *p1, const double *I, const double *dA,
                              const int32_t *const *cell, double *const *Q, int32_t *created, double *const *nHatPointsOut);

#ifdef __cplusplus
}
#endif
#endif
