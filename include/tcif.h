/*
 * tcif.h -- C ABI of the B200-native interface hot path of thermocapillaryInterFoam.
 *
 * The reference has no FFI for this path: the boundary is two C++ member functions of
 * libincompressibleMultiphaseSystems.so (phasesSystem/Make/files:28)
 *     tmp<volVectorField> phaseSystem::divCapillaryStress() const   phaseSystem.H:493-494,
 *                                                                    phaseSystem.C:1038-1105
 *     void multiphaseSystem::solveAlphas()                           multiphaseSystem.H:93-94,
 *                                                                    multiphaseSystem.C:170-429
 * plus the OpenFOAM-v2106 MULES entry points solveAlphas() calls (multiphaseSystem.C:305-317,
 * 321, 365-375).  Every entry point below names the reference interface it replaces; the
 * OpenFOAM-typed binding a maintainer adds is shown in INTEGRATION.md / foam_shim/.
 *
 * Conventions = OpenFOAM's: scalar = double (WM_DP), label = int32 (WM_LABEL_SIZE=32),
 * vectors AoS (x,y,z) exactly as Field<vector>::cdata() lays them out.  A vol field is
 * internal[nCells] + boundary[nBoundaryFaces] (patch arrays concatenated in patch order,
 * `empty` patches have size 0); a surface field is internal[nInternalFaces] + boundary[..].
 * On processor patches the boundary value of a vol field is the neighbour-rank cell value
 * (processorFvPatchField, = patchNeighbourField()).  Host pointers are borrowed for the
 * duration of a call only.  All functions return 0 on success, non-zero on error with a
 * message in tcif_last_error(); there is NO CPU fallback: without a CUDA device
 * tcif_create() fails.  One context per rank/GPU; calls on one context are not re-entrant.
 */
#ifndef TCIF_H
#define TCIF_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TCIF_API __attribute__((visibility("default")))

#define TCIF_MAX_PHASES 4
#define TCIF_MAX_PAIRS 6
#define TCIF_SIGMA_MAX_COEFFS 64

/* fvPatch kinds the path distinguishes (coupled = processor). */
enum { TCIF_PATCH_GENERIC = 0, TCIF_PATCH_EMPTY = 1, TCIF_PATCH_WEDGE = 2, TCIF_PATCH_PROCESSOR = 3 };
/* alpha boundary conditions on non-coupled patches: zeroGradient | fixedValue (fixesValue()). */
enum { TCIF_BC_ZERO_GRADIENT = 0, TCIF_BC_FIXED_VALUE = 1 };
/* surfaceTensionModel: constant | temperatureDependent with Function1 polynomial / table. */
enum { TCIF_SIGMA_NONE = -1, TCIF_SIGMA_CONSTANT = 0, TCIF_SIGMA_POLYNOMIAL = 1, TCIF_SIGMA_TABLE = 2 };
/* MULES limiter: extrema contributed by non-coupled, non-fixed-value patches. */
enum { TCIF_BEXT_GLOBAL = 0, TCIF_BEXT_PATCHVALUE = 1 };

typedef struct tcif_ctx tcif_ctx;
typedef struct tcif_group tcif_group;

/* What the shim reads off an fvMesh (lduAddr().lowerAddr()/upperAddr(), Sf(), weights(),
 * V(), boundary()[p].faceCells()/Sf()/magSf()/deltaCoeffs()/weights()). */
typedef struct tcif_mesh_desc {
    int32_t nCells, nInternalFaces, nPatches, nBoundaryFaces;
    const int32_t *owner, *neighbour;      /* [nInternalFaces], upper-triangular order */
    const double *Sf;                      /* [nInternalFaces*3]                       */
    const double *weights;                 /* [nInternalFaces]                         */
    const double *V;                       /* [nCells]                                 */
    const int32_t *patchType;              /* [nPatches] TCIF_PATCH_*                  */
    const int32_t *patchStart;             /* [nPatches] offset into boundary arrays   */
    const int32_t *patchSize;              /* [nPatches] (0 for empty)                 */
    const int32_t *patchNeighbProcNo;      /* [nPatches] (-1 unless processor)         */
    const int32_t *faceCells;              /* [nBoundaryFaces]                         */
    const double *SfBoundary;              /* [nBoundaryFaces*3]                       */
    const double *magSfBoundary;           /* [nBoundaryFaces]                         */
    const double *deltaCoeffsBoundary;     /* [nBoundaryFaces]                         */
    const double *weightsBoundary;         /* [nBoundaryFaces] (1 on non-coupled)      */
} tcif_mesh_desc;

/* fvSolution solver-dict keys MULES reads (nLimiterIter, smoothLimiter, extremaCoeff) +
 * build-independent knobs. */
typedef struct tcif_config {
    int32_t nPhases;            /* phases in HashTable iteration order     */
    int32_t nLimiterIter;       /* default 3                                           */
    double smoothLimiter;       /* default 0                                           */
    double extremaCoeff;        /* default 0                                           */
    int32_t boundaryExtrema;    /* TCIF_BEXT_GLOBAL                                    */
    int32_t device;             /* CUDA ordinal, -1 = current device                   */
    int32_t flags;              /* reserved, 0                                         */
} tcif_config;

/* ---- life cycle -------------------------------------------------------------------- */
/* Replaces the mesh/field references phaseSystem holds (phaseSystem.C:207-290): uploads
 * addressing + geometry once, builds the cell-ordered face lists on the device. */
TCIF_API int tcif_create(const tcif_mesh_desc *mesh, const tcif_config *cfg, tcif_ctx **out);
TCIF_API void tcif_destroy(tcif_ctx *ctx);
/* Message of the last failed call (ctx may be NULL for tcif_create failures). */
TCIF_API const char *tcif_last_error(const tcif_ctx *ctx);
TCIF_API const char *tcif_version(void);
/* Number of visible CUDA devices (0 => every tcif_create fails). */
TCIF_API int tcif_device_count(void);
/* Run the context's kernels on a caller-owned cudaStream_t (NULL = context's own stream). */
TCIF_API int tcif_set_stream(tcif_ctx *ctx, void *cudaStream);
TCIF_API int tcif_synchronize(tcif_ctx *ctx);
/* Pinned host allocation for shim-side staging buffers (cudaHostAlloc / cudaFreeHost). */
TCIF_API void *tcif_host_alloc(uint64_t bytes);
TCIF_API void tcif_host_free(void *p);

/* lduAddressing::ownerStartAddr / losortAddr / losortStartAddr as built for the device
 * (bit-exact check of the cell-ordered face lists against OpenFOAM's own). */
TCIF_API int tcif_get_ldu_csr(tcif_ctx *ctx, int32_t *ownerStart, int32_t *losort, int32_t *losortStart);

/* ---- processor-patch communication (Pstream stand-in) ------------------------------ */
/* NCCL over NVLink: rank 0 makes the id, the host plumbing (torch.distributed / MPI)
 * broadcasts its 128 bytes, every rank joins.  patchNeighbProcNo are NCCL ranks. */
TCIF_API int tcif_comm_unique_id(void *id128);
TCIF_API int tcif_comm_init_nccl(tcif_ctx *ctx, int rank, int nRanks, const void *id128);
/* In-process group (several contexts on one GPU, one host thread each; used by tests). */
TCIF_API int tcif_group_create(int nRanks, tcif_group **out);
TCIF_API void tcif_group_destroy(tcif_group *g);
TCIF_API int tcif_comm_init_local(tcif_ctx *ctx, int rank, tcif_group *g);

/* ---- fields (device-resident cache of the host fields) ------------------------------ */
/* alpha.<phase> (phaseModel.C:41-62): boundary-condition kinds per patch, values. */
TCIF_API int tcif_set_alpha_bc(tcif_ctx *ctx, int phase, const int32_t *bcKindPerPatch);
TCIF_API int tcif_set_alpha(tcif_ctx *ctx, int phase, const double *internal, const double *boundary);
TCIF_API int tcif_get_alpha(tcif_ctx *ctx, int phase, double *internal, double *boundary);
/* psi.oldTime(): copy alpha -> alpha0 for every phase (start of a time step / sub-cycle). */
TCIF_API int tcif_store_old_alphas(tcif_ctx *ctx);
TCIF_API int tcif_set_alpha_old(tcif_ctx *ctx, int phase, const double *internal);
/* T looked up by temperatureDependent (name from dict entry "T"). */
TCIF_API int tcif_set_temperature(tcif_ctx *ctx, const double *internal, const double *boundary);
/* surfaceTension ((a and b) {type constant|temperatureDependent; sigma ...}) entry
 * (phaseSystem.C:266-273): constant: coeffs[0]; polynomial: n (a,e) pairs; table: n (T,s). */
TCIF_API int tcif_set_sigma(tcif_ctx *ctx, int phase1, int phase2, int kind, int n, const double *coeffs);
/* phi (phaseSystem.C:227-239). */
TCIF_API int tcif_set_phi(tcif_ctx *ctx, const double *internal, const double *boundary);
/* phiAlphaCorrs[phase] / phiAlpha / phase.alphaPhi() (multiphaseSystem.C:203-216, 363, 377). */
TCIF_API int tcif_set_phi_alpha(tcif_ctx *ctx, int phase, const double *internal, const double *boundary);
TCIF_API int tcif_get_phi_alpha(tcif_ctx *ctx, int phase, double *internal, double *boundary);
/* Su_/Sp_ (multiphaseSystem.C:279-292); NULL = zero field. */
TCIF_API int tcif_set_su_sp(tcif_ctx *ctx, int phase, const double *Su, const double *Sp);

/* ---- the hot path ------------------------------------------------------------------- */
/* phaseSystem::divCapillaryStress() (phaseSystem.C:1038-1105) on the resident alpha, T. */
TCIF_API int tcif_div_capillary_stress(tcif_ctx *ctx);
/* Result as volVectorField: internal [nCells*3] + patch values [nBoundaryFaces*3]. */
TCIF_API int tcif_get_force(tcif_ctx *ctx, double *internal, double *boundary);
/* One-call drop-in with host buffers (upload alpha[P], T; compute; download). */
TCIF_API int tcif_div_capillary_stress_host(tcif_ctx *ctx, const double *const *alphaInternal,
                                            const double *const *alphaBoundary, const double *Tinternal,
                                            const double *Tboundary, double *Finternal, double *Fboundary);
/* fvc::flux(fluid.divCapillaryStress()) = linear(F) & Sf, the form pEqn.H:31-38 consumes: [nInternalFaces],
 * [nBoundaryFaces].  The force itself is re-evaluated only when alpha, T or sigma changed since the last
 * evaluation: the solver asks for it 2-4 times per time step on unchanged fields (UEqn.H:29, pEqn.H:35 per
 * corrector, thermocapillaryInterFoam.C:150) and the reference recomputes it from scratch every time. */
TCIF_API int tcif_flux_of_force(tcif_ctx *ctx, double *internal, double *boundary);
/* Switch that cache off/on (default on) and read how often F was computed vs served from the cache. */
/* ---- residency: keep fields on the device between calls ---------------------------------------------------
 * The solver evaluates divCapillaryStress() 3-5 times per time step (UEqn.H:29, pEqn.H:35 per corrector,
 * thermocapillaryInterFoam.C:150) and solve() once (:105), always on the SAME registered fields alpha.<phase>
 * and T.  The device copy of a field carries the caller's modification stamp (OpenFOAM: regIOobject::eventNo()
 * of the field, which every non-const access advances); a stamped set uploads -- and invalidates the cached
 * force -- only when the stamp differs from the one the device copy carries.  stamp 0 = "unknown": always
 * uploads.  Whatever changes alpha on the device (explicitSolve, the predictor, renormalisation, an unstamped
 * set) resets the stamp to 0; after tcif_get_alpha the caller tells the device which host state now equals its
 * copy with tcif_mark_alpha.  *uploaded (may be NULL) reports whether bytes moved. */
TCIF_API int tcif_set_alpha_stamped(tcif_ctx *ctx, int phase, uint64_t stamp, const double *internal,
                                    const double *boundary, int32_t *uploaded);
TCIF_API int tcif_set_temperature_stamped(tcif_ctx *ctx, uint64_t stamp, const double *internal,
                                          const double *boundary, int32_t *uploaded);
TCIF_API int tcif_mark_alpha(tcif_ctx *ctx, int phase, uint64_t stamp);
/* uploads[0] alpha fields uploaded, [1] alpha uploads skipped (stamp matched), [2] T uploaded, [3] T skipped */
TCIF_API int tcif_transfer_counters(const tcif_ctx *ctx, int64_t uploads[4]);
TCIF_API int tcif_set_force_cache(tcif_ctx *ctx, int on);
TCIF_API int tcif_force_counters(const tcif_ctx *ctx, int64_t *evaluations, int64_t *cacheHits);
/* fvc::grad(alpha.<phase>) with Gauss linear (phaseSystem.C:1076-1077): [nCells*3], [nBF*3]. */
TCIF_API int tcif_grad_alpha(tcif_ctx *ctx, int phase, double *internal, double *boundary);

/* MULES::limit(rDeltaT, 1, alpha, phi, phiAlpha, Sp, Su, 1, 0, returnCorr)  (multiphaseSystem.C:305-317). */
TCIF_API int tcif_mules_limit(tcif_ctx *ctx, int phase, double rDeltaT, int returnCorr);
/* MULES::limitSum(phiAlphaCorrs)  (multiphaseSystem.C:321). */
TCIF_API int tcif_mules_limit_sum(tcif_ctx *ctx);
/* phiAlpha += upwind<scalar>(mesh, phi).flux(alpha)  (multiphaseSystem.C:346, the bounded
 * upwind flux; with the implicit predictor it is alpha1Eqn.flux(), :363). */
TCIF_API int tcif_add_upwind_flux(tcif_ctx *ctx, int phase);
/* MULES::explicitSolve(1, alpha, phi, phiAlpha, Sp, Su, 1, 0)  (multiphaseSystem.C:365-375). */
TCIF_API int tcif_mules_explicit_solve(tcif_ctx *ctx, int phase, double deltaT);

/* One metric stage on resident fields: divCapillaryStress over all pairs +
 * one solveAlphas corrector over all phases (limit(returnCorr) x P, limitSum, per phase
 * phiAlpha += upwind flux; explicitSolve).  alpha0 := alpha at entry. */
TCIF_API int tcif_stage(tcif_ctx *ctx, double deltaT);
/* Same through host buffers, pipelined: uploads (alpha[P], T first; phi, phiAlphaCorr[P] behind them), the
 * force, the download of F, the MULES corrector and the downloads of alpha[P], alphaPhi[P] overlap on
 * separate copy streams.  alpha is in/out; phiAlphaCorr (multiphaseSystem.C:203-216) is read-only and
 * alphaPhi (phase.alphaPhi(), :377) a separate output, as in the reference.  Use pinned buffers
 * (tcif_host_alloc) for the copies to be asynchronous.  Boundary arrays may be NULL when nBoundaryFaces == 0. */
TCIF_API int tcif_stage_host(tcif_ctx *ctx, double deltaT,
                             double *const *alphaInternal, double *const *alphaBoundary,
                             const double *Tinternal, const double *Tboundary,
                             const double *phiInternal, const double *phiBoundary,
                             const double *const *phiAlphaCorrInternal, const double *const *phiAlphaCorrBoundary,
                             double *const *alphaPhiInternal, double *const *alphaPhiBoundary,
                             double *Finternal, double *Fboundary);

/* ---- what the host-buffer entry points (tcif_stage_host, tcif_time_step_host) move over PCIe ----------------
 * One-shot stamps for the NEXT *_host call: in[p] = stamp of the state the caller's alpha array p holds (skips the
 * upload when the device copy carries the same non-zero stamp), out[p] = stamp the device assigns to its alpha when
 * the call returns (the caller's arrays then hold the same values).  A time-stepping caller whose alpha is only
 * ever changed by solve() passes out of step k as in of step k+1 and never uploads alpha again.  NULL = zeros. */
TCIF_API int tcif_host_alpha_stamps(tcif_ctx *ctx, const uint64_t *in, const uint64_t *out);
/* on: alpha, F and alphaPhi come back as DELTAS -- the entries whose value differs from what the caller's array
 * already holds (alpha: the values passed in; F, alphaPhi: what the previous call wrote to the same array) are
 * compacted on the device, copied as (index, value) and scattered into the array by the calling thread (a few host
 * threads for long lists); a field with more than a quarter of its entries changed is copied densely.  F and the
 * fractions and fluxes of dispersed phases are exactly zero away from the interfaces, so most of their bytes never
 * cross the link.  The caller must leave the F and alphaPhi arrays as the previous call returned them.  The arrays
 * equal those of the dense path entry by entry (compared by value: the sign of a zero may differ). */
TCIF_API int tcif_host_sparse_returns(tcif_ctx *ctx, int on);
/* bytes the last *_host call really copied host->device and device->host */
TCIF_API int tcif_host_bytes_moved(const tcif_ctx *ctx, int64_t *h2d, int64_t *d2h);

/* ---- rest of multiphaseSystem::solveAlphas() on the device --------- */
/* convection schemes of fvSchemes' div(phi,alpha) / div(phir,alpha,alpha) the device path supports */
enum { TCIF_SCHEME_UPWIND = 0, TCIF_SCHEME_LINEAR = 1, TCIF_SCHEME_VANLEER = 2 };
/* fvSolution "alpha" controls read by solve()/solveAlphas() (multiphaseSystem.C:130-131, 173-175) */
typedef struct tcif_alpha_controls {
    int32_t nAlphaCorr;          /* >= 1                                                  */
    int32_t nAlphaSubCycles;     /* <= 1: no sub-cycling                                  */
    int32_t schemePhi;           /* TCIF_SCHEME_* for div(phi,alpha.x)                    */
    int32_t schemePhir;          /* TCIF_SCHEME_* for div(phir,alpha.y,alpha.x)           */
    int32_t implicitPredictor;   /* 1: alpha1Eqn.solve() + .flux() as coded (:348-363);
                                    0: the bounded explicit upwind flux of :346          */
    double predictorTol;         /* max|dalpha| between Jacobi sweeps at convergence      */
    int32_t predictorMaxIter;
} tcif_alpha_controls;
/* mesh.C() [nCells*3] and fvPatch::delta() [nBoundaryFaces*3] (own -> neighbour centre on processor
 * faces): geometry the NVD/TVD limiter of vanLeer needs; not required for upwind/linear. */
TCIF_API int tcif_set_cell_centres(tcif_ctx *ctx, const double *C, const double *deltaBoundary);
/* cAlphas ((a and b) c ...) as a dense [nPhases*nPhases] table, entry [p1*nPhases+p2] (multiphaseSystem.C:226-237). */
TCIF_API int tcif_set_calpha(tcif_ctx *ctx, const double *cAlpha);
/* phase.rho() for rhoPhi_ (multiphaseSystem.C:397-406): a uniform value, or a vol field when internal != NULL. */
TCIF_API int tcif_set_rho(tcif_ctx *ctx, int phase, double uniform, const double *internal, const double *boundary);
/* phiAlphaCorrs[p] = fvc::flux(phi, alpha_p, schemePhi) + compression fluxes with nHatf, inflow fix, for all phases
 * (multiphaseSystem.C:199-277) from the resident alpha, phi; result readable with tcif_get_phi_alpha after
 * tcif_mules_limit or through tcif_get_flux_assembly. */
TCIF_API int tcif_flux_assembly(tcif_ctx *ctx, int schemePhi, int schemePhir);
TCIF_API int tcif_get_flux_assembly(tcif_ctx *ctx, int phase, double *internal, double *boundary);
/* alpha1Eqn.solve(); phiAlpha += alpha1Eqn.flux()   (multiphaseSystem.C:348-363). */
TCIF_API int tcif_upwind_predictor(tcif_ctx *ctx, int phase, double deltaT, double tol, int maxIter);
TCIF_API int tcif_predictor_sweeps(const tcif_ctx *ctx, int phase);
/* multiphaseSystem::solve() (multiphaseSystem.C:128-168): sub-cycles x correctors of flux assembly, limit,
 * limitSum, predictor, explicitSolve, rhoPhi, renormalisation, on the resident alpha, phi, Su/Sp. */
TCIF_API int tcif_solve_alphas(tcif_ctx *ctx, const tcif_alpha_controls *ctl, double deltaT);
TCIF_API int tcif_get_rho_phi(tcif_ctx *ctx, double *internal, double *boundary);
/* The reference's call surface in one pipelined host-buffer call: fluid.solve() (thermocapillaryInterFoam.C:105)
 * followed by fluid.divCapillaryStress() on the new alpha (UEqn.H:29).  alpha in/out; T, phi in; rhoPhi, F out;
 * alphaPhi optional (NULL = not downloaded).  Su/Sp, rho, cAlpha, sigma as set on the context. */
TCIF_API int tcif_time_step_host(tcif_ctx *ctx, const tcif_alpha_controls *ctl, double deltaT,
                                 double *const *alphaInternal, double *const *alphaBoundary,
                                 const double *Tinternal, const double *Tboundary,
                                 const double *phiInternal, const double *phiBoundary,
                                 double *const *alphaPhiInternal, double *const *alphaPhiBoundary,
                                 double *rhoPhiInternal, double *rhoPhiBoundary,
                                 double *Finternal, double *Fboundary);

/* ---- Su/Sp of the phase-fraction equations on the device ----------------------------
 * Replaces MassTransferPhaseSystem::alphaTransfer(Su, Sp) (MassTransferPhaseSystem.C:503-704, called through
 * multiphaseSystem::calculateSuSp(), multiphaseSystem.C:122-125, 292), correctMassSources(T) (:447-497) with
 * the Lee model (massTransferModels/Lee/Lee.C:52-167), and phaseSystem::addInterfacePorosity with
 * VollerPrakash::S() (phaseSystem.C:1148-1183, VollerPrakash.C:66-72).  Needs tcif_set_rho for every phase
 * (coeffs = 1/rho, phaseSystem.C:1139-1145). */
enum { TCIF_VAR_T = 0, TCIF_VAR_ALPHA = 1 };          /* interfaceCompositionModel::modelVariable */
typedef struct tcif_lee_model {                        /* massTransferModel ((from to) {type Lee; ...}) */
    int32_t from, to;                                  /* phase indices of the ordered pair key */
    double C, Tactivate, alphaMin;                     /* Lee.C:39-42 */
    int32_t modelVariable;                             /* TCIF_VAR_*: KSu(alpha, .) is valid only for TCIF_VAR_ALPHA */
    int32_t includeDivU;                               /* interfaceCompositionModel::includeDivU() */
} tcif_lee_model;
TCIF_API int tcif_mass_transfer_clear(tcif_ctx *ctx);
TCIF_API int tcif_add_mass_transfer_lee(tcif_ctx *ctx, const tcif_lee_model *model);
/* totalPhasePairs() in its HashTable iteration order as (phase1, phase2) index pairs; default: all i<j. */
TCIF_API int tcif_set_phase_pairs(tcif_ctx *ctx, int nPairs, const int32_t *pairs);
/* dmdt_[from to] as computed by the host (any model); NULL = zero field.  [nCells] */
TCIF_API int tcif_set_dmdt(tcif_ctx *ctx, int from, int to, const double *internal);
TCIF_API int tcif_get_dmdt(tcif_ctx *ctx, int from, int to, double *internal);
/* correctMassSources(T): dmdt_[key] = Kexp(T) for every registered model, from the resident alpha, rho, T. */
TCIF_API int tcif_correct_mass_sources(tcif_ctx *ctx);
/* Su_ = Sp_ = 0; alphaTransfer(Su_, Sp_) from the resident alpha, phi, rho, dmdt: the result becomes the Su/Sp
 * of the MULES entry points (as tcif_set_su_sp would).  ddtAlphaMax (:700-702) optional. */
TCIF_API int tcif_alpha_transfer(tcif_ctx *ctx, double *ddtAlphaMax);
/* on: tcif_solve_alphas / tcif_time_step_host call alphaTransfer inside every corrector where solveAlphas
 * calls calculateSuSp() (multiphaseSystem.C:280-292). */
TCIF_API int tcif_set_mass_transfer_in_solve(tcif_ctx *ctx, int on);
TCIF_API int tcif_get_su_sp(tcif_ctx *ctx, int phase, double *Su, double *Sp);
/* Udiag[nCells] += V * sum_models Cu*sqr(alpha_s)/(pow3(1-alpha_s)+1e-3)   (UEqn diagonal, host buffer in/out). */
TCIF_API int tcif_interface_porosity(tcif_ctx *ctx, int nModels, const int32_t *solidPhase, const double *Cu, double *Udiag);

/* MultiComponentPhaseModel::solveYi (phaseModel/MultiComponentPhaseModel/MultiComponentPhaseModel.C:182-388): the
 * species MULES of one phase on the device -- flux assembly with interface compression (:199-297), MULES::limit
 * (:299-311), implicit upwind predictor + explicitSolve (:321-378), inert species = max(1 - sum Yi, 0) (:383-384).
 * Uses the resident alpha (all phases), phi, mesh geometry; Yinternal[i] [nCells] / Yboundary[i] [nBoundaryFaces]
 * are host buffers in/out, Ybc [nSpecies*nPatches] the TCIF_BC_* kind per patch, Su/Sp[i] may be NULL (the
 * massSpeciesTransfer sources); cYi, the two div schemes, nYiSubCycles and the Yi solver tolerance are the
 * case's fvSolution/fvSchemes entries.  Overwrites nothing of the phase-fraction state. */
TCIF_API int tcif_solve_yi(tcif_ctx *ctx, int phase, int nSpecies, int inertIndex,
                           double *const *Yinternal, double *const *Yboundary, const int32_t *Ybc,
                           const double *const *Su, const double *const *Sp, double cYi,
                           int schemePhi, int schemePhir, int nYiSubCycles, double deltaT, double tol, int maxIter);

/* ---- laserDTRM: the discrete-ray laser heat source -----------------------------------------
 * Replaces radiation::laserDTRM::calculate() (laserDTRM/laserDTRM.C:542-735) with DTRMParticle::move
 * (laserDTRM/DTRMParticle/DTRMParticle.C:94-266): reflecting-cell marking from grad(alpha) (:596-652), the cell
 * walk of the cloud with absorption/emission per cell crossed and Fresnel reflection at the interface, Q /= V.
 * The seeds (initialise(), :169-331: polar disk, Function1s, findCell) stay with the solver. */
/* polyMesh data of the cell walk and of interpolationCellPoint(nHat): mesh.faceCentres() (internal [nIF*3],
 * boundary [nBF*3]), mesh.points(), mesh.faces() as CSR over internal then boundary faces, mesh.tetBasePtIs(),
 * mesh.bounds().mag().  Needs tcif_set_cell_centres first. */
TCIF_API int tcif_laser_set_polymesh(tcif_ctx *ctx, const double *Cf, const double *Cfb, int32_t nPoints,
                                     const double *points, const int32_t *fpStart, const int32_t *fpPts,
                                     const int32_t *tetBasePt, double boundsMag);
enum { TCIF_REFL_NONE = 0, TCIF_REFL_FRESNEL = 1, TCIF_REFL_FRESNEL_LASER = 2 };
/* one entry of the `reflectionModel` table (laserDTRM.C:137-166) in the table's iteration order:
 * noReflection (noReflection.C:58-73), Fresnel (Fresnel.C:57-58: nk1, nk2), FresnelLaser (FresnelLaser.C:57: epsilon) */
typedef struct tcif_reflection_model {
    int32_t kind, from, to;
    double nk1[2], nk2[2], epsilon;
} tcif_reflection_model;
TCIF_API int tcif_laser_set_reflection(tcif_ctx *ctx, int32_t nModels, const tcif_reflection_model *models, double alphaCut);
/* localDensityAbsorptionEmission (localDensityAbsorptionEmission.C:86-186): a/e/E = sum_i max(alpha_i,0)*coeff_i */
TCIF_API int tcif_laser_set_local_density_absorption(tcif_ctx *ctx, int32_t nNames, const int32_t *phases,
                                                     const double *aCoeff, const double *eCoeff, const double *ECoeff);
/* a(), e(), E() of any other absorptionEmissionModel as host fields [nCells] (laserDTRM.C:583-585) */
TCIF_API int tcif_laser_set_absorption_fields(tcif_ctx *ctx, const double *a, const double *e, const double *E);
/* calculate(): seeds p0/p1 [n*3], Ip, dA [n], cell [n] (-1: findCell missed, not seeded); sigmaSB =
 * physicoChemical::sigma; maxParticles = capacity of the cloud incl. reflected particles (0: 64 per seed).
 * Uses the resident alpha and T.  Q [nCells] (= Ru(), W/m^3) may be NULL (read it with tcif_laser_get);
 * nParticlesFinal = particles created on this rank (its seeds + the reflected particles it spawned). */
TCIF_API int tcif_laser_calculate(tcif_ctx *ctx, int32_t nParticles, const double *p0, const double *p1, const double *I,
                                  const double *dA, const int32_t *cell, double sigmaSB, int32_t maxParticles,
                                  double *Q, int32_t *nParticlesFinal);
/* Q, reflectingCellsVol (:556), nHat (:576) [nCells*3] and its point interpolation [nPoints*3]; any may be NULL */
TCIF_API int tcif_laser_get(tcif_ctx *ctx, double *Q, int32_t *reflectingCells, double *nHat, double *nHatPoints);
/* Q of the last calculate() as (cell, value) pairs of the cells that hold energy, unordered: the beam touches a small
 * part of the mesh, the solver zeroes Q_ anyway (laserDTRM.C:580) and scatters these instead of reading nCells
 * doubles back.  *n = number of such cells; fails (with *n set) when capacity is too small. */
TCIF_API int tcif_laser_get_sparse(tcif_ctx *ctx, int32_t capacity, int32_t *cells, double *values, int32_t *n);
/* Decomposed meshes: the points of processor patches are shared by ranks, and volPointInterpolation sums weights and
 * weighted values over all ranks holding a point.  localPoint [n]: every local point lying on a processor patch,
 * globalPoint [n]: its global label (globalMeshData).  Collective over the communicator; call after
 * tcif_laser_set_polymesh.  Particles reaching a processor patch are then handed to the neighbour rank between the
 * launches of the walk (hitProcessorPatch, DTRMParticle.C:268-275); tcif_laser_calculate is collective, every rank
 * passes the seeds whose cell it owns (cell = -1 for the others) and gets back Q of its own cells. */
TCIF_API int tcif_laser_set_shared_points(tcif_ctx *ctx, int32_t nShared, const int32_t *localPoint, const int64_t *globalPoint);
/* launches of the walk kernel in the last calculate() = depth of the reflection cascade + 1 */
TCIF_API int tcif_laser_generations(const tcif_ctx *ctx);
/* Decomposed meshes only.  on = 1: a reflected particle whose insertion point (0.1 cbrt(V) behind the hit point) lies
 * behind a processor patch is DROPPED, exactly as the reference's rank-local mesh().findCell(insertP) does
 * (DTRMParticle.C:189-206) -- for comparing a decomposed device run with a decomposed run of the reference.
 * Default 0: the neighbour rank locates the particle, so that the result does not depend on the decomposition. */
TCIF_API int tcif_laser_set_local_find_cell(tcif_ctx *ctx, int on);

/* Logging scalars of multiphaseSystem.C:408-425: sum(V*alpha), min, max (global over ranks). */
TCIF_API int tcif_alpha_stats(tcif_ctx *ctx, int phase, double out3[3]);
/* The same three scalars for the cells of THIS context (rank-local: no collective; on a decomposed mesh combine the
 * ranks' triples yourself), computed without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Calls no cudaMalloc / cudaFree / allocating library routine: the scratch of the
 * reduction is a file-scope `static __device__` array of the library (24.6 KB held by every GPU that loads the library
 * until exit), the result travels through the context's pinned buffer.  Inputs: the context's resident alpha[phase]
 * and cell volumes; output: out3 (host).  Bit-identical to tcif_alpha_stats on an undecomposed mesh (same reduction
 * tree).  Fails with a clear error when the context holds more than TCIF_ALPHA_STATS_STATIC_MAX_CELLS cells.
 * The first call on a device resolves the array's device address (cudaGetSymbolAddress): make it outside a capture.
 * All calls on one GPU share the array, which keeps the last call's contents (nothing needs to start at zero: every
 * slot that is read has been overwritten by the same call); calls that OVERLAP on one GPU -- other streams, contexts
 * or host threads -- race on it and must be serialised by the caller. */
#define TCIF_ALPHA_STATS_STATIC_MAX_CELLS (1 << 27)   /* 134 M cells: more than one 180 GB GPU holds of this path's fields */
TCIF_API int tcif_alpha_stats_static(tcif_ctx *ctx, int phase, double out3[3]);

/* ---- measurement -------------------------------------------------------------------- */
/* Per-kernel CUDA-event timers on the launching stream.  ids are stable (tcif_timer_name). */
TCIF_API int tcif_timers_enable(tcif_ctx *ctx, int on);
TCIF_API int tcif_timers_reset(tcif_ctx *ctx);
TCIF_API int tcif_timer_count(void);
TCIF_API const char *tcif_timer_name(int id);
/* ms[id] = accumulated device ms, launches[id] = launch count, bytes[id] = algorithmic bytes
 * accumulated over those launches (DESIGN.md states the per-cell/face figures). */
TCIF_API int tcif_timers_read(tcif_ctx *ctx, double *ms, int64_t *launches, double *bytes);
/* Kernel launches issued by this context since creation (gpu_launches evidence). */
TCIF_API int64_t tcif_launch_count(const tcif_ctx *ctx);
/* Bytes of HBM the context holds. */
TCIF_API int64_t tcif_device_bytes(const tcif_ctx *ctx);
/* Debug aid (compute-sanitizer is not available on every pool): a context created with TCIF_GUARD=1 in the environment
 * places every device array between two 512-byte guard zones; this counts the zones and those whose pattern has been
 * overwritten (an out-of-bounds write next to an array).  Fails on a context created without the switch. */
TCIF_API int tcif_debug_guard_check(tcif_ctx *ctx, int64_t *nZones, int64_t *nCorrupt);
/* How the mesh was laid out on the device (tests assert the benchmarked code path with it):
 * out[0] Ko, [1] Kn (slot counts), [2] compile-time slot count of the fixed-K sweeps (0: generic-K kernels),
 * [3] 1 when the volume sweeps walk the 256-cell tiles strip-wise, [4] L2 prefetch distance in tiles,
 * [5] cells with patch faces, [6] of those deferred to the O(surface) launches, [7] processor faces. */
TCIF_API int tcif_mesh_info(const tcif_ctx *ctx, int32_t out[8]);
/* How processor-patch values travel: 0 no communicator, 1 in-process group (device-to-device copies), 2 NCCL
 * send/recv (pack, ncclSend/ncclRecv, unpack), 3 peer memory (the producer stores straight into the neighbour rank's
 * mapped receive buffer over NVLink and releases a flag; default for NCCL ranks on one node, TCIF_PEER_HALO=0
 * selects 2). */
TCIF_API int tcif_comm_transport(const tcif_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif
