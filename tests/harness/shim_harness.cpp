// shim_harness.cpp -- a COMPILED stand-in for foam_shim/: drives the C ABI (include/tcif.h) from C++ with
// OpenFOAM-shaped containers, the way tcifFoamBridge.H / phaseSystemDevice.C / multiphaseSystemDeviceSolve.C do.
//
// OpenFOAM is not installed here, so the shim itself cannot meet a compiler; this file keeps the part of it
// that does not need OpenFOAM honest:
//   * AoS vector fields (Field<vector> = contiguous {x,y,z}) passed by reinterpret_cast, as the shim does;
//   * boundary fields stored PER PATCH and flattened in patch order with `empty` patches skipped (they own faces
//     in the polyMesh but no field values), offsets = tcifFoamBridge::patchStart_/patchSize_;
//   * stamped transfers (regIOobject::eventNo() stand-in), tcif_mark_alpha after a download;
//   * every return code checked, tcif_last_error() on failure (FatalError stand-in: exit code 1).
// It replays createFields.H:168-179 + one time step of thermocapillaryInterFoam.C:104-150 and checks the results
// against the CPU oracle (test infrastructure, oracle/tcif_oracle.h) and the residency counters.
//
// usage: shim_harness <config 1|5> <n>      exit 0 = ok, 77 = no CUDA device (skip), 1 = failure
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "tcif.h"
#include "tcif_mesh.h"
extern "C" {
#include "tcif_oracle.h"
}

namespace foamlike
{
struct vector { double x, y, z; };          // Foam::Vector<double>: three contiguous doubles
static_assert(sizeof(vector) == 3 * sizeof(double), "Field<vector> must be AoS without padding");
typedef int32_t label;
typedef double scalar;

struct fvPatch
{
    int type;                   // TCIF_PATCH_*
    label nFaces;               // faces in the polyMesh (> 0 for empty patches too)
    int neighbProcNo;
    std::vector<label> faceCells;
    std::vector<vector> Sf, delta;
    std::vector<scalar> magSf, deltaCoeffs, weights;
    bool empty() const { return type == TCIF_PATCH_EMPTY; }
    label size() const { return empty() ? 0 : nFaces; }        // what a patch FIELD carries
};

struct fvMesh
{
    label nCells = 0, nInternalFaces = 0;
    std::vector<label> owner, neighbour;
    std::vector<vector> Sf, C;
    std::vector<scalar> weights, V;
    std::vector<fvPatch> boundary;
};

template <class T> struct GeoField
{
    std::vector<T> internal;
    std::vector<std::vector<T>> patches;    // [patch][face]; zero-length on empty patches
    mutable unsigned long event = 1;        // regIOobject::eventNo()
    std::vector<T>& primitiveFieldRef() { ++event; return internal; }
    std::vector<std::vector<T>>& boundaryFieldRef() { ++event; return patches; }
    unsigned long eventNo() const { return event; }
};
typedef GeoField<scalar> volScalarField;
typedef GeoField<vector> volVectorField;
typedef GeoField<scalar> surfaceScalarField;

[[noreturn]] static void fatal(const std::string& what, const char* why)
{
    std::fprintf(stderr, "--> FOAM FATAL ERROR (harness): %s failed: %s\n", what.c_str(), why ? why : "?");
    std::exit(1);
}

// ---- the bridge, member for member as in foam_shim/tcifFoamBridge.H -------------------------------------------
class tcifBridge
{
    const fvMesh& mesh_;
    tcif_ctx* ctx_;
    label nBF_;
    std::vector<label> patchStart_, patchSize_;

public:
    void check(int rc, const char* what) const { if (rc) fatal(what, tcif_last_error(ctx_)); }

    tcifBridge(const fvMesh& mesh, label nPhases, int nLimiterIter) : mesh_(mesh), ctx_(nullptr), nBF_(0)
    {
        const size_t nP = mesh.boundary.size();
        patchStart_.resize(nP); patchSize_.resize(nP);
        std::vector<label> pType(nP), pNbr(nP, -1);
        for (size_t p = 0; p < nP; ++p)
        {
            const fvPatch& fp = mesh.boundary[p];
            patchStart_[p] = nBF_;
            pType[p] = fp.type;
            patchSize_[p] = fp.size();
            if (fp.type == TCIF_PATCH_PROCESSOR) pNbr[p] = fp.neighbProcNo;
            nBF_ += patchSize_[p];
        }
        std::vector<label> faceCells(nBF_);
        std::vector<vector> Sfb(nBF_);
        std::vector<scalar> magSfb(nBF_), dcb(nBF_), wb(nBF_);
        for (size_t p = 0; p < nP; ++p)
            for (label i = 0; i < patchSize_[p]; ++i)
            {
                const fvPatch& fp = mesh.boundary[p];
                const label k = patchStart_[p] + i;
                faceCells[k] = fp.faceCells[i]; Sfb[k] = fp.Sf[i]; magSfb[k] = fp.magSf[i];
                dcb[k] = fp.deltaCoeffs[i]; wb[k] = fp.weights[i];
            }
        tcif_mesh_desc d;
        d.nCells = mesh.nCells; d.nInternalFaces = mesh.nInternalFaces;
        d.nPatches = (label)nP; d.nBoundaryFaces = nBF_;
        d.owner = mesh.owner.data(); d.neighbour = mesh.neighbour.data();
        d.Sf = reinterpret_cast<const double*>(mesh.Sf.data());
        d.weights = mesh.weights.data(); d.V = mesh.V.data();
        d.patchType = pType.data(); d.patchStart = patchStart_.data(); d.patchSize = patchSize_.data();
        d.patchNeighbProcNo = pNbr.data(); d.faceCells = faceCells.data();
        d.SfBoundary = reinterpret_cast<const double*>(Sfb.data());
        d.magSfBoundary = magSfb.data(); d.deltaCoeffsBoundary = dcb.data(); d.weightsBoundary = wb.data();
        tcif_config cfg;
        cfg.nPhases = nPhases; cfg.nLimiterIter = nLimiterIter; cfg.smoothLimiter = 0; cfg.extremaCoeff = 0;
        cfg.boundaryExtrema = TCIF_BEXT_GLOBAL; cfg.device = -1; cfg.flags = 0;
        if (tcif_create(&d, &cfg, &ctx_)) fatal("tcif_create", tcif_last_error(nullptr));
    }
    ~tcifBridge() { tcif_destroy(ctx_); }

    tcif_ctx* ctx() const { return ctx_; }
    label nBoundaryFaces() const { return nBF_; }
    const std::vector<label>& patchStart() const { return patchStart_; }
    const std::vector<label>& patchSize() const { return patchSize_; }

    template <class T> std::vector<T> flatBoundary(const GeoField<T>& f) const
    {
        std::vector<T> b(nBF_);
        for (size_t p = 0; p < f.patches.size(); ++p)
            for (label i = 0; i < patchSize_[p]; ++i) b[patchStart_[p] + i] = f.patches[p][i];
        return b;
    }
    template <class T> void unflatBoundary(const std::vector<T>& b, GeoField<T>& f) const
    {
        std::vector<std::vector<T>>& bf = f.boundaryFieldRef();
        for (size_t p = 0; p < bf.size(); ++p)
            for (label i = 0; i < patchSize_[p]; ++i) bf[p][i] = b[patchStart_[p] + i];
    }

    void setAlphaBc(label phasei, const std::vector<label>& kind) const { check(tcif_set_alpha_bc(ctx_, phasei, kind.data()), "tcif_set_alpha_bc"); }
    void setAlpha(label phasei, const volScalarField& alpha) const
    {
        const std::vector<scalar> b(flatBoundary(alpha));
        check(tcif_set_alpha_stamped(ctx_, phasei, uint64_t(alpha.eventNo()), alpha.internal.data(), b.data(), nullptr), "tcif_set_alpha_stamped");
    }
    void getAlpha(label phasei, volScalarField& alpha) const
    {
        std::vector<scalar> b(nBF_);
        check(tcif_get_alpha(ctx_, phasei, alpha.primitiveFieldRef().data(), b.data()), "tcif_get_alpha");
        unflatBoundary(b, alpha);
        check(tcif_mark_alpha(ctx_, phasei, uint64_t(alpha.eventNo())), "tcif_mark_alpha");
    }
    void setTemperature(const volScalarField& T) const
    {
        const std::vector<scalar> b(flatBoundary(T));
        check(tcif_set_temperature_stamped(ctx_, uint64_t(T.eventNo()), T.internal.data(), b.data(), nullptr), "tcif_set_temperature_stamped");
    }
    void setPhi(const surfaceScalarField& phi) const
    {
        const std::vector<scalar> b(flatBoundary(phi));
        check(tcif_set_phi(ctx_, phi.internal.data(), b.data()), "tcif_set_phi");
    }
    void getPhiAlpha(label phasei, surfaceScalarField& f) const
    {
        std::vector<scalar> b(nBF_);
        check(tcif_get_phi_alpha(ctx_, phasei, f.primitiveFieldRef().data(), b.data()), "tcif_get_phi_alpha");
        unflatBoundary(b, f);
    }
};

// body of phaseSystem::divCapillaryStress() (foam_shim/phaseSystemDevice.C)
static volVectorField divCapillaryStress(const tcifBridge& dev, const fvMesh& mesh, const std::vector<volScalarField>& alphas,
                                         const std::vector<std::vector<label>>& bc, const volScalarField& T)
{
    volVectorField cst;
    cst.internal.assign(mesh.nCells, vector{0, 0, 0});
    cst.patches.resize(mesh.boundary.size());
    for (size_t p = 0; p < mesh.boundary.size(); ++p) cst.patches[p].assign(mesh.boundary[p].size(), vector{0, 0, 0});
    for (size_t p = 0; p < alphas.size(); ++p) { dev.setAlphaBc((label)p, bc[p]); dev.setAlpha((label)p, alphas[p]); }
    dev.setTemperature(T);
    dev.check(tcif_div_capillary_stress(dev.ctx()), "tcif_div_capillary_stress");
    std::vector<vector> b(dev.nBoundaryFaces());
    dev.check(tcif_get_force(dev.ctx(), reinterpret_cast<double*>(cst.primitiveFieldRef().data()), reinterpret_cast<double*>(b.data())), "tcif_get_force");
    dev.unflatBoundary(b, cst);
    return cst;
}
}  // namespace foamlike

using namespace foamlike;

// ---- case set-up ------------------------------------------------------------------------------------------------
struct Case
{
    fvMesh mesh;
    tcif_mesh* hm = nullptr;
    int P = 2;
    bool twoD = true;
    double dx = 0, deltaT = 0;
    std::vector<volScalarField> alpha;
    std::vector<std::vector<label>> bc;
    volScalarField T;
    surfaceScalarField phi;
};

static void alphasAt(const Case& c, const double X[3], double zc, double* out)
{
    const double wdt = 1.5 * c.dx;
    auto blob = [&](double cx, double R) {
        double d0 = X[0] - cx, d1 = X[1] - 0.5, d2 = c.twoD ? 0.0 : X[2] - zc;
        return 0.5 * (1.0 - std::tanh((std::sqrt(d0 * d0 + d1 * d1 + d2 * d2) - R) / wdt));
    };
    if (c.P == 2) { out[0] = blob(0.5, 0.25); out[1] = 1.0 - out[0]; return; }
    double a1 = blob(0.35, 0.2), a2 = blob(0.65, 0.2), a3 = std::fmax(1.0 - a1 - a2, 0.0), s = a1 + a2 + a3;
    out[0] = a1 / s; out[1] = a2 / s; out[2] = a3 / s;
}

static void buildCase(Case& c, int config, int n)
{
    c.twoD = config == 1; c.P = config == 1 ? 2 : 3;
    const int nx = n, ny = n, nz = c.twoD ? 1 : n;
    c.dx = 1.0 / n;
    const double L[3] = {1.0, 1.0, c.twoD ? c.dx : 1.0};
    std::vector<double> xs(nx + 1), ys(ny + 1), zs(nz + 1);
    for (int i = 0; i <= nx; ++i) xs[i] = L[0] * i / nx;
    for (int i = 0; i <= ny; ++i) ys[i] = L[1] * i / ny;
    for (int i = 0; i <= nz; ++i) zs[i] = L[2] * i / nz;
    const int32_t pt2[6] = {0, 0, 0, 0, TCIF_PATCH_EMPTY, TCIF_PATCH_EMPTY}, pt3[6] = {0, 0, 0, 0, 0, 0};
    if (tcif_mesh_box_block(nx, ny, nz, xs.data(), ys.data(), zs.data(), c.twoD ? pt2 : pt3, 1, 1, 1, 0, &c.hm)) fatal("tcif_mesh_box_block", tcif_mesh_last_error());
    tcif_mesh_desc d;
    if (tcif_mesh_get_desc(c.hm, &d)) fatal("tcif_mesh_get_desc", tcif_mesh_last_error());
    const double *C = nullptr, *Cf = nullptr, *Cfb = nullptr; int64_t cnt = 0;
    tcif_mesh_array_f64(c.hm, "C", &C, &cnt); tcif_mesh_array_f64(c.hm, "Cf", &Cf, &cnt); tcif_mesh_array_f64(c.hm, "Cfb", &Cfb, &cnt);
    fvMesh& m = c.mesh;
    m.nCells = d.nCells; m.nInternalFaces = d.nInternalFaces;
    m.owner.assign(d.owner, d.owner + d.nInternalFaces); m.neighbour.assign(d.neighbour, d.neighbour + d.nInternalFaces);
    m.weights.assign(d.weights, d.weights + d.nInternalFaces); m.V.assign(d.V, d.V + d.nCells);
    m.Sf.resize(d.nInternalFaces); m.C.resize(d.nCells);
    for (int f = 0; f < d.nInternalFaces; ++f) m.Sf[f] = vector{d.Sf[3 * f], d.Sf[3 * f + 1], d.Sf[3 * f + 2]};
    for (int i = 0; i < d.nCells; ++i) m.C[i] = vector{C[3 * i], C[3 * i + 1], C[3 * i + 2]};
    // patches as OpenFOAM holds them: per-patch arrays; the empty patches own nx*ny polyMesh faces and no field values
    std::vector<std::vector<vector>> CfPatch(d.nPatches);
    m.boundary.resize(d.nPatches);
    for (int p = 0; p < d.nPatches; ++p)
    {
        fvPatch& fp = m.boundary[p];
        fp.type = d.patchType[p]; fp.neighbProcNo = -1;
        fp.nFaces = fp.type == TCIF_PATCH_EMPTY ? nx * ny : d.patchSize[p];
        for (int i = 0; i < d.patchSize[p]; ++i)
        {
            const int k = d.patchStart[p] + i;
            fp.faceCells.push_back(d.faceCells[k]);
            fp.Sf.push_back(vector{d.SfBoundary[3 * k], d.SfBoundary[3 * k + 1], d.SfBoundary[3 * k + 2]});
            fp.magSf.push_back(d.magSfBoundary[k]); fp.deltaCoeffs.push_back(d.deltaCoeffsBoundary[k]); fp.weights.push_back(d.weightsBoundary[k]);
            const int fc = d.faceCells[k];
            fp.delta.push_back(vector{Cfb[3 * k] - C[3 * fc], Cfb[3 * k + 1] - C[3 * fc + 1], Cfb[3 * k + 2] - C[3 * fc + 2]});
            CfPatch[p].push_back(vector{Cfb[3 * k], Cfb[3 * k + 1], Cfb[3 * k + 2]});
        }
    }
    // fields
    const double zc = 0.5 * L[2], T0 = 300.0, G = 100.0, U = 0.1, pi = 3.14159265358979323846;
    c.alpha.resize(c.P); c.bc.assign(c.P, std::vector<label>(d.nPatches, TCIF_BC_ZERO_GRADIENT));
    for (int p = 0; p < c.P; ++p) { c.alpha[p].internal.resize(m.nCells); c.alpha[p].patches.resize(d.nPatches); }
    c.T.internal.resize(m.nCells); c.T.patches.resize(d.nPatches);
    c.phi.internal.resize(m.nInternalFaces); c.phi.patches.resize(d.nPatches);
    for (int i = 0; i < m.nCells; ++i)
    {
        double a[4]; const double X[3] = {m.C[i].x, m.C[i].y, m.C[i].z};
        alphasAt(c, X, zc, a);
        for (int p = 0; p < c.P; ++p) c.alpha[p].internal[i] = a[p];
        c.T.internal[i] = T0 + G * X[0] + 1e-3 * std::sin(12.9898 * i);
    }
    for (int p = 0; p < d.nPatches; ++p)
    {
        const fvPatch& fp = m.boundary[p];
        for (int q = 0; q < c.P; ++q) c.alpha[q].patches[p].resize(fp.size());
        c.T.patches[p].resize(fp.size()); c.phi.patches[p].assign(fp.size(), 0.0);
        for (int i = 0; i < fp.size(); ++i)
        {
            for (int q = 0; q < c.P; ++q) c.alpha[q].patches[p][i] = c.alpha[q].internal[fp.faceCells[i]];   // zeroGradient
            c.T.patches[p][i] = T0 + G * CfPatch[p][i].x;                                                      // fixedValue wall
        }
    }
    for (int f = 0; f < m.nInternalFaces; ++f)
    {
        const double x = Cf[3 * f], y = Cf[3 * f + 1];
        const double ux = U * std::sin(pi * x) * std::cos(pi * y), uy = -U * std::cos(pi * x) * std::sin(pi * y);
        c.phi.internal[f] = m.Sf[f].x * ux + m.Sf[f].y * uy;
    }
    c.deltaT = 0.25 * c.dx / U;
}

static int nFail = 0;
static void expect(bool ok, const char* what) { if (!ok) { std::fprintf(stderr, "harness: FAILED: %s\n", what); ++nFail; } }
static double relErr(const double* a, const double* b, size_t n)
{
    double scale = 0, err = 0;
    for (size_t i = 0; i < n; ++i) { scale = std::fmax(scale, std::fmax(std::fabs(a[i]), std::fabs(b[i]))); err = std::fmax(err, std::fabs(a[i] - b[i])); }
    return scale > 0 ? err / scale : err;
}

int main(int argc, char** argv)
{
    const int config = argc > 1 ? std::atoi(argv[1]) : 1, n = argc > 2 ? std::atoi(argv[2]) : 24;
    if (config != 1 && config != 5) { std::fprintf(stderr, "usage: shim_harness <1|5> <n>\n"); return 1; }
    if (tcif_device_count() < 1) { std::printf("harness: no CUDA device (the path has no CPU fallback) -- skipped\n"); return 77; }
    Case c;
    buildCase(c, config, n);
    const fvMesh& mesh = c.mesh;
    const int P = c.P, Q = P * (P - 1) / 2;
    tcifBridge dev(mesh, P, 3);
    // uploadSigma: config 1 constant, config 5 temperatureDependent polynomial ((a 0) (b 1))
    std::vector<orc_sigma> osig(Q);
    const double s0[3] = {0.07, 0.05, 0.03}, gamma = 1e-4, T0 = 300.0;
    {
        int q = 0;
        for (int p1 = 0; p1 < P; ++p1) for (int p2 = p1 + 1; p2 < P; ++p2, ++q)
        {
            orc_sigma& o = osig[q];
            for (double& x : o.c) x = 0;
            if (config == 1) { const double s = s0[0]; dev.check(tcif_set_sigma(dev.ctx(), p1, p2, TCIF_SIGMA_CONSTANT, 1, &s), "tcif_set_sigma"); o.kind = 0; o.n = 1; o.c[0] = s; }
            else
            {
                const double co[4] = {s0[q] + gamma * T0, 0.0, -gamma, 1.0};
                dev.check(tcif_set_sigma(dev.ctx(), p1, p2, TCIF_SIGMA_POLYNOMIAL, 2, co), "tcif_set_sigma");
                o.kind = 1; o.n = 2; for (int k = 0; k < 4; ++k) o.c[k] = co[k];
            }
        }
    }
    // the oracle sees the same flattened arrays the bridge built
    const int nBF = dev.nBoundaryFaces(), nP = (int)mesh.boundary.size();
    std::vector<label> pType(nP), pNbr(nP, -1), faceCells(nBF);
    std::vector<double> Sfb(3 * (size_t)nBF), magSfb(nBF), dcb(nBF), wb(nBF), deltab(3 * (size_t)nBF);
    for (int p = 0; p < nP; ++p)
    {
        pType[p] = mesh.boundary[p].type;
        for (int i = 0; i < dev.patchSize()[p]; ++i)
        {
            const int k = dev.patchStart()[p] + i; const fvPatch& fp = mesh.boundary[p];
            faceCells[k] = fp.faceCells[i]; magSfb[k] = fp.magSf[i]; dcb[k] = fp.deltaCoeffs[i]; wb[k] = fp.weights[i];
            Sfb[3 * k] = fp.Sf[i].x; Sfb[3 * k + 1] = fp.Sf[i].y; Sfb[3 * k + 2] = fp.Sf[i].z;
            deltab[3 * k] = fp.delta[i].x; deltab[3 * k + 1] = fp.delta[i].y; deltab[3 * k + 2] = fp.delta[i].z;
        }
    }
    orc_mesh om;
    om.nCells = mesh.nCells; om.nIF = mesh.nInternalFaces; om.nPatches = nP; om.nBF = nBF;
    om.owner = mesh.owner.data(); om.neighbour = mesh.neighbour.data(); om.Sf = reinterpret_cast<const double*>(mesh.Sf.data());
    om.w = mesh.weights.data(); om.V = mesh.V.data(); om.patchType = pType.data(); om.patchStart = dev.patchStart().data();
    om.patchSize = dev.patchSize().data(); om.patchNbrProc = pNbr.data(); om.faceCells = faceCells.data(); om.Sfb = Sfb.data();
    om.magSfb = magSfb.data(); om.deltaCoeffsb = dcb.data(); om.wb = wb.data();
    om.C = reinterpret_cast<const double*>(mesh.C.data()); om.deltab = deltab.data();
    std::vector<label> bcFlat;
    for (int p = 0; p < P; ++p) bcFlat.insert(bcFlat.end(), c.bc[p].begin(), c.bc[p].end());

    auto oracleForce = [&](std::vector<double>& Fi, std::vector<double>& Fb) {
        std::vector<std::vector<double>> ab(P);
        std::vector<const double*> ai(P), abp(P);
        for (int p = 0; p < P; ++p) { ab[p] = dev.flatBoundary(c.alpha[p]); ai[p] = c.alpha[p].internal.data(); abp[p] = ab[p].data(); }
        const std::vector<double> Tb(dev.flatBoundary(c.T));
        Fi.assign(3 * (size_t)mesh.nCells, 0.0); Fb.assign(3 * (size_t)nBF, 0.0);
        orc_div_capillary_stress(&om, nullptr, P, ai.data(), abp.data(), bcFlat.data(), osig.data(), c.T.internal.data(), Tb.data(), Fi.data(), Fb.data());
    };
    auto checkForce = [&](const volVectorField& F, const char* when) {
        std::vector<double> Fi, Fb;
        oracleForce(Fi, Fb);
        const std::vector<vector> fb(dev.flatBoundary(F));
        const double e1 = relErr(reinterpret_cast<const double*>(F.internal.data()), Fi.data(), Fi.size());
        const double e2 = nBF ? relErr(reinterpret_cast<const double*>(fb.data()), Fb.data(), Fb.size()) : 0.0;
        std::printf("harness: divCapillaryStress %-28s rel err internal %.2e  boundary %.2e\n", when, e1, e2);
        expect(e1 <= 1e-12 && e2 <= 1e-12, "force within 1e-12 of the oracle");
    };

    // createFields.H:168-179
    volVectorField F = divCapillaryStress(dev, mesh, c.alpha, c.bc, c.T);
    checkForce(F, "(createFields.H:178)");

    // fluid.solve() -- foam_shim/multiphaseSystemDeviceSolve.C
    const double rho[4] = {1000.0, 800.0, 1.2, 50.0};
    {
        const std::vector<vector> deltaB = [&] { volVectorField dummy; dummy.patches.resize(nP); for (int p = 0; p < nP; ++p) if (!mesh.boundary[p].empty()) dummy.patches[p] = mesh.boundary[p].delta; return dev.flatBoundary(dummy); }();
        dev.check(tcif_set_cell_centres(dev.ctx(), reinterpret_cast<const double*>(mesh.C.data()), reinterpret_cast<const double*>(deltaB.data())), "tcif_set_cell_centres");
        std::vector<double> cA(P * P, 1.0);
        dev.check(tcif_set_calpha(dev.ctx(), cA.data()), "tcif_set_calpha");
    }
    tcif_alpha_controls ctl;
    ctl.nAlphaCorr = 1; ctl.nAlphaSubCycles = 1; ctl.schemePhi = TCIF_SCHEME_VANLEER; ctl.schemePhir = TCIF_SCHEME_LINEAR;
    ctl.implicitPredictor = 1; ctl.predictorTol = 1e-12; ctl.predictorMaxIter = 500;
    // oracle first (it works on copies of the host fields)
    std::vector<std::vector<double>> oai(P), oab(P), oapi(P), oapb(P);
    std::vector<double> orpi(mesh.nInternalFaces), orpb(nBF);
    {
        std::vector<double*> ai(P), ab(P), api(P), apb(P);
        for (int p = 0; p < P; ++p)
        {
            oai[p] = c.alpha[p].internal; oab[p] = dev.flatBoundary(c.alpha[p]); oapi[p].assign(mesh.nInternalFaces, 0.0); oapb[p].assign(nBF, 0.0);
            ai[p] = oai[p].data(); ab[p] = oab[p].data(); api[p] = oapi[p].data(); apb[p] = oapb[p].data();
        }
        const std::vector<double> phib(dev.flatBoundary(c.phi));
        std::vector<double> cA(P * P, 1.0);
        orc_mules_cfg mc; mc.nLimiterIter = 3; mc.smoothLimiter = 0; mc.extremaCoeff = 0; mc.boundaryExtrema = ORC_BEXT_GLOBAL;
        orc_alpha_ctl oc; oc.nAlphaCorr = 1; oc.nAlphaSubCycles = 1; oc.schemePhi = ctl.schemePhi; oc.schemePhir = ctl.schemePhir;
        oc.implicitPredictor = 1; oc.predictorTol = ctl.predictorTol; oc.predictorMaxIter = ctl.predictorMaxIter;
        std::vector<const double*> nullP(P, nullptr);
        orc_solve(&om, nullptr, &mc, &oc, P, ai.data(), ab.data(), bcFlat.data(), c.phi.internal.data(), phib.data(), cA.data(), rho,
                  nullP.data(), nullP.data(), nullP.data(), nullP.data(), c.deltaT, api.data(), apb.data(), orpi.data(), orpb.data());
    }
    int64_t x0[4], x1[4];
    tcif_transfer_counters(dev.ctx(), x0);
    dev.setPhi(c.phi);
    for (int p = 0; p < P; ++p)
    {
        dev.setAlphaBc(p, c.bc[p]);
        dev.setAlpha(p, c.alpha[p]);
        dev.check(tcif_set_rho(dev.ctx(), p, rho[p], nullptr, nullptr), "tcif_set_rho");
        dev.check(tcif_set_su_sp(dev.ctx(), p, nullptr, nullptr), "tcif_set_su_sp");
    }
    dev.check(tcif_solve_alphas(dev.ctx(), &ctl, c.deltaT), "tcif_solve_alphas");
    std::vector<surfaceScalarField> alphaPhi(P);
    surfaceScalarField rhoPhi;
    auto shape = [&](surfaceScalarField& f) { f.internal.resize(mesh.nInternalFaces); f.patches.resize(nP); for (int p = 0; p < nP; ++p) f.patches[p].resize(mesh.boundary[p].size()); };
    for (int p = 0; p < P; ++p)
    {
        dev.getAlpha(p, c.alpha[p]);
        shape(alphaPhi[p]);
        dev.getPhiAlpha(p, alphaPhi[p]);
        double st[3];
        dev.check(tcif_alpha_stats(dev.ctx(), p, st), "tcif_alpha_stats");
        const std::vector<double> ab(dev.flatBoundary(c.alpha[p])), apb(dev.flatBoundary(alphaPhi[p]));
        const double ea = relErr(c.alpha[p].internal.data(), oai[p].data(), oai[p].size()), eb = nBF ? relErr(ab.data(), oab[p].data(), nBF) : 0.0;
        const double ef = relErr(alphaPhi[p].internal.data(), oapi[p].data(), oapi[p].size()), eg = nBF ? relErr(apb.data(), oapb[p].data(), nBF) : 0.0;
        std::printf("harness: solve() phase %d: alpha %.2e / %.2e  alphaPhi %.2e / %.2e  min %.3g max %.3g\n", p, ea, eb, ef, eg, st[1], st[2]);
        expect(ea <= 1e-11 && eb <= 1e-11 && ef <= 1e-11 && eg <= 1e-11, "solve() within 1e-11 of the oracle");
    }
    shape(rhoPhi);
    {
        std::vector<double> rb(nBF);
        dev.check(tcif_get_rho_phi(dev.ctx(), rhoPhi.primitiveFieldRef().data(), rb.data()), "tcif_get_rho_phi");
        dev.unflatBoundary(rb, rhoPhi);
        expect(relErr(rhoPhi.internal.data(), orpi.data(), orpi.size()) <= 1e-11, "rhoPhi within 1e-11 of the oracle");
    }
    tcif_transfer_counters(dev.ctx(), x1);
    expect(x1[0] == x0[0], "solve() did not re-upload alpha (the createFields evaluation had put it on the device)");

    // UEqn.H:29, pEqn.H:35 x 3, thermocapillaryInterFoam.C:150 on unchanged fields
    int64_t e0, h0, e1, h1;
    tcif_force_counters(dev.ctx(), &e0, &h0);
    for (int k = 0; k < 5; ++k) F = divCapillaryStress(dev, mesh, c.alpha, c.bc, c.T);
    checkForce(F, "(after solve(), 5th call)");
    tcif_force_counters(dev.ctx(), &e1, &h1);
    tcif_transfer_counters(dev.ctx(), x0);
    expect(e1 - e0 == 1 && h1 - h0 == 4, "5 calls on unchanged fields = 1 evaluation + 4 cache hits");
    expect(x0[0] == x1[0], "no alpha upload after solve()");
    // error path: FatalError text comes from tcif_last_error
    expect(tcif_set_sigma(dev.ctx(), 0, 0, TCIF_SIGMA_CONSTANT, 1, s0) != 0 && std::string(tcif_last_error(dev.ctx())).size() > 0, "bad pair is a loud error");
    tcif_mesh_free(c.hm);
    if (nFail) { std::printf("harness: %d check(s) FAILED\n", nFail); return 1; }
    std::printf("harness: OK (config %d, %d cells, %d phases, %d boundary faces in %d patches)\n", config, mesh.nCells, P, nBF, nP);
    return 0;
}
