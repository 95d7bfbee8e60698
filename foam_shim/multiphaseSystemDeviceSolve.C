/*---------------------------------------------------------------------------*\
  multiphaseSystemDeviceSolve.C -- whole-function drop-in for multiphaseSystem::solve()
  (replaces phasesSystem/multiphaseSystem/multiphaseSystem.C:128-168 AND makes solveAlphas() :170-429
  unused) for cases WITHOUT a massTransferModel (Su = Sp = 0: configs 1-3 and 5 of BASELINE.json).
  With mass transfer keep solve()/solveAlphas() and use multiphaseSystemDevice.C, which leaves
  calculateSuSp() on the host between the device calls.

  NOT COMPILED HERE (no OpenFOAM in the image).  Everything between the upload of alpha/phi and the
  download of alpha/alphaPhi/rhoPhi runs on the device: flux assembly (fvc::flux with the schemes of
  fvSchemes + nHatf compression), MULES::limit, limitSum, implicit upwind predictor, explicitSolve,
  rhoPhi, renormalisation, sub-cycling.
\*---------------------------------------------------------------------------*/
#include "multiphaseSystem.H"
#include "tcifFoamBridge.H"

namespace Foam
{

static int tcifScheme(const fvMesh& mesh, const word& name)
{
    ITstream& is = mesh.divScheme(name);          // e.g. "Gauss vanLeer"
    const word gauss(is), scheme(is);
    if (gauss != "Gauss") FatalErrorInFunction << "divScheme " << name << ": only Gauss schemes" << exit(FatalError);
    if (scheme == "upwind") return TCIF_SCHEME_UPWIND;
    if (scheme == "linear") return TCIF_SCHEME_LINEAR;
    if (scheme == "vanLeer") return TCIF_SCHEME_VANLEER;
    FatalErrorInFunction << "convection scheme " << scheme << " for " << name
        << " is not supported by the device path (upwind | linear | vanLeer)" << exit(FatalError);
    return -1;
}

void multiphaseSystem::solve()
{
    const dictionary& alphaControls = mesh_.solverDict("alpha");
    const tcifFoamBridge& dev = device_();                      // created by divCapillaryStress() (createFields.H:168)
    const label P = phases_.size();

    // once per bridge (static mesh, fvSolution read once): geometry for the NVD limiter and the cAlphas table
    if (!dev.staticInputsSet())
    {
        vectorField deltaB(dev.nBoundaryFaces());
        // fvPatch::delta() patch by patch, flattened like every other boundary field
        volVectorField dummy(IOobject("d", mesh_.time().timeName(), mesh_), mesh_, dimensionedVector(dimLength, Zero));
        forAll(mesh_.boundary(), p) dummy.boundaryFieldRef()[p] == mesh_.boundary()[p].delta();
        deltaB = dev.flatBoundary(dummy);
        dev.call(tcif_set_cell_centres(dev.ctx(),
                 reinterpret_cast<const double*>(mesh_.C().primitiveField().cdata()),
                 reinterpret_cast<const double*>(deltaB.cdata())), "tcif_set_cell_centres");

        alphaControls.readEntry("cAlphas", cAlphas_);           // multiphaseSystem.C:174
        scalarList cA(P*P, 1.0);
        label p1 = 0;
        for (const phaseModel& ph1 : phases_)
        {
            label p2 = 0;
            for (const phaseModel& ph2 : phases_)
            {
                if (p1 != p2)
                {
                    const phasePairKey key(ph1.name(), ph2.name());
                    if (!cAlphas_.found(key))                   // same fatal error as :228-234
                    {
                        FatalErrorInFunction << "Phase compression factor (cAlpha) not found for : "
                            << key << exit(FatalError);
                    }
                    cA[p1*P + p2] = cAlphas_.find(key)();
                }
                ++p2;
            }
            ++p1;
        }
        dev.call(tcif_set_calpha(dev.ctx(), cA.cdata()), "tcif_set_calpha");
        dev.staticInputsSet(true);
    }

    tcif_alpha_controls ctl;
    ctl.nAlphaCorr = alphaControls.get<label>("nAlphaCorr");            // :175
    ctl.nAlphaSubCycles = alphaControls.get<label>("nAlphaSubCycles");  // :131
    ctl.schemePhi = -1; ctl.schemePhir = -1;
    ctl.implicitPredictor = 1;
    ctl.predictorTol = alphaControls.getOrDefault<scalar>("tolerance", 1e-10);
    ctl.predictorMaxIter = alphaControls.getOrDefault<label>("maxIter", 100);

    dev.setPhi(phi());
    label p = 0;
    for (phaseModel& phase : phases_)
    {
        const word n(phase.name());
        // every phase must name the same scheme: the device applies one scheme to all phases
        const int s1 = tcifScheme(mesh_, "div(phi," + n + ')');
        if (ctl.schemePhi >= 0 && s1 != ctl.schemePhi)
        {
            FatalErrorInFunction << "different div(phi,alpha) schemes per phase are not supported" << exit(FatalError);
        }
        ctl.schemePhi = s1;
        for (const phaseModel& other : phases_)
        {
            if (&other == &phase) continue;
            ctl.schemePhir = tcifScheme(mesh_, "div(phir," + other.name() + ',' + n + ')');
        }
        dev.setAlphaBc(p, phase);              // re-uploaded only if a patch changed its kind
        dev.setAlpha(p, phase);                // no bytes move unless the solver touched alpha since getAlpha() below
        const volScalarField& rho = phase.rho()();
        const scalarField rb(dev.flatBoundary(rho));
        dev.call(tcif_set_rho(dev.ctx(), p, 0, rho.primitiveField().cdata(), rb.cdata()), "tcif_set_rho");
        dev.call(tcif_set_su_sp(dev.ctx(), p, nullptr, nullptr), "tcif_set_su_sp");
        ++p;
    }

    dev.call(tcif_solve_alphas(dev.ctx(), &ctl, mesh_.time().deltaT().value()), "tcif_solve_alphas");

    p = 0;
    for (phaseModel& phase : phases_)
    {
        dev.getAlpha(p, phase);
        dev.getPhiAlpha(p, phase.alphaPhi());
        Info<< "MULES: Solving for " << phase.name() << endl;
        double st[3];
        dev.call(tcif_alpha_stats(dev.ctx(), p, st), "tcif_alpha_stats");     // :419-425
        Info<< phase.name() << " volume fraction = " << st[0]/gSum(mesh_.V())
            << "  Min(alpha) = " << st[1] << "  Max(alpha) = " << st[2] << endl;
        ++p;
    }
    scalarField rb(dev.nBoundaryFaces());
    dev.call(tcif_get_rho_phi(dev.ctx(), rhoPhi_.primitiveFieldRef().data(), rb.data()), "tcif_get_rho_phi");
    dev.unflatBoundary(rb, rhoPhi_);
}

} // End namespace Foam
