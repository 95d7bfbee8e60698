/*---------------------------------------------------------------------------*\
  laserDTRMDevice.C -- device binding of radiation::laserDTRM::calculate().

  NOT COMPILED HERE (no OpenFOAM in the image).  Shows the body a maintainer puts in place of
  laserDTRM/laserDTRM.C:542-735; initialise() (:169-331) keeps seeding DTRMCloud_ on the host (Function1s,
  the `manual` interpolation2DTable, Random(1234), mesh_.findCell), the device walks the cloud.

  One-time set-up (constructor, after initialiseReflection()):
    * polyMesh data of the walk: faceCentres, points, faces, tetBasePtIs, bounds().mag()
    * the reflectionModel table in its iteration order (reflections_, :137-166)
    * the absorptionEmission model: localDensityAbsorptionEmission is evaluated on the device from the resident
      alpha fields; any other model hands its a()/e()/E() fields over every call.
\*---------------------------------------------------------------------------*/
#include "laserDTRM.H"
#include "Fresnel.H"
#include "FresnelLaser.H"
#include "noReflection.H"
#include "localDensityAbsorptionEmission.H"
#include "physicoChemicalConstants.H"
#include "globalIndex.H"
#include "syncTools.H"
#include "processorPolyPatch.H"
#include "tcifFoamBridge.H"

namespace Foam
{
namespace radiation
{

void laserDTRM::registerDevice()
{
    tcifFoamBridge& dev = tcifFoamBridge::New(mesh_);

    // faces as CSR: internal faces, then the boundary faces in the bridge's boundary order (empty patches skipped)
    const faceList& fs = mesh_.faces();
    const labelList faceOrder(dev.meshFaceOrder());          // mesh face label of every face the bridge carries
    labelList fpStart(faceOrder.size() + 1, 0), tetBase(faceOrder.size());
    DynamicList<label> fpPts(4*faceOrder.size());
    vectorField Cf(mesh_.nInternalFaces()), Cfb(faceOrder.size() - mesh_.nInternalFaces());
    forAll(faceOrder, i)
    {
        const label f = faceOrder[i];
        fpPts.append(fs[f]);
        fpStart[i + 1] = fpPts.size();
        tetBase[i] = mesh_.tetBasePtIs()[f];
        (i < mesh_.nInternalFaces() ? Cf[i] : Cfb[i - mesh_.nInternalFaces()]) = mesh_.faceCentres()[f];
    }
    dev.call
    (
        tcif_laser_set_polymesh
        (
            dev.ctx(),
            reinterpret_cast<const double*>(Cf.cdata()), reinterpret_cast<const double*>(Cfb.cdata()),
            mesh_.nPoints(), reinterpret_cast<const double*>(mesh_.points().cdata()),
            fpStart.cdata(), fpPts.cdata(), tetBase.cdata(), mesh_.bounds().mag()
        ),
        "tcif_laser_set_polymesh"
    );

    if (Pstream::parRun())
    {
        // points of the processor patches and their global labels: volPointInterpolation sums over the ranks
        // sharing a point (syncUntransformedData); the device does the same sum with these labels
        const globalIndex globalPoints(mesh_.nPoints());          // any labelling unique per physical point will do:
        labelList gid(mesh_.nPoints());                           // here the minimum over the ranks holding the point
        forAll(gid, p) { gid[p] = globalPoints.toGlobal(p); }
        syncTools::syncPointList(mesh_, gid, minEqOp<label>(), labelMax);
        bitSet onProc(mesh_.nPoints());
        for (const polyPatch& pp : mesh_.boundaryMesh())
        {
            if (isA<processorPolyPatch>(pp)) { onProc.set(pp.meshPoints()); }
        }
        const labelList localPts(onProc.toc());
        List<int64_t> globalPts(localPts.size());
        forAll(localPts, i) { globalPts[i] = gid[localPts[i]]; }
        dev.call
        (
            tcif_laser_set_shared_points(dev.ctx(), localPts.size(), localPts.cdata(), globalPts.cdata()),
            "tcif_laser_set_shared_points"
        );
    }

    // reflections_ in iteration order = the reflectionModelId of calculate() (:596-650)
    List<tcif_reflection_model> models(reflections_.size());
    label id = 0;
    forAllConstIters(reflections_, iter)
    {
        tcif_reflection_model& m = models[id++];
        m.from = dev.phaseIndex(iter.key().first());
        m.to = dev.phaseIndex(iter.key().second());
        m.nk1[0] = m.nk1[1] = m.nk2[0] = m.nk2[1] = m.epsilon = 0;
        const reflectionModel& r = iter()();
        if (isA<Fresnel>(r))
        {
            const Fresnel& fr = refCast<const Fresnel>(r);       // needs accessors for nk1_, nk2_ (Fresnel.H)
            m.kind = TCIF_REFL_FRESNEL;
            m.nk1[0] = fr.nk1()[0]; m.nk1[1] = fr.nk1()[1]; m.nk2[0] = fr.nk2()[0]; m.nk2[1] = fr.nk2()[1];
        }
        else if (isA<FresnelLaser>(r))
        {
            m.kind = TCIF_REFL_FRESNEL_LASER;
            m.epsilon = refCast<const FresnelLaser>(r).epsilon();   // accessor for epsilon_ (FresnelLaser.H)
        }
        else
        {
            m.kind = TCIF_REFL_NONE;
        }
    }
    dev.call(tcif_laser_set_reflection(dev.ctx(), models.size(), models.cdata(), alphaCut_), "tcif_laser_set_reflection");

    if (isA<localDensityAbsorptionEmission>(absorptionEmission_()))
    {
        const auto& ld = refCast<const localDensityAbsorptionEmission>(absorptionEmission_());
        labelList ph(ld.alphaNames().size());
        forAll(ph, i) { ph[i] = dev.phaseIndex(IOobject::group(ld.alphaNames()[i])); }     // "alpha.metal" -> metal
        dev.call
        (
            tcif_laser_set_local_density_absorption
            (
                dev.ctx(), ph.size(), ph.cdata(), ld.aCoeff().cdata(), ld.eCoeff().cdata(), ld.ECoeff().cdata()
            ),
            "tcif_laser_set_local_density_absorption"
        );
        deviceAbsorption_ = true;
    }
}


// replaces laserDTRM.C:542-735
void laserDTRM::calculate()
{
    tcifFoamBridge& dev = tcifFoamBridge::New(mesh_);

    if (!deviceAbsorption_)
    {
        a_ = absorptionEmission_->a(); e_ = absorptionEmission_->e(); E_ = absorptionEmission_->E();   // :583-585
        dev.call
        (
            tcif_laser_set_absorption_fields
            (
                dev.ctx(), a_.primitiveField().cdata(), e_.primitiveField().cdata(), E_.primitiveField().cdata()
            ),
            "tcif_laser_set_absorption_fields"
        );
    }
    dev.setTemperature(T_);          // alpha is resident from the alpha sub-cycle; T changes in TEqn

    // flatten the cloud initialise() seeded
    const label n = DTRMCloud_.size();
    vectorField p0(n), p1(n);
    scalarField I(n), dA(n);
    labelList cell(n);
    label k = 0;
    for (const DTRMParticle& p : DTRMCloud_)
    {
        p0[k] = p.p0(); p1[k] = p.p1(); I[k] = p.I(); dA[k] = p.dA(); cell[k] = p.cell(); ++k;
    }

    // Reset the field (:580); the device returns the cells that hold energy, a small part of the mesh
    Q_ == dimensionedScalar(Q_.dimensions(), Zero);

    label nFinal = 0;
    dev.call
    (
        tcif_laser_calculate
        (
            dev.ctx(), n,
            reinterpret_cast<const double*>(p0.cdata()), reinterpret_cast<const double*>(p1.cdata()),
            I.cdata(), dA.cdata(), cell.cdata(), constant::physicoChemical::sigma.value(), 0,
            nullptr, &nFinal
        ),
        "tcif_laser_calculate"
    );
    labelList hit(sparseCapacity_);
    scalarField val(sparseCapacity_);
    label nHit = 0;
    int rc = tcif_laser_get_sparse(dev.ctx(), hit.size(), hit.data(), val.data(), &nHit);
    if (rc && nHit > hit.size())                     // the one recoverable failure: more cells hold energy than fit
    {
        sparseCapacity_ = 2*nHit;                    // grow once, the beam footprint changes slowly
        hit.resize(sparseCapacity_); val.resize(sparseCapacity_);
        rc = tcif_laser_get_sparse(dev.ctx(), hit.size(), hit.data(), val.data(), &nHit);
    }
    dev.call(rc, "tcif_laser_get_sparse");           // anything else (CUDA error, no polyMesh) is fatal, never Q = 0
    scalarField& Q = Q_.primitiveFieldRef();
    for (label i = 0; i < nHit; ++i) { Q[hit[i]] = val[i]; }         // already divided by V (:671)

    Info<< "Move particles..." << n << " (" << nFinal << " with reflections)" << endl;

    initialise();                    // :734, seeds the cloud of the next call
}

} // End namespace radiation
} // End namespace Foam
