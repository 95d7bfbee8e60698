/*---------------------------------------------------------------------------*\
  phaseSystemDevice.C -- drop-in body of phaseSystem::divCapillaryStress()
  (replaces phasesSystem/phaseSystem/phaseSystem.C:1038-1105; signature, return type and the
  name/dimensions of the returned field are unchanged, phaseSystem.H:493-494).

  NOT COMPILED HERE (no OpenFOAM in the image).  To integrate: delete the original body, add this
  file to phasesSystem/Make/files, add `mutable autoPtr<tcifFoamBridge> device_;` to phaseSystem.H
  and the two lines of INTEGRATION.md to Make/options.
\*---------------------------------------------------------------------------*/
#include "phaseSystem.H"
#include "tcifFoamBridge.H"
#include "Function1.H"
#include "Polynomial.H"

namespace Foam
{

// surfaceTension ((a and b) { type constant|temperatureDependent; sigma ...; }) -> tcif_set_sigma.
// Re-reads the dictionary entries the constructor handed to generatePairsAndSubModels
// (phaseSystem.C:266-273) because the upstream surfaceTensionModel exposes only sigma().
static void uploadSigma
(
    const tcifFoamBridge& dev, const dictionary& phaseProperties,
    const wordList& phaseOrder       // HashTable iteration order of phaseModels_
)
{
    const List<Tuple2<phasePairKey, dictionary>> entries(phaseProperties.lookup("surfaceTension"));
    for (const auto& e : entries)
    {
        const label p1 = phaseOrder.find(e.first().first());
        const label p2 = phaseOrder.find(e.first().second());
        const dictionary& d = e.second();
        const word type(d.get<word>("type"));
        if (type == "constant")
        {
            const scalar s = d.get<dimensionedScalar>("sigma").value();
            dev.call(tcif_set_sigma(dev.ctx(), p1, p2, TCIF_SIGMA_CONSTANT, 1, &s), "tcif_set_sigma");
        }
        else if (type == "temperatureDependent")
        {
            // Function1<scalar> "sigma": constant | polynomial ((a e)...) | table ((T s)...)
            ITstream& is = d.lookup("sigma");
            token t(is);
            scalarList co;
            int kind = TCIF_SIGMA_POLYNOMIAL;
            if (t.isNumber()) { co = {t.number(), 0}; }
            else
            {
                const word f1(t.wordToken());
                const List<Tuple2<scalar, scalar>> pairs(is);
                if (f1 == "table") kind = TCIF_SIGMA_TABLE;
                else if (f1 != "polynomial" && f1 != "constant")
                {
                    FatalErrorInFunction << "Function1 type " << f1 << " for sigma is not supported by"
                        << " the device path (constant | polynomial | table)" << exit(FatalError);
                }
                co.resize(2*pairs.size());
                forAll(pairs, i) { co[2*i] = pairs[i].first(); co[2*i + 1] = pairs[i].second(); }
            }
            dev.call(tcif_set_sigma(dev.ctx(), p1, p2, kind, co.size()/2, co.cdata()), "tcif_set_sigma");
        }
        else
        {
            FatalErrorInFunction << "surfaceTensionModel " << type << " is not supported by the device path"
                << exit(FatalError);
        }
    }
}


tmp<volVectorField> phaseSystem::divCapillaryStress() const
{
    auto tcst = tmp<volVectorField>::New
    (
        IOobject("capillaryStressTensor", mesh_.time().timeName(), mesh_),
        mesh_,
        dimensionedVector({1, -2, -2, 0, 0, 0}, Zero),
        extrapolatedCalculatedFvPatchVectorField::typeName    // what fvc::div's result carries
    );
    if (!surfaceTensionModels_.size()) return tcst;

    wordList order;
    forAllConstIters(phaseModels_, iter) order.append(iter()->name());   // table order, as :1064

    if (!device_.valid())
    {
        if (mesh_.changing())
        {
            FatalErrorInFunction << "device path supports static meshes only" << exit(FatalError);
        }
        device_.reset(new tcifFoamBridge(mesh_, order.size(), mesh_.solverDict("alpha")));
        uploadSigma(device_(), *this, order);
    }
    const tcifFoamBridge& dev = device_();

    // stamped: after solve() left the newest alpha on the device (and getAlpha() marked it) nothing is uploaded
    // here, the device keeps F cached, and the 2nd..5th evaluation of a time step is one download of F
    label p = 0;
    forAllConstIters(phaseModels_, iter)
    {
        dev.setAlphaBc(p, iter()());
        dev.setAlpha(p, iter()());
        ++p;
    }
    if (mesh_.foundObject<volScalarField>("T"))        // temperatureDependent looks T up by name
    {
        dev.setTemperature(mesh_.lookupObject<volScalarField>("T"));
    }

    dev.call(tcif_div_capillary_stress(dev.ctx()), "tcif_div_capillary_stress");

    volVectorField& cst = tcst.ref();
    vectorField b(dev.nBoundaryFaces());
    dev.call
    (
        tcif_get_force
        (
            dev.ctx(),
            reinterpret_cast<double*>(cst.primitiveFieldRef().data()),
            reinterpret_cast<double*>(b.data())
        ),
        "tcif_get_force"
    );
    dev.unflatBoundary(b, cst);     // walls: adjacent-cell value; processor: neighbour-rank value
    return tcst;
}

} // End namespace Foam
