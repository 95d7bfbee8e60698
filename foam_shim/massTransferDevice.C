/*---------------------------------------------------------------------------*\
  massTransferDevice.C -- device binding of the Su/Sp sources.

  NOT COMPILED HERE (no OpenFOAM in the image).  Shows what a maintainer adds so that cases WITH a
  massTransferModel (config 4, the melt pool) keep the whole alpha sub-cycle on the device:

    MassTransferPhaseSystem<BasePhaseSystem>::correctMassSources(T)   MassTransferPhaseSystem.C:447-497
    MassTransferPhaseSystem<BasePhaseSystem>::alphaTransfer(Su, Sp)   MassTransferPhaseSystem.C:503-704
    phaseSystem::addInterfacePorosity(UEqn)                           phaseSystem.C:1148-1183

  The phase index of a name is its position in phaseModels_ iteration order, the same order
  tcifFoamBridge uses for alpha.  Models other than Lee keep their host evaluation and hand
  the stored dmdt_ field over with tcif_set_dmdt.
\*---------------------------------------------------------------------------*/
#include "MassTransferPhaseSystem.H"
#include "Lee.H"
#include "VollerPrakash.H"
#include "tcifFoamBridge.H"

namespace Foam
{

// once, after construction of the phase system (massTransferModels_ is filled at :60-98)
template<class BasePhaseSystem>
void MassTransferPhaseSystem<BasePhaseSystem>::registerDeviceModels()
{
    tcifFoamBridge& dev = tcifFoamBridge::New(this->mesh());
    dev.call(tcif_mass_transfer_clear(dev.ctx()), "tcif_mass_transfer_clear");

    forAllConstIters(massTransferModels_, iter)
    {
        const phasePairKey& key = iter.key();                       // ordered: (from to)
        const interfaceCompositionModel& mdl = iter()();
        tcif_lee_model m;
        m.from = dev.phaseIndex(key.first());
        m.to = dev.phaseIndex(key.second());
        m.includeDivU = mdl.includeDivU();
        m.modelVariable = (mdl.variable() == interfaceCompositionModel::alpha) ? TCIF_VAR_ALPHA : TCIF_VAR_T;
        if (isA<meltingEvaporationModels::LeeBase>(mdl))            // C, Tactivate, alphaMin (Lee.C:39-42)
        {
            const auto& lee = refCast<const meltingEvaporationModels::LeeBase>(mdl);
            m.C = lee.C().value(); m.Tactivate = lee.Tactivate().value(); m.alphaMin = lee.alphaMin();
        }
        else                                                        // host-evaluated model: dmdt via tcif_set_dmdt
        {
            m.C = 0; m.Tactivate = 1; m.alphaMin = 0; m.modelVariable = TCIF_VAR_T;
        }
        dev.call(tcif_add_mass_transfer_lee(dev.ctx(), &m), "tcif_add_mass_transfer_lee");
    }

    // totalPhasePairs() in its own iteration order (alphaTransfer loops over it, :514)
    labelList pairs(2*this->totalPhasePairs().size());
    label q = 0;
    forAllConstIters(this->totalPhasePairs(), iter)
    {
        pairs[2*q] = dev.phaseIndex(iter()().phase1().name());
        pairs[2*q + 1] = dev.phaseIndex(iter()().phase2().name());
        ++q;
    }
    dev.call(tcif_set_phase_pairs(dev.ctx(), q, pairs.cdata()), "tcif_set_phase_pairs");

    forAllConstIters(this->phaseModels_, iter)                     // coeffs(key) = 1/rho (phaseSystem.C:1144)
    {
        const volScalarField rho(iter()().thermo().rho());
        dev.call
        (
            tcif_set_rho(dev.ctx(), dev.phaseIndex(iter()().name()), 0.0, rho.primitiveField().cdata(),
                         scalarField(dev.boundaryValues(rho)).cdata()),
            "tcif_set_rho"
        );
    }
    // solve() then evaluates calculateSuSp() on the device inside every corrector (multiphaseSystem.C:292)
    dev.call(tcif_set_mass_transfer_in_solve(dev.ctx(), 1), "tcif_set_mass_transfer_in_solve");
}


// body of correctMassSources(T) (:447-497)
template<class BasePhaseSystem>
void MassTransferPhaseSystem<BasePhaseSystem>::correctMassSources(const volScalarField& T)
{
    tcifFoamBridge& dev = tcifFoamBridge::New(this->mesh());
    dev.setTemperature(T);
    dev.call(tcif_correct_mass_sources(dev.ctx()), "tcif_correct_mass_sources");
    forAllConstIters(massTransferModels_, iter)                    // heatTransfer()/massSpeciesTransfer() read dmdt_ on the host
    {
        const phasePairKey& key = iter.key();
        volScalarField& dmdt = *dmdt_[key];
        dev.call
        (
            tcif_get_dmdt(dev.ctx(), dev.phaseIndex(key.first()), dev.phaseIndex(key.second()),
                          dmdt.primitiveFieldRef().data()),
            "tcif_get_dmdt"
        );
        dmdt.correctBoundaryConditions();
    }
}


// body of alphaTransfer(Su, Sp) (:503-704) for callers outside solve()
template<class BasePhaseSystem>
void MassTransferPhaseSystem<BasePhaseSystem>::alphaTransfer(SuSpTable& Su, SuSpTable& Sp)
{
    tcifFoamBridge& dev = tcifFoamBridge::New(this->mesh());
    dev.setPhi(this->phi());
    label p = 0;
    forAllConstIters(this->phaseModels_, iter) { dev.setAlpha(p++, iter()()); }
    scalar ddtAlphaMax = 0;
    dev.call(tcif_alpha_transfer(dev.ctx(), &ddtAlphaMax), "tcif_alpha_transfer");
    this->ddtAlphaMax_ = ddtAlphaMax;                              // :700-702
    p = 0;
    forAllConstIters(this->phaseModels_, iter)
    {
        const word& n = iter()().name();
        scalarField su(this->mesh().nCells()), sp(this->mesh().nCells());
        dev.call(tcif_get_su_sp(dev.ctx(), p++, su.data(), sp.data()), "tcif_get_su_sp");
        Su[n].field() += su;                                       // the reference adds to the caller's tables
        Sp[n].field() += sp;
    }
}

} // namespace Foam


// body of phaseSystem::addInterfacePorosity (phaseSystem.C:1148-1183)
void Foam::phaseSystem::addInterfacePorosity(fvVectorMatrix& UEqn)
{
    tcifFoamBridge& dev = tcifFoamBridge::New(mesh_);
    DynamicList<label> solid;
    DynamicList<scalar> Cu;
    forAllConstIters(phaseModels_, iteri)                          // same double loop and key test as :1153-1181
    {
        auto iterk = iteri;
        for (++iterk; iterk != phaseModels_.cend(); ++iterk)
        {
            const phasePairKey keyik(iteri()().name(), iterk()().name(), false);
            if (interfacePorousModelTable_.found(keyik))
            {
                const auto& vp = refCast<const porousModels::VollerPrakash>(interfacePorousModelTable_[keyik]());
                solid.append(dev.phaseIndex(vp.solidPhase()));
                Cu.append(vp.Cu());
            }
        }
    }
    dev.call
    (
        tcif_interface_porosity(dev.ctx(), solid.size(), solid.cdata(), Cu.cdata(), UEqn.diag().data()),
        "tcif_interface_porosity"
    );
}
