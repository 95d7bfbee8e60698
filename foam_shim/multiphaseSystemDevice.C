/*---------------------------------------------------------------------------*\
  multiphaseSystemDevice.C -- the MULES part of multiphaseSystem::solveAlphas() on the device
  (replaces phasesSystem/multiphaseSystem/multiphaseSystem.C:294-380; the flux assembly :196-277
  and calculateSuSp() :279-292 above it stay as they are and produce phiAlphaCorrs, Su_, Sp_).

  NOT COMPILED HERE (no OpenFOAM in the image).  Paste as the body between "calculateSuSp();" and
  "if (acorr == nAlphaCorr - 1)"; `device_` is the bridge created by divCapillaryStress().
\*---------------------------------------------------------------------------*/

    // ---- begin device section (was: MULES::limit x P, MULES::limitSum, explicitSolve x P) ----
    const tcifFoamBridge& dev = device_();
    const scalar deltaT = mesh_.time().deltaT().value();     // already the sub-cycle step inside subCycle

    dev.setPhi(phi);
    {
        label phasei = 0;
        for (phaseModel& phase : phases_)
        {
            // NB phases_ and phaseModels_ iterate in the same (HashTable) order, multiphaseSystem.C:70-76
            dev.setAlphaBc(phasei, phase);
            dev.setAlpha(phasei, phase);
            dev.call
            (
                tcif_set_alpha_old(dev.ctx(), phasei, phase.oldTime().primitiveField().cdata()),
                "tcif_set_alpha_old"
            );
            dev.setPhiAlpha(phasei, phiAlphaCorrs[phasei]);
            dev.call
            (
                tcif_set_su_sp
                (
                    dev.ctx(), phasei,
                    Su_[phase.name()].field().cdata(), Sp_[phase.name()].field().cdata()
                ),
                "tcif_set_su_sp"
            );
            ++phasei;
        }
    }

    // MULES::limit(1/deltaT, 1, alpha1, phi, phiAlphaCorr, Sp, Su, 1, 0, true)   :305-317
    forAll(phases_, phasei)
    {
        dev.call(tcif_mules_limit(dev.ctx(), phasei, 1.0/deltaT, 1), "tcif_mules_limit");
    }
    // MULES::limitSum(phiAlphaCorrs)                                              :321
    dev.call(tcif_mules_limit_sum(dev.ctx()), "tcif_mules_limit_sum");

    {
        label phasei = 0;
        for (phaseModel& phase : phases_)
        {
            // implicit upwind predictor :348-363 -- on the device (tcif_upwind_predictor solves
            // (V/dt - V Sp + sum_out phi) psi - sum_in phi psi_nb = V psi0/dt + V Su to the
            // tolerance of fvSolution's alpha solver and adds alpha1Eqn.flux() to phiAlpha)
            const dictionary& sd = mesh_.solverDict(phase.name());
            dev.call
            (
                tcif_upwind_predictor
                (
                    dev.ctx(), phasei, deltaT,
                    sd.getOrDefault<scalar>("tolerance", 1e-10),
                    sd.getOrDefault<label>("maxIter", 100)
                ),
                "tcif_upwind_predictor"
            );
            // MULES::explicitSolve(1, alpha1, phi, phiAlpha, Sp, Su, 1, 0)         :365-375
            dev.call(tcif_mules_explicit_solve(dev.ctx(), phasei, deltaT), "tcif_mules_explicit_solve");

            dev.getAlpha(phasei, phase);
            dev.getPhiAlpha(phasei, phase.alphaPhi());                             // :377
            Info<< "MULES: Solving for " << phase.name() << endl;
            ++phasei;
        }
    }
    // ---- end device section ----
