// the reference's tests/TestEwald.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestEwald.h"
#define SUITE_TESTS \
    T(testEwaldExact()), T(testEwaldPME(false)), T(testEwaldPME(true)), T(testTriclinic()), T(testTriclinic2()), \
    T(testErrorTolerance(NonbondedForce::Ewald)), T(testErrorTolerance(NonbondedForce::PME)), T(testPMEParameters()), \
    T(testNeutralizingPlasmaCorrection(NonbondedForce::Ewald, false)), T(testNeutralizingPlasmaCorrection(NonbondedForce::PME, false)), \
    T(testNeutralizingPlasmaCorrection(NonbondedForce::LJPME, false)), T(testNeutralizingPlasmaCorrection(NonbondedForce::Ewald, true)), \
    T(testNeutralizingPlasmaCorrection(NonbondedForce::PME, true)), T(testNeutralizingPlasmaCorrection(NonbondedForce::LJPME, true))
#include "suite_harness.h"
