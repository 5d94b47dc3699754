// the reference's tests/TestNonbondedForce.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestNonbondedForce.h"
#define SUITE_TESTS \
    T(testCoulomb()), T(testLJ()), T(testExclusionsAnd14()), T(testCutoff()), T(testCutoff14()), T(testPeriodic()), \
    T(testPeriodicExceptions()), T(testTriclinic()), T(testLargeSystem()), T(testDispersionCorrection()), T(testChangingParameters()), \
    T(testSwitchingFunction(NonbondedForce::CutoffNonPeriodic)), T(testSwitchingFunction(NonbondedForce::PME)), T(testTwoForces()), \
    T(testParameterOffsets()), T(testEwaldExceptions()), T(testDirectAndReciprocal()), T(testReordering())
#include "suite_harness.h"
