// the reference's tests/TestDispersionPME.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestDispersionPME.h"
#define SUITE_TESTS \
    T(testConvergence()), T(testErrorTolerance()), T(testPMEParameters()), T(testCoulombAndLJ()), T(testWater2DpmeEnergiesForcesNoExclusions()), \
    T(testDMFDpmeEnergiesForcesWithExclusions()), T(testWater2DpmeEnergiesForcesWithExclusions()), \
    T(testWater125DpmeVsLongCutoffNoExclusions()), T(testWater125DpmeVsLongCutoffWithExclusions())
#include "suite_harness.h"
