// the reference's tests/TestLocalEnergyMinimizer.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestLocalEnergyMinimizer.h"
#define SUITE_TESTS T(testHarmonicBonds()), T(testLargeSystem()), T(testLargeForces()), T(testForceGroups()), T(testMasslessParticles()), T(testReporter())
#include "suite_harness.h"
