// the reference's tests/TestLangevinMiddleIntegrator.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestLangevinMiddleIntegrator.h"
#define SUITE_TESTS T(testSingleBond()), T(testTemperature()), T(testConstraints()), T(testConstrainedMasslessParticles()), T(testRandomSeed()), T(testInitialTemperature()), T(testForceGroups())
#include "suite_harness.h"
