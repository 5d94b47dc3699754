// the reference's tests/TestVerletIntegrator.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestVerletIntegrator.h"
#define SUITE_TESTS T(testSingleBond()), T(testConstraints()), T(testConstrainedClusters()), T(testConstrainedMasslessParticles()), T(testConstrainedChain(10)), T(testInitialTemperature()), T(testForceGroups())
#include "suite_harness.h"
