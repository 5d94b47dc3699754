// the reference's tests/TestHarmonicBondForce.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestHarmonicBondForce.h"
#define SUITE_TESTS T(testBonds()), T(testPeriodic())
#include "suite_harness.h"
