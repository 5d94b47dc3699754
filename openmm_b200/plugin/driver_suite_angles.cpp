// the reference's tests/TestHarmonicAngleForce.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestHarmonicAngleForce.h"
#define SUITE_TESTS T(testAngles()), T(testPeriodic())
#include "suite_harness.h"
