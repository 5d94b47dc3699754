// the reference's tests/TestPeriodicTorsionForce.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestPeriodicTorsionForce.h"
#define SUITE_TESTS T(testPeriodicTorsions()), T(testPeriodic())
#include "suite_harness.h"
