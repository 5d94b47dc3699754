// the reference's tests/TestCheckpoints.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestCheckpoints.h"
#define SUITE_TESTS T(testSetState())
#include "suite_harness.h"
