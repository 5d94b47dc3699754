// the reference's tests/TestSettle.h against the B200 platform (see suite_harness.h)
#define SUITE_HEADER "TestSettle.h"
#define SUITE_TESTS T(testConstraints())
#include "suite_harness.h"
