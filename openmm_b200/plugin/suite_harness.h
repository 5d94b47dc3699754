// suite_harness.h -- TEST INFRASTRUCTURE: runs one of the reference's own platform-independent test suites (tests/Test*.h of the
// reference tree, included where they lie, never copied) against the `B200` platform, the way
// platforms/cuda/tests/TestCudaNonbondedForce.cpp:30-110 does for the CUDA platform.  The suites define `main`, which stops at
// the first failing test; here it is renamed and every test*() is run on its own so that the report lists which ones pass.
//
// The including .cpp defines SUITE_HEADER (the reference header) and SUITE_TESTS (the list of test calls) before including this.
#pragma once
#include "OpenMM.h"
#include "openmm/Platform.h"
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>
#include <utility>
#include <vector>

static OpenMM::Platform& loadB200() {
    const char* p = getenv("B200_PLUGIN");
    if (!p) { fprintf(stderr, "set B200_PLUGIN to the path of libOpenMMB200.so\n"); exit(2); }
    OpenMM::Platform::loadPluginLibrary(p);
    return OpenMM::Platform::getPlatformByName("B200");
}
OpenMM::Platform& platform = loadB200();     // the name the suites use (CudaTests.h declares `CudaPlatform platform`)
void initializeTests(int, char**) {}

#define main reference_suite_main
#include SUITE_HEADER
#undef main
void runPlatformTests() {}

#define T(call) {#call, [] { call; }}
int main(int argc, char** argv) {
    std::vector<std::pair<std::string, std::function<void()>>> tests = { SUITE_TESTS };
    std::string only = argc > 1 ? argv[1] : "";
    int failed = 0;
    printf("{\"suite\": \"%s\", \"results\": {", SUITE_HEADER);
    bool first = true;
    for (auto& t : tests) {
        if (!only.empty() && t.first.find(only) == std::string::npos) continue;
        std::string msg = "pass";
        try { t.second(); }
        catch (const std::exception& e) { msg = e.what(); failed++; }
        catch (...) { msg = "unknown exception"; failed++; }
        for (auto& ch : msg) if (ch == '"' || ch == '\\' || ch == '\n') ch = ' ';
        if (msg.size() > 300) msg.resize(300);
        printf("%s\"%s\": \"%s\"", first ? "" : ", ", t.first.c_str(), msg.c_str());
        first = false;
        fflush(stdout);
    }
    printf("}, \"failed\": %d}\n", failed);
    return 0;
}
