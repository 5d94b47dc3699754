// B200Platform.h -- the `B200` OpenMM Platform: a from-scratch platform plugin whose kernels call the C ABI of
// include/b200_nonbonded.h.  Boundary: olla/include/openmm/{Platform,KernelFactory,KernelImpl,kernels}.h of OpenMM 8.6.
#pragma once
#include "openmm/Platform.h"
#include "openmm/KernelFactory.h"
#include "openmm/System.h"
#include "openmm/internal/ContextImpl.h"
#include "b200_nonbonded.h"
#include <map>
#include <string>

namespace B200Plugin {

class B200Platform : public OpenMM::Platform {
public:
    // per-Context state hung off ContextImpl::setPlatformData (freed in contextDestroyed)
    struct PlatformData {
        b200_ctx* ctx = nullptr;
        unsigned pendingGroups = 0;            // force groups requested by the ForceImpls between begin/finishComputation
        bool positionsValid = false;
        int bondedGroup[3] = {-1, -1, -1};      // force group of the HarmonicBond / HarmonicAngle / PeriodicTorsion terms already added
        std::map<std::string, std::string> properties;
        ~PlatformData() { if (ctx) b200_context_destroy(ctx); }
    };
    B200Platform();
    const std::string& getName() const override { static const std::string name = "B200"; return name; }
    double getSpeed() const override { return 200.0; }   // above the CUDA platform's 100 (CudaPlatform.cpp:144-146)
    bool supportsDoublePrecision() const override { return false; }
    const std::string& getPropertyValue(const OpenMM::Context& context, const std::string& property) const override;
    void setPropertyValue(OpenMM::Context& context, const std::string& property, const std::string& value) const override;
    void contextCreated(OpenMM::ContextImpl& context, const std::map<std::string, std::string>& properties) const override;
    void linkedContextCreated(OpenMM::ContextImpl& context, OpenMM::ContextImpl& originalContext) const override;
    void contextDestroyed(OpenMM::ContextImpl& context) const override;
    static PlatformData& data(OpenMM::ContextImpl& context) { return *reinterpret_cast<PlatformData*>(context.getPlatformData()); }
    static const std::string& DeviceIndex() { static const std::string k = "DeviceIndex"; return k; }
    static const std::string& Precision() { static const std::string k = "Precision"; return k; }
    static const std::string& DeterministicForces() { static const std::string k = "DeterministicForces"; return k; }
    static const std::string& DisablePmeStream() { static const std::string k = "DisablePmeStream"; return k; }
};

class B200KernelFactory : public OpenMM::KernelFactory {
public:
    OpenMM::KernelImpl* createKernelImpl(std::string name, const OpenMM::Platform& platform, OpenMM::ContextImpl& context) const override;
};

void check(int rc);   // turns a C-ABI error code into an OpenMMException carrying b200_last_error()

}  // namespace B200Plugin
