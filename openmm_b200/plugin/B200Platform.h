// B200Platform.h -- the `B200` OpenMM Platform: a from-scratch platform plugin whose kernels call the C ABI of
// include/b200_nonbonded.h.  Boundary: olla/include/openmm/{Platform,KernelFactory,KernelImpl,kernels}.h of OpenMM 8.6.
#pragma once
#include "openmm/Platform.h"
#include "openmm/KernelFactory.h"
#include "openmm/System.h"
#include "openmm/internal/ContextImpl.h"
#include "b200_nonbonded.h"
#include <condition_variable>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace B200Plugin {

void check(int rc);   // turns a C-ABI error code into an OpenMMException carrying b200_last_error()

// One persistent host thread per additional GPU of a Context (DeviceIndex=0,1,...), the counterpart of the reference's per-device
// WorkThread (platforms/common/include/openmm/common/ComputeContext.h:667-700).  Every C-ABI call of the plugin is issued for all
// ranks at once -- rank 0 on the caller's thread, rank r on worker r -- because the multi-GPU calls are collective (device-side flag
// barriers, NCCL): a rank that was not driven concurrently would leave the others waiting.
class RankPool {
public:
    ~RankPool();
    void start(int numWorkers);
    // runs f(rank) for every rank; returns normally when all returned B200_OK, otherwise throws OpenMMException with the first message
    void run(int numRanks, const std::function<int(int)>& f);
private:
    void loop(int worker);
    std::vector<std::thread> threads;
    std::mutex m; std::condition_variable cvWork, cvDone;
    const std::function<int(int)>* job = nullptr;
    long generation = 0; int pending = 0; bool stop = false;
    std::vector<int> rc; std::vector<std::string> err;
};

class B200Platform : public OpenMM::Platform {
public:
    // per-Context state hung off ContextImpl::setPlatformData (freed in contextDestroyed)
    struct PlatformData {
        b200_ctx* ctx = nullptr;               // rank 0 (the only one in a single-GPU Context)
        std::vector<b200_ctx*> ranks;          // every rank's context, ranks[0] == ctx
        RankPool pool;
        // issue a C-ABI call on every rank (collectives need all of them); f(context, rank) returns the call's return code
        void all(const std::function<int(b200_ctx*, int)>& f) {
            if (ranks.size() <= 1) { check(f(ctx, 0)); return; }
            pool.run((int)ranks.size(), [&](int r) { return f(ranks[r], r); });
        }
        unsigned pendingGroups = 0;            // force groups requested by the ForceImpls between begin/finishComputation
        // A second (third, ...) NonbondedForce of the System is evaluated in a context of its own: positions and box are copied over
        // before each of its evaluations and its forces are added to the main buffer (b200_add_forces).  Host round trips: meant for
        // the rare System that splits its nonbonded interactions over several forces, not for speed.  One device only.
        struct ExtraNonbonded { b200_ctx* ctx = nullptr; unsigned pending = 0; };
        std::vector<ExtraNonbonded*> extras;
        bool hasNonbonded = false;
        std::vector<double> masses; double defaultBox[9] = {0}; int device = 0; unsigned flags = 0;   // what an extra context is created from
        ExtraNonbonded* addExtraNonbonded() {
            b200_system_desc desc = {};
            desc.num_atoms = (int)masses.size(); desc.mass = masses.data();
            for (int i = 0; i < 9; i++) desc.box[i] = defaultBox[i];
            desc.device = device; desc.flags = flags;
            ExtraNonbonded* x = new ExtraNonbonded();
            extras.push_back(x);
            check(b200_context_create(&desc, &x->ctx));
            return x;
        }
        bool positionsValid = false;
        int bondedCount[3] = {0, 0, 0};         // bonds / angles / torsions created so far (a force's terms start at the count it found)
        int bondedPeriodic[3] = {0, 0, 0};
        int bondedForces[3] = {0, 0, 0};        // HarmonicBond / HarmonicAngle / PeriodicTorsion objects in the System
        std::vector<int> bondedActive[3];       // (first term, count) of the forces of each kind whose kernels ran in this evaluation
        bool bondedRestricted[3] = {false, false, false};   // a subset is currently selected in the library
        // a bonded kernel's execute(): this force takes part in the evaluation in progress
        void requestBonded(int kind, unsigned groupBit, int first, int count) { pendingGroups |= groupBit; bondedActive[kind].push_back(first); bondedActive[kind].push_back(count); }
        std::map<std::string, std::string> properties;
        ~PlatformData() {
            for (ExtraNonbonded* x : extras) { if (x->ctx) b200_context_destroy(x->ctx); delete x; }
            if (ranks.empty()) { if (ctx) b200_context_destroy(ctx); return; }
            if (ranks.size() > 1) { try { pool.run((int)ranks.size(), [&](int r) { b200_context_destroy(ranks[r]); return 0; }); } catch (...) {} }
            else b200_context_destroy(ranks[0]);
        }
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

}  // namespace B200Plugin
