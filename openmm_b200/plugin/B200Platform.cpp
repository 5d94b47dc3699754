// B200Platform.cpp -- registration, properties and per-Context device state of the `B200` platform.
// Reference counterpart being replaced: platforms/cuda/src/CudaPlatform.cpp (registration :57-59, properties :62-142,
// contextCreated :209-238).  Written against the plugin boundary only; no code shared with it.
#include "B200Platform.h"
#include "B200Kernels.h"
#include "openmm/OpenMMException.h"
#include "openmm/PluginInitializer.h"
#include "openmm/Context.h"
#include <algorithm>
#include <sstream>
#include <vector>

using namespace OpenMM;
using namespace std;

namespace B200Plugin {

void check(int rc) {
    if (rc != B200_OK) throw OpenMMException(string(b200_last_error()));
}

// ---- RankPool
RankPool::~RankPool() {
    { std::lock_guard<std::mutex> l(m); stop = true; generation++; }
    cvWork.notify_all();
    for (auto& t : threads) t.join();
}
void RankPool::start(int numWorkers) {
    rc.assign(numWorkers + 1, 0); err.assign(numWorkers + 1, string());
    for (int w = 0; w < numWorkers; w++) threads.emplace_back([this, w] { loop(w); });
}
void RankPool::loop(int worker) {
    long seen = 0;
    const int rank = worker + 1;
    for (;;) {
        const std::function<int(int)>* f = nullptr;
        {
            std::unique_lock<std::mutex> l(m);
            cvWork.wait(l, [&] { return generation != seen; });
            seen = generation;
            if (stop) return;
            f = job;
        }
        int r = B200_ERR_INVALID; string msg;
        try { r = (*f)(rank); if (r != B200_OK) msg = b200_last_error(); }      // the error text is thread local: fetch it here
        catch (const std::exception& e) { msg = e.what(); }
        { std::lock_guard<std::mutex> l(m); rc[rank] = r; err[rank] = msg; pending--; }
        cvDone.notify_one();
    }
}
void RankPool::run(int numRanks, const std::function<int(int)>& f) {
    { std::lock_guard<std::mutex> l(m); job = &f; pending = numRanks - 1; generation++; }
    cvWork.notify_all();
    int r0 = B200_ERR_INVALID; string msg0;
    try { r0 = f(0); if (r0 != B200_OK) msg0 = b200_last_error(); }
    catch (const std::exception& e) { msg0 = e.what(); }
    { std::unique_lock<std::mutex> l(m); cvDone.wait(l, [&] { return pending == 0; }); job = nullptr; }
    if (r0 != B200_OK) throw OpenMMException(msg0);
    for (int r = 1; r < numRanks; r++) if (rc[r] != B200_OK) throw OpenMMException(err[r]);
}

B200Platform::B200Platform() {
    B200KernelFactory* factory = new B200KernelFactory();   // owned (and deleted once) by Platform (olla/src/Platform.cpp:70-76)
    for (const string& name : {CalcForcesAndEnergyKernel::Name(), UpdateStateDataKernel::Name(), ApplyConstraintsKernel::Name(),
                               VirtualSitesKernel::Name(), MinimizeKernel::Name(), CalcNonbondedForceKernel::Name(),
                               CalcHarmonicBondForceKernel::Name(), CalcHarmonicAngleForceKernel::Name(), CalcPeriodicTorsionForceKernel::Name(),
                               RemoveCMMotionKernel::Name(), IntegrateLangevinMiddleStepKernel::Name(), IntegrateVerletStepKernel::Name(),
                               ApplyMonteCarloBarostatKernel::Name()})
        registerKernelFactory(name, factory);
    platformProperties.push_back(DeviceIndex());
    platformProperties.push_back(Precision());
    platformProperties.push_back(DeterministicForces());
    platformProperties.push_back(DisablePmeStream());
    setPropertyDefaultValue(DeviceIndex(), "0");
    setPropertyDefaultValue(Precision(), "mixed");
    setPropertyDefaultValue(DeterministicForces(), "true");
    setPropertyDefaultValue(DisablePmeStream(), "false");
}

const string& B200Platform::getPropertyValue(const Context& context, const string& property) const {
    const ContextImpl& impl = getContextImpl(context);
    const PlatformData* d = reinterpret_cast<const PlatformData*>(impl.getPlatformData());
    auto it = d->properties.find(property);
    if (it != d->properties.end()) return it->second;
    return Platform::getPropertyValue(context, property);
}

void B200Platform::setPropertyValue(Context&, const string&, const string&) const {
    throw OpenMMException("B200 platform properties cannot be changed after the Context is created");
}

static string lower(string s) { transform(s.begin(), s.end(), s.begin(), ::tolower); return s; }

void B200Platform::contextCreated(ContextImpl& context, const map<string, string>& properties) const {
    auto prop = [&](const string& key) {
        auto it = properties.find(key);
        return it == properties.end() ? getPropertyDefaultValue(key) : it->second;
    };
    const string precision = lower(prop(Precision()));
    // "single" is served by the mixed-precision kernels (float pair math and positions with a float remainder, double velocities and
    // integration, integer force sums): everything single precision asks for, at the same speed.  "double" (FP64 pair math and meshes,
    // CudaContext.cpp:127-141) is not implemented.
    if (precision != "mixed" && precision != "single") throw OpenMMException("The B200 platform implements Precision=mixed (also selected by 'single'); 'double' is not available");
    if (lower(prop(DeterministicForces())) != "true") throw OpenMMException("The B200 platform always computes deterministic forces; DeterministicForces must be 'true'");
    // DeviceIndex may list several GPUs ("0,1,2,3", CudaPlatform.cpp:116-141): one rank per device inside this process
    vector<int> deviceList;
    {
        stringstream ss(prop(DeviceIndex())); string tok;
        while (getline(ss, tok, ',')) if (!tok.empty()) deviceList.push_back(stoi(tok));
        if (deviceList.empty()) deviceList.push_back(0);
        if (deviceList.size() > 8) throw OpenMMException("The B200 platform supports at most 8 devices per Context");
    }
    const int device = deviceList[0];
    const System& system = context.getSystem();
    const int n = system.getNumParticles();
    for (int i = 0; i < n; i++)
        if (system.isVirtualSite(i)) throw OpenMMException("The B200 platform does not support virtual sites");
    vector<double> mass(n);
    for (int i = 0; i < n; i++) mass[i] = system.getParticleMass(i);
    vector<int> conAtoms; vector<double> conDist;
    for (int i = 0; i < system.getNumConstraints(); i++) {
        int a, b; double d;
        system.getConstraintParameters(i, a, b, d);
        if (mass[a] == 0 && mass[b] == 0) continue;
        conAtoms.push_back(a); conAtoms.push_back(b); conDist.push_back(d);
    }
    Vec3 a, b, c;
    system.getDefaultPeriodicBoxVectors(a, b, c);
    b200_system_desc desc = {};
    desc.num_atoms = n; desc.mass = mass.data();
    const double box[9] = {a[0], a[1], a[2], b[0], b[1], b[2], c[0], c[1], c[2]};
    copy(box, box + 9, desc.box);
    desc.num_constraints = (int)conDist.size(); desc.constraint_atoms = conAtoms.data(); desc.constraint_dist = conDist.data();
    desc.device = device;
    desc.flags = lower(prop(DisablePmeStream())) == "true" ? B200_CTX_DISABLE_PME_STREAM : 0;
    PlatformData* d = new PlatformData();
    d->masses = mass; copy(box, box + 9, d->defaultBox); d->device = device; d->flags = desc.flags;
    try {
        const int R = (int)deviceList.size();
        d->ranks.assign(R, nullptr);
        for (int r = 0; r < R; r++) { b200_system_desc dr = desc; dr.device = deviceList[r]; check(b200_context_create(&dr, &d->ranks[r])); }
        d->ctx = d->ranks[0];
        if (R > 1) {
            // the direct-space rows are partitioned over the ranks; forces are reduce-scattered and positions all-gathered over peer memory
            char id[128];
            check(b200_nccl_unique_id(id));
            d->pool.start(R - 1);
            d->all([&](b200_ctx* c, int r) { return b200_comm_init(c, r, R, id); });
        }
    } catch (...) {
        delete d;
        throw;   // lets ContextImpl's automatic selection fall through to the next platform (ContextImpl.cpp:147-163)
    }
    d->properties[DeviceIndex()] = prop(DeviceIndex());
    d->properties[Precision()] = precision;
    d->properties[DeterministicForces()] = "true";
    d->properties[DisablePmeStream()] = desc.flags ? "true" : "false";
    context.setPlatformData(d);
}

void B200Platform::linkedContextCreated(ContextImpl& context, ContextImpl& originalContext) const {
    contextCreated(context, data(originalContext).properties);
}

void B200Platform::contextDestroyed(ContextImpl& context) const {
    delete reinterpret_cast<PlatformData*>(context.getPlatformData());
}

KernelImpl* B200KernelFactory::createKernelImpl(string name, const Platform& platform, ContextImpl& context) const {
    B200Platform::PlatformData& d = B200Platform::data(context);
    if (name == CalcForcesAndEnergyKernel::Name()) return new B200CalcForcesAndEnergyKernel(name, platform, d);
    if (name == UpdateStateDataKernel::Name()) return new B200UpdateStateDataKernel(name, platform, d);
    if (name == ApplyConstraintsKernel::Name()) return new B200ApplyConstraintsKernel(name, platform, d);
    if (name == VirtualSitesKernel::Name()) return new B200VirtualSitesKernel(name, platform);
    if (name == MinimizeKernel::Name()) return new B200MinimizeKernel(name, platform);
    if (name == CalcNonbondedForceKernel::Name()) return new B200CalcNonbondedForceKernel(name, platform, d);
    if (name == CalcHarmonicBondForceKernel::Name()) return new B200CalcHarmonicBondForceKernel(name, platform, d);
    if (name == CalcHarmonicAngleForceKernel::Name()) return new B200CalcHarmonicAngleForceKernel(name, platform, d);
    if (name == CalcPeriodicTorsionForceKernel::Name()) return new B200CalcPeriodicTorsionForceKernel(name, platform, d);
    if (name == RemoveCMMotionKernel::Name()) return new B200RemoveCMMotionKernel(name, platform, d);
    if (name == IntegrateLangevinMiddleStepKernel::Name()) return new B200IntegrateLangevinMiddleStepKernel(name, platform, d);
    if (name == IntegrateVerletStepKernel::Name()) return new B200IntegrateVerletStepKernel(name, platform, d);
    if (name == ApplyMonteCarloBarostatKernel::Name()) return new B200ApplyMonteCarloBarostatKernel(name, platform, d);
    throw OpenMMException("Tried to create kernel with illegal kernel name '" + name + "'");
}

}  // namespace B200Plugin

// ---- the two C symbols the plugin loader resolves (olla/include/openmm/PluginInitializer.h:46,56; olla/src/Platform.cpp:245-256)
extern "C" void registerPlatforms() {
    // only offer the platform where it can work: a visible CUDA device is required (no CPU fallback)
    if (b200_device_count() > 0) OpenMM::Platform::registerPlatform(new B200Plugin::B200Platform());
}
extern "C" void registerKernelFactories() {
}
