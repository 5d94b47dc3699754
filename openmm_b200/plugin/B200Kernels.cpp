// B200Kernels.cpp -- the platform's kernels: thin adapters from OpenMM's kernel interfaces to the C ABI.
// Reference host code being replaced for this path: platforms/common/src/CommonCalcNonbondedForce.cpp:217-717,1119-1254
// (initialize / copyParametersToContext), platforms/common/src/CommonKernels.cpp:95-382 (state I/O, checkpoints),
// platforms/cuda/src/CudaKernels.cpp:63-91 (CalcForcesAndEnergy).
#include "B200Kernels.h"
#include "openmm/OpenMMException.h"
#include "openmm/NonbondedForce.h"
#include "openmm/HarmonicBondForce.h"
#include "openmm/HarmonicAngleForce.h"
#include "openmm/PeriodicTorsionForce.h"
#include "openmm/LangevinMiddleIntegrator.h"
#include "openmm/VerletIntegrator.h"
#include "openmm/internal/NonbondedForceImpl.h"
#include "ReferenceMinimize.h"
#include <cmath>
#include <cstring>
#include <iostream>
#include <random>

using namespace OpenMM;
using namespace std;

namespace B200Plugin {

// ---------------------------------------------------------------------------------------------- forces and energy
void B200CalcForcesAndEnergyKernel::beginComputation(ContextImpl&, bool, bool, int) {
    d.pendingGroups = 0;   // every ForceImpl that takes part in this evaluation adds its bit from its kernel's execute()
    for (auto& r : d.bondedActive) r.clear();
    for (auto* x : d.extras) x->pending = 0;
}

double B200CalcForcesAndEnergyKernel::finishComputation(ContextImpl&, bool includeForce, bool includeEnergy, int, bool& valid) {
    double energy = 0.0;
    // the force buffer is cleared by the computation even when no group is requested, as clearAutoclearBuffers does (CudaKernels.cpp:66)
    // every rank evaluates its share (multi-GPU: the energies are all-reduced inside, the forces stay partial until someone needs them)
    vector<double> e(max<size_t>(1, d.ranks.size()), 0.0);
    // Forces of one kind in different force groups: ContextImpl ran only the kernels of the requested groups, and each left its range
    // of terms; when that is not every force of the kind, the library is told which ranges to evaluate (b200_bonded_set_active).
    static const unsigned kindBit[3] = {B200_GROUP_BONDS, B200_GROUP_ANGLES, B200_GROUP_TORSIONS};
    for (int kind = 0; kind < 3; kind++) {
        if (!(d.pendingGroups & kindBit[kind])) continue;
        const bool subset = (int)d.bondedActive[kind].size() / 2 != d.bondedForces[kind];
        if (!subset && !d.bondedRestricted[kind]) continue;
        const vector<int>& r = d.bondedActive[kind];
        d.all([&](b200_ctx* c, int) { return b200_bonded_set_active(c, kind, subset ? (int)r.size() / 2 : 0, subset ? r.data() : nullptr); });
        d.bondedRestricted[kind] = subset;
    }
    d.all([&](b200_ctx* c, int r) { return b200_compute(c, d.pendingGroups, includeForce ? 1 : 0, includeEnergy ? 1 : 0, &e[r]); });
    energy = e[0];
    // further NonbondedForce objects: each in its own context, on the positions and box of the main one; forces added to the main buffer
    for (auto* x : d.extras) {
        if (!x->pending) continue;
        vector<double> buf(3 * d.masses.size());
        double box[9], ex = 0.0;
        check(b200_get_positions(d.ctx, buf.data(), 0)); check(b200_get_box(d.ctx, box));
        check(b200_set_box(x->ctx, box)); check(b200_set_positions(x->ctx, buf.data()));
        check(b200_compute(x->ctx, x->pending, includeForce ? 1 : 0, includeEnergy ? 1 : 0, &ex));
        energy += ex;
        if (includeForce) { check(b200_get_forces(x->ctx, buf.data())); check(b200_add_forces(d.ctx, buf.data())); }
        x->pending = 0;
    }
    valid = true;            // list growth/retry is handled inside the library (CudaNonbondedUtilities.cpp:441-479 does it by invalidating)
    return includeEnergy ? energy : 0.0;
}

// ---------------------------------------------------------------------------------------------- state
double B200UpdateStateDataKernel::getTime(const ContextImpl&) const { double t; int64_t s; check(b200_get_time(d.ctx, &t, &s)); return t; }
void B200UpdateStateDataKernel::setTime(ContextImpl&, double time) { double t; int64_t s; check(b200_get_time(d.ctx, &t, &s)); d.all([&](b200_ctx* c, int) { return b200_set_time(c, time, s); }); }
long long B200UpdateStateDataKernel::getStepCount(const ContextImpl&) const { double t; int64_t s; check(b200_get_time(d.ctx, &t, &s)); return s; }
void B200UpdateStateDataKernel::setStepCount(const ContextImpl&, long long count) { double t; int64_t s; check(b200_get_time(d.ctx, &t, &s)); d.all([&](b200_ctx* c, int) { return b200_set_time(c, t, count); }); }

// A vector<Vec3> IS a contiguous double[3N] (Vec3 holds exactly three doubles, openmmapi/include/openmm/Vec3.h), so the C ABI
// reads and writes the caller's vectors in place: no intermediate copy on the state I/O path.
static_assert(sizeof(Vec3) == 3 * sizeof(double), "Vec3 must be three packed doubles");
static double* asDoubles(vector<Vec3>& v) { return reinterpret_cast<double*>(v.data()); }
static const double* asDoubles(const vector<Vec3>& v) { return reinterpret_cast<const double*>(v.data()); }

void B200UpdateStateDataKernel::getPositions(ContextImpl&, vector<Vec3>& positions, bool) {
    positions.resize(numParticles);
    d.all([&](b200_ctx* c, int r) { return b200_get_positions(c, r == 0 ? asDoubles(positions) : nullptr, 0); });   // rank 0 downloads, the others join the gather
}
void B200UpdateStateDataKernel::setPositions(ContextImpl&, const vector<Vec3>& positions) {
    if ((int)positions.size() != numParticles) throw OpenMMException("Called setPositions() on a Context with the wrong number of positions");
    d.all([&](b200_ctx* c, int) { return b200_set_positions(c, asDoubles(positions)); });
    d.positionsValid = true;
}
void B200UpdateStateDataKernel::getVelocities(ContextImpl&, vector<Vec3>& velocities) {
    velocities.resize(numParticles);
    d.all([&](b200_ctx* c, int r) { return b200_get_velocities(c, r == 0 ? asDoubles(velocities) : nullptr); });
}
void B200UpdateStateDataKernel::setVelocities(ContextImpl&, const vector<Vec3>& velocities) {
    if ((int)velocities.size() != numParticles) throw OpenMMException("Called setVelocities() on a Context with the wrong number of velocities");
    d.all([&](b200_ctx* c, int) { return b200_set_velocities(c, asDoubles(velocities)); });
}
void B200UpdateStateDataKernel::computeShiftedVelocities(ContextImpl& context, double timeShift, vector<Vec3>& velocities) {
    // v(t + shift) ~ v + shift * F/m with the forces currently in the buffer (IntegrationUtilities::computeShiftedVelocities;
    // unlike the kinetic-energy path above this host-side variant does not re-apply the velocity constraints)
    getVelocities(context, velocities);
    if (timeShift == 0.0) return;
    vector<Vec3> forces;
    getForces(context, forces);
    for (int i = 0; i < numParticles; i++)
        if (masses[i] > 0) velocities[i] += forces[i] * (timeShift / masses[i]);
}
void B200UpdateStateDataKernel::getForces(ContextImpl&, vector<Vec3>& forces) {
    forces.resize(numParticles);
    d.all([&](b200_ctx* c, int r) { return b200_get_forces(c, r == 0 ? asDoubles(forces) : nullptr); });
}
void B200UpdateStateDataKernel::getPeriodicBoxVectors(ContextImpl&, Vec3& a, Vec3& b, Vec3& c) const {
    double box[9];
    check(b200_get_box(d.ctx, box));
    a = Vec3(box[0], box[1], box[2]); b = Vec3(box[3], box[4], box[5]); c = Vec3(box[6], box[7], box[8]);
}
void B200UpdateStateDataKernel::setPeriodicBoxVectors(ContextImpl&, const Vec3& a, const Vec3& b, const Vec3& c) {
    const double box[9] = {a[0], a[1], a[2], b[0], b[1], b[2], c[0], c[1], c[2]};
    d.all([&](b200_ctx* c, int) { return b200_set_box(c, box); });
}
void B200UpdateStateDataKernel::createCheckpoint(ContextImpl& context, ostream& stream) {
    // platform-specific binary state, as in CommonKernels.cpp:354-382 (version, time, step, positions, velocities, box)
    const int version = 1;
    stream.write((const char*)&version, sizeof(int));
    double t; int64_t s;
    check(b200_get_time(d.ctx, &t, &s));
    stream.write((const char*)&t, sizeof(double)); stream.write((const char*)&s, sizeof(int64_t));
    vector<double> p(3 * (size_t)numParticles), v(p.size());
    d.all([&](b200_ctx* c, int r) { return b200_get_positions(c, r == 0 ? p.data() : nullptr, 0); });
    d.all([&](b200_ctx* c, int r) { return b200_get_velocities(c, r == 0 ? v.data() : nullptr); });
    stream.write((const char*)p.data(), sizeof(double) * p.size()); stream.write((const char*)v.data(), sizeof(double) * v.size());
    double box[9]; check(b200_get_box(d.ctx, box));
    stream.write((const char*)box, sizeof(box));
}
void B200UpdateStateDataKernel::loadCheckpoint(ContextImpl& context, istream& stream) {
    int version = 0;
    stream.read((char*)&version, sizeof(int));
    if (version != 1) throw OpenMMException("Checkpoint was created with a different version of OpenMM");
    double t; int64_t s;
    stream.read((char*)&t, sizeof(double)); stream.read((char*)&s, sizeof(int64_t));
    vector<double> p(3 * (size_t)numParticles), v(p.size());
    stream.read((char*)p.data(), sizeof(double) * p.size()); stream.read((char*)v.data(), sizeof(double) * v.size());
    double box[9]; stream.read((char*)box, sizeof(box));
    d.all([&](b200_ctx* c, int) { int rc = b200_set_box(c, box); if (!rc) rc = b200_set_positions(c, p.data()); if (!rc) rc = b200_set_velocities(c, v.data());
                                  if (!rc) rc = b200_set_time(c, t, s); return rc; });
}

void B200MinimizeKernel::execute(ContextImpl& context, double tolerance, int maxIterations, MinimizationReporter* reporter) {
    ReferenceMinimize::minimize(context, tolerance, maxIterations, reporter);
}

// ---------------------------------------------------------------------------------------------- NonbondedForce
struct B200CalcNonbondedForceKernel::Params {
    vector<double> q, sigma, eps;
    vector<int> exclusions, exceptions;
    vector<double> excQQ, excSigma, excEps;
    b200_nonbonded_desc desc;
};

// Effective parameters = base parameters + sum over offsets of scale * (current value of the global parameter)
// (NonbondedForce::addParticleParameterOffset / addExceptionParameterOffset; CommonCalcNonbondedForce.cpp:653-704 and the
// computeParameters kernel nonbondedParameters.cc:1-76 do this on the device; here it is a host loop that runs only when a
// global parameter actually changed).
void B200CalcNonbondedForceKernel::gather(const System& system, const NonbondedForce& force, Params& p, const map<string, double>& globals) const {
    const int n = force.getNumParticles();
    p.q.resize(n); p.sigma.resize(n); p.eps.resize(n);
    for (int i = 0; i < n; i++) force.getParticleParameters(i, p.q[i], p.sigma[i], p.eps[i]);
    auto value = [&](const string& name) { auto it = globals.find(name); if (it == globals.end()) throw OpenMMException("NonbondedForce: unknown global parameter '" + name + "'"); return it->second; };
    for (int i = 0; i < force.getNumParticleParameterOffsets(); i++) {
        string name; int particle; double cq, cs, ce;
        force.getParticleParameterOffset(i, name, particle, cq, cs, ce);
        const double v = value(name);
        p.q[particle] += cq * v; p.sigma[particle] += cs * v; p.eps[particle] += ce * v;
    }
    // exceptions split as CommonCalcNonbondedForce.cpp:239-250: all of them are exclusions; those with parameters (or with an
    // offset that can give them parameters later) are 1-4 pairs
    const int ne = force.getNumExceptions();
    vector<double> xq(ne), xs(ne), xe(ne); vector<int> xa(ne), xb(ne); vector<char> hasOffset(ne, 0);
    for (int i = 0; i < ne; i++) force.getExceptionParameters(i, xa[i], xb[i], xq[i], xs[i], xe[i]);
    for (int i = 0; i < force.getNumExceptionParameterOffsets(); i++) {
        string name; int exception; double cq, cs, ce;
        force.getExceptionParameterOffset(i, name, exception, cq, cs, ce);
        const double v = value(name);
        xq[exception] += cq * v; xs[exception] += cs * v; xe[exception] += ce * v; hasOffset[exception] = 1;
    }
    for (int i = 0; i < ne; i++) {
        p.exclusions.push_back(xa[i]); p.exclusions.push_back(xb[i]);
        if (xq[i] != 0.0 || xe[i] != 0.0 || hasOffset[i]) { p.exceptions.push_back(xa[i]); p.exceptions.push_back(xb[i]); p.excQQ.push_back(xq[i]); p.excSigma.push_back(xs[i]); p.excEps.push_back(xe[i]); }
    }
    b200_nonbonded_desc& dd = p.desc;
    memset(&dd, 0, sizeof(dd));
    const NonbondedForce::NonbondedMethod method = force.getNonbondedMethod();
    dd.method = (int)method;     // the B200_NB_* values are NonbondedForce::NonbondedMethod's
    const bool ljpme = method == NonbondedForce::LJPME;
    if (ljpme) {   // CommonCalcNonbondedForce.cpp:380-396: dispersion mesh from calcPMEParameters(lj = true), rounded to a legal FFT size
        int lx, ly, lz;
        NonbondedForceImpl::calcPMEParameters(system, force, dd.ljpme_alpha, lx, ly, lz, true);
        dd.ljpme_grid[0] = b200_find_legal_fft_dimension(max(lx, 6)); dd.ljpme_grid[1] = b200_find_legal_fft_dimension(max(ly, 6)); dd.ljpme_grid[2] = b200_find_legal_fft_dimension(max(lz, 6));
    }
    dd.cutoff = force.getCutoffDistance();
    dd.use_switch = force.getUseSwitchingFunction() ? 1 : 0;
    dd.switch_distance = force.getSwitchingDistance();
    dd.reaction_field_dielectric = force.getReactionFieldDielectric();
    if (method == NonbondedForce::PME || ljpme) {
        int nx, ny, nz;
        NonbondedForceImpl::calcPMEParameters(system, force, dd.ewald_alpha, nx, ny, nz, false);
        dd.grid[0] = b200_find_legal_fft_dimension(max(nx, 6)); dd.grid[1] = b200_find_legal_fft_dimension(max(ny, 6)); dd.grid[2] = b200_find_legal_fft_dimension(max(nz, 6));
    }
    else if (method == NonbondedForce::Ewald) {
        int kx, ky, kz;
        NonbondedForceImpl::calcEwaldParameters(system, force, dd.ewald_alpha, kx, ky, kz);     // CommonCalcNonbondedForce.cpp:333-343
        dd.ewald_kmax[0] = kx; dd.ewald_kmax[1] = ky; dd.ewald_kmax[2] = kz;
    }
    // long-range correction only with a periodic cutoff method and never with LJPME (CommonCalcNonbondedForce.cpp:305-312)
    const bool periodic = method == NonbondedForce::CutoffPeriodic || method == NonbondedForce::Ewald || method == NonbondedForce::PME;
    dd.dispersion_coefficient = force.getUseDispersionCorrection() && periodic ? NonbondedForceImpl::calcDispersionCorrection(system, force) : 0.0;
    dd.charge = p.q.data(); dd.sigma = p.sigma.data(); dd.epsilon = p.eps.data();
    dd.num_exclusions = (int)p.exclusions.size() / 2; dd.exclusion_pairs = p.exclusions.data();
    dd.num_exceptions = (int)p.excQQ.size(); dd.exception_pairs = p.exceptions.data();
    dd.exception_chargeprod = p.excQQ.data(); dd.exception_sigma = p.excSigma.data(); dd.exception_epsilon = p.excEps.data();
    dd.exceptions_use_periodic = force.getExceptionsUsePeriodicBoundaryConditions() ? 1 : 0;
}

void B200CalcNonbondedForceKernel::initialize(const System& system, const NonbondedForce& force) {
    forceRef = &force; systemRef = &system;      // both outlive the kernel (they belong to the user's System)
    hasOffsets = force.getNumParticleParameterOffsets() > 0 || force.getNumExceptionParameterOffsets() > 0;
    globals.clear();
    for (int i = 0; i < force.getNumGlobalParameters(); i++) globals[force.getGlobalParameterName(i)] = force.getGlobalParameterDefaultValue(i);
    Params p;
    gather(system, force, p, globals);
    if (d.hasNonbonded) {
        if (d.ranks.size() > 1) throw OpenMMException("The B200 platform supports a second NonbondedForce in a System on one device only");
        extra = d.addExtraNonbonded();
        check(b200_nonbonded_create(extra->ctx, &p.desc));
        return;
    }
    d.all([&](b200_ctx* c, int) { return b200_nonbonded_create(c, &p.desc); });
    d.hasNonbonded = true;
}

double B200CalcNonbondedForceKernel::execute(ContextImpl& context, bool, bool, bool includeDirect, bool includeReciprocal) {
    if (hasOffsets) {   // a global parameter that moved since the last evaluation changes the effective parameters
        bool changed = false;
        for (auto& g : globals) { const double v = context.getParameter(g.first); if (v != g.second) { g.second = v; changed = true; } }
        if (changed) {
            Params p; gather(*systemRef, *forceRef, p, globals);
            if (extra) check(b200_nonbonded_update_parameters(extra->ctx, &p.desc));
            else d.all([&](b200_ctx* c, int) { return b200_nonbonded_update_parameters(c, &p.desc); });
        }
    }
    unsigned& pending = extra ? extra->pending : d.pendingGroups;
    if (includeDirect) pending |= B200_GROUP_DIRECT;
    if (includeReciprocal) pending |= B200_GROUP_RECIPROCAL;
    return 0.0;   // the energy of this force is returned by CalcForcesAndEnergyKernel::finishComputation
}

void B200CalcNonbondedForceKernel::copyParametersToContext(ContextImpl& context, const NonbondedForce& force, int, int, int, int) {
    Params p;
    for (auto& g : globals) g.second = context.getParameter(g.first);
    gather(context.getSystem(), force, p, globals);
    if (extra) { check(b200_nonbonded_update_parameters(extra->ctx, &p.desc)); return; }
    d.all([&](b200_ctx* c, int) { return b200_nonbonded_update_parameters(c, &p.desc); });
}

void B200CalcNonbondedForceKernel::getPMEParameters(double& alpha, int& nx, int& ny, int& nz) const {
    int32_t x, y, z;
    check(b200_nonbonded_get_pme_parameters(extra ? extra->ctx : d.ctx, &alpha, &x, &y, &z));
    nx = x; ny = y; nz = z;
}
void B200CalcNonbondedForceKernel::getLJPMEParameters(double& alpha, int& nx, int& ny, int& nz) const {
    int32_t x, y, z;
    if (b200_nonbonded_get_ljpme_parameters(extra ? extra->ctx : d.ctx, &alpha, &x, &y, &z) != B200_OK)
        throw OpenMMException("getPMEParametersInContext: This Context is not using PME");   // message of CommonCalcNonbondedForce.cpp:1267
    nx = x; ny = y; nz = z;
}

// ---------------------------------------------------------------------------------------------- bonded forces
// All terms of one kind live in one device list; each force remembers its range of terms.  Forces of the same kind may sit in different
// force groups (the per-force energy decomposition idiom): an evaluation then covers the ranges of the forces whose kernels ran (see
// finishComputation).  They must still agree on the periodic-boundary setting, which is one flag per kind.
static void checkBondedForce(B200Platform::PlatformData& d, int kind, const Force& force, bool periodic, const char* name) {
    if (d.bondedForces[kind] > 0 && d.bondedPeriodic[kind] != (periodic ? 1 : 0))
        throw OpenMMException(string("The B200 platform evaluates all ") + name + " objects of a System together: they must agree on periodic boundary conditions");
    d.bondedPeriodic[kind] = periodic ? 1 : 0;
    d.all([&](b200_ctx* c, int) { return b200_bonded_set_periodic(c, kind, periodic ? 1 : 0); });
    d.bondedForces[kind]++;
}
void B200CalcHarmonicBondForceKernel::initialize(const System&, const HarmonicBondForce& force) {
    checkBondedForce(d, 0, force, force.usesPeriodicBoundaryConditions(), "HarmonicBondForce");
    const int n = force.getNumBonds();
    vector<int> atoms(2 * (size_t)n); vector<double> len(n), k(n);
    for (int i = 0; i < n; i++) force.getBondParameters(i, atoms[2 * i], atoms[2 * i + 1], len[i], k[i]);
    d.all([&](b200_ctx* c, int) { return b200_harmonic_bonds_create(c, n, atoms.data(), len.data(), k.data(), force.getForceGroup()); });
    firstTerm = d.bondedCount[0]; numTerms = n; d.bondedCount[0] += n;
}
void B200CalcHarmonicBondForceKernel::copyParametersToContext(ContextImpl&, const HarmonicBondForce& force, int, int) {
    if (force.getNumBonds() != numTerms) throw OpenMMException("updateParametersInContext: The number of bonds has changed");
    vector<double> len(numTerms), k(numTerms);
    for (int i = 0; i < numTerms; i++) { int a, b; force.getBondParameters(i, a, b, len[i], k[i]); }
    d.all([&](b200_ctx* c, int) { return b200_bonded_update_parameters(c, 0, firstTerm, numTerms, len.data(), k.data(), nullptr); });
}
void B200CalcHarmonicAngleForceKernel::initialize(const System&, const HarmonicAngleForce& force) {
    checkBondedForce(d, 1, force, force.usesPeriodicBoundaryConditions(), "HarmonicAngleForce");
    const int n = force.getNumAngles();
    vector<int> atoms(3 * (size_t)n); vector<double> ang(n), k(n);
    for (int i = 0; i < n; i++) force.getAngleParameters(i, atoms[3 * i], atoms[3 * i + 1], atoms[3 * i + 2], ang[i], k[i]);
    d.all([&](b200_ctx* c, int) { return b200_harmonic_angles_create(c, n, atoms.data(), ang.data(), k.data(), force.getForceGroup()); });
    firstTerm = d.bondedCount[1]; numTerms = n; d.bondedCount[1] += n;
}
void B200CalcHarmonicAngleForceKernel::copyParametersToContext(ContextImpl&, const HarmonicAngleForce& force, int, int) {
    if (force.getNumAngles() != numTerms) throw OpenMMException("updateParametersInContext: The number of angles has changed");
    vector<double> ang(numTerms), k(numTerms);
    for (int i = 0; i < numTerms; i++) { int a, b, c2; force.getAngleParameters(i, a, b, c2, ang[i], k[i]); }
    d.all([&](b200_ctx* c, int) { return b200_bonded_update_parameters(c, 1, firstTerm, numTerms, ang.data(), k.data(), nullptr); });
}
void B200CalcPeriodicTorsionForceKernel::initialize(const System&, const PeriodicTorsionForce& force) {
    checkBondedForce(d, 2, force, force.usesPeriodicBoundaryConditions(), "PeriodicTorsionForce");
    const int n = force.getNumTorsions();
    vector<int> atoms(4 * (size_t)n), per(n); vector<double> phase(n), k(n);
    for (int i = 0; i < n; i++) force.getTorsionParameters(i, atoms[4 * i], atoms[4 * i + 1], atoms[4 * i + 2], atoms[4 * i + 3], per[i], phase[i], k[i]);
    d.all([&](b200_ctx* c, int) { return b200_periodic_torsions_create(c, n, atoms.data(), per.data(), phase.data(), k.data(), force.getForceGroup()); });
    firstTerm = d.bondedCount[2]; numTerms = n; d.bondedCount[2] += n;
}
void B200CalcPeriodicTorsionForceKernel::copyParametersToContext(ContextImpl&, const PeriodicTorsionForce& force, int, int) {
    if (force.getNumTorsions() != numTerms) throw OpenMMException("updateParametersInContext: The number of torsions has changed");
    vector<double> phase(numTerms), k(numTerms); vector<int> per(numTerms);
    for (int i = 0; i < numTerms; i++) { int a, b, c2, e; force.getTorsionParameters(i, a, b, c2, e, per[i], phase[i], k[i]); }
    d.all([&](b200_ctx* c, int) { return b200_bonded_update_parameters(c, 2, firstTerm, numTerms, phase.data(), k.data(), per.data()); });
}

// ---------------------------------------------------------------------------------------------- Monte Carlo barostat
void B200ApplyMonteCarloBarostatKernel::initialize(const System& system, const Force&, int comps, bool rigidMolecules) {
    components = comps; rigid = rigidMolecules; numParticles = system.getNumParticles();
    masses.resize(numParticles);
    for (int i = 0; i < numParticles; i++) masses[i] = system.getParticleMass(i);
}
void B200ApplyMonteCarloBarostatKernel::computeKineticEnergy(ContextImpl& context, vector<double>& ke) {
    // kinetic energy of the molecular centres of mass (rigid molecules) or of all particles, split as monteCarloBarostat.cc does:
    // 1 component = total, 3 = per axis, 6 = all elements of the box tensor.  A cold path (instantaneous pressure): done on the host.
    vector<double> v(3 * (size_t)numParticles);
    d.all([&](b200_ctx* c, int r) { return b200_get_velocities(c, r == 0 ? v.data() : nullptr); });
    vector<vector<int> > molecules;
    if (rigid) molecules = context.getMolecules();
    else { molecules.resize(numParticles); for (int i = 0; i < numParticles; i++) molecules[i].push_back(i); }
    ke.assign(components, 0.0);
    for (const auto& mol : molecules) {
        double m = 0, p[3] = {0, 0, 0};
        for (int a : mol) { m += masses[a]; for (int k = 0; k < 3; k++) p[k] += masses[a] * v[3 * (size_t)a + k]; }
        if (m == 0) continue;
        const double vx = p[0] / m, vy = p[1] / m, vz = p[2] / m;
        if (components == 1) ke[0] += 0.5 * m * (vx * vx + vy * vy + vz * vz);
        else {
            ke[0] += 0.5 * m * vx * vx; ke[1] += 0.5 * m * vy * vy; ke[2] += 0.5 * m * vz * vz;
            if (components == 6) { ke[3] += 0.5 * m * vx * vy; ke[4] += 0.5 * m * vx * vz; ke[5] += 0.5 * m * vy * vz; }
        }
    }
}

// ---------------------------------------------------------------------------------------------- integrators
void B200IntegrateLangevinMiddleStepKernel::initialize(const System&, const LangevinMiddleIntegrator& integrator) {
    int s = integrator.getRandomNumberSeed();
    seed = s != 0 ? (unsigned long long)s : (unsigned long long)random_device{}();   // 0 means "pick one" (osrngseed in the reference)
}
void B200IntegrateLangevinMiddleStepKernel::execute(ContextImpl&, const LangevinMiddleIntegrator& integrator) {
    b200_integrator_desc desc = {integrator.getStepSize(), integrator.getTemperature(), integrator.getFriction(), integrator.getConstraintTolerance(), seed, 0};
    d.all([&](b200_ctx* c, int) { const int rc = b200_integrator_create(c, &desc); return rc ? rc : b200_integrate(c); });
}
double B200IntegrateLangevinMiddleStepKernel::computeKineticEnergy(ContextImpl&, const LangevinMiddleIntegrator&) {
    vector<double> ke(max<size_t>(1, d.ranks.size()), 0.0);
    d.all([&](b200_ctx* c, int r) { return b200_kinetic_energy(c, &ke[r]); });
    return ke[0];
}
void B200IntegrateVerletStepKernel::execute(ContextImpl&, const VerletIntegrator& integrator) {
    b200_integrator_desc desc = {integrator.getStepSize(), 0.0, 0.0, integrator.getConstraintTolerance(), 0, 0};
    d.all([&](b200_ctx* c, int) { const int rc = b200_integrator_create(c, &desc); return rc ? rc : b200_integrate(c); });
}
double B200IntegrateVerletStepKernel::computeKineticEnergy(ContextImpl& context, const VerletIntegrator& integrator) {
    // leapfrog: the stored velocities are half a step behind the positions; report the energy of velocities shifted forward by dt/2
    // with the forces that are in the buffer (Context::getState evaluates them first because kineticEnergyRequiresForce() is true;
    // they must not be recomputed here: getState reads them afterwards) and constrained again (CommonKernels.cpp:3117 ->
    // IntegrationUtilities::computeKineticEnergy)
    vector<double> ke(max<size_t>(1, d.ranks.size()), 0.0);
    d.all([&](b200_ctx* c, int r) { return b200_kinetic_energy_shifted(c, 0.5 * integrator.getStepSize(), &ke[r]); });
    return ke[0];
}

}  // namespace B200Plugin
