// oracle/ref_driver.cpp -- TEST / BENCH INFRASTRUCTURE, not product code.
//
// Drives the UNMODIFIED reference (OpenMM 8.6 compiled into oracle/_ref by oracle/Makefile) through its public C++ API.
// It is how golden fixtures are generated (tests/golden/make_fixtures.py), how the B200 plugin is compared with the
// reference's Reference/CPU platforms inside one process (tests/test_plugin.py), and how `bench.py --impl reference`
// times the reference's CPU platform.  It contains no algorithm of its own.
//
//   ref_driver eval  <system.bin> <out.bin> [--platform NAME] [--plugin lib.so]... [--pairs | --pair-hash]
//   ref_driver compare <system.bin> --platform B200 --plugin libOpenMMB200.so [--against Reference]
//   ref_driver bench <system.bin> --platform NAME [--plugin lib.so]... --steps K --warmup W [--threads T] [--dt ps] [--e2e] [--prop Key=Value]...
//
// system.bin layout (little endian), written by openmm_b200/system_io.py:
//   char[4] "B2SY"; int32 version=1; int32 N; double box[9]; double cutoff; double alpha; int32 grid[3]; int32 method;
//   double dispersionCorrection(0/1); [version >= 2: double ljAlpha; int32 ljGrid[3];] [version >= 3: double switchDistance;] double charge[N], sigma[N], epsilon[N], mass[N]; double pos[3N]; double vel[3N];
//   int32 nexc; {int32 i, j; double qq, sigma, eps}[nexc]; int32 ncon; {int32 i, j; double d}[ncon]
#include "OpenMM.h"
#include "openmm/internal/ContextImpl.h"
#include "ReferenceNeighborList.h"
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

using namespace OpenMM;
using namespace std;

struct SystemFile {
    int N = 0; double box[9]; double cutoff = 0, alpha = 0; int grid[3]; int method = 4; double dispersion = 0; double ljAlpha = 0; int ljGrid[3] = {0, 0, 0}; double switchDistance = -1;
    vector<double> charge, sigma, eps, mass, pos, vel;
    struct Exc { int i, j; double qq, sig, eps; }; vector<Exc> exc;
    struct Con { int i, j; double d; }; vector<Con> con;
};

template <typename T> static void rd(ifstream& f, T* p, size_t n) { f.read((char*)p, sizeof(T) * n); if (!f) throw runtime_error("system file truncated"); }

static SystemFile readSystem(const string& path) {
    ifstream f(path, ios::binary);
    if (!f) throw runtime_error("cannot open " + path);
    char magic[4]; rd(f, magic, 4);
    if (memcmp(magic, "B2SY", 4) != 0) throw runtime_error("bad magic in " + path);
    int version; rd(f, &version, 1);
    SystemFile s;
    rd(f, &s.N, 1); rd(f, s.box, 9); rd(f, &s.cutoff, 1); rd(f, &s.alpha, 1); rd(f, s.grid, 3); rd(f, &s.method, 1); rd(f, &s.dispersion, 1);
    if (version >= 2) { rd(f, &s.ljAlpha, 1); rd(f, s.ljGrid, 3); }
    if (version >= 3) rd(f, &s.switchDistance, 1);
    const size_t N = s.N;
    s.charge.resize(N); s.sigma.resize(N); s.eps.resize(N); s.mass.resize(N); s.pos.resize(3 * N); s.vel.resize(3 * N);
    rd(f, s.charge.data(), N); rd(f, s.sigma.data(), N); rd(f, s.eps.data(), N); rd(f, s.mass.data(), N); rd(f, s.pos.data(), 3 * N); rd(f, s.vel.data(), 3 * N);
    int nexc; rd(f, &nexc, 1); s.exc.resize(nexc);
    for (auto& e : s.exc) { rd(f, &e.i, 1); rd(f, &e.j, 1); rd(f, &e.qq, 1); rd(f, &e.sig, 1); rd(f, &e.eps, 1); }
    int ncon; rd(f, &ncon, 1); s.con.resize(ncon);
    for (auto& c : s.con) { rd(f, &c.i, 1); rd(f, &c.j, 1); rd(f, &c.d, 1); }
    return s;
}

static System* buildSystem(const SystemFile& s, bool cmRemover) {
    System* sys = new System();
    sys->setDefaultPeriodicBoxVectors(Vec3(s.box[0], s.box[1], s.box[2]), Vec3(s.box[3], s.box[4], s.box[5]), Vec3(s.box[6], s.box[7], s.box[8]));
    for (int i = 0; i < s.N; i++) sys->addParticle(s.mass[i]);
    for (auto& c : s.con) sys->addConstraint(c.i, c.j, c.d);
    NonbondedForce* nb = new NonbondedForce();
    nb->setNonbondedMethod((NonbondedForce::NonbondedMethod)s.method);
    nb->setCutoffDistance(s.cutoff);
    nb->setPMEParameters(s.alpha, s.grid[0], s.grid[1], s.grid[2]);   // identical grid on every platform (SURVEY 8c gotcha 1)
    if (s.method == 5) nb->setLJPMEParameters(s.ljAlpha, s.ljGrid[0], s.ljGrid[1], s.ljGrid[2]);
    if (s.switchDistance >= 0) { nb->setUseSwitchingFunction(true); nb->setSwitchingDistance(s.switchDistance); }
    nb->setUseDispersionCorrection(s.dispersion != 0);
    nb->setReciprocalSpaceForceGroup(1);                              // tests/TestNonbondedForce.h:969-1011 testDirectAndReciprocal
    for (int i = 0; i < s.N; i++) nb->addParticle(s.charge[i], s.sigma[i], s.eps[i]);
    for (auto& e : s.exc) nb->addException(e.i, e.j, e.qq, e.sig, e.eps);
    sys->addForce(nb);
    if (cmRemover) sys->addForce(new CMMotionRemover(1));
    return sys;
}

static vector<Vec3> toVec3(const vector<double>& a) { vector<Vec3> v(a.size() / 3); for (size_t i = 0; i < v.size(); i++) v[i] = Vec3(a[3 * i], a[3 * i + 1], a[3 * i + 2]); return v; }

static Platform& pickPlatform(const string& name, const vector<string>& plugins) {
    for (auto& p : plugins) Platform::loadPluginLibrary(p);
    return Platform::getPlatformByName(name);
}

struct Eval { double eDirect, eRecip; vector<Vec3> fDirect, fRecip; };

static Eval evaluate(const SystemFile& s, Platform& platform, const map<string, string>& props) {
    System* sys = buildSystem(s, false);
    VerletIntegrator integ(0.001);
    Context ctx(*sys, integ, platform, props);
    ctx.setPositions(toVec3(s.pos));
    Eval e;
    State s0 = ctx.getState(State::Forces | State::Energy, false, 1 << 0);
    State s1 = ctx.getState(State::Forces | State::Energy, false, 1 << 1);
    e.eDirect = s0.getPotentialEnergy(); e.eRecip = s1.getPotentialEnergy();
    e.fDirect = s0.getForces(); e.fRecip = s1.getForces();
    delete sys;
    return e;
}

static int cmdEval(int argc, char** argv) {
    string sysPath = argv[2], outPath = argv[3], platformName = "Reference"; vector<string> plugins; bool pairs = false, pairHash = false;
    for (int i = 4; i < argc; i++) {
        string a = argv[i];
        if (a == "--platform") platformName = argv[++i]; else if (a == "--plugin") plugins.push_back(argv[++i]); else if (a == "--pairs") pairs = true;
        else if (a == "--pair-hash") pairs = pairHash = true;   // count + order-independent hash only (million-atom systems: the list itself is > 1 GB)
    }
    SystemFile s = readSystem(sysPath);
    Platform& platform = pickPlatform(platformName, plugins);
    Eval e = evaluate(s, platform, {});
    ofstream out(outPath, ios::binary);
    out.write("B2EV", 4);
    out.write((char*)&s.N, 4); out.write((char*)&e.eDirect, 8); out.write((char*)&e.eRecip, 8);
    for (auto* fv : {&e.fDirect, &e.fRecip}) for (auto& f : *fv) { double v[3] = {f[0], f[1], f[2]}; out.write((char*)v, 24); }
    long long npairs = -1;
    if (pairs) {
        // the reference's own pair set: computeNeighborListVoxelHash with d^2 <= rc^2 (ReferenceNeighborList.cpp:208-272)
        vector<set<int>> excl(s.N);
        for (auto& x : s.exc) { excl[x.i].insert(x.j); excl[x.j].insert(x.i); }
        NeighborList nl;
        Vec3 bv[3] = {Vec3(s.box[0], s.box[1], s.box[2]), Vec3(s.box[3], s.box[4], s.box[5]), Vec3(s.box[6], s.box[7], s.box[8])};
        computeNeighborListVoxelHash(nl, s.N, toVec3(s.pos), excl, bv, true, s.cutoff, 0.0, false);
        npairs = (long long)nl.size();
        if (pairHash) {
            // same hash as tests/oracle_lib.py pair_hash: sum over pairs of splitmix64(lo*N + hi) mod 2^64
            unsigned long long h = 0;
            for (auto& p : nl) {
                unsigned long long lo = min(p.first, p.second), hi = max(p.first, p.second);
                unsigned long long z = lo * (unsigned long long)s.N + hi + 0x9E3779B97F4A7C15ULL;
                z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL; z ^= z >> 31;
                h += z;
            }
            long long marker = -2 - npairs;      // < -1: "hash follows instead of the list"; count = -2 - marker
            out.write((char*)&marker, 8); out.write((char*)&h, 8);
        } else {
            out.write((char*)&npairs, 8);
            for (auto& p : nl) { int a = min((int)p.first, (int)p.second), b = max((int)p.first, (int)p.second); out.write((char*)&a, 4); out.write((char*)&b, 4); }
        }
    } else out.write((char*)&npairs, 8);
    printf("{\"platform\": \"%s\", \"n\": %d, \"e_direct\": %.10g, \"e_recip\": %.10g, \"pairs\": %lld}\n", platformName.c_str(), s.N, e.eDirect, e.eRecip, npairs);
    return 0;
}

static double relRms(const vector<Vec3>& a, const vector<Vec3>& ref) {
    double num = 0, den = 0;
    for (size_t i = 0; i < a.size(); i++) { Vec3 d = a[i] - ref[i]; num += d.dot(d); den += ref[i].dot(ref[i]); }
    return sqrt(num / den);
}

static int cmdCompare(int argc, char** argv) {
    string sysPath = argv[2], platformName = "B200", against = "Reference"; vector<string> plugins; map<string, string> props;
    for (int i = 3; i < argc; i++) {
        string a = argv[i];
        if (a == "--platform") platformName = argv[++i]; else if (a == "--plugin") plugins.push_back(argv[++i]); else if (a == "--against") against = argv[++i];
        else if (a == "--prop") { string kv = argv[++i]; size_t eq = kv.find('='); if (eq != string::npos) props[kv.substr(0, eq)] = kv.substr(eq + 1); }
    }
    SystemFile s = readSystem(sysPath);
    Platform& p = pickPlatform(platformName, plugins);
    Platform& r = Platform::getPlatformByName(against);
    Eval a = evaluate(s, p, props), b = evaluate(s, r, {});
    vector<Vec3> ta(s.N), tb(s.N);
    for (int i = 0; i < s.N; i++) { ta[i] = a.fDirect[i] + a.fRecip[i]; tb[i] = b.fDirect[i] + b.fRecip[i]; }
    printf("{\"platform\": \"%s\", \"against\": \"%s\", \"n\": %d, \"rel_rms_force_total\": %.3e, \"rel_rms_force_direct\": %.3e, \"rel_rms_force_recip\": %.3e, "
           "\"e_direct\": %.10g, \"e_direct_ref\": %.10g, \"e_recip\": %.10g, \"e_recip_ref\": %.10g, \"rel_energy\": %.3e}\n",
           platformName.c_str(), against.c_str(), s.N, relRms(ta, tb), relRms(a.fDirect, b.fDirect), relRms(a.fRecip, b.fRecip),
           a.eDirect, b.eDirect, a.eRecip, b.eRecip, fabs((a.eDirect + a.eRecip) - (b.eDirect + b.eRecip)) / fabs(b.eDirect + b.eRecip));
    return 0;
}

static int cmdBench(int argc, char** argv) {
    string sysPath = argv[2], platformName = "CPU"; vector<string> plugins; int steps = 10, warmup = 3, threads = 0; double dt = 0.002; bool e2e = false;
    map<string, string> extraProps;
    for (int i = 3; i < argc; i++) {
        string a = argv[i];
        if (a == "--platform") platformName = argv[++i]; else if (a == "--plugin") plugins.push_back(argv[++i]);
        else if (a == "--steps") steps = atoi(argv[++i]); else if (a == "--warmup") warmup = atoi(argv[++i]);
        else if (a == "--threads") threads = atoi(argv[++i]); else if (a == "--dt") dt = atof(argv[++i]); else if (a == "--e2e") e2e = true;
        else if (a == "--prop") { string kv = argv[++i]; size_t eq = kv.find('='); if (eq != string::npos) extraProps[kv.substr(0, eq)] = kv.substr(eq + 1); }
    }
    SystemFile s = readSystem(sysPath);
    Platform& platform = pickPlatform(platformName, plugins);
    map<string, string> props = extraProps;
    if (threads > 0 && platformName == "CPU") props["Threads"] = to_string(threads);
    System* sys = buildSystem(s, true);
    // examples/benchmarks/benchmark.py:314-337: LangevinMiddle 300 K, 1/ps, constraint tolerance 1e-5
    LangevinMiddleIntegrator integ(300.0, 1.0, dt);
    integ.setConstraintTolerance(1e-5);
    integ.setRandomNumberSeed(1);
    auto t0 = chrono::steady_clock::now();
    Context ctx(*sys, integ, platform, props);
    ctx.setPositions(toVec3(s.pos));
    ctx.setVelocities(toVec3(s.vel));
    double setup = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
    integ.step(warmup);
    ctx.getState(State::Energy);
    vector<Vec3> pos;
    t0 = chrono::steady_clock::now();
    if (!e2e) { integ.step(steps); ctx.getState(State::Energy); }   // benchmark.py:99-108
    else {                                                           // host buffers in and out every step: setPositions + step + getState(Positions)
        State st = ctx.getState(State::Positions);               // (the State is kept alive: no extra copy of its position vector)
        for (int i = 0; i < steps; i++) {
            ctx.setPositions(st.getPositions());
            integ.step(1);
            st = ctx.getState(State::Positions);
        }
    }
    double secs = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
    double e = ctx.getState(State::Energy).getPotentialEnergy();
    string usedThreads = platformName == "CPU" ? platform.getPropertyValue(ctx, "Threads") : "0";
    printf("{\"platform\": \"%s\", \"n\": %d, \"steps\": %d, \"warmup\": %d, \"dt_ps\": %g, \"seconds\": %.6f, \"ms_per_step\": %.6f, \"ns_per_day\": %.6f, "
           "\"threads\": %s, \"setup_s\": %.3f, \"potential_energy\": %.8g, \"e2e\": %s}\n",
           platformName.c_str(), s.N, steps, warmup, dt, secs, 1e3 * secs / steps, dt * 1e-3 * 86400.0 * steps / secs, usedThreads.c_str(), setup, e, e2e ? "true" : "false");
    delete sys;
    return 0;
}

// createCheckpoint / loadCheckpoint round trip (UpdateStateDataKernel, kernels.h:239-245): run A steps, checkpoint, run B steps and
// remember the state; load the checkpoint, run B steps again: both runs must agree (bit for bit on a deterministic platform).
static int cmdCheckpoint(int argc, char** argv) {
    string sysPath = argv[2], platformName = "B200"; vector<string> plugins; int stepsA = 20, stepsB = 30;
    for (int i = 3; i < argc; i++) {
        string a = argv[i];
        if (a == "--platform") platformName = argv[++i]; else if (a == "--plugin") plugins.push_back(argv[++i]);
        else if (a == "--steps") { stepsA = atoi(argv[++i]); stepsB = stepsA; }
    }
    SystemFile s = readSystem(sysPath);
    Platform& platform = pickPlatform(platformName, plugins);
    System* sys = buildSystem(s, true);
    LangevinMiddleIntegrator integ(300.0, 1.0, 0.002);
    integ.setConstraintTolerance(1e-5);
    integ.setRandomNumberSeed(11);
    Context ctx(*sys, integ, platform);
    ctx.setPositions(toVec3(s.pos)); ctx.setVelocities(toVec3(s.vel));
    integ.step(stepsA);
    stringstream chk(ios::in | ios::out | ios::binary);
    ctx.createCheckpoint(chk);
    const double tChk = ctx.getState(0).getTime();
    integ.step(stepsB);
    State a = ctx.getState(State::Positions | State::Velocities);
    chk.seekg(0);
    ctx.loadCheckpoint(chk);
    const double tLoaded = ctx.getState(0).getTime();
    integ.step(stepsB);
    State b = ctx.getState(State::Positions | State::Velocities);
    double dp = 0, dv = 0;
    for (int i = 0; i < s.N; i++) {
        Vec3 d = a.getPositions()[i] - b.getPositions()[i], e = a.getVelocities()[i] - b.getVelocities()[i];
        dp = max(dp, sqrt(d.dot(d))); dv = max(dv, sqrt(e.dot(e)));
    }
    printf("{\"platform\": \"%s\", \"n\": %d, \"max_position_difference\": %.3e, \"max_velocity_difference\": %.3e, \"time_restored\": %s, \"final_time_equal\": %s}\n",
           platformName.c_str(), s.N, dp, dv, fabs(tChk - tLoaded) < 1e-12 ? "true" : "false", a.getTime() == b.getTime() ? "true" : "false");
    delete sys;
    return 0;
}

// MonteCarloBarostat on a deterministic integrator (Verlet): the sequence of box volumes depends only on the barostat's own random
// stream (seeded) and on the energies before/after each trial move, so two platforms that agree on energies produce the same sequence.
static int cmdNpt(int argc, char** argv) {
    string sysPath = argv[2], platformName = "Reference"; vector<string> plugins; int steps = 40, frequency = 2;
    for (int i = 3; i < argc; i++) {
        string a = argv[i];
        if (a == "--platform") platformName = argv[++i]; else if (a == "--plugin") plugins.push_back(argv[++i]);
        else if (a == "--steps") steps = atoi(argv[++i]); else if (a == "--frequency") frequency = atoi(argv[++i]);
    }
    SystemFile s = readSystem(sysPath);
    Platform& platform = pickPlatform(platformName, plugins);
    System* sys = buildSystem(s, false);
    MonteCarloBarostat* baro = new MonteCarloBarostat(1.0, 300.0, frequency);
    baro->setRandomNumberSeed(7);
    sys->addForce(baro);
    VerletIntegrator integ(0.001);
    integ.setConstraintTolerance(1e-6);
    Context ctx(*sys, integ, platform);
    ctx.setPositions(toVec3(s.pos)); ctx.setVelocities(toVec3(s.vel));
    ctx.applyConstraints(1e-6);
    printf("{\"platform\": \"%s\", \"n\": %d, \"volumes\": [", platformName.c_str(), s.N);
    for (int i = 0; i < steps; i += frequency) {
        integ.step(frequency);
        Vec3 a, b, c; ctx.getState(0).getPeriodicBoxVectors(a, b, c);
        printf("%s%.12g", i ? ", " : "", a[0] * b[1] * c[2]);
    }
    State st = ctx.getState(State::Positions | State::Energy, false);   // unwrapped: comparable between platforms
    double cx = 0, cy = 0, cz = 0;
    for (auto& p : st.getPositions()) { cx += p[0]; cy += p[1]; cz += p[2]; }
    printf("], \"potential_energy\": %.10g, \"position_sum\": [%.8g, %.8g, %.8g]}\n", st.getPotentialEnergy(), cx, cy, cz);
    delete sys;
    return 0;
}

int main(int argc, char** argv) {
    try {
        if (argc >= 4 && string(argv[1]) == "eval") return cmdEval(argc, argv);
        if (argc >= 3 && string(argv[1]) == "compare") return cmdCompare(argc, argv);
        if (argc >= 3 && string(argv[1]) == "bench") return cmdBench(argc, argv);
        if (argc >= 3 && string(argv[1]) == "checkpoint") return cmdCheckpoint(argc, argv);
        if (argc >= 3 && string(argv[1]) == "npt") return cmdNpt(argc, argv);
        fprintf(stderr, "usage: ref_driver eval|compare|bench ...\n");
        return 2;
    } catch (const exception& e) {
        fprintf(stderr, "ref_driver: %s\n", e.what());
        return 1;
    }
}
