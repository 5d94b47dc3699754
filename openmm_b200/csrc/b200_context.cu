// b200_context.cu -- context life cycle, state I/O, molecule/exclusion-tile set-up and the per-step orchestration.
//
// Host-side counterpart of CudaContext / ComputeContext (platforms/cuda/src/CudaContext.cpp,
// platforms/common/src/ComputeContext.cpp), CommonUpdateStateDataKernel (platforms/common/src/CommonKernels.cpp:95-297),
// CudaNonbondedUtilities::initialize (platforms/cuda/src/CudaNonbondedUtilities.cpp:161-371) and
// CudaCalcForcesAndEnergyKernel (platforms/cuda/src/CudaKernels.cpp:63-91).
#include "b200_internal.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <numeric>
#include <set>
#include <sstream>

static thread_local std::string g_lastError;
void b200_set_error(const std::string& msg) { g_lastError = msg; }
extern "C" const char* b200_last_error(void) { return g_lastError.c_str(); }

#define API_BEGIN try {
#define API_END } catch (const B200Error& e) { b200_set_error(e.what()); return e.code; } \
                  catch (const std::exception& e) { b200_set_error(e.what()); return B200_ERR_INVALID; } return B200_OK;

ScopedTimer::ScopedTimer(b200_ctx* c_, int slot_, cudaStream_t s_) : c(c_), slot(slot_), s(s_) {
    if (!c->timing) return;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a, s);
}
ScopedTimer::~ScopedTimer() {
    if (!c->timing || !a) return;
    cudaEventRecord(b, s);
    c->timers[slot].pending.emplace_back(a, b);
}

// ---------------------------------------------------------------------------------------------
void b200_box_set(b200_ctx* c, const double* box) {
    // reduced form is required (System::setDefaultPeriodicBoxVectors enforces it upstream)
    if (box[1] != 0 || box[2] != 0 || box[5] != 0) throw B200Error(B200_ERR_INVALID, "periodic box vectors must be in reduced form");
    if (!(box[0] > 0 && box[4] > 0 && box[8] > 0)) throw B200Error(B200_ERR_INVALID, "periodic box vectors must have positive diagonal");
    memcpy(c->boxHost, box, sizeof(double) * 9);
    BoxD d; d.ax = box[0]; d.bx = box[3]; d.by = box[4]; d.cx = box[6]; d.cy = box[7]; d.cz = box[8];
    d.iax = 1.0 / d.ax; d.iby = 1.0 / d.by; d.icz = 1.0 / d.cz;
    c->boxD = d;
    BoxF f; f.ax = (float)d.ax; f.bx = (float)d.bx; f.by = (float)d.by; f.cx = (float)d.cx; f.cy = (float)d.cy; f.cz = (float)d.cz;
    f.iax = (float)d.iax; f.iby = (float)d.iby; f.icz = (float)d.icz;
    f.triclinic = (d.bx != 0 || d.cx != 0 || d.cy != 0) ? 1 : 0;
    c->boxF = f;
    if (c->hasNonbonded && c->nb.method == B200_NB_EWALD && f.triclinic)
        throw B200Error(B200_ERR_UNSUPPORTED, "NonbondedForce::Ewald needs a rectangular periodic box (use PME for triclinic cells)");
    if (c->hasNonbonded && c->nb.method >= B200_NB_CUTOFF_PERIODIC) {
        // CudaNonbondedUtilities.cpp:391-396: the box must be at least twice the cutoff
        const double m = std::min(d.ax, std::min(d.by, d.cz));
        if (m < 2 * c->nb.cutoff) throw B200Error(B200_ERR_INVALID, "The periodic box size has decreased to less than twice the nonbonded cutoff.");
    }
}

extern "C" int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { b200_set_error("cudaGetDeviceCount failed"); return B200_ERR_CUDA; }
    return n;
}

extern "C" int b200_context_create(const b200_system_desc* desc, b200_ctx** out) {
    API_BEGIN
    if (!desc || !out || desc->num_atoms <= 0) throw B200Error(B200_ERR_INVALID, "b200_context_create: bad arguments");
    int ndev = 0;
    CUDA_CHECK(cudaGetDeviceCount(&ndev));
    if (desc->device < 0 || desc->device >= ndev) throw B200Error(B200_ERR_CUDA, "b200_context_create: no such CUDA device");
    CUDA_CHECK(cudaSetDevice(desc->device));
    cudaDeviceProp prop;
    CUDA_CHECK(cudaGetDeviceProperties(&prop, desc->device));
    if (prop.major < 10) throw B200Error(B200_ERR_UNSUPPORTED, "the B200 platform needs a compute capability 10.x device");
    b200_ctx* c = new b200_ctx();
    *out = nullptr;
    struct Guard { b200_ctx* c; ~Guard() { if (c) b200_context_destroy(c); } } guard{c};    // a failure below must not leak the half-built context
    c->device = desc->device; c->smCount = prop.multiProcessorCount;
    c->N = desc->num_atoms; c->Npad = ((c->N + 31) / 32) * 32; c->B = c->Npad / 32; c->numSuper = (c->B + 31) / 32;
    c->usePmeStream = !(desc->flags & B200_CTX_DISABLE_PME_STREAM);
    CUDA_CHECK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CUDA_CHECK(cudaStreamCreateWithFlags(&c->pmeStream, cudaStreamNonBlocking));
    CUDA_CHECK(cudaStreamCreateWithFlags(&c->auxStream, cudaStreamNonBlocking));
    CUDA_CHECK(cudaEventCreateWithFlags(&c->evFork, cudaEventDisableTiming));
    CUDA_CHECK(cudaEventCreateWithFlags(&c->evPmeDone, cudaEventDisableTiming));
    memset(c->partTable.owner, 0, sizeof(c->partTable.owner));
    b200_box_set(c, desc->box);
    c->mass.assign(desc->mass, desc->mass + c->N);
    c->numConstraints = desc->num_constraints;
    if (desc->num_constraints > 0) {
        c->conAtoms.assign(desc->constraint_atoms, desc->constraint_atoms + 2 * desc->num_constraints);
        c->conDist.assign(desc->constraint_dist, desc->constraint_dist + desc->num_constraints);
    }
    const size_t P = c->Npad;
    dmalloc(c->posq, P); dmalloc(c->posqCorr, P); dmalloc(c->velm, P); dmalloc(c->sigEps, P); dmalloc(c->atomIndex, P); dmalloc(c->cellOffsets, P);
    dmalloc(c->force, 3 * P); dmalloc(c->counters, 1);
    dmalloc(c->posqTmp, P); dmalloc(c->posqCorrTmp, P); dmalloc(c->velmTmp, P); dmalloc(c->atomIndexTmp, P); dmalloc(c->cellOffsetsTmp, P);
    dmalloc(c->blockCenter, (size_t)c->B); dmalloc(c->blockHalf, (size_t)c->B); dmalloc(c->superCenter, (size_t)c->numSuper); dmalloc(c->superHalf, (size_t)c->numSuper);
    dmalloc(c->refPosq, P);
    CUDA_CHECK(cudaMallocHost((void**)&c->countersHost, sizeof(DevCounters)));
    CUDA_CHECK(cudaMemsetAsync(c->posq, 0, sizeof(float4) * P, c->stream));
    CUDA_CHECK(cudaMemsetAsync(c->posqCorr, 0, sizeof(float4) * P, c->stream));
    CUDA_CHECK(cudaMemsetAsync(c->sigEps, 0, sizeof(float2) * P, c->stream));
    CUDA_CHECK(cudaMemsetAsync(c->cellOffsets, 0, sizeof(int4) * P, c->stream));
    CUDA_CHECK(cudaMemsetAsync(c->force, 0, sizeof(long long) * 3 * P, c->stream));
    CUDA_CHECK(cudaMemsetAsync(c->counters, 0, sizeof(DevCounters), c->stream));
    CUDA_CHECK(cudaMemsetAsync(c->refPosq, 0, sizeof(float4) * P, c->stream));
    std::vector<double4> v(P);
    for (size_t i = 0; i < P; i++) { v[i].x = v[i].y = v[i].z = 0; v[i].w = (i < (size_t)c->N && c->mass[i] > 0) ? 1.0 / c->mass[i] : 0.0; }
    CUDA_CHECK(cudaMemcpyAsync(c->velm, v.data(), sizeof(double4) * P, cudaMemcpyHostToDevice, c->stream));
    c->hostAtomIndex.resize(P); std::iota(c->hostAtomIndex.begin(), c->hostAtomIndex.end(), 0);
    c->hostSlotOf = c->hostAtomIndex;
    CUDA_CHECK(cudaMemcpyAsync(c->atomIndex, c->hostAtomIndex.data(), sizeof(int) * P, cudaMemcpyHostToDevice, c->stream));
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    guard.c = nullptr;
    *out = c;
    API_END
}

extern "C" void b200_context_destroy(b200_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    if (c->stepGraph) { cudaGraphExecDestroy(c->stepGraph); c->stepGraph = nullptr; }
    b200_comm_destroy(c);
    b200_pme_destroy(c);
    dfree(c->posq); dfree(c->posqCorr); dfree(c->velm); dfree(c->sigEps); dfree(c->atomIndex); dfree(c->cellOffsets); dfree(c->force); dfree(c->counters);
    dfree(c->posqTmp); dfree(c->posqCorrTmp); dfree(c->velmTmp); dfree(c->atomIndexTmp); dfree(c->cellOffsetsTmp);
    dfree(c->blockCenter); dfree(c->blockHalf); dfree(c->superCenter); dfree(c->superHalf); dfree(c->refPosq);
    dfree(c->units); dfree(c->jAtoms); dfree(c->pairOut);
    dfree(c->settleAtoms); dfree(c->settleParams); dfree(c->freeAtoms); dfree(c->shakeAtoms); dfree(c->shakeDist);
    dfree(c->gcPairs); dfree(c->gcDist); dfree(c->gcColorStart); dfree(c->gcClusterColors); dfree(c->gcAtomStart); dfree(c->gcAtomList); dfree(c->gcDelta); dfree(c->gcOld);
    dfree(c->exclTiles); dfree(c->exclFlags); dfree(c->exclRowStart); dfree(c->exclRowCols); dfree(c->dExclPairs); dfree(c->dExcPairs); dfree(c->dExcParams);
    dfree(c->bondAtoms); dfree(c->bondParams); dfree(c->angleAtoms); dfree(c->angleParams); dfree(c->torsionAtoms); dfree(c->torsionParams);
    dfree(c->molStart); dfree(c->molSlots); dfree(c->savedPosq); dfree(c->savedPosqCorr); dfree(c->savedVelm); dfree(c->savedAtomIndex); dfree(c->savedCellOffsets);
    dfree(c->ewaldS); dfree(c->sortKeys); dfree(c->sortCount); dfree(c->sortStart); dfree(c->sortOrder);
    for (auto& g : c->groups) dfree(g.dBase);
    if (c->countersHost) cudaFreeHost(c->countersHost);
    if (c->hStage) cudaFreeHost(c->hStage);
    for (auto& e : c->ioEvents) cudaEventDestroy(e);
    dfree(c->dStage);
    for (auto& t : c->timers) for (auto& p : t.pending) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    cudaEventDestroy(c->evFork); cudaEventDestroy(c->evPmeDone);
    cudaStreamDestroy(c->stream); cudaStreamDestroy(c->pmeStream); cudaStreamDestroy(c->auxStream);
    delete c;
}

// LJPME: dispersion self term + alpha_d^6/12 * sum_i c6_i^2 with c6_i = 8 (sigma_i/2)^3 2 sqrt(eps_i)
// (ReferenceLJCoulombIxn.cpp:224-227; CommonCalcNonbondedForce.cpp:401-406)
static void b200_lj_self_energy(b200_ctx* c) {
    c->ljSelfEnergy = 0;
    if (!c->ljpme) return;
    const double a6 = pow(c->nb.ljpme_alpha, 6.0);
    for (int i = 0; i < c->N; i++) c->ljSelfEnergy += a6 * 64.0 * pow(0.5 * c->sigma[i], 6.0) * 4.0 * c->epsilon[i] / 12.0;
}

// ---------------------------------------------------------------------------------------------
extern "C" int b200_nonbonded_create(b200_ctx* c, const b200_nonbonded_desc* d) {
    API_BEGIN
    if (!c || !d) throw B200Error(B200_ERR_INVALID, "b200_nonbonded_create: null argument");
    if (c->hasNonbonded) throw B200Error(B200_ERR_UNSUPPORTED, "only one NonbondedForce per context is supported");
    if (c->finalized) throw B200Error(B200_ERR_INVALID, "forces must be created before the first computation");
    if (d->method < B200_NB_NOCUTOFF || d->method > B200_NB_LJPME) throw B200Error(B200_ERR_INVALID, "NonbondedForce: unknown nonbonded method");
    if (d->method != B200_NB_NOCUTOFF && !(d->cutoff > 0)) throw B200Error(B200_ERR_INVALID, "NonbondedForce: the cutoff distance must be positive");
    if (d->use_switch && d->method != B200_NB_NOCUTOFF) {
        if (!(d->switch_distance >= 0) || !(d->switch_distance < d->cutoff))
            throw B200Error(B200_ERR_INVALID, "NonbondedForce: Switching distance must satisfy 0 <= r_switch < r_cutoff");   // NonbondedForceImpl.cpp:70-74
    }
    CUDA_CHECK(cudaSetDevice(c->device));
    c->nb = *d;
    c->nb.charge = c->nb.sigma = c->nb.epsilon = nullptr; c->nb.exclusion_pairs = c->nb.exception_pairs = nullptr;
    c->nb.exception_chargeprod = c->nb.exception_sigma = c->nb.exception_epsilon = nullptr;
    c->charge.assign(d->charge, d->charge + c->N); c->sigma.assign(d->sigma, d->sigma + c->N); c->epsilon.assign(d->epsilon, d->epsilon + c->N);
    c->exclPairs.assign(d->exclusion_pairs, d->exclusion_pairs + 2 * (size_t)d->num_exclusions);
    c->excPairs.assign(d->exception_pairs, d->exception_pairs + 2 * (size_t)d->num_exceptions);
    c->excQQ.assign(d->exception_chargeprod, d->exception_chargeprod + d->num_exceptions);
    c->excSig.assign(d->exception_sigma, d->exception_sigma + d->num_exceptions);
    c->excEps.assign(d->exception_epsilon, d->exception_epsilon + d->num_exceptions);
    const bool usesMesh = d->method == B200_NB_PME || d->method == B200_NB_LJPME;
    for (int k = 0; k < 3 && usesMesh; k++) {
        if (d->grid[k] < 6 || b200_find_legal_fft_dimension(d->grid[k]) != d->grid[k])
            throw B200Error(B200_ERR_INVALID, "PME grid dimensions must be >= 6 and factor into 2,3,5,7");
    }
    if (d->method == B200_NB_EWALD) {
        for (int k = 0; k < 3; k++) if (d->ewald_kmax[k] < 1) throw B200Error(B200_ERR_INVALID, "Ewald: kmax must be at least 1 in every direction");
    }
    if (d->method == B200_NB_NOCUTOFF) c->nb.use_switch = 0;     // no cutoff, nothing to switch (NonbondedForceImpl ignores it as well)
    c->ljpme = d->method == B200_NB_LJPME;
    if (c->ljpme) {
        if (!(d->ljpme_alpha > 0)) throw B200Error(B200_ERR_INVALID, "LJPME: the dispersion Ewald coefficient must be positive");
        for (int k = 0; k < 3; k++)
            if (d->ljpme_grid[k] < 6 || b200_find_legal_fft_dimension(d->ljpme_grid[k]) != d->ljpme_grid[k])
                throw B200Error(B200_ERR_INVALID, "LJPME grid dimensions must be >= 6 and factor into 2,3,5,7");
        c->nb.dispersion_coefficient = 0;   // no long-range correction with LJPME (CommonCalcNonbondedForce.cpp:360)
    }
    c->hasNonbonded = true;
    // list cutoff = cutoff + padding (CudaNonbondedUtilities.cpp:380-383 uses 8% of the cutoff)
    const char* padEnv = getenv("B200_NB_PADDING");
    const double padFrac = padEnv ? atof(padEnv) : 0.08;
    c->padding = padFrac * d->cutoff;
    c->paddedCutoff = d->cutoff + c->padding;
    if (d->method == B200_NB_NOCUTOFF) { c->padding = INFINITY; c->paddedCutoff = INFINITY; }   // the list holds every pair and never needs a rebuild
    const bool ewaldType = d->method >= B200_NB_EWALD;
    // the rational fit covers alpha*rc <= 3.95; the other Coulomb forms live in the scalar ("exact") instantiation of the tile kernel
    c->exactEwald = !ewaldType || d->ewald_alpha * d->cutoff > 3.95;
    if (const char* sp = getenv("B200_SPARSE_MAX")) c->sparseMax = std::max(0, std::min(32, atoi(sp)));
    b200_box_set(c, c->boxHost);
    // per-slot parameters (nonbondedParameters.cc:30-35: posq.w = q, sigmaEpsilon = (sigma/2, 2 sqrt(eps)))
    std::vector<float2> se(c->Npad, make_float2(0.f, 0.f));
    double sumQ = 0, sumQ2 = 0;
    for (int i = 0; i < c->N; i++) {
        se[i] = make_float2((float)(0.5 * c->sigma[i]), (float)(2.0 * sqrt(c->epsilon[i])));
        sumQ += c->charge[i]; sumQ2 += c->charge[i] * c->charge[i];
    }
    CUDA_CHECK(cudaMemcpyAsync(c->sigEps, se.data(), sizeof(float2) * c->Npad, cudaMemcpyHostToDevice, c->stream));
    {   // posq.w = q; the device order is still the creation order here (no reorder before the first computation)
        std::vector<float4> p(c->Npad, make_float4(0.f, 0.f, 0.f, 0.f));
        d2h(p.data(), c->posq, c->Npad, c->stream);
        CUDA_CHECK(cudaStreamSynchronize(c->stream));
        for (int i = 0; i < c->N; i++) p[i].w = (float)c->charge[i];
        CUDA_CHECK(cudaMemcpyAsync(c->posq, p.data(), sizeof(float4) * c->Npad, cudaMemcpyHostToDevice, c->stream));
    }
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    c->totalCharge = sumQ;
    // Ewald self energy (CommonCalcNonbondedForce.cpp:397-400)
    c->selfEnergy = ewaldType ? -B200_ONE_4PI_EPS0 * d->ewald_alpha * sumQ2 / sqrt(M_PI) : 0.0;
    b200_lj_self_energy(c);
    if (usesMesh) b200_pme_create(c);
    API_END
}

extern "C" int b200_nonbonded_get_pme_parameters(const b200_ctx* c, double* alpha, int32_t* nx, int32_t* ny, int32_t* nz) {
    API_BEGIN
    if (!c || !c->hasNonbonded || c->nb.method < B200_NB_PME) throw B200Error(B200_ERR_INVALID, "getPMEParametersInContext: This Context is not using PME");
    *alpha = c->nb.ewald_alpha; *nx = c->nb.grid[0]; *ny = c->nb.grid[1]; *nz = c->nb.grid[2];
    API_END
}

extern "C" int b200_nonbonded_get_ljpme_parameters(const b200_ctx* c, double* alpha, int32_t* nx, int32_t* ny, int32_t* nz) {
    API_BEGIN
    if (!c || !c->hasNonbonded || !c->ljpme) throw B200Error(B200_ERR_INVALID, "this context is not using LJPME");
    *alpha = c->nb.ljpme_alpha; *nx = c->nb.ljpme_grid[0]; *ny = c->nb.ljpme_grid[1]; *nz = c->nb.ljpme_grid[2];
    API_END
}

// CalcNonbondedForceKernel::copyParametersToContext (CommonCalcNonbondedForce.cpp:1119-1254): per-particle charge/sigma/epsilon
// and exception parameters may change; the set of exceptions may not.
extern "C" int b200_nonbonded_update_parameters(b200_ctx* c, const b200_nonbonded_desc* d) {
    b200_graph_invalidate(c);
    API_BEGIN
    if (!c || !d || !c->hasNonbonded) throw B200Error(B200_ERR_INVALID, "b200_nonbonded_update_parameters: no NonbondedForce in this context");
    if (d->num_exclusions * 2 != (int)c->exclPairs.size() || d->num_exceptions != (int)c->excQQ.size())
        throw B200Error(B200_ERR_INVALID, "updateParametersInContext: The set of non-excluded exceptions has changed");
    CUDA_CHECK(cudaSetDevice(c->device));
    c->charge.assign(d->charge, d->charge + c->N); c->sigma.assign(d->sigma, d->sigma + c->N); c->epsilon.assign(d->epsilon, d->epsilon + c->N);
    c->excQQ.assign(d->exception_chargeprod, d->exception_chargeprod + d->num_exceptions);
    c->excSig.assign(d->exception_sigma, d->exception_sigma + d->num_exceptions);
    c->excEps.assign(d->exception_epsilon, d->exception_epsilon + d->num_exceptions);
    c->nb.dispersion_coefficient = c->ljpme ? 0.0 : d->dispersion_coefficient;
    b200_sync_host_order(c);
    std::vector<float2> se(c->Npad, make_float2(0.f, 0.f));
    std::vector<float4> p(c->Npad);
    d2h(p.data(), c->posq, c->Npad, c->stream);
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    double sumQ = 0, sumQ2 = 0;
    for (int s = 0; s < c->N; s++) {
        const int o = c->hostAtomIndex[s];
        se[s] = make_float2((float)(0.5 * c->sigma[o]), (float)(2.0 * sqrt(c->epsilon[o])));
        p[s].w = (float)c->charge[o];
        sumQ += c->charge[o]; sumQ2 += c->charge[o] * c->charge[o];
    }
    h2d(c->sigEps, se.data(), c->Npad, c->stream); h2d(c->posq, p.data(), c->Npad, c->stream);
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    c->totalCharge = sumQ;
    c->selfEnergy = c->nb.method >= B200_NB_EWALD ? -B200_ONE_4PI_EPS0 * c->nb.ewald_alpha * sumQ2 / sqrt(M_PI) : 0.0;
    b200_lj_self_energy(c);
    if (c->nb.method >= B200_NB_PME) b200_pme_rescale(c);
    if (c->finalized) {
        // Molecules of a group are exchanged among the group's slots when atoms are re-sorted, which is only valid while they are
        // identical.  New parameters may have made some of them different (the reference recomputes its molecule groups:
        // ComputeContext::invalidateMolecules); a group that is no longer homogeneous is simply taken out of the sort from now on.
        std::map<int, std::string> excSig;   // molecule -> its exception parameters, in list order
        for (size_t i = 0; i < c->excQQ.size(); i++) {
            std::ostringstream p; p << c->excQQ[i] << "," << c->excSig[i] << "," << c->excEps[i] << ";";
            excSig[c->molBaseOf[c->excPairs[2 * i]]] += p.str();
        }
        for (size_t gi = 0; gi < c->groups.size();) {
            auto& g = c->groups[gi];
            bool same = true;
            const int b0 = g.base.empty() ? 0 : g.base[0];
            for (int m = 1; m < g.numMols && same; m++) {
                const int bm = g.base[m];
                for (int k = 0; k < g.atomsPerMol && same; k++)
                    same = c->charge[bm + k] == c->charge[b0 + k] && c->sigma[bm + k] == c->sigma[b0 + k] && c->epsilon[bm + k] == c->epsilon[b0 + k];
                if (same) { auto a = excSig.find(b0), b = excSig.find(bm); same = (a == excSig.end() ? std::string() : a->second) == (b == excSig.end() ? std::string() : b->second); }
            }
            if (same) gi++;
            else { dfree(g.dBase); c->groups.erase(c->groups.begin() + gi); }
        }
    }
    if (c->finalized && c->numExceptions > 0) {
        // An exception belongs to two PARTICLES; after a re-sort they sit in the slots hostSlotOf[] says (identical molecules were
        // exchanged), so both the parameters and the slot pair are written: a parameter change that concerns one of several formerly
        // identical molecules lands on the molecule it was meant for.
        std::vector<float4> xq; std::vector<int2> xp;
        for (size_t i = 0; i < c->excQQ.size(); i++) {
            xq.push_back(make_float4((float)(B200_ONE_4PI_EPS0 * c->excQQ[i]), (float)c->excSig[i], (float)(4.0 * c->excEps[i]), 0.f));
            xp.push_back(make_int2(c->hostSlotOf[c->excPairs[2 * i]], c->hostSlotOf[c->excPairs[2 * i + 1]]));
        }
        CUDA_CHECK(cudaMemcpy(c->dExcParams, xq.data(), sizeof(float4) * xq.size(), cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemcpy(c->dExcPairs, xp.data(), sizeof(int2) * xp.size(), cudaMemcpyHostToDevice));
    }
    c->forcesValid = false;
    API_END
}

extern "C" int b200_find_legal_fft_dimension(int32_t minimum) {
    if (minimum < 1) return 1;
    for (int n = minimum;; n++) {
        int m = n;
        for (int f : {2, 3, 5, 7}) while (m % f == 0) m /= f;
        if (m == 1) return n;
    }
}

extern "C" int b200_harmonic_bonds_create(b200_ctx* c, int32_t n, const int32_t* atoms, const double* length, const double* k, int32_t) {
    API_BEGIN
    if (c->finalized) throw B200Error(B200_ERR_INVALID, "forces must be created before the first computation");
    for (int i = 0; i < n; i++) { c->hBondAtoms.push_back(atoms[2 * i]); c->hBondAtoms.push_back(atoms[2 * i + 1]); c->hBondParams.push_back(length[i]); c->hBondParams.push_back(k[i]); }
    API_END
}
extern "C" int b200_harmonic_angles_create(b200_ctx* c, int32_t n, const int32_t* atoms, const double* angle, const double* k, int32_t) {
    API_BEGIN
    if (c->finalized) throw B200Error(B200_ERR_INVALID, "forces must be created before the first computation");
    for (int i = 0; i < n; i++) { for (int j = 0; j < 3; j++) c->hAngleAtoms.push_back(atoms[3 * i + j]); c->hAngleParams.push_back(angle[i]); c->hAngleParams.push_back(k[i]); }
    API_END
}
extern "C" int b200_periodic_torsions_create(b200_ctx* c, int32_t n, const int32_t* atoms, const int32_t* periodicity, const double* phase, const double* k, int32_t) {
    API_BEGIN
    if (c->finalized) throw B200Error(B200_ERR_INVALID, "forces must be created before the first computation");
    for (int i = 0; i < n; i++) { for (int j = 0; j < 4; j++) c->hTorsionAtoms.push_back(atoms[4 * i + j]); c->hTorsionPeriod.push_back(periodicity[i]); c->hTorsionParams.push_back(phase[i]); c->hTorsionParams.push_back(k[i]); }
    API_END
}

extern "C" int b200_bonded_set_active(b200_ctx* c, int32_t kind, int32_t nranges, const int32_t* firstCount) {
    if (!c || kind < 0 || kind > 2 || nranges < 0 || (nranges > 0 && !firstCount)) { b200_set_error("b200_bonded_set_active: bad arguments"); return B200_ERR_INVALID; }
    const int total = kind == 0 ? (int)c->hBondParams.size() / 2 : kind == 1 ? (int)c->hAngleParams.size() / 2 : (int)c->hTorsionPeriod.size();
    std::vector<int> r(firstCount, firstCount + 2 * (size_t)nranges);
    for (int k = 0; k < nranges; k++)
        if (r[2 * k] < 0 || r[2 * k + 1] < 0 || (long long)r[2 * k] + r[2 * k + 1] > total) { b200_set_error("b200_bonded_set_active: range outside the terms of this kind"); return B200_ERR_INVALID; }
    if (r != c->bondedActive[kind]) { b200_graph_invalidate(c); c->bondedActive[kind] = r; }     // the ranges are kernel arguments of the captured evaluation
    return B200_OK;
}

extern "C" int b200_bonded_set_periodic(b200_ctx* c, int32_t kind, int32_t periodic) {
    if (!c || kind < 0 || kind > 2) { b200_set_error("b200_bonded_set_periodic: bad arguments"); return B200_ERR_INVALID; }
    c->bondedPeriodic[kind] = periodic ? 1 : 0; b200_graph_invalidate(c);
    return B200_OK;
}

// updateParametersInContext of the bonded forces (kernels.h:359,433,507 copyParametersToContext): new parameters for the terms
// [first, first+n) of one kind (0 bonds: length,k; 1 angles: angle,k; 2 torsions: phase,k + periodicity); the atoms of a term may not change.
// Molecules whose terms changed stop being exchangeable with their former twins, so their group leaves the re-sort.
extern "C" int b200_bonded_update_parameters(b200_ctx* c, int32_t kind, int32_t first, int32_t n, const double* p0, const double* p1, const int32_t* periodicity) {
    b200_graph_invalidate(c);
    API_BEGIN
    if (!c || kind < 0 || kind > 2 || first < 0 || n < 0) throw B200Error(B200_ERR_INVALID, "b200_bonded_update_parameters: bad arguments");
    std::vector<double>& prm = kind == 0 ? c->hBondParams : kind == 1 ? c->hAngleParams : c->hTorsionParams;
    const std::vector<int>& atoms = kind == 0 ? c->hBondAtoms : kind == 1 ? c->hAngleAtoms : c->hTorsionAtoms;
    const int width = kind == 0 ? 2 : kind == 1 ? 3 : 4;
    if ((size_t)(first + n) * 2 > prm.size()) throw B200Error(B200_ERR_INVALID, "updateParametersInContext: The number of terms has changed");
    CUDA_CHECK(cudaSetDevice(c->device));
    std::set<int> changedMols;
    for (int i = 0; i < n; i++) {
        const size_t t = (size_t)first + i;
        const bool diff = prm[2 * t] != p0[i] || prm[2 * t + 1] != p1[i] || (kind == 2 && c->hTorsionPeriod[t] != periodicity[i]);
        prm[2 * t] = p0[i]; prm[2 * t + 1] = p1[i];
        if (kind == 2) c->hTorsionPeriod[t] = periodicity[i];
        if (diff && c->finalized) changedMols.insert(c->molBaseOf[atoms[width * t]]);
    }
    if (!c->finalized) return B200_OK;      // nothing is on the device yet
    for (size_t gi = 0; gi < c->groups.size() && !changedMols.empty();) {
        bool hit = false;
        for (int b : c->groups[gi].base) if (changedMols.count(b)) { hit = true; break; }
        if (hit && c->groups[gi].numMols > 1) { dfree(c->groups[gi].dBase); c->groups.erase(c->groups.begin() + gi); } else gi++;
    }
    // re-upload the parameter arrays of this kind (the index arrays stay)
    if (kind == 0) { std::vector<float2> v; for (size_t t = 0; t * 2 < prm.size(); t++) v.push_back(make_float2((float)prm[2 * t], (float)prm[2 * t + 1])); if (!v.empty()) CUDA_CHECK(cudaMemcpy(c->bondParams, v.data(), sizeof(float2) * v.size(), cudaMemcpyHostToDevice)); }
    else if (kind == 1) { std::vector<float2> v; for (size_t t = 0; t * 2 < prm.size(); t++) v.push_back(make_float2((float)prm[2 * t], (float)prm[2 * t + 1])); if (!v.empty()) CUDA_CHECK(cudaMemcpy(c->angleParams, v.data(), sizeof(float2) * v.size(), cudaMemcpyHostToDevice)); }
    else { std::vector<float4> v; for (size_t t = 0; t * 2 < prm.size(); t++) v.push_back(make_float4((float)c->hTorsionPeriod[t], (float)prm[2 * t], (float)prm[2 * t + 1], 0.f)); if (!v.empty()) CUDA_CHECK(cudaMemcpy(c->torsionParams, v.data(), sizeof(float4) * v.size(), cudaMemcpyHostToDevice)); }
    c->forcesValid = false;
    API_END
}

// ---------------------------------------------------------------------------------------------
// finalize: molecules, identical-molecule groups, SETTLE clusters, exclusion tiles
static int ufFind(std::vector<int>& p, int i) { while (p[i] != i) { p[i] = p[p[i]]; i = p[i]; } return i; }

void b200_finalize(b200_ctx* c) {
    if (c->finalized) return;
    CUDA_CHECK(cudaSetDevice(c->device));
    const int N = c->N;
    // ---- molecules = connected components over every index-based term (ComputeContext::findMoleculeGroups, ComputeContext.cpp:228-369)
    std::vector<int> parent(N); std::iota(parent.begin(), parent.end(), 0);
    auto join = [&](int a, int b) { a = ufFind(parent, a); b = ufFind(parent, b); if (a != b) parent[std::max(a, b)] = std::min(a, b); };
    for (int i = 0; i < c->numConstraints; i++) join(c->conAtoms[2 * i], c->conAtoms[2 * i + 1]);
    for (size_t i = 0; i + 1 < c->exclPairs.size(); i += 2) join(c->exclPairs[i], c->exclPairs[i + 1]);
    for (size_t i = 0; i + 1 < c->excPairs.size(); i += 2) join(c->excPairs[i], c->excPairs[i + 1]);
    for (size_t i = 0; i + 1 < c->hBondAtoms.size(); i += 2) join(c->hBondAtoms[i], c->hBondAtoms[i + 1]);
    for (size_t i = 0; i + 2 < c->hAngleAtoms.size(); i += 3) { join(c->hAngleAtoms[i], c->hAngleAtoms[i + 1]); join(c->hAngleAtoms[i], c->hAngleAtoms[i + 2]); }
    for (size_t i = 0; i + 3 < c->hTorsionAtoms.size(); i += 4) for (int j = 1; j < 4; j++) join(c->hTorsionAtoms[i], c->hTorsionAtoms[i + j]);
    std::vector<int> molOf(N), molFirst;
    std::vector<std::vector<int>> molAtoms;
    {
        std::map<int, int> rootToMol;
        for (int i = 0; i < N; i++) {
            const int r = ufFind(parent, i);
            auto it = rootToMol.find(r);
            if (it == rootToMol.end()) { it = rootToMol.emplace(r, (int)molAtoms.size()).first; molAtoms.emplace_back(); }
            molOf[i] = it->second; molAtoms[it->second].push_back(i);
        }
    }
    c->numMolecules = (int)molAtoms.size();
    {   // molecules as slot lists on the device (slot == particle index at creation; re-sorting only exchanges identical molecules)
        std::vector<int> start(molAtoms.size() + 1, 0), slots; slots.reserve(N);
        for (size_t m = 0; m < molAtoms.size(); m++) { start[m] = (int)slots.size(); slots.insert(slots.end(), molAtoms[m].begin(), molAtoms[m].end()); }
        start[molAtoms.size()] = (int)slots.size();
        dmalloc(c->molStart, start.size()); dmalloc(c->molSlots, slots.size());
        CUDA_CHECK(cudaMemcpy(c->molStart, start.data(), sizeof(int) * start.size(), cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemcpy(c->molSlots, slots.data(), sizeof(int) * slots.size(), cudaMemcpyHostToDevice));
    }
    c->molBaseOf.resize(N);
    for (int i = 0; i < N; i++) c->molBaseOf[i] = molAtoms[molOf[i]][0];
    // ---- signature of each molecule: everything that must match for two molecules to be exchangeable
    std::vector<std::ostringstream> sig(molAtoms.size());
    for (size_t m = 0; m < molAtoms.size(); m++) {
        const auto& at = molAtoms[m];
        sig[m] << at.size() << "|";
        for (int a : at) {
            sig[m] << (a - at[0]) << "," << c->mass[a];
            if (c->hasNonbonded) sig[m] << "," << c->charge[a] << "," << c->sigma[a] << "," << c->epsilon[a];
            sig[m] << ";";
        }
    }
    sig.shrink_to_fit();
    auto addTerm = [&](const char* tag, const int* atoms, int n, const std::string& params) {
        const int m = molOf[atoms[0]]; const int base = molAtoms[m][0];
        sig[m] << tag; for (int j = 0; j < n; j++) sig[m] << (atoms[j] - base) << ","; sig[m] << params << ";";
    };
    // terms are appended in list order; identical molecules created by the same builder list their terms in the same order
    for (int i = 0; i < c->numConstraints; i++) addTerm("C", &c->conAtoms[2 * i], 2, std::to_string(c->conDist[i]));
    for (size_t i = 0; i + 1 < c->exclPairs.size(); i += 2) { int a[2] = {std::min(c->exclPairs[i], c->exclPairs[i + 1]), std::max(c->exclPairs[i], c->exclPairs[i + 1])}; addTerm("X", a, 2, ""); }
    for (size_t i = 0; i + 1 < c->excPairs.size(); i += 2) { std::ostringstream p; p << c->excQQ[i / 2] << "," << c->excSig[i / 2] << "," << c->excEps[i / 2]; addTerm("E", &c->excPairs[i], 2, p.str()); }
    for (size_t i = 0; i + 1 < c->hBondAtoms.size(); i += 2) { std::ostringstream p; p << c->hBondParams[i] << "," << c->hBondParams[i + 1]; addTerm("B", &c->hBondAtoms[i], 2, p.str()); }
    for (size_t i = 0; i + 2 < c->hAngleAtoms.size(); i += 3) { std::ostringstream p; p << c->hAngleParams[2 * (i / 3)] << "," << c->hAngleParams[2 * (i / 3) + 1]; addTerm("A", &c->hAngleAtoms[i], 3, p.str()); }
    for (size_t i = 0; i + 3 < c->hTorsionAtoms.size(); i += 4) { std::ostringstream p; p << c->hTorsionPeriod[i / 4] << "," << c->hTorsionParams[2 * (i / 4)] << "," << c->hTorsionParams[2 * (i / 4) + 1]; addTerm("T", &c->hTorsionAtoms[i], 4, p.str()); }
    std::map<std::string, int> groupOf;
    c->groups.clear();
    for (size_t m = 0; m < molAtoms.size(); m++) {
        const auto& at = molAtoms[m];
        bool contiguous = true;
        for (size_t k = 1; k < at.size(); k++) contiguous &= (at[k] == at[0] + (int)k);
        std::string key = sig[m].str();
        if (!contiguous || at.size() > 128) key += "#unique" + std::to_string(m);   // only contiguous small molecules are exchanged
        auto it = groupOf.find(key);
        if (it == groupOf.end()) { it = groupOf.emplace(key, (int)c->groups.size()).first; c->groups.emplace_back(); c->groups.back().atomsPerMol = contiguous ? (int)at.size() : 0; c->groups.back().numMols = 0; }
        auto& g = c->groups[it->second];
        if (contiguous) { g.base.push_back(at[0]); g.numMols++; }
    }
    // drop groups that cannot be wrapped/sorted as contiguous runs
    c->groups.erase(std::remove_if(c->groups.begin(), c->groups.end(), [](const b200_ctx::MolGroup& g) { return g.numMols == 0 || g.atomsPerMol == 0; }), c->groups.end());
    c->maxGroupMols = 1;
    for (auto& g : c->groups) {
        dmalloc(g.dBase, (size_t)g.numMols);
        CUDA_CHECK(cudaMemcpy(g.dBase, g.base.data(), sizeof(int) * g.numMols, cudaMemcpyHostToDevice));
        c->maxGroupMols = std::max(c->maxGroupMols, g.numMols);
    }
    c->sortBits = 1;
    while (c->sortBits < 8 && (1 << (3 * c->sortBits)) < 2 * c->maxGroupMols) c->sortBits++;
    c->sortBuckets = 1 << (3 * c->sortBits);
    dmalloc(c->sortKeys, (size_t)c->maxGroupMols); dmalloc(c->sortOrder, (size_t)c->maxGroupMols);
    dmalloc(c->sortCount, (size_t)c->sortBuckets);
    dmalloc(c->sortStart, (size_t)2 * c->sortBuckets + c->sortBuckets / 1024 + 64);

    // ---- SETTLE clusters: three atoms joined by three constraints, two equal lengths (IntegrationUtilities, platforms/common/src/IntegrationUtilities.cpp)
    {
        std::vector<std::vector<std::pair<int, double>>> con(N);
        for (int i = 0; i < c->numConstraints; i++) {
            con[c->conAtoms[2 * i]].push_back({c->conAtoms[2 * i + 1], c->conDist[i]});
            con[c->conAtoms[2 * i + 1]].push_back({c->conAtoms[2 * i], c->conDist[i]});
        }
        std::vector<char> inCluster(N, 0);
        std::vector<int4> sa; std::vector<float2> sp;
        for (int i = 0; i < N; i++) {
            if (inCluster[i] || con[i].size() != 2) continue;
            const int j = con[i][0].first, k = con[i][1].first;
            if (con[j].size() != 2 || con[k].size() != 2) continue;
            double djk = -1;
            for (auto& e : con[j]) if (e.first == k) djk = e.second;
            if (djk < 0) continue;
            const double dij = con[i][0].second, dik = con[i][1].second;
            int a0, a1, a2; double d01, d12;
            if (dij == dik) { a0 = i; a1 = j; a2 = k; d01 = dij; d12 = djk; }
            else if (dij == djk) { a0 = j; a1 = i; a2 = k; d01 = dij; d12 = dik; }
            else if (dik == djk) { a0 = k; a1 = i; a2 = j; d01 = dik; d12 = dij; }
            else continue;
            if (c->mass[a0] <= 0 || c->mass[a1] <= 0 || c->mass[a2] <= 0) continue;
            inCluster[i] = inCluster[j] = inCluster[k] = 1;
            sa.push_back(make_int4(a0, a1, a2, 0)); sp.push_back(make_float2((float)d01, (float)d12));
        }
        c->numSettle = (int)sa.size();
        c->hSettleAtoms.clear();
        for (const int4& q : sa) { c->hSettleAtoms.push_back(q.x); c->hSettleAtoms.push_back(q.y); c->hSettleAtoms.push_back(q.z); c->hSettleAtoms.push_back(q.w); }
        dmalloc(c->settleAtoms, sa.size()); dmalloc(c->settleParams, sp.size());
        if (!sa.empty()) {
            CUDA_CHECK(cudaMemcpy(c->settleAtoms, sa.data(), sizeof(int4) * sa.size(), cudaMemcpyHostToDevice));
            CUDA_CHECK(cudaMemcpy(c->settleParams, sp.data(), sizeof(float2) * sp.size(), cudaMemcpyHostToDevice));
        }
        // ---- SHAKE clusters (integrationUtilities.cc:99-240 applyShakeToHydrogens; the HBonds constraints of a solvated protein):
        // a central atom whose 1-3 constrained partners are constrained to nothing else
        {
            std::vector<int4> ka; std::vector<float4> kd;
            for (int i = 0; i < N; i++) {
                if (inCluster[i] || con[i].empty() || con[i].size() > 3) continue;
                bool star = true;
                for (auto& e : con[i]) star = star && !inCluster[e.first] && con[e.first].size() == 1 && c->mass[e.first] > 0;
                if (!star || c->mass[i] <= 0) continue;
                if (con[i].size() == 1 && con[i][0].first < i) continue;     // a lone pair: its lower index is the centre
                int a[4] = {i, -1, -1, -1}; float dd[4] = {0.f, 0.f, 0.f, 0.f};
                for (size_t k = 0; k < con[i].size(); k++) { a[k + 1] = con[i][k].first; dd[k] = (float)con[i][k].second; inCluster[con[i][k].first] = 1; }
                inCluster[i] = 1;
                ka.push_back(make_int4(a[0], a[1], a[2], a[3])); kd.push_back(make_float4(dd[0], dd[1], dd[2], 0.f));
            }
            c->numShake = (int)ka.size();
            dmalloc(c->shakeAtoms, ka.size()); dmalloc(c->shakeDist, kd.size());
            if (!ka.empty()) {
                CUDA_CHECK(cudaMemcpy(c->shakeAtoms, ka.data(), sizeof(int4) * ka.size(), cudaMemcpyHostToDevice));
                CUDA_CHECK(cudaMemcpy(c->shakeDist, kd.data(), sizeof(float4) * kd.size(), cudaMemcpyHostToDevice));
            }
        }
        // ---- everything else: general clusters (connected components of the remaining constraints), solved iteratively by one CTA
        // each with the constraints processed colour by colour (CCMA's job in the reference, integrationUtilities.cc:582-804)
        {
            std::vector<int> cp(N); std::iota(cp.begin(), cp.end(), 0);
            std::vector<int> rest;
            for (int k = 0; k < c->numConstraints; k++) {
                const int a = c->conAtoms[2 * k], b = c->conAtoms[2 * k + 1];
                if (inCluster[a] || inCluster[b]) continue;
                rest.push_back(k);
                int ra = ufFind(cp, a), rb = ufFind(cp, b); if (ra != rb) cp[std::max(ra, rb)] = std::min(ra, rb);
            }
            std::map<int, std::vector<int>> clusters;      // root atom -> constraint indices
            for (int k : rest) clusters[ufFind(cp, c->conAtoms[2 * k])].push_back(k);
            std::vector<int2> pairs; std::vector<double> dist; std::vector<int> colorStart, atomStart(1, 0), atomList; std::vector<int2> clusterColors;
            for (auto& kv : clusters) {
                // greedy colouring: a constraint takes the smallest colour none of its two atoms has used yet
                std::map<int, std::set<int>> used; std::vector<int> colourOf(kv.second.size()); int ncol = 0;
                for (size_t q = 0; q < kv.second.size(); q++) {
                    const int a = c->conAtoms[2 * kv.second[q]], b = c->conAtoms[2 * kv.second[q] + 1];
                    int col = 0; while (used[a].count(col) || used[b].count(col)) col++;
                    used[a].insert(col); used[b].insert(col); colourOf[q] = col; ncol = std::max(ncol, col + 1);
                }
                clusterColors.push_back(make_int2((int)colorStart.size(), ncol));
                for (int col = 0; col < ncol; col++) {
                    colorStart.push_back((int)pairs.size());
                    for (size_t q = 0; q < kv.second.size(); q++) if (colourOf[q] == col) {
                        pairs.push_back(make_int2(c->conAtoms[2 * kv.second[q]], c->conAtoms[2 * kv.second[q] + 1])); dist.push_back(c->conDist[kv.second[q]]);
                    }
                }
                colorStart.push_back((int)pairs.size());
                for (auto& u : used) { atomList.push_back(u.first); inCluster[u.first] = 1; }
                atomStart.push_back((int)atomList.size());
            }
            c->numGc = (int)pairs.size(); c->numGcClusters = (int)clusterColors.size();
            if (c->numGc > 0) {
                dmalloc(c->gcPairs, pairs.size()); dmalloc(c->gcDist, dist.size()); dmalloc(c->gcColorStart, colorStart.size()); dmalloc(c->gcClusterColors, clusterColors.size());
                dmalloc(c->gcAtomStart, atomStart.size()); dmalloc(c->gcAtomList, atomList.size()); dmalloc(c->gcDelta, (size_t)c->Npad); dmalloc(c->gcOld, (size_t)c->Npad);
                CUDA_CHECK(cudaMemcpy(c->gcPairs, pairs.data(), sizeof(int2) * pairs.size(), cudaMemcpyHostToDevice));
                CUDA_CHECK(cudaMemcpy(c->gcDist, dist.data(), sizeof(double) * dist.size(), cudaMemcpyHostToDevice));
                CUDA_CHECK(cudaMemcpy(c->gcColorStart, colorStart.data(), sizeof(int) * colorStart.size(), cudaMemcpyHostToDevice));
                CUDA_CHECK(cudaMemcpy(c->gcClusterColors, clusterColors.data(), sizeof(int2) * clusterColors.size(), cudaMemcpyHostToDevice));
                CUDA_CHECK(cudaMemcpy(c->gcAtomStart, atomStart.data(), sizeof(int) * atomStart.size(), cudaMemcpyHostToDevice));
                CUDA_CHECK(cudaMemcpy(c->gcAtomList, atomList.data(), sizeof(int) * atomList.size(), cudaMemcpyHostToDevice));
            }
        }
        std::vector<int> freeAtoms;
        c->numOtherConstraints = 0;      // every constraint now belongs to a SETTLE molecule, a SHAKE star or a general cluster
        for (int i = 0; i < N; i++) if (!inCluster[i] && c->mass[i] > 0) freeAtoms.push_back(i);
        c->numFree = (int)freeAtoms.size();
        c->hFreeAtoms = freeAtoms;
        dmalloc(c->freeAtoms, freeAtoms.size());
        if (!freeAtoms.empty()) CUDA_CHECK(cudaMemcpy(c->freeAtoms, freeAtoms.data(), sizeof(int) * freeAtoms.size(), cudaMemcpyHostToDevice));
    }

    // ---- exclusion tiles (CudaNonbondedUtilities.cpp:178-243): every diagonal tile plus every block pair joined by an exclusion
    if (c->hasNonbonded) {
        std::map<std::pair<int, int>, std::vector<unsigned>> tiles;
        auto tileOf = [&](int x, int y) -> std::vector<unsigned>& {
            auto& t = tiles[{x, y}];
            if (t.empty()) t.assign(32, 0u);
            return t;
        };
        for (int b = 0; b < c->B; b++) {
            auto& t = tileOf(b, b);
            for (int i = 0; i < 32; i++) {
                unsigned f = (i == 31) ? 0xffffffffu : ((2u << i) - 1u);     // j <= i: each pair once, no self term
                for (int j = 0; j < 32; j++) if (b * 32 + j >= N || b * 32 + i >= N) f |= 1u << j;
                t[i] = f;
            }
        }
        for (size_t e = 0; e + 1 < c->exclPairs.size(); e += 2) {
            int a = c->exclPairs[e], b = c->exclPairs[e + 1];
            if (a == b) continue;
            if (a > b) std::swap(a, b);
            const int x = a / 32, y = b / 32;
            auto& t = tileOf(x, y);
            if (t[0] == 0 && x != y) {   // fresh off-diagonal tile: mask padding atoms
                for (int i = 0; i < 32; i++) for (int j = 0; j < 32; j++) if (x * 32 + i >= N || y * 32 + j >= N) t[i] |= 1u << j;
            }
            if (x == y) { t[a % 32] |= 1u << (b % 32); t[b % 32] |= 1u << (a % 32); }
            else t[a % 32] |= 1u << (b % 32);
        }
        // off-diagonal tiles created above may have been created by tileOf() with zero flags and no padding mask when t[0]==0 check failed: redo padding mask
        for (auto& kv : tiles) {
            const int x = kv.first.first, y = kv.first.second;
            if (x == y) continue;
            for (int i = 0; i < 32; i++) for (int j = 0; j < 32; j++) if (x * 32 + i >= N || y * 32 + j >= N) kv.second[i] |= 1u << j;
        }
        std::vector<int2> tl; std::vector<unsigned> fl; std::vector<int> rowStart(c->B + 1, 0), rowCols;
        for (auto& kv : tiles) { tl.push_back(make_int2(kv.first.first, kv.first.second)); fl.insert(fl.end(), kv.second.begin(), kv.second.end()); }
        for (auto& kv : tiles) if (kv.first.second > kv.first.first) rowStart[kv.first.first + 1]++;
        for (int b = 0; b < c->B; b++) rowStart[b + 1] += rowStart[b];
        rowCols.resize(rowStart[c->B]);
        { std::vector<int> fill(rowStart.begin(), rowStart.end() - 1); for (auto& kv : tiles) if (kv.first.second > kv.first.first) rowCols[fill[kv.first.first]++] = kv.first.second; }
        c->numExclTiles = (int)tl.size();
        dmalloc(c->exclTiles, tl.size()); dmalloc(c->exclFlags, fl.size()); dmalloc(c->exclRowStart, rowStart.size()); dmalloc(c->exclRowCols, rowCols.size());
        CUDA_CHECK(cudaMemcpy(c->exclTiles, tl.data(), sizeof(int2) * tl.size(), cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemcpy(c->exclFlags, fl.data(), sizeof(unsigned) * fl.size(), cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemcpy(c->exclRowStart, rowStart.data(), sizeof(int) * rowStart.size(), cudaMemcpyHostToDevice));
        if (!rowCols.empty()) CUDA_CHECK(cudaMemcpy(c->exclRowCols, rowCols.data(), sizeof(int) * rowCols.size(), cudaMemcpyHostToDevice));
        b200_nblist_alloc(c);
    }
    b200_bonded_upload(c);
    c->finalized = true;
}

// ---------------------------------------------------------------------------------------------
// state I/O
void b200_sync_host_order(b200_ctx* c) {
    if (!c->hostOrderStale) return;
    CUDA_CHECK(cudaMemcpyAsync(c->hostAtomIndex.data(), c->atomIndex, sizeof(int) * c->Npad, cudaMemcpyDeviceToHost, c->stream));
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    for (int s = 0; s < c->N; s++) c->hostSlotOf[c->hostAtomIndex[s]] = s;
    c->hostOrderStale = false;
}


// ---- state I/O.  The permutation between the caller's particle order and the device order, the float/correction split of
// mixed precision and the fixed-point -> double conversion all run on the GPU; the host only moves contiguous double arrays
// (through a pinned staging buffer filled by a few threads, so both the memcpy and the PCIe copy run near full speed).
#include <atomic>
#include <thread>
static void ensureStaging(b200_ctx* c) {
    if (c->hStage) return;
    CUDA_CHECK(cudaMallocHost((void**)&c->hStage, sizeof(double) * 3 * (size_t)c->Npad));
    dmalloc(c->dStage, 3 * (size_t)c->Npad);
}
// Host <-> device staging of the caller's double[3N] arrays.
//   * a caller buffer that is already page-locked (cudaHostAlloc / cudaHostRegister, e.g. a pinned torch tensor) is used by the
//     DMA engine directly;
//   * a pageable buffer goes through the context's pinned staging area in chunks: worker threads copy chunk k+1.. while the
//     DMA of chunk k is in flight, so the cost is max(host copy, PCIe) instead of their sum.
static bool isPinned(const void* p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
}
#define IO_CHUNKS 16
#define IO_THREADS 8
static void stageIn(b200_ctx* c, const double* xyz) {
    ensureStaging(c);
    const size_t bytes = sizeof(double) * 3 * (size_t)c->N;
    if (isPinned(xyz)) { CUDA_CHECK(cudaMemcpyAsync(c->dStage, xyz, bytes, cudaMemcpyHostToDevice, c->stream)); return; }
    if (bytes < (4u << 20)) {
        memcpy(c->hStage, xyz, bytes);
        CUDA_CHECK(cudaMemcpyAsync(c->dStage, c->hStage, bytes, cudaMemcpyHostToDevice, c->stream));
        return;
    }
    const size_t chunk = ((bytes + IO_CHUNKS - 1) / IO_CHUNKS + 4095) & ~(size_t)4095;
    std::atomic<int> done[IO_CHUNKS];
    for (auto& d : done) d.store(0);
    std::vector<std::thread> th;
    for (int t = 0; t < IO_THREADS; t++) th.emplace_back([&, t] {
        for (int k = t; k < IO_CHUNKS; k += IO_THREADS) {
            const size_t off = (size_t)k * chunk;
            if (off < bytes) memcpy((char*)c->hStage + off, (const char*)xyz + off, std::min(chunk, bytes - off));
            done[k].store(1, std::memory_order_release);
        }
    });
    cudaError_t err = cudaSuccess;
    for (int k = 0; k < IO_CHUNKS; k++) {
        while (!done[k].load(std::memory_order_acquire)) std::this_thread::yield();
        const size_t off = (size_t)k * chunk;
        if (off < bytes && err == cudaSuccess)
            err = cudaMemcpyAsync((char*)c->dStage + off, (const char*)c->hStage + off, std::min(chunk, bytes - off), cudaMemcpyHostToDevice, c->stream);
    }
    for (auto& x : th) x.join();
    CUDA_CHECK(err);
}
static void stageOut(b200_ctx* c, double* xyz) {
    const size_t bytes = sizeof(double) * 3 * (size_t)c->N;
    if (isPinned(xyz)) {
        CUDA_CHECK(cudaMemcpyAsync(xyz, c->dStage, bytes, cudaMemcpyDeviceToHost, c->stream));
        CUDA_CHECK(cudaStreamSynchronize(c->stream));
        return;
    }
    if (bytes < (4u << 20)) {
        CUDA_CHECK(cudaMemcpyAsync(c->hStage, c->dStage, bytes, cudaMemcpyDeviceToHost, c->stream));
        CUDA_CHECK(cudaStreamSynchronize(c->stream));
        memcpy(xyz, c->hStage, bytes);
        return;
    }
    const size_t chunk = ((bytes + IO_CHUNKS - 1) / IO_CHUNKS + 4095) & ~(size_t)4095;
    if (c->ioEvents.empty()) { c->ioEvents.resize(IO_CHUNKS); for (auto& e : c->ioEvents) CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); }
    for (int k = 0; k < IO_CHUNKS; k++) {
        const size_t off = (size_t)k * chunk;
        if (off < bytes) CUDA_CHECK(cudaMemcpyAsync((char*)c->hStage + off, (const char*)c->dStage + off, std::min(chunk, bytes - off), cudaMemcpyDeviceToHost, c->stream));
        CUDA_CHECK(cudaEventRecord(c->ioEvents[k], c->stream));
    }
    std::vector<std::thread> th;
    std::atomic<int> failed(0);
    for (int t = 0; t < IO_THREADS; t++) th.emplace_back([&, t] {
        for (int k = t; k < IO_CHUNKS; k += IO_THREADS) {
            if (cudaEventSynchronize(c->ioEvents[k]) != cudaSuccess) { failed.store(1); continue; }
            const size_t off = (size_t)k * chunk;
            if (off < bytes) memcpy((char*)xyz + off, (const char*)c->hStage + off, std::min(chunk, bytes - off));
        }
    });
    for (auto& x : th) x.join();
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    if (failed.load()) throw B200Error(B200_ERR_CUDA, "device to host copy failed");
}

// CommonKernels.cpp:182-227: float cast + correction in the device order, cell offsets reset
__global__ void ioSetPositions(const double* __restrict__ stage, const int* __restrict__ atomIndex, int N, float4* posq, float4* posqCorr, int4* cellOffsets) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= N) return;
    const int o = atomIndex[s];
    const double x = stage[3 * (size_t)o], y = stage[3 * (size_t)o + 1], z = stage[3 * (size_t)o + 2];
    const float w = posq[s].w;   // the charge belongs to the slot (identical molecules are exchanged), not to the moving particle
    posq[s] = make_float4((float)x, (float)y, (float)z, w);
    posqCorr[s] = make_float4((float)(x - (double)(float)x), (float)(y - (double)(float)y), (float)(z - (double)(float)z), 0.f);
    cellOffsets[s] = make_int4(0, 0, 0, 0);
}
// CommonKernels.cpp:95-177: positions in the caller's order, optionally with the periodic wrapping undone
__global__ void ioGetPositions(double* __restrict__ stage, const int* __restrict__ atomIndex, int N, const float4* __restrict__ posq, const float4* __restrict__ posqCorr,
                               const int4* __restrict__ cellOffsets, BoxD box, int unwrap) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= N) return;
    const int o = atomIndex[s];
    const float4 p = posq[s], q = posqCorr[s];
    double x = (double)p.x + q.x, y = (double)p.y + q.y, z = (double)p.z + q.z;
    if (unwrap) {
        const int4 off = cellOffsets[s];
        x += off.x * box.ax + off.y * box.bx + off.z * box.cx;
        y += off.y * box.by + off.z * box.cy;
        z += off.z * box.cz;
    }
    stage[3 * (size_t)o] = x; stage[3 * (size_t)o + 1] = y; stage[3 * (size_t)o + 2] = z;
}
__global__ void ioSetVelocities(const double* __restrict__ stage, const int* __restrict__ atomIndex, int N, double4* velm) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= N) return;
    const int o = atomIndex[s];
    double4 v = velm[s];
    v.x = stage[3 * (size_t)o]; v.y = stage[3 * (size_t)o + 1]; v.z = stage[3 * (size_t)o + 2];
    velm[s] = v;
}
__global__ void ioGetVelocities(double* __restrict__ stage, const int* __restrict__ atomIndex, int N, const double4* __restrict__ velm) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= N) return;
    const int o = atomIndex[s];
    const double4 v = velm[s];
    stage[3 * (size_t)o] = v.x; stage[3 * (size_t)o + 1] = v.y; stage[3 * (size_t)o + 2] = v.z;
}
// CommonKernels.cpp:286-297
__global__ void ioGetForces(double* __restrict__ stage, const int* __restrict__ atomIndex, int N, int Npad, const long long* __restrict__ force) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= N) return;
    const int o = atomIndex[s];
    const double sc = 1.0 / B200_FORCE_SCALE;
    stage[3 * (size_t)o] = sc * (double)force[s]; stage[3 * (size_t)o + 1] = sc * (double)force[s + Npad]; stage[3 * (size_t)o + 2] = sc * (double)force[s + 2 * (size_t)Npad];
}

__global__ void ioAddForces(const double* __restrict__ stage, const int* __restrict__ atomIndex, int N, int Npad, long long* __restrict__ force) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= N) return;
    const int o = atomIndex[s];
    force[s] += __double2ll_rn(stage[3 * (size_t)o] * B200_FORCE_SCALE);
    force[s + Npad] += __double2ll_rn(stage[3 * (size_t)o + 1] * B200_FORCE_SCALE);
    force[s + 2 * (size_t)Npad] += __double2ll_rn(stage[3 * (size_t)o + 2] * B200_FORCE_SCALE);
}

static void afterSetPositions(b200_ctx* c) {
    c->forcesValid = false; c->dispChecked = false; c->listVerified = false;
    // Like ComputeContext::reorderAtoms the spatial sort is periodic, not forced by every setPositions: the displacement
    // check against the last list build decides on the device whether the neighbour list must be rebuilt.
    if (!c->positionsSet) c->stepsSinceReorder = -1;
    c->positionsSet = true;
}

// Multi-GPU, collective: rank 0 uploads the host array over PCIe once and stores the converted positions into every rank's arrays
// over NVLink; the other ranks (xyz ignored, may be NULL) only wait.  The device order is the same on every rank (the re-sort is
// deterministic), so slots correspond.  Without peer mapping every rank needs the array itself (b200_set_positions).
extern "C" int b200_set_positions_root(b200_ctx* c, const double* xyz) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    if (!(c->peerMode && c->commRanks > 1)) {
        if (!xyz) throw B200Error(B200_ERR_INVALID, "b200_set_positions_root: without peer mapping every rank must pass the array");
        return b200_set_positions(c, xyz);
    }
    if (c->commRank == 0) {
        if (!xyz) throw B200Error(B200_ERR_INVALID, "b200_set_positions_root: rank 0 needs the host array");
        stageIn(c, xyz);
        ioSetPositions<<<(c->N + 255) / 256, 256, 0, c->stream>>>(c->dStage, c->atomIndex, c->N, c->posq, c->posqCorr, c->cellOffsets);
        LAUNCH_COUNT(c, 1);
        b200_peer_broadcast_positions(c);
    } else {
        CUDA_CHECK(cudaMemsetAsync(c->cellOffsets, 0, sizeof(int4) * (size_t)c->Npad, c->stream));
    }
    b200_peer_barrier(c, PEER_SLOT_STATE);
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    afterSetPositions(c);
    API_END
}

extern "C" int b200_set_positions(b200_ctx* c, const double* xyz) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    if (!xyz) throw B200Error(B200_ERR_INVALID, "b200_set_positions: null array");
    stageIn(c, xyz);
    ioSetPositions<<<(c->N + 255) / 256, 256, 0, c->stream>>>(c->dStage, c->atomIndex, c->N, c->posq, c->posqCorr, c->cellOffsets);
    LAUNCH_COUNT(c, 1);
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    afterSetPositions(c);
    API_END
}

extern "C" int b200_get_positions(b200_ctx* c, double* xyz, int32_t enforcePeriodic) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    b200_peer_sync_state(c);
    if (!xyz) {     // multi-GPU: a rank that does not need the host copy still takes part in the collective gather above
        if (c->commRanks > 1) { CUDA_CHECK(cudaStreamSynchronize(c->stream)); return B200_OK; }
        throw B200Error(B200_ERR_INVALID, "b200_get_positions: null array");
    }
    ensureStaging(c);
    ioGetPositions<<<(c->N + 255) / 256, 256, 0, c->stream>>>(c->dStage, c->atomIndex, c->N, c->posq, c->posqCorr, c->cellOffsets, c->boxD, enforcePeriodic ? 0 : 1);
    LAUNCH_COUNT(c, 1);
    stageOut(c, xyz);
    API_END
}

extern "C" int b200_set_velocities(b200_ctx* c, const double* xyz) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    stageIn(c, xyz);
    c->cmValid = false; c->cmDeferred = false;    // a pending COM removal concerned the velocities that are being replaced
    ioSetVelocities<<<(c->N + 255) / 256, 256, 0, c->stream>>>(c->dStage, c->atomIndex, c->N, c->velm);
    LAUNCH_COUNT(c, 1);
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    API_END
}

extern "C" int b200_get_velocities(b200_ctx* c, double* xyz) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    b200_peer_sync_state(c);
    b200_flush_cm(c);
    if (!xyz) {     // multi-GPU: collective part only
        if (c->commRanks > 1) { CUDA_CHECK(cudaStreamSynchronize(c->stream)); return B200_OK; }
        throw B200Error(B200_ERR_INVALID, "b200_get_velocities: null array");
    }
    ensureStaging(c);
    ioGetVelocities<<<(c->N + 255) / 256, 256, 0, c->stream>>>(c->dStage, c->atomIndex, c->N, c->velm);
    LAUNCH_COUNT(c, 1);
    stageOut(c, xyz);
    API_END
}

extern "C" int b200_get_forces(b200_ctx* c, double* xyz) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    b200_ensure_forces_reduced(c);
    if (!xyz) {     // multi-GPU: a rank that does not need the host copy still takes part in the collective sum above
        if (c->commRanks > 1) return B200_OK;
        throw B200Error(B200_ERR_INVALID, "b200_get_forces: null array");
    }
    ensureStaging(c);
    ioGetForces<<<(c->N + 255) / 256, 256, 0, c->stream>>>(c->dStage, c->atomIndex, c->N, c->Npad, c->force);
    LAUNCH_COUNT(c, 1);
    stageOut(c, xyz);
    API_END
}

extern "C" int b200_add_forces(b200_ctx* c, const double* xyz) {
    API_BEGIN
    if (!xyz) throw B200Error(B200_ERR_INVALID, "b200_add_forces: null array");
    if (c->commRanks > 1) throw B200Error(B200_ERR_UNSUPPORTED, "b200_add_forces: not available with a communicator (the force arrays hold partial sums)");
    CUDA_CHECK(cudaSetDevice(c->device));
    stageIn(c, xyz);
    ioAddForces<<<(c->N + 255) / 256, 256, 0, c->stream>>>(c->dStage, c->atomIndex, c->N, c->Npad, c->force);
    LAUNCH_COUNT(c, 1);
    CUDA_CHECK(cudaStreamSynchronize(c->stream));     // the caller's buffer may be pageable and reused at once
    API_END
}

extern "C" int b200_set_box(b200_ctx* c, const double* box9) {
    b200_graph_invalidate(c);
    API_BEGIN
    b200_box_set(c, box9);
    c->forcesValid = false; c->dispChecked = false; c->listVerified = false;
    if (c->finalized) c->stepsSinceReorder = -1;
    API_END
}
extern "C" int b200_get_box(const b200_ctx* c, double* box9) { memcpy(box9, c->boxHost, sizeof(double) * 9); return B200_OK; }
extern "C" int b200_set_time(b200_ctx* c, double t, int64_t steps) {
    API_BEGIN
    if (steps != c->stepCount) { CUDA_CHECK(cudaSetDevice(c->device)); b200_flush_cm(c); c->cmValid = false; }   // the momentum slot is indexed by the step count
    c->timePs = t; c->stepCount = steps;
    API_END
}
extern "C" int b200_get_time(const b200_ctx* c, double* t, int64_t* steps) { if (t) *t = c->timePs; if (steps) *steps = c->stepCount; return B200_OK; }

// ---------------------------------------------------------------------------------------------
// Monte Carlo barostat support: ApplyMonteCarloBarostatKernel (kernels.h:1723-1788; CommonKernels.cpp:4255-4400, monteCarloBarostat.cc)
// One thread per molecule: centre = plain average of the atom positions, displaced by the upper-triangular deformation; all atoms of
// the molecule move by the same vector (rigid molecules) or every atom is its own "molecule".
__global__ void barostatScale(int numMols, const int* __restrict__ molStart, const int* __restrict__ molSlots, int perAtom, int N,
                              float4* posq, float4* posqCorr, double sx, double sy, double sz, double sxy, double sxz, double syz) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= (perAtom ? N : numMols)) return;
    const int first = perAtom ? m : molStart[m], last = perAtom ? m + 1 : molStart[m + 1];
    double cx = 0, cy = 0, cz = 0;
    for (int k = first; k < last; k++) {
        const int s = perAtom ? k : molSlots[k];
        const float4 p = posq[s], q = posqCorr[s];
        cx += (double)p.x + q.x; cy += (double)p.y + q.y; cz += (double)p.z + q.z;
    }
    const double inv = 1.0 / (last - first);
    cx *= inv; cy *= inv; cz *= inv;
    const double dx = cx * (sx - 1.0) + cy * sxy + cz * sxz, dy = cy * (sy - 1.0) + cz * syz, dz = cz * (sz - 1.0);
    for (int k = first; k < last; k++) {
        const int s = perAtom ? k : molSlots[k];
        const float4 p = posq[s], q = posqCorr[s];
        const double x = (double)p.x + q.x + dx, y = (double)p.y + q.y + dy, z = (double)p.z + q.z + dz;
        posq[s] = make_float4((float)x, (float)y, (float)z, p.w);
        posqCorr[s] = make_float4((float)(x - (double)(float)x), (float)(y - (double)(float)y), (float)(z - (double)(float)z), 0.f);
    }
}

static void invalidateAfterMove(b200_ctx* c) {
    c->forcesValid = false; c->dispChecked = false; c->listVerified = false;
    c->stepsSinceReorder = -1;      // the next evaluation re-sorts and rebuilds the list
}

extern "C" int b200_barostat_save(b200_ctx* c) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    b200_peer_sync_state(c);
    const size_t P = c->Npad;
    if (!c->savedPosq) { dmalloc(c->savedPosq, P); dmalloc(c->savedPosqCorr, P); dmalloc(c->savedVelm, P); dmalloc(c->savedAtomIndex, P); dmalloc(c->savedCellOffsets, P); }
    cudaStream_t s = c->stream;
    CUDA_CHECK(cudaMemcpyAsync(c->savedPosq, c->posq, sizeof(float4) * P, cudaMemcpyDeviceToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c->savedPosqCorr, c->posqCorr, sizeof(float4) * P, cudaMemcpyDeviceToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c->savedVelm, c->velm, sizeof(double4) * P, cudaMemcpyDeviceToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c->savedAtomIndex, c->atomIndex, sizeof(int) * P, cudaMemcpyDeviceToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c->savedCellOffsets, c->cellOffsets, sizeof(int4) * P, cudaMemcpyDeviceToDevice, s));
    c->savedValid = true; c->savedStepsSinceReorder = c->stepsSinceReorder;
    API_END
}

extern "C" int b200_barostat_scale(b200_ctx* c, const double* scale6, int32_t rigidMolecules) {
    b200_graph_invalidate(c);
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    if (!c->positionsSet) throw B200Error(B200_ERR_INVALID, "Particle positions have not been set");
    b200_finalize(c);
    b200_peer_sync_state(c);
    const int n = rigidMolecules ? c->numMolecules : c->N;
    barostatScale<<<(n + 127) / 128, 128, 0, c->stream>>>(c->numMolecules, c->molStart, c->molSlots, rigidMolecules ? 0 : 1, c->N, c->posq, c->posqCorr,
                                                         scale6[0], scale6[1], scale6[2], scale6[3], scale6[4], scale6[5]);
    LAUNCH_COUNT(c, 1);
    CUDA_CHECK(cudaGetLastError());
    invalidateAfterMove(c);
    API_END
}

extern "C" int b200_barostat_restore(b200_ctx* c) {
    b200_graph_invalidate(c);
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    if (!c->savedValid) throw B200Error(B200_ERR_INVALID, "b200_barostat_restore: nothing was saved");
    const size_t P = c->Npad;
    cudaStream_t s = c->stream;
    CUDA_CHECK(cudaMemcpyAsync(c->posq, c->savedPosq, sizeof(float4) * P, cudaMemcpyDeviceToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c->posqCorr, c->savedPosqCorr, sizeof(float4) * P, cudaMemcpyDeviceToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c->velm, c->savedVelm, sizeof(double4) * P, cudaMemcpyDeviceToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c->atomIndex, c->savedAtomIndex, sizeof(int) * P, cudaMemcpyDeviceToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c->cellOffsets, c->savedCellOffsets, sizeof(int4) * P, cudaMemcpyDeviceToDevice, s));
    c->hostOrderStale = true;
    invalidateAfterMove(c);
    API_END
}

// ---------------------------------------------------------------------------------------------
// force evaluation
__global__ void zeroEnergy(DevCounters* c) { for (int i = 0; i < 8; i++) c->energy[i] = 0; c->pairCount = 0; c->pairHash = 0; }

static void checkOverflow(b200_ctx* c) {
    // sticky device flag set by the list builder when a capacity was exceeded (the analogue of
    // CudaNonbondedUtilities::updateNeighborListSize, CudaNonbondedUtilities.cpp:441-479)
    CUDA_CHECK(cudaMemcpyAsync(c->countersHost, c->counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
}

__global__ void nbClearOverflow(DevCounters* c) { c->overflow = 0; c->rebuild = 1; }

static void prepareLists(b200_ctx* c, bool verify) {
    // Runs the (device-decided) rebuild.  When `verify` is set the host also reads the list counters: an overflow flag that is
    // ALREADY set belongs to an earlier, unverified evaluation whose forces were incomplete -> error, never cleared silently.
    // Otherwise the storage follows the measured need: it grows (25 % head room) when the build overflowed and also as soon as
    // the list fills more than 90 % of it, so that an overflow between two checks would take an 11 % jump within 64 steps.  (The
    // threshold used to be 80 %, i.e. exactly the fill of freshly sized storage: an equilibrating system whose list grew by half a
    // per cent re-allocated at its first check, ~10 ms of cudaFree/cudaMalloc inside a timed run.)
    // Unverified evaluations stay memory-safe (the builder writes an empty descriptor for a dropped unit, the tile kernel
    // clamps the unit count).  (Host-side analogue of CudaNonbondedUtilities::updateNeighborListSize, CudaNonbondedUtilities.cpp:441-479.)
    if (verify) {
        checkOverflow(c);
        if (c->countersHost->overflow) {
            nbClearOverflow<<<1, 1, 0, c->stream>>>(c->counters);
            if (c->countersHost->overflow == 2) throw B200Error(B200_ERR_OVERFLOW, "a wait for another rank's flag timed out (20 s) during an earlier step; the state computed since the last check is not trustworthy");
            throw B200Error(B200_ERR_OVERFLOW, "the neighbour list overflowed during an earlier, unverified force evaluation; the state computed since the last check is not trustworthy");
        }
    }
    for (int attempt = 0; attempt < 8; attempt++) {
        b200_nblist_prepare(c);
        c->dispChecked = true;     // the positions have been compared with the list's reference positions
        if (!verify) return;
        checkOverflow(c);
        const DevCounters& h = *c->countersHost;
        const bool over = h.overflow != 0;
        const bool tight = (long long)h.numTiles * 10 > c->tileCapacity * 9 || (long long)h.numUnits * 10 > c->unitCapacity * 9;
        c->listVerified = true;
        if (!over && !tight) return;
        c->tileCapacity = std::max(c->tileCapacity, (long long)h.numTiles * 5 / 4 + 1024);
        c->unitCapacity = std::max(c->unitCapacity, (long long)h.numUnits * 5 / 4 + 1024);
        b200_nblist_alloc(c);
        nbClearOverflow<<<1, 1, 0, c->stream>>>(c->counters);     // only these two fields: energy sums and momenta are live on other streams
        LAUNCH_COUNT(c, 1);
    }
    throw B200Error(B200_ERR_OVERFLOW, "neighbour list kept overflowing");
}

void b200_graph_invalidate(b200_ctx* c) {
    c->stepGraphStale = true;      // the next plain evaluation is recorded again; the executable graph is updated in place when the topology still matches
}

static void computeImpl(b200_ctx* c, unsigned groups, bool forces, bool energy, double* energyOut, bool verify) {
    CUDA_CHECK(cudaSetDevice(c->device));
    if (!c->positionsSet) throw B200Error(B200_ERR_INVALID, "Particle positions have not been set");
    b200_finalize(c);
    cudaStream_t s = c->stream;
    // a pushed evaluation nobody consumed: its PME stream may still be reading the arrays the work below rewrites
    if (c->recipPending && c->partRank == 0 && c->usePmeStream && !c->timing) CUDA_CHECK(cudaStreamWaitEvent(s, c->evPmeDone, 0));
    if (c->stepsSinceReorder < 0 || c->stepsSinceReorder >= c->reorderInterval) { b200_peer_sync_state(c); b200_reorder(c); }   // the sort moves whole-state rows
    c->computeCount++;
    // The plain force evaluation of an MD step (forces only, no host read-back, one rank) is a fixed sequence of launches on two
    // streams with constant arguments -- the list kernels decide on the device whether they have work -- so it is captured once as a
    // CUDA graph and replayed: one launch per evaluation instead of ~15 (the reference submits each kernel, CommonKernels.cpp:3139-3206).
    // Anything that changes an argument (box, parameters, list storage) marks the capture stale, see b200_graph_invalidate.
    static const bool graphOff = getenv("B200_NO_GRAPH") != nullptr;
    const bool graphable = !graphOff && forces && !energy && !verify && c->commRanks == 1 && c->partRanks == 1 && !c->timing && c->hasNonbonded
                           && c->computeCount > 4 && c->graphCaptures < 32 && c->computeCount >= c->graphSkipUntil;
    const bool dispAtEntry = c->dispChecked;
    if (graphable && c->stepGraph && !c->stepGraphStale && c->stepGraphGroups == groups && c->stepGraphDisp == dispAtEntry) {
        CUDA_CHECK(cudaGraphLaunch(c->stepGraph, s));
        c->graphReplays++;
        c->launchCount += c->stepGraphLaunches;
        c->dispChecked = true; c->forcesPartial = false; c->recipPending = false;
        c->forcesValid = groups == B200_GROUP_ALL;
        return;
    }
    const bool capturing = graphable;
    const long long launchesAtEntry = c->launchCount;
    if (capturing) {
        // something that invalidates the capture every few evaluations (parameters changed every step, say) would make recording cost
        // more than replaying saves: after three short-lived captures in a row the launch-by-launch path is used for a while
        c->graphPenalty = (c->stepGraph && c->graphReplays < 8) ? c->graphPenalty + 1 : 0;
        c->graphReplays = 0;
        if (c->graphPenalty >= 3) { c->graphPenalty = 0; c->graphSkipUntil = c->computeCount + 256; }
    }
    const bool record = capturing && c->computeCount >= c->graphSkipUntil;
    if (record) { CUDA_CHECK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal)); c->capturing = true; }
    const bool doRecip = c->hasNonbonded && c->nb.method >= B200_NB_EWALD && (groups & B200_GROUP_RECIPROCAL) && c->partRank == 0;
    const bool doDirect = c->hasNonbonded && (groups & B200_GROUP_DIRECT);
    // per-family event timing needs the families serialised: overlapping streams would charge waiting time to the wrong family
    const bool usePmeStream = c->usePmeStream && !c->timing;
    cudaStream_t ps = usePmeStream ? c->pmeStream : s;
    // Multi-GPU over peer memory, plain PME, forces only: rank 0 keeps the reciprocal forces out of its fixed-point force array; its
    // interpolation kernel stores them as floats in a hand-over buffer and raises a flag, and every rank fetches the slice it
    // integrates after the reduce-scatter.  Rank 0's compute stream does not wait for the PME stream, so the barrier / reduce-scatter
    // of the direct-space forces overlaps the end of the reciprocal pipeline, and the last transfer is 12 instead of 24 bytes per atom.
    // Every rank takes the same decision (same arguments, same host tables) and counts the evaluation.
    bool push = false;
    if (forces && !energy && c->peerMode && c->commRanks > 1 && c->hasNonbonded && c->nb.method == B200_NB_PME && (groups & B200_GROUP_RECIPROCAL)) {
        // Opt-in (B200_RECIP_PUSH=1; the 2-GPU tests run it both ways): measured on STMV it does not pay.  2 GPUs 0.927-0.943 against
        // 0.906 ms/step (rank 0 also owns direct-space rows, its PME stream finishes first anyway); 8 GPUs 0.516 against 0.512-0.521
        // (whatever runs beside the interpolation kernel on the PME rank -- its own barrier spin and reduce, the peers' reads -- delays
        // the reciprocal pipeline by about what the overlap saves).
        static const char* env = getenv("B200_RECIP_PUSH");
        const bool on = env && atoi(env) != 0;
        if (on) { b200_peer_plan_shard(c); push = c->pushPlanOk; }
    }
    if (push) { ++c->peerSeq[PEER_SLOT_RECIP]; c->recipPending = true; }
    else if (forces) c->recipPending = false;
    static const bool bondedMain = getenv("B200_BONDED_MAIN") != nullptr;      // A/B switch
    const bool bondedOnPme = !bondedMain && doRecip && usePmeStream && c->commRanks == 1 && c->partRanks == 1 && c->B < 6144;   // (STMV size: 1 % slower there)
    try {
    if (forces) CUDA_CHECK(cudaMemsetAsync(c->force, 0, sizeof(long long) * 3 * (size_t)c->Npad, s));
    if (energy) { zeroEnergy<<<1, 1, 0, s>>>(c->counters); LAUNCH_COUNT(c, 1); }
    if (doRecip) {
        c->recipPushNow = push;
        if (usePmeStream) { CUDA_CHECK(cudaEventRecord(c->evFork, s)); CUDA_CHECK(cudaStreamWaitEvent(ps, c->evFork, 0)); }
        // one rank, small systems (latency-bound kernels): the exclusion corrections and bonded terms (short kernels, integer force sums)
        // ride on the PME stream, off the longer list -> tile chain of the compute stream
        if (bondedOnPme) b200_bonded_launch(c, groups, forces, energy, ps);
        if (c->nb.method == B200_NB_EWALD) b200_ewald_launch(c, ps, forces, energy); else b200_pme_launch(c, ps, forces, energy);
        if (usePmeStream) CUDA_CHECK(cudaEventRecord(c->evPmeDone, ps));
        c->recipPushNow = false;
    }
    if (doDirect) {
        prepareLists(c, verify);
        b200_tile_launch(c, forces, energy, false);
    }
    if (!bondedOnPme && (c->partRank == 0 || c->commRanks > 1)) b200_bonded_launch(c, groups, forces, energy, s);
    static const bool pushJoin = getenv("B200_PUSH_JOIN") != nullptr;     // debug: keep the data path of the push, give up the overlap
    if (doRecip && usePmeStream && (!push || pushJoin)) CUDA_CHECK(cudaStreamWaitEvent(s, c->evPmeDone, 0));
    } catch (...) {
        if (record) { c->capturing = false; cudaGraph_t g = nullptr; cudaStreamEndCapture(s, &g); if (g) cudaGraphDestroy(g); cudaGetLastError(); c->graphCaptures = 1 << 20; }
        throw;
    }
    if (record) {
        cudaGraph_t g = nullptr;
        c->capturing = false;
        CUDA_CHECK(cudaStreamEndCapture(s, &g));
        // same topology as the executable graph at hand (a box or parameter change): update it in place, far cheaper than instantiating
        bool reused = false;
        if (c->stepGraph) {
            cudaGraphExecUpdateResultInfo info;
            if (cudaGraphExecUpdate(c->stepGraph, g, &info) == cudaSuccess) reused = true;
            else { cudaGetLastError(); cudaStreamSynchronize(s); cudaGraphExecDestroy(c->stepGraph); c->stepGraph = nullptr; }
        }
        if (!reused) { c->graphCaptures++; CUDA_CHECK(cudaGraphInstantiate(&c->stepGraph, g, 0)); }
        cudaGraphDestroy(g);
        c->stepGraphStale = false;
        c->stepGraphLaunches = c->launchCount - launchesAtEntry; c->stepGraphGroups = groups; c->stepGraphDisp = dispAtEntry;
        CUDA_CHECK(cudaGraphLaunch(c->stepGraph, s));
    }
    // With a communicator every rank now holds PARTIAL forces.  They are summed when somebody needs the total: the integration step
    // reduce-scatters them over peer memory (b200_advance), b200_get_forces and the shifted kinetic energy all-reduce them.
    c->forcesPartial = forces && c->commRanks > 1;
    if (energy && c->commRanks > 1) b200_comm_allreduce_energy(c);
    c->forcesValid = forces && groups == B200_GROUP_ALL;
    if (energy) {
        checkOverflow(c);
        const DevCounters& h = *c->countersHost;
        double e = 0;
        const double inv = 1.0 / B200_ENERGY_SCALE;
// without a communicator a rank reports only what it computed (N logical ranks on one device are summed by the
        // caller); with one, the integer all-reduce above gave every rank the full device sums and every rank adds the host terms
        const bool all = c->commRanks > 1;
        const bool wantRecip = c->hasNonbonded && c->nb.method >= B200_NB_EWALD && (groups & B200_GROUP_RECIPROCAL) && (c->partRank == 0 || all);
        const bool extras = c->partRank == 0 || all;
        if (doDirect) e += inv * (double)h.energy[E_DIRECT];
        if (doDirect && extras) e += inv * ((double)h.energy[E_EXCLCORR] + (double)h.energy[E_EXCEPT]);
        if (wantRecip) {
            e += inv * (double)h.energy[E_RECIP];
            const double* b = c->boxHost; const double V = b[0] * b[4] * b[8];
            // host-side constants (CommonCalcNonbondedForce.cpp:893-898): self energy and neutralising plasma
            e += c->selfEnergy + c->ljSelfEnergy;
            e -= c->totalCharge * c->totalCharge * M_PI * B200_ONE_4PI_EPS0 / (2.0 * V * c->nb.ewald_alpha * c->nb.ewald_alpha);
        }
        if (doDirect && extras && c->nb.dispersion_coefficient != 0) {
            const double* b = c->boxHost;
            e += c->nb.dispersion_coefficient / (b[0] * b[4] * b[8]);   // CommonCalcNonbondedForce.cpp:1111-1115
        }
        if ((groups & B200_GROUP_BONDED) && extras) e += inv * (double)h.energy[E_BONDED];
        if (energyOut) *energyOut = e;
        if (h.overflow) throw B200Error(B200_ERR_OVERFLOW, "neighbour list overflow detected after the computation");
    }
}

// The host reads the list counters (one small copy + stream sync) when new positions came from outside, on every 64th
// evaluation and whenever an energy is returned anyway; in between, evaluations are queued without any host synchronisation.
void b200_compute_impl(b200_ctx* c, unsigned groups, bool forces, bool energy, double* energyOut) {
    const bool verify = !c->listVerified || energy || (c->computeCount & 63) == 0;
    computeImpl(c, groups, forces, energy, energyOut, verify);
}

extern "C" int b200_compute(b200_ctx* c, uint32_t groups, int32_t includeForces, int32_t includeEnergy, double* energy) {
    API_BEGIN
    b200_compute_impl(c, groups, includeForces != 0, includeEnergy != 0, energy);
    API_END
}

extern "C" int b200_compute_host(b200_ctx* c, const double* xyzIn, uint32_t groups, double* forcesOut, double* energy) {
    API_BEGIN
    int rc = b200_set_positions(c, xyzIn); if (rc) return rc;
    b200_compute_impl(c, groups, true, energy != nullptr, energy);
    rc = b200_get_forces(c, forcesOut); if (rc) return rc;
    API_END
}

extern "C" int b200_direct_pairs(b200_ctx* c, int64_t* count, int32_t* pairs, int64_t capacity) {
    b200_graph_invalidate(c);
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    if (!c->positionsSet || !c->hasNonbonded) throw B200Error(B200_ERR_INVALID, "b200_direct_pairs: positions and a NonbondedForce are required");
    b200_finalize(c);
    if (c->stepsSinceReorder < 0) b200_reorder(c);
    dfree(c->pairOut); c->pairCapacity = 0;
    if (pairs && capacity > 0) { dmalloc(c->pairOut, (size_t)capacity); c->pairCapacity = capacity; }
    zeroEnergy<<<1, 1, 0, c->stream>>>(c->counters);
    prepareLists(c, true);
    b200_tile_launch(c, false, false, true);
    checkOverflow(c);
    *count = (int64_t)c->countersHost->pairCount;
    if (pairs && capacity > 0) {
        const int64_t n = std::min<int64_t>(*count, capacity);
        CUDA_CHECK(cudaMemcpy(pairs, c->pairOut, sizeof(int2) * (size_t)n, cudaMemcpyDeviceToHost));
    }
    dfree(c->pairOut); c->pairCapacity = 0;
    API_END
}

extern "C" int b200_direct_pairs_hash(b200_ctx* c, int64_t* count, uint64_t* hash) {
    b200_graph_invalidate(c);
    API_BEGIN
    int rc = b200_direct_pairs(c, count, nullptr, 0); if (rc) return rc;
    *hash = (uint64_t)c->countersHost->pairHash;
    API_END
}

extern "C" int b200_nblist_get_stats(b200_ctx* c, b200_nblist_stats* out) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    checkOverflow(c);
    out->num_blocks = c->B; out->num_tiles = c->countersHost->numTiles; out->num_units = c->countersHost->numUnits;
    out->num_exclusion_tiles = c->numExclTiles; out->rebuilds = c->countersHost->rebuildCount; out->steps = c->stepCount;
    out->reorders = 0; out->num_sparse_tiles = c->countersHost->numSparseTiles; out->tile_capacity = c->tileCapacity; out->unit_capacity = c->unitCapacity;
    API_END
}

// ---------------------------------------------------------------------------------------------
// stepping
extern "C" int b200_integrator_create(b200_ctx* c, const b200_integrator_desc* d) {
    API_BEGIN
    c->integ = *d; c->hasIntegrator = true;
    API_END
}

void b200_ensure_forces_reduced(b200_ctx* c) {
    if (!c->forcesPartial || c->commRanks < 2) return;
    b200_peer_add_recip(c);             // reciprocal forces handed over separately: each rank adds its slice to its own partial forces first
    b200_comm_allreduce_forces(c);      // collective: every rank reaches this point (int64 sums: exact, order independent)
    c->forcesPartial = false;
}

// Whether the integration runs sharded over peer memory (collective decision, cached): needs the peer mappings, slices that are
// contiguous slot ranges, and no constraint kind that is not sharded yet.
static bool peerStepPlan(b200_ctx* c) {
    { static const bool fp = getenv("B200_FUSED_PUBLISH") != nullptr; c->fusedPublish = fp; }
    bool active = c->peerMode && c->commRanks > 1 && c->integ.remove_cm_every <= 1 && c->hasNonbonded && c->numShake == 0 && c->numGc == 0;
    if (active) {
        b200_peer_plan_shard(c);
        // Slices whose atoms are not one contiguous range of slots (units listed out of slot order) would need the per-atom peer
        // accesses inside the integrator; that variant is compiled but has no multi-GPU test yet, so such systems take the
        // all-reduce + replicated integration path (every rank reaches the same decision: the flag is all-reduced).
        if (!c->shardContiguous) active = false;
    }
    return active;
}

// Everything of an MD step that follows the force evaluation: (multi-GPU: flag barrier, reduce-scatter of the partial forces) ->
// integration with constraints -> (publication of the new positions, flag barrier) -> step bookkeeping.  Shared by b200_step and
// by b200_integrate (the plugin's IntegrateLangevinMiddleStepKernel::execute), so both take the same path.
void b200_advance(b200_ctx* c, cudaEvent_t* ev) {
    auto mark = [&](int k) { if (ev) { cudaEventCreate(&ev[k]); cudaEventRecord(ev[k], c->stream); } };
    c->peerStepActive = peerStepPlan(c);
    if (c->peerStepActive) {
        static const bool separate = getenv("B200_ORDERED_REDUCE") == nullptr;     // default: barrier kernel, then one-pass reduce kernel
        if (separate && c->recipPending && c->partRank == 0 && c->usePmeStream && !c->timing) {
            // The PME rank with the reciprocal forces handed over separately: its fixed-point array is complete, so the others are told
            // at once, but its own barrier spin and reduce kernel wait for the PME stream -- measured on 8 GPUs, running them beside the
            // interpolation kernel delayed the reciprocal pipeline, which everybody waits for, by ~45 us.
            b200_peer_signal_forces_early(c);
            CUDA_CHECK(cudaStreamWaitEvent(c->stream, c->evPmeDone, 0));
        }
        if (separate) {
            b200_peer_barrier(c, PEER_SLOT_FORCES);                       // every rank's partial forces are complete
            mark(2);
            b200_peer_reduce_forces(c);                                   // reduce-scatter over NVLink into the local array
        } else {
            mark(2);
            b200_peer_signal_and_reduce_forces(c);                        // signal, then add every rank's slice as its signal arrives
        }
        b200_peer_add_recip(c);                                           // + this rank's slice of the reciprocal forces from rank 0's hand-over buffer
    } else {
        if (c->commRanks > 1) { b200_peer_sync_state(c); b200_ensure_forces_reduced(c); }     // replicated integration on every rank
        mark(2);
    }
    mark(3);
    b200_integrate_step(c);
    mark(4);
    if (c->peerStepActive) {
        if (!c->fusedPublish) b200_peer_publish_positions(c);             // all-gather
        mark(5);
        b200_peer_barrier(c, PEER_SLOT_POSITIONS);                        // every rank's slice of new positions has arrived everywhere
    } else mark(5);
    mark(6);
    c->stepCount++; c->timePs += c->integ.dt; c->stepsSinceReorder++;
    c->forcesValid = false; c->forcesPartial = false;
    c->peerStepActive = false;
}

extern "C" int b200_step(b200_ctx* c, int32_t nsteps) {
    API_BEGIN
    if (!c->hasIntegrator) throw B200Error(B200_ERR_INVALID, "b200_step: no integrator");
    CUDA_CHECK(cudaSetDevice(c->device));
    b200_finalize(c);
    if (c->numOtherConstraints > 0) throw B200Error(B200_ERR_UNSUPPORTED, "constraints that are neither rigid 3-site molecules (SETTLE) nor clusters of one central atom with up to three singly-constrained partners (SHAKE) are not implemented");
    // B200_PEER_TRACE=1: event time line of the last step of this call (debug aid for the multi-GPU step)
    static const bool trace = getenv("B200_PEER_TRACE") != nullptr;
    cudaEvent_t ev[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    for (int i = 0; i < nsteps; i++) {
        const bool tr = trace && i == nsteps - 1 && nsteps > 1;
        if (tr) { cudaEventCreate(&ev[0]); cudaEventRecord(ev[0], c->stream); }
        if (c->integ.remove_cm_every > 0 && c->stepCount % c->integ.remove_cm_every == 0) b200_request_cm_removal(c);
        computeImpl(c, B200_GROUP_ALL, true, false, nullptr, !c->listVerified || (c->stepCount & 63) == 0);
        if (tr) { cudaEventCreate(&ev[1]); cudaEventRecord(ev[1], c->stream); }
        b200_advance(c, tr ? ev : nullptr);
        if (tr) {
            cudaStreamSynchronize(c->stream);
            float t[7] = {0, 0, 0, 0, 0, 0, 0};
            for (int k = 1; k < 7; k++) if (ev[k] && ev[0]) cudaEventElapsedTime(&t[k], ev[0], ev[k]);
            fprintf(stderr, "[b200 rank %d] step time line (us): forces %.0f | barrier %.0f | reduce %.0f | integrate %.0f | publish %.0f | barrier %.0f\n", c->commRank,
                    1e3 * t[1], 1e3 * (t[2] - t[1]), 1e3 * (t[3] - t[2]), 1e3 * (t[4] - t[3]), 1e3 * (t[5] - t[4]), 1e3 * (t[6] - t[5]));
            for (auto& e : ev) if (e) { cudaEventDestroy(e); e = nullptr; }
        }
    }
    checkOverflow(c);
    if (c->countersHost->overflow == 2) throw B200Error(B200_ERR_OVERFLOW, "a wait for another rank's flag timed out (20 s) during b200_step; state is not trustworthy");
    if (c->countersHost->overflow) throw B200Error(B200_ERR_OVERFLOW, "neighbour list overflowed during b200_step; state is not trustworthy");
    API_END
}

extern "C" int b200_synchronize(b200_ctx* c) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    CUDA_CHECK(cudaStreamSynchronize(c->pmeStream));
    API_END
}

// ---------------------------------------------------------------------------------------------
// measurement
extern "C" int b200_timing_enable(b200_ctx* c, int32_t on) { c->timing = on != 0; return B200_OK; }
extern "C" int b200_timing_reset(b200_ctx* c) {
    for (auto& t : c->timers) { for (auto& p : t.pending) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); } t.pending.clear(); t.ms = 0; t.launches = 0; }
    return B200_OK;
}
extern "C" int b200_timing_read(b200_ctx* c, double* ms, int64_t* launches) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    CUDA_CHECK(cudaDeviceSynchronize());
    for (int i = 0; i < B200_T_COUNT; i++) {
        auto& t = c->timers[i];
        for (auto& p : t.pending) { float m = 0; cudaEventElapsedTime(&m, p.first, p.second); t.ms += m; t.launches++; cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
        t.pending.clear();
        if (ms) ms[i] = t.ms;
        if (launches) launches[i] = t.launches;
    }
    API_END
}
extern "C" int b200_timed_steps(b200_ctx* c, int32_t nsteps, double* totalMs) {
    API_BEGIN
    CUDA_CHECK(cudaSetDevice(c->device));
    cudaEvent_t a, b; CUDA_CHECK(cudaEventCreate(&a)); CUDA_CHECK(cudaEventCreate(&b));
    CUDA_CHECK(cudaDeviceSynchronize());
    CUDA_CHECK(cudaEventRecord(a, c->stream));
    int rc = b200_step(c, nsteps);
    CUDA_CHECK(cudaEventRecord(b, c->stream));
    CUDA_CHECK(cudaEventSynchronize(b));
    float ms = 0; cudaEventElapsedTime(&ms, a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    if (rc) return rc;
    *totalMs = ms;
    API_END
}
extern "C" int64_t b200_launch_count(const b200_ctx* c) { return c->launchCount; }
