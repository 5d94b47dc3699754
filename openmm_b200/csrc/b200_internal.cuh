// b200_internal.cuh -- shared declarations of the B200 NonbondedForce/PME library (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include <stdexcept>
#include "../../include/b200_nonbonded.h"

#define B200_TILE 32
#ifndef B200_UNIT_TILES
#define B200_UNIT_TILES 8            // tiles per work unit (fixed: the float partial-sum grouping must not depend on #GPUs)
#endif
#define B200_ONE_4PI_EPS0 138.93545764438198
#define B200_FORCE_SCALE 4294967296.0   // 2^32, platforms/cuda/src/kernels/common.cu:29-31
#define B200_ENERGY_SCALE 4294967296.0
#define B200_GRID_SCALE 4294967296.0

struct B200Error : public std::runtime_error {
    int code;
    B200Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
void b200_set_error(const std::string& msg);

#define CUDA_CHECK(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) \
    throw B200Error(B200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")"); } while (0)

// ---------------------------------------------------------------------------------------------
// POD parameter blocks passed to kernels by value
struct BoxF {            // reduced-form periodic box, float
    float ax, bx, by, cx, cy, cz;
    float iax, iby, icz; // 1/ax, 1/by, 1/cz
    int   triclinic;
};
struct BoxD {
    double ax, bx, by, cx, cy, cz;
    double iax, iby, icz;
};

struct DevCounters {
    unsigned int numUnits;      // work units in the neighbour list (after build)
    unsigned int numTiles;      // 32-wide j tiles in the list
    unsigned int workNext;      // dynamic work queue head of the tile kernel
    unsigned int overflow;      // sticky: list capacity exceeded
    int          rebuild;       // 1 => the list must be rebuilt before the next tile kernel
    int          rebuildCount;
    unsigned int maxHalfBits[3];// max block half extent per dim, float bits (positive floats order as uints)
    unsigned int numSparseTiles;// sparse tiles (each occupies two 32-int slots of jAtoms) among numTiles
    unsigned long long pairCount;
    unsigned long long pairHash;   // order-independent hash of the emitted pair set (parity probe)
    long long    energy[8];     // fixed point 2^32: 0 direct tiles, 1 exclusion corr, 2 exceptions, 3 reciprocal, 4 bonded, 5 kinetic*2
    long long    cm[4][4];      // fixed point 2^20: sum m v (x,y,z); slots 0-2 rotate through the fused stepping loop, slot 3 is the stand-alone pass
    int          rebuildNow;    // the list phase in progress rebuilds (copy of `rebuild` taken by nbBegin, which clears the request)
    int          pad0;
};
#define E_DIRECT 0
#define E_EXCLCORR 1
#define E_EXCEPT 2
#define E_RECIP 3
#define E_BONDED 4
#define E_KINETIC 5

struct PmeArgs {
    int nx, ny, nz, nzc;
    double r0x, r1x, r1y, r2x, r2y, r2z;   // reciprocal box (rows as in CommonCalcNonbondedForce.cpp:950-957)
    float alphaFactor;                      // pi^2/alpha^2
    float scale;                            // ONE_4PI_EPS0/(pi V); dispersion: -2 pi sqrt(pi)/(6 V)
    int dispersion;                         // 1: dispersion mesh (reciprocalConvolution with USE_LJPME, pme.cc:137-200)
    float bfac, fac1, fac2, fac3;           // dispersion convolution constants (ReferencePME.cpp:336-410)
};

struct FFTPlan1D {
    int n;
    int nstages;
    int radix[12];
};

// peer-memory view of the other ranks' state arrays (CUDA IPC over NVLink), see b200_comm.cu
#define B200_MAX_PEERS 8
struct PeerSet {
    long long* force[B200_MAX_PEERS]; float4* posq[B200_MAX_PEERS]; float4* posqCorr[B200_MAX_PEERS]; double4* velm[B200_MAX_PEERS];
    struct DevCounters* counters[B200_MAX_PEERS]; unsigned long long* flags[B200_MAX_PEERS];
    float* recip[B200_MAX_PEERS];         // per-rank buffer of reciprocal-space forces (only rank 0's, the PME rank's, is filled)
    int n, rank;
};
// where the interpolation kernel stores the reciprocal forces: dst == nullptr -> added to the local fixed-point force array;
// otherwise stored as plain floats (x | y | z, Npad apart) for the owning ranks to fetch (multi-GPU, see computeImpl)
struct PmePush { float* dst; };
#define PEER_SLOT_FORCES 0      // "my forces of step seq are complete"
#define PEER_SLOT_POSITIONS 1   // "my shard of step seq is integrated and written to every rank"
#define PEER_SLOT_STATE 2       // "my shard's velocities / position remainders are written to every rank"
#define PEER_SLOT_RECIP 3       // rank 0: "the reciprocal-space forces of evaluation seq are in my hand-over buffer"
#define PEER_NUM_SLOTS 4

// one reciprocal-space mesh with its transforms
struct PmeMesh {
    int gridN[3] = {0, 0, 0}; int nzc = 0;
    int* gridI = nullptr;           // coefficient grid (spread target), 32-bit fixed point
    float spreadScale = 67108864.f; // fixed-point scale of the shared-memory accumulation in the spread kernel
    float* gridReal = nullptr;      // potential grid after the inverse transform
    float2* gridC = nullptr;        // half-complex spectrum, transformed in place
    float* moduli[3] = {nullptr, nullptr, nullptr};
    float2* twiddle[3] = {nullptr, nullptr, nullptr};
    FFTPlan1D plan[3];
    double alpha = 0;
    bool dispersion = false;        // spread 8 (sigma/2)^3 2 sqrt(eps) instead of the charge; dispersion convolution
};

struct TimerSlot {
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
    double ms = 0; int64_t launches = 0;
};

// ---------------------------------------------------------------------------------------------
struct b200_ctx {
    int device = 0;
    int N = 0, Npad = 0, B = 0, numSuper = 0;
    cudaStream_t stream = nullptr, pmeStream = nullptr, auxStream = nullptr;   // aux: records the conditional body of the captured evaluation
    cudaEvent_t evFork = nullptr, evPmeDone = nullptr;
    bool usePmeStream = true;
    BoxD boxD; BoxF boxF;
    double boxHost[9];

    // per-slot state
    float4* posq = nullptr;        // x,y,z,q
    float4* posqCorr = nullptr;    // mixed-precision remainder
    double4* velm = nullptr;       // vx,vy,vz, 1/m
    float2* sigEps = nullptr;      // sigma/2, 2 sqrt(eps)
    int*    atomIndex = nullptr;   // slot -> original particle
    int4*   cellOffsets = nullptr; // box vectors subtracted by wrapping
    long long* force = nullptr;    // [3*Npad] fixed point
    DevCounters* counters = nullptr;
    DevCounters* countersHost = nullptr;  // pinned
    double* hStage = nullptr; double* dStage = nullptr;   // pinned host / device staging ([3N] doubles, caller's order) for state I/O
    std::vector<cudaEvent_t> ioEvents;                    // one per chunk of a pipelined device-to-host copy
    bool chargesUploaded = false;
    // scratch for reorder
    float4* posqTmp = nullptr; float4* posqCorrTmp = nullptr; double4* velmTmp = nullptr; int* atomIndexTmp = nullptr; int4* cellOffsetsTmp = nullptr;

    std::vector<int> hostAtomIndex;   // slot -> orig (host mirror)
    std::vector<int> hostSlotOf;      // orig -> slot
    bool hostOrderStale = false;
    std::vector<double> mass;         // original order
    bool positionsSet = false;

    // constraints (slot indices == original indices at creation; permutation keeps them valid)
    int numConstraints = 0;
    std::vector<int> conAtoms; std::vector<double> conDist;
    int numSettle = 0; int4* settleAtoms = nullptr; float2* settleParams = nullptr;  // (a0,a1,a2,-), (dist01, dist12)
    int numFree = 0; int* freeAtoms = nullptr;      // unconstrained, mobile atoms
    int numShake = 0; int4* shakeAtoms = nullptr; float4* shakeDist = nullptr;   // SHAKE clusters: (central, up to three partners or -1), their distances
    double totalMassCached = 0;
    std::vector<int> hSettleAtoms, hFreeAtoms;      // host copies (4 ints per cluster), used to plan the multi-GPU integration slices
    int numOtherConstraints = 0;
    // general constraint clusters (anything that is neither SETTLE nor a SHAKE star): one CTA per connected cluster, constraints
    // ordered by colour (no two constraints of a colour share an atom) -- b200_integrate.cu generalCluster*
    int numGc = 0, numGcClusters = 0; int2* gcPairs = nullptr; double* gcDist = nullptr; int* gcColorStart = nullptr; int2* gcClusterColors = nullptr;
    int* gcAtomStart = nullptr; int* gcAtomList = nullptr; double4* gcDelta = nullptr; double4* gcOld = nullptr;

    // molecules
    bool finalized = false;
    int numMolecules = 0;
    struct MolGroup { int atomsPerMol; int numMols; int* dBase = nullptr; std::vector<int> base; };
    std::vector<MolGroup> groups;
    std::vector<int> molBaseOf;     // particle -> first particle of its molecule (creation order), for re-checking group homogeneity
    int* sortKeys = nullptr; int* sortCount = nullptr; int* sortStart = nullptr; int* sortOrder = nullptr; int sortBuckets = 0; int sortBits = 0;
    int maxGroupMols = 0;
    // every molecule as a list of slots (CSR), for the Monte Carlo barostat's rigid-molecule scaling
    int* molStart = nullptr; int* molSlots = nullptr;
    // state saved by b200_barostat_save (rejected Monte Carlo move): positions, remainders, velocities, order, wrap counters
    float4* savedPosq = nullptr; float4* savedPosqCorr = nullptr; double4* savedVelm = nullptr; int* savedAtomIndex = nullptr; int4* savedCellOffsets = nullptr;
    bool savedValid = false; int64_t savedStepsSinceReorder = 0;
    int reorderInterval = 250;
    int64_t stepsSinceReorder = 0;

    // nonbonded
    bool hasNonbonded = false;
    b200_nonbonded_desc nb;            // scalars only are valid after create (pointers nulled)
    std::vector<double> charge, sigma, epsilon;
    std::vector<int> exclPairs;        // all excluded pairs
    std::vector<int> excPairs; std::vector<double> excQQ, excSig, excEps;
    double paddedCutoff = 0, padding = 0;
    double selfEnergy = 0, totalCharge = 0;
    bool exactEwald = false;
    // exclusion tiles
    int numExclTiles = 0; int2* exclTiles = nullptr; unsigned int* exclFlags = nullptr;
    int* exclRowStart = nullptr; int* exclRowCols = nullptr;
    // exclusion correction list + exceptions
    int numExclPairs = 0; int2* dExclPairs = nullptr;
    int numExceptions = 0; int2* dExcPairs = nullptr; float4* dExcParams = nullptr;
    // neighbour list
    float4* blockCenter = nullptr; float4* blockHalf = nullptr; float4* superCenter = nullptr; float4* superHalf = nullptr;
    float4* refPosq = nullptr;
    int4* units = nullptr; int* jAtoms = nullptr; long long unitCapacity = 0, tileCapacity = 0;
    int2* pairOut = nullptr; long long pairCapacity = 0;
    int partRank = 0, partRanks = 1;
    struct OwnerTable { unsigned char owner[64]; } partTable;   // row X belongs to rank owner[X & 63] (weighted round-robin)
    double partShareBegin = 0.0, partShareEnd = 1.0;           // this rank's slice of the exclusion-tile index range
    int sparseMax = 5;             // j atoms reaching at most this many atoms of X are handled as sparse pairs

    // fused reduce + integrate + gather over peer memory (b200_comm.cu, b200_integrate.cu)
    bool peerInProcess = false;     // the peers are contexts of this process (raw pointers + cudaDeviceEnablePeerAccess instead of IPC)
    bool peerMode = false;          // IPC mappings established on every rank
    bool fusedPublish = false;      // B200_FUSED_PUBLISH=1: the integration kernel stores the new positions to the peers itself
    bool peerStepActive = false;    // inside b200_step in peer mode: no NCCL all-reduce of the forces
    bool shardsDirty = false;       // velm / posqCorr are current only for this rank's shard of the integration units
    PeerSet peers;
    unsigned long long* peerFlags = nullptr;
    unsigned long long peerSeq[PEER_NUM_SLOTS] = {0, 0, 0, 0};
    int shardBegin = 0, shardEnd = 0;   // this rank's range of integration units (SETTLE clusters, then free atoms)
    int shardAtomBegin = 0, shardAtomEnd = 0; bool shardContiguous = false; int shardUnitsFor = -1;   // atoms of those units, when they form one slot range
    // Reciprocal forces handed over separately: rank 0's interpolation kernel stores them as floats in recipF and raises a flag; every
    // rank fetches the slice it integrates (NVLink loads, 12 bytes per atom instead of the 24 of the fixed-point array) after the
    // reduce-scatter of the direct-space forces, which therefore no longer waits for the PME stream.
    float* recipF = nullptr;        // 3 * Npad floats
    int shardAtomBounds[B200_MAX_PEERS + 1] = {0}; bool pushPlanOk = false;   // every rank's atom range; valid when the ranges tile [0, N) in rank order
    bool recipPushNow = false;      // set around b200_pme_launch: this evaluation pushes
    bool recipPending = false;      // the force arrays lack the reciprocal part; it is (or will be) in rank 0's hand-over buffer

    // the captured force evaluation (computeImpl): replayed while nothing that feeds a kernel argument has changed
    bool capturing = false;         // computeImpl is recording the evaluation: b200_nblist_prepare puts the rebuild kernels into a conditional node
    cudaGraphExec_t stepGraph = nullptr; long long stepGraphLaunches = 0; unsigned stepGraphGroups = 0; bool stepGraphDisp = false; int graphCaptures = 0;
    bool stepGraphStale = false;                 // an argument changed: record again and update the executable graph in place (cudaGraphExecUpdate)
    int graphReplays = 0, graphPenalty = 0; long long graphSkipUntil = 0;   // back-off when something invalidates the capture every few evaluations

    // PME: mesh[0] = Coulomb, mesh[1] = dispersion (LJPME only)
    PmeMesh mesh[2];
    bool ljpme = false;
    double2* ewaldS = nullptr; int ewaldNumK = 0;   // NonbondedForce::Ewald: structure factors of the k vectors
    double ljSelfEnergy = 0;        // + alpha_d^6/12 * sum c6_i^2 (ReferenceLJCoulombIxn.cpp:224-227)

    // bonded
    std::vector<int> bondedActive[3];       // (first, count) pairs selected by b200_bonded_set_active; empty = every term of the kind
    int numBonds = 0; int2* bondAtoms = nullptr; float2* bondParams = nullptr;
    int numAngles = 0; int4* angleAtoms = nullptr; float2* angleParams = nullptr;
    int numTorsions = 0; int4* torsionAtoms = nullptr; float4* torsionParams = nullptr;
    std::vector<int> hBondAtoms, hAngleAtoms, hTorsionAtoms; std::vector<double> hBondParams, hAngleParams, hTorsionParams; std::vector<int> hTorsionPeriod;

    int bondedPeriodic[3] = {0, 0, 0};   // bonds / angles / torsions use the nearest periodic image of every displacement

    // integrator
    bool hasIntegrator = false;
    b200_integrator_desc integ;
    double timePs = 0; int64_t stepCount = 0;
    bool forcesValid = false;
    bool forcesPartial = false;    // multi-GPU: the force buffer holds this rank's contribution only (summed on demand)
    bool cmValid = false;          // device momentum buffer cm[stepCount % 3] matches the current velocities
    bool cmDeferred = false;       // a COM-motion removal was requested; the next integration kernel subtracts it (b200_flush_cm applies it early)
    bool dispChecked = false;      // the last integration kernel already ran the displacement check on the current positions
    bool listVerified = false;     // the host has read the list counters since the positions were last set from outside
    int64_t computeCount = 0;      // force evaluations since creation (lazy overflow check every 64th)

    // multi-GPU
    void* ncclComm = nullptr; int commRank = 0, commRanks = 1;

    // measurement
    bool timing = false;
    TimerSlot timers[B200_T_COUNT];
    int64_t launchCount = 0;
    int smCount = 148;
};

// RAII-less helpers
template <typename T> static inline void dmalloc(T*& p, size_t n) { CUDA_CHECK(cudaMalloc((void**)&p, (n ? n : 1) * sizeof(T))); }
template <typename T> static inline void dfree(T*& p) { if (p) cudaFree(p); p = nullptr; }
template <typename T> static inline void h2d(T* d, const T* h, size_t n, cudaStream_t s) { if (n) CUDA_CHECK(cudaMemcpyAsync(d, h, n * sizeof(T), cudaMemcpyHostToDevice, s)); }
template <typename T> static inline void d2h(T* h, const T* d, size_t n, cudaStream_t s) { if (n) CUDA_CHECK(cudaMemcpyAsync(h, d, n * sizeof(T), cudaMemcpyDeviceToHost, s)); }

struct ScopedTimer {   // brackets a kernel family with events when ctx->timing is on
    b200_ctx* c; int slot; cudaStream_t s; cudaEvent_t a = nullptr, b = nullptr;
    ScopedTimer(b200_ctx* c_, int slot_, cudaStream_t s_);
    ~ScopedTimer();
};

// module entry points (host side)
void b200_box_set(b200_ctx* c, const double* box9);
void b200_finalize(b200_ctx* c);                       // molecules, exclusion tiles, clusters (lazy, once)
void b200_reorder(b200_ctx* c);                        // device-side Hilbert sort of identical molecules
void b200_sync_host_order(b200_ctx* c);
void b200_nblist_alloc(b200_ctx* c);
void b200_nblist_prepare(b200_ctx* c);                 // displacement check + conditional rebuild (device-side decision)
void b200_nblist_force_rebuild(b200_ctx* c);
void b200_tile_launch(b200_ctx* c, bool forces, bool energy, bool emitPairs);
void b200_pme_create(b200_ctx* c);
void b200_pme_destroy(b200_ctx* c);
void b200_pme_rescale(b200_ctx* c);
void b200_pme_launch(b200_ctx* c, cudaStream_t s, bool forces, bool energy);
void b200_ewald_launch(b200_ctx* c, cudaStream_t s, bool forces, bool energy);   // NonbondedForce::Ewald reciprocal sum (b200_ewald.cu)
void b200_bonded_upload(b200_ctx* c);
void b200_bonded_launch(b200_ctx* c, unsigned groups, bool forces, bool energy, cudaStream_t s);
void b200_integrate_setup(b200_ctx* c);
void b200_integrate_step(b200_ctx* c);
void b200_advance(b200_ctx* c, cudaEvent_t* traceEvents);   // after the forces: (barrier, reduce) integrate (publish, barrier), bookkeeping
void b200_ensure_forces_reduced(b200_ctx* c);          // multi-GPU: all-reduce partial forces if nobody has summed them yet (collective)
void b200_request_cm_removal(b200_ctx* c);           // RemoveCMMotion: deferred into the next integration kernel when possible
void b200_flush_cm(b200_ctx* c);                        // apply a deferred COM-motion removal now (before anything reads the velocities)
void b200_comm_allreduce_forces(b200_ctx* c);
void b200_comm_allreduce_energy(b200_ctx* c);
void b200_comm_destroy(b200_ctx* c);
void b200_graph_invalidate(b200_ctx* c);              // drop the captured force evaluation (an argument of one of its kernels changed)
void b200_peer_barrier(b200_ctx* c, int slot);        // signal + wait on the compute stream (device-side, no host sync)
void b200_peer_sync_state(b200_ctx* c);
void b200_peer_plan_shard(b200_ctx* c);               // unit slice of this rank and whether its atoms are one contiguous slot range
void b200_peer_reduce_forces(b200_ctx* c);
void b200_peer_signal_forces_early(b200_ctx* c);
void b200_peer_signal_recip(b200_ctx* c, cudaStream_t s);   // rank 0, after the interpolation kernel: "hand-over buffer is written"
void b200_peer_add_recip(b200_ctx* c);                // wait for that signal, fetch this rank's slice from rank 0 and add it to the force array
void b200_peer_signal_and_reduce_forces(b200_ctx* c); // barrier + reduce-scatter in one kernel, peers added in the order they finish            // sum of all ranks' partial forces for this rank's atom range -> local force array
void b200_peer_broadcast_positions(b200_ctx* c);      // rank 0's whole posq / posqCorr -> every rank
void b200_peer_publish_positions(b200_ctx* c);        // this rank's slice of posq -> every rank               // all-gather of the sharded velocities / position remainders (collective)
void b200_fft_make_plan(int n, FFTPlan1D& plan);
void b200_fft_make_twiddle(int n, float2*& dTw, cudaStream_t s);
void b200_fft3d_reciprocal(b200_ctx* c, PmeMesh& m, cudaStream_t s, const PmeArgs& a, bool energy, bool forces);   // gridI -> (convolution, energy) -> gridReal
void b200_fft3d_exec_r2c(int nx, int ny, int nz, const FFTPlan1D* plan, float2* const* tw, const float* inReal, const int* inFix, float fixScaleInv, float2* out, float2* scratch, cudaStream_t s);
void b200_fft3d_exec_c2r(int nx, int ny, int nz, const FFTPlan1D* plan, float2* const* tw, float2* in, float2* scratch, float* out, cudaStream_t s);
// register-resident two-stage passes (b200_fft2.cu); used when every dimension has a plan, else the generic passes of b200_fft.cu run
bool b200_fft2_supported(int nx, int ny, int nz);
void b200_fft2_strided(bool inverse, float2* data, int n, int innerCount, int outerCount, size_t outerStride, size_t stride, const float2* tw, cudaStream_t s);
void b200_fft2_xconv(float2* data, int nx, int innerCount, size_t stride, const float2* tw, const PmeArgs& a, const float* modX, const float* modY, const float* modZ,
                     DevCounters* counters, bool energy, bool inverseToo, cudaStream_t s);
void b200_fft2_z_forward(int numLines, int nz, const float* inReal, const int* inFix, float fixScaleInv, float2* out, const float2* twZ, cudaStream_t s);
void b200_fft2_z_inverse(int numLines, int nz, const float2* in, float* out, const float2* twZ, cudaStream_t s);
void b200_compute_impl(b200_ctx* c, unsigned groups, bool forces, bool energy, double* energyOut);

#define LAUNCH_COUNT(c, n) ((c)->launchCount += (n))

// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ void imageDelta(float& dx, float& dy, float& dz, const BoxF& b) {
    // nearest periodic image of a displacement in a reduced triclinic cell (z, then y, then x);
    // same reduction order as the reference's APPLY_PERIODIC_TO_DELTA (platforms/common/src/kernels, triclinic branch)
    float s = rintf(dz * b.icz); dx -= s * b.cx; dy -= s * b.cy; dz -= s * b.cz;
    s = rintf(dy * b.iby); dx -= s * b.bx; dy -= s * b.by;
    s = rintf(dx * b.iax); dx -= s * b.ax;
}
__device__ __forceinline__ void imageDeltaD(double& dx, double& dy, double& dz, const BoxD& b) {
    double s = rint(dz * b.icz); dx -= s * b.cx; dy -= s * b.cy; dz -= s * b.cz;
    s = rint(dy * b.iby); dx -= s * b.bx; dy -= s * b.by;
    s = rint(dx * b.iax); dx -= s * b.ax;
}
// f * 2^32 is exact in float (a power-of-two scaling), so truncating it directly gives the same integer as the reference's
// (long long)((double)f * 2^32) with one FMUL + F2I instead of F2F + DMUL + F2I
__device__ __forceinline__ long long toFixed(float f) { return __float2ll_rz(f * 4294967296.0f); }
__device__ __forceinline__ long long toFixedD(double f) { return (long long)(f * B200_FORCE_SCALE); }
__device__ __forceinline__ void atomicAddFixed(long long* p, long long v) {
    atomicAdd((unsigned long long*)p, (unsigned long long)v);
}
__device__ __forceinline__ float warpSum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warpSumD(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
#endif
