// b200_comm.cu -- multi-GPU: one process per GPU, direct-space rows dealt round-robin to ranks.  Inside the stepping loop the
// fixed-point forces are reduced, the system integrated and the positions exchanged by kernels working on peer memory (see
// "Peer-memory step" below); force/energy evaluations outside it (getState) sum the forces with an NCCL all-reduce over NVLink.
//
// Replaces the reference's single-process multi-GPU path: peer-to-peer cuMemcpyAsync of posq and of each device's
// force buffer into a slot on device 0, followed by the sumForces kernel
// (platforms/cuda/src/CudaParallelKernels.cpp:183-269, platforms/cuda/src/kernels/parallel.cu:5-13).
// Because the force sum is an integer sum it is exact and order independent: results are bit-identical to a one-rank run
// whichever way the sum is taken.  PME and the bonded terms run on rank 0 only (CommonCalcNonbondedForce.cpp:394,437).
// NCCL is resolved at run time (dlopen) so the library has no hard dependency on a particular libnccl build;
// inside a torch process this binds to the NCCL that torch already loaded.
#include "b200_internal.cuh"
#include <dlfcn.h>
#include <unistd.h>
#include <cstdio>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclInt64 = 4, ncclSum = 0 };

static struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
} g_nccl;

static void loadNccl() {
    if (g_nccl.lib) return;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (g_nccl.lib) break; }
    if (!g_nccl.lib) throw B200Error(B200_ERR_NCCL, "cannot load libnccl.so.2");
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(g_nccl.lib, "ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(g_nccl.lib, "ncclCommInitRank");
    g_nccl.AllReduce = (decltype(g_nccl.AllReduce))dlsym(g_nccl.lib, "ncclAllReduce");
    g_nccl.AllGather = (decltype(g_nccl.AllGather))dlsym(g_nccl.lib, "ncclAllGather");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(g_nccl.lib, "ncclCommDestroy");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(g_nccl.lib, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy) throw B200Error(B200_ERR_NCCL, "libnccl is missing required symbols");
}
#define NCCL_CHECK(expr) do { ncclResult_t _r = (expr); if (_r != 0) throw B200Error(B200_ERR_NCCL, std::string(#expr) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(_r) : "NCCL error")); } while (0)

extern "C" int b200_nccl_unique_id(char* id128) {
    try { loadNccl(); ncclUniqueId id; NCCL_CHECK(g_nccl.GetUniqueId(&id)); memcpy(id128, id.internal, 128); }
    catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
    return B200_OK;
}

extern "C" int b200_set_partition(b200_ctx* c, int32_t rank, int32_t nranks) {
    b200_graph_invalidate(c);
    if (!c || nranks < 1 || rank < 0 || rank >= nranks) { b200_set_error("b200_set_partition: bad rank"); return B200_ERR_INVALID; }
    c->partRank = rank; c->partRanks = nranks;
    // Rank 0 also runs PME, the exclusion corrections and the bonded terms, so it takes a smaller share s0 of the direct-space
    // rows: balance s0*D + P = (1-s0)/(R-1)*D with rho = P/D  =>  s0 = (1 - (R-1) rho)/R  (clamped at 0: a PME-only rank).
    // The reference balances dynamically from measured times (CudaParallelKernels.cpp:247-266); a fixed model suffices here
    // because rows are sampled uniformly.  B200_PME_COST_RATIO overrides rho.
    double rho = 0.38;   // measured on STMV, 2 ranks: flat optimum between 0.30 and 0.40
    if (const char* e = getenv("B200_PME_COST_RATIO")) rho = atof(e);
    std::vector<double> share(nranks, 1.0 / nranks);
    if (nranks > 1 && c->commRanks > 1) {   // only with a communicator: logical ranks on one device are plain equal shares
        const double s0 = std::max(0.0, (1.0 - (nranks - 1) * rho) / nranks);
        share[0] = s0;
        for (int r = 1; r < nranks; r++) share[r] = (1.0 - s0) / (nranks - 1);
    }
    // weighted round-robin over a period of 64 rows (largest-remainder / Bresenham spreading)
    std::vector<double> credit(nranks, 0.0);
    for (int i = 0; i < 64; i++) {
        int best = 0;
        for (int r = 0; r < nranks; r++) { credit[r] += share[r]; if (credit[r] > credit[best] + 1e-12) best = r; }
        c->partTable.owner[i] = (unsigned char)best;
        credit[best] -= 1.0;
    }
    double begin = 0; for (int r = 0; r < rank; r++) begin += share[r];
    c->partShareBegin = begin; c->partShareEnd = (rank == nranks - 1) ? 1.0 : begin + share[rank];
    if (c->finalized) c->stepsSinceReorder = -1;   // forces a list rebuild with the new row assignment
    return B200_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Peer-memory step (no counterpart in the reference, which funnels everything through device 0).
// Every rank maps the force / posq / posqCorr / velm / counter arrays of all ranks (CUDA IPC; NVLink loads and stores).
// One MD step is then
//     forces (local partial sums)  ->  barrier  ->  fused reduce + integrate + gather kernel  ->  barrier
// where the integration kernel of rank r handles the r-th slice of the integration units: it sums the fixed-point force
// contributions of all ranks for its atoms straight out of their memories (the reduce-scatter), integrates and constrains,
// and stores the new float4 positions into every rank's posq (the all-gather).  This replaces ncclAllReduce(25.6 MB for
// STMV) + R redundant integrations.  The barriers are device-side flag exchanges in peer memory; the host never waits.
// A rank in ANOTHER process is reached through CUDA IPC handles; a rank in the SAME process (one OpenMM Context driving several GPUs,
// DeviceIndex=0,1,...) through its raw pointers after cudaDeviceEnablePeerAccess (IPC handles cannot be opened by their own process).
struct IpcBundle { cudaIpcMemHandle_t h[7]; void* raw[7]; long long pid; int device; };

// signal and wait in one launch: lane r tells rank r that this rank reached `seq`, then waits for rank r's flag
__global__ void peerBarrierKernel(PeerSet p, unsigned long long* flags, int slot, unsigned long long seq, DevCounters* counters) {
    if ((int)threadIdx.x < p.n) {
        __threadfence_system();
        volatile unsigned long long* out = p.flags[threadIdx.x] + slot * B200_MAX_PEERS + p.rank;
        *out = seq;
        __threadfence_system();
        volatile unsigned long long* f = flags + slot * B200_MAX_PEERS + threadIdx.x;
        unsigned long long t0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (*f < seq) {
            unsigned long long t1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 20000000000ull) { counters->overflow = 2; break; }   // 20 s: a rank died; fail loudly instead of hanging the GPU
        }
        __threadfence_system();
    }
}

// Signal only: "my work up to here is complete" to every rank (the waiting half lives in the consumer, see peerReduceForcesOrdered)
__global__ void peerSignalKernel(PeerSet p, int slot, unsigned long long seq) {
    if ((int)threadIdx.x < p.n) {
        __threadfence_system();
        volatile unsigned long long* out = p.flags[threadIdx.x] + slot * B200_MAX_PEERS + p.rank;
        *out = seq;
        __threadfence_system();
    }
}

void b200_peer_barrier(b200_ctx* c, int slot) {
    ScopedTimer tm(c, B200_T_COMM, c->stream);
    const unsigned long long seq = ++c->peerSeq[slot];
    peerBarrierKernel<<<1, 32, 0, c->stream>>>(c->peers, c->peerFlags, slot, seq, c->counters);
    LAUNCH_COUNT(c, 1);
}

// velocities and position remainders of this rank's units -> every rank
__global__ void peerGatherState(PeerSet p, int unitBegin, int unitEnd, int numSettle, const int4* __restrict__ settleAtoms, const int* __restrict__ freeAtoms) {
    const int t = unitBegin + blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= unitEnd) return;
    int idx[3]; int cnt;
    if (t < numSettle) { const int4 a = settleAtoms[t]; idx[0] = a.x; idx[1] = a.y; idx[2] = a.z; cnt = 3; }
    else { idx[0] = freeAtoms[t - numSettle]; cnt = 1; }
    for (int k = 0; k < cnt; k++) {
        const double4 v = p.velm[p.rank][idx[k]]; const float4 q = p.posqCorr[p.rank][idx[k]];
        for (int r = 0; r < p.n; r++) if (r != p.rank) { p.velm[r][idx[k]] = v; p.posqCorr[r][idx[k]] = q; }
    }
}

void b200_peer_sync_state(b200_ctx* c) {
    if (!c->peerMode || !c->shardsDirty) return;
    const int units = c->shardEnd - c->shardBegin;
    if (units > 0) { peerGatherState<<<(units + 127) / 128, 128, 0, c->stream>>>(c->peers, c->shardBegin, c->shardEnd, c->numSettle, c->settleAtoms, c->freeAtoms); LAUNCH_COUNT(c, 1); }
    b200_peer_barrier(c, PEER_SLOT_STATE);
    c->shardsDirty = false;
}

// Slice of the integration units owned by this rank (cut at multiples of the 128-unit blocks, so block-level partial sums are
// grouped as on one rank) and, when the atoms of the slice are exactly one range of slots (always the case for solvent boxes,
// where units follow the slot order), that range: the reduce and the publication then run as fully coalesced range operations.
void b200_peer_plan_shard(b200_ctx* c) {
    const int units = c->numSettle + c->numFree;
    if (c->shardUnitsFor == units) return;
    const int R = c->commRanks, nblk = (units + 127) / 128;
    c->shardBegin = std::min(units, 128 * (int)(((long long)nblk * c->commRank) / R));
    c->shardEnd = std::min(units, 128 * (int)(((long long)nblk * (c->commRank + 1)) / R));
    long long lo = 1LL << 40, hi = -1, cnt = 0;
    for (int t = c->shardBegin; t < c->shardEnd; t++) {
        int idx[3]; int n;
        if (t < c->numSettle) { idx[0] = c->hSettleAtoms[4 * (size_t)t]; idx[1] = c->hSettleAtoms[4 * (size_t)t + 1]; idx[2] = c->hSettleAtoms[4 * (size_t)t + 2]; n = 3; }
        else { idx[0] = c->hFreeAtoms[t - c->numSettle]; n = 1; }
        for (int k = 0; k < n; k++) { lo = std::min<long long>(lo, idx[k]); hi = std::max<long long>(hi, idx[k]); cnt++; }
    }
    int contiguous = (cnt == 0) || (hi - lo + 1 == cnt);
    c->shardAtomBegin = cnt ? (int)lo : 0; c->shardAtomEnd = cnt ? (int)hi + 1 : 0;
    // Atom range of EVERY rank (for the reciprocal-force push): usable when the ranges follow each other in rank order and cover
    // all N atoms.  Each rank evaluates the same host tables, so all ranks reach the same verdict.
    bool tiles = true; int next = 0;
    for (int r = 0; r < R; r++) {
        const int ub = std::min(units, 128 * (int)(((long long)nblk * r) / R)), ue = std::min(units, 128 * (int)(((long long)nblk * (r + 1)) / R));
        long long l = 1LL << 40, h = -1, n = 0;
        for (int t = ub; t < ue; t++) {
            if (t < c->numSettle) for (int k = 0; k < 3; k++) { const long long a = c->hSettleAtoms[4 * (size_t)t + k]; l = std::min(l, a); h = std::max(h, a); n++; }
            else { const long long a = c->hFreeAtoms[t - c->numSettle]; l = std::min(l, a); h = std::max(h, a); n++; }
        }
        c->shardAtomBounds[r] = next;
        if (n > 0) { if (l != next || h - l + 1 != n) tiles = false; next = (int)h + 1; }
    }
    c->shardAtomBounds[R] = next;
    c->pushPlanOk = tiles && next == c->N;
    // every rank must take the same code path
    long long* d = nullptr; dmalloc(d, 1);
    long long v = contiguous ? 0 : 1;
    CUDA_CHECK(cudaMemcpy(d, &v, sizeof(v), cudaMemcpyHostToDevice));
    NCCL_CHECK(g_nccl.AllReduce(d, d, 1, ncclInt64, ncclSum, (ncclComm_t)c->ncclComm, c->stream));
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    CUDA_CHECK(cudaMemcpy(&v, d, sizeof(v), cudaMemcpyDeviceToHost));
    dfree(d);
    c->shardContiguous = v == 0;
    c->shardUnitsFor = units;
}

// the reduce-scatter: coalesced loads from every rank's force array over NVLink, exact integer sums, result in the local array.
// Only entries of this rank's own atom range are touched (the other ranks read theirs concurrently).
__global__ void __launch_bounds__(256) peerReduceForces(PeerSet p, int Npad, int aBegin, int aEnd) {
    const int count = aEnd - aBegin;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < 3 * count; k += gridDim.x * blockDim.x) {
        const int comp = k / count, j = k - comp * count;
        const size_t off = (size_t)comp * Npad + aBegin + j;
        long long s = p.force[p.rank][off];
        for (int r = 0; r < p.n; r++) if (r != p.rank) s += p.force[r][off];
        p.force[p.rank][off] = s;
    }
}
void b200_peer_reduce_forces(b200_ctx* c) {
    ScopedTimer tm(c, B200_T_COMM, c->stream);
    if (c->shardAtomEnd > c->shardAtomBegin) { peerReduceForces<<<c->smCount * 4, 256, 0, c->stream>>>(c->peers, c->Npad, c->shardAtomBegin, c->shardAtomEnd); LAUNCH_COUNT(c, 1); }
}

// The reciprocal forces arrive separately: rank 0's interpolation kernel stored them as floats in its recipF (PmePush) and then
// raised PEER_SLOT_RECIP.  Wait for the flag, fetch and add the slice.  Integer addition: the total equals the all-in-one sum bit for bit.
// The wait is a kernel of its own with ONE warp: a spinning grid that fills every SM (as the adding kernel would) can keep a block of
// the reciprocal pipeline that needs a whole SM's registers from ever being scheduled on this GPU -- the flag would never be raised
// (seen as a 20 s time-out on two GPUs, whenever the compute stream reached this point before the PME stream had finished).
__global__ void peerWaitRecip(const unsigned long long* flags, unsigned long long seq, DevCounters* counters) {
    if (threadIdx.x == 0) {
        volatile const unsigned long long* f = flags + PEER_SLOT_RECIP * B200_MAX_PEERS + 0;
        unsigned long long t0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (*f < seq) {
            unsigned long long t1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 20000000000ull) { counters->overflow = 2; break; }   // 20 s: a rank died; fail loudly instead of hanging the GPU
        }
        __threadfence_system();
    }
}
// fetch this rank's slice of the PME rank's float forces (remote loads for every rank but rank 0), convert exactly as the
// interpolation kernel would have (toFixed of the same float: the same bits) and add
__global__ void __launch_bounds__(256) peerAddRecip(long long* __restrict__ force, const float* __restrict__ recip, int Npad, int aBegin, int aEnd) {
    const int count = aEnd - aBegin;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < 3 * count; k += gridDim.x * blockDim.x) {
        const int comp = k / count, j = k - comp * count;
        const size_t off = (size_t)comp * Npad + aBegin + j;
        force[off] += toFixed(__ldcv(recip + off));
    }
}
// "my forces of the coming barrier are complete", raised ahead of the barrier kernel (which repeats the same value and then waits)
void b200_peer_signal_forces_early(b200_ctx* c) {
    peerSignalKernel<<<1, 32, 0, c->stream>>>(c->peers, PEER_SLOT_FORCES, c->peerSeq[PEER_SLOT_FORCES] + 1);
    LAUNCH_COUNT(c, 1);
}
void b200_peer_signal_recip(b200_ctx* c, cudaStream_t s) {
    peerSignalKernel<<<1, 32, 0, s>>>(c->peers, PEER_SLOT_RECIP, c->peerSeq[PEER_SLOT_RECIP]);
    LAUNCH_COUNT(c, 1);
}
void b200_peer_add_recip(b200_ctx* c) {
    if (!c->recipPending) return;
    ScopedTimer tm(c, B200_T_COMM, c->stream);
    const int b = c->shardAtomBounds[c->commRank], e = c->shardAtomBounds[c->commRank + 1];
    // the wait is launched even for an empty slice: every rank consumes the signal before the next evaluation overwrites the buffers
    peerWaitRecip<<<1, 32, 0, c->stream>>>(c->peerFlags, c->peerSeq[PEER_SLOT_RECIP], c->counters);
    if (e > b) peerAddRecip<<<c->smCount * 2, 256, 0, c->stream>>>(c->force, c->peers.recip[0], c->Npad, b, e);
    LAUNCH_COUNT(c, e > b ? 2 : 1);
    c->recipPending = false;
}

// Barrier and reduce-scatter in one kernel, in two phases (opt-in, B200_ORDERED_REDUCE=1): first the contributions of the
// direct-space ranks (1..n-1) are summed, as soon as all of THEM have signalled, with all their loads in flight together; then rank 0's
// slice is added when its signal arrives.  With eight ranks rank 0 only runs PME and is the last to finish, so the bulk of the traffic
// is already summed while everybody waits for it.  Integer sums: the order does not change a bit of the result.  (A peer-by-peer
// version, one dependent NVLink pass per rank, was measured slower than barrier + one-pass reduce on 4 GPUs: 0.636 vs 0.583 ms/step.)
__device__ __forceinline__ void peerWaitFlag(const unsigned long long* flags, int slot, int r, unsigned long long seq, DevCounters* counters) {
    volatile const unsigned long long* f = flags + slot * B200_MAX_PEERS + r;
    unsigned long long t0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (*f < seq) {
        unsigned long long t1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > 20000000000ull) { counters->overflow = 2; break; }   // 20 s: a rank died; fail loudly instead of hanging the GPU
    }
}
__global__ void __launch_bounds__(256) peerReduceForcesOrdered(PeerSet p, const unsigned long long* flags, int slot, unsigned long long seq, int Npad, int aBegin, int aEnd, DevCounters* counters) {
    const int count = aEnd - aBegin;
    long long* dst = p.force[p.rank];
    if ((int)threadIdx.x >= 1 && (int)threadIdx.x < p.n && (int)threadIdx.x != p.rank) peerWaitFlag(flags, slot, threadIdx.x, seq, counters);
    __threadfence_system();
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 3 * count; i += gridDim.x * blockDim.x) {
        const int comp = i / count, j = i - comp * count;
        const size_t off = (size_t)comp * Npad + aBegin + j;
        long long s = dst[off];
        for (int r = 1; r < p.n; r++) if (r != p.rank) s += p.force[r][off];
        dst[off] = s;
    }
    if (p.rank == 0) return;
    if (threadIdx.x == 0) peerWaitFlag(flags, slot, 0, seq, counters);
    __threadfence_system();
    __syncthreads();
    const long long* src = p.force[0];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 3 * count; i += gridDim.x * blockDim.x) {
        const int comp = i / count, j = i - comp * count;
        const size_t off = (size_t)comp * Npad + aBegin + j;
        dst[off] += src[off];
    }
}
// signal "my partial forces are complete", then reduce peer by peer as their signals arrive (replaces barrier + reduce)
void b200_peer_signal_and_reduce_forces(b200_ctx* c) {
    ScopedTimer tm(c, B200_T_COMM, c->stream);
    const unsigned long long seq = ++c->peerSeq[PEER_SLOT_FORCES];
    peerSignalKernel<<<1, 32, 0, c->stream>>>(c->peers, PEER_SLOT_FORCES, seq);
    peerReduceForcesOrdered<<<c->smCount * 2, 256, 0, c->stream>>>(c->peers, c->peerFlags, PEER_SLOT_FORCES, seq, c->Npad, c->shardAtomBegin, c->shardAtomEnd, c->counters);
    LAUNCH_COUNT(c, 2);
}
// the all-gather: one kernel stores this rank's range of posq into every other rank's array (coalesced 16-byte stores over
// NVLink to all peers at once; three back-to-back DMA copies were measured at 38 us for 3 x 4.3 MB, launch-latency bound)
__global__ void __launch_bounds__(256) peerPublishPositions(PeerSet p, int aBegin, int aEnd) {
    for (int i = aBegin + blockIdx.x * blockDim.x + threadIdx.x; i < aEnd; i += gridDim.x * blockDim.x) {
        const float4 v = p.posq[p.rank][i];
        for (int k = 1; k < p.n; k++) { int r = p.rank + k; r -= r >= p.n ? p.n : 0; p.posq[r][i] = v; }
    }
}
void b200_peer_publish_positions(b200_ctx* c) {
    ScopedTimer tm(c, B200_T_COMM, c->stream);
    if (c->shardAtomEnd > c->shardAtomBegin) { peerPublishPositions<<<c->smCount * 2, 256, 0, c->stream>>>(c->peers, c->shardAtomBegin, c->shardAtomEnd); LAUNCH_COUNT(c, 1); }
}

// rank 0's freshly set positions (float part + remainder) -> every other rank (b200_set_positions_root)
__global__ void __launch_bounds__(256) peerBroadcastPositions(PeerSet p, int N) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const float4 v = p.posq[p.rank][i], q = p.posqCorr[p.rank][i];
        for (int r = 0; r < p.n; r++) if (r != p.rank) { p.posq[r][i] = v; p.posqCorr[r][i] = q; }
    }
}
void b200_peer_broadcast_positions(b200_ctx* c) {
    ScopedTimer tm(c, B200_T_COMM, c->stream);
    peerBroadcastPositions<<<c->smCount * 2, 256, 0, c->stream>>>(c->peers, c->N);
    LAUNCH_COUNT(c, 1);
}

static void peerSetup(b200_ctx* c) {
    c->peerMode = false;
    const int R = c->commRanks;
    if (R < 2 || R > B200_MAX_PEERS || getenv("B200_NO_PEER") || !g_nccl.AllGather) return;
    int ok = 1;
    IpcBundle mine; memset(&mine, 0, sizeof(mine));
    dmalloc(c->peerFlags, (size_t)PEER_NUM_SLOTS * B200_MAX_PEERS);
    CUDA_CHECK(cudaMemset(c->peerFlags, 0, sizeof(unsigned long long) * PEER_NUM_SLOTS * B200_MAX_PEERS));
    dmalloc(c->recipF, (size_t)3 * c->Npad);
    CUDA_CHECK(cudaMemset(c->recipF, 0, sizeof(float) * 3 * (size_t)c->Npad));
    constexpr int NP = 7;
    void* ptrs[NP] = {c->force, c->posq, c->posqCorr, c->velm, c->counters, c->peerFlags, c->recipF};
    for (int k = 0; k < NP; k++) { mine.raw[k] = ptrs[k]; if (cudaIpcGetMemHandle(&mine.h[k], ptrs[k]) != cudaSuccess) { cudaGetLastError(); ok = 0; } }
    mine.pid = (long long)getpid(); mine.device = c->device;
    // exchange the handles (and whether everybody could export) through the communicator
    char* dSend = nullptr; char* dRecv = nullptr;
    const size_t rec = sizeof(IpcBundle) + 8;
    dmalloc(dSend, rec); dmalloc(dRecv, rec * R);
    std::vector<char> hostSend(rec, 0), hostRecv(rec * R, 0);
    memcpy(hostSend.data(), &mine, sizeof(mine)); hostSend[sizeof(mine)] = (char)ok;
    CUDA_CHECK(cudaMemcpy(dSend, hostSend.data(), rec, cudaMemcpyHostToDevice));
    NCCL_CHECK(g_nccl.AllGather(dSend, dRecv, rec, 0 /* ncclChar */, (ncclComm_t)c->ncclComm, c->stream));
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    CUDA_CHECK(cudaMemcpy(hostRecv.data(), dRecv, rec * R, cudaMemcpyDeviceToHost));
    dfree(dSend); dfree(dRecv);
    for (int r = 0; r < R; r++) if (!hostRecv[r * rec + sizeof(IpcBundle)]) ok = 0;
    PeerSet& p = c->peers; memset(&p, 0, sizeof(p));
    p.n = R; p.rank = c->commRank;
    if (ok) for (int r = 0; r < R; r++) {
        void* q[NP];
        if (r == c->commRank) { for (int k = 0; k < NP; k++) q[k] = ptrs[k]; }
        else {
            IpcBundle b; memcpy(&b, hostRecv.data() + r * rec, sizeof(b));
            if (b.pid == mine.pid) {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, c->device, b.device) != cudaSuccess || !can) { cudaGetLastError(); ok = 0; }
                else { const cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0); if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = 0; cudaGetLastError(); }
                for (int k = 0; k < NP; k++) q[k] = ok ? b.raw[k] : nullptr;
                c->peerInProcess = true;
            }
            else for (int k = 0; k < NP; k++) if (cudaIpcOpenMemHandle(&q[k], b.h[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; q[k] = nullptr; }
        }
        p.force[r] = (long long*)q[0]; p.posq[r] = (float4*)q[1]; p.posqCorr[r] = (float4*)q[2]; p.velm[r] = (double4*)q[3];
        p.counters[r] = (DevCounters*)q[4]; p.flags[r] = (unsigned long long*)q[5]; p.recip[r] = (float*)q[6];
    }
    // all ranks must agree: one more exchange of the outcome
    {
        long long* dOk = nullptr; dmalloc(dOk, 1);
        long long v = ok ? 0 : 1;
        CUDA_CHECK(cudaMemcpy(dOk, &v, sizeof(v), cudaMemcpyHostToDevice));
        NCCL_CHECK(g_nccl.AllReduce(dOk, dOk, 1, ncclInt64, ncclSum, (ncclComm_t)c->ncclComm, c->stream));
        CUDA_CHECK(cudaStreamSynchronize(c->stream));
        CUDA_CHECK(cudaMemcpy(&v, dOk, sizeof(v), cudaMemcpyDeviceToHost));
        dfree(dOk);
        ok = v == 0;
    }
    c->peerMode = ok != 0;
    if (!c->peerMode && c->commRank == 0) fprintf(stderr, "b200: CUDA IPC peer mapping unavailable, using ncclAllReduce + replicated integration\n");
}

extern "C" int b200_comm_init(b200_ctx* c, int32_t rank, int32_t nranks, const char* id128) {
    b200_graph_invalidate(c);
    try {
        loadNccl();
        CUDA_CHECK(cudaSetDevice(c->device));
        ncclUniqueId id; memcpy(id.internal, id128, 128);
        ncclComm_t comm;
        NCCL_CHECK(g_nccl.CommInitRank(&comm, nranks, id, rank));
        c->ncclComm = comm; c->commRank = rank; c->commRanks = nranks;
        const int rc = b200_set_partition(c, rank, nranks);
        if (rc != B200_OK) return rc;
        peerSetup(c);
        return B200_OK;
    } catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
}

void b200_comm_allreduce_forces(b200_ctx* c) {
    ScopedTimer tm(c, B200_T_COMM, c->stream);
    NCCL_CHECK(g_nccl.AllReduce(c->force, c->force, 3 * (size_t)c->Npad, ncclInt64, ncclSum, (ncclComm_t)c->ncclComm, c->stream));
}

// energies are fixed-point integers too: one more exact integer all-reduce when an energy is requested
void b200_comm_allreduce_energy(b200_ctx* c) {
    NCCL_CHECK(g_nccl.AllReduce(c->counters->energy, c->counters->energy, 8, ncclInt64, ncclSum, (ncclComm_t)c->ncclComm, c->stream));
}

void b200_comm_destroy(b200_ctx* c) {
    if (c->peerMode) {
        cudaDeviceSynchronize();
        PeerSet& p = c->peers;
        for (int r = 0; r < p.n && !c->peerInProcess; r++) if (r != p.rank) {
            void* q[7] = {p.force[r], p.posq[r], p.posqCorr[r], p.velm[r], p.counters[r], p.flags[r], p.recip[r]};
            for (void* x : q) if (x) cudaIpcCloseMemHandle(x);
        }
        c->peerMode = false;
    }
    dfree(c->peerFlags); dfree(c->recipF);
    if (c->ncclComm && g_nccl.CommDestroy) g_nccl.CommDestroy((ncclComm_t)c->ncclComm);
    c->ncclComm = nullptr;
}
