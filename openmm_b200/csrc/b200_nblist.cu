// b200_nblist.cu -- neighbour list: block bounding boxes, displacement (Verlet buffer) check, tile-list build.
//
// Replaces findBlockBounds / sortBoxData / findBlocksWithInteractions
// (platforms/cuda/src/kernels/findInteractingBlocks.cu:41-542) and the host logic of
// CudaNonbondedUtilities::prepareInteractions (CudaNonbondedUtilities.cpp:385-412).
// Atoms are already ordered along a Hilbert curve (b200_reorder.cu), so 32 consecutive atoms form a compact
// block and 32 consecutive blocks a compact "super block"; the build is a two-level bounding-box sweep
// followed by an exact per-atom test against the 32 atoms of block X and a ballot/popc compaction of the
// surviving j atoms into work units of up to 8 tiles.  Everything is decided on the device: the kernels
// are launched every step; nbBegin takes the rebuild request (counters->rebuild -> rebuildNow) and the
// others return immediately unless it found one.
#include "b200_internal.cuh"
#include <algorithm>

#define NB_WARPS 4
#define NB_THREADS (NB_WARPS * 32)
#define NB_BUF (B200_UNIT_TILES * 32 + 32)

// ---- displacement check (findInteractingBlocks.cu:176-185: rebuild when an atom moved more than padding/2)
__global__ void nbCheckDisplacement(const float4* __restrict__ posq, const float4* __restrict__ refPosq, int N, float maxDisp2, DevCounters* counters) {
    if (counters->rebuild) return;
    bool moved = false;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const float4 p = posq[i], r = refPosq[i];
        const float dx = p.x - r.x, dy = p.y - r.y, dz = p.z - r.z;
        moved |= (dx * dx + dy * dy + dz * dz > maxDisp2);
    }
    if (__any_sync(0xffffffffu, moved) && (threadIdx.x & 31) == 0) counters->rebuild = 1;
}

// ---- per-block bounding boxes (findBlockBounds, findInteractingBlocks.cu:41-102), one warp per block
__global__ void nbBlockBounds(const float4* __restrict__ posq, float4* __restrict__ refPosq, int N, int B, BoxF box,
                              float4* __restrict__ blockCenter, float4* __restrict__ blockHalf, DevCounters* counters) {
    if (!counters->rebuildNow) return;
    const int lane = threadIdx.x & 31;
    const int warpsPerGrid = (gridDim.x * blockDim.x) >> 5;
    if (blockIdx.x == 0 && threadIdx.x == 0) { counters->numUnits = 0; counters->numTiles = 0; counters->numSparseTiles = 0; counters->rebuildCount++; }
    for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < B; b += warpsPerGrid) {
        const int slot = b * 32 + lane;
        const float4 p = posq[slot];
        refPosq[slot] = p;
        const float x0 = __shfl_sync(0xffffffffu, p.x, 0), y0 = __shfl_sync(0xffffffffu, p.y, 0), z0 = __shfl_sync(0xffffffffu, p.z, 0);
        float dx = p.x - x0, dy = p.y - y0, dz = p.z - z0;
        imageDelta(dx, dy, dz, box);
        if (slot >= N) { dx = dy = dz = 0.f; }
        float lx = dx, ly = dy, lz = dz, hx = dx, hy = dy, hz = dz;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lx = fminf(lx, __shfl_xor_sync(0xffffffffu, lx, o)); hx = fmaxf(hx, __shfl_xor_sync(0xffffffffu, hx, o));
            ly = fminf(ly, __shfl_xor_sync(0xffffffffu, ly, o)); hy = fmaxf(hy, __shfl_xor_sync(0xffffffffu, hy, o));
            lz = fminf(lz, __shfl_xor_sync(0xffffffffu, lz, o)); hz = fmaxf(hz, __shfl_xor_sync(0xffffffffu, hz, o));
        }
        if (lane == 0) {
            blockCenter[b] = make_float4(x0 + 0.5f * (lx + hx), y0 + 0.5f * (ly + hy), z0 + 0.5f * (lz + hz), 0.f);
            // half extents are rounded up a little so that float rounding can never exclude an atom from its own box
            blockHalf[b] = make_float4(0.5f * (hx - lx) * 1.000001f + 1e-6f, 0.5f * (hy - ly) * 1.000001f + 1e-6f, 0.5f * (hz - lz) * 1.000001f + 1e-6f, 0.f);
        }
    }
}

// ---- bounding boxes of groups of 32 consecutive blocks ("large blocks", findInteractingBlocks.cu:138-175)
__global__ void nbSuperBounds(const float4* __restrict__ blockCenter, const float4* __restrict__ blockHalf, int B, int numSuper, BoxF box,
                              float4* __restrict__ superCenter, float4* __restrict__ superHalf, const DevCounters* counters) {
    if (!counters->rebuildNow) return;
    const int lane = threadIdx.x & 31;
    const int warpsPerGrid = (gridDim.x * blockDim.x) >> 5;
    for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < numSuper; s += warpsPerGrid) {
        const int b = min(s * 32 + lane, B - 1);
        const float4 c = blockCenter[b], h = blockHalf[b];
        const float x0 = __shfl_sync(0xffffffffu, c.x, 0), y0 = __shfl_sync(0xffffffffu, c.y, 0), z0 = __shfl_sync(0xffffffffu, c.z, 0);
        float dx = c.x - x0, dy = c.y - y0, dz = c.z - z0;
        imageDelta(dx, dy, dz, box);
        float lx = dx - h.x, ly = dy - h.y, lz = dz - h.z, hx = dx + h.x, hy = dy + h.y, hz = dz + h.z;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lx = fminf(lx, __shfl_xor_sync(0xffffffffu, lx, o)); hx = fmaxf(hx, __shfl_xor_sync(0xffffffffu, hx, o));
            ly = fminf(ly, __shfl_xor_sync(0xffffffffu, ly, o)); hy = fmaxf(hy, __shfl_xor_sync(0xffffffffu, hy, o));
            lz = fminf(lz, __shfl_xor_sync(0xffffffffu, lz, o)); hz = fmaxf(hz, __shfl_xor_sync(0xffffffffu, hz, o));
        }
        if (lane == 0) {
            superCenter[s] = make_float4(x0 + 0.5f * (lx + hx), y0 + 0.5f * (ly + hy), z0 + 0.5f * (lz + hz), 0.f);
            superHalf[s] = make_float4(0.5f * (hx - lx) + 1e-6f, 0.5f * (hy - ly) + 1e-6f, 0.5f * (hz - lz) + 1e-6f, 0.f);
        }
    }
}

__device__ __forceinline__ float boxGap2(float4 cA, float4 hA, float4 cB, float4 hB, const BoxF& box) {
    float dx = cB.x - cA.x, dy = cB.y - cA.y, dz = cB.z - cA.z;
    imageDelta(dx, dy, dz, box);
    dx = fmaxf(0.f, fabsf(dx) - hA.x - hB.x);
    dy = fmaxf(0.f, fabsf(dy) - hA.y - hB.y);
    dz = fmaxf(0.f, fabsf(dz) - hA.z - hB.z);
    return dx * dx + dy * dy + dz * dz;
}

// ---- tile-list build (findBlocksWithInteractions, findInteractingBlocks.cu:289-542), one warp per block X.
// Every candidate j atom gets the 32-bit mask of the atoms of X it can reach within the padded cutoff.  Atoms that reach
// more than `sparseMax` atoms of X are compacted into dense 32-atom tiles; the others go, with their mask, into sparse
// tiles (the reference's singlePairs, findInteractingBlocks.cu:188-238) that cost a few pair evaluations per atom.
#define NB_SPARSE_TILES 4
#define NB_SBUF (NB_SPARSE_TILES * 32 + 32)

__device__ __forceinline__ void flushDense(int X, int* buf, int ntiles, int lane, int4* units, int* jAtoms, unsigned unitCapacity, unsigned tileCapacity, DevCounters* counters) {
    unsigned tileBase = 0, unit = 0;
    if (lane == 0) { tileBase = atomicAdd(&counters->numTiles, (unsigned)ntiles); unit = atomicAdd(&counters->numUnits, 1u); }
    tileBase = __shfl_sync(0xffffffffu, tileBase, 0); unit = __shfl_sync(0xffffffffu, unit, 0);
    if (tileBase + ntiles <= tileCapacity && unit < unitCapacity) {
        for (int k = lane; k < ntiles * 32; k += 32) jAtoms[(size_t)tileBase * 32 + k] = buf[k];
        if (lane == 0) units[unit] = make_int4(X, (int)tileBase, ntiles, 0);
    } else if (lane == 0) { counters->overflow = 1; if (unit < unitCapacity) units[unit] = make_int4(X, 0, 0, 0); }   // dropped: leave an empty descriptor, never a stale one
}
// Sparse tiles are written sorted by the number of atoms of X each j atom reaches (descending): the tile kernel gives one lane
// to each j atom and loops over the set bits of its mask, so a tile costs as many iterations as its LARGEST bit count; with
// equal counts side by side the lanes of a warp finish together.  (Counting sort with ballots; the order is deterministic.)
__device__ __forceinline__ void flushSparse(int X, int* slots, int* masks, int ntiles, int lane, int sparseMax, int4* units, int* jAtoms, unsigned unitCapacity, unsigned tileCapacity, DevCounters* counters) {
    unsigned tileBase = 0, unit = 0;
    if (lane == 0) { tileBase = atomicAdd(&counters->numTiles, (unsigned)(2 * ntiles)); unit = atomicAdd(&counters->numUnits, 1u); atomicAdd(&counters->numSparseTiles, (unsigned)ntiles); }
    tileBase = __shfl_sync(0xffffffffu, tileBase, 0); unit = __shfl_sync(0xffffffffu, unit, 0);
    if (tileBase + 2 * ntiles <= tileCapacity && unit < unitCapacity) {
        const unsigned ltMask = (1u << lane) - 1u;
        int base = 0;
        for (int v = sparseMax; v >= 0; v--) {   // layout per sparse tile: 32 atom slots, then their 32 masks; padding entries (count 0) go last
            for (int t = 0; t < ntiles; t++) {
                const int m = masks[t * 32 + lane];
                const bool hit = __popc(m) == v;
                const unsigned b = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int d = base + __popc(b & ltMask);
                    const size_t o = (size_t)(tileBase + 2 * (d >> 5)) * 32 + (d & 31);
                    jAtoms[o] = slots[t * 32 + lane]; jAtoms[o + 32] = m;
                }
                base += __popc(b);
            }
        }
        if (lane == 0) units[unit] = make_int4(X, (int)tileBase, ntiles, 1);
    } else if (lane == 0) { counters->overflow = 1; if (unit < unitCapacity) units[unit] = make_int4(X, 0, 0, 1); }
}

__device__ __forceinline__ float2 nbFma2(float2 a, float2 b, float2 c) { unsigned long long d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c))); return *reinterpret_cast<float2*>(&d); }
__device__ __forceinline__ float2 nbMul2(float2 a, float2 b) { unsigned long long d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b))); return *reinterpret_cast<float2*>(&d); }
__device__ __forceinline__ float2 nbSub2(float2 a, float2 b) { b.x = -b.x; b.y = -b.y; unsigned long long d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b))); return *reinterpret_cast<float2*>(&d); }

// exact test of up to 32 queued candidate atoms (one per lane) against the 32 atoms of block X, then classification
struct NbWarpBufs { int* dense; int* slots; int* masks; int unitTiles; };   // unitTiles <= B200_UNIT_TILES: tiles per dense work unit
__device__ __forceinline__ void nbProcessCandidates(int X, int lane, unsigned ltMask, bool have, int slot, float rx, float ry, float rz, const float4* sXw, const float4* sXp, bool single,
                                                    const BoxF& box, float rcp2, int sparseMax, NbWarpBufs w, int& nbuf, int& nsp,
                                                    int4* units, int* jAtoms, unsigned unitCapacity, unsigned tileCapacity, DevCounters* counters) {
    unsigned imask = 0;
    if (single) {
        // packed FP32x2, atoms i and i+16 of X per step, in the expanded form |r_i|^2 - 2 r_i.r_j < rc^2 - |r_j|^2 (three packed FMAs
        // per pair of atoms; coordinates are relative to the block centre, so the cancellation error, ~1e-7 nm^2, is far inside the
        // list padding).  sXp[i] = (x_i, x_i+16, y_i, y_i+16), sXp[16+i] = (z_i, z_i+16, |r_i|^2, |r_i+16|^2)
        const float2 jx = make_float2(-2.f * rx, -2.f * rx), jy = make_float2(-2.f * ry, -2.f * ry), jz = make_float2(-2.f * rz, -2.f * rz);
        const float thr = rcp2 - (rx * rx + ry * ry + rz * rz);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const float4 a = sXp[i]; const float4 b = sXp[16 + i];
            const float2 t = nbFma2(jx, make_float2(a.x, a.y), nbFma2(jy, make_float2(a.z, a.w), nbFma2(jz, make_float2(b.x, b.y), make_float2(b.z, b.w))));
            imask |= (t.x < thr ? 1u : 0u) << i;
            imask |= (t.y < thr ? 1u : 0u) << (i + 16);
        }
    } else {
        for (int i = 0; i < 32; i++) {
            const float4 a = sXw[i];
            float dx = rx - a.x, dy = ry - a.y, dz = rz - a.z;
            imageDelta(dx, dy, dz, box);
            imask |= (dx * dx + dy * dy + dz * dz < rcp2 ? 1u : 0u) << i;
        }
    }
    if (!have) imask = 0;
    const int cnt = __popc(imask);
    const bool dense = cnt > sparseMax, sparse = cnt > 0 && !dense;
    const unsigned mD = __ballot_sync(0xffffffffu, dense), mS = __ballot_sync(0xffffffffu, sparse);
    if (dense) w.dense[nbuf + __popc(mD & ltMask)] = slot;
    if (sparse) { const int k = nsp + __popc(mS & ltMask); w.slots[k] = slot; w.masks[k] = (int)imask; }
    nbuf += __popc(mD); nsp += __popc(mS);
    __syncwarp();
    while (nbuf >= w.unitTiles * 32) {   // flush full dense work units (at most two: at most 32 atoms were appended)
        flushDense(X, w.dense, w.unitTiles, lane, units, jAtoms, unitCapacity, tileCapacity, counters);
        const int rest = nbuf - w.unitTiles * 32;
        int carry = 0;
        if (lane < rest) carry = w.dense[w.unitTiles * 32 + lane];
        __syncwarp();
        if (lane < rest) w.dense[lane] = carry;
        nbuf = rest;
        __syncwarp();
    }
    if (nsp >= NB_SPARSE_TILES * 32) {   // flush one full sparse work unit
        flushSparse(X, w.slots, w.masks, NB_SPARSE_TILES, lane, sparseMax, units, jAtoms, unitCapacity, tileCapacity, counters);
        const int rest = nsp - NB_SPARSE_TILES * 32;
        int c0 = 0, c1 = 0;
        if (lane < rest) { c0 = w.slots[NB_SPARSE_TILES * 32 + lane]; c1 = w.masks[NB_SPARSE_TILES * 32 + lane]; }
        __syncwarp();
        if (lane < rest) { w.slots[lane] = c0; w.masks[lane] = c1; }
        nsp = rest;
        __syncwarp();
    }
}

__global__ void __launch_bounds__(NB_THREADS) nbBuild(
        const float4* __restrict__ posq, int N, int B, int numSuper, BoxF box, float rcp2, int sparseMax,
        const float4* __restrict__ blockCenter, const float4* __restrict__ blockHalf,
        const float4* __restrict__ superCenter, const float4* __restrict__ superHalf,
        const int* __restrict__ exclRowStart, const int* __restrict__ exclRowCols,
        int4* __restrict__ units, int* __restrict__ jAtoms, unsigned unitCapacity, unsigned tileCapacity,
        int partRank, b200_ctx::OwnerTable table, int split, int unitTiles, DevCounters* counters) {
    if (!counters->rebuildNow) return;
    __shared__ float4 sX[NB_WARPS][32];
    __shared__ float4 sXp[NB_WARPS][32];    // the same atoms packed in pairs (i, i+16) for the FP32x2 exact test
    __shared__ float4 sQ[NB_WARPS][64];     // queue of candidate atoms that passed the box test: (rel. position, slot)
    __shared__ int sBuf[NB_WARPS][NB_BUF];
    __shared__ int sSlot[NB_WARPS][NB_SBUF];
    __shared__ int sMask[NB_WARPS][NB_SBUF];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int warpsPerGrid = (gridDim.x * blockDim.x) >> 5;
    const unsigned ltMask = (1u << lane) - 1u;
    const float rcp = sqrtf(rcp2);
    NbWarpBufs wb = {sBuf[warp], sSlot[warp], sMask[warp], unitTiles};
    // rows are dealt (weighted) round-robin to ranks: every rank gets a statistically identical sample of the sweep.  The rows
    // this rank owns are enumerated densely (k-th owned row = 64*(k/m) + res[k%m], m owned residues of the period-64 table), so
    // that every warp of the grid has work whatever the share of the rank.
    __shared__ unsigned char sRes[64];
    __shared__ int sNumRes;
    if (warp == 0) {
        const bool o0 = table.owner[lane] == partRank, o1 = table.owner[32 + lane] == partRank;
        const unsigned b0 = __ballot_sync(0xffffffffu, o0), b1 = __ballot_sync(0xffffffffu, o1);
        if (o0) sRes[__popc(b0 & ltMask)] = (unsigned char)lane;
        if (o1) sRes[__popc(b0) + __popc(b1 & ltMask)] = (unsigned char)(32 + lane);
        if (lane == 0) sNumRes = __popc(b0) + __popc(b1);
    }
    __syncthreads();
    const int numRes = sNumRes;
    if (numRes == 0) return;
    // Small systems have fewer rows than the GPU has warp slots (DHFR: 737 rows for 9472 slots) and a row's sweep is serial, so a
    // row is shared by `split` warps: warp `part` takes the candidate super blocks sb with sb % split == part and emits its own work
    // units.  `split` depends on the number of blocks only, so the unit composition is the same on any number of GPUs.
    for (int kk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; ; kk += warpsPerGrid) {
        const int k = kk / split, part = kk - k * split;
        const int period = k / numRes;
        if (period * 64 >= B) break;
        const int X = period * 64 + sRes[k - period * numRes];
        if (X >= B) continue;
        const float4 cX = blockCenter[X], hX = blockHalf[X];
        const bool single = !box.triclinic && 0.5f * box.ax - hX.x >= rcp && 0.5f * box.by - hX.y >= rcp && 0.5f * box.cz - hX.z >= rcp;
        {
            const int slot = X * 32 + lane;
            const float4 p = posq[slot];
            float dx = p.x - cX.x, dy = p.y - cX.y, dz = p.z - cX.z;
            imageDelta(dx, dy, dz, box);
            if (slot >= N) { dx = dy = dz = 1.0e9f; }   // padding atoms never attract neighbours
            __syncwarp();
            sX[warp][lane] = make_float4(dx, dy, dz, 0.f);
            {
                const float ox = __shfl_xor_sync(0xffffffffu, dx, 16), oy = __shfl_xor_sync(0xffffffffu, dy, 16), oz = __shfl_xor_sync(0xffffffffu, dz, 16);
                const float n2 = dx * dx + dy * dy + dz * dz, on2 = ox * ox + oy * oy + oz * oz;
                if (lane < 16) { sXp[warp][lane] = make_float4(dx, ox, dy, oy); sXp[warp][16 + lane] = make_float4(dz, oz, n2, on2); }
            }
            __syncwarp();
        }
        const int exBegin = exclRowStart[X], exEnd = exclRowStart[X + 1];
        int nbuf = 0, nsp = 0, nq = 0;
        for (int s0 = (X >> 5) & ~31; s0 < numSuper; s0 += 32) {   // candidates Y > X live in super blocks >= X/32
            const int s = s0 + lane;
            bool sHit = false;
            if (s < numSuper && s >= (X >> 5)) sHit = boxGap2(cX, hX, superCenter[s], superHalf[s], box) < rcp2;
            unsigned sMaskBits = __ballot_sync(0xffffffffu, sHit);
            while (sMaskBits) {
                const int sb = s0 + __ffs(sMaskBits) - 1; sMaskBits &= sMaskBits - 1;
                if (split > 1 && sb % split != part) continue;
                const int Y = sb * 32 + lane;
                bool bHit = false;
                if (Y < B && Y > X) {
                    bHit = boxGap2(cX, hX, blockCenter[Y], blockHalf[Y], box) < rcp2;
                    for (int e = exBegin; e < exEnd && bHit; e++) if (exclRowCols[e] == Y) bHit = false;  // handled as an exclusion tile
                }
                unsigned bMask = __ballot_sync(0xffffffffu, bHit);
                // the atoms of the next hit block are loaded while the current one is tested (padding slots hold finite coordinates)
                float4 pNext = make_float4(0.f, 0.f, 0.f, 0.f);
                if (bMask) pNext = posq[(sb * 32 + __ffs(bMask) - 1) * 32 + lane];
                while (bMask) {
                    const int Yb = sb * 32 + __ffs(bMask) - 1; bMask &= bMask - 1;
                    const int slot = Yb * 32 + lane;
                    const float4 p = pNext;
                    if (bMask) pNext = posq[(sb * 32 + __ffs(bMask) - 1) * 32 + lane];
                    bool cand = false;
                    float rx = 0.f, ry = 0.f, rz = 0.f;
                    if (slot < N) {
                        rx = p.x - cX.x; ry = p.y - cX.y; rz = p.z - cX.z;
                        imageDelta(rx, ry, rz, box);
                        const float gx = fmaxf(0.f, fabsf(rx) - hX.x), gy = fmaxf(0.f, fabsf(ry) - hX.y), gz = fmaxf(0.f, fabsf(rz) - hX.z);
                        cand = gx * gx + gy * gy + gz * gz < rcp2;
                    }
                    // queue the atoms that can reach X's box; the expensive per-atom test runs on full warps of them
                    const unsigned mC = __ballot_sync(0xffffffffu, cand);
                    if (cand) sQ[warp][nq + __popc(mC & ltMask)] = make_float4(rx, ry, rz, __int_as_float(slot));
                    nq += __popc(mC);
                    __syncwarp();
                    if (nq >= 32) {
                        const float4 q = sQ[warp][lane];
                        nbProcessCandidates(X, lane, ltMask, true, __float_as_int(q.w), q.x, q.y, q.z, sX[warp], sXp[warp], single, box, rcp2, sparseMax, wb, nbuf, nsp,
                                            units, jAtoms, unitCapacity, tileCapacity, counters);
                        const int rest = nq - 32;
                        float4 carry = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (lane < rest) carry = sQ[warp][32 + lane];
                        __syncwarp();
                        if (lane < rest) sQ[warp][lane] = carry;
                        nq = rest;
                        __syncwarp();
                    }
                }
            }
        }
        if (nq > 0) {
            const float4 q = lane < nq ? sQ[warp][lane] : make_float4(0.f, 0.f, 0.f, 0.f);
            nbProcessCandidates(X, lane, ltMask, lane < nq, __float_as_int(q.w), q.x, q.y, q.z, sX[warp], sXp[warp], single, box, rcp2, sparseMax, wb, nbuf, nsp,
                                units, jAtoms, unitCapacity, tileCapacity, counters);
        }
        if (nbuf > 0) {
            const int nt = (nbuf + 31) >> 5;
            for (int k = nbuf + lane; k < nt * 32; k += 32) sBuf[warp][k] = -1;
            __syncwarp();
            flushDense(X, sBuf[warp], nt, lane, units, jAtoms, unitCapacity, tileCapacity, counters);
        }
        if (nsp > 0) {
            const int nt = (nsp + 31) >> 5;
            for (int k = nsp + lane; k < nt * 32; k += 32) { sSlot[warp][k] = -1; sMask[warp][k] = 0; }
            __syncwarp();
            flushSparse(X, sSlot[warp], sMask[warp], nt, lane, sparseMax, units, jAtoms, unitCapacity, tileCapacity, counters);
        }
        __syncwarp();
    }
}

// First kernel of the list phase (every step): takes the rebuild request (set by the integration kernel's displacement check, by
// nbCheckDisplacement or by the host), clears it and rewinds the tile kernel's work queue.  Inside the captured evaluation it also
// sets the condition of the graph's IF node that holds the three rebuild kernels, so a step that keeps its list launches none of them.
__global__ void nbBegin(DevCounters* counters, cudaGraphConditionalHandle handle, int useHandle) {
    const int r = counters->rebuild;
    counters->rebuildNow = r; counters->rebuild = 0; counters->workNext = 0;
    if (useHandle) cudaGraphSetConditional(handle, r ? 1u : 0u);
}
__global__ void nbSetRebuild(DevCounters* counters) { counters->rebuild = 1; }

void b200_nblist_alloc(b200_ctx* c) {
    b200_graph_invalidate(c);     // the list arrays are kernel arguments of the captured evaluation
    dfree(c->units); dfree(c->jAtoms);
    if (c->tileCapacity == 0) {
        c->tileCapacity = (long long)c->B * 28 + 1024;
    }
    c->unitCapacity = std::max(c->unitCapacity, 2 * (long long)c->B + c->tileCapacity / 4 + 64);
    dmalloc(c->units, (size_t)c->unitCapacity);
    dmalloc(c->jAtoms, (size_t)c->tileCapacity * 32);
}

void b200_nblist_force_rebuild(b200_ctx* c) {
    nbSetRebuild<<<1, 1, 0, c->stream>>>(c->counters);
    LAUNCH_COUNT(c, 1);
}

void b200_nblist_prepare(b200_ctx* c) {
    ScopedTimer tm(c, B200_T_NBLIST, c->stream);
    const float maxDisp = 0.5f * (float)c->padding;
    const int grid = c->smCount * 4;
    // NoCutoff / CutoffNonPeriodic: no periodic images -- the builder sees a cell so large that no displacement is ever wrapped
    BoxF nbBox = c->boxF;
    if (c->nb.method == B200_NB_NOCUTOFF || c->nb.method == B200_NB_CUTOFF_NONPERIODIC) {
        nbBox.ax = nbBox.by = nbBox.cz = 1.0e18f; nbBox.bx = nbBox.cx = nbBox.cy = 0.f; nbBox.iax = nbBox.iby = nbBox.icz = 1.0e-18f; nbBox.triclinic = 0;
    }
    if (!c->dispChecked) { nbCheckDisplacement<<<grid, 256, 0, c->stream>>>(c->posq, c->refPosq, c->N, maxDisp * maxDisp, c->counters); LAUNCH_COUNT(c, 1); }
    // The three rebuild kernels.  Launched one by one they return at once unless nbBegin found a rebuild request; while computeImpl
    // records the evaluation as a CUDA graph they go into the body of a conditional IF node instead (recorded on the aux stream with
    // cudaStreamBeginCaptureToGraph), so that a replayed step which keeps its list does not launch them at all (DHFR-size: 3 + 3 + 7 us
    // of empty launches on two steps out of three).  Opt-in with B200_GRAPH_COND=1: measured, the IF node costs what it saves.
    cudaStream_t st = c->stream;
    cudaGraphConditionalHandle handle = 0;
    cudaGraph_t bodyGraph = nullptr; cudaGraphNode_t condNode = nullptr;
    static const bool useCond = getenv("B200_GRAPH_COND") != nullptr;     // opt-in, see DESIGN.md (the IF node costs about what it saves)
    const bool conditional = c->capturing && useCond;
    if (conditional) {
        cudaStreamCaptureStatus status; cudaGraph_t graph = nullptr; const cudaGraphNode_t* deps = nullptr; size_t numDeps = 0;
        CUDA_CHECK(cudaStreamGetCaptureInfo(c->stream, &status, nullptr, &graph, &deps, &numDeps));
        if (status != cudaStreamCaptureStatusActive || !graph) throw B200Error(B200_ERR_CUDA, "b200_nblist_prepare: the compute stream is not being captured");
        CUDA_CHECK(cudaGraphConditionalHandleCreate(&handle, graph, 0, cudaGraphCondAssignDefault));
        nbBegin<<<1, 1, 0, c->stream>>>(c->counters, handle, 1);
        CUDA_CHECK(cudaStreamGetCaptureInfo(c->stream, &status, nullptr, &graph, &deps, &numDeps));
        cudaGraphNodeParams np = {};
        np.type = cudaGraphNodeTypeConditional;
        np.conditional.handle = handle; np.conditional.type = cudaGraphCondTypeIf; np.conditional.size = 1;
        CUDA_CHECK(cudaGraphAddNode(&condNode, graph, deps, numDeps, &np));
        bodyGraph = np.conditional.phGraph_out[0];
        CUDA_CHECK(cudaStreamBeginCaptureToGraph(c->auxStream, bodyGraph, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
        st = c->auxStream;
    } else {
        nbBegin<<<1, 1, 0, c->stream>>>(c->counters, handle, 0);
    }
    nbBlockBounds<<<grid, 256, 0, st>>>(c->posq, c->refPosq, c->N, c->B, nbBox, c->blockCenter, c->blockHalf, c->counters);
    nbSuperBounds<<<(c->numSuper + 7) / 8, 256, 0, st>>>(c->blockCenter, c->blockHalf, c->B, c->numSuper, nbBox, c->superCenter, c->superHalf, c->counters);
    const float rcp = (float)c->paddedCutoff;
    int owned = 0;
    for (int i = 0; i < 64; i++) owned += c->partTable.owner[i] == c->partRank ? 1 : 0;
    int rows = (int)(((long long)c->B * owned + 63) / 64) + 1;
    // rows shared by several warps on small systems (see nbBuild); B200_NB_SPLIT overrides the rule
    int split = c->B < 1536 ? 8 : 1;      // measured: DHFR (737 rows) 128 -> 116 us/step with 8; ApoA1 (2882 rows) and larger are best unsplit
    { static const char* e = getenv("B200_NB_SPLIT"); if (e) split = std::max(1, std::min(16, atoi(e))); }
    rows *= split;
    // ... and smaller work units: with a few thousand units for 3552 resident warps the longest unit sets the tile kernel's time.
    // (Both depend on the block count only, so the unit composition does not depend on the number of GPUs.)
    int unitTiles = c->B < 6144 ? 2 : 4;   // measured (tile kernel, us): DHFR 62 -> 43, ApoA1 119 -> 92 with 2; Cellulose 351 -> 336, STMV 824 -> 814 with 4
    { static const char* e = getenv("B200_UNIT_TILES_RT"); if (e) unitTiles = std::max(1, std::min(B200_UNIT_TILES, atoi(e))); }
    int buildGrid = (rows + NB_WARPS - 1) / NB_WARPS;
    if (buildGrid > c->smCount * 16) buildGrid = c->smCount * 16;
    nbBuild<<<buildGrid, NB_THREADS, 0, st>>>(c->posq, c->N, c->B, c->numSuper, nbBox, rcp * rcp, c->sparseMax,
        c->blockCenter, c->blockHalf, c->superCenter, c->superHalf, c->exclRowStart, c->exclRowCols,
        c->units, c->jAtoms, (unsigned)c->unitCapacity, (unsigned)c->tileCapacity, c->partRank, c->partTable, split, unitTiles, c->counters);
    if (conditional) {
        cudaGraph_t done = nullptr;
        CUDA_CHECK(cudaStreamEndCapture(c->auxStream, &done));
        // what follows on the compute stream (the tile kernel) depends on the IF node
        CUDA_CHECK(cudaStreamUpdateCaptureDependencies(c->stream, &condNode, 1, cudaStreamSetCaptureDependencies));
        LAUNCH_COUNT(c, 1);      // the body's kernels are not counted: most replays skip them
    } else LAUNCH_COUNT(c, 4);
    CUDA_CHECK(cudaGetLastError());
}
