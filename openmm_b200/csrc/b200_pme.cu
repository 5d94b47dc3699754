// b200_pme.cu -- PME reciprocal space: B-spline charge spreading, reciprocal convolution (+energy), force interpolation.
//
// Replaces gridSpreadCharge / finishSpreadCharge / reciprocalConvolution / gridEvaluateEnergy / gridInterpolateForce
// (platforms/common/src/kernels/pme.cc:23-356) and the set-up in CommonCalcNonbondedForceKernel::commonInitialize
// (platforms/common/src/CommonCalcNonbondedForce.cpp:482-558, B-spline moduli).  The FFTs are in b200_fft.cu.
// Differences: atoms are kept in Hilbert order, so no per-step sort by grid cell is needed (findAtomGridIndex + CudaSort,
// pme.cc:1-21, are dropped) and 128 consecutive atoms cover a compact box of grid points.  Spreading accumulates that box in
// SHARED memory with 32-bit fixed-point atomics (integer sums: bit-reproducible like USE_FIXED_POINT_CHARGE_SPREADING,
// pme.cc:106-107; scale 2^26/max(1,|q|max/1.5): 1.5e-8 e resolution, head-room |sum| < 32 per point) and adds it to the
// global integer grid once; the int -> float conversion is folded into the first FFT pass.  Interpolation stages the same
// box of potential values in shared memory.  The Coulomb constant is applied once, in the convolution.
#include "b200_internal.cuh"
#include <algorithm>
#include <cmath>
#include <vector>

#define PME_ORDER 5


// order-5 cardinal B-spline weights w[k] = M5(u+4-k) and derivatives, by raising the order from 1
// (same functions as pme.cc:67-81,297-314 / ReferencePME.cpp:166-201)
__device__ __forceinline__ void bspline5(float u, float* w, float* dw) {
    float a[PME_ORDER];
    a[0] = 1.f - u; a[1] = u; a[2] = a[3] = a[4] = 0.f;
#pragma unroll
    for (int m = 3; m <= PME_ORDER; m++) {
        if (m == PME_ORDER) {
            dw[0] = -a[0];
#pragma unroll
            for (int k = 1; k < PME_ORDER; k++) dw[k] = a[k - 1] - a[k];
        }
        const float inv = 1.f / (m - 1);
        float prev = 0.f;
#pragma unroll
        for (int k = 0; k < m; k++) {
            const float cur = a[k];
            a[k] = inv * ((u + (m - 1 - k)) * prev + (1.f - u + k) * cur);
            prev = cur;
        }
    }
#pragma unroll
    for (int k = 0; k < PME_ORDER; k++) w[k] = a[k];
}

// LJPME: the quantity spread on the dispersion mesh is c6_i = 8 (sigma_i/2)^3 * 2 sqrt(eps_i)  (ReferenceLJCoulombIxn.cpp:259-260)
__device__ __forceinline__ float dispersionCoefficient(float2 se) { return 8.0f * se.x * se.x * se.x * se.y; }

__device__ __forceinline__ void gridCoords(const float4 p, const PmeArgs& a, int& ix, int& iy, int& iz, float& ux, float& uy, float& uz) {
    // fractional coordinates in double: the only error left is the float rounding of the stored position
    const double x = p.x, y = p.y, z = p.z;
    double fa = x * a.r0x + y * a.r1x + z * a.r2x, fb = y * a.r1y + z * a.r2y, fc = z * a.r2z;
    fa = (fa - floor(fa)) * a.nx; fb = (fb - floor(fb)) * a.ny; fc = (fc - floor(fc)) * a.nz;
    ix = (int)fa; iy = (int)fb; iz = (int)fc;
    ux = (float)(fa - ix); uy = (float)(fb - iy); uz = (float)(fc - iz);
    ix = ix >= a.nx ? ix - a.nx : ix; iy = iy >= a.ny ? iy - a.ny : iy; iz = iz >= a.nz ? iz - a.nz : iz;
}

// ---- chunked spreading / interpolation.  One CTA takes 128 consecutive atoms.  Because atoms are kept in Hilbert order
// their stencils cover a compact box of grid points (about 17^3 for 128 atoms at 0.11 nm spacing); when that box fits the
// shared-memory tile the CTA accumulates all 125 contributions per atom there (native 32-bit shared atomics) and then adds
// the box to the global grid row by row (one warp per (x,y) row, lanes along z: coalesced integer atomics, no index
// arithmetic per point).  Boxes that do not fit (rare: atoms that drifted far apart between two re-sorts) fall back to
// direct global atomics.  All sums are integer, so the grid is bit-reproducible.
#define CHUNK 128
#define CHUNK_WARPS (CHUNK / 32)
#define TILE_MAX 8192
struct ChunkBox { int org[3]; int lo[3]; int ext[3]; int vol; };   // org = wrapped grid index of tile entry (0,0,0)

// d = difference of two indices in [0,n) -> representative in [-n/2, n/2]; i in (-n, 2n) -> [0,n)   (no integer division)
__device__ __forceinline__ int wrapDiff(int d, int n) { const int h = n >> 1; d -= d > h ? n : 0; d += d < -h ? n : 0; return d; }
__device__ __forceinline__ int wrapIdx(int i, int n) { i += i < 0 ? n : 0; i -= i >= n ? n : 0; return i; }

// bounding box (in grid points, relative to the first atom's base index) of the stencils of the chunk's atoms
__device__ __forceinline__ void chunkBox(bool valid, int ix, int iy, int iz, const PmeArgs& a, ChunkBox& box, int* sh /*[10]*/, int& dx, int& dy, int& dz) {
    if (threadIdx.x == 0) { sh[0] = ix; sh[1] = iy; sh[2] = iz; sh[3] = sh[4] = sh[5] = 1 << 30; sh[6] = sh[7] = sh[8] = -(1 << 30); }
    __syncthreads();
    const int r0 = sh[0], r1 = sh[1], r2 = sh[2];
    dx = wrapDiff(ix - r0, a.nx); dy = wrapDiff(iy - r1, a.ny); dz = wrapDiff(iz - r2, a.nz);
    {   // warp-level min/max first: six shared atomics per warp instead of six per thread on the same addresses
        const int big = 1 << 30;
        const int mnx = __reduce_min_sync(0xffffffffu, valid ? dx : big), mny = __reduce_min_sync(0xffffffffu, valid ? dy : big), mnz = __reduce_min_sync(0xffffffffu, valid ? dz : big);
        const int mxx = __reduce_max_sync(0xffffffffu, valid ? dx : -big), mxy = __reduce_max_sync(0xffffffffu, valid ? dy : -big), mxz = __reduce_max_sync(0xffffffffu, valid ? dz : -big);
        if ((threadIdx.x & 31) == 0 && mxx >= mnx) {
            atomicMin(&sh[3], mnx); atomicMin(&sh[4], mny); atomicMin(&sh[5], mnz);
            atomicMax(&sh[6], mxx); atomicMax(&sh[7], mxy); atomicMax(&sh[8], mxz);
        }
    }
    __syncthreads();
    for (int k = 0; k < 3; k++) { box.lo[k] = sh[3 + k]; box.ext[k] = sh[6 + k] - sh[3 + k] + PME_ORDER; }
    const int n[3] = {a.nx, a.ny, a.nz};
    bool fits = sh[6] >= sh[3];
    for (int k = 0; k < 3; k++) fits = fits && box.ext[k] <= n[k];      // a box wider than the grid would alias onto itself
    const long long vol = (long long)box.ext[0] * box.ext[1] * box.ext[2];
    box.vol = (fits && vol <= TILE_MAX) ? (int)vol : 0;
    box.org[0] = wrapIdx(r0 + box.lo[0], a.nx); box.org[1] = wrapIdx(r1 + box.lo[1], a.ny); box.org[2] = wrapIdx(r2 + box.lo[2], a.nz);
}

// The tile is visited in units of half an x plane: unit u = (x = u>>1, lower or upper half of the y range); warp w takes
// units w, w+CHUNK_WARPS, ...  Inside a unit the rows (x, y), (x, y+1), ... are one grid row apart, so the global row offset
// is advanced by an add and one wrap test per row (no index arithmetic per point), lanes run along z, and rows are handled
// ROW_BATCH at a time so that a warp keeps that many independent memory operations in flight.
#define ROW_BATCH 3
struct PlaneUnits {
    int lane, warp, ey, ez, numUnits, yHalf, planeStride, gz;
    const ChunkBox& box; const PmeArgs& a;
    __device__ __forceinline__ PlaneUnits(const ChunkBox& b, const PmeArgs& args) : box(b), a(args) {
        lane = threadIdx.x & 31; warp = threadIdx.x >> 5;
        ey = b.ext[1]; ez = b.ext[2]; numUnits = 2 * b.ext[0]; yHalf = (ey + 1) >> 1; planeStride = args.ny * args.nz;
    }
    // f(tileIndex, globalIndex, rowValid) is called ROW_BATCH rows at a time, twice (phase 0: issue loads, phase 1: consume)
    template <typename F0, typename F1>
    __device__ __forceinline__ void run(F0 issue, F1 consume) {
        for (int z0 = 0; z0 < ez; z0 += 32) {
            const int z = z0 + lane;
            const bool zOk = z < ez;
            int gzl = box.org[2] + z; gzl -= gzl >= a.nz ? a.nz : 0;
            for (int u = warp; u < numUnits; u += CHUNK_WARPS) {
                const int x = u >> 1;
                const int yBegin = (u & 1) ? yHalf : 0, yEnd = (u & 1) ? ey : yHalf;
                int gx = box.org[0] + x; gx -= gx >= a.nx ? a.nx : 0;
                int gy = box.org[1] + yBegin; gy -= gy >= a.ny ? a.ny : 0;
                int g = (gx * a.ny + gy) * a.nz + gzl;            // global index of (x, yBegin, z)
                int t = (x * ey + yBegin) * ez + z;               // tile index
                for (int y = yBegin; y < yEnd; y += ROW_BATCH) {
                    int gi[ROW_BATCH];
#pragma unroll
                    for (int k = 0; k < ROW_BATCH; k++) {
                        gi[k] = g;
                        issue(k, t + k * ez, g, zOk && (y + k) < yEnd);
                        g += a.nz; gy++;
                        if (gy == a.ny) { gy = 0; g -= planeStride; }
                    }
#pragma unroll
                    for (int k = 0; k < ROW_BATCH; k++) consume(k, t + k * ez, gi[k], zOk && (y + k) < yEnd);
                    t += ROW_BATCH * ez;
                }
            }
        }
    }
};

// Fast path for the usual case of an x plane of the tile holding at most PLANE_MAX points: the CTA computes, once, the global
// offset (inside an x plane of the grid) of every entry of a tile plane into a small shared table.  A plane is then moved by a flat
// loop -- one table look-up, one shared-memory access and one global access per point, no index arithmetic or predicates per point
// (the round-1 version kept 16 offsets per lane in registers; ncu showed half of the kernel's instructions were the predicate and
// integer bookkeeping around them: ISETP/IADD3/LOP3/LEA/P2R 47 % of the interpolation kernel, 6 % useful FFMA2).
#define PLANE_MAX 640
__device__ __forceinline__ bool planeOffsets(const ChunkBox& box, const PmeArgs& a, int* goff /*[PLANE_MAX] shared*/) {
    const int ey = box.ext[1], ez = box.ext[2], planeSize = ey * ez;
    if (planeSize > PLANE_MAX) return false;
    const float invEz = 1.0f / (float)ez;
    for (int e = threadIdx.x; e < planeSize; e += CHUNK) {
        const int y = (int)(((float)e + 0.5f) * invEz), z = e - y * ez;     // exact for e < 2^20
        int gy = box.org[1] + y; gy -= gy >= a.ny ? a.ny : 0;
        int gz = box.org[2] + z; gz -= gz >= a.nz ? a.nz : 0;
        goff[e] = gy * a.nz + gz;
    }
    return true;
}

// Bank-aware assignment of the chunk's atoms to threads.  Every thread walks the same 125 stencil offsets, so the shared-memory
// bank a lane hits at each step is (its atom's base index + a common offset) mod 32: two lanes collide on ALL 125 accesses when
// their bases agree mod 32, and with random bases a warp sees about 3.9 wavefronts per access (measured).  The atoms are
// therefore counting-sorted by (base mod 32) and dealt to the four warps in turn, so that the bases of a warp's lanes are spread
// over the banks (a conflict remains only where more than four atoms of the chunk share a bank).  Returns the thread whose atom
// this thread processes.  All sums are integers (spread) or per-atom (interpolation), so the assignment does not affect results.
__device__ __forceinline__ int bankSortedSource(int base, int* binCount /*[32]*/, int* perm /*[CHUNK]*/) {
    const int tid = threadIdx.x, bank = base & 31;
    if (tid < 32) binCount[tid] = 0;
    __syncthreads();
    const int rank = atomicAdd(&binCount[bank], 1);
    __syncthreads();
    int start = 0;
    if (tid < 32) {
        const int v = binCount[tid];
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (tid >= o) incl += t; }
        start = incl - v;
    }
    __syncthreads();
    if (tid < 32) binCount[tid] = start;
    __syncthreads();
    perm[binCount[bank] + rank] = tid;
    __syncthreads();
    return perm[(tid & 31) * CHUNK_WARPS + (tid >> 5)];
}

// float -> nearest integer: F2I keeps the full 2^26 fixed-point resolution (a mantissa trick would cap it at 2^22 per
// contribution, which was measured to triple the error of the reciprocal forces)
__device__ __forceinline__ int spreadRound(float vxy, float w) { return __float2int_rn(vxy * w); }

// charge spreading (gridSpreadCharge, pme.cc:23-121)
__global__ void __launch_bounds__(CHUNK) pmeSpreadChunks(const float4* __restrict__ posq, const float2* __restrict__ sigEps, int N, PmeArgs a, int* __restrict__ grid, float fixScale) {
    __shared__ __align__(16) int tile[TILE_MAX];
    __shared__ int sh[10];
    __shared__ int binCount[32];
    __shared__ int perm[CHUNK];
    __shared__ int goff[PLANE_MAX];
    int i = blockIdx.x * CHUNK + threadIdx.x;
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < N) { p = posq[i]; if (sigEps) p.w = dispersionCoefficient(sigEps[i]); }
    int ix = 0, iy = 0, iz = 0; float ux = 0.f, uy = 0.f, uz = 0.f;
    gridCoords(p, a, ix, iy, iz, ux, uy, uz);
    bool valid = i < N && p.w != 0.f;
    ChunkBox box; int dx, dy, dz;
    chunkBox(valid, ix, iy, iz, a, box, sh, dx, dy, dz);
    if (box.vol > 0) {
        // swap atoms between the threads of the CTA so that the lanes of a warp start in different banks (see bankSortedSource)
        const int ey0 = box.ext[1], ez0 = box.ext[2];
        const int src = bankSortedSource(valid ? ((dx - box.lo[0]) * ey0 + (dy - box.lo[1])) * ez0 + (dz - box.lo[2]) : threadIdx.x, binCount, perm);
        i = blockIdx.x * CHUNK + src;
        p = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < N) { p = posq[i]; if (sigEps) p.w = dispersionCoefficient(sigEps[i]); }
        gridCoords(p, a, ix, iy, iz, ux, uy, uz);
        valid = i < N && p.w != 0.f;
        dx = wrapDiff(ix - sh[0], a.nx); dy = wrapDiff(iy - sh[1], a.ny); dz = wrapDiff(iz - sh[2], a.nz);
    }
    float wx[5], wy[5], wz[5], d[5];
    bspline5(ux, wx, d); bspline5(uy, wy, d); bspline5(uz, wz, d);
    const float q = p.w * fixScale;
    if (box.vol > 0) {
        for (int t = threadIdx.x * 4; t < box.vol; t += CHUNK * 4) *reinterpret_cast<int4*>(&tile[t]) = make_int4(0, 0, 0, 0);
        __syncthreads();
        if (valid) {
            const int ey = box.ext[1], ez = box.ext[2];
            const int base = ((dx - box.lo[0]) * ey + (dy - box.lo[1])) * ez + (dz - box.lo[2]);
#pragma unroll
            for (int jx = 0; jx < 5; jx++) {
                const float vx = q * wx[jx];
#pragma unroll
                for (int jy = 0; jy < 5; jy++) {
                    const float vxy = vx * wy[jy];
                    int* row = &tile[base + (jx * ey + jy) * ez];
#pragma unroll
                    for (int jz = 0; jz < 5; jz++) atomicAdd(&row[jz], spreadRound(vxy, wz[jz]));
                }
            }
        }
        const bool flat = planeOffsets(box, a, goff);
        __syncthreads();
        if (flat) {
            const int lane = threadIdx.x & 31, planeSize = box.ext[1] * box.ext[2], planeStride = a.ny * a.nz;
            for (int x = threadIdx.x >> 5; x < box.ext[0]; x += CHUNK_WARPS) {
                int gx = box.org[0] + x; gx -= gx >= a.nx ? a.nx : 0;
                int* gp = grid + (size_t)gx * planeStride;
                const int* tp = tile + x * planeSize;
#pragma unroll 4
                for (int e = lane; e < planeSize; e += 32) {
                    const int v = tp[e];
                    if (v != 0) atomicAdd(gp + goff[e], v);
                }
            }
        } else {
            int v[ROW_BATCH];
            PlaneUnits(box, a).run([&](int k, int t, int, bool ok) { v[k] = ok ? tile[t] : 0; },
                                   [&](int k, int, int g, bool) { if (v[k] != 0) atomicAdd(&grid[g], v[k]); });
        }
    } else if (valid) {
        int zi[5];
#pragma unroll
        for (int k = 0; k < 5; k++) { int z = iz + k; zi[k] = z >= a.nz ? z - a.nz : z; }
#pragma unroll
        for (int jx = 0; jx < 5; jx++) {
            int x = ix + jx; x = x >= a.nx ? x - a.nx : x;
            const float vx = q * wx[jx];
#pragma unroll
            for (int jy = 0; jy < 5; jy++) {
                int y = iy + jy; y = y >= a.ny ? y - a.ny : y;
                const float vxy = vx * wy[jy];
                int* row = grid + ((size_t)x * a.ny + y) * a.nz;
#pragma unroll
                for (int jz = 0; jz < 5; jz++) atomicAdd(&row[zi[jz]], spreadRound(vxy, wz[jz]));
            }
        }
    }
}

// (the reciprocal convolution and energy sum, pme.cc:137-253, are fused into the x pass of the FFT: fftXConvolve in b200_fft.cu)

// force interpolation (gridInterpolateForce, pme.cc:255-356): same chunking; the box of potential values the chunk's atoms
// read is staged in shared memory once (coalesced rows), then every atom gathers its 125 values from there.  The z sums of
// the value and of the z derivative share each loaded value (one packed FMA); the y sums are kept per x plane.
__device__ __forceinline__ float2 pmeFma2(float v, float2 w, float2 acc) {
    unsigned long long d; const float2 vv = make_float2(v, v);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*reinterpret_cast<const unsigned long long*>(&vv)), "l"(*reinterpret_cast<const unsigned long long*>(&w)), "l"(*reinterpret_cast<const unsigned long long*>(&acc)));
    return *reinterpret_cast<float2*>(&d);
}

__global__ void __launch_bounds__(CHUNK, 6) pmeInterpolateChunks(const float4* __restrict__ posq, const float2* __restrict__ sigEps, int N, int Npad, PmeArgs a, const float* __restrict__ phi, long long* __restrict__ force, PmePush push) {
    __shared__ float tile[TILE_MAX];
    __shared__ int sh[10];
    __shared__ int binCount[32];
    __shared__ int perm[CHUNK];
    __shared__ int goff[PLANE_MAX];
    int i = blockIdx.x * CHUNK + threadIdx.x;
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < N) { p = posq[i]; if (sigEps) p.w = dispersionCoefficient(sigEps[i]); }
    int ix = 0, iy = 0, iz = 0; float ux = 0.f, uy = 0.f, uz = 0.f;
    gridCoords(p, a, ix, iy, iz, ux, uy, uz);
    bool valid = i < N && p.w != 0.f;
    ChunkBox box; int dx, dy, dz;
    chunkBox(valid, ix, iy, iz, a, box, sh, dx, dy, dz);
    if (box.vol > 0) {   // same bank-aware exchange of atoms between threads as in the spread kernel
        const int ey0 = box.ext[1], ez0 = box.ext[2];
        const int src = bankSortedSource(valid ? ((dx - box.lo[0]) * ey0 + (dy - box.lo[1])) * ez0 + (dz - box.lo[2]) : threadIdx.x, binCount, perm);
        i = blockIdx.x * CHUNK + src;
        p = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < N) { p = posq[i]; if (sigEps) p.w = dispersionCoefficient(sigEps[i]); }
        gridCoords(p, a, ix, iy, iz, ux, uy, uz);
        valid = i < N && p.w != 0.f;
        dx = wrapDiff(ix - sh[0], a.nx); dy = wrapDiff(iy - sh[1], a.ny); dz = wrapDiff(iz - sh[2], a.nz);
    }
    float wx[5], wy[5], wz[5], ex[5], ey5[5], ez5[5];
    float gx = 0.f, gy = 0.f, gz = 0.f;
    if (box.vol > 0) {
        const bool flat = planeOffsets(box, a, goff);
        __syncthreads();
        if (flat) {
            const int lane = threadIdx.x & 31, planeSize = box.ext[1] * box.ext[2], planeStride = a.ny * a.nz;
            for (int x = threadIdx.x >> 5; x < box.ext[0]; x += CHUNK_WARPS) {
                int gx = box.org[0] + x; gx -= gx >= a.nx ? a.nx : 0;
                const float* gp = phi + (size_t)gx * planeStride;
                float* tp = tile + x * planeSize;
                for (int e0 = lane; e0 < planeSize; e0 += 256) {      // eight independent loads in flight per lane
                    float v[8];
#pragma unroll
                    for (int k = 0; k < 8; k++) { const int e = e0 + 32 * k; v[k] = e < planeSize ? gp[goff[e]] : 0.f; }
#pragma unroll
                    for (int k = 0; k < 8; k++) { const int e = e0 + 32 * k; if (e < planeSize) tp[e] = v[k]; }
                }
            }
        } else {
            float v[ROW_BATCH];
            PlaneUnits(box, a).run([&](int k, int, int g, bool ok) { v[k] = ok ? phi[g] : 0.f; },
                                   [&](int k, int t, int, bool ok) { if (ok) tile[t] = v[k]; });
        }
        bspline5(ux, wx, ex); bspline5(uy, wy, ey5); bspline5(uz, wz, ez5);
        __syncthreads();
        if (valid) {
            const int ey = box.ext[1], ez = box.ext[2];
            const int base = ((dx - box.lo[0]) * ey + (dy - box.lo[1])) * ez + (dz - box.lo[2]);
            float2 wd[5];
#pragma unroll
            for (int jz = 0; jz < 5; jz++) wd[jz] = make_float2(wz[jz], ez5[jz]);
#pragma unroll
            for (int jx = 0; jx < 5; jx++) {
                float ax = 0.f, ay = 0.f, az = 0.f;   // sum_y wy*s0, sum_y dwy*s0, sum_y wy*s1
#pragma unroll
                for (int jy = 0; jy < 5; jy++) {
                    const float* row = &tile[base + (jx * ey + jy) * ez];
                    float2 s = make_float2(0.f, 0.f);
#pragma unroll
                    for (int jz = 0; jz < 5; jz++) s = pmeFma2(row[jz], wd[jz], s);
                    ax = fmaf(wy[jy], s.x, ax); ay = fmaf(ey5[jy], s.x, ay); az = fmaf(wy[jy], s.y, az);
                }
                gx = fmaf(ex[jx], ax, gx); gy = fmaf(wx[jx], ay, gy); gz = fmaf(wx[jx], az, gz);
            }
        }
    } else if (valid) {
        bspline5(ux, wx, ex); bspline5(uy, wy, ey5); bspline5(uz, wz, ez5);
        int zi[5];
#pragma unroll
        for (int k = 0; k < 5; k++) { int z = iz + k; zi[k] = z >= a.nz ? z - a.nz : z; }
#pragma unroll
        for (int jx = 0; jx < 5; jx++) {
            int x = ix + jx; x = x >= a.nx ? x - a.nx : x;
#pragma unroll
            for (int jy = 0; jy < 5; jy++) {
                int y = iy + jy; y = y >= a.ny ? y - a.ny : y;
                const float* row = phi + ((size_t)x * a.ny + y) * a.nz;
                float s0 = 0.f, s1 = 0.f;
#pragma unroll
                for (int jz = 0; jz < 5; jz++) { const float v = row[zi[jz]]; s0 = fmaf(wz[jz], v, s0); s1 = fmaf(ez5[jz], v, s1); }
                gx = fmaf(ex[jx] * wy[jy], s0, gx);
                gy = fmaf(wx[jx] * ey5[jy], s0, gy);
                gz = fmaf(wx[jx] * wy[jy], s1, gz);
            }
        }
    }
    if (!valid && (push.dst == nullptr || i >= N)) return;
    const float q = p.w;
    const float tx = gx * a.nx, ty = gy * a.ny, tz = gz * a.nz;
    const float fx = -q * (tx * (float)a.r0x);
    const float fy = -q * (tx * (float)a.r1x + ty * (float)a.r1y);
    const float fz = -q * (tx * (float)a.r2x + ty * (float)a.r2y + tz * (float)a.r2z);
    if (push.dst) {
        // multi-GPU: plain float stores into this (the PME) rank's hand-over buffer; the owning ranks fetch their slices.  Atoms
        // without a coefficient store zeros, so every entry is rewritten by every evaluation.
        push.dst[i] = valid ? fx : 0.f; push.dst[i + Npad] = valid ? fy : 0.f; push.dst[i + 2 * Npad] = valid ? fz : 0.f;
        return;
    }
    atomicAddFixed(&force[i], toFixed(fx));
    atomicAddFixed(&force[i + Npad], toFixed(fy));
    atomicAddFixed(&force[i + 2 * Npad], toFixed(fz));
}

// ---------------------------------------------------------------------------------------------
// B-spline moduli |DFT of the order-5 spline|^2 (CommonCalcNonbondedForce.cpp:482-558)
static std::vector<float> splineModuli(int n) {
    // M5 at the integers 1..4 (M5(0)=M5(5)=0): 1/24, 11/24, 11/24, 1/24
    const double M[PME_ORDER + 1] = {0.0, 1.0 / 24, 11.0 / 24, 11.0 / 24, 1.0 / 24, 0.0};
    std::vector<double> mod(n);
    for (int k = 0; k < n; k++) {
        double sc = 0, ss = 0;
        for (int j = 0; j <= PME_ORDER; j++) { const double arg = 2.0 * M_PI * k * j / n; sc += M[j] * cos(arg); ss += M[j] * sin(arg); }
        mod[k] = sc * sc + ss * ss;
    }
    std::vector<float> out(n);
    for (int k = 0; k < n; k++) {
        double v = mod[k];
        if (v < 1e-7) v = 0.5 * (mod[(k + n - 1) % n] + mod[(k + 1) % n]);
        out[k] = (float)v;
    }
    return out;
}

// fixed-point scale 2^bits: |coefficient| * 2^bits <= 1.5 * 2^26, so one contribution (weight product <= 0.215) stays below
// 2^24.6 and about 80 maximal contributions of one sign would be needed to overflow a grid point
static float spreadScaleFor(double maxCoef) {
    int bits = 26; double lim = 1.5;
    while (maxCoef > lim && bits > 8) { bits--; lim *= 2; }
    while (maxCoef > 0 && maxCoef <= 0.5 * lim && bits < 40) { bits++; lim *= 0.5; }
    return (float)std::ldexp(1.0, bits);
}

static void meshCreate(b200_ctx* c, PmeMesh& m, const int* grid, double alpha, bool dispersion, double maxCoef) {
    for (int k = 0; k < 3; k++) m.gridN[k] = grid[k];
    const int nx = m.gridN[0], ny = m.gridN[1], nz = m.gridN[2];
    m.nzc = nz / 2 + 1; m.alpha = alpha; m.dispersion = dispersion;
    const size_t G = (size_t)nx * ny * nz, GC = (size_t)nx * ny * m.nzc;
    dmalloc(m.gridI, G); dmalloc(m.gridReal, G); dmalloc(m.gridC, GC);
    m.spreadScale = spreadScaleFor(maxCoef);
    for (int k = 0; k < 3; k++) {
        const std::vector<float> mod = splineModuli(m.gridN[k]);
        dmalloc(m.moduli[k], mod.size());
        CUDA_CHECK(cudaMemcpy(m.moduli[k], mod.data(), sizeof(float) * mod.size(), cudaMemcpyHostToDevice));
        b200_fft_make_plan(m.gridN[k], m.plan[k]);
        b200_fft_make_twiddle(m.gridN[k], m.twiddle[k], c->stream);
    }
}

static void meshDestroy(PmeMesh& m) {
    dfree(m.gridI); dfree(m.gridReal); dfree(m.gridC);
    for (int k = 0; k < 3; k++) { dfree(m.moduli[k]); dfree(m.twiddle[k]); }
}

static void maxCoefficients(const b200_ctx* c, double& maxQ, double& maxC6) {
    maxQ = maxC6 = 0;
    for (double q : c->charge) maxQ = std::max(maxQ, std::fabs(q));
    for (int i = 0; i < c->N; i++) maxC6 = std::max(maxC6, 8.0 * pow(0.5 * c->sigma[i], 3.0) * 2.0 * sqrt(c->epsilon[i]));
}

void b200_pme_create(b200_ctx* c) {
    double maxQ, maxC6; maxCoefficients(c, maxQ, maxC6);
    meshCreate(c, c->mesh[0], c->nb.grid, c->nb.ewald_alpha, false, maxQ);
    if (c->ljpme) meshCreate(c, c->mesh[1], c->nb.ljpme_grid, c->nb.ljpme_alpha, true, maxC6);
}

// per-particle parameters changed (updateParametersInContext): the fixed-point scales follow the new maxima
void b200_pme_rescale(b200_ctx* c) {
    double maxQ, maxC6; maxCoefficients(c, maxQ, maxC6);
    c->mesh[0].spreadScale = spreadScaleFor(maxQ);
    if (c->ljpme) c->mesh[1].spreadScale = spreadScaleFor(maxC6);
}

void b200_pme_destroy(b200_ctx* c) { meshDestroy(c->mesh[0]); meshDestroy(c->mesh[1]); }

static void meshLaunch(b200_ctx* c, PmeMesh& m, cudaStream_t s, bool forces, bool energy) {
    PmeArgs a;
    a.nx = m.gridN[0]; a.ny = m.gridN[1]; a.nz = m.gridN[2]; a.nzc = m.nzc;
    const BoxD& b = c->boxD;
    const double det = b.ax * b.by * b.cz, sc = 1.0 / det;
    a.r0x = b.by * b.cz * sc;
    a.r1x = -b.bx * b.cz * sc; a.r1y = b.ax * b.cz * sc;
    a.r2x = (b.bx * b.cy - b.by * b.cx) * sc; a.r2y = -b.ax * b.cy * sc; a.r2z = b.ax * b.by * sc;
    a.alphaFactor = (float)(M_PI * M_PI / (m.alpha * m.alpha));
    a.dispersion = m.dispersion ? 1 : 0;
    a.scale = (float)(m.dispersion ? -2.0 * M_PI * sqrt(M_PI) / (6.0 * det) : B200_ONE_4PI_EPS0 / (M_PI * det));
    a.bfac = (float)(M_PI / m.alpha); a.fac1 = (float)(2.0 * M_PI * M_PI * M_PI * sqrt(M_PI));
    a.fac2 = (float)(m.alpha * m.alpha * m.alpha); a.fac3 = (float)(-2.0 * m.alpha * M_PI * M_PI);
    const size_t G = (size_t)a.nx * a.ny * a.nz;
    const int chunks = (c->N + CHUNK - 1) / CHUNK;
    const float2* coef = m.dispersion ? c->sigEps : nullptr;
    {
        ScopedTimer tm(c, B200_T_SPREAD, s);
        CUDA_CHECK(cudaMemsetAsync(m.gridI, 0, sizeof(int) * G, s));
        pmeSpreadChunks<<<chunks, CHUNK, 0, s>>>(c->posq, coef, c->N, a, m.gridI, m.spreadScale);
        LAUNCH_COUNT(c, 1);
    }
    { ScopedTimer tm(c, B200_T_FFT, s); b200_fft3d_reciprocal(c, m, s, a, energy, forces); }
    if (forces) {
        ScopedTimer tm(c, B200_T_INTERPOLATE, s);
        PmePush push; push.dst = (c->recipPushNow && &m == &c->mesh[0]) ? c->recipF : nullptr;
        pmeInterpolateChunks<<<chunks, CHUNK, 0, s>>>(c->posq, coef, c->N, c->Npad, a, m.gridReal, c->force, push);
        LAUNCH_COUNT(c, 1);
        if (push.dst) b200_peer_signal_recip(c, s);
    }
    CUDA_CHECK(cudaGetLastError());
}

void b200_pme_launch(b200_ctx* c, cudaStream_t s, bool forces, bool energy) {
    meshLaunch(c, c->mesh[0], s, forces, energy);
    if (c->ljpme) meshLaunch(c, c->mesh[1], s, forces, energy);
}
