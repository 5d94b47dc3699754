// b200_integrate.cu -- minimal LangevinMiddle integrator with SETTLE, centre-of-mass motion removal, kinetic energy.
//
// Supports whole-system runs of the NonbondedForce path; not an optimisation target (BASELINE.json north_star).
// Mirrors IntegrateLangevinMiddleStepKernel (olla/include/openmm/kernels.h:1392-1420): the three parts of
// platforms/common/src/kernels/langevinMiddle.cc:9-95 with the velocity and position constraint passes between
// them (CommonKernels.cpp:3139-3206) are fused into ONE kernel per step: one thread integrates one rigid 3-site
// molecule (SETTLE, Miyamoto & Kollman 1992; integrationUtilities.cc:328-580) or one unconstrained atom.
// Noise comes from a counter-based Philox4x32-10 generator keyed by (seed, step, original particle index), so a
// trajectory does not depend on the atom order on the device nor on the number of GPUs.
#include "b200_internal.cuh"
#include <algorithm>
#include <cstring>

struct V3 { double x, y, z; };
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ V3 unit(V3 a) { return a * rsqrt(dot(a, a)); }

// ---- Philox4x32-10 (Salmon et al., SC'11)
__device__ __forceinline__ uint4 philox(uint4 ctr, uint2 key) {
#pragma unroll
    for (int i = 0; i < 10; i++) {
        const unsigned hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const unsigned hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u; key.y += 0xBB67AE85u;
    }
    return ctr;
}
__device__ __forceinline__ V3 gaussian3(unsigned long long seed, unsigned long long step, int particle) {
    const uint4 r = philox(make_uint4((unsigned)particle, (unsigned)step, (unsigned)(step >> 32), 0x5eedu), make_uint2((unsigned)seed, (unsigned)(seed >> 32)));
    const double u0 = ((double)r.x + 0.5) * (1.0 / 4294967296.0), u1 = ((double)r.y + 0.5) * (1.0 / 4294967296.0);
    const double u2 = ((double)r.z + 0.5) * (1.0 / 4294967296.0), u3 = ((double)r.w + 0.5) * (1.0 / 4294967296.0);
    const double m0 = sqrt(-2.0 * log(u0)), m1 = sqrt(-2.0 * log(u2));
    double s0, c0, s1, c1;
    sincospi(2.0 * u1, &s0, &c0); sincospi(2.0 * u3, &s1, &c1);
    (void)s1;
    return {m0 * c0, m0 * s0, m1 * c1};
}

struct IntegArgs {
    float4* posq; float4* posqCorr; double4* velm; const long long* force; const int* atomIndex;
    int Npad;
    double dt, vscale, noisescale;
    unsigned long long seed, step;
    // fused extras of the stepping loop: COM-motion removal (removeCM.cc) and the Verlet-buffer displacement check
    // (findInteractingBlocks.cu:176-185) ride along, so a step needs no separate passes over the atoms for them
    const float4* refPosq; float maxDisp2; DevCounters* counters;
    int fuseCM;        // 0: no momentum bookkeeping; 1: accumulate the new momentum; 2: also subtract the current COM velocity first
    int cmCur, cmNext, cmClear; double invTotalMass;
    // peer-memory step (b200_comm.cu): this rank integrates units [unitBegin, unitEnd) only, sums the force contributions of all
    // ranks out of their memories and stores the new positions into every rank's posq
    int usePeers, unitBegin, unitEnd; PeerSet peers;
};
#define CM_SCALE 1048576.0   // 2^20 fixed point for momentum sums

__device__ __forceinline__ V3 loadPosD(const IntegArgs& a, int i) { const float4 p = a.posq[i], q = a.posqCorr[i]; return {(double)p.x + q.x, (double)p.y + q.y, (double)p.z + q.z}; }
template <bool PEER = false>
__device__ __forceinline__ void storePosD(const IntegArgs& a, int i, V3 p) {
    const float w = a.posq[i].w;
    const float4 np = make_float4((float)p.x, (float)p.y, (float)p.z, w);
    a.posq[i] = np;
    if (PEER && (a.usePeers & 1)) for (int r = 0; r < a.peers.n; r++) if (r != a.peers.rank) a.peers.posq[r][i] = np;      // the all-gather, written where the position is produced
    a.posqCorr[i] = make_float4((float)(p.x - (double)(float)p.x), (float)(p.y - (double)(float)p.y), (float)(p.z - (double)(float)p.z), 0.f);
}
template <bool PEER = false>
__device__ __forceinline__ V3 loadForce(const IntegArgs& a, int i) {
    const double sc = 1.0 / B200_FORCE_SCALE;
    if (PEER && a.usePeers == 1) {   // the reduce-scatter (scattered form): exact integer sum of every rank's partial forces for this atom
        long long fx = 0, fy = 0, fz = 0;
        for (int r = 0; r < a.peers.n; r++) { const long long* f = a.peers.force[r]; fx += f[i]; fy += f[i + a.Npad]; fz += f[i + 2 * a.Npad]; }
        return {sc * (double)fx, sc * (double)fy, sc * (double)fz};
    }
    return {sc * (double)a.force[i], sc * (double)a.force[i + a.Npad], sc * (double)a.force[i + 2 * a.Npad]};
}

// velocity constraints of a rigid triangle: solve the 3x3 system for the bond impulses
__device__ __forceinline__ void settleVelocities(V3 pA, V3 pB, V3 pC, V3& vA, V3& vB, V3& vC, double iA, double iB, double iC) {
    const V3 eAB = unit(pB - pA), eBC = unit(pC - pB), eCA = unit(pA - pC);
    const double rAB = dot(vB - vA, eAB), rBC = dot(vC - vB, eBC), rCA = dot(vA - vC, eCA);
    const double cAB_BC = dot(eAB, eBC), cBC_CA = dot(eBC, eCA), cCA_AB = dot(eCA, eAB);
    // M t = r, symmetric
    const double m00 = iA + iB, m11 = iB + iC, m22 = iC + iA;
    const double m01 = -cAB_BC * iB, m12 = -cBC_CA * iC, m02 = -cCA_AB * iA;
    const double det = m00 * (m11 * m22 - m12 * m12) - m01 * (m01 * m22 - m12 * m02) + m02 * (m01 * m12 - m11 * m02);
    const double inv = 1.0 / det;
    const double tAB = inv * (rAB * (m11 * m22 - m12 * m12) - m01 * (rBC * m22 - m12 * rCA) + m02 * (rBC * m12 - m11 * rCA));
    const double tBC = inv * (m00 * (rBC * m22 - m12 * rCA) - rAB * (m01 * m22 - m12 * m02) + m02 * (m01 * rCA - rBC * m02));
    const double tCA = inv * (m00 * (m11 * rCA - rBC * m12) - m01 * (m01 * rCA - rBC * m02) + rAB * (m01 * m12 - m11 * m02));
    vA = vA + (eAB * tAB - eCA * tCA) * iA;
    vB = vB + (eBC * tBC - eAB * tAB) * iB;
    vC = vC + (eCA * tCA - eBC * tBC) * iC;
}

// SETTLE for positions: p* are the constrained old positions, d* the unconstrained displacements (updated in place)
__device__ __forceinline__ void settlePositions(V3 p0, V3 p1, V3 p2, V3& d0, V3& d1, V3& d2, double m0, double m1, double m2, double dOH, double dHH) {
    const V3 b0 = p1 - p0, c0 = p2 - p0;
    const double invM = 1.0 / (m0 + m1 + m2);
    const V3 A1 = d0, B1 = b0 + d1, C1 = c0 + d2;
    const V3 com = (A1 * m0 + B1 * m1 + C1 * m2) * invM;
    const V3 a1 = A1 - com, b1 = B1 - com, c1 = C1 - com;
    // orthonormal frame: ez normal to the old molecular plane, ex in the plane of ez and a1
    V3 ez = cross(b0, c0);
    V3 ex = cross(a1, ez);
    V3 ey = cross(ez, ex);
    ex = unit(ex); ey = unit(ey); ez = unit(ez);
    const double b0x = dot(ex, b0), b0y = dot(ey, b0), c0x = dot(ex, c0), c0y = dot(ey, c0);
    const double a1z = dot(ez, a1);
    const double b1x = dot(ex, b1), b1y = dot(ey, b1), b1z = dot(ez, b1);
    const double c1x = dot(ex, c1), c1y = dot(ey, c1), c1z = dot(ez, c1);
    // canonical triangle
    const double rc = 0.5 * dHH;
    const double rbt = sqrt(dOH * dOH - rc * rc);
    const double ra = rbt * (m1 + m2) * invM, rb = rbt - ra;
    const double sinphi = a1z / ra, cosphi = sqrt(1.0 - sinphi * sinphi);
    const double sinpsi = (b1z - c1z) / (2.0 * rc * cosphi), cospsi = sqrt(1.0 - sinpsi * sinpsi);
    const double ya2 = ra * cosphi;
    double xb2 = -rc * cospsi;
    const double yb2 = -rb * cosphi - rc * sinpsi * sinphi, yc2 = -rb * cosphi + rc * sinpsi * sinphi;
    const double hh2 = 4.0 * xb2 * xb2 + (yb2 - yc2) * (yb2 - yc2) + (b1z - c1z) * (b1z - c1z);
    xb2 -= 0.5 * (2.0 * xb2 + sqrt(4.0 * xb2 * xb2 - hh2 + dHH * dHH));   // keeps |b-c| exactly dHH for unequal masses
    // in-plane rotation
    const double al = xb2 * (b0x - c0x) + b0y * yb2 + c0y * yc2;
    const double be = xb2 * (c0y - b0y) + b0x * yb2 + c0x * yc2;
    const double ga = b0x * b1y - b1x * b0y + c0x * c1y - c1x * c0y;
    const double ab2 = al * al + be * be;
    const double sinth = (al * ga - be * sqrt(ab2 - ga * ga)) / ab2, costh = sqrt(1.0 - sinth * sinth);
    const V3 a3 = {-ya2 * sinth, ya2 * costh, a1z};
    const V3 b3 = {xb2 * costh - yb2 * sinth, xb2 * sinth + yb2 * costh, b1z};
    const V3 c3 = {-xb2 * costh - yc2 * sinth, -xb2 * sinth + yc2 * costh, c1z};
    d0 = com + ex * a3.x + ey * a3.y + ez * a3.z;
    d1 = com + ex * b3.x + ey * b3.y + ez * b3.z - b0;
    d2 = com + ex * c3.x + ey * c3.y + ez * c3.z - c0;
}

__device__ __forceinline__ void checkMoved(const IntegArgs& a, int i, V3 p, bool& moved) {
    if (!a.refPosq) return;
    const float4 r = a.refPosq[i];
    const float dx = (float)p.x - r.x, dy = (float)p.y - r.y, dz = (float)p.z - r.z;
    moved |= dx * dx + dy * dy + dz * dz > a.maxDisp2;
}

// PEER: multi-GPU instantiation (peer-memory accesses compiled in); the single-GPU kernel carries none of that code
template <bool PEER>
__global__ void __launch_bounds__(128, 4) langevinMiddleStep(IntegArgs a, int numSettle, const int4* __restrict__ settleAtoms, const float2* __restrict__ settleParams,
                                                           int numFree, const int* __restrict__ freeAtoms) {
    const int t = a.unitBegin + blockIdx.x * blockDim.x + threadIdx.x;   // unitBegin is a multiple of the block size: blocks group the same units on any rank count
    const double halfdt = 0.5 * a.dt, invDt = 1.0 / a.dt;
    V3 vcm = {0, 0, 0}, mom = {0, 0, 0};
    if (a.fuseCM) {
        const double sc = a.fuseCM == 2 ? a.invTotalMass / CM_SCALE : 0.0;
        vcm = {sc * (double)a.counters->cm[a.cmCur][0], sc * (double)a.counters->cm[a.cmCur][1], sc * (double)a.counters->cm[a.cmCur][2]};
        if (blockIdx.x == 0 && threadIdx.x == 0) { a.counters->cm[a.cmClear][0] = 0; a.counters->cm[a.cmClear][1] = 0; a.counters->cm[a.cmClear][2] = 0; }
    }
    bool moved = false;
    if (t >= a.unitEnd) {
    } else if (t < numSettle) {
        const int4 at = settleAtoms[t]; const float2 prm = settleParams[t];
        const int idx[3] = {at.x, at.y, at.z};
        V3 p[3], v[3], d[3], old[3]; double im[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            p[k] = loadPosD(a, idx[k]);
            const double4 vm = a.velm[idx[k]];
            im[k] = vm.w;
            v[k] = V3{vm.x, vm.y, vm.z} - vcm + loadForce<PEER>(a, idx[k]) * (a.dt * vm.w);    // COM removal, then part 1
        }
        settleVelocities(p[0], p[1], p[2], v[0], v[1], v[2], im[0], im[1], im[2]);
#pragma unroll
        for (int k = 0; k < 3; k++) {                                                       // part 2
            d[k] = v[k] * halfdt;
            const V3 g = gaussian3(a.seed, a.step, a.atomIndex[idx[k]]);
            v[k] = v[k] * a.vscale + g * (a.noisescale * sqrt(im[k]));
            d[k] = d[k] + v[k] * halfdt;
            old[k] = d[k];
        }
        settlePositions(p[0], p[1], p[2], d[0], d[1], d[2], 1.0 / im[0], 1.0 / im[1], 1.0 / im[2], prm.x, prm.y);
#pragma unroll
        for (int k = 0; k < 3; k++) {                                                       // part 3
            v[k] = v[k] + (d[k] - old[k]) * invDt;
            a.velm[idx[k]] = make_double4(v[k].x, v[k].y, v[k].z, im[k]);
            const V3 pn = p[k] + d[k];
            storePosD<PEER>(a, idx[k], pn);
            checkMoved(a, idx[k], pn, moved);
            mom = mom + v[k] * (1.0 / im[k]);
        }
    } else if (t < numSettle + numFree) {
        const int i = freeAtoms[t - numSettle];
        const double4 vm = a.velm[i];
        V3 v = V3{vm.x, vm.y, vm.z} - vcm + loadForce<PEER>(a, i) * (a.dt * vm.w);
        V3 d = v * halfdt;
        const V3 g = gaussian3(a.seed, a.step, a.atomIndex[i]);
        v = v * a.vscale + g * (a.noisescale * sqrt(vm.w));
        d = d + v * halfdt;
        a.velm[i] = make_double4(v.x, v.y, v.z, vm.w);
        const V3 pn = loadPosD(a, i) + d;
        storePosD<PEER>(a, i, pn);
        checkMoved(a, i, pn, moved);
        mom = mom + v * (1.0 / vm.w);
    }
    if (a.refPosq && __any_sync(0xffffffffu, moved) && (threadIdx.x & 31) == 0) {
        a.counters->rebuild = 1;
        if (PEER && a.usePeers) for (int r = 0; r < a.peers.n; r++) if (r != a.peers.rank) a.peers.counters[r]->rebuild = 1;
    }
    if (a.fuseCM) {   // momentum of the new velocities, removed at the start of the next step (fixed-order block sum, integer atomics)
        __shared__ double red[3][128];
        red[0][threadIdx.x] = mom.x; red[1][threadIdx.x] = mom.y; red[2][threadIdx.x] = mom.z;
        __syncthreads();
        for (int o = 64; o > 0; o >>= 1) {
            if (threadIdx.x < o) for (int k = 0; k < 3; k++) red[k][threadIdx.x] += red[k][threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x < 3) {
            const long long v = (long long)(red[threadIdx.x][0] * CM_SCALE);
            atomicAddFixed(&a.counters->cm[a.cmNext][threadIdx.x], v);
            if (PEER && a.usePeers) for (int r = 0; r < a.peers.n; r++) if (r != a.peers.rank) atomicAddFixed(&a.peers.counters[r]->cm[a.cmNext][threadIdx.x], v);
        }
    }
}

// ---- SHAKE clusters: a central atom (index 0) with up to three partners, each constrained to the centre only -- the HBonds
// constraints of a solvated protein (integrationUtilities.cc:99-240 applyShakeToHydrogens).  One thread owns a cluster and
// iterates the classic SHAKE / RATTLE updates over its constraints until all of them are met to the tolerance.
#define SHAKE_MAX_ITER 60
__device__ __forceinline__ void shakeVelocities(int n, const V3* p, V3* v, const double* im, double tol, double dt) {
    for (int it = 0; it < SHAKE_MAX_ITER; it++) {
        bool done = true;
        for (int k = 1; k < n; k++) {
            const V3 s = p[k] - p[0];
            const double s2 = dot(s, s), vrel = dot(v[k] - v[0], s);
            if (fabs(vrel) * dt > tol * s2) done = false;
            const double lam = -vrel / ((im[0] + im[k]) * s2);
            v[k] = v[k] + s * (lam * im[k]); v[0] = v[0] - s * (lam * im[0]);
        }
        if (done) break;
    }
}
// p: constrained old positions, d: unconstrained displacements (updated in place)
__device__ __forceinline__ void shakePositions(int n, const V3* p, V3* d, const double* im, const float* dist, double tol) {
    for (int it = 0; it < SHAKE_MAX_ITER; it++) {
        bool done = true;
        for (int k = 1; k < n; k++) {
            const V3 s = p[k] - p[0], sp = s + d[k] - d[0];
            const double d2 = (double)dist[k - 1] * (double)dist[k - 1], diff = d2 - dot(sp, sp);
            if (fabs(diff) > 2.0 * tol * d2) done = false;
            const double lam = diff / (2.0 * (im[0] + im[k]) * dot(s, sp));
            d[k] = d[k] + s * (lam * im[k]); d[0] = d[0] - s * (lam * im[0]);
        }
        if (done) break;
    }
}

// LangevinMiddle step of the SHAKE clusters (same three parts and the same fused extras as langevinMiddleStep)
__global__ void __launch_bounds__(128) langevinMiddleShake(IntegArgs a, int numShake, const int4* __restrict__ shakeAtoms, const float4* __restrict__ shakeDist, double tol) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const double halfdt = 0.5 * a.dt, invDt = 1.0 / a.dt;
    V3 vcm = {0, 0, 0}, mom = {0, 0, 0};
    if (a.fuseCM == 2) {
        const double sc = a.invTotalMass / CM_SCALE;
        vcm = {sc * (double)a.counters->cm[a.cmCur][0], sc * (double)a.counters->cm[a.cmCur][1], sc * (double)a.counters->cm[a.cmCur][2]};
    }
    bool moved = false;
    if (t < numShake) {
        const int4 at = shakeAtoms[t]; const float4 dd = shakeDist[t];
        const int idx[4] = {at.x, at.y, at.z, at.w};
        const float dist[3] = {dd.x, dd.y, dd.z};
        const int n = at.w >= 0 ? 4 : at.z >= 0 ? 3 : 2;
        V3 p[4], v[4], d[4], old[4]; double im[4];
        for (int k = 0; k < n; k++) {
            p[k] = loadPosD(a, idx[k]);
            const double4 vm = a.velm[idx[k]];
            im[k] = vm.w;
            v[k] = V3{vm.x, vm.y, vm.z} - vcm + loadForce<false>(a, idx[k]) * (a.dt * vm.w);
        }
        shakeVelocities(n, p, v, im, tol, a.dt);
        for (int k = 0; k < n; k++) {
            d[k] = v[k] * halfdt;
            const V3 g = gaussian3(a.seed, a.step, a.atomIndex[idx[k]]);
            v[k] = v[k] * a.vscale + g * (a.noisescale * sqrt(im[k]));
            d[k] = d[k] + v[k] * halfdt;
            old[k] = d[k];
        }
        shakePositions(n, p, d, im, dist, tol);
        for (int k = 0; k < n; k++) {
            v[k] = v[k] + (d[k] - old[k]) * invDt;
            a.velm[idx[k]] = make_double4(v[k].x, v[k].y, v[k].z, im[k]);
            const V3 pn = p[k] + d[k];
            storePosD<false>(a, idx[k], pn);
            checkMoved(a, idx[k], pn, moved);
            mom = mom + v[k] * (1.0 / im[k]);
        }
    }
    if (a.refPosq && __any_sync(0xffffffffu, moved) && (threadIdx.x & 31) == 0) a.counters->rebuild = 1;
    if (a.fuseCM) {
        __shared__ double red[3][128];
        red[0][threadIdx.x] = mom.x; red[1][threadIdx.x] = mom.y; red[2][threadIdx.x] = mom.z;
        __syncthreads();
        for (int o = 64; o > 0; o >>= 1) {
            if (threadIdx.x < o) for (int k = 0; k < 3; k++) red[k][threadIdx.x] += red[k][threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x < 3) atomicAddFixed(&a.counters->cm[a.cmNext][threadIdx.x], (long long)(red[threadIdx.x][0] * CM_SCALE));
    }
}

// stand-alone constraint passes over the SHAKE clusters (ApplyConstraintsKernel::apply / applyToVelocities)
__global__ void shakeOnly(IntegArgs a, int numShake, const int4* __restrict__ shakeAtoms, const float4* __restrict__ shakeDist, double tol, int velocities) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= numShake) return;
    const int4 at = shakeAtoms[t]; const float4 dd = shakeDist[t];
    const int idx[4] = {at.x, at.y, at.z, at.w};
    const float dist[3] = {dd.x, dd.y, dd.z};
    const int n = at.w >= 0 ? 4 : at.z >= 0 ? 3 : 2;
    V3 p[4], v[4], d[4]; double im[4];
    for (int k = 0; k < n; k++) { p[k] = loadPosD(a, idx[k]); const double4 vm = a.velm[idx[k]]; v[k] = {vm.x, vm.y, vm.z}; im[k] = vm.w; d[k] = {0, 0, 0}; }
    if (velocities) {
        shakeVelocities(n, p, v, im, tol, a.dt > 0 ? a.dt : 0.001);
        for (int k = 0; k < n; k++) a.velm[idx[k]] = make_double4(v[k].x, v[k].y, v[k].z, im[k]);
    } else {
        shakePositions(n, p, d, im, dist, tol);
        for (int k = 0; k < n; k++) storePosD(a, idx[k], p[k] + d[k]);
    }
}

// ---- general constraint clusters.  Whatever is neither a rigid 3-site molecule nor a SHAKE star (all-bonds or angle constraints,
// chains, rings) is solved the way CCMA is used in the reference (integrationUtilities.cc:582-804): iteratively to the tolerance.
// One CTA owns one connected cluster; its constraints are ordered by colour (no two constraints of a colour share an atom), so the
// threads of the CTA update a whole colour in parallel and a sweep over the colours is a Gauss-Seidel sweep over the cluster.  The
// working copies (velocities in velm, displacements in gcDelta) live in global memory; __syncthreads orders the colours.
#define GC_THREADS 128
#define GC_MAX_ITER 500
struct GcArgs { const int2* pairs; const double* dist; const int* colorStart; const int2* clusterColors; const int* atomStart; const int* atomList; double4* delta; double4* old; double tol; };

__device__ __forceinline__ void gcVelocitySweeps(const IntegArgs& a, const GcArgs& g, int first, int ncol, int* flag) {
    for (int it = 0; it < GC_MAX_ITER; it++) {
        __syncthreads();
        if (threadIdx.x == 0) *flag = 0;
        __syncthreads();
        for (int col = 0; col < ncol; col++) {
            for (int q = g.colorStart[first + col] + threadIdx.x; q < g.colorStart[first + col + 1]; q += blockDim.x) {
                const int2 p = g.pairs[q];
                const V3 s = loadPosD(a, p.y) - loadPosD(a, p.x);
                double4 vi = a.velm[p.x], vj = a.velm[p.y];
                const double s2 = dot(s, s), vrel = (vj.x - vi.x) * s.x + (vj.y - vi.y) * s.y + (vj.z - vi.z) * s.z;
                if (fabs(vrel) * (a.dt > 0 ? a.dt : 0.001) > g.tol * s2) *flag = 1;
                const double lam = -vrel / ((vi.w + vj.w) * s2);
                vj.x += s.x * lam * vj.w; vj.y += s.y * lam * vj.w; vj.z += s.z * lam * vj.w;
                vi.x -= s.x * lam * vi.w; vi.y -= s.y * lam * vi.w; vi.z -= s.z * lam * vi.w;
                a.velm[p.x] = vi; a.velm[p.y] = vj;
            }
            __syncthreads();
        }
        if (!*flag) break;
    }
}
__device__ __forceinline__ void gcPositionSweeps(const IntegArgs& a, const GcArgs& g, int first, int ncol, int* flag) {
    for (int it = 0; it < GC_MAX_ITER; it++) {
        __syncthreads();
        if (threadIdx.x == 0) *flag = 0;
        __syncthreads();
        for (int col = 0; col < ncol; col++) {
            for (int q = g.colorStart[first + col] + threadIdx.x; q < g.colorStart[first + col + 1]; q += blockDim.x) {
                const int2 p = g.pairs[q];
                const V3 s = loadPosD(a, p.y) - loadPosD(a, p.x);
                double4 di = g.delta[p.x], dj = g.delta[p.y];
                const V3 sp = {s.x + dj.x - di.x, s.y + dj.y - di.y, s.z + dj.z - di.z};
                const double imi = a.velm[p.x].w, imj = a.velm[p.y].w;
                const double d2 = g.dist[q] * g.dist[q], diff = d2 - dot(sp, sp);
                if (fabs(diff) > 2.0 * g.tol * d2) *flag = 1;
                const double lam = diff / (2.0 * (imi + imj) * dot(s, sp));
                dj.x += s.x * lam * imj; dj.y += s.y * lam * imj; dj.z += s.z * lam * imj;
                di.x -= s.x * lam * imi; di.y -= s.y * lam * imi; di.z -= s.z * lam * imi;
                g.delta[p.x] = di; g.delta[p.y] = dj;
            }
            __syncthreads();
        }
        if (!*flag) break;
    }
}

// mode 0: full LangevinMiddle step of the cluster's atoms; 1: velocity constraints only; 2: position constraints only
__global__ void __launch_bounds__(GC_THREADS) generalClusterStep(IntegArgs a, GcArgs g, int mode) {
    __shared__ int flag;
    const int k = blockIdx.x;
    const int2 cc = g.clusterColors[k];
    const int aBegin = g.atomStart[k], aEnd = g.atomStart[k + 1];
    const double halfdt = 0.5 * a.dt, invDt = a.dt > 0 ? 1.0 / a.dt : 0.0;
    V3 vcm = {0, 0, 0}, mom = {0, 0, 0};
    bool moved = false;
    if (mode == 0) {
        if (a.fuseCM == 2) {
            const double sc = a.invTotalMass / CM_SCALE;
            vcm = {sc * (double)a.counters->cm[a.cmCur][0], sc * (double)a.counters->cm[a.cmCur][1], sc * (double)a.counters->cm[a.cmCur][2]};
        }
        for (int q = aBegin + threadIdx.x; q < aEnd; q += blockDim.x) {      // part 1: kick
            const int i = g.atomList[q];
            double4 vm = a.velm[i];
            const V3 f = loadForce<false>(a, i);
            vm.x += -vcm.x + f.x * a.dt * vm.w; vm.y += -vcm.y + f.y * a.dt * vm.w; vm.z += -vcm.z + f.z * a.dt * vm.w;
            a.velm[i] = vm;
        }
    }
    if (mode == 0 || mode == 1) gcVelocitySweeps(a, g, cc.x, cc.y, &flag);
    if (mode == 1) return;
    __syncthreads();
    for (int q = aBegin + threadIdx.x; q < aEnd; q += blockDim.x) {          // part 2: drift, thermostat, drift (unconstrained displacement)
        const int i = g.atomList[q];
        V3 d = {0, 0, 0};
        if (mode == 0) {
            double4 vm = a.velm[i];
            V3 v = {vm.x, vm.y, vm.z};
            d = v * halfdt;
            const V3 gs = gaussian3(a.seed, a.step, a.atomIndex[i]);
            v = v * a.vscale + gs * (a.noisescale * sqrt(vm.w));
            d = d + v * halfdt;
            a.velm[i] = make_double4(v.x, v.y, v.z, vm.w);
        }
        g.delta[i] = make_double4(d.x, d.y, d.z, 0.0); g.old[i] = make_double4(d.x, d.y, d.z, 0.0);
    }
    gcPositionSweeps(a, g, cc.x, cc.y, &flag);
    __syncthreads();
    for (int q = aBegin + threadIdx.x; q < aEnd; q += blockDim.x) {          // part 3: constraint correction of the velocities, new positions
        const int i = g.atomList[q];
        const double4 d = g.delta[i], o = g.old[i];
        if (mode == 0) {
            double4 vm = a.velm[i];
            vm.x += (d.x - o.x) * invDt; vm.y += (d.y - o.y) * invDt; vm.z += (d.z - o.z) * invDt;
            a.velm[i] = vm;
            if (vm.w != 0.0) mom = mom + V3{vm.x, vm.y, vm.z} * (1.0 / vm.w);
        }
        const V3 pn = loadPosD(a, i) + V3{d.x, d.y, d.z};
        storePosD<false>(a, i, pn);
        if (mode == 0) checkMoved(a, i, pn, moved);
    }
    if (mode != 0) return;
    if (a.refPosq && __any_sync(0xffffffffu, moved) && (threadIdx.x & 31) == 0) a.counters->rebuild = 1;
    if (a.fuseCM) {
        __shared__ double red[3][GC_THREADS];
        red[0][threadIdx.x] = mom.x; red[1][threadIdx.x] = mom.y; red[2][threadIdx.x] = mom.z;
        __syncthreads();
        for (int o = GC_THREADS / 2; o > 0; o >>= 1) {
            if (threadIdx.x < o) for (int k2 = 0; k2 < 3; k2++) red[k2][threadIdx.x] += red[k2][threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x < 3) atomicAddFixed(&a.counters->cm[a.cmNext][threadIdx.x], (long long)(red[threadIdx.x][0] * CM_SCALE));
    }
}

static GcArgs makeGcArgs(b200_ctx* c, double tol) {
    GcArgs g; g.pairs = c->gcPairs; g.dist = c->gcDist; g.colorStart = c->gcColorStart; g.clusterColors = c->gcClusterColors;
    g.atomStart = c->gcAtomStart; g.atomList = c->gcAtomList; g.delta = c->gcDelta; g.old = c->gcOld; g.tol = tol;
    return g;
}

// ---- centre-of-mass motion removal (platforms/common/src/kernels/removeCM.cc)
__global__ void __launch_bounds__(256) cmMomentum(const double4* __restrict__ velm, int N, DevCounters* counters, int slot) {
    double px = 0, py = 0, pz = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const double4 v = velm[i];
        if (v.w != 0.0) { const double m = 1.0 / v.w; px += m * v.x; py += m * v.y; pz += m * v.z; }
    }
    __shared__ double red[3][256];
    red[0][threadIdx.x] = px; red[1][threadIdx.x] = py; red[2][threadIdx.x] = pz;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) for (int k = 0; k < 3; k++) red[k][threadIdx.x] += red[k][threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x < 3) atomicAddFixed(&counters->cm[slot][threadIdx.x], (long long)(red[threadIdx.x][0] * CM_SCALE));
}
__global__ void cmRemove(double4* velm, int N, double invTotalMass, DevCounters* counters, int slot) {
    const double sc = invTotalMass / CM_SCALE;
    const double vx = sc * (double)counters->cm[slot][0], vy = sc * (double)counters->cm[slot][1], vz = sc * (double)counters->cm[slot][2];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        double4 v = velm[i];
        if (v.w != 0.0) { v.x -= vx; v.y -= vy; v.z -= vz; velm[i] = v; }
    }
}
__global__ void cmZero(DevCounters* counters, int slot) { counters->cm[slot][0] = counters->cm[slot][1] = counters->cm[slot][2] = 0; }

__global__ void __launch_bounds__(256) kineticEnergy(const double4* __restrict__ velm, int N, DevCounters* counters) {
    double e = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const double4 v = velm[i];
        if (v.w != 0.0) e += (v.x * v.x + v.y * v.y + v.z * v.z) / v.w;
    }
    __shared__ double red[256];
    red[threadIdx.x] = e;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) atomicAddFixed(&counters->energy[E_KINETIC], toFixedD(0.5 * red[0] * (1.0 / 1024.0)));   // scaled: KE of 1M atoms ~ 4e6 kJ/mol
}
__global__ void keZero(DevCounters* counters) { counters->energy[E_KINETIC] = 0; }

// standalone constraint passes (ApplyConstraintsKernel::apply / applyToVelocities, kernels.h:251-278)
__global__ void settleOnly(IntegArgs a, int numSettle, const int4* __restrict__ settleAtoms, const float2* __restrict__ settleParams, int velocities) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= numSettle) return;
    const int4 at = settleAtoms[t]; const float2 prm = settleParams[t];
    const int idx[3] = {at.x, at.y, at.z};
    V3 p[3], v[3]; double im[3];
    for (int k = 0; k < 3; k++) { p[k] = loadPosD(a, idx[k]); const double4 vm = a.velm[idx[k]]; v[k] = {vm.x, vm.y, vm.z}; im[k] = vm.w; }
    if (velocities) {
        settleVelocities(p[0], p[1], p[2], v[0], v[1], v[2], im[0], im[1], im[2]);
        for (int k = 0; k < 3; k++) a.velm[idx[k]] = make_double4(v[k].x, v[k].y, v[k].z, im[k]);
    } else {
        // as ApplyConstraintsKernel::apply does: zero displacement, SETTLE, add the resulting displacement
        V3 d0 = {0, 0, 0}, d1 = {0, 0, 0}, d2 = {0, 0, 0};
        settlePositions(p[0], p[1], p[2], d0, d1, d2, 1.0 / im[0], 1.0 / im[1], 1.0 / im[2], prm.x, prm.y);
        const V3 q0 = p[0] + d0, q1 = p[1] + d1, q2 = p[2] + d2;
        storePosD(a, idx[0], q0); storePosD(a, idx[1], q1); storePosD(a, idx[2], q2);
    }
}

static IntegArgs makeArgs(b200_ctx* c) {
    IntegArgs a;
    a.posq = c->posq; a.posqCorr = c->posqCorr; a.velm = c->velm; a.force = c->force; a.atomIndex = c->atomIndex; a.Npad = c->Npad;
    a.dt = c->integ.dt;
    const double kT = 0.00831446261815324 * c->integ.temperature;   // BOLTZ (SimTKOpenMMRealType.h)
    a.vscale = exp(-c->integ.dt * c->integ.friction);
    a.noisescale = sqrt(kT * (1.0 - a.vscale * a.vscale));
    a.seed = c->integ.seed; a.step = (unsigned long long)c->stepCount;
    a.refPosq = nullptr; a.maxDisp2 = 0.f; a.counters = c->counters; a.fuseCM = 0; a.cmCur = a.cmNext = a.cmClear = 0; a.invTotalMass = 0;
    a.usePeers = 0; a.unitBegin = 0; a.unitEnd = c->numSettle + c->numFree; memset(&a.peers, 0, sizeof(a.peers));
    return a;
}

// sum of the particle masses, computed once (this is on the per-step host path)
static double totalMass(b200_ctx* c) {
    if (c->totalMassCached <= 0) { double t = 0; for (double m : c->mass) t += m; c->totalMassCached = t; }
    return c->totalMassCached;
}

static void removeCM(b200_ctx* c) {
    cudaStream_t s = c->stream;
    const double total = totalMass(c);
    cmZero<<<1, 1, 0, s>>>(c->counters, 3);
    cmMomentum<<<c->smCount * 2, 256, 0, s>>>(c->velm, c->N, c->counters, 3);
    cmRemove<<<c->smCount * 4, 256, 0, s>>>(c->velm, c->N, 1.0 / total, c->counters, 3);
    LAUNCH_COUNT(c, 3);
    c->cmValid = false; c->cmDeferred = false;
}

// The momentum of the current velocities lives in cm[stepCount % 3] (written by the previous integration kernel).  A removal
// request only raises a flag: the next integration kernel subtracts the COM velocity while it loads the velocities anyway.
// Forces do not depend on velocities, so this equals removing it before the force evaluation (CMMotionRemover::updateContextState).
void b200_request_cm_removal(b200_ctx* c) {
    cudaStream_t s = c->stream;
    if (!c->cmValid) {   // first step, or the velocities were set from outside
        cmZero<<<1, 1, 0, s>>>(c->counters, 0); cmZero<<<1, 1, 0, s>>>(c->counters, 1); cmZero<<<1, 1, 0, s>>>(c->counters, 2);
        cmMomentum<<<c->smCount * 2, 256, 0, s>>>(c->velm, c->N, c->counters, (int)(c->stepCount % 3));
        LAUNCH_COUNT(c, 4);
        c->cmValid = true;
    }
    c->cmDeferred = true;
}
// anything that reads or replaces the velocities applies a pending removal first
void b200_flush_cm(b200_ctx* c) {
    if (!c->cmDeferred) return;
    const int slot = (int)(c->stepCount % 3);
    cmRemove<<<c->smCount * 4, 256, 0, c->stream>>>(c->velm, c->N, 1.0 / totalMass(c), c->counters, slot);
    cmZero<<<1, 1, 0, c->stream>>>(c->counters, slot);     // the momentum of the corrected velocities is zero
    LAUNCH_COUNT(c, 2);
    c->cmDeferred = false;
}

// RemoveCMMotionKernel::execute (kernels.h:1793)
extern "C" int b200_remove_cm_motion(b200_ctx* c) {
    try {
        CUDA_CHECK(cudaSetDevice(c->device));
        b200_request_cm_removal(c);     // (multi-GPU: the momentum buffer is kept in step on every rank by the integration kernels)
        CUDA_CHECK(cudaGetLastError());
    }
    catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
    return B200_OK;
}

// IntegrateLangevinMiddleStepKernel::execute (kernels.h:1412): one integration step with the forces already in the buffer
extern "C" int b200_integrate(b200_ctx* c) {
    try {
        if (!c->hasIntegrator) throw B200Error(B200_ERR_INVALID, "b200_integrate: no integrator");
        CUDA_CHECK(cudaSetDevice(c->device));
        b200_finalize(c);
        if (c->numOtherConstraints > 0) throw B200Error(B200_ERR_UNSUPPORTED, "constraints that are neither rigid 3-site molecules (SETTLE) nor SHAKE clusters are not implemented");
        b200_advance(c, nullptr);          // the caller (CMMotionRemover's own kernel) decides when COM motion is removed
    } catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
    return B200_OK;
}

void b200_integrate_step(b200_ctx* c) {
    cudaStream_t s = c->stream;
    ScopedTimer tm(c, B200_T_INTEGRATE, s);
    IntegArgs a = makeArgs(c);
    // momentum bookkeeping rides in the kernel whenever the buffer is in step with the velocities; a pending removal is applied there
    if (c->cmValid) {
        a.fuseCM = c->cmDeferred ? 2 : 1; a.invTotalMass = 1.0 / totalMass(c);
        a.cmCur = (int)(c->stepCount % 3); a.cmNext = (int)((c->stepCount + 1) % 3); a.cmClear = (int)((c->stepCount + 2) % 3);
    }
    c->cmDeferred = false;
    // so does the Verlet-buffer displacement check of the next force evaluation
    if (c->hasNonbonded) { a.refPosq = c->refPosq; const float md = 0.5f * (float)c->padding; a.maxDisp2 = md * md; c->dispChecked = true; }
    const int units = c->numSettle + c->numFree;
    if (c->peerStepActive) {
        // a.usePeers = 1: per-atom loads/stores through peer memory; 2: the forces of this slice were already reduced into the local
        // array and the positions are published afterwards by range copies (slice = one contiguous range of slots)
        // usePeers bit 0: new positions are stored into every rank's posq by the integration kernel itself (fire-and-forget NVLink stores);
        // bit 1: the forces of this slice were already reduced into the local array (contiguous slice)
        a.usePeers = c->shardContiguous ? (c->fusedPublish ? 3 : 2) : 1; a.peers = c->peers; a.unitBegin = c->shardBegin; a.unitEnd = c->shardEnd;
        const int mine = c->shardEnd - c->shardBegin;
        langevinMiddleStep<true><<<std::max(1, (mine + 127) / 128), 128, 0, s>>>(a, c->numSettle, c->settleAtoms, c->settleParams, c->numFree, c->freeAtoms);
        LAUNCH_COUNT(c, 1);
        c->shardsDirty = true;
    } else if (units > 0) {
        langevinMiddleStep<false><<<(units + 127) / 128, 128, 0, s>>>(a, c->numSettle, c->settleAtoms, c->settleParams, c->numFree, c->freeAtoms);
        LAUNCH_COUNT(c, 1);
    }
    if (c->numShake > 0) {
        langevinMiddleShake<<<(c->numShake + 127) / 128, 128, 0, s>>>(a, c->numShake, c->shakeAtoms, c->shakeDist, c->integ.constraint_tol > 0 ? c->integ.constraint_tol : 1e-5);
        LAUNCH_COUNT(c, 1);
    }
    if (c->numGcClusters > 0) {
        generalClusterStep<<<c->numGcClusters, GC_THREADS, 0, s>>>(a, makeGcArgs(c, c->integ.constraint_tol > 0 ? c->integ.constraint_tol : 1e-5), 0);
        LAUNCH_COUNT(c, 1);
    }
    CUDA_CHECK(cudaGetLastError());
}

extern "C" int b200_apply_constraints(b200_ctx* c, double tol) {
    try {
        CUDA_CHECK(cudaSetDevice(c->device));
        b200_peer_sync_state(c);
        b200_finalize(c);
        if (c->numOtherConstraints > 0) throw B200Error(B200_ERR_UNSUPPORTED, "constraints that are neither rigid 3-site molecules (SETTLE) nor SHAKE clusters are not implemented");
        if (c->numSettle > 0) { settleOnly<<<(c->numSettle + 127) / 128, 128, 0, c->stream>>>(makeArgs(c), c->numSettle, c->settleAtoms, c->settleParams, 0); LAUNCH_COUNT(c, 1); }
        if (c->numShake > 0) { shakeOnly<<<(c->numShake + 127) / 128, 128, 0, c->stream>>>(makeArgs(c), c->numShake, c->shakeAtoms, c->shakeDist, tol > 0 ? tol : 1e-5, 0); LAUNCH_COUNT(c, 1); }
        if (c->numGcClusters > 0) { generalClusterStep<<<c->numGcClusters, GC_THREADS, 0, c->stream>>>(makeArgs(c), makeGcArgs(c, tol > 0 ? tol : 1e-5), 2); LAUNCH_COUNT(c, 1); }
        CUDA_CHECK(cudaStreamSynchronize(c->stream));
        c->stepsSinceReorder = -1; c->dispChecked = false; c->listVerified = false; c->forcesValid = false;
    } catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
    return B200_OK;
}
extern "C" int b200_apply_velocity_constraints(b200_ctx* c, double tol) {
    try {
        CUDA_CHECK(cudaSetDevice(c->device));
        b200_peer_sync_state(c);
        b200_finalize(c);
        b200_flush_cm(c);
        if (c->numOtherConstraints > 0) throw B200Error(B200_ERR_UNSUPPORTED, "constraints that are neither rigid 3-site molecules (SETTLE) nor SHAKE clusters are not implemented");
        if (c->numSettle > 0) { settleOnly<<<(c->numSettle + 127) / 128, 128, 0, c->stream>>>(makeArgs(c), c->numSettle, c->settleAtoms, c->settleParams, 1); LAUNCH_COUNT(c, 1); }
        if (c->numShake > 0) { shakeOnly<<<(c->numShake + 127) / 128, 128, 0, c->stream>>>(makeArgs(c), c->numShake, c->shakeAtoms, c->shakeDist, tol > 0 ? tol : 1e-5, 1); LAUNCH_COUNT(c, 1); }
        if (c->numGcClusters > 0) { generalClusterStep<<<c->numGcClusters, GC_THREADS, 0, c->stream>>>(makeArgs(c), makeGcArgs(c, tol > 0 ? tol : 1e-5), 1); LAUNCH_COUNT(c, 1); }
        c->cmValid = false;
        CUDA_CHECK(cudaStreamSynchronize(c->stream));
    } catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
    return B200_OK;
}
extern "C" int b200_kinetic_energy(b200_ctx* c, double* ke) {
    try {
        CUDA_CHECK(cudaSetDevice(c->device));
        b200_peer_sync_state(c);
        b200_flush_cm(c);
        keZero<<<1, 1, 0, c->stream>>>(c->counters);
        kineticEnergy<<<c->smCount * 2, 256, 0, c->stream>>>(c->velm, c->N, c->counters);
        LAUNCH_COUNT(c, 2);
        CUDA_CHECK(cudaMemcpyAsync(c->countersHost, c->counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
        CUDA_CHECK(cudaStreamSynchronize(c->stream));
        *ke = (double)c->countersHost->energy[E_KINETIC] * (1024.0 / B200_ENERGY_SCALE);
    } catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
    return B200_OK;
}
// Kinetic energy of velocities shifted in time by `shift` with the forces in the buffer, v + shift * F/m, after the velocity
// constraints: what a leapfrog integrator reports (IntegrationUtilities::computeKineticEnergy(0.5 * dt), CommonKernels.cpp:3117,
// called by IntegrateVerletStepKernel::computeKineticEnergy), because its stored velocities lag the positions by half a step.
__global__ void shiftVelocities(double4* velm, const long long* __restrict__ force, int N, int Npad, double shift) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    double4 v = velm[i];
    const double sc = shift * v.w / B200_FORCE_SCALE;
    v.x += sc * (double)force[i]; v.y += sc * (double)force[i + Npad]; v.z += sc * (double)force[i + 2 * (size_t)Npad];
    velm[i] = v;
}
extern "C" int b200_kinetic_energy_shifted(b200_ctx* c, double shift, double* ke) {
    try {
        if (shift == 0.0) return b200_kinetic_energy(c, ke);
        CUDA_CHECK(cudaSetDevice(c->device));
        b200_peer_sync_state(c);
        b200_finalize(c);
        b200_flush_cm(c);
        b200_ensure_forces_reduced(c);
        cudaStream_t s = c->stream;
        CUDA_CHECK(cudaMemcpyAsync(c->velmTmp, c->velm, sizeof(double4) * c->Npad, cudaMemcpyDeviceToDevice, s));
        shiftVelocities<<<(c->N + 255) / 256, 256, 0, s>>>(c->velm, c->force, c->N, c->Npad, shift);
        const double tol = c->hasIntegrator && c->integ.constraint_tol > 0 ? c->integ.constraint_tol : 1e-5;
        if (c->numSettle > 0) settleOnly<<<(c->numSettle + 127) / 128, 128, 0, s>>>(makeArgs(c), c->numSettle, c->settleAtoms, c->settleParams, 1);
        if (c->numShake > 0) shakeOnly<<<(c->numShake + 127) / 128, 128, 0, s>>>(makeArgs(c), c->numShake, c->shakeAtoms, c->shakeDist, tol, 1);
        if (c->numGcClusters > 0) generalClusterStep<<<c->numGcClusters, GC_THREADS, 0, s>>>(makeArgs(c), makeGcArgs(c, tol), 1);
        keZero<<<1, 1, 0, s>>>(c->counters);
        kineticEnergy<<<c->smCount * 2, 256, 0, s>>>(c->velm, c->N, c->counters);
        CUDA_CHECK(cudaMemcpyAsync(c->velm, c->velmTmp, sizeof(double4) * c->Npad, cudaMemcpyDeviceToDevice, s));
        LAUNCH_COUNT(c, 5);
        CUDA_CHECK(cudaMemcpyAsync(c->countersHost, c->counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        *ke = (double)c->countersHost->energy[E_KINETIC] * (1024.0 / B200_ENERGY_SCALE);
    } catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
    return B200_OK;
}
void b200_integrate_setup(b200_ctx*) {}
