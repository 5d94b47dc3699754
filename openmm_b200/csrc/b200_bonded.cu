// b200_bonded.cu -- per-pair / per-term kernels that ride along with the nonbonded path:
//   * Ewald exclusion correction, -erf(alpha r)/r for every excluded pair (platforms/common/src/kernels/pmeExclusions.cc:20-57)
//   * 1-4 exceptions, plain Coulomb + LJ (platforms/common/src/kernels/nonbondedExceptions.cc:1-19)
//   * HarmonicBond / HarmonicAngle / PeriodicTorsion (harmonicBondForce.cc, harmonicAngleForce.cc, periodicTorsionForce.cc)
// The reference compiles all of them into one computeBondedForces kernel (BondedUtilities.cpp:71-168); here they are
// small FP64 kernels (a few terms per atom, not an optimisation target) that add into the same int64 force buffer.
#include "b200_internal.cuh"
#include <algorithm>
#include <vector>

struct D3 { double x, y, z; };
__device__ __forceinline__ D3 operator+(D3 a, D3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ D3 operator-(D3 a, D3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ D3 operator*(D3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ D3 cross(D3 a, D3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ D3 loadPos(const float4* posq, int i) { const float4 p = posq[i]; return {(double)p.x, (double)p.y, (double)p.z}; }
__device__ __forceinline__ void addForce(long long* f, int Npad, int i, D3 v) {
    atomicAddFixed(&f[i], toFixedD(v.x)); atomicAddFixed(&f[i + Npad], toFixedD(v.y)); atomicAddFixed(&f[i + 2 * Npad], toFixedD(v.z));
}
__device__ __forceinline__ void blockEnergy(double e, long long* slot) {
    // fixed-order reduction inside the block, integer accumulation across blocks
    __shared__ double red[128];
    red[threadIdx.x] = e;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0 && red[0] != 0.0) atomicAddFixed(slot, toFixedD(red[0]));
}

__global__ void __launch_bounds__(128) exclusionCorrection(const float4* __restrict__ posq, const float2* __restrict__ sigEps, const int2* __restrict__ pairs, int n, BoxD box, double alpha,
                                                            double ljAlpha, int periodic, long long* force, int Npad, DevCounters* counters, int wantForce, int wantEnergy) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    double e = 0;
    if (t < n) {
        const int2 p = pairs[t];
        const float4 a = posq[p.x], b = posq[p.y];
        const double qq = B200_ONE_4PI_EPS0 * (double)a.w * (double)b.w;
        if (sigEps) {
            // LJPME: back out the reciprocal-space dispersion of the excluded pair (ReferenceLJCoulombIxn.cpp:497-513,
            // pmeExclusions.cc USE_LJPME branch)
            const float2 sa = sigEps[p.x], sb = sigEps[p.y];
            const double c6 = 64.0 * ((double)sa.x * sa.x * sa.x * sa.y) * ((double)sb.x * sb.x * sb.x * sb.y);
            if (c6 != 0.0) {
                double dx = (double)b.x - a.x, dy = (double)b.y - a.y, dz = (double)b.z - a.z;
                if (periodic) imageDeltaD(dx, dy, dz, box);
                const double r2 = dx * dx + dy * dy + dz * dz, invR2 = 1.0 / r2, invR6 = invR2 * invR2 * invR2;
                const double x = ljAlpha * ljAlpha * r2, x2 = x * x, ex = exp(-x);
                e += c6 * invR6 * (1.0 - ex * (1.0 + x + 0.5 * x2));
                if (wantForce) {
                    const double s = 6.0 * c6 * invR6 * invR2 * (1.0 - ex * (1.0 + x + 0.5 * x2 + x2 * x / 6.0));   // -(dE/dr)/r
                    addForce(force, Npad, p.x, D3{-dx * s, -dy * s, -dz * s});
                    addForce(force, Npad, p.y, D3{dx * s, dy * s, dz * s});
                }
            }
        }
        if (qq != 0.0) {
            // like the 1-4 terms, the correction of an excluded pair uses the plain displacement unless the force asks for periodic
            // exceptions (pmeExclusions.cc:3, USE_PERIODIC = getExceptionsUsePeriodicBoundaryConditions; molecules are kept whole)
            double dx = (double)b.x - a.x, dy = (double)b.y - a.y, dz = (double)b.z - a.z;
            if (periodic) imageDeltaD(dx, dy, dz, box);
            const double r2 = dx * dx + dy * dy + dz * dz, r = sqrt(r2), ar = alpha * r;
            if (ar > 1e-6) {
                const double erfAR = erf(ar), invR = 1.0 / r;
                e += -qq * invR * erfAR;
                if (wantForce) {
                    // s = -(dE/dr)/r of E = -qq erf(alpha r)/r
                    const double s = -qq * invR * invR * invR * (erfAR - ar * 1.1283791670955126 * exp(-ar * ar));
                    addForce(force, Npad, p.x, D3{-dx * s, -dy * s, -dz * s});
                    addForce(force, Npad, p.y, D3{dx * s, dy * s, dz * s});
                }
            } else {
                e += -alpha * 1.1283791670955126 * qq;   // r -> 0 limit (pmeExclusions.cc:54-56)
            }
        }
    }
    if (wantEnergy) blockEnergy(e, &counters->energy[E_EXCLCORR]);
}

__global__ void __launch_bounds__(128) exceptions14(const float4* __restrict__ posq, const int2* __restrict__ pairs, const float4* __restrict__ params, int n,
                                                     BoxD box, int periodic, long long* force, int Npad, DevCounters* counters, int wantForce, int wantEnergy) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    double e = 0;
    if (t < n) {
        const int2 p = pairs[t];
        const float4 prm = params[t];   // (k*qq, sigma, 4 eps)
        D3 d = loadPos(posq, p.y) - loadPos(posq, p.x);
        if (periodic) imageDeltaD(d.x, d.y, d.z, box);
        const double r2 = dot(d, d), invR = rsqrt(r2), sr2 = (double)prm.y * prm.y * invR * invR, sr6 = sr2 * sr2 * sr2;
        e = prm.x * invR + prm.z * sr6 * (sr6 - 1.0);
        if (wantForce) {
            const double s = (prm.x * invR + prm.z * sr6 * (12.0 * sr6 - 6.0)) * invR * invR;
            addForce(force, Npad, p.x, d * (-s));
            addForce(force, Npad, p.y, d * s);
        }
    }
    if (wantEnergy) blockEnergy(e, &counters->energy[E_EXCEPT]);
}

// `periodic`: HarmonicBondForce::setUsesPeriodicBoundaryConditions -- every displacement is replaced by its nearest periodic image
// (harmonicBondForce.cc APPLY_PERIODIC_TO_DELTA); likewise for angles and torsions below
__global__ void __launch_bounds__(128) harmonicBonds(const float4* __restrict__ posq, const int2* __restrict__ atoms, const float2* __restrict__ params, int n,
                                                      BoxD box, int periodic, long long* force, int Npad, DevCounters* counters, int wantForce, int wantEnergy) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    double e = 0;
    if (t < n) {
        const int2 a = atoms[t]; const float2 prm = params[t];   // (r0, k)
        D3 d = loadPos(posq, a.y) - loadPos(posq, a.x);
        if (periodic) imageDeltaD(d.x, d.y, d.z, box);
        const double r = sqrt(dot(d, d)), dr = r - prm.x;
        e = 0.5 * prm.y * dr * dr;
        if (wantForce && r > 0) {
            const double s = -prm.y * dr / r;     // force on atom y = s*d
            addForce(force, Npad, a.y, d * s);
            addForce(force, Npad, a.x, d * (-s));
        }
    }
    if (wantEnergy) blockEnergy(e, &counters->energy[E_BONDED]);
}

__global__ void __launch_bounds__(128) harmonicAngles(const float4* __restrict__ posq, const int4* __restrict__ atoms, const float2* __restrict__ params, int n,
                                                       BoxD box, int periodic, long long* force, int Npad, DevCounters* counters, int wantForce, int wantEnergy) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    double e = 0;
    if (t < n) {
        const int4 a = atoms[t]; const float2 prm = params[t];   // (theta0, k), atom y is the vertex
        const D3 pb = loadPos(posq, a.y);
        D3 v0 = loadPos(posq, a.x) - pb, v1 = loadPos(posq, a.z) - pb;
        if (periodic) { imageDeltaD(v0.x, v0.y, v0.z, box); imageDeltaD(v1.x, v1.y, v1.z, box); }
        const D3 cp = cross(v0, v1);
        const double rp = fmax(sqrt(dot(cp, cp)), 1e-6), r0 = dot(v0, v0), r1 = dot(v1, v1);
        double c = dot(v0, v1) / sqrt(r0 * r1);
        c = fmin(1.0, fmax(-1.0, c));
        const double theta = acos(c), dTheta = theta - prm.x;
        e = 0.5 * prm.y * dTheta * dTheta;
        if (wantForce) {
            const double dEdT = prm.y * dTheta;
            const D3 fa = cross(v0, cp) * (-dEdT / (r0 * rp));   // -dE/dtheta * dtheta/da
            const D3 fc = cross(v1, cp) * (dEdT / (r1 * rp));
            addForce(force, Npad, a.x, fa);
            addForce(force, Npad, a.z, fc);
            addForce(force, Npad, a.y, (fa + fc) * (-1.0));
        }
    }
    if (wantEnergy) blockEnergy(e, &counters->energy[E_BONDED]);
}

__global__ void __launch_bounds__(128) periodicTorsions(const float4* __restrict__ posq, const int4* __restrict__ atoms, const float4* __restrict__ params, int n,
                                                         BoxD box, int periodic, long long* force, int Npad, DevCounters* counters, int wantForce, int wantEnergy) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    double e = 0;
    if (t < n) {
        const int4 a = atoms[t]; const float4 prm = params[t];   // (periodicity, phase, k)
        const D3 p1 = loadPos(posq, a.x), p2 = loadPos(posq, a.y), p3 = loadPos(posq, a.z), p4 = loadPos(posq, a.w);
        D3 b1 = p2 - p1, b2 = p3 - p2, b3 = p4 - p3;
        if (periodic) { imageDeltaD(b1.x, b1.y, b1.z, box); imageDeltaD(b2.x, b2.y, b2.z, box); imageDeltaD(b3.x, b3.y, b3.z, box); }
        const D3 n1 = cross(b1, b2), n2 = cross(b2, b3);
        const double lb2 = sqrt(dot(b2, b2));
        // IUPAC dihedral: phi = atan2(|b2| b1.(b2 x b3), (b1 x b2).(b2 x b3))
        const double phi = atan2(lb2 * dot(b1, n2), dot(n1, n2));
        const double arg = prm.x * phi - prm.y;
        e = prm.z * (1.0 + cos(arg));
        if (wantForce) {
            const double dEdPhi = -prm.z * prm.x * sin(arg);
            const double n1sq = fmax(dot(n1, n1), 1e-12), n2sq = fmax(dot(n2, n2), 1e-12);
            // dphi/dp1 = -|b2|/|n1|^2 n1 ; dphi/dp4 = |b2|/|n2|^2 n2  (Blondel & Karplus 1996)
            const D3 f1 = n1 * (dEdPhi * lb2 / n1sq);
            const D3 f4 = n2 * (-dEdPhi * lb2 / n2sq);
            const double u = dot(b1, b2) / (lb2 * lb2), w = dot(b3, b2) / (lb2 * lb2);
            // gradients of the inner atoms follow from translation/rotation invariance (checked against finite differences)
            const D3 F2 = f1 * (-u - 1.0) + f4 * w;
            const D3 F3 = f4 * (-w - 1.0) + f1 * u;
            addForce(force, Npad, a.x, f1); addForce(force, Npad, a.y, F2); addForce(force, Npad, a.z, F3); addForce(force, Npad, a.w, f4);
        }
    }
    if (wantEnergy) blockEnergy(e, &counters->energy[E_BONDED]);
}

void b200_bonded_upload(b200_ctx* c) {
    if (c->hasNonbonded) {
        std::vector<int2> ex;
        for (size_t i = 0; i + 1 < c->exclPairs.size(); i += 2) ex.push_back(make_int2(c->exclPairs[i], c->exclPairs[i + 1]));
        c->numExclPairs = (int)ex.size();
        dmalloc(c->dExclPairs, ex.size());
        if (!ex.empty()) CUDA_CHECK(cudaMemcpy(c->dExclPairs, ex.data(), sizeof(int2) * ex.size(), cudaMemcpyHostToDevice));
        std::vector<int2> xp; std::vector<float4> xq;
        for (size_t i = 0; i < c->excQQ.size(); i++) {
            xp.push_back(make_int2(c->excPairs[2 * i], c->excPairs[2 * i + 1]));
            xq.push_back(make_float4((float)(B200_ONE_4PI_EPS0 * c->excQQ[i]), (float)c->excSig[i], (float)(4.0 * c->excEps[i]), 0.f));   // nonbondedParameters.cc:63
        }
        c->numExceptions = (int)xp.size();
        dmalloc(c->dExcPairs, xp.size()); dmalloc(c->dExcParams, xq.size());
        if (!xp.empty()) {
            CUDA_CHECK(cudaMemcpy(c->dExcPairs, xp.data(), sizeof(int2) * xp.size(), cudaMemcpyHostToDevice));
            CUDA_CHECK(cudaMemcpy(c->dExcParams, xq.data(), sizeof(float4) * xq.size(), cudaMemcpyHostToDevice));
        }
    }
    {
        std::vector<int2> at; std::vector<float2> pr;
        for (size_t i = 0; i + 1 < c->hBondAtoms.size(); i += 2) { at.push_back(make_int2(c->hBondAtoms[i], c->hBondAtoms[i + 1])); pr.push_back(make_float2((float)c->hBondParams[i], (float)c->hBondParams[i + 1])); }
        c->numBonds = (int)at.size(); dmalloc(c->bondAtoms, at.size()); dmalloc(c->bondParams, pr.size());
        if (!at.empty()) { CUDA_CHECK(cudaMemcpy(c->bondAtoms, at.data(), sizeof(int2) * at.size(), cudaMemcpyHostToDevice)); CUDA_CHECK(cudaMemcpy(c->bondParams, pr.data(), sizeof(float2) * pr.size(), cudaMemcpyHostToDevice)); }
    }
    {
        std::vector<int4> at; std::vector<float2> pr;
        for (size_t i = 0; i + 2 < c->hAngleAtoms.size(); i += 3) { at.push_back(make_int4(c->hAngleAtoms[i], c->hAngleAtoms[i + 1], c->hAngleAtoms[i + 2], 0)); pr.push_back(make_float2((float)c->hAngleParams[2 * (i / 3)], (float)c->hAngleParams[2 * (i / 3) + 1])); }
        c->numAngles = (int)at.size(); dmalloc(c->angleAtoms, at.size()); dmalloc(c->angleParams, pr.size());
        if (!at.empty()) { CUDA_CHECK(cudaMemcpy(c->angleAtoms, at.data(), sizeof(int4) * at.size(), cudaMemcpyHostToDevice)); CUDA_CHECK(cudaMemcpy(c->angleParams, pr.data(), sizeof(float2) * pr.size(), cudaMemcpyHostToDevice)); }
    }
    {
        std::vector<int4> at; std::vector<float4> pr;
        for (size_t i = 0; i + 3 < c->hTorsionAtoms.size(); i += 4) {
            at.push_back(make_int4(c->hTorsionAtoms[i], c->hTorsionAtoms[i + 1], c->hTorsionAtoms[i + 2], c->hTorsionAtoms[i + 3]));
            pr.push_back(make_float4((float)c->hTorsionPeriod[i / 4], (float)c->hTorsionParams[2 * (i / 4)], (float)c->hTorsionParams[2 * (i / 4) + 1], 0.f));
        }
        c->numTorsions = (int)at.size(); dmalloc(c->torsionAtoms, at.size()); dmalloc(c->torsionParams, pr.size());
        if (!at.empty()) { CUDA_CHECK(cudaMemcpy(c->torsionAtoms, at.data(), sizeof(int4) * at.size(), cudaMemcpyHostToDevice)); CUDA_CHECK(cudaMemcpy(c->torsionParams, pr.data(), sizeof(float4) * pr.size(), cudaMemcpyHostToDevice)); }
    }
}

void b200_bonded_launch(b200_ctx* c, unsigned groups, bool forces, bool energy, cudaStream_t s) {
    ScopedTimer tm(c, B200_T_BONDED, s);
    const int wf = forces ? 1 : 0, we = energy ? 1 : 0;
    // with a communicator the excluded-pair corrections are split over the ranks like the direct-space rows (the energies are
    // all-reduced afterwards); everything else in this file runs on rank 0
    const bool split = c->commRanks > 1;
    // (the erf correction of excluded pairs exists only where a reciprocal sum includes them: the Ewald family)
    if (c->hasNonbonded && c->nb.method >= B200_NB_EWALD && (groups & B200_GROUP_DIRECT) && c->numExclPairs > 0) {
        // slices start at multiples of the block size, so every block sums the same pairs as on one rank (bit-identical energy)
        const int nblk = (c->numExclPairs + 127) / 128;
        const int begin = split ? 128 * (int)llround(nblk * c->partShareBegin) : 0;
        const int end = split ? std::min(c->numExclPairs, 128 * (int)llround(nblk * c->partShareEnd)) : c->numExclPairs;
        if (end > begin && (split || c->partRank == 0)) {
            exclusionCorrection<<<(end - begin + 127) / 128, 128, 0, s>>>(c->posq, c->ljpme ? c->sigEps : nullptr, c->dExclPairs + begin, end - begin, c->boxD, c->nb.ewald_alpha, c->nb.ljpme_alpha, c->nb.exceptions_use_periodic, c->force, c->Npad, c->counters, wf, we);
            LAUNCH_COUNT(c, 1);
        }
    }
    if (c->partRank != 0) { CUDA_CHECK(cudaGetLastError()); return; }
    if (c->hasNonbonded && (groups & B200_GROUP_DIRECT)) {
        if (c->numExceptions > 0) { exceptions14<<<(c->numExceptions + 127) / 128, 128, 0, s>>>(c->posq, c->dExcPairs, c->dExcParams, c->numExceptions, c->boxD, c->nb.exceptions_use_periodic, c->force, c->Npad, c->counters, wf, we); LAUNCH_COUNT(c, 1); }
    }
    {
        // every term of a kind, or the ranges selected by b200_bonded_set_active (forces of one kind in different force groups)
        auto ranges = [&](int kind, int total, auto launch) {
            const std::vector<int>& r = c->bondedActive[kind];
            if (r.empty()) { if (total > 0) { launch(0, total); LAUNCH_COUNT(c, 1); } return; }
            for (size_t k = 0; k + 1 < r.size(); k += 2) if (r[k + 1] > 0) { launch(r[k], r[k + 1]); LAUNCH_COUNT(c, 1); }
        };
        if (groups & B200_GROUP_BONDS) ranges(0, c->numBonds, [&](int first, int n) {
            harmonicBonds<<<(n + 127) / 128, 128, 0, s>>>(c->posq, c->bondAtoms + first, c->bondParams + first, n, c->boxD, c->bondedPeriodic[0], c->force, c->Npad, c->counters, wf, we); });
        if (groups & B200_GROUP_ANGLES) ranges(1, c->numAngles, [&](int first, int n) {
            harmonicAngles<<<(n + 127) / 128, 128, 0, s>>>(c->posq, c->angleAtoms + first, c->angleParams + first, n, c->boxD, c->bondedPeriodic[1], c->force, c->Npad, c->counters, wf, we); });
        if (groups & B200_GROUP_TORSIONS) ranges(2, c->numTorsions, [&](int first, int n) {
            periodicTorsions<<<(n + 127) / 128, 128, 0, s>>>(c->posq, c->torsionAtoms + first, c->torsionParams + first, n, c->boxD, c->bondedPeriodic[2], c->force, c->Npad, c->counters, wf, we); });
    }
    CUDA_CHECK(cudaGetLastError());
}
