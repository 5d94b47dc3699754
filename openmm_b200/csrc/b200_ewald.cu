// b200_ewald.cu -- NonbondedForce::Ewald: explicit reciprocal-space sum over k vectors.
//
// Replaces calculateEwaldCosSinSums / calculateEwaldForces (platforms/common/src/kernels/ewald.cc) for the one method of the
// Ewald family that does not use a mesh.  Same half-space enumeration as ReferenceLJCoulombIxn.cpp:272-368: kx in [0, kmax_x),
// ky and kz in (-kmax, kmax), the k = 0 vector and the mirror half left out, every remaining vector counted twice through the
// prefactor.  It exists so that systems and tests that ask for Ewald run; PME is the method to use for speed.  Two passes:
//   ewaldStructureFactors : one CTA per k vector, S(k) = sum_j q_j exp(i k.r_j) in double, plus that vector's energy
//   ewaldForces           : one thread per atom, loop over the k vectors
#include "b200_internal.cuh"
#include <cmath>

struct EwaldArgs { int kx, ky, kz, numK; double rbx, rby, rbz, factor, coeff; };

__device__ __forceinline__ bool ewaldVector(const EwaldArgs& a, int idx, double& kx, double& ky, double& kz) {
    // idx -> (rx, ry, rz) with rx in [0,kx), ry in (-ky,ky), rz in (-kz,kz); keep the half space rx > 0 || (rx == 0 && (ry > 0 || (ry == 0 && rz > 0)))
    const int ny = 2 * a.ky - 1, nz = 2 * a.kz - 1;
    const int rx = idx / (ny * nz), rem = idx - rx * ny * nz, ry = rem / nz - (a.ky - 1), rz = rem % nz - (a.kz - 1);
    if (!(rx > 0 || ry > 0 || (ry == 0 && rz > 0))) return false;
    kx = rx * a.rbx; ky = ry * a.rby; kz = rz * a.rbz;
    return true;
}

__global__ void __launch_bounds__(128) ewaldStructureFactors(const float4* __restrict__ posq, int N, EwaldArgs a, double2* __restrict__ S, DevCounters* counters, int wantEnergy) {
    double kx, ky, kz;
    const bool live = ewaldVector(a, blockIdx.x, kx, ky, kz);
    double cs = 0, ss = 0;
    if (live) for (int i = threadIdx.x; i < N; i += blockDim.x) {
        const float4 p = posq[i];
        double s, c; sincos(kx * p.x + ky * p.y + kz * p.z, &s, &c);
        cs += p.w * c; ss += p.w * s;
    }
    __shared__ double red[2][128];
    red[0][threadIdx.x] = cs; red[1][threadIdx.x] = ss;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) { if (threadIdx.x < o) { red[0][threadIdx.x] += red[0][threadIdx.x + o]; red[1][threadIdx.x] += red[1][threadIdx.x + o]; } __syncthreads(); }
    if (threadIdx.x == 0) {
        S[blockIdx.x] = make_double2(red[0][0], red[1][0]);
        if (live && wantEnergy) {
            const double k2 = kx * kx + ky * ky + kz * kz, ak = exp(k2 * a.factor) / k2;
            atomicAddFixed(&counters->energy[E_RECIP], toFixedD(a.coeff * ak * (red[0][0] * red[0][0] + red[1][0] * red[1][0])));
        }
    }
}

__global__ void __launch_bounds__(128) ewaldForces(const float4* __restrict__ posq, int N, int Npad, EwaldArgs a, const double2* __restrict__ S, long long* force) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float4 p = posq[i];
    double fx = 0, fy = 0, fz = 0;
    for (int idx = 0; idx < a.numK; idx++) {
        double kx, ky, kz;
        if (!ewaldVector(a, idx, kx, ky, kz)) continue;
        const double2 sk = S[idx];
        double s, c; sincos(kx * p.x + ky * p.y + kz * p.z, &s, &c);
        const double k2 = kx * kx + ky * ky + kz * kz, ak = exp(k2 * a.factor) / k2;
        const double f = 2.0 * a.coeff * ak * p.w * (sk.x * s - sk.y * c);
        fx += f * kx; fy += f * ky; fz += f * kz;
    }
    atomicAddFixed(&force[i], toFixedD(fx)); atomicAddFixed(&force[i + Npad], toFixedD(fy)); atomicAddFixed(&force[i + 2 * (size_t)Npad], toFixedD(fz));
}

void b200_ewald_launch(b200_ctx* c, cudaStream_t s, bool forces, bool energy) {
    ScopedTimer tm(c, B200_T_FFT, s);
    EwaldArgs a;
    a.kx = c->nb.ewald_kmax[0]; a.ky = c->nb.ewald_kmax[1]; a.kz = c->nb.ewald_kmax[2];
    a.numK = a.kx * (2 * a.ky - 1) * (2 * a.kz - 1);
    const double* b = c->boxHost;
    a.rbx = 2.0 * M_PI / b[0]; a.rby = 2.0 * M_PI / b[4]; a.rbz = 2.0 * M_PI / b[8];
    a.factor = -1.0 / (4.0 * c->nb.ewald_alpha * c->nb.ewald_alpha);
    a.coeff = B200_ONE_4PI_EPS0 * 4.0 * M_PI / (b[0] * b[4] * b[8]);
    if (c->ewaldNumK < a.numK) { dfree(c->ewaldS); dmalloc(c->ewaldS, (size_t)a.numK); c->ewaldNumK = a.numK; }
    ewaldStructureFactors<<<a.numK, 128, 0, s>>>(c->posq, c->N, a, c->ewaldS, c->counters, energy ? 1 : 0);
    if (forces) ewaldForces<<<(c->N + 127) / 128, 128, 0, s>>>(c->posq, c->N, c->Npad, a, c->ewaldS, c->force);
    LAUNCH_COUNT(c, forces ? 2 : 1);
    CUDA_CHECK(cudaGetLastError());
}
