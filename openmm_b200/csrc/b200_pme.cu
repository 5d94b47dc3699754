// b200_pme.cu -- PME reciprocal space: B-spline charge spreading, reciprocal convolution (+energy), force interpolation.
//
// Replaces gridSpreadCharge / finishSpreadCharge / reciprocalConvolution / gridEvaluateEnergy / gridInterpolateForce
// (platforms/common/src/kernels/pme.cc:23-356) and the set-up in CommonCalcNonbondedForceKernel::commonInitialize
// (platforms/common/src/CommonCalcNonbondedForce.cpp:482-558, B-spline moduli).  The FFTs are in b200_fft.cu.
// Differences: atoms are already in Hilbert order, so no per-step sort by grid cell is needed (pme.cc:1-21 +
// CudaSort are dropped); charges are always accumulated in 64-bit fixed point so the result is bit-reproducible
// (the reference's USE_FIXED_POINT_CHARGE_SPREADING, pme.cc:106-107); the fixed-point -> float conversion is
// folded into the first FFT pass; the Coulomb constant is applied once, in the convolution.
#include "b200_internal.cuh"
#include <cmath>
#include <vector>

#define PME_ORDER 5

struct PmeArgs {
    int nx, ny, nz, nzc;
    double r0x, r1x, r1y, r2x, r2y, r2z;   // reciprocal box (rows as in CommonCalcNonbondedForce.cpp:950-957)
    float alphaFactor;                      // pi^2/alpha^2
    float scale;                            // ONE_4PI_EPS0/(pi V)
};

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

__device__ __forceinline__ void gridCoords(const float4 p, const PmeArgs& a, int& ix, int& iy, int& iz, float& ux, float& uy, float& uz) {
    // fractional coordinates in double: the only error left is the float rounding of the stored position
    const double x = p.x, y = p.y, z = p.z;
    double fa = x * a.r0x + y * a.r1x + z * a.r2x, fb = y * a.r1y + z * a.r2y, fc = z * a.r2z;
    fa = (fa - floor(fa)) * a.nx; fb = (fb - floor(fb)) * a.ny; fc = (fc - floor(fc)) * a.nz;
    ix = (int)fa; iy = (int)fb; iz = (int)fc;
    ux = (float)(fa - ix); uy = (float)(fb - iy); uz = (float)(fc - iz);
    ix = ix >= a.nx ? ix - a.nx : ix; iy = iy >= a.ny ? iy - a.ny : iy; iz = iz >= a.nz ? iz - a.nz : iz;
}

__global__ void pmeSpread(const float4* __restrict__ posq, int N, PmeArgs a, long long* __restrict__ grid) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float4 p = posq[i];
    if (p.w == 0.f) return;
    int ix, iy, iz; float ux, uy, uz;
    gridCoords(p, a, ix, iy, iz, ux, uy, uz);
    float wx[5], wy[5], wz[5], d[5];
    bspline5(ux, wx, d); bspline5(uy, wy, d); bspline5(uz, wz, d);
    int zi[5];
#pragma unroll
    for (int k = 0; k < 5; k++) { int z = iz + k; zi[k] = z >= a.nz ? z - a.nz : z; }
#pragma unroll
    for (int jx = 0; jx < 5; jx++) {
        int x = ix + jx; x = x >= a.nx ? x - a.nx : x;
        const float qx = p.w * wx[jx];
#pragma unroll
        for (int jy = 0; jy < 5; jy++) {
            int y = iy + jy; y = y >= a.ny ? y - a.ny : y;
            const float qxy = qx * wy[jy];
            long long* row = grid + ((size_t)x * a.ny + y) * a.nz;
#pragma unroll
            for (int jz = 0; jz < 5; jz++) atomicAddFixed(&row[zi[jz]], (long long)((double)(qxy * wz[jz]) * B200_GRID_SCALE));
        }
    }
}

// reciprocal convolution (pme.cc:137-189) with the energy sum (pme.cc:191-253) folded in
template <bool ENERGY>
__global__ void pmeConvolve(float2* __restrict__ grid, PmeArgs a, const float* __restrict__ modX, const float* __restrict__ modY, const float* __restrict__ modZ, DevCounters* counters) {
    const size_t total = (size_t)a.nx * a.ny * a.nzc;
    float e = 0.f;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int kz = (int)(idx % a.nzc); const size_t r = idx / a.nzc; const int ky = (int)(r % a.ny), kx = (int)(r / a.ny);
        const int mx = kx < (a.nx + 1) / 2 ? kx : kx - a.nx, my = ky < (a.ny + 1) / 2 ? ky : ky - a.ny, mz = kz;   // kz <= nz/2
        const float mhx = mx * (float)a.r0x, mhy = mx * (float)a.r1x + my * (float)a.r1y, mhz = mx * (float)a.r2x + my * (float)a.r2y + mz * (float)a.r2z;
        const float m2 = mhx * mhx + mhy * mhy + mhz * mhz;
        const float denom = m2 * modX[kx] * modY[ky] * modZ[kz];
        float eterm = a.scale * __expf(-a.alphaFactor * m2) / denom;
        if (idx == 0) eterm = 0.f;
        const float2 g = grid[idx];
        if (ENERGY) {
            const float wgt = (kz == 0 || (2 * kz == a.nz)) ? 1.f : 2.f;   // Hermitian mirror of the missing half
            e += wgt * eterm * (g.x * g.x + g.y * g.y);
        }
        grid[idx] = make_float2(g.x * eterm, g.y * eterm);
    }
    if (ENERGY) {
        // fixed-order block reduction, then an integer atomic: bit-reproducible for a fixed launch shape
        __shared__ float red[256];
        red[threadIdx.x] = e;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
        if (threadIdx.x == 0) atomicAddFixed(&counters->energy[E_RECIP], toFixedD(0.5 * (double)red[0]));
    }
}

// force interpolation (pme.cc:255-356)
__global__ void pmeInterpolate(const float4* __restrict__ posq, int N, int Npad, PmeArgs a, const float* __restrict__ phi, long long* __restrict__ force) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float4 p = posq[i];
    if (p.w == 0.f) return;
    int ix, iy, iz; float ux, uy, uz;
    gridCoords(p, a, ix, iy, iz, ux, uy, uz);
    float wx[5], wy[5], wz[5], dx[5], dy[5], dz[5];
    bspline5(ux, wx, dx); bspline5(uy, wy, dy); bspline5(uz, wz, dz);
    int zi[5];
#pragma unroll
    for (int k = 0; k < 5; k++) { int z = iz + k; zi[k] = z >= a.nz ? z - a.nz : z; }
    float gx = 0.f, gy = 0.f, gz = 0.f;
#pragma unroll
    for (int jx = 0; jx < 5; jx++) {
        int x = ix + jx; x = x >= a.nx ? x - a.nx : x;
#pragma unroll
        for (int jy = 0; jy < 5; jy++) {
            int y = iy + jy; y = y >= a.ny ? y - a.ny : y;
            const float* row = phi + ((size_t)x * a.ny + y) * a.nz;
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int jz = 0; jz < 5; jz++) { const float v = row[zi[jz]]; s0 = fmaf(wz[jz], v, s0); s1 = fmaf(dz[jz], v, s1); }
            gx = fmaf(dx[jx] * wy[jy], s0, gx);
            gy = fmaf(wx[jx] * dy[jy], s0, gy);
            gz = fmaf(wx[jx] * wy[jy], s1, gz);
        }
    }
    const float q = p.w;
    const float tx = gx * a.nx, ty = gy * a.ny, tz = gz * a.nz;
    const float fx = -q * (tx * (float)a.r0x);
    const float fy = -q * (tx * (float)a.r1x + ty * (float)a.r1y);
    const float fz = -q * (tx * (float)a.r2x + ty * (float)a.r2y + tz * (float)a.r2z);
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

void b200_pme_create(b200_ctx* c) {
    for (int k = 0; k < 3; k++) c->gridN[k] = c->nb.grid[k];
    const int nx = c->gridN[0], ny = c->gridN[1], nz = c->gridN[2];
    c->nzc = nz / 2 + 1;
    const size_t G = (size_t)nx * ny * nz, GC = (size_t)nx * ny * c->nzc;
    dmalloc(c->gridFix, G); dmalloc(c->gridReal, G); dmalloc(c->gridC, GC); dmalloc(c->gridC2, GC);
    for (int k = 0; k < 3; k++) {
        const std::vector<float> m = splineModuli(c->gridN[k]);
        dmalloc(c->moduli[k], m.size());
        CUDA_CHECK(cudaMemcpy(c->moduli[k], m.data(), sizeof(float) * m.size(), cudaMemcpyHostToDevice));
        b200_fft_make_plan(c->gridN[k], c->plan[k]);
        b200_fft_make_twiddle(c->gridN[k], c->twiddle[k], c->stream);
    }
}

void b200_pme_destroy(b200_ctx* c) {
    dfree(c->gridFix); dfree(c->gridReal); dfree(c->gridC); dfree(c->gridC2);
    for (int k = 0; k < 3; k++) { dfree(c->moduli[k]); dfree(c->twiddle[k]); }
}

void b200_pme_launch(b200_ctx* c, cudaStream_t s, bool forces, bool energy) {
    PmeArgs a;
    a.nx = c->gridN[0]; a.ny = c->gridN[1]; a.nz = c->gridN[2]; a.nzc = c->nzc;
    const BoxD& b = c->boxD;
    const double det = b.ax * b.by * b.cz, sc = 1.0 / det;
    a.r0x = b.by * b.cz * sc;
    a.r1x = -b.bx * b.cz * sc; a.r1y = b.ax * b.cz * sc;
    a.r2x = (b.bx * b.cy - b.by * b.cx) * sc; a.r2y = -b.ax * b.cy * sc; a.r2z = b.ax * b.by * sc;
    a.alphaFactor = (float)(M_PI * M_PI / (c->nb.ewald_alpha * c->nb.ewald_alpha));
    a.scale = (float)(B200_ONE_4PI_EPS0 / (M_PI * det));
    const size_t G = (size_t)a.nx * a.ny * a.nz;
    {
        ScopedTimer tm(c, B200_T_SPREAD, s);
        CUDA_CHECK(cudaMemsetAsync(c->gridFix, 0, sizeof(long long) * G, s));
        pmeSpread<<<(c->N + 127) / 128, 128, 0, s>>>(c->posq, c->N, a, c->gridFix);
    }
    { ScopedTimer tm(c, B200_T_FFT, s); b200_fft3d_forward(c, s); }
    {
        ScopedTimer tm(c, B200_T_CONVOLVE, s);
        const int grid = c->smCount * 8;
        if (energy) pmeConvolve<true><<<grid, 256, 0, s>>>(c->gridC, a, c->moduli[0], c->moduli[1], c->moduli[2], c->counters);
        else pmeConvolve<false><<<grid, 256, 0, s>>>(c->gridC, a, c->moduli[0], c->moduli[1], c->moduli[2], c->counters);
    }
    LAUNCH_COUNT(c, 2);
    if (forces) {
        { ScopedTimer tm(c, B200_T_FFT, s); b200_fft3d_inverse(c, s); }
        ScopedTimer tm(c, B200_T_INTERPOLATE, s);
        pmeInterpolate<<<(c->N + 127) / 128, 128, 0, s>>>(c->posq, c->N, c->Npad, a, c->gridReal, c->force);
        LAUNCH_COUNT(c, 1);
    }
    CUDA_CHECK(cudaGetLastError());
}
