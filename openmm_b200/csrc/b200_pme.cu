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
// their stencils cover a compact box of grid points; when that box fits the shared-memory tile the CTA accumulates all
// 125 contributions per atom there (native 32-bit shared atomics) and then adds the box to the global grid with one
// coalesced integer atomic per touched point.  Boxes that do not fit (rare: atoms that drifted far apart between two
// re-sorts) fall back to direct global atomics.  All sums are integer, so the grid is bit-reproducible.
#define CHUNK 128
#define TILE_MAX 8192
struct ChunkBox { int ref[3]; int lo[3]; int ext[3]; int vol; };

// t -> (x, y, z) in a box of extents (., ey, ez) without integer division (exact for the small integers involved)
__device__ __forceinline__ void unflatten(int t, int ey, int ez, float invEy, float invEz, int& x, int& y, int& z) {
    const int r = (int)(((float)t + 0.5f) * invEz);
    z = t - r * ez;
    x = (int)(((float)r + 0.5f) * invEy);
    y = r - x * ey;
}
__device__ __forceinline__ int wrapDiff(int d, int n) { d += n / 2; d %= n; if (d < 0) d += n; return d - n / 2; }
__device__ __forceinline__ int wrapIdx(int i, int n) { i %= n; return i < 0 ? i + n : i; }

// bounding box (in grid points, relative to the first atom's base index) of the stencils of the chunk's atoms
__device__ __forceinline__ void chunkBox(bool valid, int ix, int iy, int iz, const PmeArgs& a, ChunkBox& box, int* sh /*[10]*/, int& dx, int& dy, int& dz) {
    if (threadIdx.x == 0) { sh[0] = ix; sh[1] = iy; sh[2] = iz; sh[3] = sh[4] = sh[5] = 1 << 30; sh[6] = sh[7] = sh[8] = -(1 << 30); }
    __syncthreads();
    box.ref[0] = sh[0]; box.ref[1] = sh[1]; box.ref[2] = sh[2];
    dx = wrapDiff(ix - box.ref[0], a.nx); dy = wrapDiff(iy - box.ref[1], a.ny); dz = wrapDiff(iz - box.ref[2], a.nz);
    if (valid) {
        atomicMin(&sh[3], dx); atomicMin(&sh[4], dy); atomicMin(&sh[5], dz);
        atomicMax(&sh[6], dx); atomicMax(&sh[7], dy); atomicMax(&sh[8], dz);
    }
    __syncthreads();
    for (int k = 0; k < 3; k++) { box.lo[k] = sh[3 + k]; box.ext[k] = sh[6 + k] - sh[3 + k] + PME_ORDER; }
    const int n[3] = {a.nx, a.ny, a.nz};
    bool fits = sh[6] >= sh[3];
    for (int k = 0; k < 3; k++) fits = fits && box.ext[k] <= n[k];      // a box wider than the grid would alias onto itself
    const long long vol = (long long)box.ext[0] * box.ext[1] * box.ext[2];
    box.vol = (fits && vol <= TILE_MAX) ? (int)vol : 0;
}

// charge spreading (gridSpreadCharge, pme.cc:23-121)
__global__ void __launch_bounds__(CHUNK) pmeSpreadChunks(const float4* __restrict__ posq, const float2* __restrict__ sigEps, int N, PmeArgs a, int* __restrict__ grid, float fixScale) {
    __shared__ int tile[TILE_MAX];
    __shared__ int sh[10];
    const int i = blockIdx.x * CHUNK + threadIdx.x;
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < N) { p = posq[i]; if (sigEps) p.w = dispersionCoefficient(sigEps[i]); }
    int ix = 0, iy = 0, iz = 0; float ux = 0.f, uy = 0.f, uz = 0.f;
    gridCoords(p, a, ix, iy, iz, ux, uy, uz);
    const bool valid = i < N && p.w != 0.f;
    ChunkBox box; int dx, dy, dz;
    chunkBox(valid, ix, iy, iz, a, box, sh, dx, dy, dz);
    float wx[5], wy[5], wz[5], d[5];
    bspline5(ux, wx, d); bspline5(uy, wy, d); bspline5(uz, wz, d);
    const float q = p.w * fixScale;
    if (box.vol > 0) {
        for (int t = threadIdx.x; t < box.vol; t += CHUNK) tile[t] = 0;
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
                    for (int jz = 0; jz < 5; jz++) atomicAdd(&row[jz], __float2int_rn(vxy * wz[jz]));
                }
            }
        }
        __syncthreads();
        const int ey = box.ext[1], ez = box.ext[2];
        const int ox = wrapIdx(box.ref[0] + box.lo[0], a.nx), oy = wrapIdx(box.ref[1] + box.lo[1], a.ny), oz = wrapIdx(box.ref[2] + box.lo[2], a.nz);
        const float invEy = 1.0f / ey, invEz = 1.0f / ez;
        for (int t = threadIdx.x; t < box.vol; t += CHUNK) {
            const int v = tile[t];
            if (v == 0) continue;
            int x, y, z;
            unflatten(t, ey, ez, invEy, invEz, x, y, z);
            int gx = ox + x; gx += gx < 0 ? a.nx : 0; gx -= gx >= a.nx ? a.nx : 0;
            int gy = oy + y; gy += gy < 0 ? a.ny : 0; gy -= gy >= a.ny ? a.ny : 0;
            int gz = oz + z; gz += gz < 0 ? a.nz : 0; gz -= gz >= a.nz ? a.nz : 0;
            atomicAdd(&grid[((size_t)gx * a.ny + gy) * a.nz + gz], v);
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
                for (int jz = 0; jz < 5; jz++) atomicAdd(&row[zi[jz]], __float2int_rn(vxy * wz[jz]));
            }
        }
    }
}

// (the reciprocal convolution and energy sum, pme.cc:137-253, are fused into the x pass of the FFT: fftXConvolve in b200_fft.cu)

// force interpolation (gridInterpolateForce, pme.cc:255-356): same chunking; the box of potential values the chunk's atoms
// read is staged in shared memory once (coalesced), then every atom gathers its 125 values from there
__global__ void __launch_bounds__(CHUNK) pmeInterpolateChunks(const float4* __restrict__ posq, const float2* __restrict__ sigEps, int N, int Npad, PmeArgs a, const float* __restrict__ phi, long long* __restrict__ force) {
    __shared__ float tile[TILE_MAX];
    __shared__ int sh[10];
    const int i = blockIdx.x * CHUNK + threadIdx.x;
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < N) { p = posq[i]; if (sigEps) p.w = dispersionCoefficient(sigEps[i]); }
    int ix = 0, iy = 0, iz = 0; float ux = 0.f, uy = 0.f, uz = 0.f;
    gridCoords(p, a, ix, iy, iz, ux, uy, uz);
    const bool valid = i < N && p.w != 0.f;
    ChunkBox box; int dx, dy, dz;
    chunkBox(valid, ix, iy, iz, a, box, sh, dx, dy, dz);
    float wx[5], wy[5], wz[5], ex[5], ey5[5], ez5[5];
    bspline5(ux, wx, ex); bspline5(uy, wy, ey5); bspline5(uz, wz, ez5);
    float gx = 0.f, gy = 0.f, gz = 0.f;
    if (box.vol > 0) {
        const int ey = box.ext[1], ez = box.ext[2];
        const int ox = wrapIdx(box.ref[0] + box.lo[0], a.nx), oy = wrapIdx(box.ref[1] + box.lo[1], a.ny), oz = wrapIdx(box.ref[2] + box.lo[2], a.nz);
        const float invEy = 1.0f / ey, invEz = 1.0f / ez;
        for (int t = threadIdx.x; t < box.vol; t += CHUNK) {
            int x, y, z;
            unflatten(t, ey, ez, invEy, invEz, x, y, z);
            int gx = ox + x; gx += gx < 0 ? a.nx : 0; gx -= gx >= a.nx ? a.nx : 0;
            int gy = oy + y; gy += gy < 0 ? a.ny : 0; gy -= gy >= a.ny ? a.ny : 0;
            int gz = oz + z; gz += gz < 0 ? a.nz : 0; gz -= gz >= a.nz ? a.nz : 0;
            tile[t] = phi[((size_t)gx * a.ny + gy) * a.nz + gz];
        }
        __syncthreads();
        if (valid) {
            const int base = ((dx - box.lo[0]) * ey + (dy - box.lo[1])) * ez + (dz - box.lo[2]);
#pragma unroll
            for (int jx = 0; jx < 5; jx++) {
#pragma unroll
                for (int jy = 0; jy < 5; jy++) {
                    const float* row = &tile[base + (jx * ey + jy) * ez];
                    float s0 = 0.f, s1 = 0.f;
#pragma unroll
                    for (int jz = 0; jz < 5; jz++) { const float v = row[jz]; s0 = fmaf(wz[jz], v, s0); s1 = fmaf(ez5[jz], v, s1); }
                    gx = fmaf(ex[jx] * wy[jy], s0, gx);
                    gy = fmaf(wx[jx] * ey5[jy], s0, gy);
                    gz = fmaf(wx[jx] * wy[jy], s1, gz);
                }
            }
        }
    } else if (valid) {
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
    if (!valid) return;
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

static void meshCreate(b200_ctx* c, PmeMesh& m, const int* grid, double alpha, bool dispersion, double maxCoef) {
    for (int k = 0; k < 3; k++) m.gridN[k] = grid[k];
    const int nx = m.gridN[0], ny = m.gridN[1], nz = m.gridN[2];
    m.nzc = nz / 2 + 1; m.alpha = alpha; m.dispersion = dispersion;
    const size_t G = (size_t)nx * ny * nz, GC = (size_t)nx * ny * m.nzc;
    dmalloc(m.gridI, G); dmalloc(m.gridReal, G); dmalloc(m.gridC, GC);
    {   // fixed-point scale: |coefficient| * scale <= 1.5 * 2^26 ~ 2^26.6, leaving a factor >= 21 of head-room for the sum at one point
        int bits = 26; double lim = 1.5;
        while (maxCoef > lim && bits > 12) { bits--; lim *= 2; }
        while (maxCoef > 0 && maxCoef <= 0.5 * lim && bits < 36) { bits++; lim *= 0.5; }
        if (dispersion) bits -= 2;   // every coefficient is positive: sums do not cancel, keep more head-room
        m.spreadScale = (float)std::ldexp(1.0, bits);
    }
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

void b200_pme_create(b200_ctx* c) {
    double maxQ = 0; for (double q : c->charge) maxQ = std::max(maxQ, std::fabs(q));
    meshCreate(c, c->mesh[0], c->nb.grid, c->nb.ewald_alpha, false, std::max(maxQ, 0.75 + 1e-9));
    if (c->ljpme) {
        double maxC6 = 0;
        for (int i = 0; i < c->N; i++) maxC6 = std::max(maxC6, 8.0 * pow(0.5 * c->sigma[i], 3.0) * 2.0 * sqrt(c->epsilon[i]));
        meshCreate(c, c->mesh[1], c->nb.ljpme_grid, c->nb.ljpme_alpha, true, maxC6);
    }
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
        pmeInterpolateChunks<<<chunks, CHUNK, 0, s>>>(c->posq, coef, c->N, c->Npad, a, m.gridReal, c->force);
        LAUNCH_COUNT(c, 1);
    }
    CUDA_CHECK(cudaGetLastError());
}

void b200_pme_launch(b200_ctx* c, cudaStream_t s, bool forces, bool energy) {
    meshLaunch(c, c->mesh[0], s, forces, energy);
    if (c->ljpme) meshLaunch(c, c->mesh[1], s, forces, energy);
}
