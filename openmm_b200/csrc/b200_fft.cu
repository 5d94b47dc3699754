// b200_fft.cu -- hand-written batched 3D real-to-complex / complex-to-real FFT for 2,3,5,7-smooth sizes.
//
// Replaces the cuFFT calls of CudaFFT3D::execFFT (platforms/cuda/src/CudaFFT3D.cpp:30-101).  Layout is cuFFT's:
// real grid [nx][ny][nz] (z fastest), half-complex grid [nx][ny][nz/2+1]; both transforms are unnormalised.
// This file holds the host drivers and the GENERIC passes: each 1D pass loads a batch of lines into shared memory, runs a
// Stockham auto-sort mixed-radix FFT (radix 4/2 specialised, 3/5/7 through the twiddle table) entirely in shared memory and
// writes the lines back, so a 3D transform is three passes over the (L2-resident) grid.  The drivers use the register-resident
// two-stage passes of b200_fft2.cu whenever every dimension has a two-stage plan (b200_fft2_supported) and these generic
// passes otherwise.  The z pass of the forward transform reads the 32-bit fixed-point charge grid directly (folding
// finishSpreadCharge, pme.cc:125-133, into the FFT).
#include "b200_internal.cuh"
#include <cmath>
#include <map>
#include <vector>

#define FFT_THREADS 256

void b200_fft_make_plan(int n, FFTPlan1D& plan) {
    plan.n = n; plan.nstages = 0;
    int m = n;
    while (m % 4 == 0) { plan.radix[plan.nstages++] = 4; m /= 4; }
    for (int f : {2, 3, 5, 7}) while (m % f == 0) { plan.radix[plan.nstages++] = f; m /= f; }
    if (m != 1) throw B200Error(B200_ERR_INVALID, "FFT size must factor into 2,3,5,7");
}

void b200_fft_make_twiddle(int n, float2*& dTw, cudaStream_t s) {
    std::vector<float2> w(n);
    for (int k = 0; k < n; k++) { const double a = -2.0 * M_PI * k / n; w[k] = make_float2((float)cos(a), (float)sin(a)); }
    dmalloc(dTw, (size_t)n);
    CUDA_CHECK(cudaMemcpy(dTw, w.data(), sizeof(float2) * n, cudaMemcpyHostToDevice));
}

__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }

// One Stockham stage over `batch` (a power of two) lines held as buf[j*pitch + l].  W = exp(-2 pi i k/n) table (conjugated
// for the inverse).  Odd radices use the conjugate symmetry of the roots: with a_t = x_t + x_{R-t}, b_t = x_t - x_{R-t},
//   y_q = x_0 + sum_t (a_t cos_qt + i b_t sin_qt),   y_{R-q} = x_0 + sum_t (a_t cos_qt - i b_t sin_qt),  t = 1..(R-1)/2
// which needs about a quarter of the multiplications of the plain R x R product.
template <bool INVERSE, int R>
__device__ __forceinline__ void fftStage(const float2* __restrict__ in, float2* __restrict__ out, int n, int Ns, int batch, int pitch, const float2* __restrict__ W) {
    const int nOverR = n / R;
    const int twStep = n / (Ns * R);
    const int logB = __ffs(batch) - 1;
    const float invNs = 1.0f / (float)Ns;
    float2 root[R];   // exp(-+2 pi i m/R)
#pragma unroll
    for (int m = 0; m < R; m++) { root[m] = W[m * nOverR]; if (INVERSE) root[m].y = -root[m].y; }
    for (int idx = threadIdx.x; idx < (nOverR << logB); idx += blockDim.x) {
        const int l = idx & (batch - 1), j = idx >> logB;
        const int k = j - (int)(((float)j + 0.5f) * invNs) * Ns;   // j % Ns without an integer division (exact for j < 2^22)
        const int j0 = (j - k) * R + k;
        float2 x[R];
#pragma unroll
        for (int t = 0; t < R; t++) {
            float2 v = in[(j + t * nOverR) * pitch + l];
            if (t > 0 && k > 0) { float2 w = W[t * k * twStep]; if (INVERSE) w.y = -w.y; v = cmul(v, w); }
            x[t] = v;
        }
        if (R == 2) {
            out[j0 * pitch + l] = cadd(x[0], x[1]);
            out[(j0 + Ns) * pitch + l] = csub(x[0], x[1]);
        } else if (R == 4) {
            const float2 a = cadd(x[0], x[2]), b = csub(x[0], x[2]), c = cadd(x[1], x[3]), d = csub(x[1], x[3]);
            const float2 di = INVERSE ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);   // -i d (forward), +i d (inverse)
            out[j0 * pitch + l] = cadd(a, c);
            out[(j0 + Ns) * pitch + l] = cadd(b, di);
            out[(j0 + 2 * Ns) * pitch + l] = csub(a, c);
            out[(j0 + 3 * Ns) * pitch + l] = csub(b, di);
        } else {
            constexpr int H = (R - 1) / 2;
            float2 sa[H + 1], sb[H + 1];
            float2 y0 = x[0];
#pragma unroll
            for (int t = 1; t <= H; t++) { sa[t] = cadd(x[t], x[R - t]); sb[t] = csub(x[t], x[R - t]); y0 = cadd(y0, sa[t]); }
            out[j0 * pitch + l] = y0;
#pragma unroll
            for (int q = 1; q <= H; q++) {
                float2 C = x[0], D = make_float2(0.f, 0.f);
#pragma unroll
                for (int t = 1; t <= H; t++) {
                    const float2 w = root[(q * t) % R];
                    C.x = fmaf(sa[t].x, w.x, C.x); C.y = fmaf(sa[t].y, w.x, C.y);
                    D.x = fmaf(sb[t].x, w.y, D.x); D.y = fmaf(sb[t].y, w.y, D.y);
                }
                out[(j0 + q * Ns) * pitch + l] = make_float2(C.x - D.y, C.y + D.x);          // C + i D
                out[(j0 + (R - q) * Ns) * pitch + l] = make_float2(C.x + D.y, C.y - D.x);    // C - i D
            }
        }
    }
}

template <bool INVERSE>
__device__ float2* fftLinesInSmem(float2* a, float2* b, const FFTPlan1D& plan, int batch, int pitch, const float2* W) {
    int Ns = 1;
    for (int s = 0; s < plan.nstages; s++) {
        __syncthreads();
        switch (plan.radix[s]) {
            case 2: fftStage<INVERSE, 2>(a, b, plan.n, Ns, batch, pitch, W); break;
            case 3: fftStage<INVERSE, 3>(a, b, plan.n, Ns, batch, pitch, W); break;
            case 4: fftStage<INVERSE, 4>(a, b, plan.n, Ns, batch, pitch, W); break;
            case 5: fftStage<INVERSE, 5>(a, b, plan.n, Ns, batch, pitch, W); break;
            default: fftStage<INVERSE, 7>(a, b, plan.n, Ns, batch, pitch, W); break;
        }
        Ns *= plan.radix[s];
        float2* t = a; a = b; b = t;
    }
    __syncthreads();
    return a;   // result
}

// z pass, forward: real (float or fixed point) lines -> nzc complex
__global__ void __launch_bounds__(FFT_THREADS) fftZForward(const float* __restrict__ inReal, const int* __restrict__ inFix, float fixScaleInv, float2* __restrict__ out,
                                                            int numLines, FFTPlan1D plan, int batch, int pitch, const float2* __restrict__ W) {
    extern __shared__ float2 smem[];
    const int n = plan.n, nzc = n / 2 + 1;
    float2* a = smem; float2* b = smem + (size_t)n * pitch;
    const int line0 = blockIdx.x * batch;
    const int nl = min(batch, numLines - line0);
    const float invN = 1.0f / (float)n, invNzc = 1.0f / (float)nzc;
    for (int idx = threadIdx.x; idx < batch * n; idx += blockDim.x) {
        const int l = (int)(((float)idx + 0.5f) * invN), j = idx - l * n;
        float v = 0.f;
        if (l < nl) {
            const size_t g = (size_t)(line0 + l) * n + j;
            v = inFix ? (float)inFix[g] * fixScaleInv : inReal[g];
        }
        a[j * pitch + l] = make_float2(v, 0.f);
    }
    float2* r = fftLinesInSmem<false>(a, b, plan, batch, pitch, W);
    for (int idx = threadIdx.x; idx < batch * nzc; idx += blockDim.x) {
        const int l = (int)(((float)idx + 0.5f) * invNzc), j = idx - l * nzc;
        if (l < nl) out[(size_t)(line0 + l) * nzc + j] = r[j * pitch + l];
    }
}

// z pass, inverse: nzc complex lines -> n real
__global__ void __launch_bounds__(FFT_THREADS) fftZInverse(const float2* __restrict__ in, float* __restrict__ out, int numLines, FFTPlan1D plan, int batch, int pitch, const float2* __restrict__ W) {
    extern __shared__ float2 smem[];
    const int n = plan.n, nzc = n / 2 + 1;
    float2* a = smem; float2* b = smem + (size_t)n * pitch;
    const int line0 = blockIdx.x * batch;
    const int nl = min(batch, numLines - line0);
    const float invN = 1.0f / (float)n, invNzc = 1.0f / (float)nzc;
    for (int idx = threadIdx.x; idx < batch * nzc; idx += blockDim.x) {
        const int l = (int)(((float)idx + 0.5f) * invNzc), j = idx - l * nzc;
        float2 v = make_float2(0.f, 0.f);
        if (l < nl) v = in[(size_t)(line0 + l) * nzc + j];
        a[j * pitch + l] = v;
        if (j > 0 && j < n - j) a[(n - j) * pitch + l] = make_float2(v.x, -v.y);   // Hermitian half
    }
    float2* r = fftLinesInSmem<true>(a, b, plan, batch, pitch, W);
    for (int idx = threadIdx.x; idx < batch * n; idx += blockDim.x) {
        const int l = (int)(((float)idx + 0.5f) * invN), j = idx - l * n;
        if (l < nl) out[(size_t)(line0 + l) * n + j] = r[j * pitch + l].x;
    }
}

// strided complex pass (y or x direction), in place.  Lines are addressed as outer*outerStride + inner + j*stride.
template <bool INVERSE>
__global__ void __launch_bounds__(FFT_THREADS) fftStrided(float2* __restrict__ data, int innerCount, size_t outerStride, size_t stride,
                                                           FFTPlan1D plan, int batch, int pitch, const float2* __restrict__ W) {
    extern __shared__ float2 smem[];
    const int n = plan.n;
    const int logB = __ffs(batch) - 1;
    float2* a = smem; float2* b = smem + (size_t)n * pitch;
    const int inner0 = blockIdx.x * batch;
    const int nl = min(batch, innerCount - inner0);
    float2* base = data + (size_t)blockIdx.y * outerStride + inner0;
    for (int idx = threadIdx.x; idx < batch * n; idx += blockDim.x) {
        const int j = idx >> logB, l = idx & (batch - 1);
        a[j * pitch + l] = l < nl ? base[(size_t)j * stride + l] : make_float2(0.f, 0.f);
    }
    float2* r = fftLinesInSmem<INVERSE>(a, b, plan, batch, pitch, W);
    for (int idx = threadIdx.x; idx < batch * n; idx += blockDim.x) {
        const int j = idx >> logB, l = idx & (batch - 1);
        if (l < nl) base[(size_t)j * stride + l] = r[j * pitch + l];
    }
}

// x pass fused with the reciprocal convolution: forward FFT along x, multiply by the influence function (and sum the energy),
// inverse FFT along x, all on lines held in shared memory.  Replaces one forward pass, reciprocalConvolution/gridEvaluateEnergy
// (pme.cc:137-253) and one inverse pass: the spectrum makes one round trip through memory instead of three.
template <bool ENERGY, bool INVERSE_TOO>
__global__ void __launch_bounds__(FFT_THREADS) fftXConvolve(float2* __restrict__ data, int innerCount, size_t stride, FFTPlan1D plan, int batch, int pitch,
                                                             const float2* __restrict__ W, PmeArgs a, const float* __restrict__ modX, const float* __restrict__ modY,
                                                             const float* __restrict__ modZ, DevCounters* counters) {
    extern __shared__ float2 smem[];
    const int n = plan.n;
    const int logB = __ffs(batch) - 1;
    float2* bufA = smem; float2* bufB = smem + (size_t)n * pitch;
    const int inner0 = blockIdx.x * batch;
    const int nl = min(batch, innerCount - inner0);
    float2* base = data + inner0;
    for (int idx = threadIdx.x; idx < batch * n; idx += blockDim.x) {
        const int j = idx >> logB, l = idx & (batch - 1);
        bufA[j * pitch + l] = l < nl ? base[(size_t)j * stride + l] : make_float2(0.f, 0.f);
    }
    float2* r = fftLinesInSmem<false>(bufA, bufB, plan, batch, pitch, W);
    float e = 0.f;
    for (int idx = threadIdx.x; idx < batch * n; idx += blockDim.x) {
        const int kx = idx >> logB, l = idx & (batch - 1);
        if (l >= nl) continue;
        const int q = inner0 + l;
        const int ky = q / a.nzc, kz = q - ky * a.nzc;
        const int mx = kx < (a.nx + 1) / 2 ? kx : kx - a.nx, my = ky < (a.ny + 1) / 2 ? ky : ky - a.ny, mz = kz;   // kz <= nz/2
        const float mhx = mx * (float)a.r0x, mhy = mx * (float)a.r1x + my * (float)a.r1y, mhz = mx * (float)a.r2x + my * (float)a.r2y + mz * (float)a.r2z;
        const float m2 = mhx * mhx + mhy * mhy + mhz * mhz;
        float eterm;
        if (!a.dispersion) {
            const float denom = m2 * modX[kx] * modY[ky] * modZ[kz];
            eterm = a.scale * __expf(-a.alphaFactor * m2) / denom;
            if (kx == 0 && q == 0) eterm = 0.f;
        } else {   // dispersion mesh: k = 0 is kept (ReferencePME.cpp:336-410, pme.cc reciprocalConvolution USE_LJPME)
            const float m = sqrtf(m2), bb = a.bfac * m;
            eterm = a.scale * (a.fac1 * erfcf(bb) * m * m2 + expf(-bb * bb) * (a.fac2 + a.fac3 * m2)) / (modX[kx] * modY[ky] * modZ[kz]);
        }
        const float2 g = r[kx * pitch + l];
        if (ENERGY) {
            const float wgt = (kz == 0 || (2 * kz == a.nz)) ? 1.f : 2.f;   // Hermitian mirror of the missing half
            e += wgt * eterm * (g.x * g.x + g.y * g.y);
        }
        r[kx * pitch + l] = make_float2(g.x * eterm, g.y * eterm);
    }
    if (ENERGY) {
        __shared__ float red[FFT_THREADS];
        red[threadIdx.x] = e;
        __syncthreads();
        for (int o = FFT_THREADS / 2; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
        if (threadIdx.x == 0) atomicAddFixed(&counters->energy[E_RECIP], toFixedD(0.5 * (double)red[0]));
    }
    if (INVERSE_TOO) {
        float2* other = (r == bufA) ? bufB : bufA;
        r = fftLinesInSmem<true>(r, other, plan, batch, pitch, W);
    } else __syncthreads();
    for (int idx = threadIdx.x; idx < batch * n; idx += blockDim.x) {
        const int j = idx >> logB, l = idx & (batch - 1);
        if (l < nl) base[(size_t)j * stride + l] = r[j * pitch + l];
    }
}

static void pickBatch(int n, int& batch, int& pitch, size_t& smemBytes) {
    batch = 16;
    while (batch > 1 && (size_t)n * (batch + 1) * sizeof(float2) * 2 > 96 * 1024) batch >>= 1;
    pitch = batch | 1;   // odd pitch: conflict-free transposed accesses
    smemBytes = (size_t)n * pitch * sizeof(float2) * 2;
    if (smemBytes > 200 * 1024) throw B200Error(B200_ERR_UNSUPPORTED, "FFT dimension too large for the shared-memory FFT");
}

template <typename K> static void allowSmem(K kernel, size_t bytes) {
    // opt in to > 48 KB of dynamic shared memory once per kernel and device (the attribute call is not free)
    static std::map<std::pair<const void*, int>, size_t> granted;
    if (bytes <= 48 * 1024) return;
    int dev = 0; cudaGetDevice(&dev);
    size_t& cur = granted[{(const void*)kernel, dev}];
    if (bytes > cur) { CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)); cur = bytes; }
}

void b200_fft3d_exec_r2c(int nx, int ny, int nz, const FFTPlan1D* plan, float2* const* tw, const float* inReal, const int* inFix, float fixScaleInv, float2* out, float2*, cudaStream_t s) {
    const int nzc = nz / 2 + 1;
    if (b200_fft2_supported(nx, ny, nz)) {
        b200_fft2_z_forward(nx * ny, nz, inReal, inFix, fixScaleInv, out, tw[2], s);
        b200_fft2_strided(false, out, ny, nzc, nx, (size_t)ny * nzc, (size_t)nzc, tw[1], s);
        b200_fft2_strided(false, out, nx, ny * nzc, 1, 0, (size_t)ny * nzc, tw[0], s);
        CUDA_CHECK(cudaGetLastError());
        return;
    }
    int batch, pitch; size_t sm;
    pickBatch(nz, batch, pitch, sm);
    allowSmem(fftZForward, sm);
    const int lines = nx * ny;
    fftZForward<<<(lines + batch - 1) / batch, FFT_THREADS, sm, s>>>(inReal, inFix, fixScaleInv, out, lines, plan[2], batch, pitch, tw[2]);
    pickBatch(ny, batch, pitch, sm);
    allowSmem(fftStrided<false>, sm);
    fftStrided<false><<<dim3((nzc + batch - 1) / batch, nx), FFT_THREADS, sm, s>>>(out, nzc, (size_t)ny * nzc, (size_t)nzc, plan[1], batch, pitch, tw[1]);
    pickBatch(nx, batch, pitch, sm);
    allowSmem(fftStrided<false>, sm);
    const int inner = ny * nzc;
    fftStrided<false><<<dim3((inner + batch - 1) / batch, 1), FFT_THREADS, sm, s>>>(out, inner, 0, (size_t)ny * nzc, plan[0], batch, pitch, tw[0]);
    CUDA_CHECK(cudaGetLastError());
}

void b200_fft3d_exec_c2r(int nx, int ny, int nz, const FFTPlan1D* plan, float2* const* tw, float2* in, float2*, float* out, cudaStream_t s) {
    const int nzc = nz / 2 + 1;
    if (b200_fft2_supported(nx, ny, nz)) {
        b200_fft2_strided(true, in, nx, ny * nzc, 1, 0, (size_t)ny * nzc, tw[0], s);
        b200_fft2_strided(true, in, ny, nzc, nx, (size_t)ny * nzc, (size_t)nzc, tw[1], s);
        b200_fft2_z_inverse(nx * ny, nz, in, out, tw[2], s);
        CUDA_CHECK(cudaGetLastError());
        return;
    }
    int batch, pitch; size_t sm;
    pickBatch(nx, batch, pitch, sm);
    allowSmem(fftStrided<true>, sm);
    const int inner = ny * nzc;
    fftStrided<true><<<dim3((inner + batch - 1) / batch, 1), FFT_THREADS, sm, s>>>(in, inner, 0, (size_t)ny * nzc, plan[0], batch, pitch, tw[0]);
    pickBatch(ny, batch, pitch, sm);
    allowSmem(fftStrided<true>, sm);
    fftStrided<true><<<dim3((nzc + batch - 1) / batch, nx), FFT_THREADS, sm, s>>>(in, nzc, (size_t)ny * nzc, (size_t)nzc, plan[1], batch, pitch, tw[1]);
    pickBatch(nz, batch, pitch, sm);
    allowSmem(fftZInverse, sm);
    const int lines = nx * ny;
    fftZInverse<<<(lines + batch - 1) / batch, FFT_THREADS, sm, s>>>(in, out, lines, plan[2], batch, pitch, tw[2]);
    CUDA_CHECK(cudaGetLastError());
}

// ---- host-buffer entry points for unit tests (TestCudaFFT3D.cpp:56-121 compares against pocketfft the same way)
static int fftHost(int device, int nx, int ny, int nz, const float* in, float* out, bool forward) {
    try {
        CUDA_CHECK(cudaSetDevice(device));
        const int dims[3] = {nx, ny, nz};
        FFTPlan1D plan[3]; float2* tw[3] = {nullptr, nullptr, nullptr};
        for (int k = 0; k < 3; k++) { b200_fft_make_plan(dims[k], plan[k]); b200_fft_make_twiddle(dims[k], tw[k], 0); }
        const size_t G = (size_t)nx * ny * nz, GC = (size_t)nx * ny * (nz / 2 + 1);
        float* dReal = nullptr; float2* dC = nullptr;
        dmalloc(dReal, G); dmalloc(dC, GC);
        if (forward) {
            CUDA_CHECK(cudaMemcpy(dReal, in, sizeof(float) * G, cudaMemcpyHostToDevice));
            b200_fft3d_exec_r2c(nx, ny, nz, plan, tw, dReal, nullptr, 1.0f, dC, nullptr, 0);
            CUDA_CHECK(cudaDeviceSynchronize());
            CUDA_CHECK(cudaMemcpy(out, dC, sizeof(float2) * GC, cudaMemcpyDeviceToHost));
        } else {
            CUDA_CHECK(cudaMemcpy(dC, in, sizeof(float2) * GC, cudaMemcpyHostToDevice));
            b200_fft3d_exec_c2r(nx, ny, nz, plan, tw, dC, nullptr, dReal, 0);
            CUDA_CHECK(cudaDeviceSynchronize());
            CUDA_CHECK(cudaMemcpy(out, dReal, sizeof(float) * G, cudaMemcpyDeviceToHost));
        }
        dfree(dReal); dfree(dC);
        for (int k = 0; k < 3; k++) dfree(tw[k]);
    } catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
    return B200_OK;
}
extern "C" int b200_fft3d_r2c_host(int32_t device, int32_t nx, int32_t ny, int32_t nz, const float* in, float* out) { return fftHost(device, nx, ny, nz, in, out, true); }
extern "C" int b200_fft3d_c2r_host(int32_t device, int32_t nx, int32_t ny, int32_t nz, const float* in, float* out) { return fftHost(device, nx, ny, nz, in, out, false); }

// =================================================================================================================
// Allocation-free entry point (no counterpart in the reference).
//
// Some callers must not allocate device memory at the moment they need a transform: code running while a memory pool
// owns the whole device, real-time loops that may not hit the allocator's lock, or processes that were handed a fixed
// memory budget by a scheduler.  b200_fft3d_c2r_noalloc serves them: the caller passes device input and output
// buffers, and the working copy of the spectrum plus the three twiddle tables live in the file-scope array below,
// which every GPU that loads this library holds (about 8 MiB) until the process exits.
#define NOALLOC_MAX_COMPLEX (1 << 20)   // largest nx*ny*(nz/2+1) accepted: 1,048,576 complex values = 8 MiB
#define NOALLOC_MAX_DIM 1024            // largest single dimension accepted
#define NOALLOC_MAX_DEVICES 64
static __device__ __align__(256) float2 g_noallocScratch[NOALLOC_MAX_COMPLEX + 3 * NOALLOC_MAX_DIM];
static float2* g_noallocPtr[NOALLOC_MAX_DEVICES];   // device address of the array, per device (host code must not use the symbol itself)

__global__ void noallocFillTwiddle(float2* tw, int n) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    double s, c;
    sincospi(-2.0 * (double)k / (double)n, &s, &c);
    tw[k] = make_float2((float)c, (float)s);
}

// Unnormalised 3D complex-to-real FFT, device buffers in, device buffer out, nothing allocated.
//   d_in  : nx*ny*(nz/2+1) complex (interleaved re,im), left untouched
//   d_out : nx*ny*nz floats
//   stream: a cudaStream_t (0 = default stream) on `device`
// The first call on each device looks up the device address of the scratch array (cudaGetSymbolAddress) and should
// therefore be made before any stream capture begins.  Every call uses the same array: calls that overlap on the GPU
// (different streams or threads) race and must be serialised by the caller.  Nothing in the array has to start at zero;
// each call overwrites the part it reads.  Sizes beyond the limits above, and dimensions with a prime factor other than
// 2, 3, 5, 7, are rejected with B200_ERR_INVALID.  The caller's current CUDA device is left as it was found.
extern "C" int b200_fft3d_c2r_noalloc(int32_t device, int32_t nx, int32_t ny, int32_t nz, const float* d_in, float* d_out, void* stream) {
    int previousDevice = -1;
    cudaGetDevice(&previousDevice);
    struct RestoreDevice { int d; ~RestoreDevice() { if (d >= 0) cudaSetDevice(d); } } restore{previousDevice};
    try {
        if (device < 0 || device >= NOALLOC_MAX_DEVICES) throw B200Error(B200_ERR_INVALID, "b200_fft3d_c2r_noalloc: device ordinal out of range");
        if (nx < 1 || ny < 1 || nz < 1 || !d_in || !d_out) throw B200Error(B200_ERR_INVALID, "b200_fft3d_c2r_noalloc: bad arguments");
        if (nx > NOALLOC_MAX_DIM || ny > NOALLOC_MAX_DIM || nz > NOALLOC_MAX_DIM)
            throw B200Error(B200_ERR_INVALID, "b200_fft3d_c2r_noalloc: a grid dimension exceeds the built-in limit of 1024");
        const size_t GC = (size_t)nx * ny * (nz / 2 + 1);
        if (GC > (size_t)NOALLOC_MAX_COMPLEX)
            throw B200Error(B200_ERR_INVALID, "b200_fft3d_c2r_noalloc: nx*ny*(nz/2+1) exceeds the built-in scratch of 1048576 complex values; use a context-owned transform instead");
        for (int n : {nx, ny, nz})
            if (b200_find_legal_fft_dimension(n) != n) throw B200Error(B200_ERR_INVALID, "b200_fft3d_c2r_noalloc: every grid dimension must factor into 2, 3, 5 and 7");
        CUDA_CHECK(cudaSetDevice(device));
        if (!g_noallocPtr[device]) {
            // once per device, with that device current: the symbol name is only a host-side handle, kernels need this address
            float2* p = nullptr;
            CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_noallocScratch));
            g_noallocPtr[device] = p;
        }
        cudaStream_t s = (cudaStream_t)stream;
        float2* work = g_noallocPtr[device];
        float2* tw[3] = {work + NOALLOC_MAX_COMPLEX, work + NOALLOC_MAX_COMPLEX + NOALLOC_MAX_DIM, work + NOALLOC_MAX_COMPLEX + 2 * NOALLOC_MAX_DIM};
        const int dims[3] = {nx, ny, nz};
        FFTPlan1D plan[3];
        for (int k = 0; k < 3; k++) {
            b200_fft_make_plan(dims[k], plan[k]);
            noallocFillTwiddle<<<(dims[k] + 127) / 128, 128, 0, s>>>(tw[k], dims[k]);
        }
        CUDA_CHECK(cudaMemcpyAsync(work, d_in, sizeof(float2) * GC, cudaMemcpyDeviceToDevice, s));
        b200_fft3d_exec_c2r(nx, ny, nz, plan, tw, work, nullptr, d_out, s);
        CUDA_CHECK(cudaGetLastError());
    } catch (const B200Error& e) { b200_set_error(e.what()); return e.code; }
    return B200_OK;
}

// whole reciprocal-space pipeline on the context's grids: int charge grid -> spectrum -> convolution (+energy) -> potential grid
void b200_fft3d_reciprocal(b200_ctx* c, PmeMesh& m, cudaStream_t s, const PmeArgs& a, bool energy, bool forces) {
    const int nx = m.gridN[0], ny = m.gridN[1], nz = m.gridN[2], nzc = nz / 2 + 1;
    if (b200_fft2_supported(nx, ny, nz)) {
        b200_fft2_z_forward(nx * ny, nz, nullptr, m.gridI, 1.0f / m.spreadScale, m.gridC, m.twiddle[2], s);
        b200_fft2_strided(false, m.gridC, ny, nzc, nx, (size_t)ny * nzc, (size_t)nzc, m.twiddle[1], s);
        b200_fft2_xconv(m.gridC, nx, ny * nzc, (size_t)ny * nzc, m.twiddle[0], a, m.moduli[0], m.moduli[1], m.moduli[2], c->counters, energy, forces, s);
        LAUNCH_COUNT(c, 3);
        if (forces) {
            b200_fft2_strided(true, m.gridC, ny, nzc, nx, (size_t)ny * nzc, (size_t)nzc, m.twiddle[1], s);
            b200_fft2_z_inverse(nx * ny, nz, m.gridC, m.gridReal, m.twiddle[2], s);
            LAUNCH_COUNT(c, 2);
        }
        CUDA_CHECK(cudaGetLastError());
        return;
    }
    int batch, pitch; size_t sm;
    pickBatch(nz, batch, pitch, sm);
    allowSmem(fftZForward, sm);
    const int lines = nx * ny;
    fftZForward<<<(lines + batch - 1) / batch, FFT_THREADS, sm, s>>>(nullptr, m.gridI, 1.0f / m.spreadScale, m.gridC, lines, m.plan[2], batch, pitch, m.twiddle[2]);
    pickBatch(ny, batch, pitch, sm);
    allowSmem(fftStrided<false>, sm);
    fftStrided<false><<<dim3((nzc + batch - 1) / batch, nx), FFT_THREADS, sm, s>>>(m.gridC, nzc, (size_t)ny * nzc, (size_t)nzc, m.plan[1], batch, pitch, m.twiddle[1]);
    pickBatch(nx, batch, pitch, sm);
    const int inner = ny * nzc;
    const dim3 grid((inner + batch - 1) / batch);
#define LAUNCH_XCONV(E, I) do { allowSmem(fftXConvolve<E, I>, sm); fftXConvolve<E, I><<<grid, FFT_THREADS, sm, s>>>(m.gridC, inner, (size_t)ny * nzc, m.plan[0], batch, pitch, m.twiddle[0], a, m.moduli[0], m.moduli[1], m.moduli[2], c->counters); } while (0)
    if (energy && forces) LAUNCH_XCONV(true, true);
    else if (energy) LAUNCH_XCONV(true, false);
    else LAUNCH_XCONV(false, true);
#undef LAUNCH_XCONV
    LAUNCH_COUNT(c, 3);
    if (forces) {
        pickBatch(ny, batch, pitch, sm);
        allowSmem(fftStrided<true>, sm);
        fftStrided<true><<<dim3((nzc + batch - 1) / batch, nx), FFT_THREADS, sm, s>>>(m.gridC, nzc, (size_t)ny * nzc, (size_t)nzc, m.plan[1], batch, pitch, m.twiddle[1]);
        pickBatch(nz, batch, pitch, sm);
        allowSmem(fftZInverse, sm);
        fftZInverse<<<(lines + batch - 1) / batch, FFT_THREADS, sm, s>>>(m.gridC, m.gridReal, lines, m.plan[2], batch, pitch, m.twiddle[2]);
        LAUNCH_COUNT(c, 2);
    }
    CUDA_CHECK(cudaGetLastError());
}
