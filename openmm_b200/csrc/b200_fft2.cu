// b200_fft2.cu -- register-resident two-stage FFT passes (the fast path of the 3D transform in b200_fft.cu).
//
// A 1D transform of length n = R1*R2 (2 <= R <= 16, factors 2,3,5,7; R2 = 1 for n <= 16) is done by two Stockham stages whose
// radix-R butterflies live entirely in registers (b200_fft_dft.cuh): stage 1 loads its inputs straight from global memory,
// the two stages exchange data once through shared memory, stage 2 writes straight back to global memory.  That is one
// shared-memory round trip per 1D transform instead of one per radix-2..7 factor, and no staging copies.
//   * strided passes (y, x): consecutive threads take consecutive lines, so every global access is a run of 8-byte
//     elements; the x pass keeps the spectrum in registers between the forward transform, the multiplication by the
//     influence function (reciprocalConvolution / gridEvaluateEnergy, pme.cc:137-253) and the first stage of the inverse.
//   * z passes use the real-data trick: the nz real values of a line are transformed as nz/2 complex values and
//     untangled (forward) / pre-tangled (inverse), which halves the butterfly work of the contiguous direction.
// Same conventions as cuFFT R2C/C2R (CudaFFT3D.cpp:66-101): unnormalised, half-complex layout [nx][ny][nz/2+1].
#include "b200_internal.cuh"
#include "b200_fft_dft.cuh"
#include <cstdlib>
#include <map>

namespace fft2 {

struct Plan { int n, R1, R2; };   // n = R1 * R2; R2 == 1: single stage

static bool radixOk(int r) { return r == 2 || r == 3 || r == 4 || r == 5 || r == 6 || r == 7 || r == 8 || r == 9 || r == 10 || r == 12 || r == 14 || r == 15 || r == 16; }

static bool makePlan(int n, Plan& p) {
    p.n = n;
    if (n <= 16) { p.R1 = n; p.R2 = 1; return radixOk(n); }
    int best = 0;
    for (int r1 = 2; r1 <= 16; r1++) {
        if (n % r1 != 0) continue;
        const int r2 = n / r1;
        if (r2 > 16 || !radixOk(r1) || !radixOk(r2) || r1 > r2) continue;
        best = r1;   // the most balanced pair has the largest r1 <= r2
    }
    if (!best) return false;
    p.R1 = best; p.R2 = n / best;
    return true;
}

// dispatch on a warp-uniform runtime radix
#define FFT2_SWITCH(R, CALL) \
    switch (R) { \
        case 2: CALL(2); break; case 3: CALL(3); break; case 4: CALL(4); break; case 5: CALL(5); break; case 6: CALL(6); break; \
        case 7: CALL(7); break; case 8: CALL(8); break; case 9: CALL(9); break; case 10: CALL(10); break; case 12: CALL(12); break; \
        case 14: CALL(14); break; case 15: CALL(15); break; case 16: CALL(16); break; default: break; }

// first stage (Ns = 1): butterfly j of n/R takes inputs j + t*items and produces outputs j*R + t
template <int R, bool INV, class LD, class ST>
__device__ __forceinline__ void stageFirst(int j, int items, LD ld, ST st) {
    float2 x[R];
#pragma unroll
    for (int t = 0; t < R; t++) x[t] = ld(j + t * items);
    Dft<R, INV>::run(x);
#pragma unroll
    for (int t = 0; t < R; t++) st(j * R + t, x[t]);
}

// last stage of a two-stage transform (Ns = R1 = number of butterflies): butterfly k takes inputs k + t*Ns multiplied by
// W_n^(t k) and produces outputs k + t*Ns.  tw[m] = exp(-2 pi i m / n).
template <int R, bool INV, class LD>
__device__ __forceinline__ void loadTwiddled(float2 (&x)[R], int k, int Ns, const float2* tw, int twStride, LD ld) {
    x[0] = ld(k);
#pragma unroll
    for (int t = 1; t < R; t++) {
        float2 w = tw[t * k * twStride];
        if (INV) w.y = -w.y;
        x[t] = cmul(ld(k + t * Ns), w);
    }
}
template <int R, bool INV, class LD, class ST>
__device__ __forceinline__ void stageLast(int k, int Ns, const float2* tw, int twStride, LD ld, ST st) {
    float2 x[R];
    loadTwiddled<R, INV>(x, k, Ns, tw, twStride, ld);
    Dft<R, INV>::run(x);
#pragma unroll
    for (int t = 0; t < R; t++) st(k + t * Ns, x[t]);
}

__device__ __forceinline__ void loadTwiddleTable(float2* tw, const float2* __restrict__ W, int n, int srcStride) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) tw[i] = W[(size_t)i * srcStride];
}

// -------------------------------------------------------------------------------------------------------------------
// strided complex pass (y or x direction), in place.  Lines are addressed as outer*outerStride + inner + idx*stride;
// thread -> (line l = tid % B, butterfly j = tid / B).
template <bool INV>
__global__ void __launch_bounds__(512) stridedPass(float2* __restrict__ data, int innerCount, size_t outerStride, size_t stride, Plan p, int B, int pitch,
                                                   const float2* __restrict__ W) {
    extern __shared__ float2 smem[];
    float2* tw = smem; float2* buf = smem + p.n;
    const int l = threadIdx.x % B, j = threadIdx.x / B;
    const int inner0 = blockIdx.x * B;
    const bool lineOk = inner0 + l < innerCount;
    float2* base = data + (size_t)blockIdx.y * outerStride + inner0 + l;
    const int items1 = p.n / p.R1;
    auto ldG = [&](int idx) { return base[(size_t)idx * stride]; };
    auto stG = [&](int idx, float2 v) { base[(size_t)idx * stride] = v; };
    auto ldS = [&](int idx) { return buf[idx * pitch + l]; };
    auto stS = [&](int idx, float2 v) { buf[idx * pitch + l] = v; };
    if (p.R2 == 1) {
#define CALL(RR) stageFirst<RR, INV>(0, 1, ldG, stG)
        if (lineOk && j == 0) FFT2_SWITCH(p.R1, CALL)
#undef CALL
        return;
    }
    loadTwiddleTable(tw, W, p.n, 1);
#define CALL(RR) stageFirst<RR, INV>(j, items1, ldG, stS)
    if (lineOk && j < items1) FFT2_SWITCH(p.R1, CALL)
#undef CALL
    __syncthreads();
#define CALL(RR) stageLast<RR, INV>(j, p.R1, tw, 1, ldS, stG)
    if (lineOk && j < p.R1) FFT2_SWITCH(p.R2, CALL)
#undef CALL
}

// -------------------------------------------------------------------------------------------------------------------
// x pass fused with the reciprocal convolution.  After the forward transform thread (l, k) holds X[k + t*R1], t < R2, of its
// line in registers; it multiplies them by the influence function, accumulates the energy, and -- because those are exactly
// the inputs of butterfly k of the inverse transform's first stage when that stage has radix R2 -- continues with the inverse
// transform without going through memory.
struct ConvLine {      // per-line (ky, kz) part of the influence function
    float mhyY, mhzYZ, modYZ, wgt; bool zeroK0;
};

__device__ __forceinline__ float influence(const PmeArgs& a, const ConvLine& c, int kx, const float* __restrict__ modX) {
    const int mx = kx < (a.nx + 1) / 2 ? kx : kx - a.nx;
    const float mhx = mx * (float)a.r0x, mhy = fmaf(mx, (float)a.r1x, c.mhyY), mhz = fmaf(mx, (float)a.r2x, c.mhzYZ);
    const float m2 = mhx * mhx + mhy * mhy + mhz * mhz;
    if (!a.dispersion) {
        const float denom = m2 * modX[kx] * c.modYZ;
        const float eterm = a.scale * __expf(-a.alphaFactor * m2) / denom;
        return (c.zeroK0 && kx == 0) ? 0.f : eterm;
    }
    const float m = sqrtf(m2), bb = a.bfac * m;   // dispersion mesh: k = 0 is kept (ReferencePME.cpp:336-410)
    return a.scale * (a.fac1 * erfcf(bb) * m * m2 + expf(-bb * bb) * (a.fac2 + a.fac3 * m2)) / (modX[kx] * c.modYZ);
}

template <int R, bool ENERGY>
__device__ __forceinline__ void convolve(float2 (&x)[R], int k, int Ns, const PmeArgs& a, const ConvLine& c, const float* __restrict__ modX, float& e) {
#pragma unroll
    for (int t = 0; t < R; t++) {
        const float eterm = influence(a, c, k + t * Ns, modX);
        if (ENERGY) e = fmaf(c.wgt * eterm, x[t].x * x[t].x + x[t].y * x[t].y, e);
        x[t].x *= eterm; x[t].y *= eterm;
    }
}

// middle of the x pass for a two-stage plan: forward last stage (radix R2) -> convolution -> inverse first stage (radix R2).
// Every thread of the block calls this (it contains a barrier); `active` guards the memory accesses.
template <int R, bool ENERGY>
__device__ __forceinline__ void xMiddle(bool active, bool inverseToo, int k, int l, int Ns, float2* buf, int pitch, const float2* tw, float2* base, size_t stride,
                                        const PmeArgs& a, const ConvLine& c, const float* __restrict__ modX, float& e) {
    float2 x[R];
    if (active) {
        loadTwiddled<R, false>(x, k, Ns, tw, 1, [&](int idx) { return buf[idx * pitch + l]; });
        Dft<R, false>::run(x);
        convolve<R, ENERGY>(x, k, Ns, a, c, modX, e);
        if (inverseToo) Dft<R, true>::run(x);   // inverse first stage: inputs k + t*Ns (items = Ns), outputs k*R + t
    }
    __syncthreads();                            // everybody has read its forward inputs from buf
    if (active) {
        if (inverseToo) {
#pragma unroll
            for (int t = 0; t < R; t++) buf[(k * R + t) * pitch + l] = x[t];
        } else {
#pragma unroll
            for (int t = 0; t < R; t++) base[(size_t)(k + t * Ns) * stride] = x[t];   // energy-only call: store the convolved spectrum
        }
    }
}

template <int R, bool ENERGY>
__device__ __forceinline__ void xSingle(bool inverseToo, float2* base, size_t stride, const PmeArgs& a, const ConvLine& c, const float* __restrict__ modX, float& e) {
    float2 x[R];
#pragma unroll
    for (int t = 0; t < R; t++) x[t] = base[(size_t)t * stride];
    Dft<R, false>::run(x);
    convolve<R, ENERGY>(x, 0, 1, a, c, modX, e);
    if (inverseToo) Dft<R, true>::run(x);
#pragma unroll
    for (int t = 0; t < R; t++) base[(size_t)t * stride] = x[t];
}

template <bool ENERGY>
__global__ void __launch_bounds__(512) xConvolvePass(float2* __restrict__ data, int innerCount, size_t stride, Plan p, int B, int pitch, const float2* __restrict__ W,
                                                     PmeArgs a, const float* __restrict__ modX, const float* __restrict__ modY, const float* __restrict__ modZ,
                                                     DevCounters* counters, int inverseToo) {
    extern __shared__ float2 smem[];
    __shared__ float red[16];
    float2* tw = smem; float2* buf = smem + p.n;
    const int l = threadIdx.x % B, j = threadIdx.x / B;
    const int q = blockIdx.x * B + l;
    const bool lineOk = q < innerCount;
    float2* base = data + q;
    ConvLine c; c.mhyY = c.mhzYZ = c.modYZ = c.wgt = 0.f; c.zeroK0 = false;
    if (lineOk) {
        const int ky = q / a.nzc, kz = q - ky * a.nzc;
        const int my = ky < (a.ny + 1) / 2 ? ky : ky - a.ny, mz = kz;   // kz <= nz/2
        c.mhyY = my * (float)a.r1y; c.mhzYZ = my * (float)a.r2y + mz * (float)a.r2z;
        c.modYZ = modY[ky] * modZ[kz];
        c.wgt = (kz == 0 || 2 * kz == a.nz) ? 1.f : 2.f;                 // Hermitian mirror of the missing half
        c.zeroK0 = q == 0;
    }
    float e = 0.f;
    if (p.R2 == 1) {
#define CALL(RR) xSingle<RR, ENERGY>(inverseToo != 0, base, stride, a, c, modX, e)
        if (lineOk && j == 0) FFT2_SWITCH(p.R1, CALL)
#undef CALL
    } else {
        loadTwiddleTable(tw, W, p.n, 1);
        const int items1 = p.n / p.R1;
        auto ldG = [&](int idx) { return base[(size_t)idx * stride]; };
        auto stG = [&](int idx, float2 v) { base[(size_t)idx * stride] = v; };
        auto ldS = [&](int idx) { return buf[idx * pitch + l]; };
        auto stS = [&](int idx, float2 v) { buf[idx * pitch + l] = v; };
#define CALL(RR) stageFirst<RR, false>(j, items1, ldG, stS)
        if (lineOk && j < items1) FFT2_SWITCH(p.R1, CALL)
#undef CALL
        __syncthreads();
#define CALL(RR) xMiddle<RR, ENERGY>(lineOk && j < p.R1, inverseToo != 0, j, l, p.R1, buf, pitch, tw, base, stride, a, c, modX, e)
        FFT2_SWITCH(p.R2, CALL)
#undef CALL
        if (inverseToo) {
            __syncthreads();
            // inverse last stage: radix R1, Ns = R2 butterflies
#define CALL(RR) stageLast<RR, true>(j, p.R2, tw, 1, ldS, stG)
            if (lineOk && j < p.R2) FFT2_SWITCH(p.R1, CALL)
#undef CALL
        }
    }
    if (ENERGY) {   // fixed-order block sum, then one fixed-point atomic: bit-reproducible
        for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = e;
        __syncthreads();
        if (threadIdx.x == 0) {
            float s = 0.f;
            for (int w = 0; w < (int)(blockDim.x + 31) / 32; w++) s += red[w];
            if (s != 0.f) atomicAddFixed(&counters->energy[E_RECIP], toFixedD(0.5 * (double)s));
        }
    }
}

// -------------------------------------------------------------------------------------------------------------------
// z pass, forward (real -> half complex) with the real-data trick: z[m] = x[2m] + i x[2m+1], Z = FFT_{n}(z), n = nz/2,
//   X[k] = (Z[k] + conj(Z[n-k]))/2 - i W_nz^k (Z[k] - conj(Z[n-k]))/2,   k = 0..n   (Z[n] = Z[0]).
// thread -> (butterfly j = tid % items, line l = tid / items).  Wz[m] = exp(-2 pi i m / nz); the table of the half-length
// transform is every second entry.
__global__ void __launch_bounds__(512) zForwardPass(const float* __restrict__ inReal, const int* __restrict__ inFix, float fixScaleInv, float2* __restrict__ out,
                                                    int numLines, int nz, Plan p, int B, int pitch, int items, const float2* __restrict__ Wz) {
    extern __shared__ float2 smem[];
    const int n = p.n, nzc = n + 1;
    float2* tw = smem; float2* bufA = smem + n; float2* bufB = bufA + (size_t)n * pitch;
    const int j = threadIdx.x % items, l = threadIdx.x / items;
    const int line0 = blockIdx.x * B;
    const bool lineOk = l < B && line0 + l < numLines;
    const size_t lineBase = (size_t)(line0 + l) * nz;
    auto ldG = [&](int idx) {
        if (inFix) { const int2 v = *reinterpret_cast<const int2*>(inFix + lineBase + 2 * idx); return make_float2((float)v.x * fixScaleInv, (float)v.y * fixScaleInv); }
        return *reinterpret_cast<const float2*>(inReal + lineBase + 2 * idx);
    };
    auto ldA = [&](int idx) { return bufA[idx * pitch + l]; };
    auto stA = [&](int idx, float2 v) { bufA[idx * pitch + l] = v; };
    auto stB = [&](int idx, float2 v) { bufB[idx * pitch + l] = v; };
    loadTwiddleTable(tw, Wz, n, 2);
    if (p.R2 == 1) {
#define CALL(RR) stageFirst<RR, false>(0, 1, ldG, stB)
        if (lineOk && j == 0) FFT2_SWITCH(p.R1, CALL)
#undef CALL
    } else {
        const int items1 = n / p.R1;
#define CALL(RR) stageFirst<RR, false>(j, items1, ldG, stA)
        if (lineOk && j < items1) FFT2_SWITCH(p.R1, CALL)
#undef CALL
        __syncthreads();
#define CALL(RR) stageLast<RR, false>(j, p.R1, tw, 1, ldA, stB)
        if (lineOk && j < p.R1) FFT2_SWITCH(p.R2, CALL)
#undef CALL
    }
    __syncthreads();
    // untangle and store: one warp per line, lanes along k (coalesced 8-byte stores)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (int ll = warp; ll < B && line0 + ll < numLines; ll += nwarps) {
        float2* o = out + (size_t)(line0 + ll) * nzc;
        for (int k = lane; k <= n; k += 32) {
            const float2 zk = bufB[(k == n ? 0 : k) * pitch + ll];
            const float2 zc = bufB[(k == 0 ? 0 : n - k) * pitch + ll];
            const float2 ev = make_float2(0.5f * (zk.x + zc.x), 0.5f * (zk.y - zc.y));     // (Z[k] + conj Z[n-k]) / 2
            const float2 od = make_float2(0.5f * (zk.x - zc.x), 0.5f * (zk.y + zc.y));     // (Z[k] - conj Z[n-k]) / 2
            const float2 w = Wz[k];                                                        // exp(-2 pi i k / nz)
            const float2 t = cmul(od, w);
            o[k] = make_float2(ev.x + t.y, ev.y - t.x);                                    // ev - i t
        }
    }
}

// z pass, inverse (half complex -> real):  Z[k] = (X[k] + conj X[n-k]) + i conj(W_nz^k) (X[k] - conj X[n-k]),  z = IFFT_n(Z),
// x[2m] = Re z[m], x[2m+1] = Im z[m]   (unnormalised, like cuFFT C2R).
__global__ void __launch_bounds__(512) zInversePass(const float2* __restrict__ in, float* __restrict__ out, int numLines, int nz, Plan p, int B, int pitch, int items,
                                                    const float2* __restrict__ Wz) {
    extern __shared__ float2 smem[];
    const int n = p.n, nzc = n + 1;
    float2* tw = smem; float2* twFull = smem + n; float2* bufA = twFull + nzc; float2* bufB = bufA + (size_t)n * pitch;
    const int j = threadIdx.x % items, l = threadIdx.x / items;
    const int line0 = blockIdx.x * B;
    const bool lineOk = l < B && line0 + l < numLines;
    const float2* src = in + (size_t)(line0 + l) * nzc;
    for (int i = threadIdx.x; i < nzc; i += blockDim.x) twFull[i] = Wz[i];
    loadTwiddleTable(tw, Wz, n, 2);
    __syncthreads();
    auto ldG = [&](int idx) {
        const float2 xk = src[idx], xc = src[n - idx];
        const float2 s = make_float2(xk.x + xc.x, xk.y - xc.y);      // X[k] + conj X[n-k]
        const float2 d = make_float2(xk.x - xc.x, xk.y + xc.y);      // X[k] - conj X[n-k]
        const float2 w = twFull[idx];                                 // exp(-2 pi i k/nz); need i * conj(w) * d
        const float2 t = make_float2(fmaf(d.x, w.x, d.y * w.y), fmaf(d.y, w.x, -d.x * w.y));   // conj(w) * d
        return make_float2(s.x - t.y, s.y + t.x);                    // s + i t
    };
    auto ldA = [&](int idx) { return bufA[idx * pitch + l]; };
    auto stA = [&](int idx, float2 v) { bufA[idx * pitch + l] = v; };
    auto stB = [&](int idx, float2 v) { bufB[idx * pitch + l] = v; };
    if (p.R2 == 1) {
#define CALL(RR) stageFirst<RR, true>(0, 1, ldG, stB)
        if (lineOk && j == 0) FFT2_SWITCH(p.R1, CALL)
#undef CALL
    } else {
        const int items1 = n / p.R1;
#define CALL(RR) stageFirst<RR, true>(j, items1, ldG, stA)
        if (lineOk && j < items1) FFT2_SWITCH(p.R1, CALL)
#undef CALL
        __syncthreads();
#define CALL(RR) stageLast<RR, true>(j, p.R1, tw, 1, ldA, stB)
        if (lineOk && j < p.R1) FFT2_SWITCH(p.R2, CALL)
#undef CALL
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (int ll = warp; ll < B && line0 + ll < numLines; ll += nwarps) {
        float2* o = reinterpret_cast<float2*>(out + (size_t)(line0 + ll) * nz);
        for (int m = lane; m < n; m += 32) o[m] = bufB[m * pitch + ll];
    }
}

}   // namespace fft2

// -------------------------------------------------------------------------------------------------------------------
// host side
using fft2::Plan;

template <typename K> static void allowSmem2(K kernel, size_t bytes) {
    static std::map<std::pair<const void*, int>, size_t> granted;
    if (bytes <= 48 * 1024) return;
    int dev = 0; cudaGetDevice(&dev);
    size_t& cur = granted[{(const void*)kernel, dev}];
    if (bytes > cur) { CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)); cur = bytes; }
}

static int roundUp32(int x) { return (x + 31) & ~31; }

bool b200_fft2_supported(int nx, int ny, int nz) {
    static const bool legacy = getenv("B200_FFT_LEGACY") != nullptr;
    if (legacy) return false;
    Plan p;
    return nz % 2 == 0 && fft2::makePlan(nz / 2, p) && fft2::makePlan(ny, p) && fft2::makePlan(nx, p);
}

// lines per block for the strided passes: as close to 32 as divides the work evenly
static void stridedGeometry(const Plan& p, int innerCount, int& B, int& pitch, int& threads, size_t& smem, int& blocks) {
    const int nb = (innerCount + 31) / 32;
    B = (innerCount + nb - 1) / nb;
    pitch = B | 1;
    const int items = p.R2 == 1 ? 1 : std::max(p.n / p.R1, p.R1);
    threads = roundUp32(B * items);
    smem = sizeof(float2) * ((size_t)p.n + (size_t)p.n * pitch);
    blocks = (innerCount + B - 1) / B;
}

void b200_fft2_strided(bool inverse, float2* data, int n, int innerCount, int outerCount, size_t outerStride, size_t stride, const float2* tw, cudaStream_t s) {
    Plan p; fft2::makePlan(n, p);
    int B, pitch, threads, blocks; size_t smem;
    stridedGeometry(p, innerCount, B, pitch, threads, smem, blocks);
    if (inverse) { allowSmem2(fft2::stridedPass<true>, smem); fft2::stridedPass<true><<<dim3(blocks, outerCount), threads, smem, s>>>(data, innerCount, outerStride, stride, p, B, pitch, tw); }
    else { allowSmem2(fft2::stridedPass<false>, smem); fft2::stridedPass<false><<<dim3(blocks, outerCount), threads, smem, s>>>(data, innerCount, outerStride, stride, p, B, pitch, tw); }
}

void b200_fft2_xconv(float2* data, int nx, int innerCount, size_t stride, const float2* tw, const PmeArgs& a, const float* modX, const float* modY, const float* modZ,
                     DevCounters* counters, bool energy, bool inverseToo, cudaStream_t s) {
    Plan p; fft2::makePlan(nx, p);
    int B, pitch, threads, blocks; size_t smem;
    stridedGeometry(p, innerCount, B, pitch, threads, smem, blocks);
    if (energy) { allowSmem2(fft2::xConvolvePass<true>, smem); fft2::xConvolvePass<true><<<blocks, threads, smem, s>>>(data, innerCount, stride, p, B, pitch, tw, a, modX, modY, modZ, counters, inverseToo ? 1 : 0); }
    else { allowSmem2(fft2::xConvolvePass<false>, smem); fft2::xConvolvePass<false><<<blocks, threads, smem, s>>>(data, innerCount, stride, p, B, pitch, tw, a, modX, modY, modZ, counters, inverseToo ? 1 : 0); }
}

static void zGeometry(const Plan& p, int& B, int& pitch, int& items, int& threads) {
    items = p.R2 == 1 ? 1 : std::max(p.n / p.R1, p.R1);
    B = std::max(1, std::min(32, 448 / items));
    if (B >= 8) B &= ~1;
    pitch = B | 1;
    threads = std::max(64, roundUp32(B * items));
}

void b200_fft2_z_forward(int numLines, int nz, const float* inReal, const int* inFix, float fixScaleInv, float2* out, const float2* twZ, cudaStream_t s) {
    Plan p; fft2::makePlan(nz / 2, p);
    int B, pitch, items, threads; zGeometry(p, B, pitch, items, threads);
    const size_t smem = sizeof(float2) * ((size_t)p.n + 2 * (size_t)p.n * pitch);
    allowSmem2(fft2::zForwardPass, smem);
    fft2::zForwardPass<<<(numLines + B - 1) / B, threads, smem, s>>>(inReal, inFix, fixScaleInv, out, numLines, nz, p, B, pitch, items, twZ);
}

void b200_fft2_z_inverse(int numLines, int nz, const float2* in, float* out, const float2* twZ, cudaStream_t s) {
    Plan p; fft2::makePlan(nz / 2, p);
    int B, pitch, items, threads; zGeometry(p, B, pitch, items, threads);
    const size_t smem = sizeof(float2) * ((size_t)p.n + (size_t)p.n + 1 + 2 * (size_t)p.n * pitch);
    allowSmem2(fft2::zInversePass, smem);
    fft2::zInversePass<<<(numLines + B - 1) / B, threads, smem, s>>>(in, out, numLines, nz, p, B, pitch, items, twZ);
}
