// b200_fft_dft.cuh -- small DFTs held entirely in registers (radices 2..16 with factors 2,3,5,7), the building block of
// the two-stage FFT in b200_fft2.cu.  Everything is unrolled at compile time; the roots of unity are literals
// (b200_fft_roots.h), so a radix-R butterfly is a straight line of FADD/FFMA on R complex registers.
#pragma once
#include "b200_fft_roots.h"

namespace fft2 {

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x)); }

// a * (c -+ i s): forward transforms use exp(-i theta), inverse ones exp(+i theta).  w = (cos, sin) is a compile-time
// constant after unrolling, so the special cases below are resolved statically.
template <bool INV>
__device__ __forceinline__ float2 cmulRoot(float2 a, float2 w) {
    const float s = INV ? w.y : -w.y;
    if (s == 0.f) { if (w.x == 1.f) return a; if (w.x == -1.f) return make_float2(-a.x, -a.y); return make_float2(a.x * w.x, a.y * w.x); }
    if (w.x == 0.f) return make_float2(-a.y * s, a.x * s);
    return make_float2(fmaf(a.x, w.x, -a.y * s), fmaf(a.x, s, a.y * w.x));
}

template <int R, bool INV> struct Dft;   // in place, natural order: x[k] <- sum_n x[n] exp(-+2 pi i n k / R)

template <bool INV> struct Dft<1, INV> { static __device__ __forceinline__ void run(float2 (&)[1]) {} };

template <bool INV> struct Dft<2, INV> {
    static __device__ __forceinline__ void run(float2 (&x)[2]) { const float2 a = cadd(x[0], x[1]), b = csub(x[0], x[1]); x[0] = a; x[1] = b; }
};

template <bool INV> struct Dft<4, INV> {
    static __device__ __forceinline__ void run(float2 (&x)[4]) {
        const float2 a = cadd(x[0], x[2]), b = csub(x[0], x[2]), c = cadd(x[1], x[3]), d = csub(x[1], x[3]);
        const float2 di = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);   // -i d (forward), +i d (inverse)
        x[0] = cadd(a, c); x[1] = cadd(b, di); x[2] = csub(a, c); x[3] = csub(b, di);
    }
};

// odd prime radix through the conjugate symmetry of the roots: with a_t = x_t + x_{R-t}, b_t = x_t - x_{R-t},
//   y_q, y_{R-q} = (x_0 + sum_t a_t cos(2 pi q t/R)) -+ i (sum_t b_t sin(2 pi q t/R))      (forward: - for y_q)
template <int R, bool INV>
__device__ __forceinline__ void dftOddPrime(float2 (&x)[R]) {
    constexpr int H = (R - 1) / 2;
    float2 sa[H + 1], sb[H + 1];
    float2 y0 = x[0];
#pragma unroll
    for (int t = 1; t <= H; t++) { sa[t] = cadd(x[t], x[R - t]); sb[t] = csub(x[t], x[R - t]); y0 = cadd(y0, sa[t]); }
    const float2 x0 = x[0];
    x[0] = y0;
#pragma unroll
    for (int q = 1; q <= H; q++) {
        float2 C = x0, D = make_float2(0.f, 0.f);
#pragma unroll
        for (int t = 1; t <= H; t++) {
            const float2 w = fftRoot<R>((q * t) % R);
            C.x = fmaf(sa[t].x, w.x, C.x); C.y = fmaf(sa[t].y, w.x, C.y);
            D.x = fmaf(sb[t].x, w.y, D.x); D.y = fmaf(sb[t].y, w.y, D.y);
        }
        // forward: y_q = C - iD, y_{R-q} = C + iD ; inverse: the other way round
        const float2 lo = make_float2(C.x + D.y, C.y - D.x), hi = make_float2(C.x - D.y, C.y + D.x);
        x[q] = INV ? hi : lo;
        x[R - q] = INV ? lo : hi;
    }
}
template <bool INV> struct Dft<3, INV> { static __device__ __forceinline__ void run(float2 (&x)[3]) { dftOddPrime<3, INV>(x); } };
template <bool INV> struct Dft<5, INV> { static __device__ __forceinline__ void run(float2 (&x)[5]) { dftOddPrime<5, INV>(x); } };
template <bool INV> struct Dft<7, INV> { static __device__ __forceinline__ void run(float2 (&x)[7]) { dftOddPrime<7, INV>(x); } };

// composite radix R = P*Q by one Cooley-Tukey step in registers: with n = P n2 + n1 and k = k2 + Q k1,
//   X[k2 + Q k1] = sum_{n1} W_P^{n1 k1} . W_R^{n1 k2} . ( sum_{n2} x[P n2 + n1] W_Q^{n2 k2} )
template <int P, int Q, bool INV>
__device__ __forceinline__ void dftComposite(float2 (&x)[P * Q]) {
    constexpr int R = P * Q;
    float2 y[P][Q];
#pragma unroll
    for (int n1 = 0; n1 < P; n1++) {
        float2 sub[Q];
#pragma unroll
        for (int n2 = 0; n2 < Q; n2++) sub[n2] = x[P * n2 + n1];
        Dft<Q, INV>::run(sub);
#pragma unroll
        for (int k2 = 0; k2 < Q; k2++) y[n1][k2] = (n1 == 0 || k2 == 0) ? sub[k2] : cmulRoot<INV>(sub[k2], fftRoot<R>((n1 * k2) % R));
    }
#pragma unroll
    for (int k2 = 0; k2 < Q; k2++) {
        float2 col[P];
#pragma unroll
        for (int n1 = 0; n1 < P; n1++) col[n1] = y[n1][k2];
        Dft<P, INV>::run(col);
#pragma unroll
        for (int k1 = 0; k1 < P; k1++) x[k2 + Q * k1] = col[k1];
    }
}
template <bool INV> struct Dft<6, INV> { static __device__ __forceinline__ void run(float2 (&x)[6]) { dftComposite<2, 3, INV>(x); } };
template <bool INV> struct Dft<8, INV> { static __device__ __forceinline__ void run(float2 (&x)[8]) { dftComposite<2, 4, INV>(x); } };
template <bool INV> struct Dft<9, INV> { static __device__ __forceinline__ void run(float2 (&x)[9]) { dftComposite<3, 3, INV>(x); } };
template <bool INV> struct Dft<10, INV> { static __device__ __forceinline__ void run(float2 (&x)[10]) { dftComposite<2, 5, INV>(x); } };
template <bool INV> struct Dft<12, INV> { static __device__ __forceinline__ void run(float2 (&x)[12]) { dftComposite<3, 4, INV>(x); } };
template <bool INV> struct Dft<14, INV> { static __device__ __forceinline__ void run(float2 (&x)[14]) { dftComposite<2, 7, INV>(x); } };
template <bool INV> struct Dft<15, INV> { static __device__ __forceinline__ void run(float2 (&x)[15]) { dftComposite<3, 5, INV>(x); } };
template <bool INV> struct Dft<16, INV> { static __device__ __forceinline__ void run(float2 (&x)[16]) { dftComposite<4, 4, INV>(x); } };

}   // namespace fft2
