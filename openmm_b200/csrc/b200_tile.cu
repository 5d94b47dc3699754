// b200_tile.cu -- direct-space 32x32 tile kernel (Coulomb erfc + Lennard-Jones, exclusions, int64 forces).
//
// Replaces computeNonbonded (platforms/cuda/src/kernels/nonbonded.cu:104-524) with the pair snippet
// platforms/common/src/kernels/coulombLennardJones.cc:2-76.  Design differences (B200-first):
//   * persistent CTAs, one warp per work unit taken from a device-side queue; a dense work unit is one block X
//     plus up to 8 compacted 32-atom j tiles, so the i-force is flushed once per unit instead of once per tile;
//   * j positions/parameters are staged in shared memory and read at a rotating index (conflict free),
//     only the three j-force accumulators travel by warp shuffle (3 SHFL per pair step instead of 9);
//   * j atoms that touch only a few atoms of X go to "sparse" units instead (the reference's singlePairs,
//     nonbonded.cu:479-519): one lane per j atom walks the bits of its i-mask, so they cost a few pair evaluations
//     instead of a 32-wide row each;
//   * coordinates are re-expressed relative to the block centre once per tile ("single periodic copy" done
//     per atom, nonbonded.cu:298-301,353-358), the per-pair minimum image is only used for small boxes;
//   * erfc/exp are replaced by a rational function of r^2 (b200_ewald_coeffs.h);
//   * pairs within 3e-5 (relative, in r^2) of the cutoff are re-decided in FP64 from the stored coordinates, which
//     makes the interacting pair set identical to the Reference platform's (ReferenceNeighborList.cpp:208-214);
//   * energy is accumulated in fixed point so it is bit-reproducible.
#include "b200_internal.cuh"
#include "b200_ewald_coeffs.h"

#ifndef TILE_WARPS
#define TILE_WARPS 8
#endif
#ifndef TILE_MIN_BLOCKS
#define TILE_MIN_BLOCKS 3
#endif
#ifndef TILE_UNROLL
#define TILE_UNROLL 16
#endif
#define TILE_STR2(x) #x
#define TILE_STR(x) TILE_STR2(x)
#define TILE_THREADS (TILE_WARPS * 32)

struct TileArgs {
    const float4* __restrict__ posq;
    const float2* __restrict__ sigEps;
    const float4* __restrict__ blockCenter;
    const float4* __restrict__ blockHalf;
    const int4* __restrict__ units;
    const int* __restrict__ jAtoms;
    const int2* __restrict__ exclTiles;
    const unsigned int* __restrict__ exclFlags;
    const int* __restrict__ atomIndex;
    long long* force;
    DevCounters* counters;
    int2* pairOut; long long pairCapacity;
    int N, Npad;
    unsigned unitCapacity;
    int exclBegin, exclEnd;      // exclusion tiles handled by this rank
    BoxF box; BoxD boxD;
    float rc2, rcPadded, alpha2, alpha3, alpha, kcoul;
    double rc2d;
    float borderTol;             // |r2-rc2| below this -> FP64 decision
    float switchDist, switchInvWidth;   // LJMODE 2: switching distance, 1/(cutoff - switching distance)
    float ljAlpha2, invCut6, ljShift;   // LJPME: alpha_d^2, rc^-6, rc^-6 (1 - exp(-x)(1 + x + x^2/2)) at x = (alpha_d rc)^2
    int periodic;
    int coulombMode;                    // 0 Ewald direct space (erfc), 1 reaction field (CutoffPeriodic / CutoffNonPeriodic), 2 plain Coulomb (NoCutoff)
    float krf, crf;                     // reaction field constants (ReferenceLJCoulombIxn.cpp:70-80)
};

__device__ __forceinline__ float fast_rsqrt(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// s = -(dE/dr)/r, e = pair energy
// LJPME (coulombLennardJones.cc:30-62, ReferenceLJCoulombIxn.cpp:352-376): sprod = (sigma_i/2)(sigma_j/2), so that
// c6_i c6_j = 64 sprod^3 * eps with eps = 2 sqrt(eps_i) 2 sqrt(eps_j)
template <bool ENERGY, bool EXACT, int LJMODE = 0>
__device__ __forceinline__ void pairTerms(const TileArgs& a, float r2, float qq, float sig, float eps, float& s, float& e, float sprod = 0.f) {
    const float invR = fast_rsqrt(r2);
    const float invR2 = invR * invR;
    float sc, ec = 0.f;
    if (!EXACT) {
        const float t = a.alpha2 * r2;
        float num = fmaf(F_N6, t, F_N5); num = fmaf(num, t, F_N4); num = fmaf(num, t, F_N3); num = fmaf(num, t, F_N2);
        num = fmaf(num, t, F_N1); num = fmaf(num, t, F_N0);
        float den = fmaf(F_D5, t, F_D4); den = fmaf(den, t, F_D3); den = fmaf(den, t, F_D2); den = fmaf(den, t, F_D1);
        den = fmaf(den, t, 1.0f);
        const float Fz = num * fast_rcp(den);
        sc = qq * fmaf(-a.alpha3, Fz, invR * invR2);
        if (ENERGY) {
            float vn = fmaf(V_N6, t, V_N5); vn = fmaf(vn, t, V_N4); vn = fmaf(vn, t, V_N3); vn = fmaf(vn, t, V_N2);
            vn = fmaf(vn, t, V_N1); vn = fmaf(vn, t, V_N0);
            float vd = fmaf(V_D6, t, V_D5); vd = fmaf(vd, t, V_D4); vd = fmaf(vd, t, V_D3); vd = fmaf(vd, t, V_D2);
            vd = fmaf(vd, t, V_D1); vd = fmaf(vd, t, 1.0f);
            ec = qq * fmaf(-a.alpha, vn / vd, invR);
        }
    } else if (a.coulombMode == 1) {   // reaction field (coulombLennardJones.cc, USE_CUTOFF branch)
        sc = qq * fmaf(-2.0f * a.krf, 1.0f, invR * invR2);
        if (ENERGY) ec = qq * (invR + a.krf * r2 - a.crf);
    } else if (a.coulombMode == 2) {   // no cutoff: plain Coulomb
        sc = qq * invR * invR2;
        if (ENERGY) ec = qq * invR;
    } else {   // alpha*rc beyond the range of the fit: libm-quality erfc/exp
        const float r = r2 * invR;
        const float ar = a.alpha * r;
        const float erfcAR = erfcf(ar);
        const float ex = expf(-ar * ar);
        sc = qq * invR * invR2 * fmaf(1.1283791670955126f * ar, ex, erfcAR);
        if (ENERGY) ec = qq * invR * erfcAR;
    }
    const float sr2 = sig * sig * invR2;
    const float sr6 = sr2 * sr2 * sr2;
    const float es6 = eps * sr6;
    // LJMODE bits: 1 = LJPME (damped multiplicative r^-6 term + potential shift), 2 = switching function.  With both, the order of
    // the Reference platform is kept (ReferenceLJCoulombIxn.cpp:385-440): the switch value scales the plain LJ force, the LJPME terms
    // are added, then the switch derivative acts on the whole van der Waals energy and the switch value scales that energy.
    float sLj = es6 * invR2 * fmaf(12.0f, sr6, -6.0f), eLj = es6 * (sr6 - 1.0f);
    float sv = 1.0f, sd = 0.0f;
    if (LJMODE & 2) {
        // S(t) = 1 - 10 t^3 + 15 t^4 - 6 t^5 on t = (r - r_switch)/(r_cut - r_switch)   (coulombLennardJones.cc USE_LJ_SWITCH)
        const float r = r2 * invR;
        if (r > a.switchDist) {
            const float t = (r - a.switchDist) * a.switchInvWidth;
            sv = fmaf(t * t * t, fmaf(t, fmaf(t, -6.0f, 15.0f), -10.0f), 1.0f);
            sd = t * t * fmaf(t, fmaf(t, -30.0f, 60.0f), -30.0f) * a.switchInvWidth;
        }
        sLj *= sv;
    }
    if (LJMODE & 1) {
        const float dar2 = a.ljAlpha2 * r2, dar4 = dar2 * dar2;
        const float ex = __expf(-dar2);
        const float c6 = 64.0f * sprod * sprod * sprod * eps;
        const float coef = c6 * invR2 * invR2 * invR2;
        const float epre = 1.0f + dar2 + 0.5f * dar4;
        const float dpre = fmaf(dar4 * dar2, 1.0f / 6.0f, epre);
        sLj = fmaf(6.0f * coef * invR2, fmaf(-ex, dpre, 1.0f), sLj);
        if (ENERGY || (LJMODE & 2)) {
            const float sg2 = sig * sig, sc6 = sg2 * sg2 * sg2 * a.invCut6;
            eLj += coef * fmaf(-ex, epre, 1.0f) + eps * (1.0f - sc6) * sc6 - c6 * a.ljShift;
        }
    }
    if (LJMODE & 2) { sLj = fmaf(-eLj * sd, invR, sLj); eLj *= sv; }
    s = sc + sLj;
    if (ENERGY) e = ec + eLj;
}

// FP64 decision for pairs that fall within rounding distance of the cutoff (rare)
__device__ __forceinline__ bool borderlineInside(const TileArgs& a, int slotI, int slotJ) {
    const float4 pi = a.posq[slotI], pj = a.posq[slotJ];
    double dx = (double)pj.x - (double)pi.x, dy = (double)pj.y - (double)pi.y, dz = (double)pj.z - (double)pi.z;
    if (a.periodic) imageDeltaD(dx, dy, dz, a.boxD);
    return dx * dx + dy * dy + dz * dz <= a.rc2d;
}

// count (.x) and order-independent hash (.y) of the emitted pairs: sum over pairs of splitmix64(lo * N + hi) mod 2^64 on the
// original particle indices (the test side computes the same number from the reference's pair set)
typedef ulonglong2 PairAcc;
template <bool EMIT>
__device__ __forceinline__ void emitPair(const TileArgs& a, int slotI, int slotJ, PairAcc& acc) {
    if (!EMIT) return;
    acc.x++;
    const int oi = a.atomIndex[slotI], oj = a.atomIndex[slotJ];
    if (a.pairOut) {
        const unsigned long long slot = atomicAdd(&a.counters->pairCount, 1ull);
        if ((long long)slot < a.pairCapacity) a.pairOut[slot] = make_int2(min(oi, oj), max(oi, oj));
    } else {
        unsigned long long z = (unsigned long long)min(oi, oj) * (unsigned long long)a.N + (unsigned long long)max(oi, oj) + 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
        acc.y += z;
    }
}

struct WarpSmem {
    float4 pos[64];   // j positions (+charge), stored twice so that index lane+k never wraps
    float2 par[64];   // j (sigma/2, 2 sqrt(eps)), stored twice
    int idx[32];      // j atom slots
    // packed layout for the FFMA2 path: entry m pairs atom m with atom m^16; entries 32..47 repeat 0..15 so lane+k never wraps
    float4 xy[48];    // (x_m, x_m', y_m, y_m')
    float4 zq[48];    // (z_m, z_m', q_m, q_m')
    float4 se[48];    // (s_m, s_m', e_m, e_m')
};

// ---- packed FP32 pairs (sm_100 FFMA2/FMUL2/FADD2: two FP32 operations per issue slot; the kernel is issue bound)
__device__ __forceinline__ unsigned long long asU64(float2 v) { return *reinterpret_cast<unsigned long long*>(&v); }
__device__ __forceinline__ float2 asF2(unsigned long long v) { return *reinterpret_cast<float2*>(&v); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { unsigned long long d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(asU64(a)), "l"(asU64(b)), "l"(asU64(c))); return asF2(d); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { unsigned long long d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(asU64(a)), "l"(asU64(b))); return asF2(d); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { unsigned long long d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(asU64(a)), "l"(asU64(b))); return asF2(d); }
__device__ __forceinline__ float2 dup(float v) { return make_float2(v, v); }

// Rare: a pair whose r^2 is within rounding distance of the cutoff was decided in float inside the tile loop; re-decide it in
// FP64 and, if the decision changes, add or remove that single pair (kept out of line so the hot loop stays small).
template <bool FORCES, bool ENERGY, bool EXACT, bool EXCL, int LJMODE>
__device__ __forceinline__ void fixBorderline(const TileArgs& a, const WarpSmem& sm, int lane, unsigned borderMask, unsigned exclRot, int slotI, bool validI,
                                           float xi, float yi, float zi, float qik, float2 pri, float& fix, float& fiy, float& fiz, float& eTile) {
    while (borderMask) {
        const int k = __ffs(borderMask) - 1;
        borderMask &= borderMask - 1;
        if (EXCL && ((exclRot >> k) & 1u)) continue;
        const int jj = (lane + k) & 31;
        const int sj = sm.idx[jj];
        if (!validI || sj < 0 || sj >= a.N) continue;
        const float4 q = sm.pos[jj];
        const float2 pr = sm.par[jj];
        const float dx = q.x - xi, dy = q.y - yi, dz = q.z - zi;
        const float r2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
        const bool inF = r2 < a.rc2, inD = borderlineInside(a, slotI, sj);
        if (inF == inD) continue;
        float s, e;
        pairTerms<ENERGY, EXACT, LJMODE>(a, r2, qik * q.w, pri.x + pr.x, pri.y * pr.y, s, e, pri.x * pr.x);
        const float sg = inD ? 1.f : -1.f;
        if (FORCES) {
            const float gx = sg * dx * s, gy = sg * dy * s, gz = sg * dz * s;
            fix -= gx; fiy -= gy; fiz -= gz;
            atomicAddFixed(&a.force[sj], toFixed(gx));
            atomicAddFixed(&a.force[sj + a.Npad], toFixed(gy));
            atomicAddFixed(&a.force[sj + 2 * a.Npad], toFixed(gz));
        }
        if (ENERGY) eTile += sg * e;
    }
}

// one dense 32x32 tile: lane owns atom i = lane of block X; at step k it meets j = (lane+k)&31.
// FIXUP: decide the cutoff in float inside the loop, remember borderline pairs in a bit mask and correct them afterwards
// (single-copy force/energy variants); otherwise decide exactly inside the loop.
template <bool FORCES, bool ENERGY, bool EMIT, bool EXACT, bool SINGLE, bool EXCL, int LJMODE>
__device__ __forceinline__ void denseTile(const TileArgs& a, const WarpSmem& sm, int lane, int slotI, bool validI, float xi, float yi, float zi, float qik, float2 pri,
                                          unsigned exclBits, float& fix, float& fiy, float& fiz, float& fjx, float& fjy, float& fjz, float& eTile, PairAcc& nPairs) {
    constexpr bool FIXUP = SINGLE && !EMIT;
    const float4* __restrict__ pp = &sm.pos[lane];
    const float2* __restrict__ pq = &sm.par[lane];
    const unsigned exclRot = EXCL ? __funnelshift_r(exclBits, exclBits, lane) : 0u;   // bit k <-> j = (lane+k)&31
    unsigned borderMask = 0;
_Pragma(TILE_STR(unroll TILE_UNROLL))
    for (int k = 0; k < 32; k++) {
        const float4 q = pp[k];
        const float2 pr = pq[k];
        float dx = q.x - xi, dy = q.y - yi, dz = q.z - zi;
        bool ok = true;
        if (!SINGLE) {
            if (a.periodic) imageDelta(dx, dy, dz, a.box);
            const int sj = sm.idx[(lane + k) & 31];
            ok = validI && sj >= 0 && sj < a.N;
        }
        const float r2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
        bool in = r2 < a.rc2;
        if (FIXUP) {
            if (fabsf(r2 - a.rc2) < a.borderTol) borderMask |= 1u << k;
        } else {
            if (fabsf(r2 - a.rc2) < a.borderTol) { if (ok) in = borderlineInside(a, slotI, sm.idx[(lane + k) & 31]); }
        }
        if (EXCL) in = in && !((exclRot >> k) & 1u);
        if (!SINGLE) in = in && ok;
        float s, e;
        pairTerms<ENERGY, EXACT, LJMODE>(a, r2, qik * q.w, pri.x + pr.x, pri.y * pr.y, s, e, pri.x * pr.x);
        s = in ? s : 0.f;
        if (FORCES) {
            fix = fmaf(-dx, s, fix); fiy = fmaf(-dy, s, fiy); fiz = fmaf(-dz, s, fiz);
            fjx = fmaf(dx, s, fjx); fjy = fmaf(dy, s, fjy); fjz = fmaf(dz, s, fjz);
            const int src = (lane + 1) & 31;
            fjx = __shfl_sync(0xffffffffu, fjx, src);
            fjy = __shfl_sync(0xffffffffu, fjy, src);
            fjz = __shfl_sync(0xffffffffu, fjz, src);
        }
        if (ENERGY) eTile += in ? e : 0.f;
        if (EMIT && in) emitPair<EMIT>(a, slotI, sm.idx[(lane + k) & 31], nPairs);
    }
    if (FIXUP) {
        if (__any_sync(0xffffffffu, borderMask != 0u))
            fixBorderline<FORCES, ENERGY, EXACT, EXCL, LJMODE>(a, sm, lane, borderMask, exclRot, slotI, validI, xi, yi, zi, qik, pri, fix, fiy, fiz, eTile);
    }
}


// Borderline fix-up for the packed path: only a flag is kept in the loop, so re-scan the lane's 32 pairs (rare).
template <bool FORCES, bool ENERGY, bool EXCL, int LJMODE>
__device__ __forceinline__ void fixBorderlinePacked(const TileArgs& a, const WarpSmem& sm, int lane, unsigned exclRot, int slotI, bool validI,
                                                    float xi, float yi, float zi, float qik, float2 pri, float2& gix, float2& giy, float2& giz, float& eTile) {
    if (!validI) return;
    for (int k = 0; k < 32; k++) {
        const int m = (lane + k) & 31;
        const float4 A = sm.xy[m], B = sm.zq[m], C = sm.se[m];
        const float dx = A.x - xi, dy = A.z - yi, dz = B.x - zi;
        const float r2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
        if (!(fabsf(r2 - a.rc2) < a.borderTol)) continue;
        if (EXCL && ((exclRot >> k) & 1u)) continue;
        const int sj = sm.idx[m];
        if (sj < 0 || sj >= a.N) continue;
        const bool inF = r2 < a.rc2, inD = borderlineInside(a, slotI, sj);
        if (inF == inD) continue;
        float s, e;
        pairTerms<ENERGY, false, LJMODE>(a, r2, qik * B.z, pri.x + C.x, pri.y * C.z, s, e, pri.x * C.x);
        const float sg = inD ? 1.f : -1.f;
        if (FORCES) {
            const float gx = sg * dx * s, gy = sg * dy * s, gz = sg * dz * s;
            gix.x += gx; giy.x += gy; giz.x += gz;
            atomicAddFixed(&a.force[sj], toFixed(gx));
            atomicAddFixed(&a.force[sj + a.Npad], toFixed(gy));
            atomicAddFixed(&a.force[sj + 2 * a.Npad], toFixed(gz));
        }
        if (ENERGY) eTile += sg * e;
    }
}

// Dense 32x32 tile, two pairs per lane and step: at step k lane l meets j = (l+k)&31 and j' = j^16, all pair arithmetic in
// packed FP32x2.  gi* accumulate +d*s for atom i (the caller negates), fj*.x / fj*.y travel with j / j'.
template <bool FORCES, bool ENERGY, bool EXCL, int LJMODE>
__device__ __forceinline__ void denseTilePacked(const TileArgs& a, const WarpSmem& sm, int lane, int slotI, bool validI, float xi, float yi, float zi, float qik, float2 pri,
                                                unsigned exclBits, float2& gix, float2& giy, float2& giz, float& fjxOut, float& fjyOut, float& fjzOut, float& eTile) {
    const float4* __restrict__ pxy = &sm.xy[lane];
    const float4* __restrict__ pzq = &sm.zq[lane];
    const float4* __restrict__ pse = &sm.se[lane];
    const unsigned exclRot = EXCL ? __funnelshift_r(exclBits, exclBits, lane) : 0u;   // bit k <-> j = (lane+k)&31
    const float2 nxi = dup(-xi), nyi = dup(-yi), nzi = dup(-zi), qi2 = dup(qik), si2 = dup(pri.x), ei2 = dup(pri.y);
    const float2 al2 = dup(a.alpha2), nal3 = dup(-a.alpha3), nrc2 = dup(-a.rc2);
    float2 fjx = dup(0.f), fjy = dup(0.f), fjz = dup(0.f), e2 = dup(0.f);
    bool border = false;
    const int src = (lane + 1) & 31;
_Pragma(TILE_STR(unroll TILE_UNROLL))
    for (int k = 0; k < 16; k++) {
        const float4 A = pxy[k], B = pzq[k], C = pse[k];
        const float2 dx = add2(make_float2(A.x, A.y), nxi), dy = add2(make_float2(A.z, A.w), nyi), dz = add2(make_float2(B.x, B.y), nzi);
        const float2 r2 = fma2(dx, dx, fma2(dy, dy, mul2(dz, dz)));
        const float2 dcut = add2(r2, nrc2);
        bool in0 = dcut.x < 0.f, in1 = dcut.y < 0.f;
        border = border || (fabsf(dcut.x) < a.borderTol) || (fabsf(dcut.y) < a.borderTol);
        if (EXCL) { in0 = in0 && !((exclRot >> k) & 1u); in1 = in1 && !((exclRot >> (k + 16)) & 1u); }
        const float2 invR = make_float2(fast_rsqrt(r2.x), fast_rsqrt(r2.y));
        const float2 invR2 = mul2(invR, invR);
        const float2 t = mul2(al2, r2);
        float2 num = fma2(dup(F_N6), t, dup(F_N5)); num = fma2(num, t, dup(F_N4)); num = fma2(num, t, dup(F_N3)); num = fma2(num, t, dup(F_N2));
        num = fma2(num, t, dup(F_N1)); num = fma2(num, t, dup(F_N0));
        float2 den = fma2(dup(F_D5), t, dup(F_D4)); den = fma2(den, t, dup(F_D3)); den = fma2(den, t, dup(F_D2)); den = fma2(den, t, dup(F_D1));
        den = fma2(den, t, dup(1.0f));
        const float2 Fz = mul2(num, make_float2(fast_rcp(den.x), fast_rcp(den.y)));
        const float2 qq = mul2(qi2, make_float2(B.z, B.w));
        const float2 sc = mul2(qq, fma2(nal3, Fz, mul2(invR, invR2)));
        const float2 sig = add2(si2, make_float2(C.x, C.y));
        const float2 sr2 = mul2(mul2(sig, sig), invR2);
        const float2 sr6 = mul2(mul2(sr2, sr2), sr2);
        const float2 eps = mul2(ei2, make_float2(C.z, C.w));
        const float2 es6 = mul2(eps, sr6);
        float2 s = fma2(mul2(es6, invR2), fma2(dup(12.0f), sr6, dup(-6.0f)), sc);
        float2 eLj = dup(0.f);   // LJPME: damped multiplicative r^-6 term + potential shift (same arithmetic as pairTerms<.., LJMODE>)
        if (LJMODE == 1) {
            const float2 dar2 = mul2(dup(a.ljAlpha2), r2), dar4 = mul2(dar2, dar2);
            const float2 ex = make_float2(__expf(-dar2.x), __expf(-dar2.y));
            const float2 sp = mul2(si2, make_float2(C.x, C.y));
            const float2 c6 = mul2(mul2(dup(64.0f), mul2(mul2(sp, sp), sp)), eps);
            const float2 coef = mul2(mul2(c6, invR2), mul2(invR2, invR2));
            const float2 epre = add2(add2(dup(1.0f), dar2), mul2(dup(0.5f), dar4));
            const float2 dpre = fma2(mul2(dar4, dar2), dup(1.0f / 6.0f), epre);
            const float2 nex = make_float2(-ex.x, -ex.y);
            s = fma2(mul2(mul2(dup(6.0f), coef), invR2), fma2(nex, dpre, dup(1.0f)), s);
            if (ENERGY) {
                const float2 sg2 = mul2(sig, sig);
                const float2 sc6 = mul2(mul2(mul2(sg2, sg2), sg2), dup(a.invCut6));
                eLj = add2(fma2(coef, fma2(nex, epre, dup(1.0f)), mul2(mul2(eps, add2(dup(1.0f), make_float2(-sc6.x, -sc6.y))), sc6)), mul2(c6, dup(-a.ljShift)));
            }
        }
        s.x = in0 ? s.x : 0.f; s.y = in1 ? s.y : 0.f;
        if (ENERGY) {
            float2 vn = fma2(dup(V_N6), t, dup(V_N5)); vn = fma2(vn, t, dup(V_N4)); vn = fma2(vn, t, dup(V_N3)); vn = fma2(vn, t, dup(V_N2));
            vn = fma2(vn, t, dup(V_N1)); vn = fma2(vn, t, dup(V_N0));
            float2 vd = fma2(dup(V_D6), t, dup(V_D5)); vd = fma2(vd, t, dup(V_D4)); vd = fma2(vd, t, dup(V_D3)); vd = fma2(vd, t, dup(V_D2));
            vd = fma2(vd, t, dup(V_D1)); vd = fma2(vd, t, dup(1.0f));
            const float2 ec = mul2(qq, fma2(dup(-a.alpha), make_float2(vn.x / vd.x, vn.y / vd.y), invR));
            float2 e = fma2(es6, add2(sr6, dup(-1.0f)), ec);
            if (LJMODE == 1) e = add2(e, eLj);
            e.x = in0 ? e.x : 0.f; e.y = in1 ? e.y : 0.f;
            e2 = add2(e2, e);
        }
        if (FORCES) {
            gix = fma2(dx, s, gix); giy = fma2(dy, s, giy); giz = fma2(dz, s, giz);
            fjx = fma2(dx, s, fjx); fjy = fma2(dy, s, fjy); fjz = fma2(dz, s, fjz);
            fjx.x = __shfl_sync(0xffffffffu, fjx.x, src); fjx.y = __shfl_sync(0xffffffffu, fjx.y, src);
            fjy.x = __shfl_sync(0xffffffffu, fjy.x, src); fjy.y = __shfl_sync(0xffffffffu, fjy.y, src);
            fjz.x = __shfl_sync(0xffffffffu, fjz.x, src); fjz.y = __shfl_sync(0xffffffffu, fjz.y, src);
        }
    }
    if (ENERGY) eTile += e2.x + e2.y;
    if (FORCES) {
        // after 16 rotations the .y chain (j' = j^16) is home in lane j'; the .x chain sits 16 lanes away
        fjxOut = fjx.y + __shfl_xor_sync(0xffffffffu, fjx.x, 16);
        fjyOut = fjy.y + __shfl_xor_sync(0xffffffffu, fjy.x, 16);
        fjzOut = fjz.y + __shfl_xor_sync(0xffffffffu, fjz.x, 16);
    }
    if (__any_sync(0xffffffffu, border))
        fixBorderlinePacked<FORCES, ENERGY, EXCL, LJMODE>(a, sm, lane, exclRot, slotI, validI, xi, yi, zi, qik, pri, gix, giy, giz, eTile);
}

template <bool FORCES, bool ENERGY, bool EMIT, bool EXACT, int LJMODE>
__global__ void __launch_bounds__(TILE_THREADS, TILE_MIN_BLOCKS) tileKernel(const TileArgs a) {
    __shared__ WarpSmem smAll[TILE_WARPS];
    constexpr bool PACKED = !EMIT && !EXACT && !(LJMODE & 2);   // (the switching-function variant uses the scalar loop)   // FFMA2 path for the single-periodic-copy force/energy variants
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpSmem& sm = smAll[warp];
    const unsigned numExclUnits = (unsigned)(a.exclEnd - a.exclBegin);
    const unsigned total = numExclUnits + min(a.counters->numUnits, a.unitCapacity);   // (an overflowed build counts past its storage)
    long long eAcc = 0; PairAcc nPairs = make_ulonglong2(0ull, 0ull);

    // Work units are taken from a device-side queue.  The descriptor of the NEXT unit is fetched while the current one is
    // computed, and the lines its first loads will touch are pulled towards L1, so a unit starts with one global-latency hop
    // (the j atom data) instead of three dependent ones (descriptor -> j index -> j atom data).
    auto loadDesc = [&](unsigned w) -> int4 {
        if (w >= total) return make_int4(0, 0, 0, 0);
        if (w < numExclUnits) { const int2 t = a.exclTiles[a.exclBegin + w]; return make_int4(t.x, t.y, 1, 0); }
        return a.units[w - numExclUnits];
    };
    unsigned u = 0;
    if (lane == 0) u = atomicAdd(&a.counters->workNext, 1u);
    u = __shfl_sync(0xffffffffu, u, 0);
    int4 desc = loadDesc(u);
    while (u < total) {
        unsigned uNext = 0;
        if (lane == 0) uNext = atomicAdd(&a.counters->workNext, 1u);   // fetch ahead: latency hidden by this unit's work
        bool haveNext = false; int4 descNext = make_int4(0, 0, 0, 0);
        auto fetchNext = [&]() {
            if (haveNext) return;
            haveNext = true;
            uNext = __shfl_sync(0xffffffffu, uNext, 0);
            descNext = loadDesc(uNext);
        };
        auto hintNext = [&]() {
            if (uNext >= total) return;
            const bool ex = uNext < numExclUnits;
            const void* p0 = &a.posq[descNext.x * 32 + lane];
            const void* p1 = ex ? (const void*)&a.posq[descNext.y * 32 + lane] : (const void*)&a.jAtoms[(size_t)descNext.y * 32 + lane];
            asm volatile("prefetch.global.L1 [%0];" :: "l"(p0));
            asm volatile("prefetch.global.L1 [%0];" :: "l"(p1));
        };
        const bool excl = u < numExclUnits;
        const int x = desc.x, tileStart = excl ? 0 : desc.y, nTiles = desc.z, kind = desc.w, yBlock = desc.y;
        unsigned exclBits = 0;
        if (excl) exclBits = a.exclFlags[(size_t)(a.exclBegin + u) * 32 + lane];
        const int slotI = x * 32 + lane;
        const float4 pi = a.posq[slotI];
        const float2 pri = a.sigEps[slotI];
        const float4 cX = a.blockCenter[x];
        const float4 hX = a.blockHalf[x];
        // single periodic copy allowed when every interacting image is the one nearest to the block centre
        const bool single = a.periodic && !a.box.triclinic &&
            0.5f * a.box.ax - hX.x >= a.rcPadded && 0.5f * a.box.by - hX.y >= a.rcPadded && 0.5f * a.box.cz - hX.z >= a.rcPadded;
        const bool validI = slotI < a.N;
        float xi = pi.x, yi = pi.y, zi = pi.z;
        if (single) {
            xi -= cX.x; yi -= cX.y; zi -= cX.z;
            imageDelta(xi, yi, zi, a.box);
            if (!validI) { xi = 3.0e4f; yi = 3.0e4f; zi = 3.0e4f; }
        }
        const float qik = a.kcoul * pi.w;

        if (kind == 0) {
            // ---------------- dense unit (or one exclusion tile)
            float fix = 0.f, fiy = 0.f, fiz = 0.f, eTile = 0.f;
            float2 gix = dup(0.f), giy = dup(0.f), giz = dup(0.f);   // packed path: +sum of d*s for atom i (two partial sums)
            // two-deep software pipeline over the unit's tiles: the index of tile t+2 and the atom data of tile t+1 are in
            // flight while tile t is computed, so neither load latency is exposed (padding slots read atom 0 and are masked)
            int ja = excl ? yBlock * 32 + lane : a.jAtoms[(size_t)tileStart * 32 + lane];
            int jaNext = (nTiles > 1) ? a.jAtoms[(size_t)(tileStart + 1) * 32 + lane] : -1;
            float4 pj = a.posq[max(ja, 0)]; float2 prj = a.sigEps[max(ja, 0)];
            for (int t = 0; t < nTiles; t++) {
                const int jaCur = ja;
                if (jaCur < 0) { pj.w = 0.f; prj = make_float2(0.f, 0.f); }
                {   // stage tile t
                    float xj = pj.x, yj = pj.y, zj = pj.z;
                    if (single) {
                        xj -= cX.x; yj -= cX.y; zj -= cX.z;
                        imageDelta(xj, yj, zj, a.box);
                        if (jaCur < 0 || jaCur >= a.N) { xj = -3.0e4f; yj = -3.0e4f; zj = -3.0e4f; }
                    }
                    __syncwarp();
                    if (PACKED && single) {
                        const float xo = __shfl_xor_sync(0xffffffffu, xj, 16), yo = __shfl_xor_sync(0xffffffffu, yj, 16), zo = __shfl_xor_sync(0xffffffffu, zj, 16);
                        const float qo = __shfl_xor_sync(0xffffffffu, pj.w, 16), so = __shfl_xor_sync(0xffffffffu, prj.x, 16), eo = __shfl_xor_sync(0xffffffffu, prj.y, 16);
                        const float4 A = make_float4(xj, xo, yj, yo), B = make_float4(zj, zo, pj.w, qo), C = make_float4(prj.x, so, prj.y, eo);
                        sm.xy[lane] = A; sm.zq[lane] = B; sm.se[lane] = C;
                        if (lane < 16) { sm.xy[lane + 32] = A; sm.zq[lane + 32] = B; sm.se[lane + 32] = C; }
                    } else {
                        sm.pos[lane] = sm.pos[lane + 32] = make_float4(xj, yj, zj, pj.w);
                        sm.par[lane] = sm.par[lane + 32] = prj;
                    }
                    sm.idx[lane] = jaCur;
                    __syncwarp();
                }
                if (t + 1 < nTiles) {   // next tile's loads fly while this one is computed
                    ja = jaNext;
                    pj = a.posq[max(ja, 0)]; prj = a.sigEps[max(ja, 0)];
                    if (t + 2 < nTiles) jaNext = a.jAtoms[(size_t)(tileStart + t + 2) * 32 + lane];
                }
                if (t == 0) fetchNext();                         // the queue atomic has long returned by now
                if (t == nTiles - 1) hintNext();
                float fjx = 0.f, fjy = 0.f, fjz = 0.f;
                if (PACKED && single) {
                    if (excl) denseTilePacked<FORCES, ENERGY, true, LJMODE>(a, sm, lane, slotI, validI, xi, yi, zi, qik, pri, exclBits, gix, giy, giz, fjx, fjy, fjz, eTile);
                    else denseTilePacked<FORCES, ENERGY, false, LJMODE>(a, sm, lane, slotI, validI, xi, yi, zi, qik, pri, 0u, gix, giy, giz, fjx, fjy, fjz, eTile);
                } else if (excl) {
                    if (single) denseTile<FORCES, ENERGY, EMIT, EXACT, true, true, LJMODE>(a, sm, lane, slotI, validI, xi, yi, zi, qik, pri, exclBits, fix, fiy, fiz, fjx, fjy, fjz, eTile, nPairs);
                    else denseTile<FORCES, ENERGY, EMIT, EXACT, false, true, LJMODE>(a, sm, lane, slotI, validI, xi, yi, zi, qik, pri, exclBits, fix, fiy, fiz, fjx, fjy, fjz, eTile, nPairs);
                } else {
                    if (single) denseTile<FORCES, ENERGY, EMIT, EXACT, true, false, LJMODE>(a, sm, lane, slotI, validI, xi, yi, zi, qik, pri, 0u, fix, fiy, fiz, fjx, fjy, fjz, eTile, nPairs);
                    else denseTile<FORCES, ENERGY, EMIT, EXACT, false, false, LJMODE>(a, sm, lane, slotI, validI, xi, yi, zi, qik, pri, 0u, fix, fiy, fiz, fjx, fjy, fjz, eTile, nPairs);
                }
                if (FORCES && jaCur >= 0 && jaCur < a.N) {
                    // after 32 rotations the accumulator for j = lane is back in this lane
                    atomicAddFixed(&a.force[jaCur], toFixed(fjx));
                    atomicAddFixed(&a.force[jaCur + a.Npad], toFixed(fjy));
                    atomicAddFixed(&a.force[jaCur + 2 * a.Npad], toFixed(fjz));
                }
                if (ENERGY) {   // per-tile float sum, reduced in a fixed order, then integer accumulation
                    const float es = warpSum(eTile);
                    eAcc += toFixed(es);
                    eTile = 0.f;
                }
            }
            if (PACKED) { fix -= gix.x + gix.y; fiy -= giy.x + giy.y; fiz -= giz.x + giz.y; }
            if (FORCES && validI) {
                atomicAddFixed(&a.force[slotI], toFixed(fix));
                atomicAddFixed(&a.force[slotI + a.Npad], toFixed(fiy));
                atomicAddFixed(&a.force[slotI + 2 * a.Npad], toFixed(fiz));
            }
        } else {
            // ---------------- sparse unit: block X in shared memory, one lane per j atom, loop over the bits of its i-mask
            __syncwarp();
            sm.pos[lane] = make_float4(xi, yi, zi, qik);
            sm.par[lane] = pri;
            __syncwarp();
            for (int t = 0; t < nTiles; t++) {
                if (t == nTiles - 1) { fetchNext(); hintNext(); }
                const size_t base = (size_t)(tileStart + 2 * t) * 32;
                const int ja = a.jAtoms[base + lane];
                unsigned mask = ja >= 0 ? (unsigned)a.jAtoms[base + 32 + lane] : 0u;
                float4 pj = make_float4(0.f, 0.f, 0.f, 0.f); float2 prj = make_float2(0.f, 0.f);
                if (ja >= 0) { pj = a.posq[ja]; prj = a.sigEps[ja]; }
                float xj = pj.x, yj = pj.y, zj = pj.z;
                if (single) { xj -= cX.x; yj -= cX.y; zj -= cX.z; imageDelta(xj, yj, zj, a.box); }
                float fjx = 0.f, fjy = 0.f, fjz = 0.f, eLane = 0.f;
                while (mask) {
                    const int i = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const float4 p = sm.pos[i];
                    const float2 pr = sm.par[i];
                    float dx = xj - p.x, dy = yj - p.y, dz = zj - p.z;
                    if (!single && a.periodic) imageDelta(dx, dy, dz, a.box);
                    const float r2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
                    const float dcut = r2 - a.rc2;
                    bool in = dcut < 0.f;
                    if (fabsf(dcut) < a.borderTol) in = borderlineInside(a, x * 32 + i, ja);
                    float s, e;
                    pairTerms<ENERGY, EXACT, LJMODE>(a, r2, p.w * pj.w, pr.x + prj.x, pr.y * prj.y, s, e, pr.x * prj.x);
                    s = in ? s : 0.f;
                    if (FORCES && in) {
                        const float gx = dx * s, gy = dy * s, gz = dz * s;
                        fjx += gx; fjy += gy; fjz += gz;
                        const int si = x * 32 + i;
                        atomicAddFixed(&a.force[si], toFixed(-gx));
                        atomicAddFixed(&a.force[si + a.Npad], toFixed(-gy));
                        atomicAddFixed(&a.force[si + 2 * a.Npad], toFixed(-gz));
                    }
                    if (ENERGY && in) eLane += e;
                    if (EMIT && in) emitPair<EMIT>(a, x * 32 + i, ja, nPairs);
                }
                if (FORCES && ja >= 0) {
                    atomicAddFixed(&a.force[ja], toFixed(fjx));
                    atomicAddFixed(&a.force[ja + a.Npad], toFixed(fjy));
                    atomicAddFixed(&a.force[ja + 2 * a.Npad], toFixed(fjz));
                }
                if (ENERGY) eAcc += toFixed(warpSum(eLane));
            }
        }
        fetchNext();
        u = uNext; desc = descNext;
    }
    if (ENERGY && lane == 0 && eAcc != 0) atomicAddFixed(&a.counters->energy[E_DIRECT], eAcc);
    if (EMIT && !a.pairOut) {
        for (int o = 16; o > 0; o >>= 1) { nPairs.x += __shfl_xor_sync(0xffffffffu, nPairs.x, o); nPairs.y += __shfl_xor_sync(0xffffffffu, nPairs.y, o); }
        if (lane == 0 && nPairs.x) { atomicAdd(&a.counters->pairCount, nPairs.x); atomicAdd(&a.counters->pairHash, nPairs.y); }
    }
}


template <bool EXACT, int LJMODE>
static void launchTile(b200_ctx* c, const TileArgs& a, bool forces, bool energy, bool emitPairs) {
    const int grid = c->smCount * TILE_MIN_BLOCKS;
    if (emitPairs) tileKernel<false, false, true, EXACT, 0><<<grid, TILE_THREADS, 0, c->stream>>>(a);
    else if (forces && energy) tileKernel<true, true, false, EXACT, LJMODE><<<grid, TILE_THREADS, 0, c->stream>>>(a);
    else if (forces) tileKernel<true, false, false, EXACT, LJMODE><<<grid, TILE_THREADS, 0, c->stream>>>(a);
    else tileKernel<false, true, false, EXACT, LJMODE><<<grid, TILE_THREADS, 0, c->stream>>>(a);
}

void b200_tile_launch(b200_ctx* c, bool forces, bool energy, bool emitPairs) {
    TileArgs a;
    a.posq = c->posq; a.sigEps = c->sigEps; a.blockCenter = c->blockCenter; a.blockHalf = c->blockHalf;
    a.units = c->units; a.jAtoms = c->jAtoms; a.exclTiles = c->exclTiles; a.exclFlags = c->exclFlags;
    a.atomIndex = c->atomIndex; a.force = c->force; a.counters = c->counters;
    a.pairOut = c->pairOut; a.pairCapacity = c->pairCapacity;
    a.N = c->N; a.Npad = c->Npad; a.unitCapacity = (unsigned)c->unitCapacity;
    // exclusion tiles are split by index over ranks (CudaNonbondedUtilities.cpp:704-708)
    a.exclBegin = (int)llround(c->numExclTiles * c->partShareBegin);
    a.exclEnd = (int)llround(c->numExclTiles * c->partShareEnd);
    a.box = c->boxF; a.boxD = c->boxD;
    const double rc = c->nb.cutoff, al = c->nb.ewald_alpha;
    a.rc2 = (float)(rc * rc); a.rc2d = rc * rc; a.rcPadded = (float)c->paddedCutoff;
    a.alpha = (float)al; a.alpha2 = (float)(al * al); a.alpha3 = (float)(al * al * al);
    a.kcoul = (float)B200_ONE_4PI_EPS0;
    a.borderTol = (float)(3e-5 * rc * rc);
    // NoCutoff / CutoffNonPeriodic: no periodic images; NoCutoff: every non-excluded pair interacts
    const int method = c->nb.method;
    a.periodic = (method == B200_NB_NOCUTOFF || method == B200_NB_CUTOFF_NONPERIODIC) ? 0 : 1;
    a.coulombMode = method == B200_NB_NOCUTOFF ? 2 : (method == B200_NB_CUTOFF_NONPERIODIC || method == B200_NB_CUTOFF_PERIODIC) ? 1 : 0;
    a.krf = a.crf = 0.f;
    if (a.coulombMode == 1) {
        const double eps = c->nb.reaction_field_dielectric;
        a.krf = (float)(pow(rc, -3.0) * (eps - 1.0) / (2.0 * eps + 1.0)); a.crf = (float)((1.0 / rc) * 3.0 * eps / (2.0 * eps + 1.0));
    }
    if (method == B200_NB_NOCUTOFF) { a.rc2 = INFINITY; a.rc2d = INFINITY; a.rcPadded = INFINITY; a.borderTol = 0.f; }
    // (the work queue head was rewound by nbBegin, the first kernel of b200_nblist_prepare, which precedes every tile launch)
    ScopedTimer tm(c, B200_T_TILE, c->stream);
    a.ljAlpha2 = a.invCut6 = a.ljShift = 0.f;
    if (c->ljpme) {
        const double x = c->nb.ljpme_alpha * c->nb.ljpme_alpha * rc * rc, ic6 = 1.0 / (rc * rc * rc * rc * rc * rc);
        a.ljAlpha2 = (float)(c->nb.ljpme_alpha * c->nb.ljpme_alpha); a.invCut6 = (float)ic6;
        a.ljShift = (float)(ic6 * (1.0 - exp(-x) * (1.0 + x + 0.5 * x * x)));
    }
    a.switchDist = 0.f; a.switchInvWidth = 0.f;
    if (c->nb.use_switch) { a.switchDist = (float)c->nb.switch_distance; a.switchInvWidth = (float)(1.0 / (rc - c->nb.switch_distance)); }
    if (c->ljpme && c->nb.use_switch) launchTile<true, 3>(c, a, forces, energy, emitPairs);   // (scalar loop either way; one instantiation)
    else if (c->ljpme) { if (c->exactEwald) launchTile<true, 1>(c, a, forces, energy, emitPairs); else launchTile<false, 1>(c, a, forces, energy, emitPairs); }
    else if (c->nb.use_switch) { if (c->exactEwald) launchTile<true, 2>(c, a, forces, energy, emitPairs); else launchTile<false, 2>(c, a, forces, energy, emitPairs); }
    else if (c->exactEwald) launchTile<true, 0>(c, a, forces, energy, emitPairs);
    else launchTile<false, 0>(c, a, forces, energy, emitPairs);
    LAUNCH_COUNT(c, 1);
    CUDA_CHECK(cudaGetLastError());
}
