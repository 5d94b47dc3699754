/*
 * nonbonded_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, double-precision CPU restatement of the reference's NonbondedForce PME algorithm, used only by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg to CHECK the CUDA path.  Nothing under openmm_b200/ may
 * import, link or call it.
 *
 * Parity is PINNED: tests/test_oracle.py checks this file against (a) the Gromacs known answers the reference keeps in
 * tests/TestEwald.h:220-269 (testTriclinic), (b) the Madelung constant of tests/TestEwald.h:47-94 and (c) forces and
 * energies produced here by the reference's own Reference platform compiled into oracle/_ref (fixtures in
 * tests/golden/, generator tests/golden/make_fixtures.py + oracle/ref_eval.cpp).
 *
 * Each function names the reference code it restates (paths relative to the OpenMM 8.6 tree):
 *   oracle_direct               platforms/reference/src/SimTKReference/ReferenceLJCoulombIxn.cpp:379-456 (pair loop) with the
 *                               pair set of ReferenceNeighborList.cpp:208-214 (d^2 <= rc^2, not excluded)
 *   oracle_exclusion_correction ReferenceLJCoulombIxn.cpp:462-520
 *   oracle_exceptions           platforms/reference/src/SimTKReference/ReferenceLJCoulomb14.cpp (plain Coulomb + LJ)
 *   oracle_reciprocal           platforms/reference/src/SimTKReference/ReferencePME.cpp:166-201 (B-splines), :203-259 (spread),
 *                               :261-333 (convolution), :413-476 (interpolation), :534-591 (exec)
 *   oracle_self_energy          ReferenceLJCoulombIxn.cpp:217-238
 */
#include "nonbonded_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define ONE_4PI_EPS0 138.93545764438198
#define PME_ORDER 5

/* ---- periodic displacement in a reduced triclinic cell (ReferenceForce::getDeltaRPeriodic, ReferenceForce.cpp) */
static void delta_periodic(const double* a, const double* b, const double* box, double* d) {
    d[0] = b[0] - a[0]; d[1] = b[1] - a[1]; d[2] = b[2] - a[2];
    double s = floor(d[2] / box[8] + 0.5);
    d[0] -= s * box[6]; d[1] -= s * box[7]; d[2] -= s * box[8];
    s = floor(d[1] / box[4] + 0.5);
    d[0] -= s * box[3]; d[1] -= s * box[4];
    s = floor(d[0] / box[0] + 0.5);
    d[0] -= s * box[0];
}

static int cmp_pair(const void* a, const void* b) {
    const int* x = (const int*)a; const int* y = (const int*)b;
    if (x[0] != y[0]) return x[0] < y[0] ? -1 : 1;
    return x[1] < y[1] ? -1 : (x[1] > y[1] ? 1 : 0);
}

long long oracle_direct(const oracle_system* s, double* forces, double* energy, int* pairs, long long capacity) {
    const int n = s->n;
    const double rc2 = s->cutoff * s->cutoff;
    /* sorted exclusion list for lookup */
    int* ex = (int*)malloc(sizeof(int) * 2 * (size_t)(s->nexcl > 0 ? s->nexcl : 1));
    for (int i = 0; i < s->nexcl; i++) {
        int a = s->excl[2 * i], b = s->excl[2 * i + 1];
        ex[2 * i] = a < b ? a : b; ex[2 * i + 1] = a < b ? b : a;
    }
    qsort(ex, (size_t)s->nexcl, 2 * sizeof(int), cmp_pair);
    /* cell list over fractional coordinates (any algorithm that yields the same pair set is a valid restatement of
       computeNeighborListVoxelHash; the set is defined by d^2 <= rc^2 with the minimum-image displacement) */
    const double* box = s->box;
    int nc[3];
    nc[0] = (int)floor(box[0] / s->cutoff); nc[1] = (int)floor(box[4] / s->cutoff); nc[2] = (int)floor(box[8] / s->cutoff);
    const int triclinic = box[3] != 0 || box[6] != 0 || box[7] != 0;
    for (int k = 0; k < 3; k++) { if (nc[k] < 3 || triclinic) nc[k] = 1; if (nc[k] > 64) nc[k] = 64; }
    const int ncell = nc[0] * nc[1] * nc[2];
    int* cellOf = (int*)malloc(sizeof(int) * (size_t)n);
    int* start = (int*)calloc((size_t)ncell + 1, sizeof(int));
    int* order = (int*)malloc(sizeof(int) * (size_t)n);
    for (int i = 0; i < n; i++) {
        int c[3];
        const double L[3] = {box[0], box[4], box[8]};
        for (int k = 0; k < 3; k++) {
            double f = s->pos[3 * i + k] / L[k]; f -= floor(f);
            c[k] = (int)(f * nc[k]); if (c[k] >= nc[k]) c[k] = nc[k] - 1;
        }
        cellOf[i] = (c[0] * nc[1] + c[1]) * nc[2] + c[2];
        start[cellOf[i] + 1]++;
    }
    for (int c = 0; c < ncell; c++) start[c + 1] += start[c];
    {
        int* fill = (int*)malloc(sizeof(int) * (size_t)ncell);
        memcpy(fill, start, sizeof(int) * (size_t)ncell);
        for (int i = 0; i < n; i++) order[fill[cellOf[i]]++] = i;
        free(fill);
    }
    long long npairs = 0;
    double esum = 0;
    for (int ci = 0; ci < ncell; ci++) {
        const int cx = ci / (nc[1] * nc[2]), cy = (ci / nc[2]) % nc[1], cz = ci % nc[2];
        /* visit each unordered pair of neighbouring cells once */
        int seen[27], nseen = 0;
        for (int dx = -1; dx <= 1; dx++) for (int dy = -1; dy <= 1; dy++) for (int dz = -1; dz <= 1; dz++) {
            if ((nc[0] == 1 && dx != 0) || (nc[1] == 1 && dy != 0) || (nc[2] == 1 && dz != 0)) continue;
            const int ox = (cx + dx + nc[0]) % nc[0], oy = (cy + dy + nc[1]) % nc[1], oz = (cz + dz + nc[2]) % nc[2];
            const int cj = (ox * nc[1] + oy) * nc[2] + oz;
            if (cj < ci) continue;
            int dup = 0;
            for (int k = 0; k < nseen; k++) if (seen[k] == cj) dup = 1;
            if (dup) continue;
            seen[nseen++] = cj;
            for (int a = start[ci]; a < start[ci + 1]; a++) {
                const int i = order[a];
                for (int b = (cj == ci ? a + 1 : start[cj]); b < start[cj + 1]; b++) {
                    const int j = order[b];
                    double d[3];
                    delta_periodic(&s->pos[3 * i], &s->pos[3 * j], box, d);   /* d = r_j - r_i */
                    const double r2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
                    if (r2 > rc2) continue;
                    int key[2] = {i < j ? i : j, i < j ? j : i};
                    if (s->nexcl > 0 && bsearch(key, ex, (size_t)s->nexcl, 2 * sizeof(int), cmp_pair)) continue;
                    if (pairs && npairs < capacity) { pairs[2 * npairs] = key[0]; pairs[2 * npairs + 1] = key[1]; }
                    npairs++;
                    const double r = sqrt(r2), invR = 1.0 / r, ar = s->alpha * r;
                    const double qq = ONE_4PI_EPS0 * s->charge[i] * s->charge[j];
                    double dEdR = qq * invR * invR * invR * (erfc(ar) + 2.0 * ar * exp(-ar * ar) / sqrt(M_PI));
                    const double sig = 0.5 * (s->sigma[i] + s->sigma[j]);
                    const double eps = 4.0 * sqrt(s->epsilon[i] * s->epsilon[j]);
                    double sig2 = invR * sig; sig2 *= sig2;
                    const double sig6 = sig2 * sig2 * sig2;
                    /* switching function on the Lennard-Jones part, ReferenceLJCoulombIxn.cpp:385-390,436-439 */
                    double switchValue = 1.0, switchDeriv = 0.0;
                    if (s->use_switch && r > s->switch_distance) {
                        const double w = s->cutoff - s->switch_distance, t = (r - s->switch_distance) / w;
                        switchValue = 1.0 + t * t * t * (-10.0 + t * (15.0 - t * 6.0));
                        switchDeriv = t * t * (-30.0 + t * (60.0 - t * 30.0)) / w;
                    }
                    dEdR += switchValue * eps * (12.0 * sig6 - 6.0) * sig6 * invR * invR;
                    double vdw = eps * (sig6 - 1.0) * sig6;
                    esum += qq * invR * erfc(ar);
                    if (s->ljpme) {
                        /* ReferenceLJCoulombIxn.cpp:404-431: remove the multiplicative C6 term the dispersion grid adds, plus the
                           potential shift between the additive and multiplicative forms at the cutoff */
                        const double dar2 = s->lj_alpha * s->lj_alpha * r2, dar4 = dar2 * dar2, dar6 = dar4 * dar2;
                        const double invR2 = invR * invR;
                        const double c6i = 8.0 * pow(0.5 * s->sigma[i], 3.0) * 2.0 * sqrt(s->epsilon[i]);
                        const double c6j = 8.0 * pow(0.5 * s->sigma[j], 3.0) * 2.0 * sqrt(s->epsilon[j]);
                        const double emult = c6i * c6j * invR2 * invR2 * invR2 * (1.0 - exp(-dar2) * (1.0 + dar2 + 0.5 * dar4));
                        dEdR += 6.0 * c6i * c6j * invR2 * invR2 * invR2 * invR2 * (1.0 - exp(-dar2) * (1.0 + dar2 + 0.5 * dar4 + dar6 / 6.0));
                        const double invCut2 = 1.0 / rc2, invCut6 = invCut2 * invCut2 * invCut2;
                        double s2 = sig * sig; const double s6 = s2 * s2 * s2;
                        double shift = eps * (1.0 - s6 * invCut6) * s6 * invCut6;
                        const double dc2 = s->lj_alpha * s->lj_alpha * rc2, dc4 = dc2 * dc2;
                        shift -= c6i * c6j * invCut6 * (1.0 - exp(-dc2) * (1.0 + dc2 + 0.5 * dc4));
                        vdw += emult + shift;
                    }
                    if (s->use_switch) { dEdR -= vdw * switchDeriv * invR; vdw *= switchValue; }
                    esum += vdw;
                    if (forces) for (int k = 0; k < 3; k++) { forces[3 * i + k] -= dEdR * d[k]; forces[3 * j + k] += dEdR * d[k]; }
                }
            }
        }
    }
    if (energy) *energy += esum;
    free(ex); free(cellOf); free(start); free(order);
    return npairs;
}

void oracle_exclusion_correction(const oracle_system* s, double* forces, double* energy) {
    double esum = 0;
    for (int e = 0; e < s->nexcl; e++) {
        const int i = s->excl[2 * e], j = s->excl[2 * e + 1];
        if (i == j) continue;
        double d[3];
        if (s->exceptions_periodic) delta_periodic(&s->pos[3 * i], &s->pos[3 * j], s->box, d);
        else for (int k = 0; k < 3; k++) d[k] = s->pos[3 * j + k] - s->pos[3 * i + k];
        const double r = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]), invR = 1.0 / r, ar = s->alpha * r;
        const double qq = ONE_4PI_EPS0 * s->charge[i] * s->charge[j];
        if (erf(ar) > 1e-6) {
            const double dEdR = qq * invR * invR * invR * (erf(ar) - 2.0 * ar * exp(-ar * ar) / sqrt(M_PI));
            if (forces) for (int k = 0; k < 3; k++) { forces[3 * i + k] += dEdR * d[k]; forces[3 * j + k] -= dEdR * d[k]; }
            esum += qq * invR * erf(ar);
        } else {
            esum += s->alpha * 2.0 / sqrt(M_PI) * qq;
        }
        if (s->ljpme) {
            /* ReferenceLJCoulombIxn.cpp:497-513: back out the reciprocal-space dispersion of the excluded pair */
            const double dar2 = s->lj_alpha * s->lj_alpha * r * r, dar4 = dar2 * dar2, dar6 = dar4 * dar2, invR2 = invR * invR;
            const double c6i = 8.0 * pow(0.5 * s->sigma[i], 3.0) * 2.0 * sqrt(s->epsilon[i]);
            const double c6j = 8.0 * pow(0.5 * s->sigma[j], 3.0) * 2.0 * sqrt(s->epsilon[j]);
            esum -= c6i * c6j * invR2 * invR2 * invR2 * (1.0 - exp(-dar2) * (1.0 + dar2 + 0.5 * dar4));
            const double dEdR = -6.0 * c6i * c6j * invR2 * invR2 * invR2 * invR2 * (1.0 - exp(-dar2) * (1.0 + dar2 + 0.5 * dar4 + dar6 / 6.0));
            if (forces) for (int k = 0; k < 3; k++) { forces[3 * i + k] += dEdR * d[k]; forces[3 * j + k] -= dEdR * d[k]; }
        }
    }
    if (energy) *energy -= esum;
}

void oracle_exceptions(const oracle_system* s, double* forces, double* energy) {
    double esum = 0;
    for (int e = 0; e < s->nexc; e++) {
        const int i = s->exc[2 * e], j = s->exc[2 * e + 1];
        double d[3];
        if (s->exceptions_periodic) delta_periodic(&s->pos[3 * i], &s->pos[3 * j], s->box, d);
        else for (int k = 0; k < 3; k++) d[k] = s->pos[3 * j + k] - s->pos[3 * i + k];
        const double r2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2], invR = 1.0 / sqrt(r2);
        const double qq = ONE_4PI_EPS0 * s->exc_qq[e];
        double sig2 = invR * s->exc_sigma[e]; sig2 *= sig2;
        const double sig6 = sig2 * sig2 * sig2, eps = 4.0 * s->exc_eps[e];
        const double dEdR = (qq * invR + eps * (12.0 * sig6 - 6.0) * sig6) * invR * invR;
        esum += qq * invR + eps * (sig6 - 1.0) * sig6;
        if (forces) for (int k = 0; k < 3; k++) { forces[3 * i + k] -= dEdR * d[k]; forces[3 * j + k] += dEdR * d[k]; }
    }
    if (energy) *energy += esum;
}

double oracle_self_energy(const oracle_system* s) {
    double q2 = 0, q = 0;
    for (int i = 0; i < s->n; i++) { q2 += s->charge[i] * s->charge[i]; q += s->charge[i]; }
    const double V = s->box[0] * s->box[4] * s->box[8];
    const double eps0 = 1.0 / (4.0 * M_PI * ONE_4PI_EPS0);
    double e = -ONE_4PI_EPS0 * q2 * s->alpha / sqrt(M_PI) - q * q / (8.0 * eps0 * V * s->alpha * s->alpha);
    if (s->ljpme)   /* dispersion self term, ReferenceLJCoulombIxn.cpp:224-227 */
        for (int i = 0; i < s->n; i++) e += pow(s->lj_alpha, 6.0) * 64.0 * pow(0.5 * s->sigma[i], 6.0) * 4.0 * s->epsilon[i] / 12.0;
    return e;
}

/* ---- FFT: recursive mixed radix (any n; O(n * sum of prime factors)) ------------------------------------------ */
static void fft_rec(const double* in, double* out, int n, int stride, int sign) {
    if (n == 1) { out[0] = in[0]; out[1] = in[1]; return; }
    int p = 2;
    while (n % p) p++;
    const int m = n / p;
    double* sub = (double*)malloc(sizeof(double) * 2 * (size_t)n);
    for (int r = 0; r < p; r++) fft_rec(in + 2 * (size_t)r * stride, sub + 2 * (size_t)r * m, m, stride * p, sign);
    for (int k = 0; k < n; k++) {
        double re = 0, im = 0;
        const int km = k % m;
        for (int r = 0; r < p; r++) {
            const double ang = sign * 2.0 * M_PI * (double)((long long)r * k % n) / n;
            const double c = cos(ang), sn = sin(ang);
            const double xr = sub[2 * ((size_t)r * m + km)], xi = sub[2 * ((size_t)r * m + km) + 1];
            re += xr * c - xi * sn; im += xr * sn + xi * c;
        }
        out[2 * k] = re; out[2 * k + 1] = im;
    }
    free(sub);
}

void oracle_fft1d(double* data, int n, int stride, int inverse) {
    double* out = (double*)malloc(sizeof(double) * 2 * (size_t)n);
    fft_rec(data, out, n, stride, inverse ? 1 : -1);
    for (int k = 0; k < n; k++) { data[2 * (size_t)k * stride] = out[2 * k]; data[2 * (size_t)k * stride + 1] = out[2 * k + 1]; }
    free(out);
}

static void fft3d(double* g, int nx, int ny, int nz, int inverse) {
    for (int x = 0; x < nx; x++) for (int y = 0; y < ny; y++) oracle_fft1d(g + 2 * ((size_t)x * ny + y) * nz, nz, 1, inverse);
    for (int x = 0; x < nx; x++) for (int z = 0; z < nz; z++) oracle_fft1d(g + 2 * ((size_t)x * ny * nz + z), ny, nz, inverse);
    for (int y = 0; y < ny; y++) for (int z = 0; z < nz; z++) oracle_fft1d(g + 2 * ((size_t)y * nz + z), nx, ny * nz, inverse);
}

void oracle_fft3d_r2c(const double* in, double* out, int nx, int ny, int nz) {
    const size_t G = (size_t)nx * ny * nz;
    double* g = (double*)calloc(2 * G, sizeof(double));
    for (size_t i = 0; i < G; i++) g[2 * i] = in[i];
    fft3d(g, nx, ny, nz, 0);
    const int nzc = nz / 2 + 1;
    for (int x = 0; x < nx; x++) for (int y = 0; y < ny; y++) for (int z = 0; z < nzc; z++) {
        const size_t src = ((size_t)x * ny + y) * nz + z, dst = ((size_t)x * ny + y) * nzc + z;
        out[2 * dst] = g[2 * src]; out[2 * dst + 1] = g[2 * src + 1];
    }
    free(g);
}

/* ---- smooth PME ------------------------------------------------------------------------------------------------ */
static void bspline(double u, double* th, double* dth) {
    /* Essmann et al. recursion, ReferencePME.cpp:166-201 */
    th[PME_ORDER - 1] = 0; th[1] = u; th[0] = 1 - u;
    for (int k = 3; k < PME_ORDER; k++) {
        const double div = 1.0 / (k - 1.0);
        th[k - 1] = div * u * th[k - 2];
        for (int l = 1; l < k - 1; l++) th[k - l - 1] = div * ((u + l) * th[k - l - 2] + (k - l - u) * th[k - l - 1]);
        th[0] = div * (1 - u) * th[0];
    }
    dth[0] = -th[0];
    for (int k = 1; k < PME_ORDER; k++) dth[k] = th[k - 1] - th[k];
    const double div = 1.0 / (PME_ORDER - 1);
    th[PME_ORDER - 1] = div * u * th[PME_ORDER - 2];
    for (int l = 1; l < PME_ORDER - 1; l++) th[PME_ORDER - l - 1] = div * ((u + l) * th[PME_ORDER - l - 2] + (PME_ORDER - l - u) * th[PME_ORDER - l - 1]);
    th[0] = div * (1 - u) * th[0];
}

static void moduli(int n, double* mod) {
    /* ReferencePME.cpp pme_calculate_bsplines_moduli: |DFT of the spline sampled at the integers|^2 */
    double th[PME_ORDER], dth[PME_ORDER];
    bspline(0.0, th, dth);
    double* data = (double*)calloc((size_t)n, sizeof(double));
    for (int i = 0; i < PME_ORDER; i++) data[(i + 1) % n] += th[i];   /* n >= 6 in practice */
    for (int k = 0; k < n; k++) {
        double sc = 0, ss = 0;
        for (int j = 0; j < n; j++) { const double arg = 2.0 * M_PI * k * j / n; sc += data[j] * cos(arg); ss += data[j] * sin(arg); }
        mod[k] = sc * sc + ss * ss;
    }
    for (int k = 0; k < n; k++) if (mod[k] < 1e-7) mod[k] = 0.5 * (mod[(k + n - 1) % n] + mod[(k + 1) % n]);
    free(data);
}

/* one smooth-PME reciprocal sum: Coulomb (dispersion = 0; ReferencePME::exec) or dispersion (dispersion = 1; exec_dpme,
   ReferencePME.cpp:336-410,594-674) with the given per-particle coefficients, Ewald coefficient and mesh */
static void pme_sum(const oracle_system* s, const double* coef, double alpha, const int* mesh, int dispersion, double* forces, double* energy) {
    const int n = s->n, nx = mesh[0], ny = mesh[1], nz = mesh[2];
    const double* b = s->box;
    /* reciprocal box vectors (ReferencePME.cpp invert_box_vectors) */
    const double det = b[0] * b[4] * b[8], sc = 1.0 / det;
    const double R[3][3] = {{b[4] * b[8] * sc, 0, 0}, {-b[3] * b[8] * sc, b[0] * b[8] * sc, 0},
                            {(b[3] * b[7] - b[4] * b[6]) * sc, -b[0] * b[7] * sc, b[0] * b[4] * sc}};
    const size_t G = (size_t)nx * ny * nz;
    double* grid = (double*)calloc(2 * G, sizeof(double));
    int* idx = (int*)malloc(sizeof(int) * 3 * (size_t)n);
    double* th = (double*)malloc(sizeof(double) * 3 * PME_ORDER * (size_t)n);
    double* dth = (double*)malloc(sizeof(double) * 3 * PME_ORDER * (size_t)n);
    const int ng[3] = {nx, ny, nz};
    for (int i = 0; i < n; i++) {
        const double* p = &s->pos[3 * i];
        for (int d = 0; d < 3; d++) {
            double t = p[0] * R[0][d] + p[1] * R[1][d] + p[2] * R[2][d];
            t = (t - floor(t)) * ng[d];
            const int ti = (int)t;
            idx[3 * i + d] = ti % ng[d];
            bspline(t - ti, &th[(3 * (size_t)i + d) * PME_ORDER], &dth[(3 * (size_t)i + d) * PME_ORDER]);
        }
    }
    for (int i = 0; i < n; i++) {
        const double q = coef[i];
        const double *tx = &th[(3 * (size_t)i) * PME_ORDER], *ty = tx + PME_ORDER, *tz = ty + PME_ORDER;
        for (int ix = 0; ix < PME_ORDER; ix++) { const int x = (idx[3 * i] + ix) % nx;
            for (int iy = 0; iy < PME_ORDER; iy++) { const int y = (idx[3 * i + 1] + iy) % ny;
                for (int iz = 0; iz < PME_ORDER; iz++) { const int z = (idx[3 * i + 2] + iz) % nz;
                    grid[2 * (((size_t)x * ny + y) * nz + z)] += q * tx[ix] * ty[iy] * tz[iz]; } } }
    }
    fft3d(grid, nx, ny, nz, 0);
    double* mx_ = (double*)malloc(sizeof(double) * nx); double* my_ = (double*)malloc(sizeof(double) * ny); double* mz_ = (double*)malloc(sizeof(double) * nz);
    moduli(nx, mx_); moduli(ny, my_); moduli(nz, mz_);
    const double factor = M_PI * M_PI / (alpha * alpha), boxfactor = M_PI * det;
    const double dboxfactor = -2.0 * M_PI * sqrt(M_PI) / (6.0 * det), bfac = M_PI / alpha;
    const double fac1 = 2.0 * M_PI * M_PI * M_PI * sqrt(M_PI), fac2 = alpha * alpha * alpha, fac3 = -2.0 * alpha * M_PI * M_PI;
    double esum = 0;
    for (int kx = 0; kx < nx; kx++) { const double mx = kx < (nx + 1) / 2 ? kx : kx - nx;
        for (int ky = 0; ky < ny; ky++) { const double my = ky < (ny + 1) / 2 ? ky : ky - ny;
            for (int kz = 0; kz < nz; kz++) { const double mz = kz < (nz + 1) / 2 ? kz : kz - nz;
                double* g = &grid[2 * (((size_t)kx * ny + ky) * nz + kz)];
                if (!dispersion && kx == 0 && ky == 0 && kz == 0) { g[0] = g[1] = 0; continue; }
                const double mhx = mx * R[0][0], mhy = mx * R[1][0] + my * R[1][1], mhz = mx * R[2][0] + my * R[2][1] + mz * R[2][2];
                const double m2 = mhx * mhx + mhy * mhy + mhz * mhz;
                double eterm;
                if (!dispersion) eterm = ONE_4PI_EPS0 * exp(-factor * m2) / (m2 * boxfactor * mx_[kx] * my_[ky] * mz_[kz]);
                else {
                    const double m = sqrt(m2), bb = bfac * m;
                    eterm = (fac1 * erfc(bb) * m * m2 + exp(-bb * bb) * (fac2 + fac3 * m2)) * dboxfactor / (mx_[kx] * my_[ky] * mz_[kz]);
                }
                esum += eterm * (g[0] * g[0] + g[1] * g[1]);
                g[0] *= eterm; g[1] *= eterm; } } }
    if (energy) *energy += 0.5 * esum;
    if (forces) {
        fft3d(grid, nx, ny, nz, 1);
        for (int i = 0; i < n; i++) {
            const double q = coef[i];
            const double *tx = &th[(3 * (size_t)i) * PME_ORDER], *ty = tx + PME_ORDER, *tz = ty + PME_ORDER;
            const double *dx = &dth[(3 * (size_t)i) * PME_ORDER], *dy = dx + PME_ORDER, *dz = dy + PME_ORDER;
            double fx = 0, fy = 0, fz = 0;
            for (int ix = 0; ix < PME_ORDER; ix++) { const int x = (idx[3 * i] + ix) % nx;
                for (int iy = 0; iy < PME_ORDER; iy++) { const int y = (idx[3 * i + 1] + iy) % ny;
                    for (int iz = 0; iz < PME_ORDER; iz++) { const int z = (idx[3 * i + 2] + iz) % nz;
                        const double v = grid[2 * (((size_t)x * ny + y) * nz + z)];
                        fx += dx[ix] * ty[iy] * tz[iz] * v; fy += tx[ix] * dy[iy] * tz[iz] * v; fz += tx[ix] * ty[iy] * dz[iz] * v; } } }
            forces[3 * i]     -= q * (fx * nx * R[0][0]);
            forces[3 * i + 1] -= q * (fx * nx * R[1][0] + fy * ny * R[1][1]);
            forces[3 * i + 2] -= q * (fx * nx * R[2][0] + fy * ny * R[2][1] + fz * nz * R[2][2]);
        }
    }
    free(grid); free(idx); free(th); free(dth); free(mx_); free(my_); free(mz_);
}

/* NonbondedForce::Ewald: explicit half-space sum over k vectors, ReferenceLJCoulombIxn.cpp:272-368 (rectangular boxes) */
static void ewald_sum(const oracle_system* s, double* forces, double* energy) {
    const int n = s->n, nrx = s->ewald_kmax[0], nry = s->ewald_kmax[1], nrz = s->ewald_kmax[2];
    const double rb[3] = {2.0 * M_PI / s->box[0], 2.0 * M_PI / s->box[4], 2.0 * M_PI / s->box[8]};
    const double factor = -1.0 / (4.0 * s->alpha * s->alpha);
    const double coeff = ONE_4PI_EPS0 * 4.0 * M_PI / (s->box[0] * s->box[4] * s->box[8]);
    double* cr = (double*)malloc(sizeof(double) * (size_t)n), *ci = (double*)malloc(sizeof(double) * (size_t)n);
    double esum = 0;
    int lowry = 0, lowrz = 1;
    for (int rx = 0; rx < nrx; rx++) {
        const double kx = rx * rb[0];
        for (int ry = lowry; ry < nry; ry++) {
            const double ky = ry * rb[1];
            for (int rz = lowrz; rz < nrz; rz++) {
                const double kz = rz * rb[2];
                double cs = 0, ss = 0;
                for (int i = 0; i < n; i++) {
                    const double ph = kx * s->pos[3 * i] + ky * s->pos[3 * i + 1] + kz * s->pos[3 * i + 2];
                    cr[i] = s->charge[i] * cos(ph); ci[i] = s->charge[i] * sin(ph);
                    cs += cr[i]; ss += ci[i];
                }
                const double k2 = kx * kx + ky * ky + kz * kz, ak = exp(k2 * factor) / k2;
                if (forces) for (int i = 0; i < n; i++) {
                    const double f = 2.0 * coeff * ak * (cs * ci[i] - ss * cr[i]);
                    forces[3 * i] += f * kx; forces[3 * i + 1] += f * ky; forces[3 * i + 2] += f * kz;
                }
                esum += coeff * ak * (cs * cs + ss * ss);
                lowrz = 1 - nrz;
            }
            lowry = 1 - nry;
        }
    }
    free(cr); free(ci);
    if (energy) *energy += esum;
}

void oracle_reciprocal(const oracle_system* s, double* forces, double* energy) {
    if (s->ewald_kmax[0] > 0) { ewald_sum(s, forces, energy); return; }
    pme_sum(s, s->charge, s->alpha, s->grid, 0, forces, energy);
    if (s->ljpme) {
        double* c6 = (double*)malloc(sizeof(double) * (size_t)s->n);
        for (int i = 0; i < s->n; i++) c6[i] = 8.0 * pow(0.5 * s->sigma[i], 3.0) * 2.0 * sqrt(s->epsilon[i]);   /* ReferenceLJCoulombIxn.cpp:259-260 */
        pme_sum(s, c6, s->lj_alpha, s->lj_grid, 1, forces, energy);
        free(c6);
    }
}

/* ReferenceLJCoulombIxn.cpp:70-80 (krf, crf) and :582-639 (calculateOneIxn): plain or reaction-field Coulomb + Lennard-Jones with
   optional switching function; excluded pairs are skipped, pairs beyond the cutoff too (methods 2, 3) */
void oracle_direct_cutoff(const oracle_system* s, double* forces, double* energy) {
    const int n = s->n, cutoff = s->nb_method >= 2, periodic = s->nb_method == 3;
    const double rc2 = s->cutoff * s->cutoff;
    const double krf = cutoff ? pow(s->cutoff, -3.0) * (s->rf_dielectric - 1.0) / (2.0 * s->rf_dielectric + 1.0) : 0.0;
    const double crf = cutoff ? (1.0 / s->cutoff) * (3.0 * s->rf_dielectric) / (2.0 * s->rf_dielectric + 1.0) : 0.0;
    int* ex = (int*)malloc(sizeof(int) * 2 * (size_t)(s->nexcl > 0 ? s->nexcl : 1));
    for (int k = 0; k < s->nexcl; k++) {
        const int a = s->excl[2 * k], b = s->excl[2 * k + 1];
        ex[2 * k] = a < b ? a : b; ex[2 * k + 1] = a < b ? b : a;
    }
    qsort(ex, (size_t)s->nexcl, 2 * sizeof(int), cmp_pair);
    double esum = 0;
    for (int i = 0; i < n; i++) for (int j = i + 1; j < n; j++) {
        double d[3];
        if (periodic) delta_periodic(&s->pos[3 * i], &s->pos[3 * j], s->box, d);
        else for (int k = 0; k < 3; k++) d[k] = s->pos[3 * j + k] - s->pos[3 * i + k];
        const double r2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
        if (cutoff && r2 > rc2) continue;
        int key[2] = {i, j};
        if (s->nexcl > 0 && bsearch(key, ex, (size_t)s->nexcl, 2 * sizeof(int), cmp_pair)) continue;
        const double r = sqrt(r2), invR = 1.0 / r;
        double switchValue = 1.0, switchDeriv = 0.0;
        if (s->use_switch && r > s->switch_distance) {
            const double w = s->cutoff - s->switch_distance, t = (r - s->switch_distance) / w;
            switchValue = 1.0 + t * t * t * (-10.0 + t * (15.0 - t * 6.0));
            switchDeriv = t * t * (-30.0 + t * (60.0 - t * 30.0)) / w;
        }
        const double qq = ONE_4PI_EPS0 * s->charge[i] * s->charge[j];
        const double sig = 0.5 * (s->sigma[i] + s->sigma[j]), eps = 4.0 * sqrt(s->epsilon[i] * s->epsilon[j]);
        double sig2 = invR * sig; sig2 *= sig2;
        const double sig6 = sig2 * sig2 * sig2;
        double dEdR = switchValue * eps * (12.0 * sig6 - 6.0) * sig6;
        dEdR += cutoff ? qq * (invR - 2.0 * krf * r2) : qq * invR;
        dEdR *= invR * invR;
        double e = eps * (sig6 - 1.0) * sig6;
        if (s->use_switch) { dEdR -= e * switchDeriv * invR; e *= switchValue; }
        e += cutoff ? qq * (invR + krf * r2 - crf) : qq * invR;
        esum += e;
        if (forces) for (int k = 0; k < 3; k++) { forces[3 * i + k] -= dEdR * d[k]; forces[3 * j + k] += dEdR * d[k]; }
    }
    free(ex);
    if (energy) *energy += esum;
}

void oracle_nonbonded(const oracle_system* s, int groups, double* forces, double* energy) {
    if (s->nb_method != 0) {
        if (groups & 1) { oracle_direct_cutoff(s, forces, energy); oracle_exceptions(s, forces, energy); }
        return;
    }
    if (groups & 1) {
        oracle_direct(s, forces, energy, 0, 0);
        oracle_exclusion_correction(s, forces, energy);
        oracle_exceptions(s, forces, energy);
    }
    if (groups & 2) {
        oracle_reciprocal(s, forces, energy);
        if (energy) *energy += oracle_self_energy(s);
    }
}
