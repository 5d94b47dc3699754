/* nonbonded_oracle.h -- TEST INFRASTRUCTURE ONLY (see nonbonded_oracle.c). */
#ifndef NONBONDED_ORACLE_H_
#define NONBONDED_ORACLE_H_
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int n;
    const double* pos;      /* [3n] nm */
    const double* charge;   /* [n] */
    const double* sigma;    /* [n] nm */
    const double* epsilon;  /* [n] kJ/mol */
    double box[9];          /* a, b, c (reduced form) */
    double cutoff;
    double alpha;
    int grid[3];
    int nexcl; const int* excl;          /* excluded pairs [2*nexcl] */
    int nexc;  const int* exc;           /* exceptions with parameters [2*nexc] */
    const double* exc_qq; const double* exc_sigma; const double* exc_eps;
    int exceptions_periodic;
    int ljpme;              /* 1: LJPME (dispersion treated with a second PME grid) */
    double lj_alpha;        /* dispersion Ewald coefficient */
    int lj_grid[3];         /* dispersion mesh */
    int use_switch;         /* 1: Lennard-Jones switching function between switch_distance and cutoff */
    double switch_distance;
    int nb_method;          /* 0: PME / LJPME (everything above); 1: NoCutoff; 2: CutoffNonPeriodic; 3: CutoffPeriodic (reaction field) */
    double rf_dielectric;   /* reaction-field dielectric of the cutoff methods (NonbondedForce default 78.3) */
    int ewald_kmax[3];      /* non-zero: NonbondedForce::Ewald, explicit k-space sum with these limits instead of the mesh */
} oracle_system;

/* direct space, pairs with r <= cutoff that are not excluded; forces [3n] are ADDED to; returns pair count.
   If pairs != NULL, up to `capacity` pairs (i<j) are written in (i,j) order of discovery. */
long long oracle_direct(const oracle_system* s, double* forces, double* energy, int* pairs, long long capacity);
/* -erf(alpha r)/r correction of every excluded pair */
void oracle_exclusion_correction(const oracle_system* s, double* forces, double* energy);
/* 1-4 exceptions: plain Coulomb + LJ, no cutoff */
void oracle_exceptions(const oracle_system* s, double* forces, double* energy);
/* smooth PME reciprocal space, order 5 */
void oracle_reciprocal(const oracle_system* s, double* forces, double* energy);
/* Ewald self energy + neutralising plasma */
double oracle_self_energy(const oracle_system* s);
/* NoCutoff / CutoffNonPeriodic / CutoffPeriodic direct sum (ReferenceLJCoulombIxn::calculateOneIxn), O(N^2): small systems only */
void oracle_direct_cutoff(const oracle_system* s, double* forces, double* energy);
/* everything above; `groups` bit 0 = direct(+exclusion corr+exceptions), bit 1 = reciprocal(+self) */
void oracle_nonbonded(const oracle_system* s, int groups, double* forces, double* energy);
/* complex 1D FFT used by the reciprocal sum (exposed for the FFT unit test): data = interleaved re,im */
void oracle_fft1d(double* data, int n, int stride, int inverse);
/* unnormalised 3D r2c FFT (out: nx*ny*(nz/2+1) complex) -- checker for b200_fft3d_r2c_host */
void oracle_fft3d_r2c(const double* in, double* out, int nx, int ny, int nz);

#ifdef __cplusplus
}
#endif
#endif
