/*
 * b200_nonbonded.h -- C ABI of the B200-native NonbondedForce/PME path.
 *
 * This is the drop-in boundary below the OpenMM `B200` Platform plugin
 * (openmm_b200/plugin/): plain pointers and sizes, no C++/torch types.  Every entry
 * point names the reference interface it replaces (paths relative to the OpenMM
 * 8.6 tree).  All functions return B200_OK (0) or a negative error code; the text of
 * the last error on the calling thread is available from b200_last_error().
 *
 * Conventions (same as the reference's device path):
 *   units nm, ps, amu, kJ/mol, e;  ONE_4PI_EPS0 = 138.935456 (SimTKOpenMMRealType.h:84-89)
 *   forces accumulate as int64 fixed point, value * 2^32 (platforms/cuda/src/kernels/common.cu:29-31),
 *   SoA [x0..x_{Npad-1} | y.. | z..], Npad = 32*ceil(N/32) (CudaContext.cpp:209)
 *   mixed precision: float positions + float correction, double velocities,
 *   float pair math, int64 force sums, fixed-point energy sums.
 */
#ifndef B200_NONBONDED_H_
#define B200_NONBONDED_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK                 0
#define B200_ERR_INVALID       -1
#define B200_ERR_CUDA          -2
#define B200_ERR_OVERFLOW      -3
#define B200_ERR_UNSUPPORTED   -4
#define B200_ERR_NCCL          -5

/* NonbondedForce::NonbondedMethod / CalcNonbondedForceKernel::NonbondedMethod (olla/include/openmm/kernels.h:628-635) */
#define B200_NB_NOCUTOFF          0
#define B200_NB_CUTOFF_NONPERIODIC 1
#define B200_NB_CUTOFF_PERIODIC   2
#define B200_NB_EWALD             3
#define B200_NB_PME               4
#define B200_NB_LJPME             5

typedef struct b200_ctx b200_ctx;

/* ---- System-level description: what ContextImpl hands a platform at contextCreated()
 *      (openmmapi/src/ContextImpl.cpp:149-156) plus UpdateStateDataKernel::initialize (kernels.h:144). */
typedef struct {
    int32_t       num_atoms;
    const double* mass;              /* [N]; 0 => immobile particle */
    double        box[9];            /* a,b,c box vectors, reduced form (a=(ax,0,0), b=(bx,by,0), c=(cx,cy,cz)) */
    int32_t       num_constraints;   /* System::getConstraintParameters */
    const int32_t* constraint_atoms; /* [2*num_constraints] */
    const double* constraint_dist;   /* [num_constraints] */
    int32_t       device;            /* CUDA device ordinal ("DeviceIndex" platform property, CudaPlatform.cpp:116-141) */
    uint32_t      flags;             /* B200_CTX_* */
} b200_system_desc;

#define B200_CTX_DISABLE_PME_STREAM  1u   /* "DisablePmeStream" property */

/* ---- NonbondedForce description: what CalcNonbondedForceKernel::initialize(system, force)
 *      (kernels.h:647; CommonCalcNonbondedForce.cpp:217-717) extracts from the force object. */
typedef struct {
    int32_t       method;            /* B200_NB_* */
    double        cutoff;            /* nm */
    double        ewald_alpha;       /* nm^-1 (NonbondedForceImpl::calcPMEParameters, NonbondedForceImpl.cpp:192-216) */
    int32_t       grid[3];           /* PME / LJPME: mesh, each dimension must factor into 2,3,5,7 */
    double        dispersion_coefficient; /* C of E = C/V (NonbondedForceImpl::calcDispersionCorrection, :268-342); 0 = off */
    int32_t       use_switch;        /* LJ switching function (coulombLennardJones.cc:32-40) */
    double        switch_distance;
    double        reaction_field_dielectric; /* CutoffPeriodic/NonPeriodic only */
    const double* charge;            /* [N] e */
    const double* sigma;             /* [N] nm */
    const double* epsilon;           /* [N] kJ/mol */
    int32_t       num_exclusions;    /* every exception (i,j): excluded from the direct-space loop */
    const int32_t* exclusion_pairs;  /* [2*num_exclusions] */
    int32_t       num_exceptions;    /* exceptions with non-zero chargeProd or epsilon ("1-4" pairs, nonbondedExceptions.cc:1-19) */
    const int32_t* exception_pairs;  /* [2*num_exceptions] */
    const double* exception_chargeprod; /* [num_exceptions] e^2 */
    const double* exception_sigma;   /* nm */
    const double* exception_epsilon; /* kJ/mol */
    int32_t       exceptions_use_periodic; /* NonbondedForce::getExceptionsUsePeriodicBoundaryConditions */
    double        ljpme_alpha;       /* LJPME only */
    int32_t       ljpme_grid[3];
    int32_t       ewald_kmax[3];     /* Ewald only: number of k vectors per direction (NonbondedForceImpl::calcEwaldParameters, NonbondedForceImpl.cpp:159-190) */
} b200_nonbonded_desc;

/* ---- context life cycle (Platform::contextCreated / contextDestroyed, olla/include/openmm/Platform.h:232-252) */
int  b200_context_create(const b200_system_desc* desc, b200_ctx** out);
void b200_context_destroy(b200_ctx* ctx);
const char* b200_last_error(void);
/* number of CUDA devices visible (Platform::getDevices, Platform.h:127-150); <0 on error */
int  b200_device_count(void);

/* CalcNonbondedForceKernel::initialize (kernels.h:647) */
int  b200_nonbonded_create(b200_ctx* ctx, const b200_nonbonded_desc* desc);
/* CalcNonbondedForceKernel::copyParametersToContext (kernels.h:669): new per-particle and exception parameters */
int  b200_nonbonded_update_parameters(b200_ctx* ctx, const b200_nonbonded_desc* desc);
/* CalcNonbondedForceKernel::getPMEParameters (kernels.h:678) */
int  b200_nonbonded_get_pme_parameters(const b200_ctx* ctx, double* alpha, int32_t* nx, int32_t* ny, int32_t* nz);
/* CalcNonbondedForceKernel::getLJPMEParameters (kernels.h:687; CommonCalcNonbondedForce.cpp:1269-1281): dispersion mesh,
   B200_ERR_INVALID unless the force was created with method B200_NB_LJPME */
int  b200_nonbonded_get_ljpme_parameters(const b200_ctx* ctx, double* alpha, int32_t* nx, int32_t* ny, int32_t* nz);

/* CalcHarmonicBondForceKernel::initialize (kernels.h:347), harmonicBondForce.cc: E = k/2 (r-r0)^2 */
int  b200_harmonic_bonds_create(b200_ctx* ctx, int32_t n, const int32_t* atoms /*[2n]*/, const double* length, const double* k, int32_t group);
/* CalcHarmonicAngleForceKernel::initialize (kernels.h:421), harmonicAngleForce.cc: E = k/2 (theta-theta0)^2 */
int  b200_harmonic_angles_create(b200_ctx* ctx, int32_t n, const int32_t* atoms /*[3n]*/, const double* angle, const double* k, int32_t group);
/* CalcPeriodicTorsionForceKernel::initialize (kernels.h:495), periodicTorsionForce.cc: E = k (1+cos(n phi - phi0)) */
int  b200_periodic_torsions_create(b200_ctx* ctx, int32_t n, const int32_t* atoms /*[4n]*/, const int32_t* periodicity, const double* phase, const double* k, int32_t group);

/* HarmonicBondForce/HarmonicAngleForce/PeriodicTorsionForce::setUsesPeriodicBoundaryConditions for all terms of one kind (0, 1, 2) */
int  b200_bonded_set_periodic(b200_ctx* ctx, int32_t kind, int32_t periodic);
/* Which terms of one kind the following b200_compute calls evaluate: nranges pairs (first, count) over the creation order, or
 * nranges = 0 for all of them.  Lets several forces of one kind live in different force groups (ContextImpl::calcForcesAndEnergy
 * calls only the kernels of the requested groups, ContextImpl.cpp; the reference keeps one kernel object per force). */
int  b200_bonded_set_active(b200_ctx* ctx, int32_t kind, int32_t nranges, const int32_t* first_count /*[2 nranges]*/);
/* copyParametersToContext of the three bonded kernels (kernels.h:359,433,507): new parameters for terms [first, first+n) of kind
 * 0 = bonds (p0 length, p1 k), 1 = angles (p0 angle, p1 k), 2 = torsions (p0 phase, p1 k, periodicity); term indices count over all
 * forces of that kind in the order they were created */
int  b200_bonded_update_parameters(b200_ctx* ctx, int32_t kind, int32_t first, int32_t n, const double* p0, const double* p1, const int32_t* periodicity);

/* ---- state I/O: UpdateStateDataKernel (kernels.h:132-246; CommonKernels.cpp:95-297).
 *      Host arrays are [3N] doubles in the caller's (original) particle order. */
int  b200_set_positions(b200_ctx* ctx, const double* xyz);
/* multi-GPU, collective: rank 0 passes the host array, uploads it once over PCIe and stores the converted positions into every rank's
 * arrays over NVLink; the other ranks pass NULL.  (The reference's multi-GPU path uploads to every device from the host thread,
 * CudaParallelKernels.cpp:183-207.)  b200_get_positions(ctx, NULL, ...) lets a rank take part in the collective gather without a download. */
int  b200_set_positions_root(b200_ctx* ctx, const double* xyz);
int  b200_get_positions(b200_ctx* ctx, double* xyz, int32_t enforce_periodic);
int  b200_set_velocities(b200_ctx* ctx, const double* xyz);
int  b200_get_velocities(b200_ctx* ctx, double* xyz);
int  b200_get_forces(b200_ctx* ctx, double* xyz);
/* Adds forces computed elsewhere (kJ/mol/nm, particle order, doubles) to the force buffer, the way every ForceImpl of a Context adds
 * into the shared fixed-point buffer (ContextImpl::calcForcesAndEnergy, ContextImpl.cpp; CudaContext's long long force array).  The
 * plugin uses it for a second NonbondedForce, which is evaluated in a context of its own.  One rank only. */
int  b200_add_forces(b200_ctx* ctx, const double* xyz);
int  b200_set_box(b200_ctx* ctx, const double* box9);
int  b200_get_box(const b200_ctx* ctx, double* box9);
int  b200_set_time(b200_ctx* ctx, double time_ps, int64_t step_count);
int  b200_get_time(const b200_ctx* ctx, double* time_ps, int64_t* step_count);

/* ---- force evaluation: CalcForcesAndEnergyKernel::beginComputation/finishComputation (kernels.h:110,125)
 *      + CalcNonbondedForceKernel::execute (kernels.h:658).  `groups` bits as below.  Energy (kJ/mol) may be NULL. */
#define B200_GROUP_DIRECT      1u   /* tile kernel + exclusion corrections... (includeDirect) */
#define B200_GROUP_RECIPROCAL  2u   /* PME reciprocal space + self/plasma terms (includeReciprocal) */
#define B200_GROUP_BONDS       4u   /* HarmonicBondForce */
#define B200_GROUP_ANGLES      8u   /* HarmonicAngleForce */
#define B200_GROUP_TORSIONS   16u   /* PeriodicTorsionForce */
#define B200_GROUP_BONDED     28u   /* bonds | angles | torsions */
#define B200_GROUP_ALL        31u
int  b200_compute(b200_ctx* ctx, uint32_t groups, int32_t include_forces, int32_t include_energy, double* energy);

/* Host-buffer convenience used by the end-to-end measurement: upload positions, evaluate, download forces.
 * Same semantics as Context::setPositions + getState(Forces|Energy) (openmmapi/src/Context.cpp:87-227). */
int  b200_compute_host(b200_ctx* ctx, const double* xyz_in, uint32_t groups, double* forces_out, double* energy);

/* ---- integrator: IntegrateLangevinMiddleStepKernel (kernels.h:1392-1420; langevinMiddle.cc:9-95),
 *      constraints ApplyConstraintsKernel (kernels.h:251-278; SETTLE integrationUtilities.cc:328-580, CCMA :582-804),
 *      RemoveCMMotionKernel (kernels.h:1793; removeCM.cc). friction=0 gives velocity Verlet (IntegrateVerletStepKernel). */
typedef struct {
    double   dt;            /* ps */
    double   temperature;   /* K */
    double   friction;      /* 1/ps */
    double   constraint_tol;
    uint64_t seed;
    int32_t  remove_cm_every; /* CMMotionRemover frequency; 0 = never */
} b200_integrator_desc;
int  b200_integrator_create(b200_ctx* ctx, const b200_integrator_desc* desc);
int  b200_step(b200_ctx* ctx, int32_t nsteps);                      /* nsteps x (forces of all groups + COM removal + integration) */
int  b200_integrate(b200_ctx* ctx);                                  /* IntegrateLangevinMiddleStepKernel::execute: integrate once with the forces in the buffer */
int  b200_remove_cm_motion(b200_ctx* ctx);                           /* RemoveCMMotionKernel::execute (kernels.h:1793) */
int  b200_apply_constraints(b200_ctx* ctx, double tol);              /* ApplyConstraintsKernel::apply */
int  b200_apply_velocity_constraints(b200_ctx* ctx, double tol);     /* ApplyConstraintsKernel::applyToVelocities */
int  b200_kinetic_energy(b200_ctx* ctx, double* ke);                 /* IntegrateLangevinMiddleStepKernel::computeKineticEnergy */
/* kinetic energy of v + shift*F/m (forces of the last b200_compute) after the velocity constraints: what leapfrog integrators report,
   IntegrateVerletStepKernel::computeKineticEnergy -> IntegrationUtilities::computeKineticEnergy(0.5*dt) (CommonKernels.cpp:3117) */
int  b200_kinetic_energy_shifted(b200_ctx* ctx, double shift_ps, double* ke);
int  b200_synchronize(b200_ctx* ctx);

/* ---- Monte Carlo barostat: ApplyMonteCarloBarostatKernel (kernels.h:1723-1788; CommonKernels.cpp:4255-4400, monteCarloBarostat.cc).
 *      save: keep the state for a possible rejection; scale: move every molecule (rigid_molecules != 0) or every atom so that its centre
 *      r goes to (sx x + sxy y + sxz z, sy y + syz z, sz z), scale6 = {sx, sy, sz, sxy, sxz, syz}; restore: undo a rejected move.
 *      The box itself is changed with b200_set_box (the PME reciprocal vectors and the neighbour list follow on the next evaluation). */
int  b200_barostat_save(b200_ctx* ctx);
int  b200_barostat_scale(b200_ctx* ctx, const double* scale6, int32_t rigid_molecules);
int  b200_barostat_restore(b200_ctx* ctx);

/* ---- parity probes (no counterpart kernel in the reference; they expose what its tests measure) */
/* Set of unordered atom pairs (original indices, i<j) that the direct-space kernel treats as interacting
 * (r < cutoff, not excluded).  Mirrors what computeNeighborListVoxelHash yields (ReferenceNeighborList.cpp:208-214).
 * Call with pairs==NULL to get the count only. */
int  b200_direct_pairs(b200_ctx* ctx, int64_t* count, int32_t* pairs /*[2*capacity]*/, int64_t capacity);
/* The same set reduced on the device to its size and an order-independent 64-bit hash, sum over pairs of splitmix64(i*N + j)
 * mod 2^64 with i<j original indices (what tests/TestNonbondedForce.h testLargeSystem needs at a million atoms, where the list
 * itself is > 1 GB).  oracle/ref_driver.cpp --pair-hash computes the same number from computeNeighborListVoxelHash. */
int  b200_direct_pairs_hash(b200_ctx* ctx, int64_t* count, uint64_t* hash);
/* Neighbour-list statistics after the last build: tiles, work units, j-slots, exclusion tiles, rebuild count. */
typedef struct {
    int64_t num_blocks, num_tiles, num_units, num_exclusion_tiles, rebuilds, steps, reorders;
    int64_t tile_capacity, unit_capacity;
    int64_t num_sparse_tiles;   /* of num_tiles (each sparse tile occupies two slots) */
} b200_nblist_stats;
int  b200_nblist_get_stats(b200_ctx* ctx, b200_nblist_stats* out);

/* ---- stage-level entry points for unit tests (CudaFFT3D::execFFT, CudaFFT3D.cpp:66-101; TestCudaFFT3D.cpp:56-121) */
/* out-of-place unnormalised 3D real-to-complex / complex-to-real FFT on host buffers (row-major, z fastest;
 * complex layout nx*ny*(nz/2+1) interleaved re,im). */
int  b200_fft3d_r2c_host(int32_t device, int32_t nx, int32_t ny, int32_t nz, const float* in, float* out);
int  b200_fft3d_c2r_host(int32_t device, int32_t nx, int32_t ny, int32_t nz, const float* in, float* out);
/* Allocation-free variant of the complex-to-real transform.  It has no counterpart in the reference.  For callers that
 * cannot call cudaMalloc when they need the transform (a memory pool owns the device, a latency-critical loop must stay
 * off the allocator, a fixed memory budget): device pointers in and out, the working storage is a static ~8 MiB device
 * array inside the library.  Limits: nx*ny*(nz/2+1) <= 1048576, each dimension <= 1024 and a product of 2, 3, 5, 7 only,
 * otherwise B200_ERR_INVALID.  The caller's current CUDA device is restored before returning.
 * The first call on a device performs a one-time address look-up and should precede any stream capture; calls share the
 * scratch, so calls that overlap on the GPU race.  d_in is not modified.  `stream` is a cudaStream_t (NULL = default). */
int  b200_fft3d_c2r_noalloc(int32_t device, int32_t nx, int32_t ny, int32_t nz, const float* d_in, float* d_out, void* stream);
/* smallest n' >= n whose only prime factors are 2,3,5,7 (ComputeContext::findLegalFFTDimension, ComputeContext.cpp:717-732) */
int  b200_find_legal_fft_dimension(int32_t minimum);

/* ---- measurement hooks: cumulative CUDA-event time (ms) and launch count per kernel family since the last reset */
#define B200_T_TILE        0
#define B200_T_NBLIST      1
#define B200_T_SPREAD      2
#define B200_T_FFT         3
#define B200_T_CONVOLVE    4
#define B200_T_INTERPOLATE 5
#define B200_T_BONDED      6
#define B200_T_INTEGRATE   7
#define B200_T_REORDER     8
#define B200_T_COMM        9
#define B200_T_COUNT       10
int  b200_timing_enable(b200_ctx* ctx, int32_t on);      /* when on, every family is bracketed by events (adds syncs at read time only) */
int  b200_timing_read(b200_ctx* ctx, double* ms /*[B200_T_COUNT]*/, int64_t* launches /*[B200_T_COUNT]*/);
int  b200_timing_reset(b200_ctx* ctx);
/* time `nsteps` MD steps with CUDA events on the context's own stream; returns total ms */
int  b200_timed_steps(b200_ctx* ctx, int32_t nsteps, double* total_ms);
/* total kernels launched by this context since creation (the bench's gpu_launches claim) */
int64_t b200_launch_count(const b200_ctx* ctx);

/* ---- multi-GPU: direct-space rows are partitioned over ranks; forces are all-reduced as int64 over NCCL
 *      (replaces CudaParallelKernels.cpp:183-269 + parallel.cu:5-13 sumForces). */
int  b200_nccl_unique_id(char* id128);                                           /* rank 0 creates, caller broadcasts */
int  b200_comm_init(b200_ctx* ctx, int32_t rank, int32_t nranks, const char* id128);
int  b200_set_partition(b200_ctx* ctx, int32_t rank, int32_t nranks);            /* partition without NCCL (single-device emulation of N ranks) */

#ifdef __cplusplus
}
#endif
#endif /* B200_NONBONDED_H_ */
