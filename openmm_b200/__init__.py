"""openmm_b200 -- host-side mirror of the B200 NonbondedForce/PME platform (ctypes over the C ABI in include/b200_nonbonded.h).

The C++ OpenMM plugin (openmm_b200/plugin -> lib/libOpenMMB200.so) is the drop-in used from OpenMM itself; this module
is the same C ABI seen from Python, used by the tests and by bench.py.  Names follow the reference's kernel interface
(olla/include/openmm/kernels.h): a `Context` owns the device state (UpdateStateDataKernel), `NonbondedForce` parameters go
in through `add_nonbonded_force` (CalcNonbondedForceKernel::initialize), forces/energy come from `compute`
(CalcForcesAndEnergyKernel + CalcNonbondedForceKernel::execute), `step` is IntegrateLangevinMiddleStepKernel::execute.

There is no CPU fallback: importing works anywhere, but creating a Context without the CUDA library or without a GPU raises.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_LIB_PATH", os.path.join(_HERE, "lib", "libb200nb.so"))   # override: kernel-variant sweeps
PLUGIN_PATH = os.path.join(_HERE, "lib", "libOpenMMB200.so")

NB_NOCUTOFF, NB_CUTOFF_NONPERIODIC, NB_CUTOFF_PERIODIC, NB_EWALD, NB_PME, NB_LJPME = range(6)
GROUP_DIRECT, GROUP_RECIPROCAL, GROUP_BONDS, GROUP_ANGLES, GROUP_TORSIONS, GROUP_BONDED, GROUP_ALL = 1, 2, 4, 8, 16, 28, 31
TIMER_NAMES = ["tile", "nblist", "spread", "fft", "convolve", "interpolate", "bonded", "integrate", "reorder", "comm"]

c_double_p = ctypes.POINTER(ctypes.c_double)
c_int32_p = ctypes.POINTER(ctypes.c_int32)


class B200Error(RuntimeError):
    pass


class SystemDesc(ctypes.Structure):
    _fields_ = [("num_atoms", ctypes.c_int32), ("mass", c_double_p), ("box", ctypes.c_double * 9),
                ("num_constraints", ctypes.c_int32), ("constraint_atoms", c_int32_p), ("constraint_dist", c_double_p),
                ("device", ctypes.c_int32), ("flags", ctypes.c_uint32)]


class NonbondedDesc(ctypes.Structure):
    _fields_ = [("method", ctypes.c_int32), ("cutoff", ctypes.c_double), ("ewald_alpha", ctypes.c_double), ("grid", ctypes.c_int32 * 3),
                ("dispersion_coefficient", ctypes.c_double), ("use_switch", ctypes.c_int32), ("switch_distance", ctypes.c_double),
                ("reaction_field_dielectric", ctypes.c_double),
                ("charge", c_double_p), ("sigma", c_double_p), ("epsilon", c_double_p),
                ("num_exclusions", ctypes.c_int32), ("exclusion_pairs", c_int32_p),
                ("num_exceptions", ctypes.c_int32), ("exception_pairs", c_int32_p),
                ("exception_chargeprod", c_double_p), ("exception_sigma", c_double_p), ("exception_epsilon", c_double_p),
                ("exceptions_use_periodic", ctypes.c_int32), ("ljpme_alpha", ctypes.c_double), ("ljpme_grid", ctypes.c_int32 * 3),
                ("ewald_kmax", ctypes.c_int32 * 3)]


class IntegratorDesc(ctypes.Structure):
    _fields_ = [("dt", ctypes.c_double), ("temperature", ctypes.c_double), ("friction", ctypes.c_double),
                ("constraint_tol", ctypes.c_double), ("seed", ctypes.c_uint64), ("remove_cm_every", ctypes.c_int32)]


class NblistStats(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int64) for n in ("num_blocks", "num_tiles", "num_units", "num_exclusion_tiles", "rebuilds", "steps",
                                              "reorders", "tile_capacity", "unit_capacity", "num_sparse_tiles")]


# every symbol include/b200_nonbonded.h declares (tests/test_abi.py checks the header against this list and the .so)
EXPORTED_SYMBOLS = [
    "b200_context_create", "b200_context_destroy", "b200_last_error", "b200_device_count", "b200_nonbonded_create",
    "b200_nonbonded_update_parameters", "b200_nonbonded_get_pme_parameters", "b200_nonbonded_get_ljpme_parameters", "b200_harmonic_bonds_create",
    "b200_harmonic_angles_create", "b200_periodic_torsions_create", "b200_bonded_update_parameters", "b200_bonded_set_periodic", "b200_bonded_set_active", "b200_set_positions", "b200_set_positions_root", "b200_get_positions",
    "b200_set_velocities", "b200_get_velocities", "b200_get_forces", "b200_add_forces", "b200_set_box", "b200_get_box", "b200_set_time",
    "b200_get_time", "b200_compute", "b200_compute_host", "b200_integrator_create", "b200_step", "b200_integrate",
    "b200_remove_cm_motion", "b200_apply_constraints",
    "b200_apply_velocity_constraints", "b200_kinetic_energy", "b200_kinetic_energy_shifted", "b200_synchronize", "b200_barostat_save", "b200_barostat_scale", "b200_barostat_restore", "b200_direct_pairs", "b200_direct_pairs_hash", "b200_nblist_get_stats",
    "b200_fft3d_r2c_host", "b200_fft3d_c2r_host", "b200_fft3d_c2r_noalloc", "b200_find_legal_fft_dimension", "b200_timing_enable", "b200_timing_read",
    "b200_timing_reset", "b200_timed_steps", "b200_launch_count", "b200_nccl_unique_id", "b200_comm_init", "b200_set_partition",
]

_lib = None


def load_library():
    """dlopen the CUDA library; raises if it was not built (no silent fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error(f"{LIB_PATH} is missing: run `python -m openmm_b200.build` (there is no CPU fallback)")
    lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    lib.b200_last_error.restype = ctypes.c_char_p
    lib.b200_launch_count.restype = ctypes.c_int64
    lib.b200_context_destroy.restype = None
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        raise B200Error(f"b200 error {rc}: {load_library().b200_last_error().decode()}")


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(c_double_p)


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(c_int32_p)


def find_legal_fft_dimension(n):
    """ComputeContext::findLegalFFTDimension (platforms/common/src/ComputeContext.cpp:717-732), host-only."""
    n = max(int(n), 1)
    while True:
        m = n
        for f in (2, 3, 5, 7):
            while m % f == 0:
                m //= f
        if m == 1:
            return n
        n += 1


def pme_parameters(cutoff, box, tol=5e-4, lj=False):
    """alpha and mesh as NonbondedForceImpl::calcPMEParameters (openmmapi/src/NonbondedForceImpl.cpp:192-216) followed by
    the 2,3,5,7-smooth rounding of the GPU platforms."""
    alpha = np.sqrt(-np.log(2 * tol)) / cutoff
    grid = [find_legal_fft_dimension(int(np.ceil((1 if lj else 2) * alpha * L / (3 * tol ** 0.2)))) for L in (box[0], box[4], box[8])]
    return float(alpha), [max(g, 6) for g in grid]


class Context:
    """Device state + kernels of one simulation (the B200 analogue of ContextImpl's platform data)."""

    def __init__(self, masses, box, constraints=None, device=0, disable_pme_stream=False):
        self.lib = load_library()
        self._keep = []
        masses, mp = _d(masses)
        self.n = len(masses)
        d = SystemDesc()
        d.num_atoms = self.n
        d.mass = mp
        box = np.asarray(box, dtype=np.float64).reshape(9)
        for k in range(9):
            d.box[k] = box[k]
        if constraints is not None and len(constraints[0]) > 0:
            ca, cap = _i(np.asarray(constraints[0]).reshape(-1))
            cd, cdp = _d(constraints[1])
            d.num_constraints = len(cd)
            d.constraint_atoms = cap
            d.constraint_dist = cdp
        d.device = device
        d.flags = 1 if disable_pme_stream else 0
        self.handle = ctypes.c_void_p()
        _check(self.lib.b200_context_create(ctypes.byref(d), ctypes.byref(self.handle)))

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b200_context_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # CalcNonbondedForceKernel::initialize
    def add_nonbonded_force(self, charge, sigma, epsilon, cutoff, alpha, grid, exclusions=(), exceptions=None,
                            method=NB_PME, dispersion_coefficient=0.0, exceptions_use_periodic=False, lj_alpha=None, lj_grid=None, switch_distance=None,
                            reaction_field_dielectric=78.3, ewald_kmax=None):
        d = NonbondedDesc()
        d.method = method
        d.reaction_field_dielectric = reaction_field_dielectric
        if ewald_kmax is not None:    # NonbondedForce::Ewald: explicit sum over k vectors
            for k in range(3):
                d.ewald_kmax[k] = int(ewald_kmax[k])
        if switch_distance is not None:    # NonbondedForce::setUseSwitchingFunction(true) + setSwitchingDistance
            d.use_switch, d.switch_distance = 1, switch_distance
        if lj_alpha is not None:      # NonbondedForce::LJPME with setLJPMEParameters(lj_alpha, *lj_grid)
            d.method = NB_LJPME
            d.ljpme_alpha = lj_alpha
            for k in range(3):
                d.ljpme_grid[k] = int(lj_grid[k])
        d.cutoff = cutoff
        d.ewald_alpha = alpha
        for k in range(3):
            d.grid[k] = int(grid[k])
        d.dispersion_coefficient = dispersion_coefficient
        q, d.charge = _d(charge)
        s, d.sigma = _d(sigma)
        e, d.epsilon = _d(epsilon)
        ex, d.exclusion_pairs = _i(np.asarray(exclusions, dtype=np.int32).reshape(-1))
        d.num_exclusions = len(ex) // 2
        if exceptions is not None and len(exceptions[0]) > 0:
            xp, d.exception_pairs = _i(np.asarray(exceptions[0]).reshape(-1))
            xq, d.exception_chargeprod = _d(exceptions[1])
            xs, d.exception_sigma = _d(exceptions[2])
            xe, d.exception_epsilon = _d(exceptions[3])
            d.num_exceptions = len(xq)
        d.exceptions_use_periodic = 1 if exceptions_use_periodic else 0
        _check(self.lib.b200_nonbonded_create(self.handle, ctypes.byref(d)))

    def update_nonbonded_parameters(self, charge, sigma, epsilon, num_exclusions, exceptions=None, dispersion_coefficient=0.0):
        """NonbondedForce::updateParametersInContext (CalcNonbondedForceKernel::copyParametersToContext, kernels.h:669): new per-particle
        parameters and exception parameters; the SET of exclusions/exceptions must be the one the force was created with."""
        d = NonbondedDesc()
        d.dispersion_coefficient = dispersion_coefficient
        q, d.charge = _d(charge)
        s, d.sigma = _d(sigma)
        e, d.epsilon = _d(epsilon)
        d.num_exclusions = int(num_exclusions)
        if exceptions is not None and len(exceptions[0]) > 0:
            xp, d.exception_pairs = _i(np.asarray(exceptions[0]).reshape(-1))
            xq, d.exception_chargeprod = _d(exceptions[1])
            xs, d.exception_sigma = _d(exceptions[2])
            xe, d.exception_epsilon = _d(exceptions[3])
            d.num_exceptions = len(xq)
        _check(self.lib.b200_nonbonded_update_parameters(self.handle, ctypes.byref(d)))

    def add_harmonic_bonds(self, atoms, length, k):
        a, ap = _i(np.asarray(atoms).reshape(-1)); l, lp = _d(length); kk, kp = _d(k)
        _check(self.lib.b200_harmonic_bonds_create(self.handle, len(l), ap, lp, kp, 0))

    def add_harmonic_angles(self, atoms, angle, k):
        a, ap = _i(np.asarray(atoms).reshape(-1)); l, lp = _d(angle); kk, kp = _d(k)
        _check(self.lib.b200_harmonic_angles_create(self.handle, len(l), ap, lp, kp, 0))

    def add_periodic_torsions(self, atoms, periodicity, phase, k):
        a, ap = _i(np.asarray(atoms).reshape(-1)); n, np_ = _i(periodicity); p, pp = _d(phase); kk, kp = _d(k)
        _check(self.lib.b200_periodic_torsions_create(self.handle, len(p), ap, np_, pp, kp, 0))

    def bonded_set_active(self, kind, ranges=()):
        """terms of kind 0 bonds / 1 angles / 2 torsions evaluated by the following compute() calls: (first, count) pairs over the
        creation order, or nothing for all (several forces of one kind in different force groups)"""
        r = np.asarray(ranges, dtype=np.int32).reshape(-1)
        _check(self.lib.b200_bonded_set_active(self.handle, int(kind), r.size // 2, r.ctypes.data_as(c_int32_p) if r.size else None))

    # UpdateStateDataKernel
    def set_positions(self, xyz):
        a, p = _d(np.asarray(xyz).reshape(-1))
        assert a.size == 3 * self.n
        _check(self.lib.b200_set_positions(self.handle, p))

    def set_positions_root(self, xyz=None):
        """collective over the ranks of a communicator: rank 0 passes the array, the others pass None (positions travel over NVLink)"""
        if xyz is None:
            _check(self.lib.b200_set_positions_root(self.handle, None))
            return
        a, p = _d(np.asarray(xyz).reshape(-1))
        _check(self.lib.b200_set_positions_root(self.handle, p))

    def join_get_positions(self):
        """collective counterpart of get_positions() for ranks that do not need the host copy"""
        _check(self.lib.b200_get_positions(self.handle, None, 0))

    def get_positions(self, enforce_periodic=False, out=None):
        if out is None:
            out = np.empty((self.n, 3))
        assert out.dtype == np.float64 and out.size == 3 * self.n and out.flags.c_contiguous
        _check(self.lib.b200_get_positions(self.handle, out.ctypes.data_as(c_double_p), 1 if enforce_periodic else 0))
        return out.reshape(-1, 3)

    def set_velocities(self, xyz):
        a, p = _d(np.asarray(xyz).reshape(-1))
        _check(self.lib.b200_set_velocities(self.handle, p))

    def get_velocities(self):
        out = np.empty(3 * self.n)
        _check(self.lib.b200_get_velocities(self.handle, out.ctypes.data_as(c_double_p)))
        return out.reshape(-1, 3)

    def add_forces(self, xyz):
        a, p = _d(np.asarray(xyz).reshape(-1))
        assert a.size == 3 * self.n
        _check(self.lib.b200_add_forces(self.handle, p))

    def get_forces(self):
        out = np.empty(3 * self.n)
        _check(self.lib.b200_get_forces(self.handle, out.ctypes.data_as(c_double_p)))
        return out.reshape(-1, 3)

    def set_box(self, box):
        a, p = _d(np.asarray(box).reshape(9))
        _check(self.lib.b200_set_box(self.handle, p))

    # CalcForcesAndEnergyKernel / CalcNonbondedForceKernel::execute
    def compute(self, groups=GROUP_ALL, forces=True, energy=True):
        e = ctypes.c_double(0.0)
        _check(self.lib.b200_compute(self.handle, groups, 1 if forces else 0, 1 if energy else 0, ctypes.byref(e)))
        return e.value if energy else None

    def compute_host(self, xyz, groups=GROUP_ALL, want_energy=True):
        a, p = _d(np.asarray(xyz).reshape(-1))
        f = np.empty(3 * self.n)
        e = ctypes.c_double(0.0)
        _check(self.lib.b200_compute_host(self.handle, p, groups, f.ctypes.data_as(c_double_p), ctypes.byref(e) if want_energy else None))
        return f.reshape(-1, 3), e.value

    def direct_pairs(self, capacity=0):
        cnt = ctypes.c_int64(0)
        if capacity > 0:
            buf = np.empty((capacity, 2), dtype=np.int32)
            _check(self.lib.b200_direct_pairs(self.handle, ctypes.byref(cnt), buf.ctypes.data_as(c_int32_p), capacity))
            return cnt.value, buf[:min(cnt.value, capacity)]
        _check(self.lib.b200_direct_pairs(self.handle, ctypes.byref(cnt), None, 0))
        return cnt.value, None

    def direct_pairs_hash(self):
        """(count, hash) of the interacting pair set, reduced on the device (sum of splitmix64(i*N+j) over the pairs, mod 2^64)"""
        cnt, h = ctypes.c_int64(0), ctypes.c_uint64(0)
        _check(self.lib.b200_direct_pairs_hash(self.handle, ctypes.byref(cnt), ctypes.byref(h)))
        return cnt.value, h.value

    def nblist_stats(self):
        s = NblistStats()
        _check(self.lib.b200_nblist_get_stats(self.handle, ctypes.byref(s)))
        return {n: getattr(s, n) for n, _ in s._fields_}

    # IntegrateLangevinMiddleStepKernel
    def set_integrator(self, dt=0.002, temperature=300.0, friction=1.0, constraint_tol=1e-5, seed=1, remove_cm_every=1):
        d = IntegratorDesc(dt, temperature, friction, constraint_tol, seed, remove_cm_every)
        _check(self.lib.b200_integrator_create(self.handle, ctypes.byref(d)))

    def step(self, n):
        _check(self.lib.b200_step(self.handle, int(n)))

    def timed_steps(self, n):
        ms = ctypes.c_double(0.0)
        _check(self.lib.b200_timed_steps(self.handle, int(n), ctypes.byref(ms)))
        return ms.value

    def kinetic_energy(self):
        e = ctypes.c_double(0.0)
        _check(self.lib.b200_kinetic_energy(self.handle, ctypes.byref(e)))
        return e.value

    def kinetic_energy_shifted(self, shift_ps):
        """kinetic energy of v + shift*F/m with the forces of the last compute() (leapfrog integrators: shift = dt/2)"""
        e = ctypes.c_double(0.0)
        _check(self.lib.b200_kinetic_energy_shifted(self.handle, ctypes.c_double(shift_ps), ctypes.byref(e)))
        return e.value

    def apply_constraints(self, tol=1e-5):
        _check(self.lib.b200_apply_constraints(self.handle, ctypes.c_double(tol)))

    def apply_velocity_constraints(self, tol=1e-5):
        _check(self.lib.b200_apply_velocity_constraints(self.handle, ctypes.c_double(tol)))

    # ApplyMonteCarloBarostatKernel
    def barostat_save(self):
        _check(self.lib.b200_barostat_save(self.handle))

    def barostat_scale(self, sx, sy, sz, sxy=0.0, sxz=0.0, syz=0.0, rigid_molecules=True):
        a = (ctypes.c_double * 6)(sx, sy, sz, sxy, sxz, syz)
        _check(self.lib.b200_barostat_scale(self.handle, a, 1 if rigid_molecules else 0))

    def barostat_restore(self):
        _check(self.lib.b200_barostat_restore(self.handle))

    def synchronize(self):
        _check(self.lib.b200_synchronize(self.handle))

    # measurement
    def timing_enable(self, on=True):
        _check(self.lib.b200_timing_enable(self.handle, 1 if on else 0))

    def timing_reset(self):
        _check(self.lib.b200_timing_reset(self.handle))

    def timing_read(self):
        ms = (ctypes.c_double * len(TIMER_NAMES))()
        n = (ctypes.c_int64 * len(TIMER_NAMES))()
        _check(self.lib.b200_timing_read(self.handle, ms, n))
        return {name: (ms[i], n[i]) for i, name in enumerate(TIMER_NAMES)}

    def launch_count(self):
        return int(self.lib.b200_launch_count(self.handle))

    # multi-GPU
    def set_partition(self, rank, nranks):
        _check(self.lib.b200_set_partition(self.handle, rank, nranks))

    def comm_init(self, rank, nranks, unique_id):
        buf = ctypes.create_string_buffer(bytes(unique_id), 128)
        _check(self.lib.b200_comm_init(self.handle, rank, nranks, buf))


def nccl_unique_id():
    buf = ctypes.create_string_buffer(128)
    _check(load_library().b200_nccl_unique_id(buf))
    return buf.raw


def fft3d_r2c(real, device=0):
    """Unnormalised 3D R2C FFT of a float32 array [nx, ny, nz] on the GPU (CudaFFT3D::execFFT forward)."""
    real = np.ascontiguousarray(real, dtype=np.float32)
    nx, ny, nz = real.shape
    out = np.empty((nx, ny, nz // 2 + 1), dtype=np.complex64)
    _check(load_library().b200_fft3d_r2c_host(device, nx, ny, nz, real.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), out.ctypes.data_as(ctypes.POINTER(ctypes.c_float))))
    return out


def fft3d_c2r(spec, nz, device=0):
    spec = np.ascontiguousarray(spec, dtype=np.complex64)
    nx, ny, nzc = spec.shape
    assert nzc == nz // 2 + 1
    out = np.empty((nx, ny, nz), dtype=np.float32)
    _check(load_library().b200_fft3d_c2r_host(device, nx, ny, nz, spec.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), out.ctypes.data_as(ctypes.POINTER(ctypes.c_float))))
    return out


def fft3d_c2r_noalloc(d_in_ptr, d_out_ptr, nx, ny, nz, device=0, stream=0):
    """Allocation-free complex-to-real transform on DEVICE pointers (integers).  For callers that may not allocate device
    memory; limits and sharing rules are in include/b200_nonbonded.h."""
    lib = load_library()
    lib.b200_fft3d_c2r_noalloc.argtypes = [ctypes.c_int32] * 4 + [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    _check(lib.b200_fft3d_c2r_noalloc(device, nx, ny, nz, ctypes.c_void_p(d_in_ptr), ctypes.c_void_p(d_out_ptr), ctypes.c_void_p(stream)))
