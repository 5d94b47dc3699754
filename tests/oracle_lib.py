"""ctypes binding of oracle/liboracle_nb.so (the plain-C restatement).  TEST INFRASTRUCTURE: imported only from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "liboracle_nb.so")
REF_DRIVER = os.path.join(ORACLE_DIR, "_ref", "ref_driver")
REF_PLUGINS = os.path.join(ORACLE_DIR, "_ref", "plugins")

c_double_p = ctypes.POINTER(ctypes.c_double)
c_int_p = ctypes.POINTER(ctypes.c_int)


class OracleSystem(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int), ("pos", c_double_p), ("charge", c_double_p), ("sigma", c_double_p), ("epsilon", c_double_p),
                ("box", ctypes.c_double * 9), ("cutoff", ctypes.c_double), ("alpha", ctypes.c_double), ("grid", ctypes.c_int * 3),
                ("nexcl", ctypes.c_int), ("excl", c_int_p), ("nexc", ctypes.c_int), ("exc", c_int_p),
                ("exc_qq", c_double_p), ("exc_sigma", c_double_p), ("exc_eps", c_double_p), ("exceptions_periodic", ctypes.c_int),
                ("ljpme", ctypes.c_int), ("lj_alpha", ctypes.c_double), ("lj_grid", ctypes.c_int * 3),
                ("use_switch", ctypes.c_int), ("switch_distance", ctypes.c_double),
                ("nb_method", ctypes.c_int), ("rf_dielectric", ctypes.c_double), ("ewald_kmax", ctypes.c_int * 3)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "restatement"])
        _lib = ctypes.CDLL(ORACLE_SO)
        _lib.oracle_direct.restype = ctypes.c_longlong
        _lib.oracle_self_energy.restype = ctypes.c_double
    return _lib


class Oracle:
    def __init__(self, sysd, cutoff, alpha, grid, lj_alpha=None, lj_grid=None, switch_distance=None, method=None, rf_dielectric=78.3, ewald_kmax=None):
        self.keep = []
        s = OracleSystem()

        def dbl(a):
            a = np.ascontiguousarray(a, dtype=np.float64); self.keep.append(a); return a.ctypes.data_as(c_double_p)

        def i32(a):
            a = np.ascontiguousarray(a, dtype=np.int32); self.keep.append(a); return a.ctypes.data_as(c_int_p)

        self.n = s.n = int(sysd["n"])
        s.pos = dbl(np.asarray(sysd["positions"]).reshape(-1))
        s.charge, s.sigma, s.epsilon = dbl(sysd["charge"]), dbl(sysd["sigma"]), dbl(sysd["epsilon"])
        for k in range(9):
            s.box[k] = float(sysd["box"][k])
        s.cutoff, s.alpha = cutoff, alpha
        for k in range(3):
            s.grid[k] = int(grid[k])
        ex = np.asarray(sysd.get("exclusions", np.zeros((0, 2))), dtype=np.int32).reshape(-1)
        s.nexcl, s.excl = len(ex) // 2, i32(ex)
        exc = sysd.get("exceptions")
        if exc is not None and len(exc[1]) > 0:
            s.nexc, s.exc = len(exc[1]), i32(np.asarray(exc[0]).reshape(-1))
            s.exc_qq, s.exc_sigma, s.exc_eps = dbl(exc[1]), dbl(exc[2]), dbl(exc[3])
        s.exceptions_periodic = 0
        if lj_alpha is not None:
            s.ljpme, s.lj_alpha = 1, lj_alpha
            for k in range(3):
                s.lj_grid[k] = int(lj_grid[k])
        if switch_distance is not None:
            s.use_switch, s.switch_distance = 1, switch_distance
        if method is not None:       # NonbondedForce::NonbondedMethod 0 NoCutoff, 1 CutoffNonPeriodic, 2 CutoffPeriodic (3, 4, 5: the Ewald family)
            s.nb_method = {0: 1, 1: 2, 2: 3}.get(int(method), 0)
            s.rf_dielectric = rf_dielectric
        if ewald_kmax is not None:   # NonbondedForce::Ewald: explicit k-space sum
            for k in range(3):
                s.ewald_kmax[k] = int(ewald_kmax[k])
        self.s = s

    def nonbonded(self, groups=3):
        f = np.zeros(3 * self.n)
        e = ctypes.c_double(0.0)
        lib().oracle_nonbonded(ctypes.byref(self.s), groups, f.ctypes.data_as(c_double_p), ctypes.byref(e))
        return f.reshape(-1, 3), e.value

    def pairs(self):
        cap = max(1024, int(self.n) * 400)
        buf = np.empty((cap, 2), dtype=np.int32)
        cnt = lib().oracle_direct(ctypes.byref(self.s), None, None, buf.ctypes.data_as(c_int_p), ctypes.c_longlong(cap))
        assert cnt <= cap
        return buf[:cnt]


def ewald_parameters(cutoff, box, tol=5e-4):
    """alpha and kmax as NonbondedForceImpl::calcEwaldParameters (openmmapi/src/NonbondedForceImpl.cpp:159-190,218-230)"""
    alpha = np.sqrt(-np.log(2.0 * tol)) / cutoff

    def find_zero(width):
        f = lambda a: tol - 0.05 * np.sqrt(width * alpha) * a * np.exp(-(a * np.pi / (width * alpha)) ** 2)
        arg = 10
        v = f(arg)
        if v > 0.0:
            while v > 0.0 and arg > 0:
                arg -= 1
                v = f(arg)
            return arg + 1
        while v < 0.0:
            arg += 1
            v = f(arg)
        return arg

    kmax = [find_zero(w) for w in (box[0], box[4], box[8])]
    return float(alpha), [k + 1 if k % 2 == 0 else k for k in kmax]


def pair_hash(pairs, n):
    """order-independent 64-bit hash of a pair set: sum over pairs of splitmix64(i*n+j) mod 2^64, plus the count"""
    p = np.asarray(pairs, dtype=np.uint64)
    lo = np.minimum(p[:, 0], p[:, 1]); hi = np.maximum(p[:, 0], p[:, 1])
    z = lo * np.uint64(n) + hi
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
        return int(len(p)), int(np.sum(z, dtype=np.uint64))


def oracle_fft3d_r2c(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    nx, ny, nz = a.shape
    out = np.empty((nx, ny, nz // 2 + 1), dtype=np.complex128)
    lib().oracle_fft3d_r2c(a.ctypes.data_as(c_double_p), out.ctypes.data_as(c_double_p), nx, ny, nz)
    return out


def rel_rms(a, ref):
    a = np.asarray(a, dtype=np.float64); ref = np.asarray(ref, dtype=np.float64)
    return float(np.sqrt(np.sum((a - ref) ** 2) / np.sum(ref ** 2)))
