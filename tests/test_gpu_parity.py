"""GPU parity tests: the CUDA path (through the C ABI) against the oracle restatement and the committed reference fixtures.
Tolerances are BASELINE.json's: forces 1e-5 relative RMS, energy 1e-5 relative, pair set bit-exact, run-to-run bit-identical."""
import os

import numpy as np
import pytest

import openmm_b200 as b
from openmm_b200 import workloads
from oracle_lib import Oracle, pair_hash, rel_rms, oracle_fft3d_r2c

pytestmark = pytest.mark.gpu

FORCE_TOL = 1e-5
ENERGY_TOL = 1e-5


def make(n, seed=None, **kw):
    sysd = workloads.water_box(n, seed=n if seed is None else seed)
    ctx = workloads.make_context(sysd, **kw)
    ctx.set_positions(sysd["positions"])
    return sysd, ctx


# the last four are the meshes of the BASELINE.json configurations: STMV 196^3 (14 x 14 two-stage plan), Cellulose 144^3, ApoA1 90^3 and
# its LJPME dispersion mesh 45^3 (odd nz: generic shared-memory passes)
@pytest.mark.parametrize("shape", [(28, 25, 30), (25, 25, 28), (21, 25, 27), (56, 56, 56), (90, 98, 70), (144, 36, 48),
                                   (196, 196, 196), (144, 144, 144), (90, 90, 90), (45, 45, 45)])
def test_fft3d_r2c_c2r(shape):
    # platforms/cuda/tests/TestCudaFFT3D.cpp:56-121 (same sizes, vs pocketfft there; vs the oracle FFT + numpy here)
    rng = np.random.default_rng(1)
    a = rng.normal(size=shape).astype(np.float32)
    got = b.fft3d_r2c(a)
    ref = np.fft.rfftn(a.astype(np.float64))
    scale = np.abs(ref).max()
    assert np.abs(got - ref).max() / scale < 2e-6
    if np.prod(shape) <= 30000:
        assert np.abs(oracle_fft3d_r2c(a.astype(np.float64)) - ref).max() < 1e-8
    back = b.fft3d_c2r(got, shape[2])
    assert np.abs(back / np.prod(shape) - a).max() < 5e-6


@pytest.mark.parametrize("n", [1536, 6000, 23558])
def test_forces_and_energy_vs_reference_fixture(golden_dir, n):
    g = np.load(os.path.join(golden_dir, f"ref_water_{n}.npz"))
    sysd, ctx = make(n)
    assert np.array_equal(sysd["positions"].astype(np.float32), g["positions"])
    assert list(ctx.pme[1]) == [int(x) for x in g["grid"]]
    e_dir = ctx.compute(b.GROUP_DIRECT)
    f_dir = ctx.get_forces()
    e_rec = ctx.compute(b.GROUP_RECIPROCAL)
    f_rec = ctx.get_forces()
    e_tot = ctx.compute(b.GROUP_ALL)
    f_tot = ctx.get_forces()
    ref_tot = g["f_direct"].astype(np.float64) + g["f_recip"].astype(np.float64)
    print(n, "relRMS direct/recip/total", rel_rms(f_dir, g["f_direct"]), rel_rms(f_rec, g["f_recip"]), rel_rms(f_tot, ref_tot))
    assert rel_rms(f_tot, ref_tot) < FORCE_TOL
    assert rel_rms(f_dir, g["f_direct"]) < FORCE_TOL
    assert rel_rms(f_rec, g["f_recip"]) < 2e-5   # reciprocal part alone is ~8x smaller than the total force
    e_ref = float(g["e_direct"]) + float(g["e_recip"])
    assert abs(e_tot - e_ref) / abs(e_ref) < ENERGY_TOL
    assert abs(e_dir - float(g["e_direct"])) / abs(float(g["e_direct"])) < ENERGY_TOL
    assert abs(e_rec - float(g["e_recip"])) / abs(float(g["e_recip"])) < ENERGY_TOL
    assert abs((e_dir + e_rec) - e_tot) / abs(e_ref) < 1e-6     # testDirectAndReciprocal: groups add up
    ctx.close()


@pytest.mark.parametrize("n", [1536, 6000, 23558])
def test_pair_set_bit_exact(golden_dir, n):
    g = np.load(os.path.join(golden_dir, f"ref_water_{n}.npz"))
    sysd, ctx = make(n)
    cnt, pairs = ctx.direct_pairs(capacity=int(g["pair_count"]) + 1024)
    assert cnt == int(g["pair_count"])
    c2, h = pair_hash(pairs, n)
    assert (c2, h) == (int(g["pair_count"]), int(g["pair_hash"]))
    if n <= 6000:
        ref = Oracle(sysd, 0.9, *ctx.pme).pairs()
        key = lambda p: np.sort(p[:, 0].astype(np.int64) * n + p[:, 1])
        assert np.array_equal(key(pairs), key(ref))
    ctx.close()


LARGE = [("apoa1", 92224, False), ("apoa1_ljpme", 92224, True), ("cellulose", 408609, False), ("stmv", 1066628, False)]


@pytest.mark.parametrize("name,n,ljpme", LARGE, ids=[x[0] for x in LARGE])
def test_baseline_configs_vs_reference_fixture(golden_dir, name, n, ljpme):
    """The BASELINE.json configurations themselves (tests/TestNonbondedForce.h:447-537 testLargeSystem at the benchmarked sizes): forces,
    energies and the pair set against one Reference-platform evaluation stored compactly by tests/golden/make_fixtures.py large()
    (forces of every 16th atom; count + hash of the pair set)."""
    import hashlib
    g = np.load(os.path.join(golden_dir, f"ref_large_ljpme_{n}.npz" if ljpme else f"ref_large_{n}.npz"))
    sysd = workloads.cached_water_box(n, ljpme=ljpme)
    assert hashlib.sha256(sysd["positions"].astype(np.float32).tobytes()).hexdigest() == str(g["positions_sha256"])
    ctx = workloads.make_context(sysd, ljpme=ljpme)
    assert list(ctx.pme[1]) == [int(x) for x in g["grid"]]
    if ljpme:
        assert list(ctx.lj_pme[1]) == [int(x) for x in g["lj_grid"]]
    ctx.set_positions(sysd["positions"])
    st = int(g["stride"])
    e_dir = ctx.compute(b.GROUP_DIRECT); f_dir = ctx.get_forces()[::st]
    e_rec = ctx.compute(b.GROUP_RECIPROCAL); f_rec = ctx.get_forces()[::st]
    e_tot = ctx.compute(b.GROUP_ALL); f_tot = ctx.get_forces()[::st]
    rd, rr = g["f_direct"].astype(np.float64), g["f_recip"].astype(np.float64)
    # the stored forces are float32 (6e-8 relative): far below the tolerance
    errs = rel_rms(f_dir, rd), rel_rms(f_rec, rr), rel_rms(f_tot, rd + rr)
    print(name, "relRMS direct/recip/total", errs)
    assert errs[0] < FORCE_TOL and errs[2] < FORCE_TOL
    assert errs[1] < 2e-5      # reciprocal part alone is ~8x smaller than the total force
    e_ref = float(g["e_direct"]) + float(g["e_recip"])
    assert abs(e_tot - e_ref) / abs(e_ref) < ENERGY_TOL
    assert abs(e_dir - float(g["e_direct"])) / abs(float(g["e_direct"])) < ENERGY_TOL
    assert abs(e_rec - float(g["e_recip"])) / abs(float(g["e_recip"])) < ENERGY_TOL
    if not ljpme:   # same positions and exclusions as the PME case
        cnt, h = ctx.direct_pairs_hash()
        assert (cnt, h) == (int(g["pair_count"]), int(g["pair_hash"]))
    # a second evaluation is bit-identical (fixed-point sums) at this size too
    e2 = ctx.compute(b.GROUP_ALL)
    assert e2 == e_tot and np.array_equal(ctx.get_forces()[::st], f_tot)
    ctx.close()


def test_pair_hash_on_device_matches_list(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_water_6000.npz"))
    sysd, ctx = make(6000)
    assert ctx.direct_pairs_hash() == (int(g["pair_count"]), int(g["pair_hash"]))
    ctx.close()


def test_run_to_run_bit_identical():
    # platforms/cuda/tests/TestCudaNonbondedForce.cpp:35-71 testDeterministicForces
    sysd, ctx = make(6000)
    e1 = ctx.compute()
    f1 = ctx.get_forces().copy()
    e2 = ctx.compute()
    f2 = ctx.get_forces()
    assert np.array_equal(f1, f2) and e1 == e2
    ctx.close()
    sysd, ctx2 = make(6000)
    e3 = ctx2.compute()
    assert np.array_equal(f1, ctx2.get_forces()) and e1 == e3
    ctx2.close()


@pytest.mark.parametrize("n", [1538, 6002])
def test_ljpme_vs_reference_fixture(golden_dir, n):
    # NonbondedForce::LJPME: dispersion on a second mesh (tests/TestNonbondedForce.h testLJPME; benchmark "apoa1pme" with LJPME)
    g = np.load(os.path.join(golden_dir, f"ref_ljpme_{n}.npz"))
    sysd = workloads.ljpme_water_box(n)
    assert np.array_equal(sysd["positions"].astype(np.float32), g["positions"])
    ctx = workloads.make_context(sysd, ljpme=True)
    assert list(ctx.pme[1]) == [int(x) for x in g["grid"]] and list(ctx.lj_pme[1]) == [int(x) for x in g["lj_grid"]]
    ctx.set_positions(sysd["positions"])
    e_dir = ctx.compute(b.GROUP_DIRECT); f_dir = ctx.get_forces()
    e_rec = ctx.compute(b.GROUP_RECIPROCAL); f_rec = ctx.get_forces()
    e_tot = ctx.compute(b.GROUP_ALL); f_tot = ctx.get_forces()
    ref_tot = g["f_direct"] + g["f_recip"]
    print(n, "LJPME relRMS direct/recip/total", rel_rms(f_dir, g["f_direct"]), rel_rms(f_rec, g["f_recip"]), rel_rms(f_tot, ref_tot))
    assert rel_rms(f_tot, ref_tot) < FORCE_TOL
    assert rel_rms(f_dir, g["f_direct"]) < FORCE_TOL
    assert rel_rms(f_rec, g["f_recip"]) < 2e-5
    e_ref = float(g["e_direct"]) + float(g["e_recip"])
    assert abs(e_tot - e_ref) / abs(e_ref) < ENERGY_TOL
    assert abs(e_dir - float(g["e_direct"])) / abs(float(g["e_direct"])) < ENERGY_TOL
    assert abs(e_rec - float(g["e_recip"])) / abs(float(g["e_recip"])) < ENERGY_TOL
    # energy-only and forces+energy kernels agree; forces-only kernel gives the same forces
    ctx.compute(b.GROUP_ALL, energy=False)
    assert rel_rms(ctx.get_forces(), f_tot) < 1e-7
    ctx.close()


def test_ljpme_oracle_on_fresh_seed_and_determinism():
    sysd = workloads.ljpme_water_box(3002)
    ctx = workloads.make_context(sysd, ljpme=True)
    ctx.set_positions(sysd["positions"])
    e = ctx.compute(); f = ctx.get_forces()
    fo, eo = Oracle(sysd, 0.9, ctx.pme[0], ctx.pme[1], ctx.lj_pme[0], ctx.lj_pme[1]).nonbonded()
    assert rel_rms(f, fo) < FORCE_TOL and abs(e - eo) / abs(eo) < ENERGY_TOL
    e2 = ctx.compute(); f2 = ctx.get_forces()
    assert e2 == e and np.array_equal(f, f2)
    ctx.close()


def test_switching_function_vs_reference_fixture(golden_dir):
    # setUseSwitchingFunction(true), switching distance 0.75 nm: direct-space forces/energy against the Reference platform,
    # the energy-only kernel agrees with forces+energy, and a larger system (single periodic copy path) against the oracle
    g = np.load(os.path.join(golden_dir, "ref_switch_1536.npz"))
    sysd = workloads.water_box(1536, seed=1536)
    ctx = workloads.make_context(sysd, switch_distance=float(g["switch_distance"]))
    ctx.set_positions(sysd["positions"])
    e_dir = ctx.compute(b.GROUP_DIRECT); f_dir = ctx.get_forces()
    assert rel_rms(f_dir, g["f_direct"]) < FORCE_TOL
    assert abs(e_dir - float(g["e_direct"])) / abs(float(g["e_direct"])) < ENERGY_TOL
    e_only = ctx.compute(b.GROUP_DIRECT, forces=False)
    assert abs(e_only - e_dir) / abs(e_dir) < 1e-7
    ctx.close()
    sysd = workloads.water_box(6000, seed=11)
    ctx = workloads.make_context(sysd, switch_distance=0.8)
    ctx.set_positions(sysd["positions"])
    e = ctx.compute(); f = ctx.get_forces()
    fo, eo = Oracle(sysd, 0.9, ctx.pme[0], ctx.pme[1], switch_distance=0.8).nonbonded()
    assert rel_rms(f, fo) < FORCE_TOL and abs(e - eo) / abs(eo) < ENERGY_TOL
    ctx.close()


def test_triclinic_ions_gromacs_and_reference(golden_dir):
    # tests/TestEwald.h:220-269 testTriclinic: 8 ions in a triclinic cell, setPMEParameters(3.45891, 32, 40, 48), cutoff 1 nm
    pos = np.array([[1.744, 2.788, 3.162], [1.048, 0.762, 2.340], [2.489, 1.570, 2.817], [1.027, 1.893, 3.271],
                    [0.937, 0.825, 0.009], [2.290, 1.887, 3.352], [1.266, 1.111, 2.894], [0.933, 1.862, 3.490]])
    sysd = dict(n=8, positions=pos, box=np.array([2.5, 0, 0, 0.5, 3.0, 0, 0.7, 0.9, 3.5]), charge=np.array([-1.0] * 4 + [1.0] * 4),
                sigma=np.array([0.440104] * 4 + [0.332840] * 4), epsilon=np.array([0.4184] * 4 + [0.0115897] * 4), mass=np.ones(8),
                exclusions=np.zeros((0, 2), dtype=np.int32))
    ctx = workloads.make_context(sysd, cutoff=1.0, alpha=3.45891, grid=[32, 40, 48])
    ctx.set_positions(pos)
    e = ctx.compute(); f = ctx.get_forces()
    g = np.load(os.path.join(golden_dir, "ref_ions_triclinic.npz"))
    ref_f = g["f_direct"] + g["f_recip"]; ref_e = float(g["e_direct"]) + float(g["e_recip"])
    # ions 3 and 7 sit 0.24 nm apart, where a 4e3 kJ/mol/nm LJ repulsion cancels most of the Coulomb attraction: the float pair
    # arithmetic of mixed precision leaves 1.07e-5 here (measured), the reference's own tolerance for this case is 1e-4
    print("triclinic relRMS", rel_rms(f, ref_f))
    assert rel_rms(f, ref_f) < 2e-5 and abs(e - ref_e) / abs(ref_e) < ENERGY_TOL
    gromacs_f = np.array([[4.25253e+01, -1.23503e+02, 1.22139e+02], [9.74752e+01, 1.68213e+02, 1.93169e+02], [-1.50348e+02, 1.29165e+02, 3.70435e+02],
                          [9.18644e+02, -3.52571e+00, -1.34772e+03], [-1.61193e+02, 9.01528e+01, -7.12904e+01], [2.82630e+02, 2.78029e+01, -3.72864e+02],
                          [-1.47454e+02, -2.14448e+02, -3.55789e+02], [-8.82195e+02, -7.39132e+01, 1.46202e+03]])
    assert abs(e - (-963.370)) / 963.370 < 1e-4
    for i in range(8):
        assert np.linalg.norm(f[i] - gromacs_f[i]) / max(np.linalg.norm(gromacs_f[i]), 1.0) < 1e-4
    ctx.close()


def test_nacl_known_answers(golden_dir):
    # tests/TestEwald.h:47-94: rock-salt crystal, Madelung energy (tolerance 1e-3 there); every ion is equivalent, so forces vanish
    idx = np.array([(i, j, k) for i in range(10) for j in range(10) for k in range(10)])
    q = np.where(idx.sum(1) % 2 == 0, 1.0, -1.0)
    sysd = dict(n=1000, positions=0.141 + 0.282 * idx, box=np.array([2.82, 0, 0, 0, 2.82, 0, 0, 0, 2.82]), charge=q, sigma=np.ones(1000),
                epsilon=np.zeros(1000), mass=np.ones(1000), exclusions=np.zeros((0, 2), dtype=np.int32))
    ctx = workloads.make_context(sysd, cutoff=1.0, alpha=float(np.sqrt(-np.log(2e-5)) / 1.0), grid=[60, 60, 60])
    ctx.set_positions(sysd["positions"])
    e = ctx.compute(); f = ctx.get_forces()
    exact = -(1.7476 * 1.6022e-19 * 1.6022e-19 * 6.02214076e23 * 1000) / (1.112e-10 * 0.282e-9 * 2 * 1000)
    assert abs(e - exact) / abs(exact) < 1e-3
    assert np.abs(f).max() < 1e-2 * 138.935 / 0.282 ** 2     # << the nearest-neighbour pair force
    ctx.close()
    # tests/TestEwald.h:96-141: 894 amorphous ions, cutoff 1.2 nm, Gromacs energy -3.82047e5; forces against the oracle
    pos = np.load(os.path.join(golden_dir, "nacl_amorph.npy")).astype(np.float64)
    n, L = 894, 3.00646
    sysd = dict(n=n, positions=pos, box=np.array([L, 0, 0, 0, L, 0, 0, 0, L]), charge=np.array([1.0] * (n // 2) + [-1.0] * (n // 2)), sigma=np.ones(n),
                epsilon=np.zeros(n), mass=np.ones(n), exclusions=np.zeros((0, 2), dtype=np.int32))
    alpha = float(np.sqrt(-np.log(2e-5)) / 1.2)
    ctx = workloads.make_context(sysd, cutoff=1.2, alpha=alpha, grid=[96, 96, 96])
    ctx.set_positions(pos)
    e = ctx.compute(); f = ctx.get_forces()
    assert abs(e - (-3.82047e5)) / 3.82047e5 < 2e-5
    fo, eo = Oracle(sysd, 1.2, alpha, [96, 96, 96]).nonbonded()
    assert rel_rms(f, fo) < FORCE_TOL and abs(e - eo) / abs(eo) < ENERGY_TOL
    ctx.close()


def test_ljpme_gromacs_known_answers(golden_dir):
    # tests/TestDispersionPME.h:812-965, 1221-1274 (uncharged CHARMM water dimer, tolerances 1e-4 / 5e-4 there) and :1773-2267
    # (125 waters with exclusions, Gromacs energy, 5e-5), same data as tests/test_oracle.py
    from test_oracle import _water2_ljpme, _assert_vec
    alpha = 4.0124063605
    for excl, e_ref, tol, f_ref in (
            (False, 1.34804e+03, 1e-4, [[3.15301e-01, -3.91114e-03, -6.68965e+04], [1.67241e+04, -3.79846e-02, 3.34487e+04], [-1.67243e+04, -9.55931e-02, 3.34486e+04],
                                        [-1.71643e+00, -1.76696e+00, -6.68993e+04], [1.67254e+04, 1.59080e+00, 3.34495e+04], [-1.67239e+04, 3.13897e-01, 3.34491e+04]]),
            (True, -7.78060e-01, 5e-4, [[3.15301e-01, -3.91114e-03, 9.48373e-01], [-4.74745e-02, -3.79846e-02, -5.69616e-02], [-7.16866e-02, -9.55931e-02, -1.43348e-01],
                                        [-1.78136e+00, -1.76696e+00, -1.70022e+00], [1.19309e+00, 1.59080e+00, 7.95443e-01], [3.92366e-01, 3.13897e-01, 1.56962e-01]])):
        sysd = _water2_ljpme(excl)
        ctx = workloads.make_context(sysd, cutoff=0.7, alpha=alpha, grid=[32] * 3, lj_alpha=alpha, lj_grid=[32] * 3)
        ctx.set_positions(sysd["positions"])
        e = ctx.compute(); f = ctx.get_forces()
        assert abs(e - e_ref) / max(abs(e_ref), 1.0) < tol
        _assert_vec(np.array(f_ref), f, tol)
        ctx.close()
    pos = np.load(os.path.join(golden_dir, "water125_dpme.npy")) * 0.1
    ow = 3 * np.arange(125)
    ex = np.stack([np.stack([ow, ow + 1], 1), np.stack([ow, ow + 2], 1), np.stack([ow + 1, ow + 2], 1)], 1).reshape(-1, 2).astype(np.int32)
    sysd = dict(n=375, positions=pos, box=np.array([2.301, 0, 0, 0, 2.301, 0, 0, 0, 2.301]), charge=np.zeros(375),
                sigma=2.0 * 2.0 ** (-1.0 / 6.0) * np.array([1.7682, 0.2245, 0.2245] * 125) * 0.1, epsilon=np.array([0.1521, 0.046, 0.046] * 125) * 4.184,
                mass=np.ones(375), exclusions=ex)
    ctx = workloads.make_context(sysd, cutoff=1.15, alpha=4.5, grid=[60] * 3, lj_alpha=4.2760443169, lj_grid=[60] * 3)
    ctx.set_positions(pos)
    e = ctx.compute()
    assert abs(e - (-2.17481e+02)) / 2.17481e+02 < 5e-5
    ctx.close()


def test_update_parameters_in_context():
    # updateParametersInContext (tests/TestNonbondedForce.h testChangingParameters): new charges and LJ parameters after a few
    # steps (the device order has been re-sorted by then), compared with the oracle evaluated on the new parameters
    sysd, ctx = make(3000, seed=21)
    ctx.set_velocities(sysd["velocities"])
    ctx.set_integrator(dt=0.002, seed=3)
    ctx.step(5)
    pos = ctx.get_positions()
    new = dict(sysd)
    new["charge"] = sysd["charge"] * 0.9
    new["sigma"] = sysd["sigma"] * 1.05
    new["epsilon"] = sysd["epsilon"] * 1.2
    ctx.update_nonbonded_parameters(new["charge"], new["sigma"], new["epsilon"], len(sysd["exclusions"]))
    e = ctx.compute(); f = ctx.get_forces()
    new["positions"] = pos

    def check(f, e):
        # with the enlarged sigma the direct (+1e5) and reciprocal (-1e5) energies nearly cancel, so the energy is judged
        # against the size of the parts rather than against their small sum
        o = Oracle(new, 0.9, *ctx.pme)
        (fd, ed), (fr, er) = o.nonbonded(1), o.nonbonded(2)
        assert rel_rms(f, fd + fr) < FORCE_TOL and abs(e - (ed + er)) / (abs(ed) + abs(er)) < ENERGY_TOL

    check(f, e)
    # one molecule made different from all others: its group must drop out of the re-sort (identical molecules only are exchanged),
    # so after the next re-sort (every 250 steps) every particle still carries its own parameters
    new["charge"] = new["charge"].copy(); new["charge"][0:3] *= 1.1
    ctx.update_nonbonded_parameters(new["charge"], new["sigma"], new["epsilon"], len(sysd["exclusions"]))
    ctx.step(260)
    new["positions"] = ctx.get_positions()
    e = ctx.compute(); f = ctx.get_forces()
    check(f, e)
    # changing the number of exclusions is refused with the reference's message
    with pytest.raises(b.B200Error, match="set of non-excluded exceptions has changed"):
        ctx.update_nonbonded_parameters(new["charge"], new["sigma"], new["epsilon"], len(sysd["exclusions"]) + 1)
    ctx.close()


def test_oracle_vs_gpu_on_fresh_seed():
    # same seeded inputs through the oracle restatement (not a stored fixture)
    sysd, ctx = make(3000, seed=77)
    e = ctx.compute()
    f = ctx.get_forces()
    fo, eo = Oracle(sysd, 0.9, *ctx.pme).nonbonded()
    assert rel_rms(f, fo) < FORCE_TOL and abs(e - eo) / abs(eo) < ENERGY_TOL
    ctx.close()


def test_partition_emulation_matches_single_rank():
    # tests/TestNonbondedForce.h:1013-1083 testParallelComputation: N logical ranks on one device, integer sums => bit-identical
    sysd, ctx = make(6000)
    ctx.compute(forces=True, energy=False)
    f1 = ctx.get_forces().copy()
    total = np.zeros_like(f1)
    for r in range(3):
        c = workloads.make_context(sysd)
        c.set_partition(r, 3)
        c.set_positions(sysd["positions"])
        c.compute(forces=True, energy=False)
        total += c.get_forces()
        c.close()
    assert np.abs(total - f1).max() < 1e-6 * np.abs(f1).max()
    ctx.close()


def test_dynamics_conserves_constraints_and_temperature():
    sysd, ctx = make(6000)
    ctx.set_velocities(sysd["velocities"])
    ctx.set_integrator(dt=0.002, temperature=300.0, friction=1.0, seed=3)
    ctx.step(300)
    pos = ctx.get_positions()
    nw = sysd["num_waters"]
    o, h1, h2 = pos[0:3 * nw:3], pos[1:3 * nw:3], pos[2:3 * nw:3]
    assert np.abs(np.linalg.norm(h1 - o, axis=1) - workloads.D_OH).max() < 1e-5
    assert np.abs(np.linalg.norm(h2 - o, axis=1) - workloads.D_OH).max() < 1e-5
    assert np.abs(np.linalg.norm(h2 - h1, axis=1) - workloads.D_HH).max() < 1e-5
    ke = ctx.kinetic_energy()
    dof = 6 * nw - 3
    T = 2 * ke / (dof * 0.00831446261815324)
    assert 250 < T < 350, T
    # the neighbour list stayed valid through rebuilds and a reorder: forces still match the oracle on the new coordinates
    sys2 = dict(sysd); sys2["positions"] = ctx.get_positions(enforce_periodic=True)
    e = ctx.compute()
    f = ctx.get_forces()
    fo, eo = Oracle(sys2, 0.9, *ctx.pme).nonbonded()
    assert rel_rms(f, fo) < FORCE_TOL and abs(e - eo) / abs(eo) < ENERGY_TOL
    st = ctx.nblist_stats()
    assert st["rebuilds"] >= 2
    ctx.close()
