"""GPU tests of the small kernels that ride along with the nonbonded path: HarmonicBond / HarmonicAngle / PeriodicTorsion
(known answers of the reference's tests/TestHarmonicBondForce.h, TestHarmonicAngleForce.h, TestPeriodicTorsionForce.h and a
double-precision numpy restatement differentiated numerically), the 1-4 exception kernel (tests/TestNonbondedForce.h:98-179
testExclusionsAnd14, and a water box with non-zero exception parameters against the oracle), the Verlet limit of the
integrator, and ApplyConstraintsKernel::apply / applyToVelocities."""
import numpy as np
import pytest

import openmm_b200 as b
from openmm_b200 import workloads
from oracle_lib import Oracle, rel_rms

pytestmark = pytest.mark.gpu

BOX = np.array([10.0, 0, 0, 0, 10.0, 0, 0, 0, 10.0])
TOL = 1e-5           # the reference's TOL in these tests


def bare_context(n, mass=1.0):
    return b.Context(np.full(n, mass), BOX)


def test_harmonic_bond_known_answers():
    # tests/TestHarmonicBondForce.h:44-67 testBonds
    ctx = bare_context(3)
    ctx.add_harmonic_bonds([[0, 1], [1, 2]], [1.5, 1.2], [0.8, 0.7])
    ctx.set_positions(np.array([[0, 2, 0], [0, 0, 0], [1, 0, 0]], dtype=float))
    e = ctx.compute(b.GROUP_BONDS)
    f = ctx.get_forces()
    assert np.allclose(f[0], [0, -0.8 * 0.5, 0], atol=TOL) and np.allclose(f[2], [0.7 * 0.2, 0, 0], atol=TOL)
    assert np.allclose(f[1], -f[0] - f[2], atol=TOL)
    assert abs(e - (0.5 * 0.8 * 0.25 + 0.5 * 0.7 * 0.04)) < TOL
    ctx.close()


def test_harmonic_angle_known_answers():
    # tests/TestHarmonicAngleForce.h:44-77 testAngles
    ctx = bare_context(4)
    ctx.add_harmonic_angles([[0, 1, 2], [1, 2, 3]], [np.pi / 3, np.pi / 2], [1.1, 1.2])
    ctx.set_positions(np.array([[0, 1, 0], [0, 0, 0], [1, 0, 0], [2, 1, 0]], dtype=float))
    e = ctx.compute(b.GROUP_ANGLES)
    f = ctx.get_forces()
    t1, t2 = 1.1 * np.pi / 6, 1.2 * np.pi / 4
    assert np.allclose(f[0], [t1, 0, 0], atol=TOL) and np.allclose(f[3], [-0.5 * t2, 0.5 * t2, 0], atol=TOL)
    assert np.allclose(f.sum(0), 0, atol=TOL)
    assert abs(e - (0.5 * 1.1 * (np.pi / 6) ** 2 + 0.5 * 1.2 * (np.pi / 4) ** 2)) < TOL
    ctx.close()


def test_periodic_torsion_known_answers():
    # tests/TestPeriodicTorsionForce.h:44-72 testPeriodicTorsions
    ctx = bare_context(4)
    ctx.add_periodic_torsions([[0, 1, 2, 3]], [2], [np.pi / 3], [1.1])
    ctx.set_positions(np.array([[0, 1, 0], [0, 0, 0], [1, 0, 0], [1, 0, 2]], dtype=float))
    e = ctx.compute(b.GROUP_TORSIONS)
    f = ctx.get_forces()
    torque = -2 * 1.1 * np.sin(2 * np.pi / 3)
    assert np.allclose(f[0], [0, 0, torque], atol=TOL) and np.allclose(f[3], [0, 0.5 * torque, 0], atol=TOL)
    assert np.allclose(f.sum(0), 0, atol=TOL)
    assert abs(e - 1.1 * (1 + np.cos(2 * np.pi / 3))) < TOL
    ctx.close()


# ---- numpy restatement (double precision) of harmonicBondForce.cc, harmonicAngleForce.cc, periodicTorsionForce.cc
def bonded_energy(pos, bonds, angles, torsions):
    e = 0.0
    for (i, j), r0, k in zip(*bonds):
        e += 0.5 * k * (np.linalg.norm(pos[j] - pos[i]) - r0) ** 2
    for (i, j, k_), t0, k in zip(*angles):
        v0, v1 = pos[i] - pos[j], pos[k_] - pos[j]
        c = np.dot(v0, v1) / np.sqrt(np.dot(v0, v0) * np.dot(v1, v1))
        e += 0.5 * k * (np.arccos(np.clip(c, -1, 1)) - t0) ** 2
    for (a, bb, c, d), n, phase, k in zip(*torsions):
        b1, b2, b3 = pos[bb] - pos[a], pos[c] - pos[bb], pos[d] - pos[c]
        n1, n2 = np.cross(b1, b2), np.cross(b2, b3)
        phi = np.arctan2(np.linalg.norm(b2) * np.dot(b1, n2), np.dot(n1, n2))
        e += k * (1 + np.cos(n * phi - phase))
    return e


def test_bonded_terms_vs_numpy_restatement():
    rng = np.random.default_rng(4)
    nm, L = 40, 6          # 40 chains of 6 atoms
    pos = []
    for m in range(nm):
        p = rng.random(3) * 6 + 2
        chain = [p]
        for _ in range(L - 1):
            d = rng.normal(size=3); d *= 0.15 / np.linalg.norm(d)
            chain.append(chain[-1] + d)
        pos += chain
    pos = np.array(pos).astype(np.float32).astype(np.float64)
    idx = np.arange(nm * L).reshape(nm, L)
    bond_atoms = np.concatenate([np.stack([idx[:, k], idx[:, k + 1]], 1) for k in range(L - 1)])
    angle_atoms = np.concatenate([np.stack([idx[:, k], idx[:, k + 1], idx[:, k + 2]], 1) for k in range(L - 2)])
    tors_atoms = np.concatenate([np.stack([idx[:, k], idx[:, k + 1], idx[:, k + 2], idx[:, k + 3]], 1) for k in range(L - 3)])
    bonds = (bond_atoms, rng.uniform(0.1, 0.2, len(bond_atoms)), rng.uniform(1e4, 3e5, len(bond_atoms)))
    angles = (angle_atoms, rng.uniform(1.5, 2.2, len(angle_atoms)), rng.uniform(100, 800, len(angle_atoms)))
    torsions = (tors_atoms, rng.integers(1, 5, len(tors_atoms)), rng.uniform(0, np.pi, len(tors_atoms)), rng.uniform(1, 30, len(tors_atoms)))
    ctx = bare_context(nm * L, 12.0)
    ctx.add_harmonic_bonds(*bonds); ctx.add_harmonic_angles(*angles); ctx.add_periodic_torsions(*torsions)
    ctx.set_positions(pos)
    e = ctx.compute(b.GROUP_BONDED)
    f = ctx.get_forces()
    e_ref = bonded_energy(pos, bonds, angles, torsions)
    assert abs(e - e_ref) / abs(e_ref) < 1e-6
    # per-kind energies add up (force groups of the plugin map to these bits)
    parts = [ctx.compute(g) for g in (b.GROUP_BONDS, b.GROUP_ANGLES, b.GROUP_TORSIONS)]
    assert abs(sum(parts) - e) / abs(e) < 1e-9
    ctx.compute(b.GROUP_BONDED)
    # forces against central differences of the restatement for the atoms of the first three chains
    h = 1e-6
    fd = np.zeros((3 * L, 3))
    for a in range(3 * L):
        for k in range(3):
            pp, pm = pos.copy(), pos.copy()
            pp[a, k] += h; pm[a, k] -= h
            fd[a, k] = -(bonded_energy(pp, bonds, angles, torsions) - bonded_energy(pm, bonds, angles, torsions)) / (2 * h)
    assert rel_rms(f[:3 * L], fd) < 1e-5
    assert np.abs(f.sum(0)).max() < 1e-4 * np.abs(f).max()
    ctx.close()


def test_add_forces_accumulates_into_the_fixed_point_buffer():
    # b200_add_forces: forces computed elsewhere (the plugin's second NonbondedForce) join the int64 buffer at its 2^-32 resolution
    ctx = bare_context(3)
    ctx.add_harmonic_bonds([[0, 1], [1, 2]], [1.5, 1.2], [0.8, 0.7])
    ctx.set_positions(np.array([[0, 2, 0], [0, 0, 0], [1, 0, 0]], dtype=float))
    ctx.compute(b.GROUP_BONDS, energy=False)
    f0 = ctx.get_forces().copy()
    extra = np.array([[1.25, -3.5, 0.125], [1e-3, 2e-3, -7e-3], [100.0, -200.0, 300.0]])
    ctx.add_forces(extra)
    assert np.abs(ctx.get_forces() - (f0 + extra)).max() < 1e-9
    ctx.add_forces(-extra)
    assert np.abs(ctx.get_forces() - f0).max() < 1e-9
    ctx.close()


def test_bonded_forces_of_one_kind_in_different_groups():
    # two HarmonicBondForce objects (and two torsion forces) created one after the other, as the plugin does for forces of the same kind
    # in different force groups; b200_bonded_set_active selects the ranges of the forces whose kernels ran in an evaluation
    rng = np.random.default_rng(9)
    n = 60
    pos = (rng.random((n, 3)) * 3 + 3).astype(np.float32).astype(np.float64)
    b1 = (np.stack([np.arange(0, 20), np.arange(1, 21)], 1), rng.uniform(0.5, 1.5, 20), rng.uniform(10, 50, 20))
    b2 = (np.stack([np.arange(30, 55), np.arange(31, 56)], 1), rng.uniform(0.5, 1.5, 25), rng.uniform(10, 50, 25))
    t1 = (np.stack([np.arange(0, 10), np.arange(1, 11), np.arange(2, 12), np.arange(3, 13)], 1), rng.integers(1, 4, 10), rng.uniform(0, 3, 10), rng.uniform(1, 9, 10))
    t2 = (np.stack([np.arange(20, 27), np.arange(21, 28), np.arange(22, 29), np.arange(23, 30)], 1), rng.integers(1, 4, 7), rng.uniform(0, 3, 7), rng.uniform(1, 9, 7))
    none = (np.zeros((0, 2), int), np.zeros(0), np.zeros(0)); nona = (np.zeros((0, 3), int), np.zeros(0), np.zeros(0)); nont = (np.zeros((0, 4), int), np.zeros(0, int), np.zeros(0), np.zeros(0))
    ctx = bare_context(n, 12.0)
    ctx.add_harmonic_bonds(*b1); ctx.add_harmonic_bonds(*b2); ctx.add_periodic_torsions(*t1); ctx.add_periodic_torsions(*t2)
    ctx.set_positions(pos)
    e_all = ctx.compute(b.GROUP_BONDS | b.GROUP_TORSIONS)
    assert abs(e_all - (bonded_energy(pos, b1, nona, t1) + bonded_energy(pos, b2, nona, t2))) / abs(e_all) < 1e-6
    ctx.bonded_set_active(0, [(0, 20)])
    e = ctx.compute(b.GROUP_BONDS); f = ctx.get_forces()
    assert abs(e - bonded_energy(pos, b1, nona, nont)) / abs(e) < 1e-6
    assert np.all(f[22:] == 0) and np.abs(f[:21]).max() > 0       # only the first force's atoms feel anything
    ctx.bonded_set_active(0, [(20, 25)])
    assert abs(ctx.compute(b.GROUP_BONDS) - bonded_energy(pos, b2, nona, nont)) / abs(e) < 1e-6
    ctx.bonded_set_active(2, [(10, 7)])
    e = ctx.compute(b.GROUP_BONDS | b.GROUP_TORSIONS)
    assert abs(e - bonded_energy(pos, b2, nona, t2)) / abs(e) < 1e-6
    ctx.bonded_set_active(0, [(0, 20), (20, 25)]); ctx.bonded_set_active(2)
    assert abs(ctx.compute(b.GROUP_BONDS | b.GROUP_TORSIONS) - e_all) / abs(e_all) < 1e-9
    with pytest.raises(b.B200Error, match="outside the terms"):
        ctx.bonded_set_active(0, [(40, 10)])
    ctx.close()


def test_exclusions_and_14_known_answers():
    # tests/TestNonbondedForce.h:98-179 testExclusionsAnd14 restated for the PME method: five particles in a row; 1-2 and 1-3 pairs are
    # excluded, 1-4 pairs carry scaled parameters.  Checked against the oracle (exceptions are plain Coulomb + LJ without cutoff,
    # nonbondedExceptions.cc:1-19) and, for the exception term alone, against its closed form.
    n = 5
    pos = np.zeros((n, 3)); pos[:, 0] = np.arange(n) * 0.31 + 1.0; pos[:, 1] = 1.0; pos[:, 2] = 1.0
    pos = pos.astype(np.float32).astype(np.float64)
    q = np.array([1.0, -0.5, 0.7, -1.0, 0.3]); sig = np.full(n, 0.25); eps = np.full(n, 0.8)
    excl, exc_pairs, exc_qq, exc_sig, exc_eps = [], [], [], [], []
    for i in range(n):
        for j in range(i + 1, n):
            if j - i <= 2:
                excl.append((i, j))
            elif j - i == 3:
                excl.append((i, j))
                exc_pairs.append((i, j)); exc_qq.append(q[i] * q[j] / 1.2); exc_sig.append(0.25); exc_eps.append(0.8 / 2.0)
    sysd = dict(n=n, positions=pos, box=np.array([4.0, 0, 0, 0, 4.0, 0, 0, 0, 4.0]), charge=q, sigma=sig, epsilon=eps, mass=np.ones(n),
                exclusions=np.array(excl, dtype=np.int32), exceptions=(np.array(exc_pairs, dtype=np.int32), np.array(exc_qq), np.array(exc_sig), np.array(exc_eps)))
    ctx = workloads.make_context(sysd, cutoff=1.0, alpha=3.0, grid=[40, 40, 40])
    ctx.set_positions(pos)
    e = ctx.compute(); f = ctx.get_forces()
    fo, eo = Oracle(sysd, 1.0, 3.0, [40, 40, 40]).nonbonded()
    assert rel_rms(f, fo) < 1e-5 and abs(e - eo) / abs(eo) < 1e-5
    # the same system without exception parameters differs by exactly the closed-form 1-4 terms
    plain = dict(sysd); plain.pop("exceptions")
    ctx2 = workloads.make_context(plain, cutoff=1.0, alpha=3.0, grid=[40, 40, 40])
    ctx2.set_positions(pos)
    e2 = ctx2.compute(); f2 = ctx2.get_forces()
    e14, f14 = 0.0, np.zeros((n, 3))
    for (i, j), qq, s, ep in zip(exc_pairs, exc_qq, exc_sig, exc_eps):
        d = pos[j] - pos[i]; r = np.linalg.norm(d); x = (s / r) ** 6
        e14 += 138.93545764438198 * qq / r + 4 * ep * x * (x - 1)
        dEdr = -138.93545764438198 * qq / r ** 2 - 4 * ep * (12 * x * x - 6 * x) / r
        f14[j] += -dEdr * d / r; f14[i] -= -dEdr * d / r
    assert abs((e - e2) - e14) < 1e-4 * max(1.0, abs(e14))
    assert np.abs((f - f2) - f14).max() < 1e-4 * np.abs(f14).max()
    ctx.close(); ctx2.close()


def test_water_box_with_exception_parameters_vs_oracle():
    # a 3000-atom water box in which 300 close intermolecular O...H pairs are made exceptions with their own charge product and LJ
    # parameters (so the molecule graph, the exclusion tiles, the erf correction and the 1-4 kernel all see them), stepped past a
    # re-sort so that the exception list is exercised in the permuted device order as well
    from scipy.spatial import cKDTree
    sysd = workloads.water_box(3000, seed=9)
    L = sysd["box"][0]
    pos = np.mod(sysd["positions"], L)
    pairs = cKDTree(pos, boxsize=L).query_pairs(0.3, output_type="ndarray")
    pairs = pairs[(pairs[:, 0] // 3 != pairs[:, 1] // 3) & (pairs[:, 0] % 3 == 0) & (pairs[:, 1] % 3 != 0)]
    # each molecule takes part in at most one exception, so that "molecules" stay small (two waters)
    used, chosen = set(), []
    for i, j in pairs:
        if i // 3 not in used and j // 3 not in used:
            used.add(i // 3); used.add(j // 3); chosen.append((i, j))
        if len(chosen) == 300:
            break
    chosen = np.array(chosen, dtype=np.int32)
    sysd["exclusions"] = np.vstack([sysd["exclusions"], chosen]).astype(np.int32)
    qq = 0.5 * sysd["charge"][chosen[:, 0]] * sysd["charge"][chosen[:, 1]]
    sysd["exceptions"] = (chosen, qq, np.full(len(chosen), 0.2), np.full(len(chosen), 0.3))
    ctx = workloads.make_context(sysd)
    ctx.set_positions(sysd["positions"])
    e = ctx.compute(); f = ctx.get_forces()
    fo, eo = Oracle(sysd, 0.9, *ctx.pme).nonbonded()
    assert rel_rms(f, fo) < 1e-5 and abs(e - eo) / abs(eo) < 1e-5
    ctx.set_velocities(sysd["velocities"])
    ctx.set_integrator(dt=0.001, seed=2)
    ctx.step(260)
    s2 = dict(sysd); s2["positions"] = ctx.get_positions()
    e = ctx.compute(); f = ctx.get_forces()
    fo, eo = Oracle(s2, 0.9, *ctx.pme).nonbonded()
    assert rel_rms(f, fo) < 1e-5 and abs(e - eo) / abs(eo) < 1e-5
    ctx.close()


def test_verlet_limit_conserves_energy():
    # friction 0, temperature 0: the LangevinMiddle update is the leapfrog/velocity Verlet scheme (IntegrateVerletStepKernel); the total
    # energy of rigid water at 1 fs must not drift.  Ewald tolerance 1e-5: with the default 5e-4 the pairs crossing the cutoff (where
    # erfc(alpha r)/r is 2e-4 of the bare Coulomb term) make the total energy random-walk by ~25 kJ/mol in 400 steps (measured), which
    # would hide an integrator fault of the same size
    sysd = workloads.water_box(3000, seed=12)
    ctx = workloads.make_context(sysd, tol=1e-5)
    ctx.set_positions(sysd["positions"]); ctx.set_velocities(sysd["velocities"])
    ctx.set_integrator(dt=0.002, temperature=300.0, friction=5.0, seed=1)
    ctx.step(200)            # equilibrate the synthetic start with a thermostat first
    ctx.set_integrator(dt=0.001, temperature=0.0, friction=0.0, seed=1, remove_cm_every=0)
    # the stored velocities lag the positions by half a step; the kinetic energy that belongs to the positions is that of velocities
    # shifted by dt/2 with the current forces (what IntegrateVerletStepKernel::computeKineticEnergy reports).  With the unshifted
    # velocities the sum fluctuates by ~20 kJ/mol (first order in dt; measured), which would hide a drift.
    tot, raw = [], []
    for _ in range(8):
        ctx.step(50)
        pe = ctx.compute()
        tot.append(pe + ctx.kinetic_energy_shifted(0.0005)); raw.append(pe + ctx.kinetic_energy())
    tot = np.array(tot)
    print("total energy, shifted KE:", tot - tot[0], " unshifted:", np.array(raw) - raw[0])
    assert np.abs(tot - tot[0]).max() < 1e-4 * abs(tot[0]), tot
    ctx.close()


def test_apply_constraints_and_velocity_constraints():
    sysd = workloads.water_box(3000, seed=3)
    ctx = workloads.make_context(sysd)
    rng = np.random.default_rng(0)
    ctx.set_positions(sysd["positions"] + rng.normal(scale=0.003, size=sysd["positions"].shape))
    vel = rng.normal(size=sysd["positions"].shape)
    ctx.set_velocities(vel)
    ctx.apply_constraints(1e-6)
    pos = ctx.get_positions()
    nw = sysd["num_waters"]
    o, h1, h2 = pos[0:3 * nw:3], pos[1:3 * nw:3], pos[2:3 * nw:3]
    for a, c, d in ((o, h1, workloads.D_OH), (o, h2, workloads.D_OH), (h1, h2, workloads.D_HH)):
        assert np.abs(np.linalg.norm(a - c, axis=1) - d).max() < 2e-6
    ctx.apply_velocity_constraints(1e-6)
    v = ctx.get_velocities()
    vo, v1, v2 = v[0:3 * nw:3], v[1:3 * nw:3], v[2:3 * nw:3]
    for (a, va), (c, vc) in (((o, vo), (h1, v1)), ((o, vo), (h2, v2)), ((h1, v1), (h2, v2))):
        rel = np.einsum("ij,ij->i", a - c, va - vc)            # d/dt |r_ab|^2 / 2 must vanish
        assert np.abs(rel).max() < 1e-9 * 1.0 + 1e-6
    # momentum of each molecule is conserved by the velocity constraints
    m = sysd["mass"][:3 * nw].reshape(-1, 3, 1)
    assert np.abs((m * vel[:3 * nw].reshape(-1, 3, 3)).sum(1) - (m * v[:3 * nw].reshape(-1, 3, 3)).sum(1)).max() < 1e-9
    ctx.close()


def _shake_system(nside=8, spacing=0.5):
    """nside^3 five-atom molecules X-C(H3): the carbon and its three hydrogens form a SHAKE cluster (three constraints to one centre, the
    HBonds constraint pattern of a protein), X is bonded to C by a harmonic bond; neutral molecules with small partial charges"""
    rng = np.random.default_rng(17)
    tet = np.array([[1, 1, 1], [1, -1, -1], [-1, 1, -1], [-1, -1, 1]], dtype=float) / np.sqrt(3.0)
    pos, charge, sigma, eps, mass, con_a, con_d, bonds, excl = [], [], [], [], [], [], [], [], []
    for ix in range(nside):
        for iy in range(nside):
            for iz in range(nside):
                base = len(pos)
                c = (np.array([ix, iy, iz]) + 0.5) * spacing + rng.normal(scale=0.01, size=3)
                q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
                dirs = tet @ q.T
                pos.append(c)
                for k in range(3):
                    pos.append(c + 0.109 * dirs[k])
                pos.append(c + 0.15 * dirs[3])
                charge += [-0.27, 0.09, 0.09, 0.09, 0.0]
                sigma += [0.34, 0.25, 0.25, 0.25, 0.33]; eps += [0.36, 0.066, 0.066, 0.066, 0.4]; mass += [12.011, 1.008, 1.008, 1.008, 35.45]
                for k in range(3):
                    con_a.append((base, base + 1 + k)); con_d.append(0.109)
                bonds.append((base, base + 4))
                for a in range(5):
                    for b2 in range(a + 1, 5):
                        excl.append((base + a, base + b2))
    n = len(pos); L = nside * spacing
    pos = np.array(pos).astype(np.float32).astype(np.float64)
    return dict(n=n, positions=pos, box=np.array([L, 0, 0, 0, L, 0, 0, 0, L]), charge=np.array(charge), sigma=np.array(sigma), epsilon=np.array(eps),
                mass=np.array(mass), exclusions=np.array(excl, dtype=np.int32), constraints=(np.array(con_a, dtype=np.int32), np.array(con_d)),
                bonds=(np.array(bonds, dtype=np.int32), np.full(len(bonds), 0.15), np.full(len(bonds), 2.0e5)))


def test_shake_clusters_hold_their_constraints_and_conserve_energy():
    sysd = _shake_system()
    ctx = workloads.make_context(sysd)
    ctx.add_harmonic_bonds(*sysd["bonds"])
    ctx.set_positions(sysd["positions"])
    rng = np.random.default_rng(1)
    ctx.set_velocities(rng.normal(size=sysd["positions"].shape) * np.sqrt(2.494 / sysd["mass"])[:, None])
    ctx.apply_velocity_constraints(1e-7)
    v = ctx.get_velocities(); p0 = ctx.get_positions()
    ca = sysd["constraints"][0]
    assert np.abs(np.einsum("ij,ij->i", p0[ca[:, 0]] - p0[ca[:, 1]], v[ca[:, 0]] - v[ca[:, 1]])).max() < 1e-5
    ctx.set_integrator(dt=0.002, temperature=300.0, friction=2.0, constraint_tol=1e-6, seed=4)
    ctx.step(300)
    p = ctx.get_positions()
    d = np.linalg.norm(p[ca[:, 0]] - p[ca[:, 1]], axis=1)
    assert np.abs(d - 0.109).max() < 2e-6 * 0.109 * 2, np.abs(d - 0.109).max()
    dof = 3 * sysd["n"] - len(ca) - 3
    T = 2 * ctx.kinetic_energy() / (dof * 0.00831446261815324)
    assert 240 < T < 360, T
    # forces after the dynamics (re-sorted device order, rebuilt lists) still match the oracle
    s2 = dict(sysd); s2["positions"] = ctx.get_positions(enforce_periodic=True)
    e = ctx.compute(b.GROUP_DIRECT | b.GROUP_RECIPROCAL); f = ctx.get_forces()
    fo, eo = Oracle(s2, 0.9, *ctx.pme).nonbonded()
    assert rel_rms(f, fo) < 1e-5 and abs(e - eo) / abs(eo) < 1e-5
    # Verlet limit: total energy does not drift
    ctx.set_integrator(dt=0.001, temperature=0.0, friction=0.0, constraint_tol=1e-7, seed=4, remove_cm_every=0)
    tot = []
    for _ in range(6):
        ctx.step(50)
        tot.append(ctx.compute() + ctx.kinetic_energy_shifted(0.0005))
    tot = np.array(tot)
    scale = abs(ctx.kinetic_energy())
    assert np.abs(tot - tot[0]).max() < 2e-3 * scale, tot
    # ApplyConstraintsKernel::apply on perturbed coordinates
    ctx.set_positions(ctx.get_positions() + rng.normal(scale=0.002, size=p.shape))
    ctx.apply_constraints(1e-7)
    p = ctx.get_positions()
    assert np.abs(np.linalg.norm(p[ca[:, 0]] - p[ca[:, 1]], axis=1) - 0.109).max() < 1e-6
    ctx.close()


def test_general_constraint_clusters_chains():
    # chains of six atoms with ALL bonds constrained (five constraints in a row: neither SETTLE nor a SHAKE star) and harmonic angles:
    # the general cluster solver (one CTA per chain, colour-ordered Gauss-Seidel sweeps; the reference uses CCMA here,
    # tests/TestVerletIntegrator.h testConstrainedChain) must hold every distance to the tolerance and conserve the energy
    rng = np.random.default_rng(8)
    nm, L = 30, 6
    pos = []
    for m in range(nm):
        p = rng.random(3) * 6 + 2
        chain = [p]
        for k in range(L - 1):
            d = rng.normal(size=3) + np.array([1.5, 0, 0]); d *= 0.15 / np.linalg.norm(d)
            chain.append(chain[-1] + d)
        pos += chain
    pos = np.array(pos)
    idx = np.arange(nm * L).reshape(nm, L)
    con = np.concatenate([np.stack([idx[:, k], idx[:, k + 1]], 1) for k in range(L - 1)]).astype(np.int32)
    angle_atoms = np.concatenate([np.stack([idx[:, k], idx[:, k + 1], idx[:, k + 2]], 1) for k in range(L - 2)])
    masses = rng.uniform(1.0, 16.0, nm * L)
    ctx = b.Context(masses, BOX, constraints=(con, np.full(len(con), 0.15)))
    ctx.add_harmonic_angles(angle_atoms, np.full(len(angle_atoms), 1.9), np.full(len(angle_atoms), 400.0))
    ctx.set_positions(pos)
    ctx.set_velocities(rng.normal(size=pos.shape) * np.sqrt(2.494 / masses)[:, None])
    ctx.apply_constraints(1e-7); ctx.apply_velocity_constraints(1e-7)
    ctx.set_integrator(dt=0.001, temperature=0.0, friction=0.0, constraint_tol=1e-7, seed=1, remove_cm_every=0)
    tot = []
    for _ in range(6):
        ctx.step(50)
        tot.append(ctx.compute(b.GROUP_ANGLES) + ctx.kinetic_energy_shifted(0.0005))
    p = ctx.get_positions()
    d = np.linalg.norm(p[con[:, 0]] - p[con[:, 1]], axis=1)
    assert np.abs(d - 0.15).max() < 1e-6
    tot = np.array(tot)
    print("general clusters: total energy", tot)
    assert np.abs(tot - tot[0]).max() < 2e-3 * np.abs(tot).max()
    ctx.close()
