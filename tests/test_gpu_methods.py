"""GPU parity of the nonbonded methods outside the PME family: NoCutoff, CutoffNonPeriodic and CutoffPeriodic
(reaction field; coulombLennardJones.cc:77-116) and NonbondedForce::Ewald (explicit k-space sum, ewald.cc), against the
Reference-platform fixtures made by tests/golden/make_fixtures.py and against the oracle on other seeds; plus the switching
function together with LJPME."""
import os

import numpy as np
import pytest

import openmm_b200 as b
from openmm_b200 import workloads
from oracle_lib import Oracle, ewald_parameters, rel_rms

pytestmark = pytest.mark.gpu
FORCE_TOL, ENERGY_TOL = 1e-5, 1e-5


@pytest.mark.parametrize("method,name", [(b.NB_NOCUTOFF, "nocutoff"), (b.NB_CUTOFF_NONPERIODIC, "cutoff_nonperiodic"), (b.NB_CUTOFF_PERIODIC, "cutoff_periodic")])
def test_cutoff_methods_vs_reference_fixture(golden_dir, method, name):
    g = np.load(os.path.join(golden_dir, "ref_cutoff_methods_768.npz"))
    sysd = workloads.water_box(768, seed=768)
    assert np.array_equal(sysd["positions"].astype(np.float32), g["positions"])
    ctx = workloads.make_context(sysd, method=method)
    ctx.set_positions(sysd["positions"])
    e = ctx.compute(); f = ctx.get_forces()
    print(name, rel_rms(f, g["f_" + name]), e, float(g["e_" + name]))
    assert rel_rms(f, g["f_" + name]) < FORCE_TOL
    assert abs(e - float(g["e_" + name])) / abs(float(g["e_" + name])) < ENERGY_TOL
    # energy-only and force-only evaluations agree with the combined one
    assert abs(ctx.compute(forces=False) - e) / abs(e) < 1e-7
    ctx.compute(energy=False)
    assert rel_rms(ctx.get_forces(), f) < 1e-7
    ctx.close()


@pytest.mark.parametrize("method", [b.NB_CUTOFF_NONPERIODIC, b.NB_CUTOFF_PERIODIC])
def test_cutoff_methods_larger_box_vs_oracle(method):
    # 6000 atoms: the periodic case takes the single-periodic-copy tile path, the non-periodic one runs without any image
    sysd = workloads.water_box(6000, seed=31)
    ctx = workloads.make_context(sysd, method=method, reaction_field_dielectric=60.0)
    ctx.set_positions(sysd["positions"])
    e = ctx.compute(); f = ctx.get_forces()
    fo, eo = Oracle(sysd, 0.9, 1.0, [8, 8, 8], method=method, rf_dielectric=60.0).nonbonded()
    assert rel_rms(f, fo) < FORCE_TOL and abs(e - eo) / abs(eo) < ENERGY_TOL
    # dynamics keep the list valid (rebuilds, re-sort): forces still match on the new coordinates
    ctx.set_velocities(sysd["velocities"]); ctx.set_integrator(dt=0.002, seed=5)
    ctx.step(260)
    # identical coordinates on both sides: the float positions the mixed-precision kernels compute with.  Without periodic
    # wrapping the droplet spreads out and the float remainder of a coordinate (up to 2.4e-7 nm at 4 nm) alone is worth 1.4e-5 in the forces.
    s2 = dict(sysd); s2["positions"] = ctx.get_positions().astype(np.float32).astype(np.float64)
    e = ctx.compute(); f = ctx.get_forces()
    fo, eo = Oracle(s2, 0.9, 1.0, [8, 8, 8], method=method, rf_dielectric=60.0).nonbonded()
    assert rel_rms(f, fo) < FORCE_TOL and abs(e - eo) / abs(eo) < ENERGY_TOL
    ctx.close()


def test_ewald_vs_reference_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_ewald_768.npz"))
    sysd = workloads.water_box(768, seed=768)
    alpha, kmax = ewald_parameters(0.9, sysd["box"])
    ctx = workloads.make_context(sysd, method=b.NB_EWALD, alpha=alpha, ewald_kmax=kmax)
    ctx.set_positions(sysd["positions"])
    e_dir = ctx.compute(b.GROUP_DIRECT); f_dir = ctx.get_forces()
    e_rec = ctx.compute(b.GROUP_RECIPROCAL); f_rec = ctx.get_forces()
    e_tot = ctx.compute(); f_tot = ctx.get_forces()
    print("ewald", rel_rms(f_dir, g["f_direct"]), rel_rms(f_rec, g["f_recip"]))
    assert rel_rms(f_dir, g["f_direct"]) < FORCE_TOL and rel_rms(f_rec, g["f_recip"]) < FORCE_TOL
    assert rel_rms(f_tot, g["f_direct"] + g["f_recip"]) < FORCE_TOL
    assert abs(e_dir - float(g["e_direct"])) / abs(float(g["e_direct"])) < ENERGY_TOL
    assert abs(e_rec - float(g["e_recip"])) / abs(float(g["e_recip"])) < ENERGY_TOL
    assert abs(e_tot - float(g["e_direct"]) - float(g["e_recip"])) / abs(float(g["e_direct"]) + float(g["e_recip"])) < ENERGY_TOL
    ctx.close()


def test_switching_function_with_ljpme_vs_oracle():
    sysd = workloads.ljpme_water_box(3002)
    ctx = workloads.make_context(sysd, ljpme=True, switch_distance=0.8)
    ctx.set_positions(sysd["positions"])
    e = ctx.compute(); f = ctx.get_forces()
    fo, eo = Oracle(sysd, 0.9, ctx.pme[0], ctx.pme[1], ctx.lj_pme[0], ctx.lj_pme[1], switch_distance=0.8).nonbonded()
    assert rel_rms(f, fo) < FORCE_TOL and abs(e - eo) / abs(eo) < ENERGY_TOL
    ctx.close()
