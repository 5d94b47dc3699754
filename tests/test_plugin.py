"""The OpenMM `B200` Platform plugin (openmm_b200/lib/libOpenMMB200.so) driven through the reference's own public API by
oracle/_ref/ref_driver (unmodified OpenMM 8.6): an unmodified System/Context selects the platform by name."""
import ctypes
import json
import os
import subprocess
import tempfile

import numpy as np
import pytest

import openmm_b200 as b
from openmm_b200 import workloads
from oracle_lib import REF_DRIVER, REF_PLUGINS


def test_plugin_exports_the_loader_symbols():
    assert os.path.exists(b.PLUGIN_PATH), "run python -m openmm_b200.build"
    lib = ctypes.CDLL(b.PLUGIN_PATH)
    assert hasattr(lib, "registerPlatforms") and hasattr(lib, "registerKernelFactories")   # PluginInitializer.h:46,56


def _driver(args):
    out = subprocess.run([REF_DRIVER] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert out.returncode == 0, out.stderr
    return json.loads(out.stdout.strip().splitlines()[-1])


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1536, 6000])
def test_context_on_b200_platform_matches_reference_platform(n):
    sysd = workloads.water_box(n, seed=n)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    with tempfile.TemporaryDirectory() as td:
        sp = os.path.join(td, "s.bin")
        workloads.write_system_file(sp, sysd, 0.9, alpha, grid, dispersion=True)
        r = _driver(["compare", sp, "--platform", "B200", "--plugin", b.PLUGIN_PATH, "--against", "Reference"])
    print(r)
    assert r["rel_rms_force_total"] < 1e-5 and r["rel_rms_force_direct"] < 1e-5 and r["rel_rms_force_recip"] < 2e-5
    assert r["rel_energy"] < 1e-5
    assert abs(r["e_direct"] - r["e_direct_ref"]) / abs(r["e_direct_ref"]) < 1e-5
    assert abs(r["e_recip"] - r["e_recip_ref"]) / abs(r["e_recip_ref"]) < 1e-5


@pytest.mark.gpu
def test_ljpme_context_on_b200_platform_matches_reference_platform():
    # NonbondedForce::LJPME with explicit setLJPMEParameters, evaluated by both platforms inside one process
    sysd = workloads.ljpme_water_box(3002)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    lj_alpha, lj_grid = b.pme_parameters(0.9, sysd["box"], lj=True)
    with tempfile.TemporaryDirectory() as td:
        sp = os.path.join(td, "s.bin")
        workloads.write_system_file(sp, sysd, 0.9, alpha, grid, method=5, lj_alpha=lj_alpha, lj_grid=lj_grid)
        r = _driver(["compare", sp, "--platform", "B200", "--plugin", b.PLUGIN_PATH, "--against", "Reference"])
    print(r)
    assert r["rel_rms_force_total"] < 1e-5 and r["rel_rms_force_direct"] < 1e-5 and r["rel_rms_force_recip"] < 2e-5
    assert r["rel_energy"] < 1e-5


@pytest.mark.gpu
def test_switching_function_context_on_b200_platform_matches_reference_platform():
    # setUseSwitchingFunction + dispersion correction (calcDispersionCorrection integrates the switched tail on the host)
    sysd = workloads.water_box(3000, seed=5)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    with tempfile.TemporaryDirectory() as td:
        sp = os.path.join(td, "s.bin")
        workloads.write_system_file(sp, sysd, 0.9, alpha, grid, dispersion=True, switch_distance=0.75)
        r = _driver(["compare", sp, "--platform", "B200", "--plugin", b.PLUGIN_PATH, "--against", "Reference"])
    print(r)
    assert r["rel_rms_force_total"] < 1e-5 and r["rel_rms_force_direct"] < 1e-5
    assert r["rel_energy"] < 1e-5


@pytest.mark.gpu
def test_langevin_middle_steps_through_the_plugin():
    sysd = workloads.water_box(6000, seed=6000)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    with tempfile.TemporaryDirectory() as td:
        sp = os.path.join(td, "s.bin")
        workloads.write_system_file(sp, sysd, 0.9, alpha, grid)
        g = _driver(["bench", sp, "--platform", "B200", "--plugin", b.PLUGIN_PATH, "--steps", "200", "--warmup", "50"])
        c = _driver(["bench", sp, "--platform", "CPU", "--plugin", os.path.join(REF_PLUGINS, "libOpenMMCPU.so"), "--plugin",
                     os.path.join(REF_PLUGINS, "libOpenMMPME.so"), "--steps", "200", "--warmup", "50"])
    print(g, c)
    assert np.isfinite(g["potential_energy"])
    # same ensemble after 250 steps: potential energies of the two platforms agree to within thermal fluctuations (~1%)
    assert abs(g["potential_energy"] - c["potential_energy"]) / abs(c["potential_energy"]) < 0.03


@pytest.mark.gpu
def test_checkpoint_round_trip_through_the_plugin():
    # Context::createCheckpoint / loadCheckpoint on the B200 platform: the continuation from a loaded checkpoint repeats the original
    # continuation (positions, velocities, time and the step counter that keys the integrator's noise are all restored).  The
    # checkpoint holds unwrapped double-precision coordinates, so the restored float positions and the first list build differ
    # from the original run in the last bit: the two continuations agree to rounding, not bit for bit.
    sysd = workloads.water_box(3000, seed=8)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    with tempfile.TemporaryDirectory() as td:
        sp = os.path.join(td, "s.bin")
        workloads.write_system_file(sp, sysd, 0.9, alpha, grid)
        r = _driver(["checkpoint", sp, "--platform", "B200", "--plugin", b.PLUGIN_PATH, "--steps", "25"])
    print(r)
    assert r["time_restored"] and r["final_time_equal"]
    assert r["max_position_difference"] < 1e-5 and r["max_velocity_difference"] < 1e-3


@pytest.mark.gpu
def test_monte_carlo_barostat_through_the_plugin():
    # MonteCarloBarostat (ApplyMonteCarloBarostatKernel: save / scale molecule centres / restore) with a Verlet integrator on the B200
    # platform and on the reference's Reference platform: same barostat seed, so the accepted and rejected volume moves must follow
    # the same sequence as long as the energies agree (tests/TestMonteCarloBarostat.h checks the ensemble; this checks the mechanics)
    sysd = workloads.water_box(1536, seed=3)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    with tempfile.TemporaryDirectory() as td:
        sp = os.path.join(td, "s.bin")
        workloads.write_system_file(sp, sysd, 0.9, alpha, grid)
        g = _driver(["npt", sp, "--platform", "B200", "--plugin", b.PLUGIN_PATH, "--steps", "24", "--frequency", "2"])
        r = _driver(["npt", sp, "--platform", "Reference", "--steps", "24", "--frequency", "2"])
    print(g["volumes"], r["volumes"])
    v0 = float(np.prod([sysd["box"][0], sysd["box"][4], sysd["box"][8]]))
    assert any(abs(v - v0) / v0 > 1e-6 for v in r["volumes"]), "no volume move was accepted: the test would be vacuous"
    assert np.allclose(g["volumes"], r["volumes"], rtol=1e-7)
    assert abs(g["potential_energy"] - r["potential_energy"]) / abs(r["potential_energy"]) < 1e-4
    assert np.allclose(g["position_sum"], r["position_sum"], atol=2e-2)


@pytest.mark.gpu
def test_several_gpus_in_one_context_through_the_plugin():
    # DeviceIndex=0,1 (platforms/cuda/src/CudaPlatform.cpp:116-141, CudaParallelKernels.cpp): one OpenMM Context drives two B200s from one
    # process.  Forces and energies against the Reference platform, then dynamics against the one-GPU run of the same plugin (the
    # integer force sums make the trajectories identical, so the potential energies after 100 steps agree to the last digit printed).
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    sysd = workloads.water_box(23558, seed=23558)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    with tempfile.TemporaryDirectory() as td:
        sp = os.path.join(td, "s.bin")
        workloads.write_system_file(sp, sysd, 0.9, alpha, grid, dispersion=True)
        r = _driver(["compare", sp, "--platform", "B200", "--plugin", b.PLUGIN_PATH, "--against", "Reference", "--prop", "DeviceIndex=0,1"])
        g2 = _driver(["bench", sp, "--platform", "B200", "--plugin", b.PLUGIN_PATH, "--steps", "60", "--warmup", "40", "--prop", "DeviceIndex=0,1"])
        g1 = _driver(["bench", sp, "--platform", "B200", "--plugin", b.PLUGIN_PATH, "--steps", "60", "--warmup", "40"])
    print(r, g2, g1)
    assert r["rel_rms_force_total"] < 1e-5 and r["rel_energy"] < 1e-5
    assert g2["potential_energy"] == g1["potential_energy"]
