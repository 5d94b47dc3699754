#!/usr/bin/env python
"""bench.py -- ns/day of the NonbondedForce PME path (PME, 2 fs, mixed precision, LangevinMiddle, rigid TIP3P water).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload stmv|cellulose|apoa1|dhfr] [--impl reference]

One MD "step" = neighbour-list check/rebuild + direct-space tile kernel + PME (spread, 3D FFT, convolution, inverse FFT,
interpolation) + exclusion correction + LangevinMiddle/SETTLE + COM-motion removal.  Default workload: the STMV-sized box
(1,066,628 atoms, the configuration BASELINE.json's multi-GPU metric is quoted on; it fits one GPU).
`value`   : whole-job ns/day with the state resident in HBM (CUDA events on the context's stream, max over ranks).
`e2e`     : ns/day through the C ABI with HOST buffers: every step uploads the positions from host memory
            (b200_set_positions), runs one step, and downloads the positions (b200_get_positions).
`roofline`: the dominant kernel (direct-space tile kernel) against the FP32 pipe peak; PME kernels against measured HBM.
`cpu_baseline` / `--impl reference`: the reference's own CPU platform (oracle/_ref, all host cores) on the same system.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

FLOP_PER_PAIR = 56.0          # force-only direct-space work per in-cutoff pair
DT_PS = 0.002


def ns_per_day(ms_per_step):
    return DT_PS * 1e-3 * 86400.0 / (ms_per_step * 1e-3)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("sm_max_mhz", 1965.0), "measured"
    return 6650.0, 1965.0, "fallback"


class ClockSampler:
    def __init__(self, device):
        self.path = os.path.join(tempfile.gettempdir(), f"b200_clocks_{os.getpid()}.csv")
        self.proc = None
        self.device = device

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device),
                                          "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
                                          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
                                          "clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(np.max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def measured_traffic(workload):
    """DRAM bytes per launch of the tile kernel (dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture), read
    from the profile summary the capture was extracted into; None when no capture of this workload is committed."""
    p = os.path.join(ROOT, "profiles", "tile_kernel_dram_traffic.json")
    try:
        return json.load(open(p)).get(workload, {})
    except Exception:
        return {}


def build_system(workload):
    """the synthetic TIP3P box of a BASELINE.json configuration; "apoa1ljpme" = the ApoA1-size box with NonbondedForce::LJPME (second, 45^3 mesh)"""
    from openmm_b200 import workloads
    if workload == "apoa1ljpme":
        sysd = workloads.cached_water_box(workloads.WORKLOADS["apoa1"], ljpme=True)
        sysd["ljpme"] = True
        return sysd
    return workloads.cached_water_box(workloads.WORKLOADS[workload])


def sysfile_kw(sysd):
    """extra arguments of write_system_file for this system (the LJPME method with its dispersion mesh)"""
    if not sysd.get("ljpme"):
        return {}
    import openmm_b200 as b
    lj_alpha, lj_grid = b.pme_parameters(0.9, sysd["box"], lj=True)
    return dict(method=5, lj_alpha=lj_alpha, lj_grid=lj_grid)


def workload_string(name, n, alpha, grid):
    """one description of the system, shared by both arms (the driver compares the two `config.workload` strings)"""
    return (f"{name}: {n} atoms TIP3P water box at 100 atoms/nm^3, {'LJPME' if name.endswith('ljpme') else 'PME'} cutoff 0.9 nm, alpha {alpha:.5f}, grid {list(grid)}, "
            "LangevinMiddle 300 K 1/ps 2 fs, rigid water, CMMotionRemover")


def run_ref_driver(sysd, alpha, grid, steps, warmup, platform="CPU", threads=0, e2e=False, system_file=None, extra_plugins=(), props=()):
    """Runs oracle/_ref/ref_driver (the UNMODIFIED reference built from /root/reference, driven through its public API):
    platform "CPU" times the reference's own CPU platform, platform "B200" times this repository's plugin inside the
    reference's Context/Integrator machinery (the drop-in path an OpenMM user takes), "CUDA" the reference's CUDA platform."""
    from openmm_b200 import workloads
    from oracle_lib import REF_DRIVER, REF_PLUGINS
    if not os.path.exists(REF_DRIVER):
        return None
    with tempfile.TemporaryDirectory() as td:
        sp = system_file
        if sp is None:
            sp = os.path.join(td, "system.bin")
            workloads.write_system_file(sp, sysd, 0.9, alpha, grid, **sysfile_kw(sysd))
        cmd = [REF_DRIVER, "bench", sp, "--platform", platform, "--steps", str(steps), "--warmup", str(warmup), "--dt", str(DT_PS)]
        if platform == "CPU":
            cmd += ["--plugin", os.path.join(REF_PLUGINS, "libOpenMMCPU.so"), "--plugin", os.path.join(REF_PLUGINS, "libOpenMMPME.so")]
        for pl in extra_plugins:
            cmd += ["--plugin", pl]
        for k, v in props:
            cmd += ["--prop", f"{k}={v}"]
        if threads:
            cmd += ["--threads", str(threads)]
        if e2e:
            cmd += ["--e2e"]
        out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        if out.returncode != 0:
            sys.stderr.write(out.stderr[-2000:])
            return None
        return json.loads(out.stdout.strip().splitlines()[-1])


def run_reference_cpu(sysd, alpha, grid, steps, warmup, threads=0, e2e=False):
    return run_ref_driver(sysd, alpha, grid, steps, warmup, "CPU", threads, e2e)


def cpu_steps_for(n_atoms, budget_s=20.0):
    est_ms = 1.3e-3 * n_atoms        # ~1.2 s/step for 1M atoms on 8 cores; rescaled below
    cores = os.cpu_count() or 8
    est_ms *= 8.0 / max(cores, 8) * 1.5
    return max(3, min(200, int(budget_s * 1e3 / est_ms)))


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import openmm_b200 as b
    sysd = build_system(args.workload)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    n = sysd["n"]
    steps = min(args.steps, cpu_steps_for(n, 60.0))
    warm = min(args.warmup, max(2, steps // 3))
    r = run_reference_cpu(sysd, alpha, grid, steps, warm)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_driver missing or failed (reference not built)"}))
        return 0
    line = {
        "impl": "reference", "metric": "ns/day (PME, 2 fs, mixed precision)", "value": r["ns_per_day"], "unit": "ns/day",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32 pair math / f64 integration (reference CPU platform)", "data": "synthetic",
        "config": {"workload": workload_string(args.workload, n, alpha, grid), "atoms": n,
                   "platform": "reference CPU platform + CPU-PME plugin, unmodified OpenMM 8.6 (oracle/_ref)"},
        "cpu_baseline": {"value": r["ns_per_day"], "unit": "ns/day", "cores": int(r["threads"]), "kind": "reference",
                         "sample": f"{r['steps']} MD steps after {r['warmup']} warm-up steps of the full {n}-atom system"},
        "e2e": {"value": r["ns_per_day"], "unit": "ns/day", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


def multi_rank_identity_check(sysd, rank, world, local, dist, torch, b, workloads, steps=12):
    def make(comm):
        c = workloads.make_context(sysd, device=local, ljpme=bool(sysd.get("ljpme")))
        if comm:
            ids = [b.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            c.comm_init(rank, world, ids[0])
        c.set_positions(sysd["positions"]); c.set_velocities(sysd["velocities"])
        c.set_integrator(dt=DT_PS, temperature=300.0, friction=1.0, constraint_tol=1e-5, seed=1, remove_cm_every=1)
        return c
    ok = True
    try:
        single, multi = make(False), make(True)
        e1 = single.compute(); f1 = single.get_forces().copy()
        e2 = multi.compute(); f2 = multi.get_forces()
        ok = bool(e1 == e2 and np.array_equal(f1, f2))
        single.step(steps); multi.step(steps)
        ok = ok and bool(np.array_equal(single.get_positions(), multi.get_positions()))
        single.close(); multi.close()
    except Exception as ex:   # a failed check is reported, not hidden
        sys.stderr.write(f"[rank {rank}] identity check failed with {ex!r}\n")
        ok = False
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


def plugin_records(sysd, alpha, grid, args, n, devices=None):
    """The same system through an unmodified OpenMM Context + LangevinMiddleIntegrator + CMMotionRemover (oracle/_ref/ref_driver):
    `plugin` = this repository's libOpenMMB200.so selected by name ("B200"), `reference_gpu` = the reference's own CUDA platform
    when oracle/_ref holds a build of it.  Wall-clock around integ.step(K) + one getState, as examples/benchmarks/benchmark.py:99-108."""
    from openmm_b200 import workloads, PLUGIN_PATH
    from oracle_lib import REF_DRIVER, REF_PLUGINS
    if not os.path.exists(REF_DRIVER):
        return None, None
    plugin = ref_gpu = None
    with tempfile.TemporaryDirectory() as td:
        sp = os.path.join(td, "system.bin")
        workloads.write_system_file(sp, sysd, 0.9, alpha, grid, **sysfile_kw(sysd))
        k, w = max(args.steps, 100), max(args.warmup, 20)
        dev = [] if not devices else [("DeviceIndex", ",".join(str(x) for x in devices))]
        if os.path.exists(PLUGIN_PATH):
            r = run_ref_driver(sysd, alpha, grid, k, w, "B200", system_file=sp, extra_plugins=[PLUGIN_PATH], props=dev)
            r2 = run_ref_driver(sysd, alpha, grid, max(5, min(args.e2e_steps, 20)), 3, "B200", e2e=True, system_file=sp, extra_plugins=[PLUGIN_PATH], props=dev)
            if r is not None:
                plugin = {"value": r["ns_per_day"], "unit": "ns/day", "ms_per_step": r["ms_per_step"], "steps": r["steps"],
                          "api": "OpenMM::Context(system, LangevinMiddleIntegrator, Platform::getPlatformByName(\"B200\")" +
                                 (", {DeviceIndex=" + dev[0][1] + "}" if dev else "") + "); integrator.step(K)",
                          "e2e": None if r2 is None else {"value": r2["ns_per_day"], "unit": "ns/day", "ms_per_step": r2["ms_per_step"],
                                                           "h2d_bytes_per_step": 24 * n, "d2h_bytes_per_step": 24 * n,
                                                           "api": "per step: Context::setPositions(vector<Vec3>) + step(1) + getState(Positions)"}}
        cuda_lib = os.path.join(REF_PLUGINS, "libOpenMMCUDA.so")
        if os.path.exists(cuda_lib):
            r = run_ref_driver(sysd, alpha, grid, k, w, "CUDA", system_file=sp, extra_plugins=[cuda_lib], props=[("Precision", "mixed")] + dev)
            if r is not None:
                ref_gpu = {"value": r["ns_per_day"], "unit": "ns/day", "ms_per_step": r["ms_per_step"], "steps": r["steps"],
                           "what": "the reference's own CUDA platform (unmodified OpenMM 8.6, Precision=mixed) on this GPU, same system file"}
    return plugin, ref_gpu


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=30)
    ap.add_argument("--workload", default="stmv", choices=["dhfr", "apoa1", "apoa1ljpme", "cellulose", "stmv"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--no-plugin", action="store_true", help="skip the runs through libOpenMMB200.so / the reference's CUDA platform")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    import torch
    import openmm_b200 as b
    from openmm_b200 import workloads
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sysd = build_system(args.workload)
    n = sysd["n"]
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    ctx = workloads.make_context(sysd, device=local, ljpme=bool(sysd.get("ljpme")))
    if world > 1:
        ids = [b.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(rank, world, ids[0])
    ctx.set_positions(sysd["positions"])
    ctx.set_velocities(sysd["velocities"])
    ctx.set_integrator(dt=DT_PS, temperature=300.0, friction=1.0, constraint_tol=1e-5, seed=1, remove_cm_every=1)
    ctx.compute(forces=True, energy=False)      # first list build + capacity sizing outside the timed region

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up, then the timed region
    ctx.step(args.warmup)
    ctx.synchronize()
    launches0 = ctx.launch_count()
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    total_ms = ctx.timed_steps(args.steps)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.launch_count() - launches0
    if dist is not None:
        t = torch.tensor([total_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps

    # ---- per-kernel-family timing (CUDA events on the launching streams) over a second window of the same steps
    ctx.timing_enable(True); ctx.timing_reset()
    prof_steps = min(args.steps, 50)
    ctx.step(prof_steps)
    fam = ctx.timing_read()
    ctx.timing_enable(False)
    if os.environ.get("B200_BENCH_VERBOSE"):
        singles = [ctx.timed_steps(1) for _ in range(24)]      # step-by-step times: shows rebuild steps and any one-off stall
        sys.stderr.write(f"[rank {rank}] single steps (ms): " + " ".join(f"{x:.2f}" for x in singles) + "\n")
        sys.stderr.write(f"[rank {rank}] ms/step {ms_per_step:.4f} | " + " ".join(f"{k}={v[0] / prof_steps * 1e3:.0f}us" for k, v in fam.items() if v[0] > 0) + "\n")
        if dist is not None:   # raw NCCL all-reduce of a force-sized int64 buffer, for comparison with the library's own collective
            buf = torch.zeros(3 * (((n + 31) // 32) * 32), dtype=torch.int64, device="cuda")
            for _ in range(5):
                dist.all_reduce(buf)
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                dist.all_reduce(buf)
            e1.record(); torch.cuda.synchronize()
            sys.stderr.write(f"[rank {rank}] torch all_reduce of {buf.numel() * 8 / 1e6:.1f} MB: {e0.elapsed_time(e1) / 20 * 1e3:.0f} us\n")
    stats = ctx.nblist_stats()
    pairs, _ = ctx.direct_pairs()
    pairs *= 1 if world == 1 else 1   # each rank counts its own rows
    if dist is not None:
        t = torch.tensor([pairs], device="cuda", dtype=torch.int64)
        dist.all_reduce(t)
        pairs_total = int(t.item())
    else:
        pairs_total = pairs
    sparse_tiles = int(stats["num_sparse_tiles"])
    dense_tiles = int(stats["num_tiles"]) - 2 * sparse_tiles
    dense_slots = (dense_tiles + int(stats["num_exclusion_tiles"])) * 1024
    tile_ms, tile_n = fam["tile"]
    tile_avg_ms = tile_ms / max(tile_n, 1)
    hbm_gbs, sm_max_mhz, peak_kind = measured_peaks()
    fp32_peak = ctx_sm = None
    sm_count = torch.cuda.get_device_properties(local).multi_processor_count
    fp32_peak = sm_count * 128 * 2 * sm_max_mhz * 1e6 / 1e12
    achieved_tflops = FLOP_PER_PAIR * pairs / (tile_avg_ms * 1e-3) / 1e12 if tile_avg_ms > 0 else 0.0
    G = grid[0] * grid[1] * grid[2]
    pme_bytes = 36.0 * G + 72.0 * n
    if sysd.get("ljpme"):      # LJPME adds the same traffic for its own (smaller) mesh, SURVEY.md 8d
        lg = sysfile_kw(sysd)["lj_grid"]
        pme_bytes += 36.0 * lg[0] * lg[1] * lg[2] + 72.0 * n
    pme_ms = sum(fam[k][0] for k in ("spread", "fft", "convolve", "interpolate")) / max(prof_steps, 1)
    tile_traffic = measured_traffic(args.workload)
    roofline = {"kernel": "tileKernel<forces>", "bound": "fp32", "achieved": achieved_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
                "frac": achieved_tflops / fp32_peak if fp32_peak else None,
                "traffic": tile_traffic.get("bytes") if world == 1 else None, "traffic_source": tile_traffic.get("source") if world == 1 else None,
                "peak_source": f"{sm_count} SMs x 128 lanes x 2 x {sm_max_mhz:.0f} MHz ({peak_kind} sm_max_mhz)",
                "algorithmic": f"{FLOP_PER_PAIR:.0f} FLOP x {pairs} in-cutoff pairs per launch (this rank)",
                "executed_pair_slots_per_launch": dense_slots,
                "avg_launch_ms": tile_avg_ms,
                # second roofline, flattened so that a parser that keeps only scalar keys still sees it: all PME kernels against measured HBM
                "pme_bound": "hbm", "pme_achieved": pme_bytes / (pme_ms * 1e-3) / 1e9 if pme_ms > 0 else None, "pme_peak": hbm_gbs, "pme_unit": "GB/s",
                "pme_frac": (pme_bytes / (pme_ms * 1e-3) / 1e9 / hbm_gbs) if pme_ms > 0 else None, "pme_bytes_per_step": pme_bytes, "pme_ms_per_step": pme_ms}

    # ---- end to end through the C ABI with host buffers
    e2e_steps = max(1, min(args.e2e_steps, args.steps))

    def e2e_loop(pos, k):
        # multi-rank: ONE rank owns the host buffers (as one OpenMM Context would); it uploads once over PCIe and the positions reach
        # the other GPUs over NVLink (b200_set_positions_root), the download happens on that rank only
        root_io = world > 1
        ctx.get_positions(enforce_periodic=True, out=pos)
        barrier()
        t0 = time.perf_counter()
        for _ in range(k):
            if root_io:
                ctx.set_positions_root(pos if rank == 0 else None)
            else:
                ctx.set_positions(pos)
            ctx.step(1)
            if root_io and rank != 0:
                ctx.join_get_positions()
            else:
                ctx.get_positions(enforce_periodic=True, out=pos)
        barrier()
        return (time.perf_counter() - t0) * 1e3 / k

    pinned = torch.empty((n, 3), dtype=torch.float64, pin_memory=True)      # the caller's host buffers: page-locked (contract) ...
    e2e_ms = e2e_loop(pinned.numpy(), e2e_steps)
    e2e_pageable_ms = e2e_loop(np.empty((n, 3)), max(1, e2e_steps // 4))      # ... and, for reference, ordinary pageable memory
    if dist is not None:
        t = torch.tensor([e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    npad = ((n + 31) // 32) * 32
    e2e = {"value": ns_per_day(e2e_ms), "unit": "ns/day", "ms_per_step": e2e_ms, "h2d_bytes_per_step": 24 * n, "d2h_bytes_per_step": 24 * n,
           "ms_per_step_pageable_buffers": e2e_pageable_ms,
           "api": ("b200_set_positions(host doubles, page-locked caller buffer) + b200_step(1) + b200_get_positions(host doubles); permutation and "
                   "precision split on the GPU; pageable caller buffers go through a chunk-pipelined pinned staging area") if world == 1 else
                  ("rank 0: b200_set_positions_root(host doubles) [PCIe once, NVLink to the other ranks] + b200_step(1) + b200_get_positions(host doubles); "
                   "other ranks take part in the collectives without host buffers")}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        k = cpu_steps_for(n, 20.0)
        r = run_reference_cpu(sysd, alpha, grid, k, max(2, k // 3))
        if r is not None:
            cpu = {"value": r["ns_per_day"], "unit": "ns/day", "cores": int(r["threads"]), "kind": "reference", "ms_per_step": r["ms_per_step"],
                   "sample": f"{r['steps']} MD steps after {r['warmup']} warm-up steps of the full {n}-atom system on the reference CPU platform"}
    # ---- multi-rank runs carry a correctness bit: the same start state evolved by a plain one-rank context and by a fresh
    # context with the communicator must give bit-identical forces, energy and positions (tests/TestNonbondedForce.h:1013-1083
    # testParallelComputation; here with integer sums the agreement is exact, not within a tolerance)
    bit_identical = None
    if dist is not None:
        bit_identical = multi_rank_identity_check(sysd, rank, world, local, dist, torch, b, workloads)
    ctx.close()
    plugin = ref_gpu = None
    if world > 1 and not args.no_plugin:
        # the same N GPUs driven by ONE process through the plugin (DeviceIndex=0,...,N-1): the other ranks have released their GPUs
        # and wait on the host (gloo) while rank 0 runs the reference's own Context/Integrator machinery on the B200 platform
        torch.cuda.synchronize()
        host_group = dist.new_group(backend="gloo")
        dist.barrier(group=host_group)
        if rank == 0:
            plugin, ref_gpu = plugin_records(sysd, alpha, grid, args, n, devices=list(range(world)))
        dist.barrier(group=host_group)
    elif rank == 0 and not args.no_plugin:
        plugin, ref_gpu = plugin_records(sysd, alpha, grid, args, n)
    if rank == 0:
        line = {
            "metric": "ns/day (PME, 2 fs, mixed precision)", "value": ns_per_day(ms_per_step), "unit": "ns/day", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32 pair math, int64 fixed-point force sums, f64 integration (mixed)", "data": "synthetic",
            "config": {"workload": workload_string(args.workload, n, alpha, grid), "atoms": n,
                       "parallelism": (f"direct-space rows dealt weighted round-robin over {world} ranks, PME on rank 0; forces reduce-scattered, integration sharded and "
                                       "positions all-gathered by peer-memory kernels over NVLink (int64 sums)") if world > 1 else "one GPU",
                       "l2": "per-step working set ~285 MB for stmv (> 126 MB L2); steady-state MD loop, no explicit flush"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu,
            "plugin": plugin, "reference_gpu": ref_gpu, "bit_identical_to_one_rank": bit_identical,
            "pair_interactions_per_s": pairs_total / (tile_avg_ms * 1e-3) if tile_avg_ms > 0 else None,
            "in_cutoff_pairs": pairs_total,
            "kernel_ms_per_step": {k: v[0] / prof_steps for k, v in fam.items()},
            "nblist": {"tiles": dense_tiles, "sparse_tiles": sparse_tiles, "units": int(stats["num_units"]), "exclusion_tiles": int(stats["num_exclusion_tiles"]),
                       "rebuilds": int(stats["rebuilds"]), "steps": int(stats["steps"]),
                       "tile_efficiency": pairs / max(1.0, float(dense_slots))},
        }
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
