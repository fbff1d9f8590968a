#!/usr/bin/env python
"""Benchmark of the Pixel-simulator hot path (see DESIGN.md "Measurement").

A "step" is one RK212 step (2 fused RHS+RK stages) of the gray-scott-3d model
over a full-box compartment, driven through the product's public call
(``PixelSim.run`` -> C-ABI ``smeb200_run``) in the reference benchmark's
fixed-step configuration (benchmark/benchmark.cpp:214-216: maxErr = DBL_MAX,
maxTimestep = dt): the controller still reads the fused error norm every step.

With --gpus N (torchrun, one rank per GPU) the SAME global box (512^3 by
default: BASELINE config 4) is cut into N z-slabs — strong scaling; --weak keeps
--nz planes per GPU instead (the box grows with N).

  python bench.py --gpus N --steps K --warmup W            (own arm)
  python bench.py --impl reference --gpus N --steps K ...  (CPU reference arm)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pixel_voxel_species_rhs_evals_per_sec"
UNIT = "evals/s"
DT = 0.25  # below the forward-Euler bound 1/(2*3*D/dx^2) = 3.18 for voxel size 2


def gray_scott_3d_state(nx, ny, nz, z0=0):
    """U = 1, V = separable Gaussian spots with period 100 voxels; [N, 2] in
    voxel-list order (z, x, y)"""
    def axis(n, c, off=0):
        p = ((np.arange(n) + off) % 100).astype(np.float64)
        return np.exp(-((p - c) ** 2) / 50.0)
    v = (axis(nz, 25.0, z0)[:, None, None] * axis(nx, 33.0)[None, :, None]
         * axis(ny, 40.0)[None, None, :]
         + axis(nz, 60.0, z0)[:, None, None] * axis(nx, 70.0)[None, :, None]
         * axis(ny, 66.0)[None, None, :])
    out = np.empty((nx * ny * nz, 2), dtype=np.float64)
    out[:, 0] = 1.0
    out[:, 1] = v.reshape(-1)
    return out


def make_problem(nx, ny, nz, integrator, device=-1, z0=0):
    import spatial_model_editor_b200 as smeb
    from spatial_model_editor_b200 import models as M
    p = M.box_problem(M.GRAY_SCOTT, nx, ny, nz, conc0=gray_scott_3d_state(nx, ny, nz, z0),
                      voxel_size=(2.0, 2.0, 2.0), integrator=integrator,
                      max_err_rel=smeb.DBL_MAX, max_err_abs=smeb.DBL_MAX, max_timestep=DT)
    p.device = device
    return p


def make_listed_problem(args, integrator, device):
    """--workload c2disc / c3: voxel-list geometries (index-list kernels).
    c2disc: gray-scott on the disc of the 100x100 example upscaled to size^2
    (N ~ 0.315 size^2); c3: two brusselator compartments (outer frame / inner
    square) coupled by a membrane flux, size^2 voxels in total."""
    import spatial_model_editor_b200 as smeb
    from spatial_model_editor_b200 import models as M
    n = args.size
    tol = dict(max_err_rel=0.005) if args.adaptive else dict(
        max_err_rel=smeb.DBL_MAX, max_err_abs=smeb.DBL_MAX, max_timestep=args.dt)
    if args.workload == "c4blob":
        # BASELINE config 4's irregular variant: the blob of the gray-scott-3d example
        # (11 789 of 50^3 voxels) upscaled to size^3 by nearest neighbour — a 3-D
        # voxel list that z-slabs can cut
        from spatial_model_editor_b200 import geometry as G
        z = np.load(os.path.join(ROOT, "tests", "golden", "gray-scott-3d.npz"))
        small = z["labels"].astype(np.int32)
        src = (np.arange(n) * small.shape[0]) // n
        lab = small[np.ix_(src, (np.arange(n) * small.shape[1]) // n,
                           (np.arange(n) * small.shape[2]) // n)]
        label = int(np.bincount(small.ravel())[1:].argmax()) + 1
        lab = (lab == label).astype(np.int32)
        vox = G.compartment_voxels(lab, 1)
        px, py, pz = ((vox[:, k] % 100).astype(np.float64) for k in range(3))
        conc0 = np.ones((len(vox), 2))
        conc0[:, 1] = (np.exp(-((px - 33.0) ** 2 + (py - 40.0) ** 2 + (pz - 25.0) ** 2) / 50.0)
                       + np.exp(-((px - 70.0) ** 2 + (py - 66.0) ** 2 + (pz - 60.0) ** 2) / 50.0))
        p = M.labelled_problem(lab, [dict(label=1, id="blob", kin=M.GRAY_SCOTT, conc0=conc0)],
                               voxel_size=(2.0, 2.0, 2.0), integrator=integrator, **tol)
        text = f"gray-scott kinetics on the gray-scott-3d blob upscaled to {n}^3"
    elif args.workload == "c2disc":
        lab = M.disc_labels(n, n)
        from spatial_model_editor_b200 import geometry as G
        vox = G.compartment_voxels(lab, 1)
        px, py = (vox[:, 0] % 100).astype(np.float64), (vox[:, 1] % 100).astype(np.float64)
        conc0 = np.ones((len(vox), 2))
        conc0[:, 1] = (np.exp(-((px - 33.0) ** 2 + (py - 40.0) ** 2) / 50.0)
                       + np.exp(-((px - 70.0) ** 2 + (py - 66.0) ** 2) / 50.0))
        p = M.labelled_problem(lab, [dict(label=1, id="disc", kin=M.GRAY_SCOTT, conc0=conc0)],
                               integrator=integrator, **tol)
        text = f"gray-scott kinetics on the example disc upscaled to {n}x{n}"
    else:
        p = M.two_compartment_brusselator(n, n, integrator=integrator, **tol)
        text = (f"brusselator kinetics in two compartments (outer frame / inner square) + "
                f"membrane flux km*(X_a-X_b), {n}x{n} pixels")
    p.device = device
    sizes = ", ".join(f"{c.id}: N={c.n_voxels} S={c.n_species}" for c in p.compartments)
    return p, f"{text}; voxel-list geometry ({sizes})"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                smax.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for k, n in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def host_threads():
    """the host cores this process may use; torchrun exports OMP_NUM_THREADS=1 to
    its workers, which must not decide how many threads the CPU arm gets"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_sample_problem(args, integrator, nz):
    """a slab of the bench workload for the CPU arm: same kinetics, same x-y
    extent (capped at 512), nz planes"""
    nx = ny = min(args.size, 512)
    nz = 1 if (args.nz == 1) else nz
    return make_problem(nx, ny, nz, integrator), (nx, ny, nz)


def time_oracle(sim, evals_per_step, budget_s, max_steps):
    """run fixed steps of DT until the time budget or the step cap is reached"""
    t0 = time.perf_counter()
    steps = 0
    while True:
        steps += sim.run(DT)
        el = time.perf_counter() - t0
        if el > budget_s or steps >= max_steps:
            break
    return evals_per_step * steps / el, steps, el


def cpu_baseline(args, integrator, threads, budget_s=12.0):
    """the CPU restatement of the reference TBB Pixel path on a bounded sample of
    the same workload (a 32-plane slab of the same x-y extent, about 10-20 s of
    CPU work), all host threads"""
    from oracle.oracle_sim import OraclePixelSim
    if args.workload in ("c2disc", "c3", "c4blob"):
        import copy
        a2 = copy.copy(args)
        a2.size, a2.adaptive = min(args.size, 192 if args.workload == "c4blob" else 2048), False
        p, text = make_listed_problem(a2, integrator, -1)
        sim = OraclePixelSim(p, threads=threads)
        sim.nextTimestep = args.dt
        sim.run(args.dt)
        t0, steps = time.perf_counter(), 0
        while time.perf_counter() - t0 < budget_s:
            steps += sim.run(args.dt)
        el = time.perf_counter() - t0
        ncores = threads
        return {"value": p.evals_per_step() * steps / el, "unit": UNIT, "cores": ncores,
                "kind": "port", "sample": f"{text}, {args.integrator.upper()} fixed step, "
                                          f"{steps} steps, {el:.1f} s, OpenMP threads={ncores}"}
    p, (nx, ny, nz) = cpu_sample_problem(args, integrator, 32)
    sim = OraclePixelSim(p, threads=threads)
    sim.nextTimestep = DT  # skip the controller's 1e-7 start-up step
    sim.run(DT)    # warm-up
    value, steps, el = time_oracle(sim, p.evals_per_step(), budget_s, 100000)
    ncores = threads
    return {"value": value, "unit": UNIT, "cores": ncores,
            "kind": "port",
            "sample": f"gray-scott-3d {args.integrator.upper()} fixed step, {nx}x{ny}x{nz} "
                      f"slab ({nx * ny * nz} voxels x 2 species), {steps} steps, {el:.1f} s, "
                      f"OpenMP threads={ncores}"}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path — here the
    oracle port (the reference cannot be built: see DESIGN.md), all host threads.
    Each step is one RK step over a bounded sample (a 32-plane slab) of the
    workload."""
    if rank != 0:
        return
    from spatial_model_editor_b200.problem import INTEGRATOR_NAMES
    from oracle.oracle_sim import OraclePixelSim
    integrator = INTEGRATOR_NAMES[args.integrator]
    p, (nx, ny, nz) = cpu_sample_problem(args, integrator, 32)
    cores = host_threads()
    sim = OraclePixelSim(p, threads=cores)
    sim.nextTimestep = DT
    for _ in range(max(args.warmup, 1)):
        sim.run(DT)
    t0 = time.perf_counter()
    steps = 0
    for _ in range(args.steps):
        steps += sim.run(DT)
        if time.perf_counter() - t0 > 120.0:
            break
    el = time.perf_counter() - t0
    value = p.evals_per_step() * steps / el
    sample = (f"{nx}x{ny}x{nz} slab of the gray-scott-3d workload per step "
              f"({nx * ny * nz} voxels x 2 species), OpenMP threads={cores}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * el / max(steps, 1), "higher_is_better": True,
        "scaling": "weak" if args.weak else "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"gray-scott-3d {args.integrator.upper()} fixed-step dt={DT}, "
                               f"{sample}"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def run_refcuda_port(args, rank):
    """--impl refcuda-port (not a driver arm; SURVEY §8(d) "Reference-CUDA baseline"):
    the unfused kernel sequence of the reference's own GPU backend, restated in
    baselines/refcuda_port.cu, on the headline workload on ONE GPU — "the kernel to
    beat", recompiled for sm_100a.  Same metric, same timing rules."""
    if rank != 0:
        return
    from baselines.refcuda_port import RefCudaPort
    nx = ny = args.size
    nz = args.nz or args.size
    W, K = max(args.warmup, 3), args.steps
    sim = RefCudaPort(nx, ny, nz, gray_scott_3d_state(nx, ny, nz, 0))
    sim.steps(args.dt, W)
    l0 = sim.launches()
    sampler = ClockSampler(0)
    sampler.start()
    ms, _ = sim.steps(args.dt, K)
    clocks = sampler.stop()
    launches = sim.launches() - l0
    evals = 2 * nx * ny * nz * 2  # RK212: 2 stages x voxels x species
    value = evals * K / (ms * 1e-3)
    print(json.dumps({
        "impl": "reference-cuda-port", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": ms / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"gray-scott-3d kinetics, {nx}x{ny}x{nz} full-box compartment, "
                               f"S=2, RK212 fixed step dt={args.dt}",
                   "l2": "inputs larger than L2" if nx * ny * nz * 16 > 2 ** 28 else "small",
                   "kernels": "reference GPU backend restated: reaction, diffusion (uint32 nn "
                              "gather), rk212 substep 1/2, error, 1-block reduce, clamp; "
                              "blocking error download per step"},
        "gpu_launches": launches, "clocks": clocks}))


def run_ensemble(args, rank, world, local_rank):
    """--workload ensemble: BASELINE config 5 — liver-simplified (7800 x 8 + 481 x 6
    voxel-species, one membrane), 64 members scanning one rate constant
    log-uniformly +-10x, member m on rank m mod N, members of one GPU concurrent
    on their own streams.  Not the headline line: reported for completeness."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import load_golden
    from spatial_model_editor_b200 import ensemble as EN
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    p, _, _ = load_golden("liver-simplified-scan")
    vals = EN.scan_values(p.params[0], args.members)
    if args.ensemble_impl == "batched":
        # members as gridDim.y of one launch per stage (smeb200_ensemble_*)
        ens = EN.GroupedEnsemble(p, vals, rank=rank, world=world, device=local_rank,
                                 groups=args.groups)
        impl_text = (f"member-batched launches (smeb200_ensemble_run), {args.groups} "
                     f"group(s) per GPU")
    else:
        ens = EN.Ensemble(p, vals, rank=rank, world=world, device=local_rank,
                          concurrency=args.concurrency)
        impl_text = f"{args.concurrency} concurrent single simulations per GPU"
    if args.ensemble_impl == "batched":
        # warm-up on a throw-away ensemble: run() continues from the members' states
        warm = EN.BatchedEnsemble(p, vals[:max(world, 1) * 2], rank=rank, world=world,
                                  device=local_rank)
        warm.run(args.sim_time / 10)
        warm.close()
    else:
        ens.run(args.sim_time / 10)  # warm-up: module compile, allocator, clocks
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = ens.run(args.sim_time)
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
    steps = sum(r.steps + r.discarded for r in res)
    if world > 1:
        t = torch.tensor([el, steps], dtype=torch.float64, device="cuda")
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        el, steps = float(tm[0]), int(t[1])
    evals = steps * p.evals_per_step()
    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": evals / el, "unit": UNIT, "n_gpus": world,
            "steps": steps, "ms_per_step": 1e3 * el / max(steps, 1), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"liver-simplified parameter scan, {args.members} members x "
                                   f"t={args.sim_time}, RK212 adaptive (model settings), "
                                   f"{impl_text}",
                       "parallelism": f"one simulation per GPU slot x{world}"},
            "wall_s": el, "errors": sorted({r.error for r in res if r.error})}))
    if world > 1:
        dist.destroy_process_group()


def slab_planes(nz_total, rank, world):
    from spatial_model_editor_b200.multigpu import slab_range
    z0, z1 = slab_range(nz_total, rank, world)
    return z1 - z0


def parity_vs_oracle(args, integrator, local_rank, steps=3):
    """the CUDA path and the CPU oracle on the slab the CPU arm is timed on
    (512 x 512 x 32 of the headline workload, all host threads): max relative
    difference after `steps` fixed steps"""
    import spatial_model_editor_b200 as smeb
    from oracle.oracle_sim import OraclePixelSim
    p, (nx, ny, nz) = cpu_sample_problem(args, integrator, 32)
    p.device = local_rank
    gpu = smeb.PixelSim(p)
    cpu = OraclePixelSim(p, threads=host_threads())
    gpu.nextTimestep = cpu.nextTimestep = DT
    ng, nc = gpu.run(steps * DT), cpu.run(steps * DT)
    a, b = gpu.getConcentrations(0), cpu.getConcentrations(0)
    rel = float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
    gpu.close()
    return {"max_rel": rel, "steps": int(ng), "oracle_steps": int(nc),
            "box": f"{nx}x{ny}x{nz}", "tolerance": 1e-12}


def parity_slabs_vs_one_gpu(integrator, rank, world, local_rank, dist, torch, steps=4):
    """a reduced box cut into `world` z-slabs against ONE GPU on the same box: the
    gathered slabs must be bit-identical (max |difference| = 0)"""
    import spatial_model_editor_b200 as smeb
    from spatial_model_editor_b200.multigpu import SlabPixelSim, default_comm_name
    nx, ny, nz = 128, 64, 16 * world
    pg = make_problem(nx, ny, nz, integrator, device=local_rank)
    sim = SlabPixelSim(pg, rank, world, comm_name=default_comm_name("parity"))
    if sim.errorMessage():
        raise SystemExit("bench parity: " + sim.errorMessage())
    sim.nextTimestep = DT
    n = sim.run(steps * DT)
    mine = torch.from_numpy(np.ascontiguousarray(sim.getConcentrations(0))).cuda()
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    out = None
    if rank == 0:
        one = smeb.PixelSim(make_problem(nx, ny, nz, integrator, device=local_rank))
        one.nextTimestep = DT
        n1 = one.run(steps * DT)
        ref = one.getConcentrations(0)
        got = torch.cat(parts).cpu().numpy()
        out = {"max_abs": float(np.max(np.abs(got - ref))), "steps": int(n),
               "one_gpu_steps": int(n1), "box": f"{nx}x{ny}x{nz}", "slabs": world,
               "halo_exchanges": sim.haloExchanges()}
        one.close()
    dist.barrier()
    sim.close()
    return out


def refcuda_baseline(args, nx, ny, nz, steps=8):
    """"the kernel to beat" beside the product, same box, same process: the
    reference GPU backend's unfused kernel sequence restated
    (baselines/refcuda_port.cu; cudapixelsim.cpp:164-349, 1549-1555, 1584-1616)"""
    from baselines.refcuda_port import RefCudaPort
    sim = RefCudaPort(nx, ny, nz, gray_scott_3d_state(nx, ny, nz, 0))
    sim.steps(args.dt, 3)
    sampler = ClockSampler(0)
    sampler.start()
    ms, _ = sim.steps(args.dt, steps)
    clocks = sampler.stop()
    evals = 2 * nx * ny * nz * 2
    return {"ms_per_step": ms / steps, "value": evals * steps / (ms * 1e-3), "unit": UNIT,
            "steps": steps, "kind": "port", "source": "baselines/refcuda_port.cu",
            "clocks": clocks}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--size", type=int, default=512, help="box edge (voxels)")
    ap.add_argument("--nz", type=int, default=0,
                    help="planes of the global box (default: size); with --weak: per GPU")
    ap.add_argument("--weak", action="store_true",
                    help="N GPUs: --nz planes PER GPU (the box grows with N) instead of "
                         "one global box cut N ways")
    ap.add_argument("--integrator", default="rk212")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the parity checks, the e2e interval sweep and the "
                         "reference-GPU-kernel baseline (A/B runs)")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--workload", default="c4",
                    help="c4 (headline) | c4blob | c2disc | c3 | ensemble")
    ap.add_argument("--adaptive", action="store_true",
                    help="c2disc / c3: adaptive steps at relTol 0.005 instead of fixed dt")
    ap.add_argument("--dt", type=float, default=DT)
    ap.add_argument("--members", type=int, default=64)
    ap.add_argument("--concurrency", type=int, default=8)
    ap.add_argument("--groups", type=int, default=2,
                    help="ensemble workload: member-batched ensembles per GPU, one "
                         "host thread each")
    ap.add_argument("--ensemble-impl", default="batched",
                    help="ensemble workload: batched (members = gridDim.y) | threads")
    ap.add_argument("--sim-time", type=float, default=1.0)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.impl == "refcuda-port":
        run_refcuda_port(args, rank)
        return
    if args.workload == "ensemble":
        run_ensemble(args, rank, world, local_rank)
        return

    import torch
    import spatial_model_editor_b200 as smeb
    from spatial_model_editor_b200.problem import INTEGRATOR_NAMES, STAGES
    import torch.distributed as dist

    integrator = INTEGRATOR_NAMES[args.integrator]
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    nx = ny = args.size
    nz_arg = args.nz or args.size
    # strong scaling (default): one global box; weak: nz_arg planes per GPU
    nz_total = nz_arg * world if args.weak else nz_arg
    W, K = max(args.warmup, 3), args.steps
    listed = args.workload in ("c2disc", "c3", "c4blob")
    DTs = args.dt
    if listed:
        from spatial_model_editor_b200.multigpu import SlabPixelSim
        p, workload_text = make_listed_problem(args, integrator, local_rank)
        p.kernel_variant = args.variant
        evals_per_step = p.evals_per_step()
        dom = max(p.compartments, key=lambda c: c.n_voxels * c.n_species)
        dom_index = p.compartments.index(dom)
        if world > 1 and args.workload != "c4blob":
            raise SystemExit("bench: c2disc / c3 are 2-D images (one z plane): z-slabs have "
                             "nothing to cut; the 3-D voxel-list workload is c4blob")
        if world > 1:  # the voxel list and its membranes cut at plane boundaries
            sim = SlabPixelSim(p, rank, world, variant=args.variant)
        else:
            sim = smeb.PixelSim(p)
        values_per_rank = dom.n_voxels * dom.n_species // world
        nn_bytes = sim.neighbourBytesPerVoxel(dom_index) * dom.n_voxels
        workload_text += (f"; neighbour data {sim.neighbourBytesPerVoxel(dom_index)} B/voxel"
                          f" ({sim.neighbourRuns(dom_index)} runs)")
        nz_local = 1
    elif world > 1:
        from spatial_model_editor_b200.multigpu import SlabPixelSim
        # the global box is only described (no global array is ever built);
        # every rank generates its own planes, ghosts included
        p = make_problem(nx, ny, 1, integrator, device=local_rank)
        p.compartments[0].box = (nx, ny, nz_total)
        sim = SlabPixelSim(p, rank, world, make_slab=lambda z0, z1: gray_scott_3d_state(
            nx, ny, z1 - z0, z0), variant=args.variant)
        evals_per_step = STAGES[integrator] * nx * ny * nz_total * 2
        nz_local = slab_planes(nz_total, rank, world)
    else:
        p = make_problem(nx, ny, nz_total, integrator, device=local_rank)
        p.kernel_variant = args.variant
        sim = smeb.PixelSim(p)
        evals_per_step = p.evals_per_step()
        nz_local = nz_total
    if sim.errorMessage():
        raise SystemExit("bench: " + sim.errorMessage())
    if not listed:
        values_per_rank = nx * ny * nz_local * 2
        nn_bytes = 0
        workload_text = (f"gray-scott-3d kinetics, {nx}x{ny}x{nz_total} full-box "
                         f"compartment ({nz_local} planes on rank 0), S=2")

    # warm-up: W steps of DT (the controller's 1e-7 start-up step is skipped by
    # seeding nextTimestep, which run() otherwise raises to DT after one step)
    adaptive = listed and args.adaptive
    if adaptive:
        sim.run(W * DTs)  # start-up transient of the step-size controller
    else:
        sim.nextTimestep = DTs
        assert sim.run(W * DTs) == W
    launches0 = sim.launchCount()
    disc0 = sim.discardedSteps
    # dominant-kernel durations by CUDA events inside the timed region, on every
    # 4th step: the two event records per launch cost a few microseconds of stream
    # time and keep that launch from overlapping its predecessor's tail
    timing_stride = 1 if K < 8 else 4
    sim.setKernelTiming(timing_stride)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    steps = sim.run(K * DTs)
    barrier()
    ms = sim.lastRunMs()
    clocks = sampler.stop() if rank == 0 else None
    kms, klaunches = sim.kernelTiming()
    sim.setKernelTiming(False)
    launches = sim.launchCount() - launches0
    if adaptive:  # a step = one attempted RK step (accepted or discarded)
        K = steps + (sim.discardedSteps - disc0)
    else:
        assert steps == K, (steps, K, sim.errorMessage())
    if world > 1:
        ms = allmax(ms)
        tl = torch.tensor([launches], dtype=torch.float64, device="cuda")
        dist.all_reduce(tl, op=dist.ReduceOp.SUM)
        launches = int(tl.item())
    value = evals_per_step * K / (ms * 1e-3)

    # End to end through the public call with HOST buffers, the way Simulation
    # drives a simulator (simulate.cpp:397-462): the state is uploaded once from
    # pinned host memory (a restart: setConcentrations), then image intervals of Ke
    # steps, each followed by the download of the whole state (getConcentrations).
    # The headline e2e uses Ke = --steps; the sweep shows the share of the copies.
    ncomp = len(p.compartments)
    hosts = [sim.getConcentrationsWithGhosts(0)] if world > 1 else [
        np.array(sim.getConcentrations(i)) if ncomp > 1 else sim.getConcentrations(i)
        for i in range(ncomp)]
    state_bytes = sum(h.nbytes for h in hosts)

    sim.setAsyncDownload(True)  # what the SME adapter does: the D2H starts in run()

    def e2e_run(ke, intervals):
        barrier()
        t0 = time.perf_counter()
        for i, h in enumerate(hosts):
            sim.setConcentrations(i, h)
        done = 0
        checksum = 0.0
        for _ in range(intervals):
            n = sim.run(ke * DTs)
            done += ke if adaptive else n
            for i in range(ncomp):
                out = sim.getConcentrations(i)
                checksum += float(out[::4097, -1].sum())
        barrier()
        el = allmax(time.perf_counter() - t0)
        return {"value": evals_per_step * done / el, "steps": done, "intervals": intervals,
                "steps_per_interval": ke, "seconds": el,
                "h2d_bytes_per_step": state_bytes / max(done, 1),
                "d2h_bytes_per_step": state_bytes * intervals / max(done, 1),
                "checksum": checksum}

    head = e2e_run(args.steps, 2)
    e2e_sweep = None
    if not args.no_extras and not adaptive:
        e2e_sweep = {str(ke): {k: v for k, v in e2e_run(ke, n).items()
                               if k in ("value", "seconds", "intervals")}
                     for ke, n in ((1, 4), (20, 2), (200, 1))}

    # roofline of the dominant kernel (the fused stage kernel): algorithmic bytes
    # per launch = 8 B x values x (reads + writes of whole arrays) averaged over
    # the stages of one step: RK212 16 + 24 = 40 B per value per step
    bytes_per_value_step = {0: 16, 1: 40, 2: 64, 3: 152}[integrator]
    alg_bytes_per_launch = (bytes_per_value_step * values_per_rank / STAGES[integrator]
                            + nn_bytes)
    launch_scope = "whole compartment"
    if world > 1 and listed:
        nn_bytes //= world
        alg_bytes_per_launch = (bytes_per_value_step * values_per_rank / STAGES[integrator]
                                + nn_bytes)
        launch_scope = ("rank 0: its share of the voxel list (the two boundary planes are "
                        "launched apart: an upper bound of the launch's bytes)")
    elif world > 1:
        timed = sim.timedLaunchPlanes()
        alg_bytes_per_launch *= timed / nz_local
        launch_scope = f"rank 0: {timed} of its {nz_local} owned planes"
    peak, peak_src = measured_peak()
    traffic = traffic_src = None
    if (nx, ny, nz_total, args.integrator, world) == (512, 512, 512, "rk212", 1) and not listed:
        for name in ("r2_traffic.json", "r1_v4_traffic.json"):
            try:  # ncu dram__bytes of the same kernel on this config (builder-run)
                with open(os.path.join(ROOT, "profiles", name)) as f:
                    traffic = json.load(f)["dram_bytes_per_launch"]["mean"]
                traffic_src = f"profiles/{name} (ncu --set full, builder-run; not measured by this run)"
                break
            except Exception:
                continue
    roofline = None
    if klaunches:
        achieved = alg_bytes_per_launch / (kms / klaunches * 1e-3) / 1e9
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": peak_src,
                    "kernel": sim.dominantKernel(), "launches_timed": klaunches,
                    "avg_launch_ms": kms / klaunches,
                    "algorithmic_bytes_per_launch": alg_bytes_per_launch,
                    "launch_scope": launch_scope,
                    # the same bytes over the whole step (launch gaps included, no
                    # event records around the launches): a lower bound that does
                    # not carry the few microseconds the timing events add to a
                    # short launch; single-compartment workloads only
                    "frac_of_step": (bytes_per_value_step * values_per_rank
                                     + nn_bytes * STAGES[integrator])
                    / (ms / K * 1e-3) / 1e9 / peak if (ncomp == 1 and world == 1) else None,
                    "timed_steps": f"every {timing_stride}" if timing_stride > 1 else "all"}
    sim.close()
    del hosts

    parity = {}
    extras = not args.no_extras and not listed
    if extras and world > 1:
        parity["slabs_vs_one_gpu"] = parity_slabs_vs_one_gpu(integrator, rank, world,
                                                             local_rank, dist, torch)
    if extras and rank == 0:
        parity["vs_oracle"] = parity_vs_oracle(args, integrator, local_rank)
    cpu = refcuda = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(args, integrator, threads=host_threads())
    if extras and rank == 0 and world == 1 and args.integrator == "rk212":
        try:
            refcuda = refcuda_baseline(args, nx, ny, nz_total)
            refcuda["product_over_refcuda"] = (ms / K) and refcuda["ms_per_step"] / (ms / K)
        except Exception as e:  # e.g. not enough memory beside another process
            refcuda = {"unavailable": str(e)}
    # every stage streams its whole input and output arrays: nothing of a step is
    # left in L2 for the next one when the rotating buffers together exceed it
    nbuf = {0: 2, 1: 3, 2: 3, 3: 4}[integrator]
    working_set = nbuf * values_per_rank * 8
    l2_text = ("rotating state buffers on rank 0: %d x %.3f GiB %s the 126 MB L2" % (
        nbuf, values_per_rank * 8 / 2 ** 30,
        "— larger than" if working_set > 126e6 * 1.5 else
        "— NOT larger than (L2-resident numbers: no flush between steps)"))
    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "weak" if args.weak else "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {
                "workload": workload_text + f", {args.integrator.upper()} " + (
                    "adaptive relTol=0.005 (a step = one attempted step)" if adaptive else
                    f"fixed step dt={DTs} (maxErr=DBL_MAX like benchmark.cpp:214-216)"),
                "l2": l2_text,
                "parallelism": (f"z-slabs x{world}, " + ("weak: box grows with N" if args.weak
                                else "strong: one global box")) if world > 1 else "single GPU",
                "kernel_variant": args.variant},
            "e2e": {"value": head["value"], "unit": UNIT,
                    "h2d_bytes_per_step": head["h2d_bytes_per_step"],
                    "d2h_bytes_per_step": head["d2h_bytes_per_step"],
                    "note": "Simulation's usage (simulate.cpp:397-462): one pinned H2D of the "
                            "state (restart), then image intervals of --steps steps, each "
                            "followed by the D2H of the whole state; bytes are totals / steps",
                    "intervals": head["intervals"], "steps_per_interval": head["steps_per_interval"],
                    "seconds": head["seconds"], "checksum": head["checksum"],
                    "by_steps_per_interval": e2e_sweep},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
            "parity": parity or None, "cpu_baseline": cpu, "refcuda_baseline": refcuda}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
