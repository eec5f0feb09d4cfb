#!/usr/bin/env python
"""bench.py — DoF-timesteps/s of the PFC Newton-Krylov hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # our arm (CUDA, through the C ABI)
  python bench.py --impl reference --steps K --warmup W    # CPU arm: the oracle port on the host cores

A "step" is one iterate_timestep (mpi.cpp:576-593): Newton loop with an assembly / residual evaluation and a
GMRES(28) solve per iteration.  Workloads (box scale S = degree * 2^(n_refines - 6) keeps the node
spacing of the shipped parameter file):

  M4 (default; BASELINE.json config 4): Q1, n_refines 12, 50 356 227 DoF, polycrystalline initial condition
      (64 periodic Voronoi grains, numpy PCG64 seed 12345).  --gpus N is STRONG scaling: the same mesh at every N,
      cut along the Morton curve into N chunks (p4est partition).
  M2 (config 2): Q2, n_refines 9, 3 151 875 DoF, perfect hexagonal lattice.
  M3 (config 3): Q1, n_refines 11, 12 595 203 DoF, dislocation dipole at (+-2 a S, 0), stress projection after every step.
  M5 (config 5): M4 mesh, perfect lattice; matrix-free and assembled operator run back to back, psi compared.

Prints ONE JSON line on rank 0.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

A = 4.0 * math.pi / math.sqrt(3.0)
DT, THETA, EPS, PSI0 = 0.1, 1.0, -0.8, -0.43  # parameter-files/hexagonal_lattice.toml:12-29

WORKLOADS = {
    "M4": dict(degree=1, n_refines=12, ic="polycrystal", config="BASELINE.json config 4"),
    "M2": dict(degree=2, n_refines=9, ic="perfect", config="BASELINE.json config 2"),
    "M3": dict(degree=1, n_refines=11, ic="dipole", config="BASELINE.json config 3", stress=True),
    "M5": dict(degree=1, n_refines=12, ic="perfect", config="BASELINE.json config 5", both_operators=True),
}
N_GRAINS, GRAIN_SEED = 64, 12345


def host_cores():
    """cores this process may use; torch.distributed.run exports OMP_NUM_THREADS=1, which must not decide it"""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def box(degree, n_refines, rep=(1, 1)):
    """commensurate box: width 12 a S, height 6 sqrt(3) a S = 6 S * 4 pi (SURVEY.md §7 hard part 7)"""
    S = degree * 2 ** (n_refines - 6)
    w, h = 6.0 * A * S * rep[0], 3.0 * math.sqrt(3.0) * A * S * rep[1]
    return S, (-w, -h), (w, h)


def resolve_workload(args):
    wl = dict(WORKLOADS[args.workload])
    if args.degree:
        wl["degree"] = args.degree
    if args.n_refines:
        wl["n_refines"] = args.n_refines
    if args.ic:
        wl["ic"] = args.ic
    wl["name"] = args.workload
    wl["custom"] = bool(args.degree or args.n_refines or args.ic)
    return wl


def workload_string(wl, n_dofs):
    """identical in both arms (the driver compares the config of the two lines)"""
    ic = {"perfect": "perfect hexagonal lattice",
          "dipole": "dislocation dipole at (+-2 a S, 0), b = (+-a, 0)",
          "polycrystal": f"polycrystal of {N_GRAINS} periodic Voronoi grains (PCG64 seed {GRAIN_SEED})"}[wl["ic"]]
    return (f"{wl['name']}{' (custom size)' if wl['custom'] else ''}: Q{wl['degree']}, n_refines {wl['n_refines']}, "
            f"{n_dofs} DoF, {ic}, commensurate box at the shipped file's h; same mesh at every N (strong scaling)")


def initial_condition(obj, wl, p1, p2, S, A0):
    """obj: the stand-in (pfc_b200.StandIn) or the oracle mesh (oracle.Mesh); same generators, bit-equal results"""
    hex_fn = getattr(obj, "hexagonal_lattice", None) or obj.initial_condition
    if wl["ic"] == "perfect":
        return hex_fn(A0, PSI0)
    if wl["ic"] == "dipole":
        return hex_fn(A0, PSI0, [[2 * A * S, 0.0], [-2 * A * S, 0.0]], [[A, 0.0], [-A, 0.0]])
    seeds, alpha, offset = polycrystal_grains(p1, p2, N_GRAINS, GRAIN_SEED)
    return obj.polycrystal(A0, PSI0, seeds, alpha, offset)


def polycrystal_grains(p1, p2, n_grains, seed):
    """grain table of M4 (same as pfc_b200.polycrystal_grains; repeated here so that the CPU arm imports nothing of
    the product): seeds uniform in the box, rotation in [0, pi/3), offset in [0, a)^2 from numpy PCG64(seed)"""
    rng = np.random.Generator(np.random.PCG64(seed))
    uv = rng.random((n_grains, 2))
    seeds = np.array([p1[0], p1[1]]) + uv * np.array([p2[0] - p1[0], p2[1] - p1[1]])
    alpha = rng.random(n_grains) * (np.pi / 3.0)
    offset = rng.random((n_grains, 2)) * A
    return seeds, alpha, offset


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self._stop = index, [], threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)

    def _nvml_open(self):
        """NVML handle of the GPU (a sample costs microseconds: a 0.2 s timed region still gets several), or None"""
        try:
            import pynvml as N
            N.nvmlInit()
            h = N.nvmlDeviceGetHandleByIndex(self.index)
            reasons = getattr(N, "nvmlDeviceGetCurrentClocksEventReasons", None) or N.nvmlDeviceGetCurrentClocksThrottleReasons
            N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM), reasons(h)  # fail here, not in the loop
            return N, h, reasons
        except Exception:
            return None

    def _run(self):
        nv = self._nvml_open()
        while not self._stop.is_set():
            try:
                if nv is not None:
                    N, h, reasons = nv
                    r = int(reasons(h))
                    act = lambda bit: "Active" if r & bit else "Not Active"  # noqa: E731
                    # bits of nvml.h: SwPowerCap 0x4, HwSlowdown 0x8, SwThermalSlowdown 0x20, HwThermalSlowdown 0x40
                    self.samples.append([str(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)),
                                         str(N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)),
                                         act(0x8), act(0x40), act(0x20), act(0x4)])
                    self._stop.wait(0.02)
                    continue
            except Exception:
                nv = None  # fall back to nvidia-smi for the rest of the run
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(s[2 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(sm)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ---- CPU arm: the oracle port of the reference algorithm (unpreconditioned GMRES(28), mpi.cpp:471-492) ----
def _oracle(cores):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    O.set_threads(cores)  # explicit: OMP_NUM_THREADS=1 under torch.distributed.run must not shrink the arm
    return O


def replica_counts(O, wl, steps, warmup, cores):
    """(assemblies, GMRES iterations) of every time step of a full oracle run of the same physics on a small
    replica at the same h (the counts do not depend on the domain size at fixed h, SURVEY.md §0 [probe]).
    A polycrystal replica keeps the grain SIZE of the workload, i.e. max(1, 64 / 4^(dn)) grains."""
    nr = min(wl["n_refines"], 7 if wl["degree"] == 1 else 6)
    S, p1, p2 = box(wl["degree"], nr)
    om = O.Mesh(wl["degree"], nr, p1, p2)
    A0 = O.default_A0(PSI0, EPS)
    if wl["ic"] == "polycrystal":
        k = max(1, N_GRAINS // 4 ** (wl["n_refines"] - nr))
        Pn = om.polycrystal(A0, PSI0, *polycrystal_grains(p1, p2, k, GRAIN_SEED))
    else:
        Pn = initial_condition(om, wl, p1, p2, S, A0)
    Pn1 = Pn.copy()
    out = []
    for _ in range(warmup + steps):
        r = om.newton_step(Pn, Pn1, DT, THETA, EPS, threads=cores)
        out.append((r["n_assemblies"], sum(r["gmres_iters"])))
    return nr, out


def cpu_sample(O, wl, cores, sample_iters):
    """Bounded sample at FULL size, measured: mesh + initial condition, one assembly (mpi.cpp:300-467), the first
    `sample_iters` GMRES(28) iterations of the first Newton solve.  Small workloads (<= 4 M DoF) also run time step 1
    completely (all Newton iterations, all GMRES iterations)."""
    S, p1, p2 = box(wl["degree"], wl["n_refines"])
    om = O.Mesh(wl["degree"], wl["n_refines"], p1, p2)
    Psi = initial_condition(om, wl, p1, p2, S, O.default_A0(PSI0, EPS))
    t0 = time.perf_counter()
    vals, rhs = om.assemble(Psi, Psi, DT, THETA, EPS, threads=cores)
    t_asm = time.perf_counter() - t0
    om.gmres(vals, rhs, stop_after=2)  # touch the pages
    t0 = time.perf_counter()
    _, its, _, _ = om.gmres(vals, rhs, stop_after=sample_iters)
    t_gm = time.perf_counter() - t0
    s = {"om": om, "Psi": Psi, "vals": vals, "rhs": rhs, "n_dofs": int(om.n_dofs), "t_assembly_s": t_asm,
         "t_gmres_iter_s": t_gm / max(its, 1), "sample_iters": int(its), "measured_s": t_asm + t_gm, "real_step": None}
    if om.n_dofs <= 4_000_000:
        Pn, Pn1 = Psi.copy(), Psi.copy()
        t0 = time.perf_counter()
        r = om.newton_step(Pn, Pn1, DT, THETA, EPS, threads=cores)
        t_real = time.perf_counter() - t0
        its_real = sum(r["gmres_iters"])
        s["real_step"] = {"t_s": t_real, "n_assemblies": r["n_assemblies"], "gmres_iters": its_real, "Psi_after": Pn}
        s["t_gmres_iter_s"] = max(t_real - r["n_assemblies"] * t_asm, 0.0) / max(its_real, 1)
        s["measured_s"] += t_real
    return s


def cpu_model(O, wl, cores, steps, warmup, sample_iters):
    """DoF-steps/s of the CPU arm over the `steps` timed steps: per-step time = assemblies * t_assembly + GMRES
    iterations * t_iteration with both unit times MEASURED at full size and the counts of each step from the replica
    run (rescaled by the full-size count of step 1 when that step was run for real)."""
    s = cpu_sample(O, wl, cores, sample_iters)
    nr, counts = replica_counts(O, wl, steps, warmup, cores)
    scale = 1.0
    if s["real_step"]:
        scale = s["real_step"]["gmres_iters"] / max(counts[0][1], 1)  # counts[0] is step 1 from the same initial condition
    timed = counts[warmup:]
    per_step = [a * s["t_assembly_s"] + scale * g * s["t_gmres_iter_s"] for a, g in timed]
    t = sum(per_step)
    s["value"] = s["n_dofs"] * len(per_step) / t
    s["ms_per_step"] = 1e3 * t / len(per_step)
    s["modelled_s"] = t
    mean_a, mean_g = float(np.mean([c[0] for c in timed])), float(np.mean([c[1] for c in timed])) * scale
    what = (f"measured at full size on {cores} threads ({s['measured_s']:.1f} s): 1 assembly ({s['t_assembly_s']:.2f} s) + "
            f"{s['sample_iters']} GMRES(28) iterations (MGS, OpenMP)")
    if s["real_step"]:
        what += (f" + time step 1 run completely ({s['real_step']['t_s']:.1f} s, {s['real_step']['n_assemblies']} assemblies, "
                 f"{s['real_step']['gmres_iters']} GMRES iterations)")
    s["sample"] = (what + f"; modelled ({t:.0f} s for {len(per_step)} steps): assemblies * t_assembly + iterations * "
                   f"{s['t_gmres_iter_s']:.3f} s with the per-step counts of a full oracle run on a same-h replica "
                   f"(Q{wl['degree']}, n_refines {nr}; mean {mean_a:.1f} assemblies, {mean_g:.0f} iterations per step"
                   + (f", rescaled by {scale:.2f} = full-size / replica iterations of step 1" if s["real_step"] else "") + ")")
    return s


def run_reference(args):
    """--impl reference: the reference's algorithm on the host cores.  The reference binary needs
    deal.II + Trilinos + p4est + MPI (absent) so this is the oracle port, all threads."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    wl = resolve_workload(args)
    cores = host_cores()
    O = _oracle(cores)
    s = cpu_model(O, wl, cores, args.steps, args.warmup, sample_iters=28)
    line = {"impl": "reference", "metric": "DoF-timesteps/s", "value": s["value"], "unit": "DoF-steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": s["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_string(wl, s["n_dofs"]), "n_dofs": s["n_dofs"],
                       "solver": "GMRES(28) unpreconditioned, modified Gram-Schmidt, tol 1e-10 abs, Newton tol 1e-8 (reference algorithm)"},
            "cpu_baseline": {"value": s["value"], "unit": "DoF-steps/s", "cores": cores, "kind": "port", "sample": s["sample"],
                             "measured_s": s["measured_s"], "modelled_s": s["modelled_s"],
                             "t_assembly_s": s["t_assembly_s"], "t_gmres_iter_s": s["t_gmres_iter_s"]},
            "e2e": {"value": s["value"], "unit": "DoF-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def kernel_legs(solver, stats, sweeps, precond, mf, peak, n_nodes, ms_per_step):
    """Per-kernel legs on resident data (device time per launch, replayed from CUDA graphs, operands rotated through the
    Krylov basis so that they do not stay in L2), their launches per time step from the iteration counts of the timed
    steps, and their share of the step.  Algorithmic bytes are DESIGN.md §4's per-node figures times the nodes."""
    K = len(stats)
    its = [g for s in stats for g in s["gmres_iters"]]
    n_it = sum(its) / K
    n_solve = sum(s["n_solves"] for s in stats) / K
    n_asm = sum(s["n_assemblies"] for s in stats) / K
    legs = {}

    def add(name, ms, nbytes, per_step):
        legs[name] = {"ms": ms, "algorithmic_bytes": nbytes, "gbs": nbytes / ms / 1e6, "frac": nbytes / ms / 1e6 / peak,
                      "launches_per_step": per_step, "share_of_step": ms * per_step / ms_per_step}

    reduced = "reduced solve active" in solver.operator_info()[1]
    vec = (8.0 if reduced else 24.0) * n_nodes  # bytes of one Krylov vector
    if reduced:
        # psi-only reduced solve (csrc/pfc_reduced.cuh): per iteration one operator T y (two kernels) and one preconditioner
        # application (one exchange + sweeps - 1 sweep kernels); + the x update at the end of every cycle, the start of a
        # restarted cycle and the Newton update (three fields once per solve)
        rd = solver.time_reduced(reps=20)
        ms, by = rd["ms"], rd["bytes"]
        n_app = n_it + n_solve
        add("reduced operator, the three fields M^2 y, -K M y, K^2 y: lat_pstage_kernel", ms[1], by[1], n_app + n_solve)
        add("reduced operator, row 0 of J on the three fields: lat_j1_kernel<ROW0>", ms[2], by[2], n_app)
        if sweeps >= 2:
            add("Chebyshev steps 0 and 1 on the Schur stand-in in one pass: ring_sweep1a_kernel<FIRST>", ms[4], by[4], n_app)
        if sweeps >= 3:
            add("later Chebyshev sweep: ring_sweep1a_kernel", ms[5], by[5], n_app * (sweeps - 2))
        legs["reduced operator T y (all its kernels)"] = {"ms": ms[0], "launches_per_step": n_app, "share_of_step": ms[0] * n_app / ms_per_step}
        legs["reduced preconditioner (exchange + all its kernels)"] = {"ms": ms[3], "launches_per_step": n_app,
                                                                       "share_of_step": ms[3] * n_app / ms_per_step}
        precond = False
    else:
        ms_op, by_op = solver.time_operator(reps=20)
        add(("J apply: lat_j1_kernel" if solver.params.degree == 1 else "J apply: lat_kernel<LAT_J>") if mf else "J apply: spmv_sell_kernel",
            ms_op, by_op, n_it + n_solve)
    if precond:
        pc = solver.time_precond(reps=20)
        n_pc = n_it + n_solve  # one application per iteration + the x update at the end of a cycle
        ms, by = pc["ms"], pc["bytes"]
        if mf:
            q1 = ms[3] == 0.0  # Q1: closed-form kernels (csrc/pfc_ring_q1.cuh), the t kernel is fused into the right-hand side
            add("Chebyshev sweep on the Schur complement: " + ("ring_sweep1a_kernel" if q1 else "ring_sweep_kernel"), ms[1], by[0],
                n_pc * (sweeps - 1))
            if q1:
                add("t kernel + Schur right-hand side + first Chebyshev step: ring_rhs1a_kernel", ms[2], by[1], n_pc)
            else:
                add("Schur right-hand side + first Chebyshev step: lat_kernel<LAT_S>", ms[2], by[1], n_pc)
                add("t = M_L^-1 (r2 - K M_L^-1 r1): ring_t_kernel", ms[3], by[2], n_pc)
            add("back substitution chi, phi: " + ("ring_back1a_kernel" if q1 else "ring_back_kernel"), ms[4], by[3], n_pc)
        else:
            add("K product: kmv_kernel", ms[1], by[0], n_pc * (2 * sweeps + 1))
            add("Schur row: bcd_kernel", ms[2], by[1], n_pc * sweeps)
        legs["preconditioner application (all its kernels)"] = {"ms": pc["ms_apply"], "launches_per_step": n_pc,
                                                                "share_of_step": pc["ms_apply"] * n_pc / ms_per_step}
    kry = solver.time_krylov(dim=14, reps=10)
    dot_bytes = sum((d + 1) * vec for n in its for d in range(1, n + 1)) / K
    axpy_bytes = sum((d + 2) * vec for n in its for d in range(1, n + 1)) / K
    for nm, ms14, b14, tot in (("multidot", kry["ms_dot"], kry["bytes_dot"], dot_bytes), ("multiaxpy", kry["ms_axpy"], kry["bytes_axpy"], axpy_bytes)):
        legs[f"Gram-Schmidt: {nm}_kernel<dim> (timed at dim 14)"] = {
            "ms": ms14, "algorithmic_bytes": b14, "gbs": b14 / ms14 / 1e6, "frac": b14 / ms14 / 1e6 / peak,
            "launches_per_step": n_it, "share_of_step": (tot / (b14 / ms14)) / ms_per_step}
    ms_asm, by_asm = solver.time_assembly(reps=3)
    add("residual -R + diag(B): lat_kernel<LAT_R>" if mf else "assembly of B, C, -R (all colours)", ms_asm, by_asm, n_asm)
    return legs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="M4", choices=sorted(WORKLOADS))
    ap.add_argument("--degree", type=int, default=0, help="override the workload's element degree")
    ap.add_argument("--n-refines", dest="n_refines", type=int, default=0, help="override the workload's n_refines")
    ap.add_argument("--ic", default="", choices=["", "perfect", "dipole", "polycrystal"])
    ap.add_argument("--ortho", type=int, default=0)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--precond", type=int, default=-1,
                    help="0 none (reference algorithm), 1 block-Schur right preconditioner, 2 psi-only reduced solve "
                         "(Q1 matrix-free; elsewhere it acts as 1); default: 2 for Q1 matrix-free workloads, else 1")
    ap.add_argument("--assembled", action="store_true", help="assembled shared-pattern SELL-32 operator instead of matrix-free")
    ap.add_argument("--sweeps", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-legs", action="store_true")
    args = ap.parse_args()
    args.mf = not args.assembled
    if args.impl == "reference":
        return run_reference(args)

    # libraries (NCCL banner, torchrun notices) write to fd 1; the contract is ONE JSON line on stdout,
    # so everything else goes to stderr and the line is written to the saved descriptor at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # host-side OpenMP loops (stand-in, initial condition): share the cores between the ranks of the node
    os.environ["OMP_NUM_THREADS"] = str(max(1, host_cores() // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", world)))))

    import torch
    import pfc_b200 as P

    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N")
        args.gpus = world
    if world not in (1, 2, 4, 8):
        raise SystemExit("--gpus must be 1, 2, 4 or 8 (Morton chunks that are rectangles)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    wl = resolve_workload(args)
    degree, n_refines = wl["degree"], wl["n_refines"]
    if args.precond < 0:
        args.precond = 2 if (degree == 1 and args.mf) else 1
    S, p1, p2 = box(degree, n_refines)
    A0 = P.default_A0(PSI0, EPS)
    mesh = P.StandIn(degree, n_refines, p1, p2, rank=rank, n_ranks=world)
    n_dofs, N = mesh.n_dofs, mesh.N

    def make_solver(operator_kind, comm=True):
        s = P.Solver(degree=degree, dt=DT, theta=THETA, eps=EPS, device=local, orthogonalization=args.ortho,
                     use_cuda_graph=0 if args.no_graph else 1, preconditioner=args.precond, precond_sweeps=args.sweeps,
                     operator_kind=operator_kind)
        if world > 1 and comm:
            uid = [P.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            s.comm_init(uid[0], rank, world)
        return s

    solver = make_solver(1 if args.mf else 0)
    solver.setup(mesh)
    reduced_active = "reduced solve active" in solver.operator_info()[1]
    Psi0 = initial_condition(mesh, wl, p1, p2, S, A0)
    solver.set_state(Psi0, Psi0)
    n_own = solver.n_owned_nodes()
    own0 = int(mesh.node_offsets[rank])

    def pinned(n):
        return torch.empty(n, dtype=torch.float64).pin_memory().numpy()

    # ---- device-resident loop: W warm-up steps, then K timed steps (CUDA events inside pfc_step) ----
    for _ in range(args.warmup):
        solver.step()
    start_own = solver.get_state_owned(out_n=pinned(3 * n_own)).copy()  # the e2e loop replays the same K steps from here
    stats = []
    stress_buf = pinned(4 * n_own).reshape(4, n_own) if wl.get("stress") else None
    barrier()
    with ClockSampler(local) as clk:
        t_wall0 = time.perf_counter()
        ms_stress = []
        for _ in range(args.steps):
            stats.append(solver.step())
            if wl.get("stress"):  # calculate_stress after every time step (mpi.cpp:690-693); not part of the metric
                ts = time.perf_counter()
                _, cg_its = solver.stress(EPS, out=stress_buf, owned=True)
                ms_stress.append((time.perf_counter() - ts) * 1e3)
        barrier()
        t_wall = time.perf_counter() - t_wall0 - sum(ms_stress) * 1e-3
    ms_dev = sum(s["ms_total"] for s in stats)
    launches = sum(s["kernel_launches"] for s in stats)
    end_own = solver.get_state_owned().copy()

    # ---- config 3: stress projection after every step (mpi.cpp:690-697), timed beside the stepper ----
    stress_leg = None
    if args.mf:
        sbuf = pinned(4 * n_own).reshape(4, n_own)
        solver.stress(EPS, out=sbuf, owned=True)
        barrier()
        t0 = time.perf_counter()
        _, cg_its = solver.stress(EPS, out=sbuf, owned=True)
        barrier()
        stress_leg = {"ms_wall_incl_d2h": (time.perf_counter() - t0) * 1e3, "cg_iters": [int(i) for i in cg_its],
                      "d2h_bytes": int(sbuf.nbytes), "max_abs": float(np.abs(sbuf).max())}

    # ---- e2e: host buffers in, host buffers out, every step (state upload + step + state download) ----
    # Every rank moves what a rank of the MPI original holds: its owned ranges of the three blocks
    # (owned_partitioning, mpi.cpp:167-176; ghost values come over NVLink like update_ghost_values).
    # pinned ping-pong buffers: the state downloaded after step k is the state uploaded for step k + 1
    # (Psi_n_1 = Psi_n after a step, mpi.cpp:592)
    outs = [pinned(3 * n_own) for _ in range(2)]
    outs[0][:] = start_own
    solver.set_state_owned(outs[0], outs[0])
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        src, dst = outs[k & 1], outs[(k + 1) & 1]
        solver.set_state_owned(src, src)
        solver.step()
        solver.get_state_owned(out_n=dst)
    barrier()
    t_e2e = time.perf_counter() - t0
    replay = outs[args.steps & 1]
    replay_bits = bool(np.array_equal(replay, end_own))

    times = torch.tensor([ms_dev, t_wall * 1e3, t_e2e * 1e3], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_dev, ms_wall, ms_e2e = (float(x) for x in times.cpu())

    def owned_rel_l2(mine_own, ref_full):
        """relative L2 distance of this job's distributed state ([3][n_own] per rank) to a full reference state
        [n_dofs] known to every rank: (psi block, all three blocks)"""
        acc = np.zeros(4)
        for c in range(3):
            d = mine_own[c * n_own:(c + 1) * n_own] - ref_full[c * N + own0:c * N + own0 + n_own]
            r = ref_full[c * N + own0:c * N + own0 + n_own]
            acc[2] += float(d @ d)
            acc[3] += float(r @ r)
            if c == 0:
                acc[0], acc[1] = float(d @ d), float(r @ r)
        t = torch.tensor(acc, dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t)
        a = t.cpu().numpy()
        return math.sqrt(a[0] / a[1]), math.sqrt(a[2] / a[3])

    # ---- parity of the state that was timed -------------------------------------------------------------
    # the e2e loop replays the timed steps from the same state through host buffers: same result up to the solver
    # tolerance (the Chebyshev interval of the preconditioner is re-estimated every 16 solves, so not bit for bit)
    acc = torch.tensor([float((replay[:n_own] - end_own[:n_own]) @ (replay[:n_own] - end_own[:n_own])),
                        float(end_own[:n_own] @ end_own[:n_own])], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(acc)
    parity = {"e2e_replay_vs_timed_psi_rel_l2": math.sqrt(float(acc[0]) / float(acc[1])), "e2e_replay_bit_identical": replay_bits}
    if world > 1 and not args.no_parity:
        # the N-rank state after the W + K steps against a 1-rank run of the same mesh, initial condition and steps
        # on rank 0's GPU (no communicator): the only difference is the partition (halo exchange, reduction order)
        ref = torch.empty(n_dofs, dtype=torch.float64, device="cuda")
        t0 = time.perf_counter()
        if rank == 0:
            mesh1 = P.StandIn(degree, n_refines, p1, p2)
            s1 = make_solver(1 if args.mf else 0, comm=False)
            s1.setup(mesh1)
            Psi1 = initial_condition(mesh1, wl, p1, p2, S, A0)
            s1.set_state(Psi1, Psi1)
            its1 = [s1.step()["gmres_iters_total"] for _ in range(args.warmup + args.steps)]
            ref.copy_(torch.from_numpy(s1.get_state()))
            s1.close()
            mesh1.close()
        dist.broadcast(ref, src=0)
        e_psi, e_all = owned_rel_l2(end_own, ref.cpu().numpy())
        del ref
        parity.update({"against": f"1-rank run of the same mesh and steps on rank 0's GPU ({args.warmup + args.steps} steps)",
                       "psi_rel_l2": e_psi, "all_blocks_rel_l2": e_all, "bar": 1e-9 * (args.warmup + args.steps),
                       "bar_source": "north_star: <= 1e-9 per step", "seconds": time.perf_counter() - t0})
        if rank == 0:
            parity["gmres_iters_1rank_timed_steps"] = its1[args.warmup:]
        parity["ok"] = bool(e_psi <= parity["bar"])

    if world == 1 and reduced_active and not args.no_parity:
        # the timed (reduced) solve against the 3-block solve with the block-Schur preconditioner: same mesh, initial
        # condition and steps; both solve every Newton system to 1e-10 on the residual of the 3-block system
        t0 = time.perf_counter()
        s3 = P.Solver(degree=degree, dt=DT, theta=THETA, eps=EPS, device=local, orthogonalization=args.ortho,
                      use_cuda_graph=0 if args.no_graph else 1, preconditioner=1, operator_kind=1)
        s3.setup(mesh)
        s3.set_state(Psi0, Psi0)
        its3 = [s3.step()["gmres_iters_total"] for _ in range(args.warmup + args.steps)]
        ref3 = s3.get_state_owned()
        s3.close()
        d = end_own - ref3
        parity["reduced_vs_3block_solve"] = {
            "psi_rel_l2": float(np.linalg.norm(d[:n_own]) / np.linalg.norm(ref3[:n_own])),
            "all_blocks_rel_l2": float(np.linalg.norm(d) / np.linalg.norm(ref3)), "steps": args.warmup + args.steps,
            "bar": 1e-9 * (args.warmup + args.steps), "gmres_iters_3block_timed_steps": its3[args.warmup:],
            "seconds": time.perf_counter() - t0}
        parity["reduced_vs_3block_solve"]["ok"] = bool(parity["reduced_vs_3block_solve"]["psi_rel_l2"] <= parity["reduced_vs_3block_solve"]["bar"])

    # ---- config 5: the other operator path on the same run ----------------------------------------------
    other = None
    if wl.get("both_operators"):
        s2 = make_solver(0 if args.mf else 1)
        s2.setup(mesh)
        s2.set_state(Psi0, Psi0)
        for _ in range(args.warmup):
            s2.step()
        barrier()
        st2 = [s2.step() for _ in range(args.steps)]
        barrier()
        t2 = torch.tensor([sum(x["ms_total"] for x in st2)], dtype=torch.float64, device="cuda")
        full = torch.zeros(n_dofs, dtype=torch.float64, device="cuda")
        full.copy_(torch.from_numpy(s2.get_state()))  # owned entries, zeros elsewhere
        if dist is not None:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
            dist.all_reduce(full)
        e_psi, e_all = owned_rel_l2(end_own, full.cpu().numpy())
        del full
        other = {"operator": "assembled (shared-pattern SELL-32)" if args.mf else "matrix-free lattice kernels",
                 "value": n_dofs * args.steps / (float(t2.item()) * 1e-3), "ms_per_step": float(t2.item()) / args.steps,
                 "gmres_iters_per_step": [x["gmres_iters_total"] for x in st2],
                 "psi_rel_l2_between_paths": e_psi, "all_blocks_rel_l2_between_paths": e_all, "bar": 1e-7,
                 "steps_compared": args.warmup + args.steps}
        s2.close()

    # ---- roofline legs on resident data (rank-local kernels) ----
    peak, peak_src = measured_peak()
    ms_per_step = ms_dev / args.steps
    sweeps = args.sweeps if args.sweeps > 0 else ((6 if degree == 2 else 4) if args.mf else (5 if degree == 2 else 3))  # the library's defaults
    if reduced_active and args.sweeps <= 0:
        sweeps = 2  # the library's default of the reduced solve
    legs = None
    if not args.no_legs:
        solver.assemble()
        legs = kernel_legs(solver, stats, sweeps, args.precond >= 1, args.mf, peak, n_own, ms_per_step)
    comm_legs = solver.time_comm(reps=50) if world > 1 else None

    if rank == 0:
        its = [s["gmres_iters_total"] for s in stats]
        line = {
            "metric": "DoF-timesteps/s", "value": n_dofs * args.steps / (ms_dev * 1e-3), "unit": "DoF-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_string(wl, n_dofs), "n_dofs": int(n_dofs), "baseline_config": wl["config"],
                       "partition": f"{world} contiguous chunk(s) of the Morton curve (p4est), {n_own} owned nodes per component on rank 0",
                       "solver": ("GMRES(28) unpreconditioned, tol 1e-10 abs, Newton tol 1e-8 (reference algorithm)" if args.precond == 0 else
                                  (f"psi-only Schur-reduced Newton system (serial.cpp:319-369) with dpsi = M^2 y instead of inner mass solves: "
                                   f"GMRES(28) on vectors of N doubles, right preconditioner = {sweeps} Jacobi-Chebyshev steps on the "
                                   "constant-coefficient Schur stand-in, same tol 1e-10 abs on the residual of the 3-block system, Newton tol 1e-8"
                                   if reduced_active else
                                   f"GMRES(28), right block-Schur preconditioner (lumped mass + {sweeps} Jacobi-Chebyshev sweeps), same tol 1e-10 abs on the true residual, Newton tol 1e-8")),
                       "operator": ("matrix-free, sum-factorised lattice cell kernels (no stored matrix)" if args.mf
                                    else "assembled (shared-pattern SELL-32)"),
                       "orthogonalization": ["CGS+Kelley", "CGS2"][args.ortho],
                       "comm": ["single rank", "NCCL", "peer-memory kernels over NVLink (CUDA IPC) + NCCL bootstrap"][solver.comm_mode()],
                       "cuda_graph": not args.no_graph,
                       "l2": ("timed steps: every GMRES iteration streams the Krylov basis, "
                              f"{(29 * (8.0 if reduced_active else 24.0) * n_own) / 1e6:.0f} MB per rank against the 126 MB L2, no flush; "
                              "kernel legs: operands rotated through the basis vectors"),
                       "newton_assemblies_per_step": [s["n_assemblies"] for s in stats], "gmres_iters_per_step": its},
            "ms_wall_per_step": ms_wall / args.steps,
            "e2e": {"value": n_dofs * args.steps / (ms_e2e * 1e-3), "unit": "DoF-steps/s",
                    "h2d_bytes_per_step": int(n_dofs * 8), "d2h_bytes_per_step": int(n_dofs * 8),
                    "what": "per step and rank: owned ranges of Psi_n uploaded from pinned host memory, pfc_step, owned ranges downloaded"},
            "gpu_launches": int(launches),
            "parity": parity,
            "clocks": clk.summary(),
        }
        if legs:
            timed = {k: v for k, v in legs.items() if "frac" in v}
            top = max(timed, key=lambda k: timed[k]["share_of_step"])
            t = timed[top]
            line["roofline"] = {"bound": "hbm", "kernel": top, "achieved": t["gbs"], "peak": peak, "unit": "GB/s",
                                "frac": t["frac"], "traffic": None, "peak_source": peak_src, "ms_per_launch": t["ms"],
                                "algorithmic_bytes_per_launch": t["algorithmic_bytes"], "share_of_step": t["share_of_step"],
                                "chosen": "the kernel leg with the largest share of the timed step"}
            line["kernels"] = legs
            accounted = sum(v["share_of_step"] for k, v in legs.items() if "all its kernels" not in k)
            line["step"] = {"ms": ms_per_step, "share_accounted_by_legs": accounted,
                            "algorithmic_gbs": sum(v["gbs"] * v["share_of_step"] for v in timed.values()) / max(accounted, 1e-30),
                            "note": "share-weighted mean of the legs' algorithmic GB/s; the rest of the step is small "
                                    "vector kernels, device-side control and (N > 1) halo exchanges"}
            line["step"]["frac"] = line["step"]["algorithmic_gbs"] / peak
            try:  # dram traffic of the roofline kernel from the committed ncu capture of this workload
                with open(os.path.join(ROOT, "profiles", "r04_traffic.json")) as f:
                    tr = json.load(f)
                ent = tr.get(wl["name"] + ("/reduced" if reduced_active else ""), {}).get(top)
                if ent and world == 1 and not wl["custom"]:
                    line["roofline"]["traffic"] = ent["bytes_per_launch"]
                    line["roofline"]["traffic_source"] = ent["source"]
            except (OSError, KeyError, ValueError):
                pass
        if comm_legs is not None:
            line["comm_legs_us"] = comm_legs
        if stress_leg is not None:
            line["stress_projection"] = stress_leg
            if wl.get("stress"):
                line["stress_projection"]["ms_per_step_in_loop"] = float(np.mean(ms_stress)) if ms_stress else None
        if other is not None:
            line["config5_other_path"] = other
        if world == 1 and not args.no_cpu_baseline:
            cores = host_cores()
            O = _oracle(cores)
            c = cpu_model(O, wl, cores, args.steps, args.warmup, sample_iters=28)
            line["cpu_baseline"] = {"value": c["value"], "unit": "DoF-steps/s", "cores": cores, "kind": "port", "sample": c["sample"],
                                    "measured_s": c["measured_s"], "modelled_s": c["modelled_s"],
                                    "t_assembly_s": c["t_assembly_s"], "t_gmres_iter_s": c["t_gmres_iter_s"]}
            if not args.no_parity:
                # full-size parity against the oracle on the timed mesh: -R of the initial condition and J x (random x)
                solver.set_state(c["Psi"], c["Psi"])
                solver.assemble()
                rhs = solver.get_rhs()
                x = np.random.default_rng(3).standard_normal(n_dofs)
                y = solver.apply_operator(x)
                yo = c["om"].spmv(c["vals"], x)
                cons = c["om"].master >= 0
                parity.update({"against": "oracle at full size on the timed mesh (assembly of the initial condition, monolithic CSR product)",
                               "rhs_max_rel": float(np.abs(rhs - c["rhs"]).max() / np.abs(c["rhs"]).max()),
                               "Jx_max_rel": float(np.abs(y - yo)[~cons].max() / np.abs(yo).max()), "bar": 1e-11,
                               "bar_note": "max-norm relative to max |R| resp. max |J x|; -R of the initial condition is a difference of O(1) terms"})
                parity["ok"] = bool(parity["rhs_max_rel"] <= 1e-11 and parity["Jx_max_rel"] <= 1e-11)
                if c["real_step"]:  # M2: time step 1 of the oracle at full size exists -> psi after one step
                    Pn = c["real_step"]["Psi_after"]
                    solver.step()
                    got = solver.get_state()
                    parity["psi_rel_l2_after_step_1"] = float(np.linalg.norm(got[:N] - Pn[:N]) / np.linalg.norm(Pn[:N]))
                    parity["ok"] = bool(parity["ok"] and parity["psi_rel_l2_after_step_1"] <= 1e-9)
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    solver.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
