#!/usr/bin/env python
"""bench.py — DoF-timesteps/s of the PFC Newton-Krylov hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W          # our arm (CUDA, through the C ABI)
  python bench.py --impl reference --steps K --warmup W  # CPU arm: the oracle port on the host cores

A "step" is one iterate_timestep (mpi.cpp:576-593): Newton loop with an assembly and an
unpreconditioned GMRES(28) solve per iteration.  Workload ("M2"): Q2 elements,
n_refines 9 per GPU (3 151 875 DoF at N = 1), perfect hexagonal lattice on a commensurate box with
the node spacing of the shipped parameter file (box scale S = degree * 2^(n_refines - 6)).
N > 1 is weak scaling: the mesh grows with N (rep_x x rep_y trees of 2^n x 2^n cells, box and
lattice grow with it, h fixed), cut along the Morton curve into N chunks.
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


def workload(degree, n_refines, n_gpus):
    """box and tree repetitions for N ranks; h is that of the shipped file at every size."""
    rep = {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2)}.get(n_gpus)
    if rep is None:
        raise SystemExit("--gpus must be 1, 2, 4 or 8")
    S = degree * 2 ** (n_refines - 6)
    # commensurate box: width 12 a S, height 6 sqrt(3) a S = 6 S * 4 pi   (SURVEY.md §7 hard part 7)
    w, h = 6.0 * A * S * rep[0], 3.0 * math.sqrt(3.0) * A * S * rep[1]
    return rep, (-w, -h), (w, h)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self._stop = index, [], threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
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


def cpu_sample(degree, n_refines, counts, n_threads, sample_iters=56):
    """Oracle timed on a bounded sample of the same workload: one assembly + `sample_iters`
    GMRES(28) iterations on the full-size problem, scaled to a timestep with the iteration
    counts `counts` = (assemblies, gmres iterations) per step."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    O.set_threads(n_threads)
    _, p1, p2 = workload(degree, n_refines, 1)
    om = O.Mesh(degree, n_refines, p1, p2)
    Psi = om.initial_condition(O.default_A0(PSI0, EPS), PSI0)
    t0 = time.perf_counter()
    vals, rhs = om.assemble(Psi, Psi, DT, THETA, EPS, threads=n_threads)
    t_asm = time.perf_counter() - t0
    om.gmres(vals, rhs, stop_after=2)  # touch
    t0 = time.perf_counter()
    _, its, _, _ = om.gmres(vals, rhs, stop_after=sample_iters)
    t_it = (time.perf_counter() - t0) / max(its, 1)
    n_asm, n_its = counts
    t_step = n_asm * t_asm + n_its * t_it
    return {"n_dofs": int(om.n_dofs), "t_assembly_s": t_asm, "t_gmres_iter_s": t_it, "t_step_s": t_step,
            "value": om.n_dofs / t_step, "sample_iters": int(its)}


def replica_counts(degree, steps, warmup):
    """Newton / GMRES counts per timestep from a full oracle run of the same physics on a small
    replica at the same h (iteration counts do not depend on the domain size at fixed h)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    nr = 6 if degree == 1 else 5
    _, p1, p2 = workload(degree, nr, 1)
    om = O.Mesh(degree, nr, p1, p2)
    Pn = om.initial_condition(O.default_A0(PSI0, EPS), PSI0)
    Pn1 = Pn.copy()
    out = []
    for _ in range(warmup + steps):
        r = om.newton_step(Pn, Pn1, DT, THETA, EPS, threads=O.max_threads())
        out.append((r["n_assemblies"], sum(r["gmres_iters"])))
    return out[warmup:]


def run_reference(args):
    """--impl reference: the reference's algorithm on the host cores.  The reference binary needs
    deal.II + Trilinos + p4est + MPI (absent) so this is the oracle port, all threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    cores = O.max_threads()
    counts = replica_counts(args.degree, args.steps, args.warmup)
    per_step = []
    n_dofs = None
    t_begin = time.perf_counter()
    for k, c in enumerate(counts):
        s = cpu_sample(args.degree, args.n_refines, c, cores, sample_iters=28)
        n_dofs = s["n_dofs"]
        per_step.append(s["t_step_s"])
        if time.perf_counter() - t_begin > 240 and k + 1 < len(counts):  # keep the run bounded
            per_step += [s["t_step_s"] * cc[1] / max(c[1], 1) for cc in counts[k + 1:]]
            break
    t = sum(per_step)
    value = n_dofs * len(per_step) / t
    sample = ("per step: 1 assembly + 28 GMRES(28) iterations (MGS) of the full-size problem timed, scaled by the "
              "step's Newton/GMRES counts taken from a full oracle run on a same-h replica "
              f"(Q{args.degree}, n_refines {6 if args.degree == 1 else 5})")
    line = {"impl": "reference", "metric": "DoF-timesteps/s", "value": value, "unit": "DoF-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / len(per_step),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"M2: Q{args.degree}, n_refines {args.n_refines}, {n_dofs} DoF, perfect hexagonal lattice, "
                                   "commensurate box at the shipped file's h (CPU arm always runs the N=1 mesh)"},
            "cpu_baseline": {"value": value, "unit": "DoF-steps/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "DoF-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--degree", type=int, default=2)
    ap.add_argument("--n-refines", dest="n_refines", type=int, default=9)
    ap.add_argument("--ortho", type=int, default=0)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--precond", type=int, default=1)
    ap.add_argument("--mf", action="store_true", help="(default) matrix-free operator (PFC_OPERATOR_MATRIX_FREE)")
    ap.add_argument("--assembled", action="store_true", help="assembled shared-pattern SELL-32 operator instead")
    ap.add_argument("--sweeps", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.mf = not args.assembled
    if args.impl == "reference":
        return run_reference(args)

    # libraries (NCCL banner, torchrun notices) write to fd 1; the contract is ONE JSON line on stdout,
    # so everything else goes to stderr and the line is written to the saved descriptor at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import pfc_b200 as P

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N")
        args.gpus = world
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

    rep, p1, p2 = workload(args.degree, args.n_refines, world)
    mesh = P.StandIn(args.degree, args.n_refines, p1, p2, rank=rank, n_ranks=world, rep=rep)
    solver = P.Solver(degree=args.degree, dt=DT, theta=THETA, eps=EPS, device=local, orthogonalization=args.ortho,
                      use_cuda_graph=0 if args.no_graph else 1, preconditioner=args.precond, precond_sweeps=args.sweeps,
                      operator_kind=1 if args.mf else 0)
    if world > 1:
        uid = [P.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        solver.comm_init(uid[0], rank, world)
    solver.setup(mesh)
    n_dofs = mesh.n_dofs
    Psi0 = mesh.hexagonal_lattice(P.default_A0(PSI0, EPS), PSI0)  # perfect lattice (dislocation_positions = [])
    pin = torch.empty((2, n_dofs), dtype=torch.float64).pin_memory()
    pin_np = pin.numpy()
    pin_np[0] = Psi0
    pin_np[1] = Psi0
    solver.set_state(pin_np[0], pin_np[1])

    # ---- device-resident loop: W warm-up steps, then K timed steps (CUDA events inside pfc_step) ----
    for _ in range(args.warmup):
        solver.step()
    start_state = solver.get_state()  # the e2e loop below replays the same K steps from here
    stats = []
    barrier()
    with ClockSampler(local) as clk:
        t_wall0 = time.perf_counter()
        for _ in range(args.steps):
            stats.append(solver.step())
        barrier()
        t_wall = time.perf_counter() - t_wall0
    ms_dev = sum(s["ms_total"] for s in stats)
    launches = sum(s["kernel_launches"] for s in stats)
    # ---- e2e: host buffers in, host buffers out, every step (state upload + step + state download) ----
    # Every rank moves what a rank of the MPI original holds: its owned ranges of the three blocks
    # (owned_partitioning, mpi.cpp:167-176; ghost values come over NVLink like update_ghost_values).
    # pinned ping-pong buffers: the state downloaded after step k is the state uploaded for step k + 1
    # (Psi_n_1 = Psi_n after a step, mpi.cpp:592)
    n_own3 = 3 * solver.n_owned_nodes()
    outs = [torch.empty(n_own3, dtype=torch.float64).pin_memory().numpy() for _ in range(2)]
    solver.set_state(start_state, start_state)
    solver.get_state_owned(out_n=outs[0])
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        src, dst = outs[k & 1], outs[(k + 1) & 1]
        solver.set_state_owned(src, src)
        st = solver.step()
        solver.get_state_owned(out_n=dst)
    barrier()
    t_e2e = time.perf_counter() - t0

    times = torch.tensor([ms_dev, t_wall * 1e3, t_e2e * 1e3], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_dev, ms_wall, ms_e2e = (float(x) for x in times.cpu())

    # ---- roofline legs on resident data (rank-local kernels) ----
    solver.assemble()
    ms_op, bytes_op = solver.time_operator(reps=20)
    kry = solver.time_krylov(dim=14, reps=10)
    ms_asm, bytes_asm = solver.time_assembly(reps=3)
    comm_legs = solver.time_comm(reps=50) if world > 1 else None
    stress_leg = None
    if args.mf:  # SURVEY.md 8f-1 (outside the timed region): stress projection of the final state, result to the host
        sbuf = torch.empty((4, solver.n_owned_nodes()), dtype=torch.float64).pin_memory().numpy()
        solver.stress(EPS, out=sbuf, owned=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _, cg_its = solver.stress(EPS, out=sbuf, owned=True)
        stress_leg = {"ms_wall_incl_d2h": (time.perf_counter() - t0) * 1e3, "cg_iters": [int(i) for i in cg_its],
                      "d2h_bytes": int(sbuf.nbytes)}
    peak, peak_src = measured_peak()
    N = mesh.N
    nnz_csr = None
    if rank == 0:
        its = [s["gmres_iters_total"] for s in stats]
        line = {
            "metric": "DoF-timesteps/s", "value": n_dofs * args.steps / (ms_dev * 1e-3), "unit": "DoF-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"M2: Q{args.degree}, n_refines {args.n_refines} per GPU, trees {rep[0]}x{rep[1]}, "
                                   f"{n_dofs} DoF total, perfect hexagonal lattice, commensurate box at the shipped file's h",
                       "n_dofs": int(n_dofs), "solver": ("GMRES(28) unpreconditioned, tol 1e-10 abs, Newton tol 1e-8 (reference algorithm)" if args.precond == 0 else
                                  "GMRES(28), right block-Schur preconditioner (lumped mass + Jacobi-Chebyshev), same tol 1e-10 abs on the true residual, Newton tol 1e-8"),
                       "operator": ("matrix-free, sum-factorised lattice cell kernels (no stored matrix)" if args.mf
                                    else "assembled (shared-pattern SELL-32)"),
                       "orthogonalization": ["CGS+Kelley", "CGS2"][args.ortho],
                       "comm": ["single rank", "NCCL", "peer-memory kernels over NVLink (CUDA IPC) + NCCL bootstrap"][solver.comm_mode()], "cuda_graph": not args.no_graph,
                       "l2": "working set (matrix + Krylov basis) exceeds the 126 MB L2; no flush needed",
                       "newton_assemblies_per_step": [s["n_assemblies"] for s in stats], "gmres_iters_per_step": its},
            "ms_wall_per_step": ms_wall / args.steps,
            "e2e": {"value": n_dofs * args.steps / (ms_e2e * 1e-3), "unit": "DoF-steps/s",
                    "h2d_bytes_per_step": int(n_dofs * 8), "d2h_bytes_per_step": int(n_dofs * 8)},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": ("lat_kernel<LAT_J> (y = J x matrix-free on the tensor lattice)" if args.mf else
                                    "spmv_sell_kernel (y = J x, shared-pattern SELL-32 block layout)"),
                         "achieved": bytes_op / (ms_op * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": bytes_op / (ms_op * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                         "ms_per_launch": ms_op, "algorithmic_bytes_per_launch": bytes_op,
                         "csr_equivalent_gbs": None},
            "kernels": {
                "multidot<14>": {"ms": kry["ms_dot"], "gbs": kry["bytes_dot"] / kry["ms_dot"] / 1e6,
                                 "frac": kry["bytes_dot"] / kry["ms_dot"] / 1e6 / peak},
                "multiaxpy<14>": {"ms": kry["ms_axpy"], "gbs": kry["bytes_axpy"] / kry["ms_axpy"] / 1e6,
                                  "frac": kry["bytes_axpy"] / kry["ms_axpy"] / 1e6 / peak},
                "assembly (all colours)": {"ms": ms_asm, "gbs": bytes_asm / ms_asm / 1e6,
                                           "frac": bytes_asm / ms_asm / 1e6 / peak}},
            "clocks": clk.summary(),
        }
        # dram traffic of the roofline kernel from the committed ncu capture of the same workload (None for other sizes)
        try:
            with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
                tr = json.load(f)["lat_kernel<LAT_J>" if args.mf else "spmv_sell_kernel"]
            if args.degree == 2 and args.n_refines == 9 and world == 1:
                line["roofline"]["traffic"] = tr["bytes_per_launch"]
                line["roofline"]["traffic_source"] = tr["source"]
        except (OSError, KeyError, ValueError):
            pass
        if comm_legs is not None:
            line["comm_legs_us"] = comm_legs
        if stress_leg is not None:
            line["stress_projection"] = stress_leg
        # the reference's monolithic CSR would move 12 nnz + 12 n_rows + 8 n_cols bytes for the same product
        if not args.mf:
            nnz_mono = 7 * (bytes_op - 48.0 * N) / 36.0  # approx: 7 blocks per stored scalar entry
            line["roofline"]["csr_equivalent_gbs"] = (12.0 * nnz_mono + 60.0 * N) / (ms_op * 1e-3) / 1e9
        if world == 1 and not args.no_cpu_baseline:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import oracle as O
            cores = O.max_threads()
            # the CPU arm runs the reference's algorithm (unpreconditioned GMRES(28)): its per-step counts come from
            # a full oracle run of the same physics on a same-h replica
            rc = replica_counts(args.degree, args.steps, args.warmup)
            mean_counts = (float(np.mean([x[0] for x in rc])), float(np.mean([x[1] for x in rc])))
            c = cpu_sample(args.degree, args.n_refines, mean_counts, cores, sample_iters=56)
            line["cpu_baseline"] = {"value": c["value"], "unit": "DoF-steps/s", "cores": cores, "kind": "port",
                                    "sample": "1 assembly + 56 GMRES(28) iterations (MGS, OpenMP) of this workload timed on the "
                                              "host, scaled by the reference algorithm's per-step counts on a same-h replica "
                                              f"({mean_counts[0]:.1f} assemblies, {mean_counts[1]:.0f} GMRES iterations)",
                                    "t_assembly_s": c["t_assembly_s"], "t_gmres_iter_s": c["t_gmres_iter_s"]}
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    solver.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
