"""Per-kernel legs of one GMRES iteration on the bench workload (M2), timed back to back on resident data
with CUDA events on the handle's stream (no profiler).  Usage: python profiles/legs.py [--degree 2] [--n-refines 9]
[--mf 1] [--sweeps 2 3 4]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402
import pfc_b200 as P  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--degree", type=int, default=2)
    ap.add_argument("--n-refines", dest="n_refines", type=int, default=9)
    ap.add_argument("--mf", type=int, default=1)
    ap.add_argument("--sweeps", type=int, nargs="*", default=[4])
    ap.add_argument("--steps", type=int, default=3)
    a = ap.parse_args()
    _, p1, p2 = B.box(a.degree, a.n_refines)
    mesh = P.StandIn(a.degree, a.n_refines, p1, p2)
    Psi0 = mesh.hexagonal_lattice(P.default_A0(B.PSI0, B.EPS), B.PSI0)
    for sw in a.sweeps:
        s = P.Solver(degree=a.degree, dt=B.DT, theta=B.THETA, eps=B.EPS, preconditioner=1, precond_sweeps=sw,
                     operator_kind=a.mf)
        s.setup(mesh)
        s.set_state(Psi0, Psi0)
        for _ in range(2):
            s.step()
        st = [s.step() for _ in range(a.steps)]
        s.assemble()
        ms_op, by_op = s.time_operator(reps=50)
        pc = s.time_precond(reps=50)
        kr = s.time_krylov(dim=10, reps=20)
        ms_asm, by_asm = s.time_assembly(reps=10)
        out = {"sweeps": sw, "ms_per_step": sum(x["ms_total"] for x in st) / len(st),
               "gmres_iters": [x["gmres_iters_total"] for x in st], "launches": st[-1]["kernel_launches"],
               "ms_solver": st[-1]["ms_solver"], "ms_assembly": st[-1]["ms_assembly"],
               "us_operator": 1e3 * ms_op, "gbs_operator": by_op / ms_op / 1e6,
               "us_precond_apply": 1e3 * pc["ms_apply"],
               "us_precond_legs [sweep | rhs | t | back] (assembled: [K product | Schur row])": [1e3 * x for x in pc["ms"][1:]],
               "gbs_precond_legs": [b / m / 1e6 if m > 0 else None for m, b in zip(pc["ms"][1:], pc["bytes"])],
               "us_multidot10": 1e3 * kr["ms_dot"], "us_multiaxpy10": 1e3 * kr["ms_axpy"],
               "us_assembly": 1e3 * ms_asm}
        print(json.dumps(out), flush=True)
        s.close()


if __name__ == "__main__":
    main()
