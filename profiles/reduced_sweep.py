"""Tuning aid: the reduced solve (PFC_PRECOND_SCHUR_REDUCED) against the block-Schur solve on a Q1 bench mesh, Chebyshev
degree by degree.  Usage: python profiles/reduced_sweep.py --n-refines 12 --precond 2 --sweeps 1 2 3"""
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
    ap.add_argument("--n-refines", dest="n_refines", type=int, default=12)
    ap.add_argument("--precond", type=int, default=2)
    ap.add_argument("--sweeps", type=int, nargs="*", default=[2])
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--ic", default="polycrystal")
    a = ap.parse_args()
    S, p1, p2 = B.box(1, a.n_refines)
    mesh = P.StandIn(1, a.n_refines, p1, p2)
    wl = dict(B.WORKLOADS["M4"], ic=a.ic)
    Psi0 = B.initial_condition(mesh, wl, p1, p2, S, P.default_A0(B.PSI0, B.EPS))
    for sw in a.sweeps:
        s = P.Solver(degree=1, dt=B.DT, theta=B.THETA, eps=B.EPS, preconditioner=a.precond, precond_sweeps=sw, operator_kind=1)
        s.setup(mesh)
        s.set_state(Psi0, Psi0)
        first = s.step()
        s.step()
        st = [s.step() for _ in range(a.steps)]
        print(json.dumps({"n_refines": a.n_refines, "precond": a.precond, "sweeps": sw,
                          "ms_per_step": round(sum(x["ms_total"] for x in st) / len(st), 2),
                          "ms_first_step": round(first["ms_total"], 2), "gmres_first": first["gmres_iters"],
                          "gmres_iters": [x["gmres_iters"] for x in st], "ms_solver": round(st[-1]["ms_solver"], 2),
                          "ms_assembly": round(st[-1]["ms_assembly"], 2), "launches": st[-1]["kernel_launches"]}), flush=True)
        s.close()


if __name__ == "__main__":
    main()
