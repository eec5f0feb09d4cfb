"""Tuning aid: variants of the Q1 J apply kernel (PFC_J1_VARIANT, read at pfc_setup; csrc/pfc_lattice.cu) in ONE process
per mesh size: self-check against the general cell kernel (pfc_selfcheck_q1), then the operator time with operands
rotating through the Krylov basis (pfc_time_operator).  n_refines 11 / 10 stand for the subdomain of a rank at 4 / 16
ranks of M4.  Usage: python profiles/j1p_sweep.py --n-refines 12 --variants 6 20 22"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402
import pfc_b200 as P  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-refines", dest="n_refines", type=int, default=12)
    ap.add_argument("--variants", type=int, nargs="*", default=[6, 20])
    ap.add_argument("--reps", type=int, default=50)
    a = ap.parse_args()
    _, p1, p2 = B.box(1, a.n_refines)
    t0 = time.time()
    mesh = P.StandIn(1, a.n_refines, p1, p2)
    Psi0 = mesh.hexagonal_lattice(P.default_A0(B.PSI0, B.EPS), B.PSI0)
    print(json.dumps({"n_refines": a.n_refines, "s_mesh": round(time.time() - t0, 2)}), flush=True)
    for v in a.variants:
        os.environ["PFC_J1_VARIANT"] = str(v)
        t0 = time.time()
        s = P.Solver(degree=1, dt=B.DT, theta=B.THETA, eps=B.EPS, preconditioner=1, operator_kind=1)
        s.setup(mesh)
        s.set_state(Psi0, Psi0)
        s.assemble()
        ej, _ = s.selfcheck_q1()
        s.assemble()
        ms, by = s.time_operator(reps=a.reps)
        ms2, _ = s.time_operator(reps=a.reps)
        print(json.dumps({"n_refines": a.n_refines, "variant": v, "selfcheck_J": ej, "us_operator": round(1e3 * min(ms, ms2), 2),
                          "gbs": round(by / min(ms, ms2) / 1e6, 1), "s_wall": round(time.time() - t0, 2)}), flush=True)
        s.close()


if __name__ == "__main__":
    main()
