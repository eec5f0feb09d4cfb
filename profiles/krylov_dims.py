"""Gram-Schmidt kernels at several Krylov dimensions on the bench mesh (M4 by default): GB/s of multidot / multiaxpy.
Usage: python profiles/krylov_dims.py [--degree 1] [--n-refines 12]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402
import pfc_b200 as P  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--degree", type=int, default=1)
ap.add_argument("--n-refines", dest="n_refines", type=int, default=12)
a = ap.parse_args()
_, p1, p2 = B.box(a.degree, a.n_refines)
mesh = P.StandIn(a.degree, a.n_refines, p1, p2)
Psi0 = mesh.hexagonal_lattice(P.default_A0(B.PSI0, B.EPS), B.PSI0)
s = P.Solver(degree=a.degree, dt=B.DT, theta=B.THETA, eps=B.EPS, preconditioner=1, operator_kind=1)
s.setup(mesh)
s.set_state(Psi0, Psi0)
if not os.environ.get("SKIP_STEP"):
    s.step()
else:
    s.assemble()
for dim in [int(x) for x in os.environ.get("DIMS", "4 8 12 14 16 20 24 27").split()]:
    k = s.time_krylov(dim=dim, reps=10)
    print(json.dumps({"dim": dim, "us_dot": 1e3 * k["ms_dot"], "gbs_dot": k["bytes_dot"] / k["ms_dot"] / 1e6,
                      "us_axpy": 1e3 * k["ms_axpy"], "gbs_axpy": k["bytes_axpy"] / k["ms_axpy"] / 1e6}), flush=True)
s.close()
