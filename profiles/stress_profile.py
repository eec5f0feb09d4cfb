"""One stress projection on the M2 mesh (Q2, n_refines 9) for an ncu launch list:
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:'stress|lat_mass|lat_kernel' --csv --log-file gpurun_out/stress_launches.csv python profiles/stress_profile.py"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pfc_b200 as P  # noqa: E402

a = P.LATTICE_A
S = 16
mesh = P.StandIn(2, 9, (-6 * a * S, -3 * np.sqrt(3.0) * a * S), (6 * a * S, 3 * np.sqrt(3.0) * a * S))
s = P.Solver(degree=2, operator_kind=1, preconditioner=1)
s.setup(mesh)
Psi = mesh.hexagonal_lattice(P.default_A0(-0.43, -0.8), -0.43)
s.set_state(Psi, Psi)  # the initial condition: no lat_kernel launches but the stress right-hand side
out = np.zeros((4, mesh.N))
s.stress(-0.8, out=out)
t0 = time.perf_counter()
_, its = s.stress(-0.8, out=out)
print(f"stress: {1e3 * (time.perf_counter() - t0):.2f} ms wall, CG iterations {its}")
s.close()
