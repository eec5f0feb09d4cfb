"""Generates tests/golden/*.npz from the CPU oracle (the reference cannot run here; see oracle/pfc_oracle.h:
PARITY UNPINNED).  These fixtures pin regressions of the oracle and travel to the GPU box, where the CUDA path
is compared with them without the oracle being re-run.   python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import oracle as O  # noqa: E402

A = O.LATTICE_A


def run(degree, n_refines, steps, disl, name):
    om = O.Mesh(degree, n_refines, (-6 * A, -5 * A), (6 * A, 5 * A))
    pos, bur = ([[2 * A, 0], [-2 * A, 0]], [[A, 0], [-A, 0]]) if disl else ((), ())
    Pn = om.initial_condition(O.default_A0(-0.43, -0.8), -0.43, pos, bur)
    Pn1 = Pn.copy()
    its, res = [], []
    for _ in range(steps):
        r = om.newton_step(Pn, Pn1, 0.1, 1.0, -0.8)
        its.append(r["gmres_iters"])
        res.append(r["residuals"])
    np.savez_compressed(os.path.join(HERE, name), psi=Pn[:om.N], chi=Pn[om.N:2 * om.N], phi=Pn[2 * om.N:],
                        gmres_iters=np.array(its[-1]), residuals=np.array(res[-1]),
                        all_gmres_iters=np.array([sum(i) for i in its]), steps=steps)
    print(name, its)


def run_stress(degree, n_refines, name_in, name_out):
    """stress projection (stress_calculator_mpi.cpp:345-353) of a committed state, as the reference computes it after
    every timestep (mpi.cpp:690-693)"""
    om = O.Mesh(degree, n_refines, (-6 * A, -5 * A), (6 * A, 5 * A))
    g = np.load(os.path.join(HERE, name_in))
    r = om.stress(np.concatenate([g["psi"], g["chi"], g["phi"]]), -0.8)
    assert r["status"] == 0
    np.savez_compressed(os.path.join(HERE, name_out), stress=r["stress"], cg_iters=np.array(r["cg_iters"]))
    print(name_out, r["cg_iters"])


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "stress":  # only the stress fixtures (the state fixtures stay as committed)
        run_stress(1, 6, "config1_step1.npz", "config1_step1_stress.npz")
        run_stress(2, 4, "q2_r4_step3.npz", "q2_r4_step3_stress.npz")
        sys.exit(0)
    run(1, 6, 1, True, "config1_step1.npz")      # shipped TOML, first timestep
    run(1, 6, 10, True, "config1_step10.npz")    # shipped TOML, all 10 timesteps (n_timesteps = 10)
    run(2, 4, 3, True, "q2_r4_step3.npz")
