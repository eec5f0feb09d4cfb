"""Host-side mirror of the reference's driver around the hot path: the class `PhaseFieldCrystalSystemMPI`
(src/phase_field_crystal_systems/phase_field_crystal_system_mpi.hpp:30-51: same constructor arguments, `run()`) and the
parameter-file front end of app/simulations/phase_field_crystal_simulation_mpi.cpp:15-183 (same TOML keys, same error
texts).  Everything numerical is one C-ABI call (pfc_setup / pfc_step / pfc_stress / pfc_output); this file only orders
the calls the way `run()` does (mpi.cpp:650-711) and prints what the reference prints through pcout.

    python -m pfc_b200.simulation parameter-files/hexagonal_lattice.toml
"""
import math
import os
import sys
from dataclasses import dataclass, field

import numpy as np

from . import LATTICE_A, PfcError, Solver, StandIn, default_A0, vtu_write


@dataclass
class HexagonalLattice:
    """HexagonalLattice<dim>(A_0, psi_0, dislocation_positions, burgers_vectors), hexagonal_lattice.hpp / .cpp:14-21"""
    A_0: float
    psi_0: float
    dislocation_positions: list = field(default_factory=list)
    burgers_vectors: list = field(default_factory=list)
    a = LATTICE_A


class PhaseFieldCrystalSystemMPI:
    """mpi.hpp:30-51.  Extra keyword arguments (not in the reference): rank / n_ranks / nccl_id for one process per GPU,
    `log` (pcout), and Solver options (operator_kind, preconditioner, device ...); the default is the matrix-free lattice
    path with the block-Schur preconditioner, `operator_kind=0, preconditioner=0` is the reference's algorithm,
    `preconditioner=2` the psi-only reduced solve (Q1; include/pfc_b200.h PFC_PRECOND_SCHUR_REDUCED)."""

    def __init__(self, degree, data_folder, configuration_filename, rhs_filename, output_interval, eps, dt, n_timesteps,
                 theta, simulation_tol, simulation_max_iters, n_refines, lower_left, upper_right, initial_condition, *,
                 rank=0, n_ranks=1, nccl_id=None, log=None, **solver_kw):
        self.degree, self.eps, self.dt, self.theta = int(degree), float(eps), float(dt), float(theta)
        self.data_folder = str(data_folder)
        self.configuration_stem = os.path.splitext(os.path.basename(str(configuration_filename)))[0]
        self.rhs_stem = os.path.splitext(os.path.basename(str(rhs_filename)))[0]
        self.output_interval, self.n_timesteps = int(output_interval), int(n_timesteps)
        self.simulation_tol, self.simulation_max_iters = float(simulation_tol), int(simulation_max_iters)
        self.n_refines, self.lower_left, self.upper_right = int(n_refines), tuple(lower_left), tuple(upper_right)
        self.initial_condition = initial_condition
        self.rank, self.n_ranks, self.nccl_id = rank, n_ranks, nccl_id
        self.pcout = (log if log is not None else print) if rank == 0 else (lambda *a, **k: None)
        solver_kw.setdefault("operator_kind", 1)
        solver_kw.setdefault("preconditioner", 1)
        self.solver_kw = solver_kw
        self.mesh = self.solver = None
        self.timer = {"assembly": 0.0, "solver": 0.0, "total": 0.0}
        self.stats = []

    # mpi.cpp:115-146 make_grid, 150-274 setup_dofs: the stand-in produces the flattened arrays, pfc_setup the plan
    def make_grid_and_setup_dofs(self):
        self.mesh = StandIn(self.degree, self.n_refines, self.lower_left, self.upper_right, rank=self.rank, n_ranks=self.n_ranks)
        self.solver = Solver(degree=self.degree, dt=self.dt, theta=self.theta, eps=self.eps, newton_tol=self.simulation_tol,
                             newton_max_iters=self.simulation_max_iters, **self.solver_kw)
        if self.n_ranks > 1:
            self.solver.comm_init(self.nccl_id, self.rank, self.n_ranks)
        self.solver.setup(self.mesh)

    def initialize_fe_field(self):  # mpi.cpp:278-296
        ic = self.initial_condition
        Psi = self.mesh.hexagonal_lattice(ic.A_0, ic.psi_0, ic.dislocation_positions, ic.burgers_vectors)
        self.solver.set_state(Psi, Psi)

    def iterate_timestep(self):  # mpi.cpp:576-593
        st = self.solver.step()
        for i, r in enumerate(st["residuals"]):
            self.pcout(f"Residual is: {r}")  # :585
            if i < len(st["gmres_iters"]):
                self.pcout(f"Number of iterations: {st['gmres_iters'][i]}")  # :483
        self.timer["assembly"] += st.get("ms_assembly", 0.0)
        self.timer["solver"] += st.get("ms_solver", 0.0)
        self.timer["total"] += st.get("ms_total", 0.0)
        self.stats.append(st)
        return st

    def output_configuration(self, iteration):  # mpi.cpp:597-619
        self.solver.output(self.mesh, 0, self.data_folder, self.configuration_stem, iteration)

    def output_rhs(self, iteration):  # mpi.cpp:624-645; before the first assembly system_rhs is all zero (mpi.cpp:271)
        if iteration == 0 and not self.stats:
            vtu_write(self.mesh, self.data_folder, self.rhs_stem, 0, ["psi", "chi", "phi"], np.zeros((3, self.mesh.N)))
        else:
            self.solver.output(self.mesh, 1, self.data_folder, self.rhs_stem, iteration)

    def calculate_and_output_stress(self, announce=False):  # mpi.cpp:673-683, 690-697
        try:
            _, its = self.solver.stress(self.eps, owned=True)
        except PfcError as e:
            if e.code != 7:  # PFC_ERR_UNSUPPORTED: no lattice path -> no stress projection
                raise
            if announce:
                self.pcout("stress projection needs the lattice path (operator_kind=1): skipped")
            return None
        if announce:
            for i in its:
                self.pcout(f"Number of stress iterations: {i}")  # :677-678
        self.solver.output(self.mesh, 2, self.data_folder, "stress", 0)
        return its

    def run(self):  # mpi.cpp:650-711
        if self.data_folder and self.rank == 0:
            os.makedirs(self.data_folder, exist_ok=True)  # (deal.II would fail on a missing folder)
        self.pcout("Making grid!")
        self.pcout("Setting up dofs!")
        self.make_grid_and_setup_dofs()
        self.pcout("Initializing fe field!")
        self.initialize_fe_field()
        self.pcout("Outputting initial configuration!")
        self.output_configuration(0)
        self.pcout("Outputting initial right-hand side!")
        self.output_rhs(0)
        self.pcout("Initializing stress_calculator!")
        self.calculate_and_output_stress(announce=True)
        for timestep in range(1, self.n_timesteps + 1):
            self.pcout(f"Iterating timestep: {timestep}")
            self.iterate_timestep()
            self.calculate_and_output_stress()
            if timestep % self.output_interval == 0:
                self.pcout("Outputting configuration!")
                self.output_configuration(timestep)
                self.pcout("Outputting right-hand side!")
                self.output_rhs(timestep)
            self.pcout("")
        # timer.print_summary() (mpi.cpp:711): the sections the hot path owns
        self.pcout(f"assembly {self.timer['assembly']:.3f} ms | solver {self.timer['solver']:.3f} ms | "
                   f"timesteps total {self.timer['total']:.3f} ms")

    def close(self):
        if self.solver is not None:
            self.solver.close()
            self.solver = None


def parse_simulation_parameters(tbl, **kw):
    """parse_simulation_parameters<2> (app_mpi.cpp:15-141): same keys, same std::invalid_argument texts (ValueError)."""
    def need(key, what):
        if key not in tbl:
            raise ValueError(f"No {what} in parameter file")
        return tbl[key]

    data_folder = need("data_folder", "data folder")
    configuration_filename = need("configuration_filename", "configuration filename")
    rhs_filename = need("rhs_filename", "rhs filename")
    output_interval = need("output_interval", "output_interval")
    vals = {k: need(k, k) for k in ("degree", "eps", "dt", "n_timesteps", "theta", "simulation_tol", "simulation_max_iters", "n_refines")}
    for key in ("p1", "p2"):
        if not isinstance(tbl.get(key), list):
            raise ValueError("p1 is not an array!")  # (the reference prints p1 for both, app_mpi.cpp:61,69)
    a = HexagonalLattice.a
    scale = bool(tbl.get("scale_by_lattice_constant", False))
    s = a if scale else 1.0
    p1, p2 = [float(v) * s for v in tbl["p1"]], [float(v) * s for v in tbl["p2"]]
    psi_0 = float(tbl["psi_0"])
    A_0 = tbl.get("A_0")
    if isinstance(A_0, (int, float)) and not isinstance(A_0, bool):
        A_0 = float(A_0)
    elif A_0 == "default":
        A_0 = 0.2 * (abs(psi_0) + (1.0 / 3.0) * math.sqrt(-15 * vals["eps"] - 36 * psi_0 * psi_0))  # app_mpi.cpp:87-88
    else:
        raise ValueError("Incorrect input for A_0")
    if not isinstance(tbl.get("dislocation_positions"), list):
        raise ValueError("Incorrect input for dislocation positions")
    if not isinstance(tbl.get("burgers_vectors"), list):
        raise ValueError("Incorrect input for burgers vectors")
    pos = [[float(c) * s for c in p] for p in tbl["dislocation_positions"]]
    bur = [[float(c) * s for c in b] for b in tbl["burgers_vectors"]]
    assert abs(A_0 - default_A0(psi_0, vals["eps"])) < 1e-15 or tbl.get("A_0") != "default"
    ic = HexagonalLattice(A_0, psi_0, pos, bur)
    # optional [gpu] table (not in the reference; its files have none and parse unchanged): pfc_params fields such as
    # operator_kind, preconditioner, precond_sweeps, orthogonalization, use_cuda_graph, device
    for k, v in dict(tbl.get("gpu", {})).items():
        kw.setdefault(k, v)
    return PhaseFieldCrystalSystemMPI(vals["degree"], data_folder, configuration_filename, rhs_filename, output_interval,
                                      vals["eps"], vals["dt"], vals["n_timesteps"], vals["theta"], vals["simulation_tol"],
                                      vals["simulation_max_iters"], vals["n_refines"], p1, p2, ic, **kw)


def main(argv):
    """main (app_mpi.cpp:143-183): errors are printed, the exit code stays 0; a parse error returns 1."""
    import tomllib
    try:
        try:
            with open(argv[1], "rb") as f:
                tbl = tomllib.load(f)
        except tomllib.TOMLDecodeError as err:
            print(f"Parsing failed:\n{err}", file=sys.stderr)
            return 1
        if int(tbl.get("dim", 0)) == 2:
            system = parse_simulation_parameters(tbl)
            system.run()
            system.close()
        elif int(tbl.get("dim", 0)) == 3:
            raise ValueError("dim = 3 is instantiated upstream (mpi.cpp:716) but not supported by this path")
    except Exception as err:  # noqa: BLE001 - mirrors catch (const std::exception &)
        print("--------------------------\nFound error in program!\n" + str(err) + "--------------------------", file=sys.stderr)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
