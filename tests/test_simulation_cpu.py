"""Parameter-file front end (app/simulations/phase_field_crystal_simulation_mpi.cpp:15-141) — host only."""
import tomllib

import numpy as np
import pytest

# the shipped parameter-files/hexagonal_lattice.toml, verbatim values
SHIPPED = """
dim = 2
degree = 1
data_folder = "./temp-data/"
configuration_filename = "phase_field.vtu"
rhs_filename = "phase_field_rhs.vtu"
output_interval = 10
eps = -0.8
dt = 0.1
n_timesteps = 10
theta = 1.0
simulation_tol = 1e-8
simulation_max_iters = 200
n_refines = 6
p1 = [-6.0, -5.0]
p2 = [6.0, 5.0]
scale_by_lattice_constant = true
psi_0 = -0.43
A_0 = "default"
dislocation_positions = [[2.0, 0.0], [-2.0, 0.0]]
burgers_vectors = [[1.0, 0.0], [-1.0, 0.0]]
"""


def test_shipped_parameter_file(built):
    import pfc_b200.simulation as S
    sys_ = S.parse_simulation_parameters(tomllib.loads(SHIPPED))
    a = 4 * np.pi / np.sqrt(3.0)
    assert (sys_.degree, sys_.n_refines, sys_.n_timesteps, sys_.output_interval) == (1, 6, 10, 10)
    assert np.allclose(sys_.lower_left, (-6 * a, -5 * a)) and np.allclose(sys_.upper_right, (6 * a, 5 * a))
    ic = sys_.initial_condition
    assert ic.A_0 == pytest.approx(0.24010818710676385, abs=1e-16) and ic.psi_0 == -0.43
    assert np.allclose(ic.dislocation_positions, [[2 * a, 0], [-2 * a, 0]]) and np.allclose(ic.burgers_vectors, [[a, 0], [-a, 0]])
    assert (sys_.configuration_stem, sys_.rhs_stem, sys_.data_folder) == ("phase_field", "phase_field_rhs", "./temp-data/")
    assert sys_.solver_kw == {"operator_kind": 1, "preconditioner": 1}


@pytest.mark.parametrize("drop,msg", [("degree", "No degree in parameter file"), ("data_folder", "No data folder in parameter file"),
                                      ("n_refines", "No n_refines in parameter file"), ("p2", "p1 is not an array!"),
                                      ("burgers_vectors", "Incorrect input for burgers vectors")])
def test_missing_parameters_raise_like_the_reference(built, drop, msg):
    import pfc_b200.simulation as S
    tbl = tomllib.loads(SHIPPED)
    del tbl[drop]
    with pytest.raises(ValueError, match=msg):
        S.parse_simulation_parameters(tbl)


def test_bad_A0(built):
    import pfc_b200.simulation as S
    tbl = tomllib.loads(SHIPPED)
    tbl["A_0"] = "auto"
    with pytest.raises(ValueError, match="Incorrect input for A_0"):
        S.parse_simulation_parameters(tbl)
    tbl["A_0"] = 0.3
    assert S.parse_simulation_parameters(tbl).initial_condition.A_0 == 0.3


def test_optional_gpu_table(built):
    """Old files parse unchanged; an optional [gpu] table carries pfc_params fields."""
    import pfc_b200.simulation as S
    tbl = tomllib.loads(SHIPPED + '\n[gpu]\noperator_kind = 0\npreconditioner = 0\n')
    assert S.parse_simulation_parameters(tbl).solver_kw == {"operator_kind": 0, "preconditioner": 0}
    assert S.parse_simulation_parameters(tbl, preconditioner=1).solver_kw == {"operator_kind": 0, "preconditioner": 1}


# ---- native driver phase-field-crystals_b200/pfc_sim (csrc/pfc_sim.cpp), host-only behaviour --------------------
def _pfc_sim(built, tmp_path, text, *flags):
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(built.__file__), "pfc_sim")
    f = tmp_path / "p.toml"
    f.write_text(text)
    return subprocess.run([exe, *flags, str(f)], capture_output=True, text=True, timeout=60)


def test_pfc_sim_check_mode_parses_the_shipped_file(built, tmp_path):
    r = _pfc_sim(built, tmp_path, "# This is a test parameter file\n" + SHIPPED + "\n# dislocation_positions = []\n", "--check")
    assert r.returncode == 0, r.stderr
    assert "degree 1 n_refines 6 n_timesteps 10 output_interval 10" in r.stdout
    assert "A_0 0.24010818710676385 dislocations 2" in r.stdout
    assert "p1 -43.531184741621232 -36.275987284684355" in r.stdout
    assert "./temp-data/phase_field_NN.R.vtu ./temp-data/phase_field_rhs_NN.R.vtu" in r.stdout
    assert "gpu operator_kind 1 preconditioner 1" in r.stdout
    r = _pfc_sim(built, tmp_path, SHIPPED + "\n[gpu]\noperator_kind = 0\npreconditioner = 0\n", "--check")
    assert "gpu operator_kind 0 preconditioner 0" in r.stdout


def test_pfc_sim_errors_like_the_reference_main(built, tmp_path):
    """app_mpi.cpp:143-183: a missing key is caught and printed, exit code 0; a file that does not parse returns 1."""
    r = _pfc_sim(built, tmp_path, SHIPPED.replace("degree = 1", ""), "--check")
    assert r.returncode == 0 and "Found error in program!" in r.stderr and "No degree in parameter file" in r.stderr
    r = _pfc_sim(built, tmp_path, SHIPPED.replace('A_0 = "default"', 'A_0 = "auto"'), "--check")
    assert r.returncode == 0 and "Incorrect input for A_0" in r.stderr
    r = _pfc_sim(built, tmp_path, "dim = [2\n")
    assert r.returncode == 1 and "Parsing failed" in r.stderr


def test_pfc_sim_reader_covers_the_toml_forms_of_the_reference_files(built, tmp_path):
    """comments after values, multi-line arrays, integer-valued floats, empty arrays (the commented-out perfect-lattice
    variant of the shipped file), numeric A_0."""
    text = (SHIPPED.replace('dislocation_positions = [[2.0, 0.0], [-2.0, 0.0]]', 'dislocation_positions = []  # perfect lattice')
            .replace('burgers_vectors = [[1.0, 0.0], [-1.0, 0.0]]', 'burgers_vectors = [\n]\n')
            .replace('A_0 = "default"', 'A_0 = 0.25 # amplitude')
            .replace('p1 = [-6.0, -5.0]', 'p1 = [\n  -6.0,   # x\n  -5.0,\n]')
            .replace('n_refines = 6', 'n_refines = 7'))
    r = _pfc_sim(built, tmp_path, text, "--check")
    assert r.returncode == 0 and r.stderr == "", r.stderr
    assert "n_refines 7" in r.stdout and "A_0 0.25 dislocations 0" in r.stdout and "p1 -43.531184741621232 -36.275987284684355" in r.stdout
    # and the Python front end reads the same file the same way
    import pfc_b200.simulation as S
    s = S.parse_simulation_parameters(tomllib.loads(text))
    assert s.n_refines == 7 and s.initial_condition.A_0 == 0.25 and s.initial_condition.dislocation_positions == []
