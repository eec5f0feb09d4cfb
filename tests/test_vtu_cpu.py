"""Output writer: pfc_vtu_write against the arrays it is given — no GPU involved.
File naming follows write_vtu_with_pvtu_record(directory, basename, counter, comm, n_digits_for_counter = 2)
as called at phase_field_crystal_system_mpi.cpp:612-616, 638-642 and stress_calculator_mpi.cpp:377-381."""
import os

import numpy as np
import pytest

from vtu_reader import read_piece, read_record

A = 4 * np.pi / np.sqrt(3.0)
P1, P2 = (-6 * A, -5 * A), (6 * A, 5 * A)


def test_piece_names(built):
    assert built.vtu_piece_name("hexagonal_lattice", 0, 0, 1) == "hexagonal_lattice_00.0.vtu"
    assert built.vtu_piece_name("hexagonal_lattice", 7, 3, 8) == "hexagonal_lattice_07.3.vtu"
    assert built.vtu_piece_name("stress", 123, 11, 12) == "stress_123.11.vtu"
    assert built.vtu_piece_name("rhs", 5, -1, 4) == "rhs_05.pvtu"


@pytest.mark.parametrize("degree,n_refines", [(1, 3), (2, 3)])
def test_single_rank_piece_round_trip(built, tmp_path, degree, n_refines):
    st = built.StandIn(degree, n_refines, P1, P2)
    N = st.N
    x, y = st.support[:, 0], st.support[:, 1]
    fields = np.stack([np.sin(0.1 * x) + y, x * y, 3.0 + 0 * x])
    d = str(tmp_path) + os.sep
    built.vtu_write(st, d, "hexagonal_lattice", 4, ["psi", "chi", "phi"], fields)
    pc = read_piece(d + "hexagonal_lattice_04.0.vtu")
    nc = 4**n_refines
    assert pc["n_cells"] == nc and pc["n_points"] == 4 * nc
    assert np.all(pc["types"] == 9) and np.array_equal(pc["offsets"], 4 * np.arange(1, nc + 1))
    assert np.array_equal(pc["connectivity"].reshape(nc, 4) - 4 * np.arange(nc)[:, None], np.tile([0, 1, 3, 2], (nc, 1)))
    pts = pc["points"]
    assert np.all(pts[:, 2] == 0)
    # every written value is the field at that point (the fields are functions of the position)
    assert np.allclose(pc["point_data"]["psi"], np.sin(0.1 * pts[:, 0]) + pts[:, 1], atol=1e-12)
    assert np.allclose(pc["point_data"]["chi"], pts[:, 0] * pts[:, 1], atol=1e-10)
    assert np.all(pc["point_data"]["phi"] == 3.0)
    # quads are counter-clockwise and tile the box
    q = pts[pc["connectivity"].reshape(nc, 4)]
    area = 0.5 * np.abs(np.sum(q[:, :, 0] * np.roll(q[:, :, 1], -1, 1) - np.roll(q[:, :, 0], -1, 1) * q[:, :, 1], axis=1))
    assert np.isclose(area.sum(), (P2[0] - P1[0]) * (P2[1] - P1[1]))
    rec = read_record(d + "hexagonal_lattice_04.pvtu")
    assert rec["pieces"] == ["hexagonal_lattice_04.0.vtu"] and rec["arrays"] == {"psi": 1, "chi": 1, "phi": 1}


def test_tensor_field_and_ranks(built, tmp_path):
    """Four fields named "stress" with component_is_part_of_tensor become one 9-component array (2x2 block filled);
    each rank writes its owned cells, rank 0 the record listing all pieces."""
    d = str(tmp_path) + os.sep
    world, total_cells, seen = 4, 0, []
    for rank in range(world):
        st = built.StandIn(1, 4, P1, P2, rank=rank, n_ranks=world)
        x, y = st.support[:, 0], st.support[:, 1]
        fields = np.stack([x, y, 2 * x, 3 * y])
        built.vtu_write(st, d, "stress", 0, ["stress"] * 4, fields, interpretation=[1, 1, 1, 1])
        pc = read_piece(d + f"stress_00.{rank}.vtu")
        total_cells += pc["n_cells"]
        t, pts = pc["point_data"]["stress"], pc["points"]
        assert t.shape == (pc["n_points"], 9)
        assert np.allclose(t[:, 0], pts[:, 0]) and np.allclose(t[:, 1], pts[:, 1])
        assert np.allclose(t[:, 3], 2 * pts[:, 0]) and np.allclose(t[:, 4], 3 * pts[:, 1])
        assert np.all(t[:, [2, 5, 6, 7, 8]] == 0)
        seen.append(pts[pc["connectivity"].reshape(-1, 4)].mean(axis=1))
    assert total_cells == 4**4
    centres = np.round(np.concatenate(seen), 9)
    assert len(np.unique(centres, axis=0)) == 4**4  # every cell exactly once over the ranks
    rec = read_record(d + "stress_00.pvtu")
    assert rec["pieces"] == [f"stress_00.{r}.vtu" for r in range(world)] and rec["arrays"] == {"stress": 9}


def test_errors(built, tmp_path):
    st = built.StandIn(1, 2, P1, P2)
    with pytest.raises(built.PfcError) as e:
        built.vtu_write(st, str(tmp_path / "missing_dir") + os.sep, "x", 0, ["psi"], np.zeros((1, st.N)))
    assert "cannot open" in str(e.value)
