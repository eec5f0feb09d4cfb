"""Host detection of the uniform periodic tensor lattice (csrc/pfc_lattice_plan.cpp) — no GPU needed.

The lattice plan must (a) number the owned free nodes lexicographically, (b) map every frame position to
the resolved (periodic master) node sitting there, on every rank of the partitions the bench uses, and
(c) step aside (active = 0, reason given) when the mesh is not such a lattice."""
import numpy as np
import pytest

A = 4 * np.pi / np.sqrt(3.0)
P1, P2 = (-6 * A, -5 * A), (6 * A, 5 * A)


def _check(built, degree, n_refines, n_ranks, rep, flags=1):
    full = built.StandIn(degree, n_refines, P1, P2, rep=rep)          # coordinates of every global node
    sup = full.support
    Lx, Ly = P2[0] - P1[0], P2[1] - P1[1]
    master = np.arange(full.N)
    sl, ms = full.constraint_slave, full.constraint_master
    master[sl[sl < full.N]] = ms[sl < full.N]
    covered = np.zeros(full.N, bool)
    for rank in range(n_ranks):
        st = built.StandIn(degree, n_refines, P1, P2, rank=rank, n_ranks=n_ranks, rep=rep)
        pl = built.HostPlan(st.mesh_ptr, degree, flags)
        L = pl.lattice
        assert L["active"] == 1, L["why"]
        w, h, ex0, ey0, p = L["w"], L["h"], L["ex0"], L["ey0"], degree
        lut = L["lut"]
        assert lut.shape == (p * L["ncy"] + 1, p * L["ncx"] + 1)
        # internal id -> global node
        glob = np.full(pl.Np + pl.n_ghost, -1, np.int64)
        glob[:pl.Np] = pl.perm
        glob[pl.Np:] = pl.ghost_global
        g = glob[lut]
        assert (g >= 0).all()
        assert (master[g] == g).all()                                   # frame holds resolved nodes only
        # owned rectangle is lexicographic
        assert np.array_equal(lut[ey0:ey0 + h, ex0:ex0 + w], np.arange(w * h).reshape(h, w))
        own = glob[:w * h]
        assert not covered[own].any()
        covered[own] = True
        # geometry: frame position (ex, ey) sits at origin + (ex, ey) * h / p modulo the box
        x0 = sup[own[0]] - np.array([ex0 * L["hx"] / p, ey0 * L["hy"] / p])
        ey, ex = np.meshgrid(np.arange(lut.shape[0]), np.arange(lut.shape[1]), indexing="ij")
        dx = (sup[g, 0] - (x0[0] + ex * L["hx"] / p)) / Lx
        dy = (sup[g, 1] - (x0[1] + ey * L["hy"] / p)) / Ly
        assert np.abs(dx - np.round(dx)).max() < 1e-9 and np.abs(dy - np.round(dy)).max() < 1e-9
        # the owned slaves follow the free nodes
        rest = glob[w * h:pl.n_owned]
        assert (master[rest] != rest).all()
        pl.close()
        st.close()
    assert covered[master == np.arange(full.N)].all()                   # every free node owned exactly once
    full.close()


@pytest.mark.parametrize("degree,n_refines,n_ranks,rep", [
    (1, 1, 1, (1, 1)), (2, 1, 1, (1, 1)), (1, 4, 1, (1, 1)), (2, 3, 1, (1, 1)),
    (1, 4, 2, (1, 1)), (2, 3, 2, (1, 1)), (1, 4, 4, (1, 1)), (2, 3, 4, (1, 1)), (1, 4, 8, (1, 1)), (2, 3, 8, (1, 1)),
    (2, 3, 2, (2, 1)), (1, 3, 4, (2, 2)), (2, 3, 8, (4, 2)), (1, 3, 8, (4, 2))])
def test_lattice_detected(built, degree, n_refines, n_ranks, rep):
    _check(built, degree, n_refines, n_ranks, rep)


def test_lattice_without_tables(built):
    _check(built, 2, 3, 2, (2, 1), flags=3)


def test_lattice_declines_non_rectangular_partition(built):
    st = built.StandIn(1, 3, P1, P2, rank=1, n_ranks=3)
    pl = built.HostPlan(st.mesh_ptr, 1, 1)
    assert pl.lattice["active"] == 0 and pl.lattice["why"]
    # the general plan is untouched by the failed detection
    ref = built.HostPlan(st.mesh_ptr, 1, 0)
    assert np.array_equal(pl.perm, ref.perm) and np.array_equal(pl.col, ref.col)
