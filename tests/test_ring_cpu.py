"""Deep ghost ring of the lattice path (csrc/pfc_ring_plan.cpp) — host lists checked without a GPU.

For every rank of a partition: after the exchange described by the ring lists (self copies + peer transfers, emulated in
numpy over all ranks of the partition) every padded position holds the value of the global node that sits at that
lattice position modulo the period; the positions of the handed-over one-cell frame keep their global ids; every owned
slave finds its master's value at its own position (constraints.distribute, mpi.cpp:485)."""
import numpy as np
import pytest

A = 4 * np.pi / np.sqrt(3.0)
P1, P2 = (-6 * A, -5 * A), (6 * A, 5 * A)


def ring_index(X, Y, ex0, ey0, w, h, fw, fh):
    """numpy restatement of ring_index (csrc/pfc_plan.h)"""
    top = Y < ey0
    bot = Y >= ey0 + h
    side = (fh - h) * fw + (Y - ey0) * (fw - w) + np.where(X < ex0, X, X - w)
    return np.where(top, Y * fw + X, np.where(bot, (Y - h) * fw + X, side))


def _partition(built, degree, n_refines, n_ranks, depth, rep=(1, 1)):
    full = built.StandIn(degree, n_refines, P1, P2, rep=rep)
    sup = full.support.copy()
    N = full.N
    master = np.arange(N)
    sl, ms = full.constraint_slave, full.constraint_master
    master[sl[sl < N]] = ms[sl < N]
    p = degree
    hx = (P2[0] - P1[0]) / (rep[0] * 2 ** n_refines) / p
    hy = (P2[1] - P1[1]) / (rep[1] * 2 ** n_refines) / p
    PX, PY = p * rep[0] * 2 ** n_refines, p * rep[1] * 2 ** n_refines
    ix = np.rint((sup[:, 0] - P1[0]) / hx).astype(np.int64)
    iy = np.rint((sup[:, 1] - P1[1]) / hy).astype(np.int64)
    node_at = np.full((PY, PX), -1, np.int64)          # lattice position modulo the period -> resolved global node
    free = master == np.arange(N)
    node_at[iy[free] % PY, ix[free] % PX] = np.arange(N)[free]
    assert (node_at >= 0).all()
    sts = [built.StandIn(degree, n_refines, P1, P2, rank=r, n_ranks=n_ranks, rep=rep) for r in range(n_ranks)]
    plans = [built.HostPlan(s.mesh_ptr, degree, 3) for s in sts]
    built.ring_overlay(plans, depth)
    return full, master, (ix, iy, PX, PY), node_at, sts, plans


@pytest.mark.parametrize("degree,n_refines,n_ranks,depth,rep", [
    (1, 4, 1, 11, (1, 1)), (2, 3, 1, 34, (1, 1)), (1, 2, 1, 11, (1, 1)), (2, 1, 1, 10, (1, 1)),
    (1, 4, 2, 11, (1, 1)), (2, 3, 2, 12, (1, 1)), (1, 4, 4, 5, (1, 1)), (2, 3, 4, 34, (1, 1)),
    (1, 5, 8, 11, (1, 1)), (2, 4, 8, 16, (1, 1)), (1, 3, 8, 1, (4, 2)), (2, 3, 2, 7, (2, 1))])
def test_ring_exchange_fills_every_position(built, degree, n_refines, n_ranks, depth, rep):
    full, master, (ix, iy, PX, PY), node_at, sts, plans = _partition(built, degree, n_refines, n_ranks, depth, rep)
    f = np.random.default_rng(5).standard_normal(full.N)
    f = f[master]                                       # a periodic field
    rings = [pl.ring() for pl in plans]
    loc = []
    for pl, rg in zip(plans, rings):
        assert rg["active"] == 1 and rg["n_ghost"] == rg["n_ring"] + 1
        L = pl.lattice
        assert rg["ex0"] >= depth and rg["ey0"] >= depth and rg["fw"] - rg["ex0"] - L["w"] >= depth
        assert rg["fh"] - rg["ey0"] - L["h"] >= depth
        assert (rg["fw"] - 1) % degree == 0 and (rg["fh"] - 1) % degree == 0 and (rg["ex0"] - L["ex0"]) % degree == 0
        v = np.full(pl.Np + rg["n_ghost"], np.nan)
        own = pl.perm[:pl.n_owned]
        v[:pl.n_owned] = f[own]
        loc.append(v)
    # the exchange: self copies, then peer transfers in list order
    for r, (pl, rg) in enumerate(zip(plans, rings)):
        loc[r][pl.Np + rg["self_dst"]] = loc[r][rg["self_src"]]
    for r, (pl, rg) in enumerate(zip(plans, rings)):
        for k, q in enumerate(rg["peers"]):
            rq = rings[q]
            kq = list(rq["peers"]).index(r)             # every pair is two-way (dummy values where the geometry is not)
            send = rq["send_rows"][rq["send_ptr"][kq]:rq["send_ptr"][kq + 1]]
            pos = rg["recv_pos"][rg["recv_ptr"][k]:rg["recv_ptr"][k + 1]]
            assert len(send) == len(pos) and len(pos) > 0
            loc[r][pl.Np + pos] = loc[q][send]
    for r, (pl, rg) in enumerate(zip(plans, rings)):
        L = pl.lattice
        w, h, ex0, ey0, fw, fh = L["w"], L["h"], rg["ex0"], rg["ey0"], rg["fw"], rg["fh"]
        g0 = pl.perm[0]                                  # owned free node (0, 0)
        own = pl.perm[:pl.n_owned]
        Y, X = np.meshgrid(np.arange(fh), np.arange(fw), indexing="ij")
        want = f[node_at[(iy[g0] + Y - ey0) % PY, (ix[g0] + X - ex0) % PX]]
        inside = (X >= ex0) & (X < ex0 + w) & (Y >= ey0) & (Y < ey0 + h)
        idx = np.where(inside, (Y - ey0) * w + (X - ex0), pl.Np + ring_index(X, Y, ex0, ey0, w, h, fw, fh))
        assert len(np.unique(idx)) == fw * fh           # a bijection onto owned + ring ids
        assert idx[~inside].max() == pl.Np + rg["n_ring"] - 1
        got = loc[r][idx]
        assert not np.isnan(got).any()
        assert np.array_equal(got, want)
        # the handed-over frame keeps its global ids
        gg = rg["ghost_global"]
        known = np.nonzero(gg >= 0)[0]
        assert len(known) > 0 or n_ranks == 1
        assert not np.isin(gg[known], own).any()        # images of the rank's own nodes stay anonymous
        assert np.array_equal(loc[r][pl.Np + known], f[gg[known]])
        fx, fy = np.meshgrid(np.arange(L["fw"]), np.arange(L["fh"]))
        lx, ly = fx - L["ex0"], fy - L["ey0"]
        out = ~((lx >= 0) & (lx < w) & (ly >= 0) & (ly < h))
        at = gg[ring_index(lx[out] + ex0, ly[out] + ey0, ex0, ey0, w, h, fw, fh)]
        there = node_at[(iy[g0] + ly[out]) % PY, (ix[g0] + lx[out]) % PX]   # resolved node at the frame position
        mine = np.isin(there, own)
        assert np.array_equal(at[~mine], there[~mine]) and (at[mine] == -1).all()
        # owned slaves: internal ids after the free nodes, master value at the slave's own position
        sid, smas = rg["slave_id"], rg["slave_master"]
        assert np.array_equal(sid, np.arange(w * h, pl.n_owned))
        assert np.array_equal(loc[r][smas], f[master[pl.perm[sid]]])
    for x in plans + sts + [full]:
        x.close()


def test_ring_rejects_partitions_that_are_no_lattice(built):
    sts = [built.StandIn(1, 3, P1, P2, rank=r, n_ranks=3) for r in range(3)]
    plans = [built.HostPlan(s.mesh_ptr, 1, 1) for s in sts]
    with pytest.raises(built.PfcError):
        built.ring_overlay(plans, 3)
