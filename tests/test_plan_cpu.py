"""Host side of pfc_setup against the oracle, without a GPU (-m "not gpu").

* stand-in arrays == oracle arrays bit for bit (DoF maps, constraints, support points, tables)
* derived monolithic sparsity pattern == oracle pattern bit for bit          (mpi.cpp:220-244)
* numpy emulation of the kernels over the plan's tables (tests/emul.py) reproduces the oracle's
  assembled matrix, right-hand side and SpMV  => SELL layout, permutation, colouring and scatter
  positions are right
* P-rank plans (Morton partition + ghost layer + halo lists) reproduce the 1-rank result
"""
import numpy as np
import pytest

import emul

A = 4 * np.pi / np.sqrt(3.0)
P1, P2 = (-6 * A, -5 * A), (6 * A, 5 * A)
DISL, BURG = [[2 * A, 0], [-2 * A, 0]], [[A, 0], [-A, 0]]
DT, THETA, EPS, PSI0 = 0.1, 1.0, -0.8, -0.43


def _case(built, degree, n_refines, theta=THETA):
    import oracle as O
    om = O.Mesh(degree, n_refines, P1, P2)
    st = built.StandIn(degree, n_refines, P1, P2)
    A0 = built.default_A0(PSI0, EPS)
    Psi = st.hexagonal_lattice(A0, PSI0, DISL, BURG)
    rng = np.random.default_rng(7)
    Psi1 = Psi + 1e-2 * rng.standard_normal(Psi.size)
    om.distribute(Psi1)
    return om, st, Psi, Psi1


@pytest.mark.parametrize("degree,n_refines", [(1, 1), (1, 3), (1, 5), (2, 1), (2, 2), (2, 4)])
def test_standin_matches_oracle_bit_exact(built, degree, n_refines):
    om, st, Psi, _ = _case(built, degree, n_refines)
    assert np.array_equal(st.cell_dofs, om.cell_dofs)
    assert np.array_equal(st.cell_vertices, om.cell_vertices)
    assert np.array_equal(st.support, om.support)
    assert np.array_equal(st.shape, om.shape) and np.array_equal(st.dshape, om.dshape)
    assert np.array_equal(st.qw, om.qw)
    slaves = np.nonzero(om.master >= 0)[0]
    order = np.argsort(st.constraint_slave)
    assert np.array_equal(st.constraint_slave[order], slaves)
    assert np.array_equal(st.constraint_master[order], om.master[slaves])
    A0 = built.default_A0(PSI0, EPS)
    assert np.array_equal(Psi, om.initial_condition(A0, PSI0, DISL, BURG))


@pytest.mark.parametrize("degree,n_refines,n_grains", [(1, 5, 1), (1, 6, 7), (2, 4, 64), (2, 5, 5)])
def test_polycrystal_standin_matches_oracle_bit_exact(built, degree, n_refines, n_grains):
    """config-4 initial condition (harness extension): the stand-in's and the oracle's generators agree bit for bit,
    grains are periodic Voronoi cells, every grain is the reference's cos-sum in a rotated frame."""
    import oracle as O
    om = O.Mesh(degree, n_refines, P1, P2)
    st = built.StandIn(degree, n_refines, P1, P2)
    A0 = built.default_A0(PSI0, EPS)
    seeds, alpha, offset = built.polycrystal_grains(P1, P2, n_grains)
    assert np.array_equal(built.polycrystal_grains(P1, P2, n_grains)[0], seeds)  # deterministic (PCG64 seed 12345)
    assert 0 <= alpha.min() and alpha.max() < np.pi / 3 and seeds.shape == (n_grains, 2)
    a = st.polycrystal(A0, PSI0, seeds, alpha, offset)
    b = om.polycrystal(A0, PSI0, seeds, alpha, offset)
    assert np.array_equal(a, b)
    N = om.N
    assert np.array_equal(a[N:2 * N], -(a[:N] - PSI0)) or np.allclose(a[N:2 * N], -(a[:N] - PSI0), atol=1e-15)
    assert np.array_equal(a[2 * N:], -a[N:2 * N])
    slaves = np.nonzero(om.master >= 0)[0]
    assert np.array_equal(a[slaves], a[om.master[slaves]])  # constraints.distribute
    # independent numpy evaluation of the definition
    W, H = P2[0] - P1[0], P2[1] - P1[1]
    d = om.support[:, None, :] - seeds[None, :, :]
    d[..., 0] -= W * np.rint(d[..., 0] / W)
    d[..., 1] -= H * np.rint(d[..., 1] / H)
    g = np.argmin((d ** 2).sum(-1), axis=1)
    rho = d[np.arange(N), g]
    ca, sa = np.cos(alpha[g]), np.sin(alpha[g])
    qx, qy = ca * rho[:, 0] + sa * rho[:, 1] + offset[g, 0], ca * rho[:, 1] - sa * rho[:, 0] + offset[g, 1]
    ssum = np.cos(qy) + np.cos(0.5 * np.sqrt(3.0) * qx - 0.5 * qy) + np.cos(-0.5 * np.sqrt(3.0) * qx - 0.5 * qy)
    free = om.master[:N] < 0
    assert np.abs((PSI0 + A0 * ssum)[free] - a[:N][free]).max() <= 1e-12
    if n_grains == 1:  # one grain with alpha = 0, offset 0 is the perfect lattice seen from the seed
        z = st.polycrystal(A0, PSI0, [[0.0, 0.0]], [0.0], [[0.0, 0.0]])
        assert np.abs(z - st.hexagonal_lattice(A0, PSI0))[:N][free].max() <= 1e-10


@pytest.mark.parametrize("degree,n_refines", [(1, 1), (1, 2), (1, 5), (2, 1), (2, 3)])
def test_pattern_bit_exact(built, degree, n_refines):
    om, st, _, _ = _case(built, degree, n_refines)
    plan = built.HostPlan(st.mesh_ptr, degree)
    rp, ci, blk, pos = plan.monolithic()
    assert plan.nnz_mono == om.nnz
    assert np.array_equal(rp, om.row_ptr)
    assert np.array_equal(ci, om.col_idx)
    # SELL structure: every real entry of a row is inside its slice, padding is (row, 0-valued)
    assert plan.Np % 32 == 0 and plan.slice_ptr[-1] == plan.n_ent
    assert sorted(plan.perm[plan.perm >= 0].tolist()) == list(range(om.N))


def _mono_values(plan, vB, vC, vM, vK, dtth):
    rp, ci, blk, pos = plan.monolithic()
    D = dtth * (vK - 2.0 * vM)
    src = np.stack([vB, vC, D, vK, vM])
    return src[blk.astype(np.int64), pos]


@pytest.mark.parametrize("degree,n_refines,theta", [(1, 3, 1.0), (1, 4, 0.5), (2, 2, 1.0), (2, 3, 0.5), (1, 1, 1.0), (2, 1, 0.5)])
def test_emulated_assembly_and_spmv_match_oracle(built, degree, n_refines, theta):
    om, st, Psi, Psi1 = _case(built, degree, n_refines)
    plan = built.HostPlan(st.mesh_ptr, degree)
    N = om.N
    tab = (st.shape, st.dshape, st.qw)
    pn, pn1 = emul.to_internal(plan, Psi1, N), emul.to_internal(plan, Psi, N)
    B, C, M, K, rhs = emul.local_blocks(plan, tab, pn, pn1, DT, theta, EPS)
    vB, vC, vM, vK, r = emul.scatter(plan, B, C, M, K, rhs)
    vals = _mono_values(plan, vB, vC, vM, vK, DT * theta)
    ovals, orhs = om.assemble(Psi1, Psi, DT, theta, EPS)
    # scaled by the row infinity norm
    rown = np.maximum.reduceat(np.abs(ovals), om.row_ptr[:-1])
    scale = np.repeat(rown, np.diff(om.row_ptr))
    assert np.max(np.abs(vals - ovals) / scale) <= 1e-13
    got_rhs = emul.owned_to_api(plan, r, N, np.zeros(3 * N))
    assert np.max(np.abs(got_rhs - orhs)) <= 1e-13 * np.max(np.abs(orhs))
    assert np.all(got_rhs[om.master >= 0] == 0.0)
    # SpMV on a random vector (including non-zero entries on constrained dofs)
    x = np.random.default_rng(3).standard_normal(3 * N)
    y = emul.spmv(plan, vB, vC, vM, vK, emul.to_internal(plan, x, N), DT * theta)
    got = emul.owned_to_api(plan, y, N, np.zeros(3 * N))
    ref = om.spmv(ovals, x)
    assert np.max(np.abs(got - ref)) <= 1e-13 * np.max(np.abs(ref))
    yf = emul.spmv_fast(plan, vB, vC, vM, vK, emul.to_internal(plan, x, N), DT * theta)
    assert np.max(np.abs(yf - y)) <= 1e-13 * np.max(np.abs(ref))


@pytest.mark.parametrize("degree,n_refines,n_ranks", [(1, 3, 2), (1, 4, 4), (1, 4, 8), (2, 2, 2), (2, 3, 4), (2, 3, 8), (1, 2, 3)])
def test_multi_rank_plans_reproduce_single_rank(built, degree, n_refines, n_ranks):
    """emulated P-rank partition == 1 rank to 1e-12."""
    om, st1, Psi, Psi1 = _case(built, degree, n_refines)
    N = om.N
    ovals, orhs = om.assemble(Psi1, Psi, DT, THETA, EPS)
    x = np.random.default_rng(5).standard_normal(3 * N)
    om.distribute(x)
    ref = om.spmv(ovals, x)
    got_y, got_rhs = np.full(3 * N, np.nan), np.full(3 * N, np.nan)
    owned_seen = np.zeros(N, dtype=int)
    for r in range(n_ranks):
        st = built.StandIn(degree, n_refines, P1, P2, rank=r, n_ranks=n_ranks)
        assert np.array_equal(st.node_offsets, built.StandIn(degree, n_refines, P1, P2, 0, n_ranks).node_offsets)
        plan = built.HostPlan(st.mesh_ptr, degree)
        assert plan.own0 == st.node_offsets[r] and plan.n_owned == st.node_offsets[r + 1] - st.node_offsets[r]
        owned_seen[plan.perm[plan.perm >= 0]] += 1
        # ghosts are sorted by owner and the receive ranges tile them
        if plan.n_ghost:
            owners = np.searchsorted(st.node_offsets, plan.ghost_global, side="right") - 1
            assert np.all(np.diff(owners) >= 0) and np.all(owners != r)
            assert plan.recv_ptr[-1] == plan.n_ghost
            for k, peer in enumerate(plan.peers):
                assert np.all(owners[plan.recv_ptr[k]:plan.recv_ptr[k + 1]] == peer)
        tab = (st.shape, st.dshape, st.qw)
        # to_internal plays the halo exchange: ghost entries take the owners' values
        pn, pn1 = emul.to_internal(plan, Psi1, N), emul.to_internal(plan, Psi, N)
        B, C, M, K, rhs = emul.local_blocks(plan, tab, pn, pn1, DT, THETA, EPS)
        vB, vC, vM, vK, rr = emul.scatter(plan, B, C, M, K, rhs)
        y = emul.spmv_fast(plan, vB, vC, vM, vK, emul.to_internal(plan, x, N), DT * THETA)
        emul.owned_to_api(plan, y, N, got_y)
        emul.owned_to_api(plan, rr, N, got_rhs)
        # every slave that is local has its master local
        assert np.all(plan.slave_master >= 0)
    assert np.all(owned_seen == 1)
    assert np.max(np.abs(got_y - ref)) <= 1e-12 * np.max(np.abs(ref))
    assert np.max(np.abs(got_rhs - orhs)) <= 1e-12 * np.max(np.abs(orhs))


def test_plan_rejects_unsupported_inputs(built):
    import ctypes as C
    st = built.StandIn(1, 2, P1, P2)
    with pytest.raises(built.PfcError):
        built.HostPlan(st.mesh_ptr, 2)          # degree does not match dofs_per_cell
    m = built.MeshStruct()
    C.memmove(C.byref(m), st.mesh_ptr, C.sizeof(m))
    bad = st.cell_dofs.copy()
    bad[0, 1] += 1                               # breaks block congruence
    m.cell_dofs = bad.ctypes.data_as(C.POINTER(C.c_int32))
    with pytest.raises(built.PfcError):
        built.HostPlan(C.pointer(m), 1)
    C.memmove(C.byref(m), st.mesh_ptr, C.sizeof(m))
    v = st.cell_vertices.copy()
    v[0, 1, 1] += 0.1                            # not an axis-aligned rectangle
    m.cell_vertices = v.ctypes.data_as(C.POINTER(C.c_double))
    with pytest.raises(built.PfcError):
        built.HostPlan(C.pointer(m), 1)
    with pytest.raises(built.PfcError):
        built.StandIn(3, 2, P1, P2)
