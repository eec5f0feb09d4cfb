"""Parity of the CUDA path (through the C ABI) with the CPU oracle — run with -m gpu on a B200.

Tolerances (BASELINE.json north_star): psi relative L2 <= 1e-9 per step, <= 1e-7 after 100 steps;
sparsity pattern and DoF maps bit-exact; assembled values <= 1e-13 relative to the row inf-norm.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

A = 4 * np.pi / np.sqrt(3.0)
P1, P2 = (-6 * A, -5 * A), (6 * A, 5 * A)
DISL, BURG = [[2 * A, 0], [-2 * A, 0]], [[A, 0], [-A, 0]]
DT, EPS, PSI0 = 0.1, -0.8, -0.43


def _setup(built, degree, n_refines, theta=1.0, dislocations=True, **kw):
    import oracle as O
    om = O.Mesh(degree, n_refines, P1, P2)
    st = built.StandIn(degree, n_refines, P1, P2)
    A0 = built.default_A0(PSI0, EPS)
    Psi = st.hexagonal_lattice(A0, PSI0, DISL if dislocations else (), BURG if dislocations else ())
    s = built.Solver(degree=degree, theta=theta, **kw)
    s.setup(st)
    return om, st, s, Psi


def _rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("degree,n_refines,theta", [(1, 4, 1.0), (1, 6, 0.5), (2, 3, 1.0), (2, 5, 0.5), (1, 1, 1.0), (2, 1, 1.0)])
def test_assembly_pattern_rhs_operator(built, degree, n_refines, theta):
    om, st, s, Psi = _setup(built, degree, n_refines, theta)
    N = om.N
    Psi1 = Psi + 1e-2 * np.random.default_rng(7).standard_normal(Psi.size)
    om.distribute(Psi1)
    s.set_state(Psi1, Psi)
    s.assemble()
    rp, ci = s.get_pattern()
    assert np.array_equal(rp, om.row_ptr) and np.array_equal(ci, om.col_idx)      # bit-exact
    ovals, orhs = om.assemble(Psi1, Psi, DT, theta, EPS)
    vals = s.get_matrix_values()
    rown = np.maximum.reduceat(np.abs(ovals), om.row_ptr[:-1])
    assert np.max(np.abs(vals - ovals) / np.repeat(rown, np.diff(om.row_ptr))) <= 1e-13
    rhs = s.get_rhs()
    assert np.max(np.abs(rhs - orhs)) <= 1e-13 * np.max(np.abs(orhs))
    assert np.all(rhs[om.master >= 0] == 0.0)
    assert abs(s.residual_norm() - np.linalg.norm(orhs)) <= 1e-13 * np.linalg.norm(orhs)
    x = np.random.default_rng(3).standard_normal(3 * N)
    y, ref = s.apply_operator(x), om.spmv(ovals, x)
    assert np.max(np.abs(y - ref)) <= 1e-13 * np.max(np.abs(ref))
    # state round trip through the internal permutation
    a, b = s.get_state(want_n1=True)
    assert np.array_equal(a, Psi1) and np.array_equal(b, Psi)
    s.close()


@pytest.mark.parametrize("degree,n_refines", [(1, 5), (2, 4)])
def test_gmres_solve_matches_oracle(built, degree, n_refines):
    om, st, s, Psi = _setup(built, degree, n_refines)
    s.set_state(Psi, Psi)
    s.assemble()
    ovals, orhs = om.assemble(Psi, Psi, DT, 1.0, EPS)
    its, res = s.solve_update()
    xo, oits, conv, ores = om.gmres(ovals, orhs)
    assert conv and res <= 1e-10
    om.distribute(xo)
    got = s.get_state() - Psi
    # both solves stop at ||b - A x|| <= 1e-10 absolute: compare through the operator
    assert np.linalg.norm(om.spmv(ovals, np.where(om.master >= 0, 0.0, got)) - orhs) <= 2e-10
    assert _rel(got, xo) <= 1e-8
    assert its == oits, (its, oits)  # unpreconditioned: the same Krylov space, the same stopping decisions
    s.close()


@pytest.mark.parametrize("degree,n_refines,n_steps,kw", [
    (1, 6, 3, {}),                                   # config 1: shipped TOML (hexagonal_lattice.toml)
    (1, 6, 2, {"orthogonalization": 1}),             # CGS2
    (1, 6, 2, {"use_cuda_graph": 0}),
    (2, 4, 2, {}),
    (1, 4, 2, {"theta": 0.5}),
    (1, 6, 3, {"preconditioner": 1}),                # block-Schur right preconditioner, same stopping rule
    (2, 4, 2, {"preconditioner": 1}),
    (2, 4, 2, {"preconditioner": 1, "precond_sweeps": 6, "use_cuda_graph": 0}),
    (1, 5, 2, {"preconditioner": 1, "theta": 0.5}),
])
def test_step_parity(built, degree, n_refines, n_steps, kw):
    theta = kw.get("theta", 1.0)
    om, st, s, Psi = _setup(built, degree, n_refines, **kw)
    s.set_state(Psi, Psi)
    Pn, Pn1 = Psi.copy(), Psi.copy()
    N = om.N
    for step in range(n_steps):
        g = s.step()
        o = om.newton_step(Pn, Pn1, DT, theta, EPS)
        got = s.get_state()
        assert g["status"] == 0 and o["status"] == 0
        assert g["n_assemblies"] == o["n_assemblies"], (g["residuals"], o["residuals"])
        err = _rel(got[:N], Pn[:N])
        print(f"step {step}: gmres gpu {g['gmres_iters']} oracle {o['gmres_iters']} rel L2 psi {err:.2e}")
        assert err <= 1e-9
        assert _rel(got, Pn) <= 1e-9
    s.close()


def test_determinism_bitwise(built):
    """no atomics: two runs give bit-identical matrices and states."""
    outs = []
    for _ in range(2):
        om, st, s, Psi = _setup(built, 1, 5)
        s.set_state(Psi, Psi)
        s.assemble()
        v = s.get_matrix_values()
        s.step()
        outs.append((v, s.get_state()))
        s.close()
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("degree,n_refines,kw", [(1, 9, {}), (1, 9, {"operator_kind": 1, "preconditioner": 1}),
                                                 (2, 8, {"operator_kind": 1, "preconditioner": 1})])
def test_full_size_properties(built, degree, n_refines, kw):
    """Size-independent properties at a size the oracle would need minutes for (2^9 x 2^9 Q1 / 2^8 x 2^8 Q2 cells):
    sum(M chi) == 0, J linear, rows of K sum to 0; assembled and matrix-free lattice path."""
    st = built.StandIn(degree, n_refines, (-6 * A * 8, -5 * A * 8), (6 * A * 8, 5 * A * 8))
    A0 = built.default_A0(PSI0, EPS)
    Psi = st.hexagonal_lattice(A0, PSI0, [[16 * A, 0], [-16 * A, 0]], BURG)
    s = built.Solver(degree=degree, **kw)
    s.setup(st)
    s.set_state(Psi, Psi)
    g = s.step()
    assert g["status"] == 0
    new = s.get_state()
    N = st.N
    s.set_state(new, new)
    s.assemble()
    rng = np.random.default_rng(0)
    x, y = rng.standard_normal(3 * N), rng.standard_normal(3 * N)
    Jx, Jy, Jxy = s.apply_operator(x), s.apply_operator(y), s.apply_operator(2.0 * x - 3.0 * y)
    assert np.max(np.abs(Jxy - (2 * Jx - 3 * Jy))) <= 1e-12 * np.max(np.abs(Jxy))
    # chi equation M chi + K psi = 0 (mpi.cpp:400-405): summing its rows over the periodic mesh gives
    # sum(M chi) = -sum(K psi) = 0 because the rows of K sum to zero.  (The psi equation is NOT
    # discretely conservative: the reference expands lap(psi^3) = 3 psi^2 chi + 6 psi |grad psi|^2.)
    d = np.zeros(3 * N)
    d[N:2 * N] = new[N:2 * N]
    free = np.ones(3 * N, bool)
    free[st.constraint_slave] = False
    d[~free] = 0.0
    Md = s.apply_operator(d)[N:2 * N]
    assert abs(Md[free[N:2 * N]].sum()) <= 1e-7 * np.abs(Md).sum()
    one = np.zeros(3 * N)
    one[:N] = 1.0
    one[~free] = 0.0
    K1 = s.apply_operator(one)[N:2 * N]
    assert np.max(np.abs(K1[free[N:2 * N]])) <= 1e-12
    s.close()


def test_call_order_errors(built):
    s = built.Solver()
    for fn in (s.assemble, s.step, s.residual_norm, s.solve_update):
        with pytest.raises(built.PfcError) as e:
            fn()
        assert e.value.code == built.PFC_ERR_STATE
    with pytest.raises(built.PfcError):
        built.Solver(degree=3)
    with pytest.raises(built.PfcError):
        built.Solver(dt=-1.0)
    s.close()


@pytest.mark.parametrize("kw", [{}, {"operator_kind": 1}, {"operator_kind": 1, "preconditioner": 1}])
@pytest.mark.parametrize("name,degree,n_refines", [("config1_step10.npz", 1, 6), ("q2_r4_step3.npz", 2, 4)])
def test_against_committed_golden(built, name, degree, n_refines, kw):
    """CUDA path vs tests/golden (oracle output generated by tests/golden/make_golden.py), without the oracle:
    config 1 = the shipped TOML for all its 10 timesteps.  kw: assembled (reference algorithm), matrix-free lattice
    kernels, lattice kernels + block-Schur preconditioner (the bench default)."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", name))
    st = built.StandIn(degree, n_refines, P1, P2)
    Psi = st.hexagonal_lattice(built.default_A0(PSI0, EPS), PSI0, DISL, BURG)
    s = built.Solver(degree=degree, **kw)
    s.setup(st)
    s.set_state(Psi, Psi)
    tot = []
    for _ in range(int(g["steps"])):
        r = s.step()
        assert r["status"] == 0
        tot.append(r["gmres_iters_total"])
    got = s.get_state()
    N = st.N
    for k, (a, b) in enumerate(((got[:N], g["psi"]), (got[N:2 * N], g["chi"]), (got[2 * N:], g["phi"]))):
        assert _rel(a, b) <= 1e-9, (k, _rel(a, b))
    if not kw.get("preconditioner"):
        assert max(abs(int(x) - int(y)) for x, y in zip(tot, g["all_gmres_iters"])) <= 3, (tot, g["all_gmres_iters"])
    s.close()


@pytest.mark.parametrize("kind", [1, 2])
@pytest.mark.parametrize("degree,n_refines,theta", [(1, 4, 1.0), (1, 6, 0.5), (2, 3, 1.0), (2, 5, 0.5), (1, 1, 1.0), (2, 1, 1.0),
                                                    (1, 5, 0.0), (2, 4, 0.25)])
def test_matrix_free_operator_matches_csr(built, degree, n_refines, theta, kind):
    """north_star item 2: ||J_mf x - J_csr x|| / ||J_csr x|| <= 1e-13 on random x.
    kind 1 = lattice kernels (sum-factorised, no index tables), kind 2 = general colour-ordered cell kernel.
    The lattice path carries no rows for the periodic slaves (they are decoupled diagonal rows upstream)."""
    om, st, s, Psi = _setup(built, degree, n_refines, theta, operator_kind=kind)
    Psi1 = Psi + 1e-2 * np.random.default_rng(7).standard_normal(Psi.size)
    om.distribute(Psi1)
    s.set_state(Psi1, Psi)
    s.assemble()
    ovals, orhs = om.assemble(Psi1, Psi, DT, theta, EPS)
    rhs = s.get_rhs()
    assert np.max(np.abs(rhs - orhs)) <= 1e-13 * np.max(np.abs(orhs))
    x = np.random.default_rng(3).standard_normal(3 * om.N)
    y, ref = s.apply_operator(x), om.spmv(ovals, x)
    if kind == 1:
        slave = np.tile(om.master >= 0, 3) if om.master.size == om.N else (om.master >= 0)
        assert np.all(y[slave] == 0.0)
        ref = np.where(slave, 0.0, ref)
    assert np.linalg.norm(y - ref) <= 1e-13 * np.linalg.norm(ref)
    assert np.max(np.abs(y - ref)) <= 1e-12 * np.max(np.abs(ref))
    s.close()


@pytest.mark.parametrize("kind,precond", [(1, 0), (2, 0), (1, 1)])
@pytest.mark.parametrize("degree,n_refines,n_steps", [(1, 6, 2), (2, 4, 2)])
def test_matrix_free_step_parity(built, degree, n_refines, n_steps, kind, precond):
    om, st, s, Psi = _setup(built, degree, n_refines, operator_kind=kind, preconditioner=precond)
    s.set_state(Psi, Psi)
    Pn, Pn1 = Psi.copy(), Psi.copy()
    for step in range(n_steps):
        g = s.step()
        o = om.newton_step(Pn, Pn1, DT, 1.0, EPS)
        got = s.get_state()
        print(f"mf kind {kind} precond {precond} step {step}: gmres gpu {g['gmres_iters']} oracle {o['gmres_iters']} rel {_rel(got, Pn):.2e}")
        assert g["status"] == 0 and g["n_assemblies"] == o["n_assemblies"]
        assert _rel(got[:om.N], Pn[:om.N]) <= 1e-9 and _rel(got, Pn) <= 1e-9
    s.close()


@pytest.mark.parametrize("n_refines,n_steps,dislocations,theta", [(6, 3, True, 1.0), (5, 2, False, 1.0), (7, 2, True, 0.5), (4, 2, True, 1.0)])
def test_reduced_solve_step_parity(built, n_refines, n_steps, dislocations, theta):
    """PFC_PRECOND_SCHUR_REDUCED: the psi-only Schur-reduced solve of the serial class (serial.cpp:319-369) with
    dpsi = M^2 y instead of inner mass solves, against the oracle's 3-block Newton steps.  The initial condition has
    chi = phi = 0, so the first Newton iteration carries the two mass solves (G, H != 0); later ones do not."""
    om, st, s, Psi = _setup(built, 1, n_refines, theta=theta, dislocations=dislocations, operator_kind=1, preconditioner=2)
    kind, note = s.operator_info()
    assert kind == built.OPERATOR_MATRIX_FREE and "reduced solve active" in note, note
    s.set_state(Psi, Psi)
    Pn, Pn1 = Psi.copy(), Psi.copy()
    for step in range(n_steps):
        g = s.step()
        o = om.newton_step(Pn, Pn1, DT, theta, EPS)
        got = s.get_state()
        print(f"reduced 2^{n_refines} step {step}: gmres gpu {g['gmres_iters']} oracle {o['gmres_iters']} "
              f"final residuals {['%.1e' % r for r in g['gmres_final_residual']]} rel {_rel(got, Pn):.2e}")
        assert g["status"] == 0 and g["n_assemblies"] == o["n_assemblies"]
        assert all(r <= 1e-10 for r in g["gmres_final_residual"])
        assert _rel(got[:om.N], Pn[:om.N]) <= 1e-9 and _rel(got, Pn) <= 1e-9
    s.close()


def test_reduced_solve_against_block_schur(built):
    """Same mesh and steps with the 3-block preconditioned solve and with the reduced one: same states, fewer iterations;
    and the TRUE residual of the 3-block system for the update the reduced solve returns (J applied to it by
    pfc_apply_operator) is within the tolerance the reference asks of its GMRES (1e-10 absolute, mpi.cpp:478)."""
    res = {}
    for pc in (1, 2):
        om, st, s, Psi = _setup(built, 1, 6, operator_kind=1, preconditioner=pc)
        s.set_state(Psi, Psi)
        its = []
        for step in range(2):
            g = s.step()
            assert g["status"] == 0
            its.append(g["gmres_iters"])
        res[pc] = (s.get_state(), its)
        if pc == 2:
            X = res[pc][0]
            free = np.tile(om.master < 0, 3) if om.master.size == om.N else (om.master < 0)
            for k in range(2):                 # first and second Newton system of the third step
                s.assemble()
                rhs = s.get_rhs()
                n_it, est = s.solve_update()
                Y = s.get_state()
                s.set_state(X, res[pc][0])     # J is the Jacobian at the state before the update
                s.assemble()
                r = (rhs - s.apply_operator(Y - X))[free]
                print(f"reduced, Newton system {k}: {n_it} iterations, estimate {est:.2e}, true 3-block residual "
                      f"{np.linalg.norm(r):.2e}, |rhs| {np.linalg.norm(rhs[free]):.2e}")
                assert np.linalg.norm(r) <= 2e-10 and est <= 1e-10
                X = Y
                s.set_state(X, res[pc][0])
        s.close()
    print("gmres iterations block-Schur / reduced:", res[1][1], res[2][1])
    assert _rel(res[2][0], res[1][0]) <= 1e-9
    assert sum(sum(i) for i in res[2][1]) < sum(sum(i) for i in res[1][1])


def test_reduced_solve_falls_back_off_the_q1_lattice(built):
    """Q2, or the assembled operator: PFC_PRECOND_SCHUR_REDUCED acts as PFC_PRECOND_BLOCK_SCHUR and says so."""
    om, st, s, Psi = _setup(built, 2, 4, operator_kind=1, preconditioner=2)
    kind, note = s.operator_info()
    assert "using PFC_PRECOND_BLOCK_SCHUR" in note
    s.set_state(Psi, Psi)
    g = s.step()
    o = om.newton_step(Psi.copy(), Psi.copy(), DT, 1.0, EPS)
    assert g["status"] == 0 and g["n_assemblies"] == o["n_assemblies"]
    s.close()


def test_lattice_and_assembled_preconditioned_runs_agree(built):
    """Same physics, 3 steps, dislocation dipole: lattice matrix-free + block-Schur vs assembled + block-Schur."""
    res = []
    for kind in (0, 1):
        om, st, s, Psi = _setup(built, 2, 5, operator_kind=kind, preconditioner=1)
        s.set_state(Psi, Psi)
        its = []
        for _ in range(3):
            g = s.step()
            assert g["status"] == 0
            its.append(g["gmres_iters"])
        res.append((s.get_state(), its))
        s.close()
    print("gmres iterations assembled / lattice:", res[0][1], res[1][1])
    assert _rel(res[1][0], res[0][0]) <= 1e-9


@pytest.mark.parametrize("degree,n_refines", [(1, 1), (2, 1), (1, 5), (2, 4), (1, 8), (2, 7)])
def test_lattice_k_stencil_matches_cell_quadrature(built, degree, n_refines):
    """The node-stencil K product (assembled 1-D rows) and the cell-quadrature K product are the same operator."""
    om, st, s, Psi = _setup(built, degree, n_refines, operator_kind=1, preconditioner=1)
    err = s.selfcheck()
    print("K stencil vs cells:", err)
    assert 0.0 < err <= 1e-13
    s.close()


@pytest.mark.parametrize("degree,n_refines,kw", [(1, 5, {}), (1, 5, {"operator_kind": 1, "preconditioner": 1}),
                                                 (2, 4, {"operator_kind": 1, "preconditioner": 1})])
def test_hundred_steps_within_1e7(built, degree, n_refines, kw):
    """north_star: psi relative L2 <= 1e-7 after 100 steps (<= 1e-9 per step is checked by the step tests).
    A 4a x 4a box keeps the lattice resolved at sizes the oracle finishes 100 steps of in about a minute
    (on the 12a x 10a box of the other tests 2^4 cells are too coarse: the oracle itself blows up after ~30 steps)."""
    import oracle as O
    p1, p2 = (-2 * A, -2 * A), (2 * A, 2 * A)
    om = O.Mesh(degree, n_refines, p1, p2)
    st = built.StandIn(degree, n_refines, p1, p2)
    Psi = st.hexagonal_lattice(built.default_A0(PSI0, EPS), PSI0, [[A, 0], [-A, 0]], BURG)
    s = built.Solver(degree=degree, theta=1.0, **kw)
    s.setup(st)
    s.set_state(Psi, Psi)
    Pn, Pn1 = Psi.copy(), Psi.copy()
    worst = 0.0
    for step in range(100):
        g = s.step()
        o = om.newton_step(Pn, Pn1, DT, 1.0, EPS)
        assert g["status"] == 0 and o["status"] == 0
        if step % 10 == 9:
            worst = max(worst, _rel(s.get_state()[:om.N], Pn[:om.N]))
    got = s.get_state()
    print(f"100 steps Q{degree} 2^{n_refines} {kw}: rel L2 psi {_rel(got[:om.N], Pn[:om.N]):.2e} (worst at a multiple of 10: {worst:.2e})")
    assert _rel(got[:om.N], Pn[:om.N]) <= 1e-7 and _rel(got, Pn) <= 1e-7
    s.close()


# ---- stress projection (stress_calculator_mpi.cpp:345-353) -------------------------------------
@pytest.mark.parametrize("degree,n_refines", [(1, 5), (2, 4), (2, 5)])
def test_stress_matches_oracle(built, degree, n_refines):
    """Right-hand side kernel + mass stencil + batched CG against the oracle's assembled-matrix restatement:
    same CG iteration counts per component, nodal stress to 1e-9 of its max (CG to 1e-10 absolute on both sides)."""
    om, st, s, Psi = _setup(built, degree, n_refines, operator_kind=1, preconditioner=1)
    s.set_state(Psi, Psi)
    for _ in range(2):  # a state with chi, phi away from the initial condition
        s.step()
    state = s.get_state()
    o = om.stress(state, EPS)
    assert o["status"] == 0
    got, its = s.stress(EPS)
    print(f"stress Q{degree} 2^{n_refines}: CG its gpu {its} oracle {o['cg_iters']}, "
          f"max |diff| {np.abs(got - o['stress']).max():.2e} of {np.abs(o['stress']).max():.2e}")
    assert np.abs(got - o["stress"]).max() <= 1e-9 * np.abs(o["stress"]).max()
    assert all(abs(a - b) <= 1 for a, b in zip(its, o["cg_iters"]))
    so, its_o = s.stress(EPS, owned=True)
    assert np.array_equal(so, got) and its_o == its  # single rank: owned = all
    # off-diagonal components differ (div tau term), periodic images equal their masters
    assert np.abs(got[1] - got[2]).max() > 1e-3
    sl = np.nonzero(om.master[:om.N] >= 0)[0]
    assert np.array_equal(got[:, sl], got[:, om.master[sl]])
    s.close()


def test_stress_needs_the_lattice_path(built):
    om, st, s, Psi = _setup(built, 1, 4)
    s.set_state(Psi, Psi)
    with pytest.raises(built.PfcError) as e:
        s.stress(EPS)
    assert e.value.code == 7 and "lattice" in str(e.value)
    s.close()


@pytest.mark.parametrize("degree,n_refines", [(1, 9), (2, 8)])
def test_stress_full_size_strong_form(built, degree, n_refines):
    """Size-independent check at 2^9 x 2^9 Q1 / 2^8 x 2^8 Q2 cells (no oracle run): for psi = cos kx, chi = beta sin ly
    the projected stress converges to the strong form sigma_ab (tests/test_oracle_cpu.py pins the oracle the same way),
    including the unsymmetric xy / yx pair; constant fields are reproduced exactly."""
    p1, p2 = (-6 * A * 4, -5 * A * 4), (6 * A * 4, 5 * A * 4)
    st = built.StandIn(degree, n_refines, p1, p2)
    N = st.N
    x, y = st.support[:, 0], st.support[:, 1]
    k, l, beta = 2 * np.pi * 8 / (p2[0] - p1[0]), 2 * np.pi * 8 / (p2[1] - p1[1]), 0.6
    Psi = np.zeros(3 * N)
    Psi[:N], Psi[N:2 * N] = np.cos(k * x), beta * np.sin(l * y)
    s = built.Solver(degree=degree, operator_kind=1)
    s.setup(st)
    s.set_state(Psi, Psi)
    got, its = s.stress(EPS)
    c, sn, s2, c2 = np.cos(k * x), np.sin(k * x), np.sin(l * y), np.cos(l * y)
    F = 0.5 * (beta**2 * s2**2 + (1 + EPS) * c**2 + 0.5 * c**4)
    exact = np.stack([k * k * sn * sn + k * k * c * beta * s2 + F, 0 * x, -beta * k * l * sn * c2, F - k * k * c * c])
    err = np.abs(got - exact).max()
    print(f"stress Q{degree} 2^{n_refines}: CG its {its}, max |projected - strong form| {err:.2e}")
    assert err <= (2e-3 if degree == 1 else 5e-4) and all(0 < i < 100 for i in its[:1] + its[2:])
    Psi[:N], Psi[N:2 * N], Psi[2 * N:] = 0.3, -0.7, 5.0
    s.set_state(Psi, Psi)
    got, its = s.stress(EPS)
    f = 0.5 * (0.49 + (1 + EPS) * 0.09 + 0.5 * 0.3**4)
    assert np.abs(got[0] - f).max() <= 1e-9 and np.abs(got[3] - f).max() <= 1e-9 and np.abs(got[1:3]).max() <= 1e-9
    s.close()


# ---- output writer behind the handle (mpi.cpp:597-645, stress_calculator_mpi.cpp:358-384) -------
def test_output_writes_state_rhs_and_stress(built, tmp_path):
    import os
    from vtu_reader import read_piece, read_record
    om, st, s, Psi = _setup(built, 2, 4, operator_kind=1, preconditioner=1)
    s.set_state(Psi, Psi)
    s.step()
    d = str(tmp_path) + os.sep
    with pytest.raises(built.PfcError):
        s.output(st, 2, d, "stress", 0)  # before pfc_stress
    sg, _ = s.stress(EPS)
    s.assemble()
    s.output(st, 0, d, "hexagonal_lattice", 1)
    s.output(st, 1, d, "rhs", 1)
    s.output(st, 2, d, "stress", 0)
    state, rhs = s.get_state(), s.get_rhs()
    N = st.N
    key = {tuple(np.round(p, 9)): t for t, p in enumerate(st.support)}
    for name, ref, arrays in (("hexagonal_lattice_01", state, ("psi", "chi", "phi")), ("rhs_01", rhs, ("psi", "chi", "phi"))):
        pc = read_piece(d + name + ".0.vtu")
        node = np.array([key[tuple(np.round(p[:2], 9))] for p in pc["points"]])
        assert pc["n_cells"] == 4**4
        for c, a in enumerate(arrays):
            assert np.array_equal(pc["point_data"][a], ref[c * N + node]), (name, a)
        assert read_record(d + name + ".pvtu")["pieces"] == [name + ".0.vtu"]
    pc = read_piece(d + "stress_00.0.vtu")
    node = np.array([key[tuple(np.round(p[:2], 9))] for p in pc["points"]])
    t = pc["point_data"]["stress"]
    for c, slot in enumerate((0, 1, 3, 4)):
        assert np.array_equal(t[:, slot], sg[c, node])
    s.close()


def test_owned_range_state_transfers(built):
    """pfc_set_state_owned / pfc_get_state_owned (what a rank of the MPI original holds); single rank: owned = all."""
    om, st, s, Psi = _setup(built, 2, 3, operator_kind=1)
    assert s.n_owned_nodes() == st.N
    s.set_state_owned(Psi, Psi)
    assert np.array_equal(s.get_state(), Psi) and np.array_equal(s.get_state_owned(), Psi)
    a = s.step()
    got = s.get_state_owned()
    s.set_state(Psi, Psi)
    b = s.step()
    assert np.array_equal(got, s.get_state()) and a["gmres_iters"] == b["gmres_iters"]
    s.close()


@pytest.mark.parametrize("name,degree,n_refines", [("config1_step1", 1, 6), ("q2_r4_step3", 2, 4)])
def test_stress_against_committed_golden(built, name, degree, n_refines):
    """pfc_stress of a committed state vs the committed oracle stress (tests/golden/make_golden.py stress), without
    the oracle: config 1 = the shipped TOML after its first timestep, where the reference computes the stress too
    (mpi.cpp:690-693)."""
    import os
    gd = os.path.join(os.path.dirname(__file__), "golden")
    g, gs = np.load(os.path.join(gd, name + ".npz")), np.load(os.path.join(gd, name + "_stress.npz"))
    st = built.StandIn(degree, n_refines, P1, P2)
    s = built.Solver(degree=degree, operator_kind=1)
    s.setup(st)
    Psi = np.concatenate([g["psi"], g["chi"], g["phi"]])
    s.set_state(Psi, Psi)
    got, its = s.stress(EPS)
    assert np.abs(got - gs["stress"]).max() <= 1e-9 * np.abs(gs["stress"]).max()
    assert its == [int(i) for i in gs["cg_iters"]]
    s.close()


def test_shipped_toml_run_end_to_end(built, tmp_path):
    """The shipped parameter file through the mirror of PhaseFieldCrystalSystemMPI::run (mpi.cpp:650-711), one timestep:
    the configuration written to <data_folder>phase_field_01.0.vtu equals the committed oracle state after step 1
    (tests/golden/config1_step1.npz, psi rel L2 <= 1e-9), stress and right-hand-side files are in place."""
    import os
    import tomllib
    import pfc_b200.simulation as S
    from test_simulation_cpu import SHIPPED
    from vtu_reader import read_piece, read_record
    tbl = tomllib.loads(SHIPPED)
    tbl["data_folder"] = str(tmp_path) + os.sep
    tbl["n_timesteps"], tbl["output_interval"] = 1, 1
    lines = []
    sim = S.parse_simulation_parameters(tbl, log=lines.append)
    sim.run()
    d = tbl["data_folder"]
    for f in ("phase_field_00.0.vtu", "phase_field_00.pvtu", "phase_field_rhs_00.0.vtu", "phase_field_01.0.vtu",
              "phase_field_rhs_01.0.vtu", "stress_00.0.vtu", "stress_00.pvtu"):
        assert os.path.exists(d + f), f
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "config1_step1.npz"))
    gs = np.load(os.path.join(os.path.dirname(__file__), "golden", "config1_step1_stress.npz"))
    key = {tuple(np.round(p, 9)): t for t, p in enumerate(sim.mesh.support)}
    pc = read_piece(d + "phase_field_01.0.vtu")
    node = np.array([key[tuple(np.round(p[:2], 9))] for p in pc["points"]])
    for name in ("psi", "chi", "phi"):
        assert _rel(pc["point_data"][name], g[name][node]) <= 1e-9, name
    ps = read_piece(d + "stress_00.0.vtu")["point_data"]["stress"]
    for c, slot in enumerate((0, 1, 3, 4)):
        assert np.abs(ps[:, slot] - gs["stress"][c, node]).max() <= 1e-8 * np.abs(gs["stress"]).max()
    assert np.all(read_piece(d + "phase_field_rhs_00.0.vtu")["point_data"]["psi"] == 0)
    assert read_record(d + "stress_00.pvtu")["arrays"] == {"stress": 9}
    text = "\n".join(lines)
    for s_ in ("Making grid!", "Initializing stress_calculator!", "Number of stress iterations: ", "Iterating timestep: 1",
               "Residual is: ", "Number of iterations: ", "Outputting configuration!"):
        assert s_ in text, s_
    assert text.count("Number of stress iterations: ") == 4
    sim.close()


def test_native_driver_runs_the_shipped_toml(built, tmp_path):
    """The C++ driver csrc/pfc_sim.cpp (the stand-in for app/simulations/phase_field_crystal_simulation_mpi.cpp:15-183 +
    PhaseFieldCrystalSystemMPI::run, mpi.cpp:650-711) on the GPU: the shipped parameter file, one timestep; the VTU it
    writes equals the committed oracle state after step 1 (psi rel L2 <= 1e-9) and the pcout lines are the reference's."""
    import os
    import subprocess
    from test_simulation_cpu import SHIPPED
    from vtu_reader import read_piece
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "phase-field-crystals_b200", "pfc_sim")
    assert os.path.exists(exe), "pfc_sim is built by __graft_entry__.build()"
    d = str(tmp_path) + os.sep
    text = []
    for line in SHIPPED.splitlines():
        key = line.split("=")[0].strip()
        if key == "data_folder":
            line = f'data_folder = "{d}"'
        elif key == "n_timesteps":
            line = "n_timesteps = 1"
        elif key == "output_interval":
            line = "output_interval = 1"
        text.append(line)
    toml = os.path.join(d, "run.toml")
    with open(toml, "w") as f:
        f.write("\n".join(text) + "\n")
    r = subprocess.run([exe, toml], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    for s_ in ("Making grid!", "Iterating timestep: 1", "Residual is: ", "Number of iterations: ", "Number of stress iterations: "):
        assert s_ in r.stdout, s_
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "config1_step1.npz"))
    st = built.StandIn(1, 6, (-6 * A, -5 * A), (6 * A, 5 * A))
    key = {tuple(np.round(p, 9)): t for t, p in enumerate(st.support)}
    pc = read_piece(d + "phase_field_01.0.vtu")
    node = np.array([key[tuple(np.round(p[:2], 9))] for p in pc["points"]])
    for name in ("psi", "chi", "phi"):
        assert _rel(pc["point_data"][name], g[name][node]) <= 1e-9, name
    assert os.path.exists(d + "stress_00.0.vtu") and os.path.exists(d + "phase_field_rhs_01.0.vtu")
    st.close()


@pytest.mark.parametrize("degree,n_refines,sweeps", [(1, 6, 0), (2, 5, 0), (2, 7, 0), (1, 3, 2), (2, 2, 1), (1, 8, 4)])
def test_ring_smoother_and_selfchecks(built, degree, n_refines, sweeps):
    """The streaming Chebyshev sweep (csrc/pfc_ring.cuh) against the composition of the K-stencil and mass-stencil kernels;
    the K stencil against the cell-quadrature form; on meshes smaller than the ring as well (the ring then wraps around
    the period several times).  Then a step against the oracle: the smoother changes the preconditioner, not the answer."""
    om, st, s, Psi = _setup(built, degree, n_refines, operator_kind=1, preconditioner=1, precond_sweeps=sweeps)
    assert s.operator_info()[0] == built.OPERATOR_MATRIX_FREE
    assert s.selfcheck() <= 1e-12
    assert s.selfcheck_smoother() <= 1e-12
    if degree == 1:
        # closed-form Q1 kernels (csrc/pfc_ring_q1.cuh) against the general cell kernels, on a state with structure
        s.set_state(Psi + 1e-2 * np.random.default_rng(3).standard_normal(Psi.size), Psi)
        s.assemble()
        ej, ep = s.selfcheck_q1()
        print(f"Q1 closed-form kernels vs general: J apply {ej:.2e}, preconditioner {ep:.2e}")
        assert ej <= 1e-12 and ep <= 1e-10
    if n_refines <= 6:
        s.set_state(Psi, Psi)
        g = s.step()
        Pn, Pn1 = Psi.copy(), Psi.copy()
        o = om.newton_step(Pn, Pn1, DT, 1.0, EPS)
        assert g["status"] == 0 and o["status"] == 0 and _rel(s.get_state()[:om.N], Pn[:om.N]) <= 1e-9
        print(f"ring path Q{degree} 2^{n_refines} sweeps {sweeps}: gmres {g['gmres_iters']}")
    s.close()


# ---- parity at the sizes the bench lines are quoted on (VERDICT r01 "what's weak" 2) -------------------
def _bench_case(built, name):
    import bench as B
    import oracle as O
    O.set_threads(B.host_cores())
    wl = dict(B.WORKLOADS[name], name=name, custom=False)
    S, p1, p2 = B.box(wl["degree"], wl["n_refines"])
    A0 = built.default_A0(PSI0, EPS)
    st = built.StandIn(wl["degree"], wl["n_refines"], p1, p2)
    om = O.Mesh(wl["degree"], wl["n_refines"], p1, p2)
    Psi = B.initial_condition(st, wl, p1, p2, S, A0)
    assert np.array_equal(Psi, B.initial_condition(om, wl, p1, p2, S, A0))  # the two generators agree bit for bit
    return B, wl, st, om, Psi


@pytest.mark.parametrize("name", ["M2", "M4"])
def test_bench_size_rhs_and_operator_match_oracle(built, name):
    """-R and J x (random x) at the full bench size against the oracle's assembled monolithic CSR matrix, on the
    matrix-free lattice path and on the assembled SELL path: grids of thousands of CTAs, ring / lut handling at tile
    counts the small cases never reach.  Perturbed state so that Psi_n != Psi_n-1."""
    B, wl, st, om, Psi = _bench_case(built, name)
    N = om.N
    Psi1 = Psi + 1e-2 * np.random.default_rng(7).standard_normal(Psi.size)
    om.distribute(Psi1)
    ovals, orhs = om.assemble(Psi1, Psi, DT, 1.0, EPS, threads=B.host_cores())
    x = np.random.default_rng(3).standard_normal(3 * N)
    ref = om.spmv(ovals, x)
    free = om.master < 0
    for kw in ({"operator_kind": 1, "preconditioner": 1}, {"operator_kind": 0}):
        s = built.Solver(degree=wl["degree"], **kw)
        s.setup(st)
        s.set_state(Psi1, Psi)
        s.assemble()
        rhs = s.get_rhs()
        e_rhs = np.abs(rhs - orhs).max() / np.abs(orhs).max()
        y = s.apply_operator(x)
        e_op = np.abs(y - ref)[free].max() / np.abs(ref).max()
        e_nrm = abs(s.residual_norm() - np.linalg.norm(orhs)) / np.linalg.norm(orhs)
        print(f"{name} {kw}: -R max rel {e_rhs:.2e}, J x max rel {e_op:.2e}, ||R|| rel {e_nrm:.2e}")
        assert np.all(rhs[~free] == 0.0)
        assert e_rhs <= 1e-12 and e_op <= 1e-12 and e_nrm <= 1e-12
        s.close()


def test_bench_size_step_matches_oracle_M2(built):
    """psi after one time step at M2 (Q2 2^9, 3.15 M DoF): bench path (lattice kernels + block-Schur preconditioner)
    against the oracle's unpreconditioned reference algorithm (about 20 s of CPU)."""
    B, wl, st, om, Psi = _bench_case(built, "M2")
    N = om.N
    s = built.Solver(degree=2, operator_kind=1, preconditioner=1)
    s.setup(st)
    s.set_state(Psi, Psi)
    g = s.step()
    got = s.get_state()
    Pn, Pn1 = Psi.copy(), Psi.copy()
    o = om.newton_step(Pn, Pn1, DT, 1.0, EPS, threads=B.host_cores())
    err = _rel(got[:N], Pn[:N])
    print(f"M2 step 1: newton gpu {g['residuals']} oracle {o['residuals']}; gmres gpu {g['gmres_iters']} oracle {o['gmres_iters']}; "
          f"rel L2 psi {err:.2e}")
    assert g["status"] == 0 and o["status"] == 0 and g["n_assemblies"] == o["n_assemblies"]
    assert err <= 1e-9 and _rel(got, Pn) <= 1e-9
    s.close()
