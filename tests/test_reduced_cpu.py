"""CPU checks of the psi-only reduced solve (PFC_PRECOND_SCHUR_REDUCED, csrc/pfc_reduced.cuh) against the oracle's
assembled blocks — the algebra the CUDA path relies on, with no GPU:

  * K and M commute on the periodic Q1 lattice (free nodes), so S M^2 = B M^2 + D K^2 - C K M with the Schur complement
    S = B + (D M^-1 K - C) M^-1 K of the serial class (phase_field_crystal_system.cpp:347);
  * the update (M^2 y, u - K M y, K^2 y - w) built from a solution of T y = F - C u + D w solves the 3-block Newton system
    (mpi.cpp:388-444 blocks) — rows 1 and 2 exactly, row 0 to the accuracy of y;
  * the symbol the library takes its Jacobi scaling and Chebyshev interval from (pfc_reduced_symbol, host only) is the
    diagonal and the spectrum of the assembled S~.
"""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spl

A = 4 * np.pi / np.sqrt(3.0)
DT, EPS, PSI0 = 0.1, -0.8, -0.43


def _blocks(O, n_refines, theta, p1, p2, perturb=True):
    om = O.Mesh(1, n_refines, p1, p2)
    Psi = om.initial_condition(O.default_A0(PSI0, EPS), PSI0, [[2 * A, 0], [-2 * A, 0]], [[A, 0], [-A, 0]])
    Psi1 = Psi.copy()
    if perturb:
        Psi1 = Psi + 1e-2 * np.random.default_rng(5).standard_normal(Psi.size)
        om.distribute(Psi1)
    vals, rhs = om.assemble(Psi1, Psi, DT, theta, EPS)
    J = om.csr(vals).tocsr()
    N = om.N
    free = np.nonzero(om.master[:N] < 0)[0]
    blk = lambda r, c: J[r * N + free][:, c * N + free].tocsr()  # noqa: E731
    return om, free, blk, [rhs[c * N + free] for c in range(3)]


@pytest.fixture(scope="module")
def oracle_mod():
    import oracle as O
    return O


@pytest.mark.parametrize("n_refines,theta,box", [(4, 1.0, "test"), (5, 0.5, "test"), (5, 1.0, "bench")])
def test_reduced_operator_is_schur_complement_times_mass_squared(oracle_mod, n_refines, theta, box):
    O = oracle_mod
    if box == "test":
        p1, p2 = (-6 * A, -5 * A), (6 * A, 5 * A)
    else:  # the bench's commensurate box at the shipped file's node spacing
        S = 2.0 ** (n_refines - 6)
        p1, p2 = (-6 * A * S, -3 * np.sqrt(3.0) * A * S), (6 * A * S, 3 * np.sqrt(3.0) * A * S)
    om, free, blk, (F, G, H) = _blocks(O, n_refines, theta, p1, p2)
    B, C, D, K, M = blk(0, 0), blk(0, 1), blk(0, 2), blk(1, 0), blk(1, 1)
    assert abs(blk(2, 1) - K).max() == 0 and abs(blk(2, 2) - M).max() == 0          # L_chi = L_psi = K, M_phi = M_chi = M
    assert abs(D - DT * theta * (K - 2.0 * M)).max() <= 1e-14 * abs(D).max()        # D = dt theta (K - 2 M)
    comm = (K @ M - M @ K)
    assert abs(comm).max() <= 1e-13 * abs(K).max() * abs(M).max()                   # circulant blocks commute
    n = len(free)
    rng = np.random.default_rng(1)
    y = rng.standard_normal(n)
    Mi = spl.factorized(M.tocsc())
    x = M @ (M @ y)
    Sx = B @ x + D @ Mi(K @ Mi(K @ x)) - C @ Mi(K @ x)                              # serial.cpp:347 applied to M^2 y
    Ty = B @ (M @ (M @ y)) + D @ (K @ (K @ y)) - C @ (K @ (M @ y))
    assert np.linalg.norm(Sx - Ty) <= 1e-11 * np.linalg.norm(Ty)
    # the update built from the exact solution of T y = f solves the 3-block system
    u = Mi(G)
    w = Mi(K @ u - H)
    f = F - C @ u + D @ w
    T = (B @ M @ M + D @ K @ K - C @ K @ M).tocsc()
    ys = spl.spsolve(T, f)
    d = np.concatenate([M @ (M @ ys), u - K @ (M @ ys), K @ (K @ ys) - w])
    J3 = sp.bmat([[B, C, D], [K, M, None], [None, K, M]]).tocsr()
    b3 = np.concatenate([F, G, H])
    r = b3 - J3 @ d
    assert np.linalg.norm(r) <= 1e-11 * np.linalg.norm(b3)
    direct = spl.spsolve(J3.tocsc(), b3)
    assert np.linalg.norm(d - direct) <= 1e-8 * np.linalg.norm(direct)


@pytest.mark.parametrize("n_refines,theta", [(4, 1.0), (5, 0.5)])
def test_symbol_matches_assembled_schur_stand_in(built, oracle_mod, n_refines, theta):
    O = oracle_mod
    p1, p2 = (-6 * A, -5 * A), (6 * A, 5 * A)
    om, free, blk, _ = _blocks(O, n_refines, theta, p1, p2, perturb=False)
    K, M = blk(1, 0), blk(1, 1)
    D = DT * theta * (K - 2.0 * M)
    iml = sp.diags(1.0 / np.asarray(M.sum(axis=1)).ravel())
    St = (M + D @ iml @ K @ iml @ K).toarray()
    hx, hy = (p2[0] - p1[0]) / 2 ** n_refines, (p2[1] - p1[1]) / 2 ** n_refines
    diag, lmin, lmax = built.reduced_symbol(hx, hy, DT * theta)
    assert np.allclose(np.diag(St), diag, rtol=1e-12, atol=0)
    ev = np.linalg.eigvalsh(St) / diag
    # the lattice's frequencies are a subset of the torus the symbol is sampled on (96 x 96 angles)
    assert lmin * (1 - 2e-3) <= ev.min() and ev.max() <= lmax * (1 + 2e-3)
    assert lmin > 0 and ev.min() <= 1.35 * lmin and ev.max() >= 0.9 * lmax


def test_symbol_rejects_bad_arguments(built):
    with pytest.raises(ValueError):
        built.reduced_symbol(0.0, 1.0, 0.1)


def _serial_class_solve(B, C, D, K, M, F, G, H):
    """solve_and_update of the reference's serial class, restated line by line with scipy
    (phase_field_crystal_system.cpp:319-369): M_chi_inv / M_phi_inv = CG with Jacobi and ReductionControl(2000, 1e-18, 1e-10)
    (:321-339), psi_matrix = B + (D M_phi_inv L_chi - C) M_chi_inv L_psi (:347), psi_rhs (:353-355), GMRES with
    SolverControl(500) (default tolerance 1e-10) and PreconditionIdentity (:357-359), back substitution (:361-365).
    L_psi = L_chi = K and M_chi = M_phi = M on this mesh.  TEST INFRASTRUCTURE."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from study_reduced import Op, gmres_right_x
    jac = spl.LinearOperator(M.shape, matvec=lambda v, d=M.diagonal(): v / d)

    def m_inv(b):
        if not np.any(b):
            return np.zeros_like(b)
        x, info = spl.cg(M, b, rtol=1e-10, atol=1e-18, maxiter=2000, M=jac)
        assert info == 0
        return x
    psi_matrix = Op(lambda v: B @ v + D @ m_inv(K @ m_inv(K @ v)) - C @ m_inv(K @ v))
    psi_rhs = F - C @ m_inv(G) + D @ m_inv(K @ m_inv(G) - H)
    dpsi, its, res = gmres_right_x(psi_matrix, psi_rhs, lambda v: v, tol=1e-10, restart=28, maxit=500)
    dchi = m_inv(G - K @ dpsi)
    dphi = m_inv(H - K @ dchi)
    return np.concatenate([dpsi, dchi, dphi]), its, res


def test_reduced_update_equals_the_serial_class_solve(oracle_mod):
    """The update of the reduced formulation (dpsi = M^2 y, no mass solve inside the operator) against the serial class's own
    Schur solve with its CG mass inverses, on the same Newton system (first iteration of a step: G, H != 0)."""
    O = oracle_mod
    p1, p2 = (-6 * A, -5 * A), (6 * A, 5 * A)
    om, free, blk, (F, G, H) = _blocks(O, 4, 1.0, p1, p2, perturb=False)
    B, C, D, K, M = blk(0, 0), blk(0, 1), blk(0, 2), blk(1, 0), blk(1, 1)
    assert np.linalg.norm(G) > 1e-3  # chi = phi = 0 in the initial condition: the auxiliary rows are violated
    ref, its, res = _serial_class_solve(B, C, D, K, M, F, G, H)
    assert res <= 1e-10 and its < 500
    Mi = spl.factorized(M.tocsc())
    u = Mi(G)
    w = Mi(K @ u - H)
    T = (B @ M @ M + D @ K @ K - C @ K @ M).tocsc()
    y = spl.spsolve(T, F - C @ u + D @ w)
    mine = np.concatenate([M @ (M @ y), u - K @ (M @ y), K @ (K @ y) - w])
    n = len(free)
    assert np.linalg.norm(mine[:n] - ref[:n]) <= 1e-7 * np.linalg.norm(ref[:n])
    assert np.linalg.norm(mine - ref) <= 1e-7 * np.linalg.norm(ref)
