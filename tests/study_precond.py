"""CPU study for the next round, not a test: how many GMRES(28) iterations does the block-Schur
right preconditioner need if the Chebyshev smoother works on a CONSTANT-COEFFICIENT stand-in of the psi Schur complement
(one composite node stencil on the GPU) instead of the exact, psi-dependent one (2 K products + 1 cell-quadrature Schur
row per sweep today)?  Uses the oracle's assembled blocks of the first Newton systems of a timestep.

    python tests/study_precond.py [degree] [n_refines] [sweeps]
"""
import os
import sys

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle as O  # noqa: E402

DT, TH, EPS, PSI0 = 0.1, 1.0, -0.8, -0.43


def blocks(om, vals):
    J = om.csr(vals).tocsr()
    N = om.N
    free = np.nonzero(om.master[:N] < 0)[0]
    return (lambda r, c: J[r * N + free][:, c * N + free].tocsr()), free


def cheb_coeffs(lmin, lmax, sweeps):
    theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)
    sigma = theta / delta
    rho = 1.0 / sigma
    c = [(1.0 / theta, 0.0)]
    for _ in range(1, sweeps):
        rn = 1.0 / (2.0 * sigma - rho)
        c.append((rn * rho, 2.0 * rn / delta))
        rho = rn
    return c


def bounds(apply_dinv_S, n, iters=40):
    """what precond_refresh does: power iteration for lambda_max, shifted power iteration for lambda_min"""
    rng = np.random.default_rng(0)
    x = rng.standard_normal(n)
    x /= np.linalg.norm(x)
    for _ in range(iters):
        y = apply_dinv_S(x)
        mu = np.linalg.norm(y)
        x = y / mu
    lam = 1.1 * mu
    x = rng.standard_normal(n)
    x /= np.linalg.norm(x)
    for _ in range(iters):
        y = lam * x - apply_dinv_S(x)
        mu2 = np.linalg.norm(y)
        x = y / mu2
    return max(0.85 * (lam - mu2), lam / 200.0), lam


def make_precond(B, C, D, K, M, sweeps, smoother):
    n = B.shape[0]
    iml = 1.0 / np.asarray(M.sum(axis=1)).ravel()
    Ml = sp.diags(iml)
    S_exact = (B - C @ Ml @ K + D @ Ml @ K @ Ml @ K).tocsr()
    dB, dM = B.diagonal(), M.diagonal()
    if smoother == "exact":
        S_s, dinv = S_exact, 1.0 / dB
    elif smoother == "const":       # B -> b M, C -> c M with the mean diagonal ratios; D M_L^-1 K M_L^-1 K is constant anyway
        b, c = np.mean(dB / dM), np.mean(C.diagonal() / dM)
        S_s = (b * M - c * (M @ Ml @ K) + D @ Ml @ K @ Ml @ K).tocsr()
        dinv = 1.0 / dB
    elif smoother == "const_diagB":  # as const, but the exact diagonal of B is kept (it is a by-product of the residual kernel)
        b, c = np.mean(dB / dM), np.mean(C.diagonal() / dM)
        S_s = (b * M - c * (M @ Ml @ K) + D @ Ml @ K @ Ml @ K).tocsr()
        S_s = (S_s + sp.diags(dB - b * dM)).tocsr()
        dinv = 1.0 / dB
    elif smoother == "stiff_only":   # S ~ M + D M_L^-1 K M_L^-1 K: no psi-dependent coefficient at all
        S_s = (M + D @ Ml @ K @ Ml @ K).tocsr()
        dinv = 1.0 / dB
    else:
        raise ValueError(smoother)
    lmin, lmax = bounds(lambda v: dinv * (S_s @ v), n)
    cc = cheb_coeffs(lmin, lmax, sweeps)

    def apply(r):
        r0, r1, r2 = r[:n], r[n:2 * n], r[2 * n:]
        m1 = iml * r1
        t = iml * (r2 - K @ m1)
        rhs = r0 - C @ m1 - D @ t
        d = cc[0][0] * dinv * rhs
        x = d.copy()
        for a, bb in cc[1:]:
            d = a * d + bb * dinv * (rhs - S_s @ x)
            x = x + d
        chi = iml * (r1 - K @ x)
        phi = iml * (r2 - K @ chi)
        return np.concatenate([x, chi, phi])
    return apply, (lmin, lmax)


def gmres_right(A, b, P, tol=1e-10, restart=28, maxit=2000):
    """right-preconditioned restarted GMRES, absolute tolerance on the true residual estimate; returns iterations"""
    x = np.zeros_like(b)
    its = 0
    while True:
        r = b - A @ x
        beta = np.linalg.norm(r)
        if beta <= tol or its >= maxit:
            return its, beta
        V = [r / beta]
        H = np.zeros((restart + 1, restart))
        g = np.zeros(restart + 1)
        g[0] = beta
        cs, sn = np.zeros(restart), np.zeros(restart)
        k_used = 0
        for k in range(restart):
            w = A @ P(V[k])
            for i in range(k + 1):
                H[i, k] = V[i] @ w
                w = w - H[i, k] * V[i]
            H[k + 1, k] = np.linalg.norm(w)
            V.append(w / H[k + 1, k])
            for i in range(k):
                H[i, k], H[i + 1, k] = cs[i] * H[i, k] + sn[i] * H[i + 1, k], -sn[i] * H[i, k] + cs[i] * H[i + 1, k]
            rr = np.hypot(H[k, k], H[k + 1, k])
            cs[k], sn[k] = H[k, k] / rr, H[k + 1, k] / rr
            H[k, k], H[k + 1, k] = rr, 0.0
            g[k + 1] = -sn[k] * g[k]
            g[k] = cs[k] * g[k]
            its += 1
            k_used = k + 1
            if abs(g[k + 1]) <= tol or its >= maxit:
                break
        y = np.linalg.solve(np.triu(H[:k_used, :k_used]), g[:k_used])
        x = x + P(sum(y[i] * V[i] for i in range(k_used)))


def main():
    degree = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    n_refines = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    sweeps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    a = O.LATTICE_A
    S = degree * 2 ** (n_refines - 6)  # box-scale rule of SURVEY.md §8d: the shipped file's node spacing
    p1, p2 = (-6 * a * S, -3 * np.sqrt(3.0) * a * S), (6 * a * S, 3 * np.sqrt(3.0) * a * S)
    om = O.Mesh(degree, n_refines, p1, p2)
    disl = ([[2 * a * S, 0], [-2 * a * S, 0]], [[a, 0], [-a, 0]]) if os.environ.get("DIPOLE") else ((), ())
    Pn = om.initial_condition(O.default_A0(PSI0, EPS), PSI0, *disl)
    Pn1 = Pn.copy()
    print(f"Q{degree} 2^{n_refines}: {om.n_dofs} dofs, sweeps {sweeps}")
    for newton in range(3):
        vals, rhs = om.assemble(Pn, Pn1, DT, TH, EPS)
        blk, free = blocks(om, vals)
        B, C, D, K, M = blk(0, 0), blk(0, 1), blk(0, 2), blk(1, 0), blk(1, 1)
        n = len(free)
        A = sp.bmat([[B, C, D], [K, M, None], [None, K, M]]).tocsr()
        b = np.concatenate([rhs[c * om.N + free] for c in range(3)])
        line = f"  Newton {newton}: |rhs| {np.linalg.norm(b):.2e}"
        for sm in ("exact", "const", "const_diagB", "stiff_only"):
            P, (lmin, lmax) = make_precond(B, C, D, K, M, sweeps, sm)
            its, res = gmres_right(A, b, P)
            line += f" | {sm}: {its} its (lambda {lmin:.2f}..{lmax:.2f})"
        print(line, flush=True)
        # advance Newton with an exact solve so that all variants see the same systems
        import scipy.sparse.linalg as spl
        dx = spl.spsolve(A.tocsc(), b)
        for c in range(3):
            Pn[c * om.N + free] += dx[c * n:(c + 1) * n]
        om.distribute(Pn)


if __name__ == "__main__":
    main()
