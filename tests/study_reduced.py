"""CPU study, not a test: the psi-only reduced Newton system WITHOUT inner mass solves.

The reference's serial class solves  S dpsi = f,  S = B + (D M^-1 K - C) M^-1 K  (phase_field_crystal_system.cpp:319-369),
with CG mass inverses inside every operator application.  On the uniform periodic Q1 lattice K and M are circulant and
commute, so with dpsi = M^2 y:

    T y = (B M^2 + D K^2 - C K M) y = f,      f = F - C u + D w,  u = M^-1 G,  w = M^-1 (K u - H)
    dpsi = M^2 y,   dchi = u - K M y,   dphi = K^2 y - w

needs no inverse per Krylov iteration (u, w: two mass solves per Newton iteration, and G = H = 0 to solver tolerance after
the first step because the auxiliary equations are linear).  The symbol of T is m^3 b + dt theta (k - 2 m) k^2 + ...: its
stiff part EQUALS that of the constant-coefficient stand-in S~ = M + D M_L^-1 K M_L^-1 K the Chebyshev smoother already
works on, so Cheb(S~) preconditions T better than the block elimination with lumped masses preconditions J.

    python tests/study_reduced.py [n_refines] [sweeps]      (Q1)
Prints GMRES(28) iterations to 1e-10 (absolute, true residual at restarts) of (a) today's 3-block solve with the
block-Schur preconditioner and (b) the reduced solve, for the Newton systems of the first time step, and the
difference of the two Newton updates.
"""
import os
import sys

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spl

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from study_precond import DT, EPS, PSI0, TH, O, blocks, bounds, cheb_coeffs, gmres_right, make_precond  # noqa: E402


class Op:
    def __init__(self, f):
        self.f = f

    def __matmul__(self, v):
        return self.f(v)


def main():
    n_refines = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    sweeps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    a = O.LATTICE_A
    S = 2 ** (n_refines - 6)
    p1, p2 = (-6 * a * S, -3 * np.sqrt(3.0) * a * S), (6 * a * S, 3 * np.sqrt(3.0) * a * S)
    om = O.Mesh(1, n_refines, p1, p2)
    disl = ([[2 * a * S, 0], [-2 * a * S, 0]], [[a, 0], [-a, 0]]) if os.environ.get("DIPOLE") else ((), ())
    Pn = om.initial_condition(O.default_A0(PSI0, EPS), PSI0, *disl)
    Pn1 = Pn.copy()
    print(f"Q1 2^{n_refines}: {om.n_dofs} dofs, sweeps {sweeps}")
    for step in range(2):
        for newton in range(4):
            vals, rhs = om.assemble(Pn, Pn1, DT, TH, EPS)
            blk, free = blocks(om, vals)
            B, C, D, K, M = blk(0, 0), blk(0, 1), blk(0, 2), blk(1, 0), blk(1, 1)
            n = len(free)
            F, G, H = (rhs[c * om.N + free] for c in range(3))
            b3 = np.concatenate([F, G, H])
            if np.linalg.norm(b3) < 1e-8:
                break
            A = sp.bmat([[B, C, D], [K, M, None], [None, K, M]]).tocsr()
            P3, _ = make_precond(B, C, D, K, M, sweeps, "stiff_only")
            its3, _ = gmres_right(A, b3, P3)
            # reduced system
            Mi = spl.factorized(M.tocsc())
            u = Mi(G)
            w = Mi(K @ u - H)
            f = F - C @ u + D @ w
            T = Op(lambda y: B @ (M @ (M @ y)) + D @ (K @ (K @ y)) - C @ (K @ (M @ y)))
            iml = 1.0 / np.asarray(M.sum(axis=1)).ravel()
            Ml = sp.diags(iml)
            out = f"  step {step} Newton {newton}: |rhs| {np.linalg.norm(b3):.2e} |G| {np.linalg.norm(G):.1e} |H| {np.linalg.norm(H):.1e} | 3-block: {its3} its"
            for name, St in (("S~", (M + D @ Ml @ K @ Ml @ K).tocsr()),
                             ("S~3", (M @ Ml @ M @ Ml @ M + D @ Ml @ K @ Ml @ K).tocsr())):
                dinv = 1.0 / St.diagonal()
                lmin, lmax = bounds(lambda v: dinv * (St @ v), n)
                for sw in (sweeps, sweeps - 1, sweeps + 2):
                    cc = cheb_coeffs(lmin, lmax, sw)

                    def Q(r, cc=cc, dinv=dinv, St=St):
                        d = cc[0][0] * dinv * r
                        x = d.copy()
                        for a_, b_ in cc[1:]:
                            d = a_ * d + b_ * dinv * (r - St @ x)
                            x = x + d
                        return x
                    y, its, res = gmres_right_x(T, f, Q)
                    out += f" | {name} sw{sw}: {its}"
            # Newton update of the reduced solve against the exact one
            dpsi = M @ (M @ y)
            dchi = u - K @ (M @ y)
            dphi = K @ (K @ y) - w
            dx = spl.spsolve(A.tocsc(), b3)
            err = np.linalg.norm(np.concatenate([dpsi, dchi, dphi]) - dx) / np.linalg.norm(dx)
            r3 = np.linalg.norm(A @ np.concatenate([dpsi, dchi, dphi]) - b3)
            print(out + f" | update vs direct solve {err:.1e}, 3-block residual {r3:.1e}", flush=True)
            for c in range(3):
                Pn[c * om.N + free] += dx[c * n:(c + 1) * n]
            om.distribute(Pn)
        Pn1 = Pn.copy()


def gmres_right_x(A, b, P, tol=1e-10, restart=28, maxit=2000):
    """study_precond.gmres_right, returning the solution as well"""
    x = np.zeros_like(b)
    its = 0
    while True:
        r = b - A @ x
        beta = np.linalg.norm(r)
        if beta <= tol or its >= maxit:
            return x, its, beta
        V = [r / beta]
        Hm = np.zeros((restart + 1, restart))
        g = np.zeros(restart + 1)
        g[0] = beta
        cs, sn = np.zeros(restart), np.zeros(restart)
        k_used = 0
        for k in range(restart):
            w = A @ P(V[k])
            for i in range(k + 1):
                Hm[i, k] = V[i] @ w
                w = w - Hm[i, k] * V[i]
            Hm[k + 1, k] = np.linalg.norm(w)
            V.append(w / Hm[k + 1, k])
            for i in range(k):
                Hm[i, k], Hm[i + 1, k] = cs[i] * Hm[i, k] + sn[i] * Hm[i + 1, k], -sn[i] * Hm[i, k] + cs[i] * Hm[i + 1, k]
            rr = np.hypot(Hm[k, k], Hm[k + 1, k])
            cs[k], sn[k] = Hm[k, k] / rr, Hm[k + 1, k] / rr
            Hm[k, k], Hm[k + 1, k] = rr, 0.0
            g[k + 1] = -sn[k] * g[k]
            g[k] = cs[k] * g[k]
            its += 1
            k_used = k + 1
            if abs(g[k + 1]) <= tol or its >= maxit:
                break
        yk = np.linalg.solve(np.triu(Hm[:k_used, :k_used]), g[:k_used])
        x = x + P(sum(yk[i] * V[i] for i in range(k_used)))


if __name__ == "__main__":
    main()
