"""numpy emulation of what the CUDA kernels do with the host plan (TEST INFRASTRUCTURE).

Used by the CPU tests to check the host side of pfc_setup (SELL layout, permutation, colouring,
scatter tables, halo lists) against the oracle without a GPU.  The arithmetic mirrors
csrc/pfc_kernels.cuh statement by statement; it is not a product path.
"""
import numpy as np


def vaddr(u, c, Np):
    u = np.asarray(u, dtype=np.int64)
    return np.where(u < Np, c * Np + u, 3 * u + c)


def to_internal(plan, x_api, N):
    """API vector [3N] -> internal vector [Lv] (owned + ghost)."""
    v = np.zeros(plan.Lv)
    own = plan.perm >= 0
    rows = np.nonzero(own)[0]
    for c in range(3):
        v[c * plan.Np + rows] = x_api[c * N + plan.perm[rows]]
        if plan.n_ghost:
            g = np.arange(plan.n_ghost)
            v[3 * (plan.Np + g) + c] = x_api[c * N + plan.ghost_global]
    return v


def owned_to_api(plan, v, N, out):
    rows = np.nonzero(plan.perm >= 0)[0]
    for c in range(3):
        out[c * N + plan.perm[rows]] = v[c * plan.Np + rows]
    return out


def local_blocks(plan, tab, psi_n, psi_n1, dt, theta, eps):
    """Per-cell local B, C, M, K [cells, ns, ns] and rhs [cells, 3, ns] (assemble_kernel)."""
    shape, dshape, qw = tab
    ns = plan.n_s
    hx, hy = plan.cell_h[:, 0], plan.cell_h[:, 1]
    gx = dshape[None, :, :, 0] / hx[:, None, None]          # [cells, q, a]
    gy = dshape[None, :, :, 1] / hy[:, None, None]
    JxW = qw[None, :] * (hx * hy)[:, None]                   # [cells, q]
    u = np.empty((plan.n_cells, 6, ns))
    for c in range(3):
        ad = vaddr(plan.cell_gather, c, plan.Np)
        u[:, c, :] = psi_n[ad]
        u[:, 3 + c, :] = psi_n1[ad]
    v = np.einsum("qa,kfa->kfq", shape, u)
    dx = np.einsum("kqa,kfa->kfq", gx, u)
    dy = np.einsum("kqa,kfa->kfq", gy, u)
    psi, chi, phi, psi1, chi1, phi1 = (v[:, i] for i in range(6))
    g2 = dx[:, 0] ** 2 + dy[:, 0] ** 2
    g21 = dx[:, 3] ** 2 + dy[:, 3] ** 2
    dtth, dt1 = dt * theta, dt * (1 - theta)
    f0 = (-psi + psi1 + dtth * (2 * phi + (1 + eps) * chi + 3 * psi * psi * chi + 6 * psi * g2)
          + dt1 * (2 * phi1 + (1 + eps) * chi1 + 3 * psi1 * psi1 * chi1 + 6 * psi1 * g21)) * JxW
    g0x = dt * (theta * dx[:, 2] + (1 - theta) * dx[:, 5]) * JxW
    g0y = dt * (theta * dy[:, 2] + (1 - theta) * dy[:, 5]) * JxW
    N = shape[None]                                           # [1, q, a]
    rhs = np.empty((plan.n_cells, 3, ns))
    rhs[:, 0] = np.einsum("kqa,kq->ka", N + 0 * gx, f0) - np.einsum("kqa,kq->ka", gx, g0x) - np.einsum("kqa,kq->ka", gy, g0y)
    rhs[:, 1] = np.einsum("kqa,kq->ka", N + 0 * gx, -chi * JxW) - np.einsum("kqa,kq->ka", gx, dx[:, 0] * JxW) - np.einsum("kqa,kq->ka", gy, dy[:, 0] * JxW)
    rhs[:, 2] = np.einsum("kqa,kq->ka", N + 0 * gx, -phi * JxW) - np.einsum("kqa,kq->ka", gx, dx[:, 1] * JxW) - np.einsum("kqa,kq->ka", gy, dy[:, 1] * JxW)
    bm = (1.0 - dtth * (6 * psi * chi + 6 * g2)) * JxW
    bx = -12.0 * dtth * psi * dx[:, 0] * JxW
    by = -12.0 * dtth * psi * dy[:, 0] * JxW
    cm = -dtth * ((1 + eps) + 3 * psi * psi) * JxW
    Nf = np.broadcast_to(N, gx.shape)
    B = np.einsum("kqi,kq,kqj->kij", Nf, bm, Nf) + np.einsum("kqi,kq,kqj->kij", Nf, bx, gx) + np.einsum("kqi,kq,kqj->kij", Nf, by, gy)
    Cm = np.einsum("kqi,kq,kqj->kij", Nf, cm, Nf)
    M = np.einsum("kqi,kq,kqj->kij", Nf, JxW, Nf)
    K = np.einsum("kqi,kq,kqj->kij", gx, JxW, gx) + np.einsum("kqi,kq,kqj->kij", gy, JxW, gy)
    return B, Cm, M, K, rhs


def scatter(plan, B, Cm, M, K, rhs):
    """Colour by colour read-modify-write, as the kernel does; asserts conflict freedom."""
    vB, vC, vM, vK = (np.zeros(plan.n_ent) for _ in range(4))
    r = np.zeros(plan.Lv)
    ns = plan.n_s
    for c in range(plan.n_colors):
        b, e = plan.color_ptr[c], plan.color_ptr[c + 1]
        pos = plan.cell_pos[b:e].reshape(-1)
        ok = pos >= 0
        assert np.unique(pos[ok]).size == ok.sum(), "colour has a scatter conflict"
        for dst, src in ((vB, B), (vC, Cm), (vM, M), (vK, K)):
            dst[pos[ok]] += src[b:e].reshape(-1)[ok]
        row = plan.cell_row[b:e].reshape(-1)
        okr = row >= 0
        assert np.unique(row[okr]).size == okr.sum(), "colour has a rhs conflict"
        for comp in range(3):
            r[comp * plan.Np + row[okr]] += rhs[b:e, comp].reshape(-1)[okr]
        sp = plan.cell_spos[b:e].reshape(-1)
        oks = sp >= 0
        assert np.unique(sp[oks]).size == oks.sum()
        idx = np.arange(ns)
        vB[sp[oks]] += np.abs(B[b:e][:, idx, idx]).reshape(-1)[oks]
        vM[sp[oks]] += np.abs(M[b:e][:, idx, idx]).reshape(-1)[oks]
    return vB, vC, vM, vK, r


def spmv(plan, vB, vC, vM, vK, x, dtth):
    """spmv_sell_kernel on the internal vector x [Lv]; returns y [Lv] (owned part filled)."""
    Np = plan.Np
    y = np.zeros(plan.Lv)
    for s in range(plan.n_slices):
        base = plan.slice_ptr[s]
        width = (plan.slice_ptr[s + 1] - base) // 32
        rows = s * 32 + np.arange(32)
        y0 = np.zeros(32); y1 = np.zeros(32); y2 = np.zeros(32)
        for k in range(width):
            p = base + k * 32 + np.arange(32)
            c = plan.col[p].astype(np.int64)
            slave = c < 0
            cc = np.where(slave, ~c, c)
            x0, x1, x2 = x[vaddr(cc, 0, Np)], x[vaddr(cc, 1, Np)], x[vaddr(cc, 2, Np)]
            b, cm, m, kk = vB[p], vC[p], vM[p], vK[p]
            y0 += np.where(slave, b * x0, b * x0 + cm * x1 + dtth * (kk - 2.0 * m) * x2)
            y1 += np.where(slave, m * x1, kk * x0 + m * x1)
            y2 += np.where(slave, m * x2, kk * x1 + m * x2)
        y[rows] = y0; y[Np + rows] = y1; y[2 * Np + rows] = y2
    return y


def spmv_fast(plan, vB, vC, vM, vK, x, dtth):
    """Vectorised version of spmv() for larger meshes (same arithmetic, entry-parallel)."""
    Np = plan.Np
    ent = np.arange(plan.n_ent)
    sl = np.searchsorted(plan.slice_ptr, ent, side="right") - 1
    rows = sl * 32 + (ent - plan.slice_ptr[sl]) % 32
    c = plan.col.astype(np.int64)
    slave = c < 0
    cc = np.where(slave, ~c, c)
    x0, x1, x2 = x[vaddr(cc, 0, Np)], x[vaddr(cc, 1, Np)], x[vaddr(cc, 2, Np)]
    t0 = np.where(slave, vB * x0, vB * x0 + vC * x1 + dtth * (vK - 2.0 * vM) * x2)
    t1 = np.where(slave, vM * x1, vK * x0 + vM * x1)
    t2 = np.where(slave, vM * x2, vK * x1 + vM * x2)
    y = np.zeros(plan.Lv)
    np.add.at(y, rows, t0)
    np.add.at(y, Np + rows, t1)
    np.add.at(y, 2 * Np + rows, t2)
    return y
