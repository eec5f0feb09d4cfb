"""ctypes binding of the CPU oracle (oracle/pfc_oracle.c).

TEST INFRASTRUCTURE ONLY — importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / ``--impl reference`` legs, never from the product package.
PARITY UNPINNED: see oracle/pfc_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libpfc_oracle.so")

# lattice constant a = 4 pi / sqrt(3)   (hexagonal_lattice.hpp:35)
LATTICE_A = 4 * math.pi * 0.5773502691896257645091487805019574556476017


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (gcc, OpenMP)."""
    src = [os.path.join(_HERE, f) for f in ("pfc_oracle.c", "pfc_oracle.h", "Makefile")]
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= max(os.path.getmtime(s) for s in src)):
        return _LIB_PATH
    subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return _LIB_PATH


class _Mesh(C.Structure):
    _fields_ = [
        ("degree", C.c_int32), ("n_refines", C.c_int32), ("nx", C.c_int32), ("ny", C.c_int32),
        ("n_s", C.c_int32), ("dofs_per_cell", C.c_int32), ("n_q", C.c_int32),
        ("n_cells", C.c_int64), ("N", C.c_int64), ("n_dofs", C.c_int64),
        ("n_constrained", C.c_int64), ("nnz", C.c_int64),
        ("x0", C.c_double), ("y0", C.c_double), ("x1", C.c_double), ("y1", C.c_double),
        ("hx", C.c_double), ("hy", C.c_double),
        ("cell_dofs", C.POINTER(C.c_int32)), ("cell_vertices", C.POINTER(C.c_double)),
        ("master", C.POINTER(C.c_int32)), ("support", C.POINTER(C.c_double)),
        ("shape", C.POINTER(C.c_double)), ("dshape", C.POINTER(C.c_double)),
        ("qw", C.POINTER(C.c_double)),
        ("local_comp", C.POINTER(C.c_int32)), ("local_shape", C.POINTER(C.c_int32)),
        ("row_ptr", C.POINTER(C.c_int64)), ("col_idx", C.POINTER(C.c_int32)),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        dp = C.POINTER(C.c_double)
        L.orc_mesh_create.restype = C.POINTER(_Mesh)
        L.orc_mesh_create.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double]
        L.orc_mesh_destroy.argtypes = [C.POINTER(_Mesh)]
        L.orc_initial_condition.argtypes = [C.POINTER(_Mesh), C.c_double, C.c_double, C.c_int, dp, dp, dp]
        L.orc_polycrystal.argtypes = [C.POINTER(_Mesh), C.c_double, C.c_double, C.c_int, dp, dp, dp, dp]
        L.orc_assemble.argtypes = [C.POINTER(_Mesh), dp, dp, C.c_double, C.c_double, C.c_double, dp, dp,
                                   C.c_int, C.c_int]
        L.orc_spmv.argtypes = [C.POINTER(_Mesh), dp, dp, dp]
        L.orc_gmres.restype = C.c_int64
        L.orc_gmres.argtypes = [C.POINTER(_Mesh), dp, dp, dp, C.c_double, C.c_int64, C.c_int,
                                C.POINTER(C.c_int), dp, C.c_int64]
        L.orc_newton_step.restype = C.c_int
        L.orc_newton_step.argtypes = [C.POINTER(_Mesh), dp, dp, C.c_double, C.c_double, C.c_double,
                                      C.c_double, C.c_int, C.c_double, C.c_int, dp,
                                      C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_distribute.argtypes = [C.POINTER(_Mesh), dp]
        L.orc_stress.restype = C.c_int
        L.orc_stress.argtypes = [C.POINTER(_Mesh), dp, C.c_double, C.c_double, dp, C.POINTER(C.c_int64), dp]
        L.orc_max_threads.restype = C.c_int
        L.orc_set_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def default_A0(psi0: float, eps: float) -> float:
    """A_0 = "default" (phase_field_crystal_simulation_mpi.cpp:87-88)."""
    return 0.2 * (abs(psi0) + (1.0 / 3.0) * math.sqrt(-15 * eps - 36 * psi0 * psi0))


class Mesh:
    """Uniform periodic rectangle, deal.II enumeration (mpi.cpp:115-274)."""

    def __init__(self, degree, n_refines, p1, p2):
        self._p = lib().orc_mesh_create(degree, n_refines, p1[0], p1[1], p2[0], p2[1])
        if not self._p:
            raise ValueError("unsupported degree / n_refines")
        m = self._p.contents
        for k in ("degree", "n_refines", "nx", "ny", "n_s", "dofs_per_cell", "n_q", "n_cells", "N",
                  "n_dofs", "n_constrained", "nnz", "x0", "y0", "x1", "y1", "hx", "hy"):
            setattr(self, k, getattr(m, k))

        def arr(ptr, shape):
            return np.ctypeslib.as_array(ptr, shape=shape)

        self.cell_dofs = arr(m.cell_dofs, (self.n_cells, self.dofs_per_cell))
        self.cell_vertices = arr(m.cell_vertices, (self.n_cells, 4, 2))
        self.master = arr(m.master, (self.n_dofs,))
        self.support = arr(m.support, (self.N, 2))
        self.shape = arr(m.shape, (self.n_q, self.n_s))
        self.dshape = arr(m.dshape, (self.n_q, self.n_s, 2))
        self.qw = arr(m.qw, (self.n_q,))
        self.local_comp = arr(m.local_comp, (self.dofs_per_cell,))
        self.local_shape = arr(m.local_shape, (self.dofs_per_cell,))
        self.row_ptr = arr(m.row_ptr, (self.n_dofs + 1,))
        self.col_idx = arr(m.col_idx, (self.nnz,))

    def close(self):
        if self._p:
            lib().orc_mesh_destroy(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- reference-path functions -------------------------------------------------------
    def initial_condition(self, A0, psi0, disl_pos=(), burgers=()):
        pos = np.ascontiguousarray(np.asarray(disl_pos, dtype=np.float64).reshape(-1, 2))
        bur = np.ascontiguousarray(np.asarray(burgers, dtype=np.float64).reshape(-1, 2))
        Psi = np.empty(self.n_dofs)
        lib().orc_initial_condition(self._p, A0, psi0, len(pos), _dp(pos), _dp(bur), _dp(Psi))
        return Psi

    def polycrystal(self, A0, psi0, seeds, alpha, offset):
        """harness extension (BASELINE.json config 4): periodic Voronoi grains of rotated hexagonal lattices"""
        sd = np.ascontiguousarray(np.asarray(seeds, dtype=np.float64).reshape(-1, 2))
        al = np.ascontiguousarray(np.asarray(alpha, dtype=np.float64).reshape(-1))
        of = np.ascontiguousarray(np.asarray(offset, dtype=np.float64).reshape(-1, 2))
        Psi = np.empty(self.n_dofs)
        lib().orc_polycrystal(self._p, A0, psi0, len(al), _dp(sd), _dp(al), _dp(of), _dp(Psi))
        return Psi

    def assemble(self, Psi_n, Psi_n1, dt, theta, eps, threads=1, dead_work=False):
        vals = np.empty(self.nnz)
        rhs = np.empty(self.n_dofs)
        lib().orc_assemble(self._p, _dp(np.ascontiguousarray(Psi_n)), _dp(np.ascontiguousarray(Psi_n1)),
                           dt, theta, eps, _dp(vals), _dp(rhs), threads, int(dead_work))
        return vals, rhs

    def spmv(self, vals, x):
        y = np.empty(self.n_dofs)
        lib().orc_spmv(self._p, _dp(vals), _dp(np.ascontiguousarray(x)), _dp(y))
        return y

    def gmres(self, vals, b, tol=1e-10, max_iters=None, n_tmp_vectors=30, stop_after=0):
        x = np.empty(self.n_dofs)
        conv = C.c_int(0)
        res = C.c_double(0)
        its = lib().orc_gmres(self._p, _dp(vals), _dp(np.ascontiguousarray(b)), _dp(x), tol,
                              self.n_dofs if max_iters is None else max_iters, n_tmp_vectors,
                              C.byref(conv), C.byref(res), stop_after)
        return x, int(its), bool(conv.value), res.value

    def newton_step(self, Psi_n, Psi_n1, dt, theta, eps, newton_tol=1e-8, newton_max=200,
                    gmres_tol=1e-10, threads=1):
        """One iterate_timestep (mpi.cpp:576-593); Psi_n, Psi_n1 are updated in place."""
        assert Psi_n.flags.c_contiguous and Psi_n1.flags.c_contiguous
        res = np.zeros(newton_max + 1)
        gm = np.zeros(newton_max, dtype=np.int64)
        nasm = C.c_int(0)
        status = C.c_int(0)
        ns = lib().orc_newton_step(self._p, _dp(Psi_n), _dp(Psi_n1), dt, theta, eps, newton_tol,
                                   newton_max, gmres_tol, threads, _dp(res),
                                   gm.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(nasm),
                                   C.byref(status))
        return {"gmres_iters": gm[:ns].tolist(), "residuals": res[:nasm.value].tolist(),
                "n_assemblies": nasm.value, "status": status.value}

    def stress(self, Psi, eps, tol=1e-10):
        """StressCalculatorMPI::calculate_stress (stress.cpp:345-352): [4][N] nodal stress, c = a*2 + b."""
        out = np.zeros((4, self.N))
        rhs = np.zeros((4, self.N))
        its = np.zeros(4, dtype=np.int64)
        st = lib().orc_stress(self._p, _dp(np.ascontiguousarray(Psi)), eps, tol, _dp(out),
                              its.ctypes.data_as(C.POINTER(C.c_int64)), _dp(rhs))
        return {"stress": out, "rhs": rhs, "cg_iters": its.tolist(), "status": st}

    def distribute(self, v):
        lib().orc_distribute(self._p, _dp(v))
        return v

    def csr(self, vals):
        import scipy.sparse as sp
        return sp.csr_matrix((vals, self.col_idx, self.row_ptr), shape=(self.n_dofs, self.n_dofs))


def max_threads():
    return lib().orc_max_threads()


def set_threads(n):
    lib().orc_set_threads(n)
