"""pfc_b200 — Python harness over the C ABI of ``include/pfc_b200.h`` (ctypes).

The product is ``libpfc_b200.so`` (hand-written sm_100a CUDA kernels behind a C ABI); this
module only binds it for tests, the benchmark and the Python mirror of the reference's
``PhaseFieldCrystalSystemMPI`` interface (``system.py``).  It never computes anything itself:
without the built library, or without a GPU, every compute call raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpfc_b200.so")

PFC_OK = 0
PFC_ERR_BAD_ARGUMENT = 1
PFC_ERR_CUDA = 2
PFC_ERR_NCCL = 3
PFC_ERR_GMRES_NOT_CONVERGED = 4
PFC_ERR_NONFINITE = 5
PFC_WARN_NEWTON_MAX_ITERS = 6
PFC_ERR_UNSUPPORTED = 7
PFC_ERR_STATE = 8
OPERATOR_ASSEMBLED, OPERATOR_MATRIX_FREE, OPERATOR_MATRIX_FREE_GENERAL = 0, 1, 2
PLAN_LATTICE, PLAN_SKIP_TABLES = 1, 2
ORTHO_CGS, ORTHO_CGS2, ORTHO_MGS = 0, 1, 2
PRECOND_NONE, PRECOND_BLOCK_SCHUR, PRECOND_SCHUR_REDUCED = 0, 1, 2
MAX_NEWTON_RECORD = 256
NOALLOC_MAX_N = 134217728  # PFC_NOALLOC_MAX_N of include/pfc_b200.h

# lattice constant a = 4 pi / sqrt(3) (src/phase_field_functions/hexagonal_lattice.hpp:35)
LATTICE_A = 4.0 * np.pi / np.sqrt(3.0)


class PfcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"pfc_b200 error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [("dim", C.c_int32), ("degree", C.c_int32), ("dt", C.c_double), ("theta", C.c_double),
                ("eps", C.c_double), ("newton_tol", C.c_double), ("newton_max_iters", C.c_int32),
                ("gmres_tol", C.c_double), ("gmres_max_iters", C.c_int64),
                ("gmres_n_tmp_vectors", C.c_int32), ("operator_kind", C.c_int32),
                ("orthogonalization", C.c_int32), ("preconditioner", C.c_int32),
                ("precond_sweeps", C.c_int32), ("device", C.c_int32), ("use_cuda_graph", C.c_int32)]


_i32p, _i64p, _f64p = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_double)


class MeshStruct(C.Structure):
    _fields_ = [("n_cells", C.c_int64), ("n_owned_cells", C.c_int64), ("n_dofs", C.c_int64),
                ("dofs_per_cell", C.c_int32), ("n_q", C.c_int32),
                ("cell_vertices", _f64p), ("cell_dofs", _i32p), ("local_component", _i32p),
                ("local_scalar_index", _i32p), ("shape_value", _f64p), ("shape_grad_ref", _f64p),
                ("q_weight", _f64p), ("block_start", C.c_int64 * 3), ("block_size", C.c_int64 * 3),
                ("n_constraints", C.c_int64), ("constraint_slave", _i32p), ("constraint_master", _i32p),
                ("row_ptr", _i64p), ("col_idx", _i32p), ("rank", C.c_int32), ("n_ranks", C.c_int32),
                ("node_offsets", _i64p)]


class StepStats(C.Structure):
    _fields_ = [("status", C.c_int32), ("n_assemblies", C.c_int32), ("n_solves", C.c_int32),
                ("gmres_iters_total", C.c_int64), ("residual", C.c_double * MAX_NEWTON_RECORD),
                ("gmres_iters", C.c_int64 * MAX_NEWTON_RECORD),
                ("gmres_final_residual", C.c_double * MAX_NEWTON_RECORD),
                ("ms_assembly", C.c_double), ("ms_solver", C.c_double), ("ms_total", C.c_double),
                ("kernel_launches", C.c_int64)]

    def as_dict(self):
        na, ns = min(self.n_assemblies, MAX_NEWTON_RECORD), min(self.n_solves, MAX_NEWTON_RECORD)
        return {"status": self.status, "n_assemblies": self.n_assemblies, "n_solves": self.n_solves,
                "gmres_iters_total": self.gmres_iters_total,
                "residuals": [self.residual[i] for i in range(na)],
                "gmres_iters": [self.gmres_iters[i] for i in range(ns)],
                "gmres_final_residual": [self.gmres_final_residual[i] for i in range(ns)],
                "ms_assembly": self.ms_assembly, "ms_solver": self.ms_solver, "ms_total": self.ms_total,
                "kernel_launches": self.kernel_launches}


class PlanView(C.Structure):
    _fields_ = [("n_s", C.c_int32), ("n_colors", C.c_int32), ("rank", C.c_int32), ("n_ranks", C.c_int32)] + \
               [(k, C.c_int64) for k in ("n_cells", "N_global", "own0", "n_owned", "Np", "n_ghost", "Lv",
                                         "n_slices", "n_ent", "nnz_mono", "n_slaves", "n_peers")] + \
               [(k, _i32p) for k in ("perm", "iperm", "ghost_global", "row_len", "row_master", "col")] + \
               [("slice_ptr", _i64p), ("color_ptr", _i64p)] + \
               [(k, _i32p) for k in ("cell_src", "cell_gather", "cell_row", "cell_pos", "cell_spos")] + \
               [("cell_h", _f64p), ("slave_id", _i32p), ("slave_master", _i32p), ("peers", _i32p),
                ("recv_ptr", _i64p)]


class LatticeView(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("active", "degree", "ncx", "ncy", "ex0", "ey0", "w", "h", "fw", "fh")] + \
               [("hx", C.c_double), ("hy", C.c_double), ("lut", _i32p), ("why", C.c_char * 128)]


class RingView(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("active", "depth", "ex0", "ey0", "fw", "fh", "ncx", "ncy")] + \
               [(k, C.c_int64) for k in ("n_ring", "n_self", "n_recv", "n_send", "n_peers")] + \
               [(k, _i32p) for k in ("self_src", "self_dst", "recv_pos", "send_rows", "peers")] + \
               [("recv_ptr", _i64p), ("send_ptr", _i64p)]


# every symbol include/pfc_b200.h declares (tests check the library exports all of them)
EXPORTED_SYMBOLS = [
    "pfc_default_params", "pfc_create", "pfc_destroy", "pfc_last_error", "pfc_version", "pfc_setup",
    "pfc_nccl_unique_id", "pfc_comm_init", "pfc_comm_mode", "pfc_time_comm", "pfc_set_state", "pfc_get_state", "pfc_n_owned_nodes", "pfc_set_state_owned", "pfc_get_state_owned", "pfc_assemble",
    "pfc_residual_norm", "pfc_dot_noalloc", "pfc_solve_update", "pfc_step", "pfc_run", "pfc_get_rhs", "pfc_get_nnz",
    "pfc_get_pattern", "pfc_get_matrix_values", "pfc_apply_operator", "pfc_get_dof_permutation",
    "pfc_time_operator", "pfc_time_assembly", "pfc_time_krylov", "pfc_time_precond", "pfc_time_reduced", "pfc_reduced_symbol", "pfc_selfcheck", "pfc_selfcheck_smoother", "pfc_selfcheck_q1", "pfc_stress", "pfc_stress_owned", "pfc_vtu_write", "pfc_vtu_piece_name", "pfc_output", "pfc_standin_create",
    "pfc_standin_create_ex", "pfc_standin_destroy",
    "pfc_standin_mesh", "pfc_standin_n_nodes", "pfc_standin_support_points",
    "pfc_standin_hexagonal_lattice", "pfc_standin_polycrystal", "pfc_plan_build", "pfc_plan_build_ex", "pfc_plan_get_lattice", "pfc_plan_ring_overlay", "pfc_plan_get_ring", "pfc_operator_info", "pfc_plan_get_view", "pfc_plan_monolithic",
    "pfc_plan_destroy",
]

_lib = None


def build(force: bool = False) -> str:
    """Compile libpfc_b200.so in-tree with the committed Makefile (nvcc, sm_100a)."""
    if force:
        subprocess.run(["make", "-C", _HERE, "clean"], check=True, capture_output=True)
    r = subprocess.run(["make", "-C", _HERE, "-j8"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libpfc_b200.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return LIB_PATH


def _preload_nccl():
    """Map torch's bundled NCCL first so the library's lazy dlopen("libnccl.so.2") reuses it."""
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        if spec and spec.submodule_search_locations:
            p = os.path.join(list(spec.submodule_search_locations)[0], "lib", "libnccl.so.2")
            if os.path.exists(p):
                C.CDLL(p, mode=C.RTLD_GLOBAL)
    except Exception:
        pass


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PfcError(PFC_ERR_CUDA, f"{LIB_PATH} is not built (run __graft_entry__.build()); there is no fallback")
    L = C.CDLL(LIB_PATH)
    hp = C.c_void_p
    L.pfc_default_params.argtypes = [C.POINTER(Params)]
    L.pfc_default_params.restype = None
    L.pfc_create.argtypes = [C.POINTER(Params), C.POINTER(hp)]
    L.pfc_destroy.argtypes = [hp]
    L.pfc_destroy.restype = None
    L.pfc_last_error.argtypes = [hp]
    L.pfc_last_error.restype = C.c_char_p
    L.pfc_version.restype = C.c_char_p
    L.pfc_setup.argtypes = [hp, C.POINTER(MeshStruct)]
    L.pfc_nccl_unique_id.argtypes = [C.c_void_p]
    L.pfc_comm_init.argtypes = [hp, C.c_void_p, C.c_int, C.c_int]
    L.pfc_comm_mode.argtypes = [hp]
    L.pfc_time_comm.argtypes = [hp, C.c_int32, _f64p]
    L.pfc_set_state.argtypes = [hp, _f64p, _f64p]
    L.pfc_get_state.argtypes = [hp, _f64p, _f64p]
    L.pfc_assemble.argtypes = [hp]
    L.pfc_residual_norm.argtypes = [hp, _f64p]
    L.pfc_dot_noalloc.argtypes = [hp, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    L.pfc_solve_update.argtypes = [hp, _i64p, _f64p]
    L.pfc_step.argtypes = [hp, C.POINTER(StepStats)]
    L.pfc_run.argtypes = [hp, C.c_int32, C.POINTER(StepStats), _f64p]
    L.pfc_get_rhs.argtypes = [hp, _f64p]
    L.pfc_get_nnz.argtypes = [hp]
    L.pfc_get_nnz.restype = C.c_int64
    L.pfc_get_pattern.argtypes = [hp, _i64p, _i32p]
    L.pfc_get_matrix_values.argtypes = [hp, _f64p]
    L.pfc_apply_operator.argtypes = [hp, _f64p, _f64p]
    L.pfc_get_dof_permutation.argtypes = [hp, _i32p]
    L.pfc_time_operator.argtypes = [hp, C.c_int32, _f64p, _f64p]
    L.pfc_time_assembly.argtypes = [hp, C.c_int32, _f64p, _f64p]
    L.pfc_time_precond.argtypes = [hp, C.c_int32, _f64p, _f64p]
    L.pfc_selfcheck.argtypes = [hp, _f64p]
    L.pfc_selfcheck_smoother.argtypes = [hp, _f64p]
    L.pfc_selfcheck_q1.argtypes = [hp, _f64p]
    L.pfc_n_owned_nodes.argtypes = [hp]
    L.pfc_n_owned_nodes.restype = C.c_int64
    L.pfc_set_state_owned.argtypes = [hp, _f64p, _f64p]
    L.pfc_get_state_owned.argtypes = [hp, _f64p, _f64p]
    L.pfc_stress.argtypes = [hp, C.c_double, C.c_double, _f64p, C.POINTER(C.c_int64)]
    L.pfc_stress_owned.argtypes = [hp, C.c_double, C.c_double, _f64p, C.POINTER(C.c_int64)]
    L.pfc_vtu_write.argtypes = [C.POINTER(MeshStruct), C.c_char_p, C.c_char_p, C.c_uint32, C.c_int32, C.POINTER(C.c_char_p),
                                _i32p, _f64p, C.c_char_p, C.c_int32]
    L.pfc_vtu_piece_name.argtypes = [C.c_char_p, C.c_uint32, C.c_int32, C.c_int32, C.c_char_p, C.c_int32]
    L.pfc_output.argtypes = [hp, C.POINTER(MeshStruct), C.c_int32, C.c_char_p, C.c_char_p, C.c_uint32]
    L.pfc_standin_create.argtypes = [C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_double,
                                     C.c_int32, C.c_int32, C.POINTER(hp)]
    L.pfc_standin_create_ex.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_double,
                                        C.c_double, C.c_double, C.c_int32, C.c_int32, C.POINTER(hp)]
    L.pfc_time_krylov.argtypes = [hp, C.c_int32, C.c_int32, _f64p, _f64p, _f64p, _f64p]
    L.pfc_standin_destroy.argtypes = [hp]
    L.pfc_standin_destroy.restype = None
    L.pfc_standin_mesh.argtypes = [hp]
    L.pfc_standin_mesh.restype = C.POINTER(MeshStruct)
    L.pfc_standin_n_nodes.argtypes = [hp]
    L.pfc_standin_n_nodes.restype = C.c_int64
    L.pfc_standin_support_points.argtypes = [hp]
    L.pfc_standin_support_points.restype = _f64p
    L.pfc_standin_hexagonal_lattice.argtypes = [hp, C.c_double, C.c_double, C.c_int32, _f64p, _f64p, _f64p]
    L.pfc_standin_polycrystal.argtypes = [hp, C.c_double, C.c_double, C.c_int32, _f64p, _f64p, _f64p, _f64p]
    L.pfc_plan_build.argtypes = [C.POINTER(MeshStruct), C.c_int32, C.POINTER(hp), C.c_char_p, C.c_int32]
    L.pfc_plan_build_ex.argtypes = [C.POINTER(MeshStruct), C.c_int32, C.c_int32, C.POINTER(hp), C.c_char_p, C.c_int32]
    L.pfc_plan_get_lattice.argtypes = [hp, C.POINTER(LatticeView)]
    L.pfc_plan_get_lattice.restype = None
    L.pfc_plan_ring_overlay.argtypes = [C.POINTER(hp), C.c_int32, C.c_int32, C.c_char_p, C.c_int32]
    L.pfc_plan_get_ring.argtypes = [hp, C.POINTER(RingView)]
    L.pfc_plan_get_ring.restype = None
    L.pfc_time_reduced.argtypes = [hp, C.c_int32, _f64p, _f64p]
    L.pfc_reduced_symbol.argtypes = [C.c_double, C.c_double, C.c_double, _f64p]
    L.pfc_operator_info.argtypes = [hp, C.c_char_p, C.c_int32]
    L.pfc_plan_get_view.argtypes = [hp, C.POINTER(PlanView)]
    L.pfc_plan_get_view.restype = None
    L.pfc_plan_monolithic.argtypes = [hp, _i64p, _i32p, C.POINTER(C.c_int8), _i32p]
    L.pfc_plan_destroy.argtypes = [hp]
    L.pfc_plan_destroy.restype = None
    _lib = L
    return L


def reduced_symbol(hx, hy, dt_theta):
    """(diag(S~), lambda_min, lambda_max of diag^-1 S~): pfc_reduced_symbol, host only"""
    out = (C.c_double * 3)()
    rc = lib().pfc_reduced_symbol(hx, hy, dt_theta, out)
    if rc != 0:
        raise ValueError("pfc_reduced_symbol: bad argument")
    return tuple(out)


def _p(a, ct):
    return a.ctypes.data_as(C.POINTER(ct))


def default_A0(psi0: float, eps: float) -> float:
    """A_0 = "default" (app/simulations/phase_field_crystal_simulation_mpi.cpp:87-88)."""
    return 0.2 * (abs(psi0) + (1.0 / 3.0) * np.sqrt(-15 * eps - 36 * psi0 * psi0))


def polycrystal_grains(p1, p2, n_grains=64, seed=12345):
    """Grain table of the config-4 workload (SURVEY.md 8d "M4"): numpy PCG64(seed) -> seeds uniform in the box,
    rotation alpha_g in [0, pi/3), offset in [0, a)^2.  Returns (seeds [K][2], alpha [K], offset [K][2])."""
    rng = np.random.Generator(np.random.PCG64(seed))
    uv = rng.random((n_grains, 2))
    seeds = np.array([p1[0], p1[1]]) + uv * np.array([p2[0] - p1[0], p2[1] - p1[1]])
    alpha = rng.random(n_grains) * (np.pi / 3.0)
    offset = rng.random((n_grains, 2)) * LATTICE_A
    return seeds, alpha, offset


def vtu_piece_name(basename, counter, rank, n_ranks):
    buf = C.create_string_buffer(512)
    if lib().pfc_vtu_piece_name(basename.encode(), counter, rank, n_ranks, buf, 512) != 0:
        raise PfcError(1, "pfc_vtu_piece_name")
    return buf.value.decode()


def vtu_write(mesh, directory, basename, counter, names, values, interpretation=None):
    """pfc_vtu_write: values [n_fields][N] nodal fields -> <directory><basename>_<counter>.<rank>.vtu (+ .pvtu on rank 0)."""
    ptr = mesh.mesh_ptr if hasattr(mesh, "mesh_ptr") else mesh
    v = np.ascontiguousarray(values, dtype=np.float64)
    nm = (C.c_char_p * len(names))(*[n.encode() for n in names])
    it = None if interpretation is None else np.ascontiguousarray(interpretation, dtype=np.int32)
    err = C.create_string_buffer(400)
    rc = lib().pfc_vtu_write(ptr, directory.encode(), basename.encode(), counter, len(names), nm,
                             _p(it, C.c_int32) if it is not None else None, _p(v, C.c_double), err, 400)
    if rc != 0:
        raise PfcError(rc, err.value.decode())


class StandIn:
    """Host stand-in for make_grid + setup_dofs (mpi.cpp:115-274); see csrc/pfc_standin.cpp."""

    def __init__(self, degree, n_refines, p1, p2, rank=0, n_ranks=1, rep=(1, 1)):
        self._h = C.c_void_p()
        rc = lib().pfc_standin_create_ex(degree, n_refines, rep[0], rep[1], p1[0], p1[1], p2[0], p2[1],
                                         rank, n_ranks, C.byref(self._h))
        if rc != PFC_OK:
            raise PfcError(rc, "pfc_standin_create: invalid arguments")
        self.degree, self.n_refines = degree, n_refines
        self.mesh_ptr = lib().pfc_standin_mesh(self._h)
        m = self.mesh_ptr.contents
        self.N = lib().pfc_standin_n_nodes(self._h)
        self.n_dofs, self.n_cells, self.dofs_per_cell, self.n_q = m.n_dofs, m.n_cells, m.dofs_per_cell, m.n_q
        self.n_owned_cells = m.n_owned_cells
        self.n_s = self.dofs_per_cell // 3
        as_arr = np.ctypeslib.as_array
        self.cell_dofs = as_arr(m.cell_dofs, (self.n_cells, self.dofs_per_cell))
        self.cell_vertices = as_arr(m.cell_vertices, (self.n_cells, 4, 2))
        self.shape = as_arr(m.shape_value, (self.n_q, self.n_s))
        self.dshape = as_arr(m.shape_grad_ref, (self.n_q, self.n_s, 2))
        self.qw = as_arr(m.q_weight, (self.n_q,))
        nc = m.n_constraints
        self.constraint_slave = as_arr(m.constraint_slave, (nc,)) if nc else np.zeros(0, np.int32)
        self.constraint_master = as_arr(m.constraint_master, (nc,)) if nc else np.zeros(0, np.int32)
        self.node_offsets = as_arr(m.node_offsets, (m.n_ranks + 1,)).copy()
        self.support = as_arr(lib().pfc_standin_support_points(self._h), (self.N, 2))

    def hexagonal_lattice(self, A0, psi0, positions=(), burgers=()):
        pos = np.ascontiguousarray(np.asarray(positions, dtype=np.float64).reshape(-1, 2))
        bur = np.ascontiguousarray(np.asarray(burgers, dtype=np.float64).reshape(-1, 2))
        Psi = np.empty(self.n_dofs)
        rc = lib().pfc_standin_hexagonal_lattice(self._h, A0, psi0, len(pos), _p(pos, C.c_double),
                                                 _p(bur, C.c_double), _p(Psi, C.c_double))
        if rc != PFC_OK:
            raise PfcError(rc, "pfc_standin_hexagonal_lattice")
        return Psi

    def polycrystal(self, A0, psi0, seeds, alpha, offset):
        """pfc_standin_polycrystal: periodic Voronoi grains, grain g = the hexagonal lattice rotated by alpha[g]."""
        sd = np.ascontiguousarray(np.asarray(seeds, dtype=np.float64).reshape(-1, 2))
        al = np.ascontiguousarray(np.asarray(alpha, dtype=np.float64).reshape(-1))
        of = np.ascontiguousarray(np.asarray(offset, dtype=np.float64).reshape(-1, 2))
        assert len(sd) == len(al) == len(of)
        Psi = np.empty(self.n_dofs)
        rc = lib().pfc_standin_polycrystal(self._h, A0, psi0, len(al), _p(sd, C.c_double), _p(al, C.c_double),
                                           _p(of, C.c_double), _p(Psi, C.c_double))
        if rc != PFC_OK:
            raise PfcError(rc, "pfc_standin_polycrystal")
        return Psi

    def close(self):
        if self._h:
            lib().pfc_standin_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class HostPlan:
    """Host-only view of what pfc_setup derives (no GPU needed)."""

    def __init__(self, mesh_ptr, degree, flags=0):
        self._h = C.c_void_p()
        err = C.create_string_buffer(512)
        rc = lib().pfc_plan_build_ex(mesh_ptr, degree, flags, C.byref(self._h), err, 512)
        if rc != PFC_OK:
            raise PfcError(rc, err.value.decode())
        v = PlanView()
        lib().pfc_plan_get_view(self._h, C.byref(v))
        lv = LatticeView()
        lib().pfc_plan_get_lattice(self._h, C.byref(lv))
        self.lattice = {k: getattr(lv, k) for k, _ in LatticeView._fields_ if k not in ("lut", "why")}
        self.lattice["why"] = lv.why.decode()
        self.lattice["lut"] = (np.ctypeslib.as_array(lv.lut, (lv.fh, lv.fw)).copy() if lv.active else None)
        for k, _ in PlanView._fields_[:16]:
            setattr(self, k, getattr(v, k))
        arr = np.ctypeslib.as_array
        ns = self.n_s

        def get(ptr, shape):
            return arr(ptr, shape) if (int(np.prod(shape)) > 0 and bool(ptr)) else np.zeros(shape, dtype=np.int32)

        self.perm = get(v.perm, (self.Np,))
        self.iperm = get(v.iperm, (self.n_owned,))
        self.ghost_global = get(v.ghost_global, (self.n_ghost,))
        self.row_len = get(v.row_len, (self.Np,))
        self.row_master = get(v.row_master, (self.Np,))
        self.col = get(v.col, (self.n_ent,))
        self.slice_ptr = arr(v.slice_ptr, (self.n_slices + 1,))
        self.color_ptr = get(v.color_ptr, (self.n_colors + 1,))
        self.cell_src = get(v.cell_src, (self.n_cells,))
        self.cell_gather = get(v.cell_gather, (self.n_cells, ns))
        self.cell_row = get(v.cell_row, (self.n_cells, ns))
        self.cell_pos = get(v.cell_pos, (self.n_cells, ns, ns))
        self.cell_spos = get(v.cell_spos, (self.n_cells, ns))
        self.cell_h = get(v.cell_h, (self.n_cells, 2))
        self.slave_id = get(v.slave_id, (self.n_slaves,))
        self.slave_master = get(v.slave_master, (self.n_slaves,))
        self.peers = get(v.peers, (self.n_peers,))
        self.recv_ptr = arr(v.recv_ptr, (self.n_peers + 1,)) if self.n_peers else np.zeros(1, np.int64)

    def ring(self):
        """view of the ring overlay (after ring_overlay([...plans...], depth))"""
        v = RingView()
        lib().pfc_plan_get_ring(self._h, C.byref(v))
        out = {k: getattr(v, k) for k, _ in RingView._fields_[:13]}
        arr = np.ctypeslib.as_array

        def get(ptr, n, dt=np.int32):
            return arr(ptr, (n,)).copy() if n > 0 and bool(ptr) else np.zeros(0, dt)

        out.update(self_src=get(v.self_src, v.n_self), self_dst=get(v.self_dst, v.n_self), recv_pos=get(v.recv_pos, v.n_recv),
                   send_rows=get(v.send_rows, v.n_send), peers=get(v.peers, v.n_peers),
                   recv_ptr=get(v.recv_ptr, v.n_peers + 1, np.int64), send_ptr=get(v.send_ptr, v.n_peers + 1, np.int64))
        pv = PlanView()
        lib().pfc_plan_get_view(self._h, C.byref(pv))
        out["n_ghost"] = pv.n_ghost
        out["ghost_global"] = get(pv.ghost_global, pv.n_ghost)
        out["slave_id"], out["slave_master"] = get(pv.slave_id, pv.n_slaves), get(pv.slave_master, pv.n_slaves)
        return out

    def monolithic(self):
        n = 3 * self.N_global
        rp = np.empty(n + 1, np.int64)
        ci = np.empty(self.nnz_mono, np.int32)
        blk = np.empty(self.nnz_mono, np.int8)
        pos = np.empty(self.nnz_mono, np.int32)
        rc = lib().pfc_plan_monolithic(self._h, _p(rp, C.c_int64), _p(ci, C.c_int32), _p(blk, C.c_int8),
                                       _p(pos, C.c_int32))
        if rc != PFC_OK:
            raise PfcError(rc, "pfc_plan_monolithic")
        return rp, ci, blk, pos

    def close(self):
        if self._h:
            lib().pfc_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def ring_overlay(plans, depth):
    """pfc_plan_ring_overlay over the HostPlans of all ranks of a partition (flags 3), in rank order"""
    arr = (C.c_void_p * len(plans))(*[p._h for p in plans])
    err = C.create_string_buffer(512)
    rc = lib().pfc_plan_ring_overlay(arr, len(plans), depth, err, 512)
    if rc != PFC_OK:
        raise PfcError(rc, err.value.decode())


class Solver:
    """One handle = one GPU = one rank.  Thin wrapper: every method is one C-ABI call."""

    def __init__(self, **kw):
        L = lib()
        self.params = Params()
        L.pfc_default_params(C.byref(self.params))
        for k, v in kw.items():
            if not hasattr(self.params, k):
                raise TypeError(f"unknown parameter {k}")
            setattr(self.params, k, v)
        self._h = C.c_void_p()
        rc = L.pfc_create(C.byref(self.params), C.byref(self._h))
        if rc != PFC_OK:
            raise PfcError(rc, L.pfc_last_error(None).decode())
        self.n_dofs = 0

    def _check(self, rc, allow=()):
        if rc != PFC_OK and rc not in allow:
            raise PfcError(rc, lib().pfc_last_error(self._h).decode())
        return rc

    def comm_init(self, unique_id: bytes, rank: int, n_ranks: int):
        _preload_nccl()
        buf = C.create_string_buffer(unique_id, 128)
        self._check(lib().pfc_comm_init(self._h, buf, rank, n_ranks))

    def comm_mode(self):
        return lib().pfc_comm_mode(self._h)

    def time_comm(self, reps=50):
        ms = (C.c_double * 3)()
        self._check(lib().pfc_time_comm(self._h, reps, ms))
        return {"us_halo_scalar": 1e3 * ms[0], "us_halo_vec3": 1e3 * ms[1], "us_allreduce16": 1e3 * ms[2]}

    def setup(self, mesh):
        ptr = mesh.mesh_ptr if hasattr(mesh, "mesh_ptr") else mesh
        self._check(lib().pfc_setup(self._h, ptr))
        self.n_dofs = ptr.contents.n_dofs

    def set_state(self, Psi_n, Psi_n1):
        a = np.ascontiguousarray(Psi_n, dtype=np.float64)
        b = np.ascontiguousarray(Psi_n1, dtype=np.float64)
        assert a.size == self.n_dofs and b.size == self.n_dofs
        self._check(lib().pfc_set_state(self._h, _p(a, C.c_double), _p(b, C.c_double)))

    def get_state(self, out_n=None, out_n1=None, want_n1=False):
        a = np.empty(self.n_dofs) if out_n is None else out_n
        b = (np.empty(self.n_dofs) if out_n1 is None else out_n1) if (want_n1 or out_n1 is not None) else None
        self._check(lib().pfc_get_state(self._h, _p(a, C.c_double), _p(b, C.c_double) if b is not None else None))
        return (a, b) if b is not None else a

    def n_owned_nodes(self):
        return int(lib().pfc_n_owned_nodes(self._h))

    def set_state_owned(self, Psi_n, Psi_n1):
        """[3][n_owned] arrays (owned_partitioning, mpi.cpp:167-176); ghosts come from their owners."""
        a = np.ascontiguousarray(Psi_n, dtype=np.float64)
        b = a if Psi_n1 is Psi_n else np.ascontiguousarray(Psi_n1, dtype=np.float64)
        assert a.size == 3 * self.n_owned_nodes() and b.size == a.size
        self._check(lib().pfc_set_state_owned(self._h, _p(a, C.c_double), _p(b, C.c_double)))

    def get_state_owned(self, out_n=None):
        a = np.empty(3 * self.n_owned_nodes()) if out_n is None else out_n
        self._check(lib().pfc_get_state_owned(self._h, _p(a, C.c_double), None))
        return a

    def assemble(self):
        self._check(lib().pfc_assemble(self._h))

    def residual_norm(self):
        v = C.c_double()
        self._check(lib().pfc_residual_norm(self._h, C.byref(v)))
        return v.value

    def dot_noalloc(self, x_ptr, y_ptr, n, out_ptr, stream=None):
        """pfc_dot_noalloc: device pointers (ints), nothing allocated; out_ptr[0] = x . y over n doubles"""
        self._check(lib().pfc_dot_noalloc(self._h, C.c_void_p(x_ptr), C.c_void_p(y_ptr), n, C.c_void_p(out_ptr), C.c_void_p(stream)))

    def solve_update(self):
        its, res = C.c_int64(), C.c_double()
        self._check(lib().pfc_solve_update(self._h, C.byref(its), C.byref(res)))
        return its.value, res.value

    def step(self):
        st = StepStats()
        self._check(lib().pfc_step(self._h, C.byref(st)), allow=(PFC_WARN_NEWTON_MAX_ITERS,))
        return st.as_dict()

    def run(self, n_timesteps):
        st, ms = StepStats(), C.c_double()
        self._check(lib().pfc_run(self._h, n_timesteps, C.byref(st), C.byref(ms)), allow=(PFC_WARN_NEWTON_MAX_ITERS,))
        return st.as_dict(), ms.value

    def get_rhs(self):
        r = np.empty(self.n_dofs)
        self._check(lib().pfc_get_rhs(self._h, _p(r, C.c_double)))
        return r

    def get_pattern(self):
        nnz = lib().pfc_get_nnz(self._h)
        rp, ci = np.empty(self.n_dofs + 1, np.int64), np.empty(nnz, np.int32)
        self._check(lib().pfc_get_pattern(self._h, _p(rp, C.c_int64), _p(ci, C.c_int32)))
        return rp, ci

    def get_matrix_values(self):
        v = np.empty(lib().pfc_get_nnz(self._h))
        self._check(lib().pfc_get_matrix_values(self._h, _p(v, C.c_double)))
        return v

    def apply_operator(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(self.n_dofs)
        self._check(lib().pfc_apply_operator(self._h, _p(x, C.c_double), _p(y, C.c_double)))
        return y

    def time_operator(self, reps=20):
        ms, by = C.c_double(), C.c_double()
        self._check(lib().pfc_time_operator(self._h, reps, C.byref(ms), C.byref(by)))
        return ms.value, by.value

    def time_krylov(self, dim, reps=10):
        a, b, c, d = C.c_double(), C.c_double(), C.c_double(), C.c_double()
        self._check(lib().pfc_time_krylov(self._h, dim, reps, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return {"ms_dot": a.value, "bytes_dot": b.value, "ms_axpy": c.value, "bytes_axpy": d.value}

    def selfcheck(self):
        v = C.c_double()
        self._check(lib().pfc_selfcheck(self._h, C.byref(v)))
        return v.value

    def stress(self, eps, tol=1e-10, out=None, owned=False):
        """StressCalculatorMPI::calculate_stress (stress_calculator_mpi.cpp:345-353) of the current Psi_n:
        ([4][N] nodal stress with c = a*2 + b — [4][n_owned] if owned —, CG iterations per component)."""
        if out is None:
            out = np.zeros((4, self.n_owned_nodes() if owned else self.n_dofs // 3))
        its = (C.c_int64 * 4)()
        f = lib().pfc_stress_owned if owned else lib().pfc_stress
        self._check(f(self._h, eps, tol, _p(out, C.c_double), its))
        return out, list(its)

    def output(self, mesh, kind, directory, basename, counter):
        """kind 0: output_configuration (mpi.cpp:597-619), 1: output_rhs (:624-645), 2: output_stress (stress.cpp:358-384)."""
        ptr = mesh.mesh_ptr if hasattr(mesh, "mesh_ptr") else mesh
        self._check(lib().pfc_output(self._h, ptr, kind, directory.encode(), basename.encode(), counter))

    def selfcheck_smoother(self):
        v = C.c_double()
        self._check(lib().pfc_selfcheck_smoother(self._h, C.byref(v)))
        return v.value

    def selfcheck_q1(self):
        """(J apply, preconditioner application): closed-form Q1 kernels against the general ones (pfc_selfcheck_q1)"""
        v = (C.c_double * 2)()
        self._check(lib().pfc_selfcheck_q1(self._h, v))
        return v[0], v[1]

    def time_precond(self, reps=10):
        """legs of include/pfc_b200.h pfc_time_precond: ms[0..4], bytes[0..3] (bytes[k] belongs to ms[k + 1])"""
        ms, by = (C.c_double * 5)(), (C.c_double * 4)()
        self._check(lib().pfc_time_precond(self._h, reps, ms, by))
        return {"ms": list(ms), "bytes": list(by), "ms_apply": ms[0]}

    def time_reduced(self, reps=20):
        """legs of the reduced solve (pfc_time_reduced): ms and algorithmic bytes per launch"""
        ms, by = (C.c_double * 6)(), (C.c_double * 6)()
        self._check(lib().pfc_time_reduced(self._h, reps, ms, by))
        return {"ms": list(ms), "bytes": list(by)}

    def operator_info(self):
        buf = C.create_string_buffer(512)
        kind = lib().pfc_operator_info(self._h, buf, 512)
        return kind, buf.value.decode()

    def time_assembly(self, reps=5):
        ms, by = C.c_double(), C.c_double()
        self._check(lib().pfc_time_assembly(self._h, reps, C.byref(ms), C.byref(by)))
        return ms.value, by.value

    def close(self):
        if self._h:
            lib().pfc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nccl_unique_id() -> bytes:
    _preload_nccl()
    buf = C.create_string_buffer(128)
    rc = lib().pfc_nccl_unique_id(buf)
    if rc != PFC_OK:
        raise PfcError(rc, "pfc_nccl_unique_id: NCCL unavailable")
    return buf.raw
