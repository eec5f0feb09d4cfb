/*
 * pfc_b200.h — C ABI of the B200-native per-timestep PFC hot path.
 *
 * The reference (lucasmyers97/phase-field-crystals) has no FFI: its hot path is three private
 * methods of PhaseFieldCrystalSystemMPI<dim>
 *   (src/phase_field_crystal_systems/phase_field_crystal_system_mpi.hpp:96-104, "mpi.hpp";
 *    src/phase_field_crystal_systems/phase_field_crystal_system_mpi.cpp, "mpi.cpp").
 * This header is the boundary a maintainer binds instead of those method bodies; every entry
 * point names the reference lines it replaces.  Plain C types only; no exception crosses it.
 *
 * Ownership: the caller owns every host array; the library copies during the call and retains
 * no host pointer.  All device memory belongs to the handle.
 * Threading: one handle per GPU (per rank); a handle is not thread-safe.  With n_ranks > 1 the
 * calls marked [collective] must be made by all ranks, like the MPI original.
 * There is no CPU fallback: every compute entry point returns PFC_ERR_CUDA without a GPU.
 */
#ifndef PFC_B200_H
#define PFC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes (app/simulations/phase_field_crystal_simulation_mpi.cpp:172-180 catches C++
 * exceptions and prints them; here every failure is a code + pfc_last_error text) */
enum
{
  PFC_OK = 0,
  PFC_ERR_BAD_ARGUMENT = 1,   /* std::invalid_argument in the reference */
  PFC_ERR_CUDA = 2,           /* no device / CUDA runtime error */
  PFC_ERR_NCCL = 3,
  PFC_ERR_GMRES_NOT_CONVERGED = 4, /* deal.II SolverControl::NoConvergence at mpi.cpp:482 */
  PFC_ERR_NONFINITE = 5,      /* NaN/Inf residual */
  PFC_WARN_NEWTON_MAX_ITERS = 6, /* mpi.cpp:579-592: silently accepted by the reference */
  PFC_ERR_UNSUPPORTED = 7,    /* input outside what the path handles (see pfc_setup) */
  PFC_ERR_STATE = 8           /* call order (e.g. pfc_step before pfc_setup) */
};

enum
{
  PFC_OPERATOR_ASSEMBLED = 0,
  PFC_OPERATOR_MATRIX_FREE = 1,         /* lattice kernels (sum-factorised, no index tables) when the mesh is the
                                           uniform periodic tensor lattice the reference always builds, else general */
  PFC_OPERATOR_MATRIX_FREE_GENERAL = 2  /* colour-ordered table-driven cell kernel on any supported mesh */
};
enum
{
  PFC_ORTHO_CGS = 0,   /* classical Gram-Schmidt, one fused pass, Kelley re-orthogonalisation test */
  PFC_ORTHO_CGS2 = 1,  /* always two passes */
  PFC_ORTHO_MGS = 2    /* modified Gram-Schmidt in the reference's operation order */
};
enum
{
  PFC_PRECOND_NONE = 0,        /* mpi.cpp:482 PreconditionIdentity (the reference's active path) */
  PFC_PRECOND_BLOCK_SCHUR = 1, /* right preconditioner: lumped-mass block elimination + Chebyshev on
                                  the psi Schur complement (cf. mpi.cpp:496-572, serial.cpp:319-369) */
  PFC_PRECOND_SCHUR_REDUCED = 2 /* the psi-only Schur-reduced solve of the serial class (serial.cpp:319-369:
                                  S = B + (D M^-1 K - C) M^-1 K, GMRES on vectors of N instead of 3 N, back substitution
                                  for chi, phi) without inner mass solves: with dpsi = M^2 y the operator is
                                  T = B M^2 + D K^2 - C K M (K, M commute on the uniform periodic Q1 lattice), right
                                  preconditioned by Chebyshev sweeps on the constant-coefficient Schur stand-in.  Same
                                  tolerance on the residual of the 3-block system.  Q1 + PFC_OPERATOR_MATRIX_FREE on a
                                  lattice mesh; anywhere else it acts as PFC_PRECOND_BLOCK_SCHUR (pfc_operator_info
                                  says so) */
};

/* Constructor arguments of the hot path: the subset of the 15-argument constructor
 * (mpi.hpp:30-50) the per-timestep path reads, plus the deal.II defaults it relies on. */
typedef struct pfc_params
{
  int32_t dim;                 /* 2 (dim 3 is instantiated upstream, mpi.cpp:716, but never run) */
  int32_t degree;              /* FE_Q degree, 1 or 2 (mpi.cpp:84-86) */
  double  dt, theta, eps;      /* mpi.hpp:80-84 */
  double  newton_tol;          /* simulation_tol, absolute on ||-R||_2 (mpi.cpp:586) */
  int32_t newton_max_iters;    /* simulation_max_iters (mpi.cpp:579) */
  double  gmres_tol;           /* SolverControl default 1e-10 absolute (mpi.cpp:478) */
  int64_t gmres_max_iters;     /* <= 0: n_dofs, as SolverControl(n_dofs) (mpi.cpp:478) */
  int32_t gmres_n_tmp_vectors; /* <= 0: 30, deal.II default => restart length 28 */
  int32_t operator_kind;       /* PFC_OPERATOR_* */
  int32_t orthogonalization;   /* PFC_ORTHO_* */
  int32_t preconditioner;      /* PFC_PRECOND_* */
  int32_t precond_sweeps;      /* Chebyshev degree of the Schur solve (<= 0: default) */
  int32_t device;              /* CUDA device ordinal; < 0: current device */
  int32_t use_cuda_graph;      /* capture the GMRES restart cycle in a CUDA graph */
} pfc_params;

/* Flattened products of make_grid + setup_dofs (mpi.cpp:115-274) for the cells this rank
 * integrates.  Single rank: all active cells.  Multi rank: owned cells first, then the one-cell
 * ghost layer (mpi.cpp:178-186), so that every owned row is complete without compress(add)
 * (mpi.cpp:464-466).  dof indices are global. */
typedef struct pfc_mesh
{
  int64_t n_cells;            /* cells handed over (owned + ghost layer) */
  int64_t n_owned_cells;      /* first n_owned_cells are locally owned (mpi.cpp:353) */
  int64_t n_dofs;             /* global number of dofs (dof_handler.n_dofs(), mpi.cpp:478) */
  int32_t dofs_per_cell;      /* fe_system.n_dofs_per_cell() (mpi.cpp:317): 12 or 27 */
  int32_t n_q;                /* quadrature_formula.size() (mpi.cpp:318): 4 or 9 */
  const double  *cell_vertices;      /* [n_cells][4][2] lexicographic; cells must be axis-aligned
                                        rectangles (hyper_rectangle, mpi.cpp:119-122) */
  const int32_t *cell_dofs;          /* [n_cells][dofs_per_cell], cell->get_dof_indices (mpi.cpp:453) */
  const int32_t *local_component;    /* [dofs_per_cell] fe.system_to_component_index(i).first */
  const int32_t *local_scalar_index; /* [dofs_per_cell] fe.system_to_component_index(i).second */
  const double  *shape_value;        /* [n_q][n_s]    scalar FE_Q values at the Gauss points */
  const double  *shape_grad_ref;     /* [n_q][n_s][2] reference-cell gradients */
  const double  *q_weight;           /* [n_q] weights on the unit cell */
  int64_t block_start[3];            /* first global dof of block psi, chi, phi (component_wise,
                                        mpi.cpp:157-165); blocks must be index-congruent */
  int64_t block_size[3];
  /* AffineConstraints (mpi.cpp:188-217), closed; only identity periodic lines are supported:
   * slave = 1.0 * master, inhomogeneity 0 */
  int64_t        n_constraints;
  const int32_t *constraint_slave;   /* [n_constraints] global dof */
  const int32_t *constraint_master;  /* [n_constraints] global dof */
  /* optional sparsity pattern of the host (mpi.cpp:232-244), monolithic CSR with ascending
   * columns; if given, pfc_setup checks it is bit-identical to the pattern the library derives */
  const int64_t *row_ptr;            /* [n_dofs + 1] or NULL */
  const int32_t *col_idx;            /* [nnz] or NULL */
  /* partition (single rank: n_ranks = 1, node_offsets may be NULL).  Rank r owns, in every
   * block b, the contiguous dofs block_start[b] + [node_offsets[r], node_offsets[r+1])
   * (owned_partitioning, mpi.cpp:167-176; all ranks' counts as from
   * Utilities::MPI::all_gather of n_locally_owned_dofs_per_block). */
  int32_t rank, n_ranks;
  const int64_t *node_offsets;       /* [n_ranks + 1] */
} pfc_mesh;

/* per-call statistics of pfc_step: what the reference prints through pcout
 * ("Residual is:" mpi.cpp:585, "Number of iterations:" mpi.cpp:483) and TimerOutput sections
 * "assembly" / "solver" (mpi.cpp:303,474) */
#define PFC_MAX_NEWTON_RECORD 256
typedef struct pfc_step_stats
{
  int32_t status;              /* PFC_OK, PFC_WARN_NEWTON_MAX_ITERS, or an error */
  int32_t n_assemblies;
  int32_t n_solves;
  int64_t gmres_iters_total;
  double  residual[PFC_MAX_NEWTON_RECORD];     /* ||-R||_2 after each assembly */
  int64_t gmres_iters[PFC_MAX_NEWTON_RECORD];  /* per solve */
  double  gmres_final_residual[PFC_MAX_NEWTON_RECORD];
  double  ms_assembly, ms_solver, ms_total;    /* CUDA-event times on the handle's stream */
  int64_t kernel_launches;                     /* launches of this library's kernels */
} pfc_step_stats;

typedef struct pfc_handle pfc_handle;

/* ---- lifecycle ------------------------------------------------------------------------- */
void        pfc_default_params(pfc_params *p); /* values of parameter-files/hexagonal_lattice.toml */
int         pfc_create(const pfc_params *p, pfc_handle **out);    /* ctor, mpi.cpp:50-111 */
void        pfc_destroy(pfc_handle *h);
const char *pfc_last_error(const pfc_handle *h);  /* h may be NULL: last error of pfc_create */
const char *pfc_version(void);

/* setup_dofs (mpi.cpp:150-274): derives the scalar pattern shared by the 7 coupled blocks,
 * the periodic elimination, the colouring of the scatter; uploads; assembles the iterate-
 * independent blocks once. */
int pfc_setup(pfc_handle *h, const pfc_mesh *mesh);

/* multi-GPU: a NCCL communicator per handle.  id_bytes is the 128-byte ncclUniqueId obtained
 * by rank 0 from pfc_nccl_unique_id and broadcast by the host (MPI_COMM_WORLD at mpi.cpp:79). */
int pfc_nccl_unique_id(void *id_bytes128);
int pfc_comm_init(pfc_handle *h, const void *id_bytes128, int rank, int n_ranks);
/* after pfc_setup: 0 single rank, 1 NCCL send/recv + all-reduce, 2 peer-memory kernels over NVLink (CUDA IPC;
 * chosen when every rank can map its neighbours, PFC_NO_P2P=1 forces NCCL) */
int pfc_comm_mode(const pfc_handle *h);
/* device time (ms) of one scalar halo exchange, one 3-component halo exchange and one all-reduce of 16 doubles,
 * `reps` of each replayed from a CUDA graph [collective] */
int pfc_time_comm(pfc_handle *h, int32_t reps, double *ms3);

/* ---- state (mpi.cpp:278-296 writes Psi_n, Psi_n_1; output reads them mpi.cpp:597-619) ---- */
int pfc_set_state(pfc_handle *h, const double *Psi_n, const double *Psi_n_1); /* [n_dofs] each; [collective] on the
                                                                               * lattice path with n_ranks > 1 (ring values) */
int pfc_get_state(pfc_handle *h, double *Psi_n, double *Psi_n_1);             /* either may be NULL */
/* The same with what a rank of the MPI original holds (owned_partitioning, mpi.cpp:167-176): arrays [3][n_owned],
 * block b = the contiguous owned dofs block_start[b] + [node_offsets[rank], node_offsets[rank + 1]).
 * pfc_set_state_owned fetches the ghost values from their owners (update_ghost_values) [collective];
 * n_owned = pfc_n_owned_nodes.  Per-rank transfers are 24 n_owned bytes instead of 8 n_dofs. */
int64_t pfc_n_owned_nodes(const pfc_handle *h);
int pfc_set_state_owned(pfc_handle *h, const double *Psi_n, const double *Psi_n_1);
int pfc_get_state_owned(pfc_handle *h, double *Psi_n, double *Psi_n_1);       /* either may be NULL */

/* ---- the hot path ---------------------------------------------------------------------- */
int pfc_assemble(pfc_handle *h);                        /* assemble_system, mpi.cpp:300-467 [collective] */
int pfc_residual_norm(pfc_handle *h, double *norm);     /* system_rhs.l2_norm(), mpi.cpp:584 [collective] */
/* NO COUNTERPART IN THE REFERENCE; added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  out_dev[0] = sum_i x_dev[i] * y_dev[i] (x_dev == y_dev:
 * the squared l2 norm that pfc_residual_norm takes the root of) for n doubles in DEVICE memory, result to DEVICE memory, on
 * `stream` (a cudaStream_t; NULL = the handle's stream); rank-local, no collective.  Allocates nothing: the partial sums
 * and the block counter of the reduction live in a file-scope static __device__ array (csrc/pfc_noalloc.cu, 64 KiB + 256 B
 * on every GPU that loads the library, until exit), which bounds the length: n > PFC_NOALLOC_MAX_N returns
 * PFC_ERR_BAD_ARGUMENT with a message.  The result does not depend on the SM count (fixed chunk per block, fixed summation
 * order).  All calls share the scratch: calls that OVERLAP on the GPU (different streams, concurrently replayed graphs)
 * race; calls on one stream are ordered.  The first call on a device resolves the array's address (cudaGetSymbolAddress,
 * no allocation): make one call, e.g. with n = 0, before starting a capture. */
#define PFC_NOALLOC_MAX_N 134217728LL
int pfc_dot_noalloc(pfc_handle *h, const double *x_dev, const double *y_dev, int64_t n, double *out_dev, void *stream);
int pfc_solve_update(pfc_handle *h, int64_t *gmres_iters, double *final_residual);
                                                        /* solve_and_update_no_precondition, mpi.cpp:471-492 [collective] */
int pfc_step(pfc_handle *h, pfc_step_stats *stats);     /* iterate_timestep, mpi.cpp:576-593 [collective] */
int pfc_run(pfc_handle *h, int32_t n_timesteps, pfc_step_stats *last, double *ms_total);
                                                        /* the loop of run(), mpi.cpp:685-688, without output */

/* ---- inspection (parity tests; output_rhs mpi.cpp:623-645) ----------------------------- */
int pfc_get_rhs(pfc_handle *h, double *rhs);            /* [n_dofs] */
int64_t pfc_get_nnz(pfc_handle *h);
int pfc_get_pattern(pfc_handle *h, int64_t *row_ptr, int32_t *col_idx);  /* monolithic CSR */
int pfc_get_matrix_values(pfc_handle *h, double *values);               /* [nnz], same order */
int pfc_apply_operator(pfc_handle *h, const double *x, double *y);      /* y = J x (assembled or matrix-free) */
int pfc_get_dof_permutation(pfc_handle *h, int32_t *internal_of_node);  /* [N] internal row of scalar node */

/* timing of one operator application on resident data (bench.py roofline leg): runs `reps`
 * applications on the handle's stream between CUDA events, returns the mean in ms and the
 * algorithmic bytes of one launch for the active layout */
int pfc_time_operator(pfc_handle *h, int32_t reps, double *ms_mean, double *bytes_per_launch);
int pfc_time_assembly(pfc_handle *h, int32_t reps, double *ms_mean, double *bytes_per_launch);
/* block-Schur preconditioner legs on resident data, device time per launch replayed from a CUDA graph:
 *   ms5[0] one full application P^-1 v (lattice path: including its one halo exchange)
 *   assembled path: ms5[1] one K product (kmv_kernel), ms5[2] one Schur-row product (bcd_kernel, Chebyshev step)
 *   lattice path:   ms5[1] one Chebyshev sweep (ring_sweep_kernel), ms5[2] Schur right-hand side + first step
 *                   (lat_kernel<LAT_S>), ms5[3] t kernel (ring_t_kernel), ms5[4] back substitution (ring_back_kernel)
 * bytes4[k] = algorithmic bytes of one launch of leg k + 1 (0 where the leg does not exist) */
int pfc_time_precond(pfc_handle *h, int32_t reps, double *ms5, double *bytes4);
/* legs of the reduced solve (PFC_PRECOND_SCHUR_REDUCED active), device time per launch replayed from a CUDA graph:
 *   ms6[0] operator T y = row 0 of J (M^2 y, -K M y, K^2 y) (both kernels), ms6[1] lat_pstage_kernel (the three fields),
 *   ms6[2] lat_j1_kernel<ROW0>, ms6[3] preconditioner (one exchange + the Chebyshev sweeps), ms6[4] the first sweep
 *   (steps 0 and 1 in one pass), ms6[5] a later sweep (0 if precond_sweeps < 3); bytes6[k] = algorithmic bytes of a
 *   launch of leg k (0 for leg 3).  pfc_time_krylov times the Gram-Schmidt kernels on the reduced vectors then. */
int pfc_time_reduced(pfc_handle *h, int32_t reps, double *ms6, double *bytes6);
/* host only, no GPU: the constant-coefficient Schur stand-in S~ = M + dt theta (K - 2 M)(M_L^-1 K)^2 of the Chebyshev sweeps on
 * a periodic Q1 lattice of hx x hy cells: out3 = {its diagonal entry, min, max of the spectrum of diag^-1 S~} (what the
 * reduced solve takes its Jacobi scaling and Chebyshev interval from; checked against assembled matrices on the CPU) */
int pfc_reduced_symbol(double hx, double hy, double dt_theta, double *out3);
/* consistency of redundant kernel forms on a deterministic pseudo-random field (lattice path: K product as node
 * stencil vs cell quadrature); max_rel = max |a - b| / max |a|.  0 when the handle has no redundant form. [collective] */
int pfc_selfcheck(pfc_handle *h, double *max_rel);
/* lattice path with the preconditioner: S~ x by the one-kernel streaming Chebyshev sweep (csrc/pfc_ring.cuh) against the
 * composition of the K-stencil and mass-stencil kernels with a halo exchange before every product [collective] */
int pfc_selfcheck_smoother(pfc_handle *h, double *max_rel);
/* Q1 lattice path: the closed-form kernels (csrc/pfc_ring_q1.cuh: J apply with the constant blocks as stencils; t kernel +
 * Schur right-hand side fused; sweeps, back substitution) against the general kernels they replace, on a pseudo-random
 * vector: max_rel2[0] J apply, max_rel2[1] preconditioner application; max |a - b| / max |b| over the owned rows.  Needs
 * pfc_assemble and PFC_PRECOND_BLOCK_SCHUR.  [collective] */
int pfc_selfcheck_q1(pfc_handle *h, double *max_rel2);
/* which operator path pfc_setup chose: returns PFC_OPERATOR_ASSEMBLED, PFC_OPERATOR_MATRIX_FREE (lattice kernels) or
 * PFC_OPERATOR_MATRIX_FREE_GENERAL, -1 before pfc_setup; note (optional) receives why the lattice kernels were declined
 * when PFC_OPERATOR_MATRIX_FREE was asked for and the general kernel is in charge (empty otherwise) */
int pfc_operator_info(const pfc_handle *h, char *note, int32_t note_len);

/* ---- stress projection ------------------------------------------------------
 * Replaces StressCalculatorMPI<dim>::calculate_stress(mpi_communicator, Psi_dof_handler, Psi, eps)
 * (src/stress_tools/stress_calculator_mpi.cpp:345-353 = calculate_righthand_side :204-314 + solve
 * :317-341, with the mass matrix of calculate_mass_matrix :141-200) for Psi = the handle's current
 * Psi_n, which is read where the time stepper left it in HBM.
 *   stress   [4][N], N = n_dofs / 3: component c = a*2 + b of the L2-projected stress at scalar node t
 *            (the node carrying psi dof block_start[0] + t) at stress[c*N + t]; like pfc_get_state,
 *            a rank fills the nodes it owns and leaves the others 0.  Periodic images carry the value of their master (:336).
 *   cg_iters [4]: SolverControl::last_step() of each component's SolverCG (:329-334).
 *   tol:     absolute tolerance on ||b - M x||_2 (SolverControl default 1e-10, :320); at most 4 N steps.
 * Needs the lattice path (operator_kind = PFC_OPERATOR_MATRIX_FREE on a uniform periodic mesh);
 * PFC_ERR_UNSUPPORTED otherwise, PFC_ERR_GMRES_NOT_CONVERGED where the reference's SolverCG would throw
 * NoConvergence.  [collective] */
int pfc_stress(pfc_handle *h, double eps, double tol, double *stress, int64_t *cg_iters);
/* the same with stress [4][n_owned] over this rank's owned nodes (node_offsets[rank] + j), cf. pfc_get_state_owned */
int pfc_stress_owned(pfc_handle *h, double eps, double tol, double *stress, int64_t *cg_iters);

/* ---- output -----------------------------------------------------------------
 * Host-side writer replacing DataOut::build_patches() + write_vtu_with_pvtu_record(directory, basename,
 * counter, comm, n_digits_for_counter = 2) as called by output_configuration / output_rhs (mpi.cpp:597-645)
 * and StressCalculatorMPI::output_stress (stress_calculator_mpi.cpp:358-384): this rank's owned cells as
 * VTK_QUADs with their 4 vertex values into <directory><basename>_<counter %02u>.<rank>.vtu, and on rank 0 the
 * record <directory><basename>_<counter %02u>.pvtu naming all ranks' pieces.  directory and basename are
 * concatenated as given (like the reference; pass "out/" to write into a folder).
 *   values [n_fields][N] nodal fields by scalar node t (psi dof block_start[0] + t); the vertices of the owned
 *          cells must be filled (owned + ghost nodes).
 *   interpretation[f]: 0 component_is_scalar, 1 component_is_part_of_tensor (consecutive fields of one name
 *          form one 9-component tensor array, 2-D component a*2 + b in slot a*3 + b); NULL = all scalar.
 * No GPU involved.  Returns PFC_OK or PFC_ERR_BAD_ARGUMENT with a message in err. */
int pfc_vtu_write(const pfc_mesh *mesh, const char *directory, const char *basename, uint32_t counter,
                  int32_t n_fields, const char *const *names, const int32_t *interpretation,
                  const double *values, char *err, int32_t errlen);
/* file name of the piece of `rank` (rank < 0: of the .pvtu record), without directory */
int pfc_vtu_piece_name(const char *basename, uint32_t counter, int32_t rank, int32_t n_ranks, char *out, int32_t outlen);
/* download (owned + ghost values) and write: kind 0 = Psi_n as "psi","chi","phi" (output_configuration),
 * 1 = the right-hand side of the last assembly under the same names (output_rhs; system_rhs as it stands after
 * pfc_assemble / pfc_step) [collective],
 * 2 = the result of the last pfc_stress as tensor "stress" (output_stress).  mesh = what pfc_setup was given. */
int pfc_output(pfc_handle *h, const pfc_mesh *mesh, int32_t kind, const char *directory, const char *basename,
               uint32_t counter);
/* the two Gram-Schmidt kernels at Krylov dimension dim (1..restart): multi-dot (reads dim+1
 * vectors) and multi-axpy with norm (reads dim+1, writes 1) */
int pfc_time_krylov(pfc_handle *h, int32_t dim, int32_t reps, double *ms_dot, double *bytes_dot,
                    double *ms_axpy, double *bytes_axpy);

/* ---- host stand-in for the deal.II products (deal.II is not available in this image) ----
 * Produces exactly the arrays of pfc_mesh for the only mesh the reference builds: one
 * hyper_rectangle coarse cell refined n_refines times, periodic in x and y (mpi.cpp:115-146),
 * FESystem(FE_Q(degree)^3), component-wise numbering (mpi.cpp:155-158), periodic constraints
 * (mpi.cpp:188-217).  `repetitions` > 1 extends the coarse mesh (not reachable upstream). */
typedef struct pfc_standin pfc_standin;
int  pfc_standin_create(int32_t degree, int32_t n_refines, double x0, double y0, double x1,
                        double y1, int32_t rank, int32_t n_ranks, pfc_standin **out);
/* rep_x x rep_y coarse cells instead of one (subdivided_hyper_rectangle; a harness extension
 * for the weak-scaling meshes at 2 and 8 ranks): trees lexicographic, Morton inside */
int  pfc_standin_create_ex(int32_t degree, int32_t n_refines, int32_t rep_x, int32_t rep_y, double x0,
                           double y0, double x1, double y1, int32_t rank, int32_t n_ranks, pfc_standin **out);
void pfc_standin_destroy(pfc_standin *s);
const pfc_mesh *pfc_standin_mesh(const pfc_standin *s);
int64_t pfc_standin_n_nodes(const pfc_standin *s);             /* nodes per component */
const double *pfc_standin_support_points(const pfc_standin *s); /* [N][2], node t <-> dof c*N+t */
/* HexagonalLattice<dim>::value (src/phase_field_functions/hexagonal_lattice.cpp:52-77) interpolated
 * at the support points, then constraints.distribute (mpi.cpp:285-288).  Psi: [n_dofs]. */
int  pfc_standin_hexagonal_lattice(const pfc_standin *s, double A0, double psi0, int32_t n_dislocations,
                                   const double *positions /*[n][2]*/, const double *burgers /*[n][2]*/,
                                   double *Psi);
/* Polycrystalline initial condition for BASELINE.json config 4 (a harness extension: the reference has no such
 * function): periodic Voronoi grains around seeds[n][2]; grain g is HexagonalLattice<dim>::value (no dislocations)
 * rotated by alpha[g] and shifted by offset[g][2].  Bit-identical to the oracle's orc_polycrystal on the same inputs. */
int  pfc_standin_polycrystal(const pfc_standin *s, double A0, double psi0, int32_t n_grains, const double *seeds /*[n][2]*/,
                             const double *alpha /*[n]*/, const double *offset /*[n][2]*/, double *Psi);

/* ---- host-only view of what pfc_setup derives (tests run it without a GPU) ------------------
 * The same code path pfc_setup uses: shared scalar pattern in SELL-32-sigma order, internal
 * permutation, colouring and scatter tables.  Pointers stay valid until pfc_plan_destroy. */
typedef struct pfc_plan pfc_plan;
typedef struct pfc_plan_view
{
  int32_t n_s, n_colors, rank, n_ranks;
  int64_t n_cells, N_global, own0, n_owned, Np, n_ghost, Lv, n_slices, n_ent, nnz_mono, n_slaves, n_peers;
  const int32_t *perm, *iperm, *ghost_global, *row_len, *row_master, *col;
  const int64_t *slice_ptr, *color_ptr;
  const int32_t *cell_src, *cell_gather, *cell_row, *cell_pos, *cell_spos;
  const double  *cell_h;
  const int32_t *slave_id, *slave_master;
  const int32_t *peers;
  const int64_t *recv_ptr;
} pfc_plan_view;
int  pfc_plan_build(const pfc_mesh *mesh, int32_t degree, pfc_plan **out, char *err, int32_t err_len);
/* flags: 1 = look for the uniform periodic tensor lattice behind the arrays (what make_grid, mpi.cpp:115-146,
 * always produces) and number the owned free nodes lexicographically, as pfc_setup does for
 * PFC_OPERATOR_MATRIX_FREE; 2 = additionally skip the SELL pattern and per-cell tables */
int  pfc_plan_build_ex(const pfc_mesh *mesh, int32_t degree, int32_t flags, pfc_plan **out, char *err, int32_t err_len);
typedef struct pfc_lattice_view
{
  int32_t active, degree;            /* active = 0: `why` says what ruled the lattice out */
  int32_t ncx, ncy;                  /* frame: the cells touching an owned free node */
  int32_t ex0, ey0, w, h;            /* owned free nodes: frame rectangle [ex0, ex0+w) x [ey0, ey0+h) */
  int32_t fw, fh;                    /* frame nodes per direction: degree * ncx + 1, degree * ncy + 1 */
  double  hx, hy;
  const int32_t *lut;                /* [fh][fw] internal id of the resolved node at a frame position */
  char why[128];
} pfc_lattice_view;
void pfc_plan_get_lattice(const pfc_plan *p, pfc_lattice_view *view);
/* Deep ghost ring of the lattice path (what pfc_setup builds after the ranks exchanged their places on the lattice;
 * csrc/pfc_ring_plan.cpp).  plans[r] = the plan of rank r built with flags 3; afterwards rank r's local positions are
 * the padded rectangle [0, fw) x [0, fh) with the owned free nodes at [ex0, ex0 + w) x [ey0, ey0 + h); a position
 * outside has the ghost id Np + ring index and is filled by the halo exchange described by the lists:
 *   tail[self_dst[i]] = owned[self_src[i]]                  positions this rank owns itself (periodic wrap-around)
 *   peer k sends owned[send_rows[send_ptr[k] ..)]; the j-th value received (peer-major) goes to ring position recv_pos[j] */
typedef struct pfc_ring_view
{
  int32_t active, depth, ex0, ey0, fw, fh, ncx, ncy;
  int64_t n_ring, n_self, n_recv, n_send, n_peers;
  const int32_t *self_src, *self_dst, *recv_pos, *send_rows, *peers;
  const int64_t *recv_ptr, *send_ptr;
} pfc_ring_view;
int  pfc_plan_ring_overlay(pfc_plan *const *plans, int32_t n_ranks, int32_t depth, char *err, int32_t err_len);
void pfc_plan_get_ring(const pfc_plan *p, pfc_ring_view *view);
void pfc_plan_get_view(const pfc_plan *p, pfc_plan_view *view);
/* monolithic CSR the reference would build (mpi.cpp:220-244) + where each entry lives:
 * block_of 0 B, 1 C, 2 D, 3 K, 4 M; pos_of = entry position in the SELL arrays (single rank) */
int  pfc_plan_monolithic(const pfc_plan *p, int64_t *row_ptr, int32_t *col_idx, int8_t *block_of, int32_t *pos_of);
void pfc_plan_destroy(pfc_plan *p);

#ifdef __cplusplus
}
#endif
#endif /* PFC_B200_H */
