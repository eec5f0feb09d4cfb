/*
 * pfc_oracle.h — CPU restatement of the reference's per-timestep PFC hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * PARITY UNPINNED: the reference (lucasmyers97/phase-field-crystals) ships no golden vectors,
 * no known-answer tests and cannot be compiled here (deal.II / Trilinos / p4est / MPI absent).
 * The oracle is pinned only by the self-consistency checks of tests/test_oracle_*.py
 * (closed-form element matrices, finite-difference Jacobian, mass conservation, analytic RHS).
 *
 * Every function cites the reference file:line it follows.  Abbreviations:
 *   mpi.cpp = src/phase_field_crystal_systems/phase_field_crystal_system_mpi.cpp
 *   hex.cpp = src/phase_field_functions/hexagonal_lattice.cpp
 *   app.cpp = app/simulations/phase_field_crystal_simulation_mpi.cpp
 *   stress.cpp = src/stress_tools/stress_calculator_mpi.cpp
 * [dealii-ext] marks behaviour of deal.II (>= 9.3, not vendored in the reference) that is
 * restated from its published algorithms.
 */
#ifndef PFC_ORACLE_H
#define PFC_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_mesh
{
  int32_t degree, n_refines;
  int32_t nx, ny;         /* cells per direction (2^n_refines) */
  int32_t n_s;            /* scalar shape functions per cell (degree+1)^2 */
  int32_t dofs_per_cell;  /* 3 n_s */
  int32_t n_q;            /* (degree+1)^2 Gauss points */
  int64_t n_cells;
  int64_t N;              /* nodes per component (incl. periodic duplicates) */
  int64_t n_dofs;         /* 3 N */
  int64_t n_constrained;
  int64_t nnz;
  double  x0, y0, x1, y1, hx, hy;
  int32_t *cell_dofs;     /* [n_cells][dofs_per_cell] global dof, deal.II local order   */
  double  *cell_vertices; /* [n_cells][4][2], lexicographic vertex order                */
  int32_t *master;        /* [n_dofs] closed constraint: master dof, or -1 if free      */
  double  *support;       /* [N][2] support point of scalar node t (dof = c*N + t)      */
  double  *shape;         /* [n_q][n_s]    scalar shape values at Gauss points          */
  double  *dshape;        /* [n_q][n_s][2] reference gradients                          */
  double  *qw;            /* [n_q] Gauss weights on [0,1]^2                             */
  int32_t *local_comp;    /* [dofs_per_cell] component of system shape function i       */
  int32_t *local_shape;   /* [dofs_per_cell] scalar shape index of system shape fn i    */
  int64_t *row_ptr;       /* [n_dofs+1] monolithic CSR, columns ascending               */
  int32_t *col_idx;       /* [nnz]                                                      */
} orc_mesh;

/* mpi.cpp:115-146 (grid), 150-245 (dofs, constraints, sparsity) */
orc_mesh *orc_mesh_create(int degree, int n_refines, double x0, double y0, double x1, double y1);
void      orc_mesh_destroy(orc_mesh *m);

/* hex.cpp:52-77 + mpi.cpp:278-296; A_0 "default" per app.cpp:87-88 is computed by the caller */
void orc_initial_condition(const orc_mesh *m, double A0, double psi0, int n_disl,
                           const double *disl_pos /*[n][2]*/, const double *burgers /*[n][2]*/,
                           double *Psi /*[n_dofs]*/);

/* Polycrystal (harness extension, BASELINE.json config 4; not in the reference): periodic Voronoi grains around
 * seeds[n][2], grain g = hex.cpp:52-77's cos-sum rotated by alpha[g] and shifted by offset[g][2]. */
void orc_polycrystal(const orc_mesh *m, double A0, double psi0, int n_grains, const double *seeds,
                     const double *alpha, const double *offset, double *Psi /*[n_dofs]*/);

/* mpi.cpp:300-467.  values[nnz] and rhs[n_dofs] are overwritten.
 * threads <= 1: strict serial cell order (parity mode); > 1: 4-colour OpenMP (timed mode).
 * dead_work != 0 additionally accumulates local_preconditioner_matrix (mpi.cpp:446-448). */
void orc_assemble(const orc_mesh *m, const double *Psi_n, const double *Psi_n1, double dt,
                  double theta, double eps, double *values, double *rhs, int threads,
                  int dead_work);

/* y = A x on the monolithic CSR */
void orc_spmv(const orc_mesh *m, const double *values, const double *x, double *y);

/* [dealii-ext] SolverGMRES as called at mpi.cpp:478-482: x0 = 0, left preconditioning with
 * identity, 30 temporary vectors (restart 28), modified Gram-Schmidt with the Kelley
 * re-orthogonalisation test, absolute tolerance tol, at most max_iters steps.
 * Returns SolverControl::last_step(); *converged = 0 if NoConvergence would be thrown.
 * stop_after > 0 stops after that many inner iterations (timing samples only). */
int64_t orc_gmres(const orc_mesh *m, const double *values, const double *b, double *x, double tol,
                  int64_t max_iters, int n_tmp_vectors, int *converged, double *final_res,
                  int64_t stop_after);

/* mpi.cpp:576-593 (+471-492).  Psi_n, Psi_n1 updated in place.  res_hist[newton_max+1],
 * gmres_hist[newton_max].  Returns number of GMRES solves; *n_assemblies is set.
 * status: 0 ok, 1 Newton hit max iters (silently accepted, mpi.cpp:579-592), 2 GMRES failed. */
int orc_newton_step(const orc_mesh *m, double *Psi_n, double *Psi_n1, double dt, double theta,
                    double eps, double newton_tol, int newton_max, double gmres_tol, int threads,
                    double *res_hist, int64_t *gmres_hist, int *n_assemblies, int *status);

/* mpi.cpp:485 — slave dof := master dof */
void orc_distribute(const orc_mesh *m, double *v);

/* stress.cpp:141-200 (mass matrix), 204-314 (right-hand side), 317-341 (one unpreconditioned CG
 * per tensor component, then constraints.distribute).  stress[c*N + t]: component c = a*2 + b of
 * the L2-projected stress on scalar node t.  cg_iters[4] = SolverControl::last_step() per
 * component; rhs_out (optional, [4N]) receives the assembled right-hand side.
 * Returns 0, or 2 if a CG solve would throw NoConvergence. */
int orc_stress(const orc_mesh *m, const double *Psi, double eps, double tol, double *stress,
               int64_t *cg_iters, double *rhs_out);

int orc_max_threads(void);
void orc_set_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
