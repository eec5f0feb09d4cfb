/*
 * pfc_oracle.c — CPU restatement of the reference's per-timestep PFC hot path.
 * TEST INFRASTRUCTURE ONLY; PARITY UNPINNED (see pfc_oracle.h).
 *
 * Written to follow the reference statement by statement where the reference has code
 * (mpi.cpp:300-467, 471-492, 576-593; hex.cpp:52-77) and deal.II's documented algorithms
 * where the reference delegates ([dealii-ext]: DoF enumeration, periodic constraints,
 * sparsity, distribute_local_to_global, SolverGMRES).
 */
#include "pfc_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

int orc_max_threads(void)
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void orc_set_threads(int n)
{
#ifdef _OPENMP
  if (n > 0)
    omp_set_num_threads(n);
#else
  (void)n;
#endif
}

static void *xmalloc(size_t n)
{
  void *p = malloc(n ? n : 1);
  if (!p)
    {
      fprintf(stderr, "pfc_oracle: out of memory (%zu bytes)\n", n);
      abort();
    }
  return p;
}

/* ------------------------------------------------------------------------------------------
 * Grid.  mpi.cpp:119-122: one hyper_rectangle coarse cell; mpi.cpp:145: refine_global(n).
 * [dealii-ext] children are created in lexicographic order (x fastest) and a globally refined
 * single coarse cell therefore enumerates its active cells along the Morton (z-order) curve.
 * ------------------------------------------------------------------------------------------ */
static void cell_coords(int64_t cell, int n_refines, int *cx, int *cy)
{
  int x = 0, y = 0;
  for (int k = 0; k < n_refines; ++k)
    {
      x |= (int)((cell >> (2 * k)) & 1) << k;
      y |= (int)((cell >> (2 * k + 1)) & 1) << k;
    }
  *cx = x;
  *cy = y;
}

/* [dealii-ext] FE_Q(p) local (hierarchical) numbering on the unit square, p in {1,2}, as
 * fine-grid offsets (units of h/p): 4 vertices, then lines x=0, x=1, y=0, y=1, then interior. */
static void local_offsets(int degree, int ox[9], int oy[9])
{
  if (degree == 1)
    {
      const int X[4] = {0, 1, 0, 1}, Y[4] = {0, 0, 1, 1};
      for (int a = 0; a < 4; ++a)
        {
          ox[a] = X[a];
          oy[a] = Y[a];
        }
    }
  else
    {
      const int X[9] = {0, 2, 0, 2, 0, 2, 1, 1, 1}, Y[9] = {0, 0, 2, 2, 1, 1, 0, 2, 1};
      for (int a = 0; a < 9; ++a)
        {
          ox[a] = X[a];
          oy[a] = Y[a];
        }
    }
}

/* 1-D Lagrange basis on equidistant nodes k/p, and derivative */
static double lagrange(int p, int k, double x)
{
  double v = 1.0;
  for (int m = 0; m <= p; ++m)
    if (m != k)
      v *= (x - (double)m / p) / ((double)k / p - (double)m / p);
  return v;
}

static double dlagrange(int p, int k, double x)
{
  double s = 0.0;
  for (int l = 0; l <= p; ++l)
    {
      if (l == k)
        continue;
      double t = 1.0 / ((double)k / p - (double)l / p);
      for (int m = 0; m <= p; ++m)
        if (m != k && m != l)
          t *= (x - (double)m / p) / ((double)k / p - (double)m / p);
      s += t;
    }
  return s;
}

static int cmp_i32(const void *a, const void *b)
{
  const int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
  return (x > y) - (x < y);
}

orc_mesh *orc_mesh_create(int degree, int n_refines, double x0, double y0, double x1, double y1)
{
  if (degree < 1 || degree > 2 || n_refines < 1 || n_refines > 13)
    return NULL;
  orc_mesh *m = (orc_mesh *)calloc(1, sizeof(orc_mesh));
  const int p = degree;
  m->degree = p;
  m->n_refines = n_refines;
  m->nx = m->ny = 1 << n_refines;
  m->n_s = (p + 1) * (p + 1);
  m->dofs_per_cell = 3 * m->n_s;
  m->n_q = (p + 1) * (p + 1); /* mpi.cpp:308 QGauss(degree+1) */
  m->n_cells = (int64_t)m->nx * m->ny;
  const int64_t Gx = (int64_t)p * m->nx + 1, Gy = (int64_t)p * m->ny + 1;
  m->N = Gx * Gy;
  m->n_dofs = 3 * m->N;
  m->x0 = x0;
  m->y0 = y0;
  m->x1 = x1;
  m->y1 = y1;
  m->hx = (x1 - x0) / m->nx;
  m->hy = (y1 - y0) / m->ny;
  const int ns = m->n_s, dpc = m->dofs_per_cell, nq = m->n_q;

  /* ---- reference-cell tables: QGauss(p+1) x FE_Q(p) [dealii-ext] ---- */
  double gp[3], gw[3];
  if (p == 1)
    {
      gp[0] = 0.5 * (1.0 - 1.0 / sqrt(3.0));
      gp[1] = 0.5 * (1.0 + 1.0 / sqrt(3.0));
      gw[0] = gw[1] = 0.5;
    }
  else
    {
      gp[0] = 0.5 * (1.0 - sqrt(3.0 / 5.0));
      gp[1] = 0.5;
      gp[2] = 0.5 * (1.0 + sqrt(3.0 / 5.0));
      gw[0] = gw[2] = 5.0 / 18.0;
      gw[1] = 8.0 / 18.0;
    }
  int ox[9], oy[9];
  local_offsets(p, ox, oy);
  m->shape = (double *)xmalloc(sizeof(double) * nq * ns);
  m->dshape = (double *)xmalloc(sizeof(double) * nq * ns * 2);
  m->qw = (double *)xmalloc(sizeof(double) * nq);
  for (int qy = 0; qy <= p; ++qy)
    for (int qx = 0; qx <= p; ++qx)
      {
        const int q = qx + (p + 1) * qy; /* x fastest */
        m->qw[q] = gw[qx] * gw[qy];
        for (int a = 0; a < ns; ++a)
          {
            m->shape[q * ns + a] = lagrange(p, ox[a], gp[qx]) * lagrange(p, oy[a], gp[qy]);
            m->dshape[(q * ns + a) * 2 + 0] = dlagrange(p, ox[a], gp[qx]) * lagrange(p, oy[a], gp[qy]);
            m->dshape[(q * ns + a) * 2 + 1] = lagrange(p, ox[a], gp[qx]) * dlagrange(p, oy[a], gp[qy]);
          }
      }
  /* [dealii-ext] FESystem(FE_Q,1,FE_Q,1,FE_Q,1) interleaves the components on every geometric
   * object; with one scalar node per object (p<=2) that is i = 3*a + c.  mpi.cpp:84-86 */
  m->local_comp = (int32_t *)xmalloc(sizeof(int32_t) * dpc);
  m->local_shape = (int32_t *)xmalloc(sizeof(int32_t) * dpc);
  for (int i = 0; i < dpc; ++i)
    {
      m->local_comp[i] = i % 3;
      m->local_shape[i] = i / 3;
    }

  /* ---- distribute_dofs + component_wise (mpi.cpp:155-158) [dealii-ext] ----
   * Cell-by-cell first touch (vertices, lines, interior), 3 consecutive system dofs per node;
   * component_wise keeps the relative order inside a component, so dof(c, node) = c*N + t(node)
   * with t = first-touch rank of the node. */
  int32_t *node_id = (int32_t *)xmalloc(sizeof(int32_t) * m->N);
  for (int64_t i = 0; i < m->N; ++i)
    node_id[i] = -1;
  m->cell_dofs = (int32_t *)xmalloc(sizeof(int32_t) * m->n_cells * dpc);
  m->cell_vertices = (double *)xmalloc(sizeof(double) * m->n_cells * 8);
  m->support = (double *)xmalloc(sizeof(double) * m->N * 2);
  int32_t *cell_nodes = (int32_t *)xmalloc(sizeof(int32_t) * m->n_cells * ns);
  int32_t next = 0;
  for (int64_t cell = 0; cell < m->n_cells; ++cell)
    {
      int cx, cy;
      cell_coords(cell, n_refines, &cx, &cy);
      for (int v = 0; v < 4; ++v)
        {
          m->cell_vertices[cell * 8 + v * 2 + 0] = x0 + (cx + (v & 1)) * m->hx;
          m->cell_vertices[cell * 8 + v * 2 + 1] = y0 + (cy + (v >> 1)) * m->hy;
        }
      for (int a = 0; a < ns; ++a)
        {
          const int64_t fx = (int64_t)p * cx + ox[a], fy = (int64_t)p * cy + oy[a];
          const int64_t g = fy * Gx + fx;
          if (node_id[g] < 0)
            {
              node_id[g] = next;
              m->support[2 * (int64_t)next + 0] = x0 + fx * (m->hx / p);
              m->support[2 * (int64_t)next + 1] = y0 + fy * (m->hy / p);
              ++next;
            }
          cell_nodes[cell * ns + a] = node_id[g];
        }
      for (int i = 0; i < dpc; ++i)
        m->cell_dofs[cell * dpc + i] =
          (int32_t)(m->local_comp[i] * m->N + cell_nodes[cell * ns + m->local_shape[i]]);
    }

  /* ---- periodic constraints (mpi.cpp:188-217) [dealii-ext] ----
   * make_periodicity_constraints(face_1 = boundary id 0 / 2, face_2 = id 1 / 3): dofs on the
   * x_min (y_min) face are constrained, weight 1, to their images on x_max (y_max); after
   * close() the four corners chain to the (x_max, y_max) corner. */
  m->master = (int32_t *)xmalloc(sizeof(int32_t) * m->n_dofs);
  int32_t *node_master = (int32_t *)xmalloc(sizeof(int32_t) * m->N); /* resolved node */
  m->n_constrained = 0;
  for (int64_t fy = 0; fy < Gy; ++fy)
    for (int64_t fx = 0; fx < Gx; ++fx)
      {
        const int64_t mx = (fx == 0) ? Gx - 1 : fx, my = (fy == 0) ? Gy - 1 : fy;
        const int32_t t = node_id[fy * Gx + fx], tm = node_id[my * Gx + mx];
        node_master[t] = tm;
        for (int c = 0; c < 3; ++c)
          m->master[c * m->N + t] = (tm == t) ? -1 : (int32_t)(c * m->N + tm);
        if (tm != t)
          m->n_constrained += 3;
      }

  /* ---- sparsity (mpi.cpp:220-244) [dealii-ext make_sparsity_pattern(coupling, constraints,
   * keep_constrained_dofs=false)]: {(m(i), m(j)) : i, j on one cell, coupling(c_i, c_j)} plus
   * the diagonal of every constrained row.  All coupled blocks share the scalar pattern. ---- */
  int64_t *adj_ptr = (int64_t *)calloc((size_t)m->N + 1, sizeof(int64_t));
  for (int64_t k = 0; k < m->n_cells * ns; ++k)
    adj_ptr[node_master[cell_nodes[k]] + 1]++;
  for (int64_t i = 0; i < m->N; ++i)
    adj_ptr[i + 1] += adj_ptr[i];
  int32_t *adj = (int32_t *)xmalloc(sizeof(int32_t) * adj_ptr[m->N]);
  {
    int64_t *fill = (int64_t *)xmalloc(sizeof(int64_t) * m->N);
    memcpy(fill, adj_ptr, sizeof(int64_t) * m->N);
    for (int64_t cell = 0; cell < m->n_cells; ++cell)
      for (int a = 0; a < ns; ++a)
        adj[fill[node_master[cell_nodes[cell * ns + a]]]++] = (int32_t)cell;
    free(fill);
  }
  /* scalar rows */
  int64_t *srow_ptr = (int64_t *)xmalloc(sizeof(int64_t) * (m->N + 1));
  srow_ptr[0] = 0;
  int32_t *scol = NULL;
  {
    /* two passes: count, fill */
    int32_t *len = (int32_t *)xmalloc(sizeof(int32_t) * m->N);
#pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < m->N; ++t)
      {
        if (node_master[t] != t)
          {
            len[t] = 0;
            continue;
          }
        int32_t tmp[4 * 9];
        int n = 0;
        for (int64_t k = adj_ptr[t]; k < adj_ptr[t + 1]; ++k)
          for (int a = 0; a < ns; ++a)
            tmp[n++] = node_master[cell_nodes[(int64_t)adj[k] * ns + a]];
        qsort(tmp, n, sizeof(int32_t), cmp_i32);
        int u = 0;
        for (int i = 0; i < n; ++i)
          if (i == 0 || tmp[i] != tmp[i - 1])
            u++;
        len[t] = u;
      }
    for (int64_t t = 0; t < m->N; ++t)
      srow_ptr[t + 1] = srow_ptr[t] + len[t];
    scol = (int32_t *)xmalloc(sizeof(int32_t) * srow_ptr[m->N]);
#pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < m->N; ++t)
      {
        if (node_master[t] != t)
          continue;
        int32_t tmp[4 * 9];
        int n = 0;
        for (int64_t k = adj_ptr[t]; k < adj_ptr[t + 1]; ++k)
          for (int a = 0; a < ns; ++a)
            tmp[n++] = node_master[cell_nodes[(int64_t)adj[k] * ns + a]];
        qsort(tmp, n, sizeof(int32_t), cmp_i32);
        int64_t o = srow_ptr[t];
        for (int i = 0; i < n; ++i)
          if (i == 0 || tmp[i] != tmp[i - 1])
            scol[o++] = tmp[i];
      }
    free(len);
  }
  /* monolithic rows; coupling table mpi.cpp:220-229: (phi,psi) and (chi,phi) are none */
  static const int coupled[3][3] = {{1, 1, 1}, {1, 1, 0}, {0, 1, 1}};
  m->row_ptr = (int64_t *)xmalloc(sizeof(int64_t) * (m->n_dofs + 1));
  m->row_ptr[0] = 0;
  for (int c = 0; c < 3; ++c)
    for (int64_t t = 0; t < m->N; ++t)
      {
        int64_t l;
        if (node_master[t] != t)
          l = 1;
        else
          {
            int nb = coupled[c][0] + coupled[c][1] + coupled[c][2];
            l = nb * (srow_ptr[t + 1] - srow_ptr[t]);
          }
        m->row_ptr[c * m->N + t + 1] = m->row_ptr[c * m->N + t] + l;
      }
  m->nnz = m->row_ptr[m->n_dofs];
  m->col_idx = (int32_t *)xmalloc(sizeof(int32_t) * m->nnz);
#pragma omp parallel for schedule(static) collapse(2)
  for (int c = 0; c < 3; ++c)
    for (int64_t t = 0; t < m->N; ++t)
      {
        int64_t o = m->row_ptr[c * m->N + t];
        if (node_master[t] != t)
          {
            m->col_idx[o] = (int32_t)(c * m->N + t);
            continue;
          }
        for (int d = 0; d < 3; ++d)
          if (coupled[c][d])
            for (int64_t k = srow_ptr[t]; k < srow_ptr[t + 1]; ++k)
              m->col_idx[o++] = (int32_t)(d * m->N + scol[k]);
      }

  free(scol);
  free(srow_ptr);
  free(adj);
  free(adj_ptr);
  free(node_master);
  free(cell_nodes);
  free(node_id);
  return m;
}

void orc_mesh_destroy(orc_mesh *m)
{
  if (!m)
    return;
  free(m->cell_dofs);
  free(m->cell_vertices);
  free(m->master);
  free(m->support);
  free(m->shape);
  free(m->dshape);
  free(m->qw);
  free(m->local_comp);
  free(m->local_shape);
  free(m->row_ptr);
  free(m->col_idx);
  free(m);
}

/* ------------------------------------------------------------------------------------------
 * Initial condition.  hex.cpp:14-21 (q vectors), hex.cpp:52-77 (value), mpi.cpp:285-288
 * (nodal interpolation, then constraints.distribute).
 * ------------------------------------------------------------------------------------------ */
static double hexagonal_lattice_value(double px, double py, int component, double A0, double psi0,
                                      int n_disl, const double *pos, const double *burg)
{
  const double qv[3][2] = {{0.0, 1.0}, {0.5 * sqrt(3.0), -0.5}, {-0.5 * sqrt(3.0), -0.5}};
  double sum = 0;
  for (int n = 0; n < 3; ++n)
    {
      double phase_sum = 0;
      for (int i = 0; i < n_disl; ++i)
        {
          const double s_n_j = 1.0 / (2 * M_PI) * (qv[n][0] * burg[2 * i] + qv[n][1] * burg[2 * i + 1]);
          const double theta_j = atan2(py - pos[2 * i + 1], px - pos[2 * i]);
          phase_sum += s_n_j * theta_j;
        }
      sum += cos((qv[n][0] * px + qv[n][1] * py) - phase_sum);
    }
  if (component == 0)
    return psi0 + A0 * sum;
  else if (component == 1)
    return -A0 * sum;
  return A0 * sum;
}

void orc_distribute(const orc_mesh *m, double *v)
{
  for (int64_t i = 0; i < m->n_dofs; ++i)
    if (m->master[i] >= 0)
      v[i] = v[m->master[i]];
}

void orc_initial_condition(const orc_mesh *m, double A0, double psi0, int n_disl,
                           const double *disl_pos, const double *burgers, double *Psi)
{
  for (int c = 0; c < 3; ++c)
    for (int64_t t = 0; t < m->N; ++t)
      Psi[c * m->N + t] = hexagonal_lattice_value(m->support[2 * t], m->support[2 * t + 1], c, A0,
                                                  psi0, n_disl, disl_pos, burgers);
  orc_distribute(m, Psi);
}

/* ------------------------------------------------------------------------------------------
 * Polycrystalline initial condition (harness extension for BASELINE.json config 4; the reference has
 * no such function).  Grain g owns the periodic Voronoi cell of its seed (minimum-image distance on
 * the box); inside it the field is hex.cpp:52-77's cos-sum without dislocations, evaluated in the
 * grain's frame: p' = R(alpha_g)^T rho + o_g with rho the minimum-image displacement from the seed,
 * i.e. the lattice of hexagonal_lattice_value rotated by alpha_g and shifted by o_g.  The seeds,
 * angles and offsets come from the caller (numpy PCG64(12345) in the harness) so that this function
 * and the stand-in's copy see identical inputs; ties go to the lowest grain index.
 * ------------------------------------------------------------------------------------------ */
void orc_polycrystal(const orc_mesh *m, double A0, double psi0, int n_grains, const double *seeds,
                     const double *alpha, const double *offset, double *Psi)
{
  const double W = m->x1 - m->x0, H = m->y1 - m->y0;
#pragma omp parallel for schedule(static)
  for (int64_t t = 0; t < m->N; ++t)
    {
      const double px = m->support[2 * t], py = m->support[2 * t + 1];
      int best = 0;
      double bd = 0, bx = 0, by = 0;
      for (int g = 0; g < n_grains; ++g)
        {
          double dx = px - seeds[2 * g], dy = py - seeds[2 * g + 1];
          dx -= W * rint(dx / W);
          dy -= H * rint(dy / H);
          const double d2 = dx * dx + dy * dy;
          if (g == 0 || d2 < bd)
            {
              best = g;
              bd = d2;
              bx = dx;
              by = dy;
            }
        }
      const double ca = cos(alpha[best]), sa = sin(alpha[best]);
      const double qx = ca * bx + sa * by + offset[2 * best];
      const double qy = ca * by - sa * bx + offset[2 * best + 1];
      for (int c = 0; c < 3; ++c)
        Psi[c * m->N + t] = hexagonal_lattice_value(qx, qy, c, A0, psi0, 0, NULL, NULL);
    }
  orc_distribute(m, Psi);
}

/* ------------------------------------------------------------------------------------------
 * Assembly.  mpi.cpp:300-467, loops kept literal (dense dofs_per_cell^2 incl. structural zeros).
 * ------------------------------------------------------------------------------------------ */
static int64_t csr_find(const orc_mesh *m, int64_t row, int32_t col)
{
  int64_t lo = m->row_ptr[row], hi = m->row_ptr[row + 1] - 1;
  while (lo <= hi)
    {
      const int64_t mid = (lo + hi) >> 1;
      const int32_t c = m->col_idx[mid];
      if (c == col)
        return mid;
      if (c < col)
        lo = mid + 1;
      else
        hi = mid - 1;
    }
  return -1;
}

#define MAXD 27
#define MAXQ 9

static void assemble_cell(const orc_mesh *m, int64_t cell, const double *Psi_n, const double *Psi_n_1,
                          double dt, double theta, double eps, double *values, double *rhs,
                          int dead_work, double *dead_sink)
{
  const int dpc = m->dofs_per_cell, ns = m->n_s, nq = m->n_q;
  const int32_t *dofs = m->cell_dofs + cell * dpc;
  /* fe_values.reinit(cell) mpi.cpp:355 [dealii-ext MappingQ1 on an axis-aligned rectangle]:
   * physical gradient = reference gradient / (hx, hy); JxW = w_q hx hy */
  const double *vtx = m->cell_vertices + cell * 8;
  const double hx = vtx[2] - vtx[0], hy = vtx[5] - vtx[1];

  double local_matrix[MAXD][MAXD], local_pre[MAXD][MAXD], local_rhs[MAXD];
  memset(local_matrix, 0, sizeof(local_matrix));
  memset(local_rhs, 0, sizeof(local_rhs));
  if (dead_work)
    memset(local_pre, 0, sizeof(local_pre));

  /* mpi.cpp:360-371 — 12 get_function_* calls */
  double psi_n[MAXQ], chi_n[MAXQ], phi_n[MAXQ], psi_n_1[MAXQ], chi_n_1[MAXQ], phi_n_1[MAXQ];
  double grad_psi_n[MAXQ][2], grad_chi_n[MAXQ][2], grad_phi_n[MAXQ][2];
  double grad_psi_n_1[MAXQ][2], grad_chi_n_1[MAXQ][2], grad_phi_n_1[MAXQ][2];
  (void)grad_chi_n_1; /* gathered by the reference (mpi.cpp:370) but never used */
  for (int q = 0; q < nq; ++q)
    {
      double v[2][3] = {{0, 0, 0}, {0, 0, 0}}, g[2][3][2] = {{{0, 0}, {0, 0}, {0, 0}}, {{0, 0}, {0, 0}, {0, 0}}};
      for (int k = 0; k < dpc; ++k)
        {
          const int c = m->local_comp[k], a = m->local_shape[k];
          const double N = m->shape[q * ns + a];
          const double gx = m->dshape[(q * ns + a) * 2 + 0] / hx, gy = m->dshape[(q * ns + a) * 2 + 1] / hy;
          const double u0 = Psi_n[dofs[k]], u1 = Psi_n_1[dofs[k]];
          v[0][c] += u0 * N;
          g[0][c][0] += u0 * gx;
          g[0][c][1] += u0 * gy;
          v[1][c] += u1 * N;
          g[1][c][0] += u1 * gx;
          g[1][c][1] += u1 * gy;
        }
      psi_n[q] = v[0][0];
      chi_n[q] = v[0][1];
      phi_n[q] = v[0][2];
      psi_n_1[q] = v[1][0];
      chi_n_1[q] = v[1][1];
      phi_n_1[q] = v[1][2];
      for (int d = 0; d < 2; ++d)
        {
          grad_psi_n[q][d] = g[0][0][d];
          grad_chi_n[q][d] = g[0][1][d];
          grad_phi_n[q][d] = g[0][2][d];
          grad_psi_n_1[q][d] = g[1][0][d];
          grad_chi_n_1[q][d] = g[1][1][d];
          grad_phi_n_1[q][d] = g[1][2][d];
        }
    }

#define DOT2(a, b) ((a)[0] * (b)[0] + (a)[1] * (b)[1])
  double eta_psi[MAXD], eta_chi[MAXD], eta_phi[MAXD];
  double grad_eta_psi[MAXD][2], grad_eta_chi[MAXD][2], grad_eta_phi[MAXD][2];
  for (int q = 0; q < nq; ++q) /* mpi.cpp:373 */
    {
      const double JxW = m->qw[q] * hx * hy;
      for (int k = 0; k < dpc; ++k) /* mpi.cpp:376-384: extractor views are zero off-component */
        {
          const int c = m->local_comp[k], a = m->local_shape[k];
          const double N = m->shape[q * ns + a];
          const double gx = m->dshape[(q * ns + a) * 2 + 0] / hx, gy = m->dshape[(q * ns + a) * 2 + 1] / hy;
          eta_psi[k] = (c == 0) ? N : 0.0;
          eta_chi[k] = (c == 1) ? N : 0.0;
          eta_phi[k] = (c == 2) ? N : 0.0;
          grad_eta_psi[k][0] = (c == 0) ? gx : 0.0;
          grad_eta_psi[k][1] = (c == 0) ? gy : 0.0;
          grad_eta_chi[k][0] = (c == 1) ? gx : 0.0;
          grad_eta_chi[k][1] = (c == 1) ? gy : 0.0;
          grad_eta_phi[k][0] = (c == 2) ? gx : 0.0;
          grad_eta_phi[k][1] = (c == 2) ? gy : 0.0;
        }
      double grad_phi_mix[2];
      grad_phi_mix[0] = theta * grad_phi_n[q][0] + (1 - theta) * grad_phi_n_1[q][0];
      grad_phi_mix[1] = theta * grad_phi_n[q][1] + (1 - theta) * grad_phi_n_1[q][1];
      for (int i = 0; i < dpc; ++i)
        {
          /* mpi.cpp:388-417 */
          local_rhs[i] +=
            ((eta_psi[i] *
              (-psi_n[q] + psi_n_1[q] +
               dt * theta *
                 (2 * phi_n[q] + (1 + eps) * chi_n[q] + 3 * psi_n[q] * psi_n[q] * chi_n[q] +
                  6 * psi_n[q] * DOT2(grad_psi_n[q], grad_psi_n[q])) +
               dt * (1 - theta) *
                 (2 * phi_n_1[q] + (1 + eps) * chi_n_1[q] + 3 * psi_n_1[q] * psi_n_1[q] * chi_n_1[q] +
                  6 * psi_n_1[q] * DOT2(grad_psi_n_1[q], grad_psi_n_1[q])))) -
             (dt * DOT2(grad_eta_psi[i], grad_phi_mix)) - (eta_chi[i] * chi_n[q]) -
             (DOT2(grad_eta_chi[i], grad_psi_n[q])) - (eta_phi[i] * phi_n[q]) -
             (DOT2(grad_eta_phi[i], grad_chi_n[q]))) *
            JxW;
          for (int j = 0; j < dpc; ++j)
            {
              /* mpi.cpp:421-444 */
              local_matrix[i][j] +=
                ((eta_psi[i] *
                  (eta_psi[j] -
                   dt * theta *
                     (2 * eta_phi[j] + (1 + eps) * eta_chi[j] + 3 * psi_n[q] * psi_n[q] * eta_chi[j] +
                      6 * psi_n[q] * chi_n[q] * eta_psi[j] +
                      6 * DOT2(grad_psi_n[q], grad_psi_n[q]) * eta_psi[j] +
                      12 * psi_n[q] * DOT2(grad_psi_n[q], grad_eta_psi[j])))) +
                 (dt * theta * DOT2(grad_eta_psi[i], grad_eta_phi[j])) + (eta_chi[i] * eta_chi[j]) +
                 (DOT2(grad_eta_chi[i], grad_eta_psi[j])) + (eta_phi[i] * eta_phi[j]) +
                 (DOT2(grad_eta_phi[i], grad_eta_chi[j]))) *
                JxW;
              if (dead_work) /* mpi.cpp:446-448 */
                local_pre[i][j] += (eta_psi[i] * eta_psi[j]) * JxW;
            }
        }
    }
#undef DOT2

  /* mpi.cpp:453-458 [dealii-ext AffineConstraints::distribute_local_to_global, identity
   * constraints]: entry (i,j) goes to (m(i), m(j)); exact zeros are skipped; rhs_i goes to
   * m(i); every constrained local row adds |local(i,i)| (or the mean |diagonal| if that is
   * zero) to its own global diagonal, its rhs stays 0. */
  int n_constr = 0;
  for (int i = 0; i < dpc; ++i)
    {
      const int32_t gi = m->master[dofs[i]] >= 0 ? m->master[dofs[i]] : dofs[i];
      if (m->master[dofs[i]] >= 0)
        n_constr++;
      rhs[gi] += local_rhs[i];
      for (int j = 0; j < dpc; ++j)
        {
          if (local_matrix[i][j] == 0.0)
            continue;
          const int32_t gj = m->master[dofs[j]] >= 0 ? m->master[dofs[j]] : dofs[j];
          const int64_t pos = csr_find(m, gi, gj);
          if (pos < 0)
            {
              fprintf(stderr, "pfc_oracle: entry (%d,%d) not in sparsity pattern\n", gi, gj);
              abort();
            }
          values[pos] += local_matrix[i][j];
        }
    }
  if (n_constr > 0)
    {
      double average_diagonal = 0;
      for (int i = 0; i < dpc; ++i)
        average_diagonal += fabs(local_matrix[i][i]);
      average_diagonal /= dpc;
      for (int i = 0; i < dpc; ++i)
        if (m->master[dofs[i]] >= 0)
          {
            const double d = fabs(local_matrix[i][i]) != 0 ? fabs(local_matrix[i][i]) : average_diagonal;
            values[csr_find(m, dofs[i], dofs[i])] += d;
          }
    }
  if (dead_work)
    {
      double s = 0;
      for (int i = 0; i < dpc; ++i)
        s += local_pre[i][i];
      *dead_sink += s;
    }
}

void orc_assemble(const orc_mesh *m, const double *Psi_n, const double *Psi_n1, double dt,
                  double theta, double eps, double *values, double *rhs, int threads, int dead_work)
{
  /* mpi.cpp:305-306 */
  memset(values, 0, sizeof(double) * m->nnz);
  memset(rhs, 0, sizeof(double) * m->n_dofs);
  double sink = 0;
  if (threads <= 1)
    {
      for (int64_t cell = 0; cell < m->n_cells; ++cell) /* mpi.cpp:350 */
        assemble_cell(m, cell, Psi_n, Psi_n1, dt, theta, eps, values, rhs, dead_work, &sink);
    }
  else
    {
      /* timed mode: periodic 2x2 colouring makes the scatter conflict free (nx, ny even) */
      for (int colour = 0; colour < 4; ++colour)
        {
#pragma omp parallel for schedule(static) reduction(+ : sink) num_threads(threads)
          for (int64_t cell = 0; cell < m->n_cells; ++cell)
            {
              int cx, cy;
              cell_coords(cell, m->n_refines, &cx, &cy);
              if (((cx & 1) | ((cy & 1) << 1)) != colour)
                continue;
              double s = 0;
              assemble_cell(m, cell, Psi_n, Psi_n1, dt, theta, eps, values, rhs, dead_work, &s);
              sink += s;
            }
        }
    }
  if (dead_work && sink == 12345.678)
    fprintf(stderr, " ");
}

/* ------------------------------------------------------------------------------------------
 * Linear algebra: Epetra-style row-wise SpMV and BLAS-1 [dealii-ext].
 * ------------------------------------------------------------------------------------------ */
void orc_spmv(const orc_mesh *m, const double *values, const double *x, double *y)
{
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < m->n_dofs; ++r)
    {
      double s = 0;
      for (int64_t k = m->row_ptr[r]; k < m->row_ptr[r + 1]; ++k)
        s += values[k] * x[m->col_idx[k]];
      y[r] = s;
    }
}

static double vdot(int64_t n, const double *a, const double *b)
{
  double s = 0;
#pragma omp parallel for schedule(static) reduction(+ : s)
  for (int64_t i = 0; i < n; ++i)
    s += a[i] * b[i];
  return s;
}

/* deal.II Vector::add_and_dot(a, V, W): this += a V; return this . W */
static double add_and_dot(int64_t n, double *self, double a, const double *V, const double *W)
{
  double s = 0;
#pragma omp parallel for schedule(static) reduction(+ : s)
  for (int64_t i = 0; i < n; ++i)
    {
      self[i] += a * V[i];
      s += self[i] * W[i];
    }
  return s;
}

static void vaxpy(int64_t n, double *y, double a, const double *x)
{
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i)
    y[i] += a * x[i];
}

static void vscale(int64_t n, double *y, double a)
{
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i)
    y[i] *= a;
}

/* [dealii-ext] SolverGMRES::modified_gram_schmidt (deal.II 9.3/9.4) */
static double modified_gram_schmidt(int64_t n, double **V, int dim, double *vv, double *h,
                                    int *reorthogonalize)
{
  const int inner_iteration = dim - 1;
  double norm_vv_start = 0;
  const int consider = (*reorthogonalize == 0) && (inner_iteration % 5 == 4);
  if (consider)
    norm_vv_start = sqrt(vdot(n, vv, vv));

  h[0] = vdot(n, vv, V[0]);
  for (int i = 1; i < dim; ++i)
    h[i] = add_and_dot(n, vv, -h[i - 1], V[i - 1], V[i]);
  double norm_vv = sqrt(add_and_dot(n, vv, -h[dim - 1], V[dim - 1], vv));

  if (consider)
    {
      if (norm_vv > 10. * norm_vv_start * sqrt(2.220446049250313e-16))
        return norm_vv;
      *reorthogonalize = 1;
    }
  if (*reorthogonalize)
    {
      double htmp = vdot(n, vv, V[0]);
      h[0] += htmp;
      for (int i = 1; i < dim; ++i)
        {
          htmp = add_and_dot(n, vv, -htmp, V[i - 1], V[i]);
          h[i] += htmp;
        }
      norm_vv = sqrt(add_and_dot(n, vv, -htmp, V[dim - 1], vv));
    }
  return norm_vv;
}

/* [dealii-ext] SolverGMRES::givens_rotation */
static void givens_rotation(double *h, double *b, double *ci, double *si, int col)
{
  for (int i = 0; i < col; i++)
    {
      const double s = si[i], c = ci[i], dummy = h[i];
      h[i] = c * dummy + s * h[i + 1];
      h[i + 1] = -s * dummy + c * h[i + 1];
    }
  const double r = 1. / sqrt(h[col] * h[col] + h[col + 1] * h[col + 1]);
  si[col] = h[col + 1] * r;
  ci[col] = h[col] * r;
  h[col] = ci[col] * h[col] + si[col] * h[col + 1];
  b[col + 1] = -si[col] * b[col];
  b[col] *= ci[col];
}

/* [dealii-ext] SolverControl::check: 0 iterate, 1 success, 2 failure */
static int control_check(int64_t step, double value, double tol, int64_t max_steps)
{
  if (value <= tol)
    return 1;
  if (step >= max_steps || isnan(value))
    return 2;
  return 0;
}

int64_t orc_gmres(const orc_mesh *m, const double *values, const double *b, double *x, double tol,
                  int64_t max_iters, int n_tmp_vectors, int *converged, double *final_res,
                  int64_t stop_after)
{
  const int64_t n = m->n_dofs;
  const int nt = n_tmp_vectors; /* 30 by default => restart length nt - 2 = 28 */
  double **tmp = (double **)xmalloc(sizeof(double *) * nt);
  for (int i = 0; i < nt; ++i)
    tmp[i] = (double *)xmalloc(sizeof(double) * n);
  double *H = (double *)calloc((size_t)nt * (nt - 1), sizeof(double));
  double *gamma = (double *)calloc(nt, sizeof(double));
  double *ci = (double *)calloc(nt - 1, sizeof(double));
  double *si = (double *)calloc(nt - 1, sizeof(double));
  double *h = (double *)calloc(nt, sizeof(double));
  double *v = tmp[0], *p = tmp[nt - 1];

  memset(x, 0, sizeof(double) * n); /* mpi.cpp:481 */
  int64_t accumulated_iterations = 0;
  int dim = 0, state = 0, reorthogonalize = 0;
  double rho = 0;
  do
    {
      memset(h, 0, sizeof(double) * nt);
      /* left preconditioning with identity: v = b - A x */
      orc_spmv(m, values, x, p);
#pragma omp parallel for schedule(static)
      for (int64_t i = 0; i < n; ++i)
        v[i] = b[i] - p[i];
      rho = sqrt(vdot(n, v, v));
      state = control_check(accumulated_iterations, rho, tol, max_iters);
      if (state != 0)
        break;
      gamma[0] = rho;
      vscale(n, v, 1. / rho);
      dim = 0;
      for (int inner = 0; inner < nt - 2 && state == 0; ++inner)
        {
          ++accumulated_iterations;
          double *vv = tmp[inner + 1];
          orc_spmv(m, values, tmp[inner], vv); /* A.vmult(p, tmp[inner]); precondition = copy */
          dim = inner + 1;
          const double s = modified_gram_schmidt(n, tmp, dim, vv, h, &reorthogonalize);
          h[inner + 1] = s;
          if (s != 0)
            vscale(n, vv, 1. / s);
          givens_rotation(h, gamma, ci, si, inner);
          for (int i = 0; i < dim; ++i)
            H[i * (nt - 1) + inner] = h[i];
          rho = fabs(gamma[dim]);
          state = control_check(accumulated_iterations, rho, tol, max_iters);
          if (stop_after > 0 && accumulated_iterations >= stop_after && state == 0)
            state = 3;
        }
      /* H1.backward(h, gamma): upper-triangular back substitution */
      for (int i = dim - 1; i >= 0; --i)
        {
          double s = gamma[i];
          for (int j = i + 1; j < dim; ++j)
            s -= H[i * (nt - 1) + j] * h[j];
          h[i] = s / H[i * (nt - 1) + i];
        }
      for (int i = 0; i < dim; ++i)
        vaxpy(n, x, h[i], tmp[i]);
    }
  while (state == 0);

  *converged = (state == 1);
  *final_res = rho;
  for (int i = 0; i < nt; ++i)
    free(tmp[i]);
  free(tmp);
  free(H);
  free(gamma);
  free(ci);
  free(si);
  free(h);
  return accumulated_iterations;
}

/* ------------------------------------------------------------------------------------------
 * Newton loop.  mpi.cpp:576-593 with solve_and_update_no_precondition mpi.cpp:471-492.
 * ------------------------------------------------------------------------------------------ */
int orc_newton_step(const orc_mesh *m, double *Psi_n, double *Psi_n1, double dt, double theta,
                    double eps, double newton_tol, int newton_max, double gmres_tol, int threads,
                    double *res_hist, int64_t *gmres_hist, int *n_assemblies, int *status)
{
  double *values = (double *)xmalloc(sizeof(double) * m->nnz);
  double *rhs = (double *)xmalloc(sizeof(double) * m->n_dofs);
  double *dPsi = (double *)xmalloc(sizeof(double) * m->n_dofs);
  int n_solves = 0;
  *n_assemblies = 0;
  *status = 1;
  for (int it = 0; it < newton_max; ++it) /* mpi.cpp:579 */
    {
      orc_assemble(m, Psi_n, Psi_n1, dt, theta, eps, values, rhs, threads, 0); /* :582 */
      ++*n_assemblies;
      const double residual = sqrt(vdot(m->n_dofs, rhs, rhs)); /* :584 */
      res_hist[it] = residual;
      if (residual < newton_tol) /* :586 */
        {
          *status = 0;
          break;
        }
      int conv;
      double fr;
      /* :478 SolverControl(n_dofs) => max steps n_dofs, tol 1e-10 [dealii-ext default] */
      gmres_hist[n_solves] = orc_gmres(m, values, rhs, dPsi, gmres_tol, m->n_dofs, 30, &conv, &fr, 0);
      ++n_solves;
      if (!conv)
        {
          *status = 2;
          break;
        }
      orc_distribute(m, dPsi); /* :485 */
      vaxpy(m->n_dofs, Psi_n, 1.0, dPsi); /* :487-491 */
    }
  memcpy(Psi_n1, Psi_n, sizeof(double) * m->n_dofs); /* :592 */
  free(values);
  free(rhs);
  free(dPsi);
  return n_solves;
}

/* ------------------------------------------------------------------------------------------
 * Stress projection.  stress.cpp = src/stress_tools/stress_calculator_mpi.cpp.
 * The reference's FESystem(FE_Q(degree), dim*dim) shares the scalar FE_Q(degree) node set with
 * the (psi, chi, phi) system, so component c = a*dim + b of the stress lives on scalar node t
 * at stress[c*N + t] (dof numbering of the psi block; the reference's own block-wise
 * renumbering, stress.cpp:46-58, is a permutation of this).
 * ------------------------------------------------------------------------------------------ */
static void scalar_cell_nodes(const orc_mesh *m, int64_t cell, int32_t *node)
{
  /* node[a] = global dof of the psi component carried by scalar shape a (dof = 0*N + t) */
  const int32_t *dofs = m->cell_dofs + cell * m->dofs_per_cell;
  for (int k = 0; k < m->dofs_per_cell; ++k)
    if (m->local_comp[k] == 0)
      node[m->local_shape[k]] = dofs[k];
}

/* stress.cpp:141-200 — block-diagonal mass matrix; every diagonal block is the same scalar
 * matrix, stored here once on the rows/columns [0,N) of the monolithic pattern. */
static void stress_mass_matrix(const orc_mesh *m, double *values)
{
  const int ns = m->n_s, nq = m->n_q;
  memset(values, 0, sizeof(double) * m->nnz);
  for (int64_t cell = 0; cell < m->n_cells; ++cell)
    {
      const double *vtx = m->cell_vertices + cell * 8;
      const double hx = vtx[2] - vtx[0], hy = vtx[5] - vtx[1];
      int32_t node[MAXQ];
      double local_matrix[MAXQ][MAXQ];
      scalar_cell_nodes(m, cell, node);
      memset(local_matrix, 0, sizeof(local_matrix));
      for (int q = 0; q < nq; ++q)
        {
          const double JxW = m->qw[q] * hx * hy;
          for (int i = 0; i < ns; ++i)
            for (int j = 0; j < ns; ++j)
              local_matrix[i][j] += m->shape[q * ns + i] * m->shape[q * ns + j] * JxW; /* :183-187 */
        }
      /* stress.cpp:193-196 [dealii-ext distribute_local_to_global, identity constraints] */
      for (int i = 0; i < ns; ++i)
        {
          const int32_t gi = m->master[node[i]] >= 0 ? m->master[node[i]] : node[i];
          for (int j = 0; j < ns; ++j)
            {
              const int32_t gj = m->master[node[j]] >= 0 ? m->master[node[j]] : node[j];
              values[csr_find(m, gi, gj)] += local_matrix[i][j];
            }
          if (m->master[node[i]] >= 0)
            values[csr_find(m, node[i], node[i])] += fabs(local_matrix[i][i]);
        }
    }
}

/* stress.cpp:204-314 — right-hand side of the L2 projection, rhs[c*N + t] */
static void stress_rhs(const orc_mesh *m, const double *Psi, double eps, double *rhs)
{
  const int ns = m->n_s, nq = m->n_q, dpc = m->dofs_per_cell;
  const int64_t N = m->N;
  memset(rhs, 0, sizeof(double) * 4 * N);
  for (int64_t cell = 0; cell < m->n_cells; ++cell)
    {
      const int32_t *dofs = m->cell_dofs + cell * dpc;
      const double *vtx = m->cell_vertices + cell * 8;
      const double hx = vtx[2] - vtx[0], hy = vtx[5] - vtx[1];
      int32_t node[MAXQ];
      double local_rhs[4][MAXQ];
      scalar_cell_nodes(m, cell, node);
      memset(local_rhs, 0, sizeof(local_rhs));
      for (int q = 0; q < nq; ++q)
        {
          /* stress.cpp:262-267 */
          double psi = 0, chi = 0, grad_psi[2] = {0, 0}, grad_chi[2] = {0, 0};
          for (int k = 0; k < dpc; ++k)
            {
              const int c = m->local_comp[k], a = m->local_shape[k];
              const double u = Psi[dofs[k]];
              const double Nv = m->shape[q * ns + a];
              const double gx = m->dshape[(q * ns + a) * 2 + 0] / hx, gy = m->dshape[(q * ns + a) * 2 + 1] / hy;
              if (c == 0)
                {
                  psi += u * Nv;
                  grad_psi[0] += u * gx;
                  grad_psi[1] += u * gy;
                }
              else if (c == 1)
                {
                  chi += u * Nv;
                  grad_chi[0] += u * gx;
                  grad_chi[1] += u * gy;
                }
            }
          const double JxW = m->qw[q] * hx * hy;
          const double G[2] = {grad_psi[0] + grad_chi[0], grad_psi[1] + grad_chi[1]};
          for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 2; ++b)
              for (int i = 0; i < ns; ++i)
                {
                  /* the shape function of component (a,b) on scalar shape i:
                   * tau = N e_a (x) e_b, (div tau)_a = d_b N [dealii-ext Tensor<2> view],
                   * tr tau = delta_ab N, grad tr tau = delta_ab grad N (stress.cpp:271-280) */
                  const double Nv = m->shape[q * ns + i];
                  const double gN[2] = {m->dshape[(q * ns + i) * 2 + 0] / hx, m->dshape[(q * ns + i) * 2 + 1] / hy};
                  const double tr = (a == b) ? Nv : 0.0;
                  double v = Nv * (G[a] * grad_psi[b] + grad_psi[a] * G[b]);             /* :284-289 */
                  v += gN[b] * (psi + chi) * grad_psi[a];                                   /* :291 */
                  v += 0.5 * tr * (chi * chi + (1 + eps) * psi * psi + 0.5 * psi * psi * psi * psi); /* :293-295 */
                  if (a == b)
                    v -= (gN[0] * grad_psi[0] + gN[1] * grad_psi[1]) * psi;                 /* :297 */
                  v -= tr * (grad_psi[0] * grad_psi[0] + grad_psi[1] * grad_psi[1]);       /* :299 */
                  local_rhs[a * 2 + b][i] += v * JxW;
                }
        }
      for (int c = 0; c < 4; ++c)
        for (int i = 0; i < ns; ++i)
          {
            const int32_t gi = m->master[node[i]] >= 0 ? m->master[node[i]] : node[i];
            rhs[c * N + gi] += local_rhs[c][i];
          }
    }
}

/* [dealii-ext] SolverCG with PreconditionIdentity as called at stress.cpp:319-333: x0 = 0,
 * absolute tolerance tol on ||b - A x||, at most max_iters steps. */
static int64_t stress_cg(const orc_mesh *m, const double *values, const double *b, double *x, double tol,
                         int64_t max_iters, int *converged)
{
  const int64_t N = m->N;
  double *g = (double *)xmalloc(sizeof(double) * N), *d = (double *)xmalloc(sizeof(double) * N);
  double *h = (double *)xmalloc(sizeof(double) * N);
  memset(x, 0, sizeof(double) * N);
  for (int64_t i = 0; i < N; ++i)
    g[i] = -b[i];
  double res = sqrt(vdot(N, g, g));
  int64_t it = 0;
  int state = control_check(0, res, tol, max_iters);
  if (state == 0)
    {
      for (int64_t i = 0; i < N; ++i)
        d[i] = -g[i];
      double gh = res * res;
      while (state == 0)
        {
          it++;
          for (int64_t r = 0; r < N; ++r)
            {
              double s = 0;
              for (int64_t k = m->row_ptr[r]; k < m->row_ptr[r + 1] && m->col_idx[k] < N; ++k)
                s += values[k] * d[m->col_idx[k]];
              h[r] = s;
            }
          const double alpha = gh / vdot(N, d, h);
          vaxpy(N, x, alpha, d);
          vaxpy(N, g, alpha, h);
          res = sqrt(vdot(N, g, g));
          state = control_check(it, res, tol, max_iters);
          if (state != 0)
            break;
          const double beta = res * res / gh;
          gh = res * res;
          for (int64_t i = 0; i < N; ++i)
            d[i] = beta * d[i] - g[i];
        }
    }
  *converged = (state == 1);
  free(g);
  free(d);
  free(h);
  return it;
}

int orc_stress(const orc_mesh *m, const double *Psi, double eps, double tol, double *stress,
               int64_t *cg_iters, double *rhs_out)
{
  const int64_t N = m->N;
  double *values = (double *)xmalloc(sizeof(double) * m->nnz);
  double *rhs = (double *)xmalloc(sizeof(double) * 4 * N);
  int ok = 1;
  stress_mass_matrix(m, values);
  stress_rhs(m, Psi, eps, rhs);
  if (rhs_out)
    memcpy(rhs_out, rhs, sizeof(double) * 4 * N);
  for (int c = 0; c < 4; ++c) /* stress.cpp:327-334; SolverControl(n_dofs) of the 4N-dof handler */
    {
      int conv = 0;
      cg_iters[c] = stress_cg(m, values, rhs + c * N, stress + c * N, tol, 4 * N, &conv);
      ok &= conv;
    }
  /* stress.cpp:336 — constraints.distribute */
  for (int c = 0; c < 4; ++c)
    for (int64_t t = 0; t < N; ++t)
      if (m->master[t] >= 0)
        stress[c * N + t] = stress[c * N + m->master[t]];
  free(values);
  free(rhs);
  return ok ? 0 : 2;
}
