// pfc_standin.cpp — host stand-in for the deal.II objects the reference owns.
//
// deal.II / p4est are not available in this image, so the flattened arrays a deal.II host
// would extract from Triangulation / DoFHandler / AffineConstraints / FEValues are produced
// here for the one mesh family the reference builds:
//   make_grid   mpi.cpp:115-146  hyper_rectangle(colorize), periodic in x and y, refine_global(n)
//   setup_dofs  mpi.cpp:150-217  FESystem(FE_Q(p)^3), distribute_dofs, component_wise,
//                                make_periodicity_constraints on boundary pairs (0,1), (2,3)
// [dealii-ext] facts used (deal.II >= 9.3 documented behaviour): active cells of a globally
// refined single coarse cell are in Morton order; FE_Q local numbering is vertices, lines
// (x=0, x=1, y=0, y=1), interior; distribute_dofs numbers cell by cell on first touch;
// component_wise keeps the relative order inside a component; p4est cuts the Morton curve
// into contiguous equal chunks and a dof on a rank interface belongs to the lowest rank —
// which makes the global numbering independent of the number of ranks.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/pfc_b200.h"

struct pfc_standin
{
  pfc_mesh mesh;
  int degree, n_refines;
  int64_t N;
  double x0, y0, x1, y1;
  std::vector<double> cell_vertices, support, shape, dshape, qw;
  std::vector<int32_t> cell_dofs, local_comp, local_shape, c_slave, c_master;
  std::vector<int64_t> node_offsets;
  std::vector<int32_t> node_master; // resolved, per scalar node
};

namespace
{
inline void morton_decode(int64_t cell, int n, int &cx, int &cy)
{
  cx = cy = 0;
  for (int k = 0; k < n; ++k)
    {
      cx |= (int)((cell >> (2 * k)) & 1) << k;
      cy |= (int)((cell >> (2 * k + 1)) & 1) << k;
    }
}
inline int64_t morton_encode(int cx, int cy, int n)
{
  int64_t c = 0;
  for (int k = 0; k < n; ++k)
    c |= ((int64_t)((cx >> k) & 1) << (2 * k)) | ((int64_t)((cy >> k) & 1) << (2 * k + 1));
  return c;
}
// Lagrange polynomial k of degree p on equidistant nodes, and its derivative
double lag(int p, int k, double x)
{
  double v = 1;
  for (int m = 0; m <= p; ++m)
    if (m != k)
      v *= (p * x - m) / double(k - m);
  return v;
}
double dlag(int p, int k, double x)
{
  double s = 0;
  for (int l = 0; l <= p; ++l)
    if (l != k)
      {
        double t = p / double(k - l);
        for (int m = 0; m <= p; ++m)
          if (m != k && m != l)
            t *= (p * x - m) / double(k - m);
        s += t;
      }
  return s;
}
} // namespace

extern "C" int pfc_standin_create_ex(int32_t degree, int32_t n_refines, int32_t rep_x, int32_t rep_y, double x0,
                                     double y0, double x1, double y1, int32_t rank, int32_t n_ranks,
                                     pfc_standin **out);

extern "C" int pfc_standin_create(int32_t degree, int32_t n_refines, double x0, double y0, double x1,
                                  double y1, int32_t rank, int32_t n_ranks, pfc_standin **out)
{
  return pfc_standin_create_ex(degree, n_refines, 1, 1, x0, y0, x1, y1, rank, n_ranks, out);
}

// rep_x x rep_y coarse cells (GridGenerator::subdivided_hyper_rectangle; not reachable upstream, where
// the coarse mesh is one cell): trees in lexicographic order, Morton order inside each tree, which is
// the order p4est walks its forest.
extern "C" int pfc_standin_create_ex(int32_t degree, int32_t n_refines, int32_t rep_x, int32_t rep_y, double x0,
                                     double y0, double x1, double y1, int32_t rank, int32_t n_ranks,
                                     pfc_standin **out)
{
  if (!out)
    return PFC_ERR_BAD_ARGUMENT;
  *out = nullptr;
  if (degree < 1 || degree > 2 || n_refines < 1 || n_refines > 13 || !(x1 > x0) || !(y1 > y0) || rep_x < 1 ||
      rep_y < 1 || rep_x > 64 || rep_y > 64)
    return PFC_ERR_BAD_ARGUMENT;
  if (n_ranks < 1)
    n_ranks = 1;
  if (rank < 0 || rank >= n_ranks)
    return PFC_ERR_BAD_ARGUMENT;
  const int p = degree, ns = (p + 1) * (p + 1), dpc = 3 * ns;
  const int nt = 1 << n_refines; // cells per tree and direction
  const int nx = rep_x * nt, ny = rep_y * nt;
  const int64_t cells_per_tree = (int64_t)nt * nt;
  const int64_t n_cells_g = (int64_t)nx * ny;
  auto cell_xy = [&](int64_t cell, int &cx, int &cy) {
    const int64_t tree = cell / cells_per_tree;
    morton_decode(cell % cells_per_tree, n_refines, cx, cy);
    cx += (int)(tree % rep_x) * nt;
    cy += (int)(tree / rep_x) * nt;
  };
  auto cell_of = [&](int cx, int cy) {
    const int64_t tree = (int64_t)(cy / nt) * rep_x + cx / nt;
    return tree * cells_per_tree + morton_encode(cx % nt, cy % nt, n_refines);
  };
  if (n_ranks > n_cells_g)
    return PFC_ERR_BAD_ARGUMENT;
  const int64_t G = (int64_t)p * nx + 1, Gy = (int64_t)p * ny + 1; // fine-grid points per direction
  const int64_t N = G * Gy;
  pfc_standin *S = new (std::nothrow) pfc_standin();
  if (!S)
    return PFC_ERR_BAD_ARGUMENT;
  S->degree = p;
  S->n_refines = n_refines;
  S->N = N;
  S->x0 = x0, S->y0 = y0, S->x1 = x1, S->y1 = y1;
  const double hx = (x1 - x0) / nx, hy = (y1 - y0) / ny;

  // FE_Q(p) hierarchical local nodes as fine-grid offsets
  int ox[9], oy[9];
  if (p == 1)
    {
      const int X[4] = {0, 1, 0, 1}, Y[4] = {0, 0, 1, 1};
      for (int a = 0; a < 4; ++a)
        ox[a] = X[a], oy[a] = Y[a];
    }
  else
    {
      const int X[9] = {0, 2, 0, 2, 0, 2, 1, 1, 1}, Y[9] = {0, 0, 2, 2, 1, 1, 0, 2, 1};
      for (int a = 0; a < 9; ++a)
        ox[a] = X[a], oy[a] = Y[a];
    }

  // QGauss(p+1) x FE_Q(p) tables, quadrature point index x-fastest (mpi.cpp:308-315)
  double gp[3], gw[3];
  if (p == 1)
    {
      gp[0] = 0.5 - 0.5 / std::sqrt(3.0);
      gp[1] = 0.5 + 0.5 / std::sqrt(3.0);
      gw[0] = gw[1] = 0.5;
    }
  else
    {
      gp[0] = 0.5 - 0.5 * std::sqrt(0.6);
      gp[1] = 0.5;
      gp[2] = 0.5 + 0.5 * std::sqrt(0.6);
      gw[0] = gw[2] = 5.0 / 18.0;
      gw[1] = 8.0 / 18.0;
    }
  S->shape.resize(ns * ns);
  S->dshape.resize(ns * ns * 2);
  S->qw.resize(ns);
  for (int qy = 0; qy <= p; ++qy)
    for (int qx = 0; qx <= p; ++qx)
      {
        const int q = qx + (p + 1) * qy;
        S->qw[q] = gw[qx] * gw[qy];
        for (int a = 0; a < ns; ++a)
          {
            S->shape[q * ns + a] = lag(p, ox[a], gp[qx]) * lag(p, oy[a], gp[qy]);
            S->dshape[(q * ns + a) * 2 + 0] = dlag(p, ox[a], gp[qx]) * lag(p, oy[a], gp[qy]);
            S->dshape[(q * ns + a) * 2 + 1] = lag(p, ox[a], gp[qx]) * dlag(p, oy[a], gp[qy]);
          }
      }
  // FESystem interleaves components per geometric object: local i = 3 a + c (mpi.cpp:84-86)
  S->local_comp.resize(dpc);
  S->local_shape.resize(dpc);
  for (int i = 0; i < dpc; ++i)
    {
      S->local_comp[i] = i % 3;
      S->local_shape[i] = i / 3;
    }

  // global first-touch numbering over the Morton curve + ownership offsets
  std::vector<int32_t> id(N, -1);
  S->support.resize((size_t)N * 2);
  S->node_offsets.assign(n_ranks + 1, 0);
  auto chunk_begin = [&](int r) { return n_cells_g * r / n_ranks; };
  {
    int32_t next = 0;
    int r = 0;
    for (int64_t cell = 0; cell < n_cells_g; ++cell)
      {
        while (r + 1 <= n_ranks && cell == chunk_begin(r + 1))
          S->node_offsets[++r] = next;
        int cx, cy;
        cell_xy(cell, cx, cy);
        for (int a = 0; a < ns; ++a)
          {
            const int64_t fx = (int64_t)p * cx + ox[a], fy = (int64_t)p * cy + oy[a];
            int32_t &t = id[fy * G + fx];
            if (t < 0)
              {
                t = next++;
                S->support[2 * (int64_t)t] = x0 + fx * (hx / p);
                S->support[2 * (int64_t)t + 1] = y0 + fy * (hy / p);
              }
          }
      }
    for (int q = r + 1; q <= n_ranks; ++q)
      S->node_offsets[q] = next;
  }

  // periodic constraints: x_min -> x_max, y_min -> y_max, corners chain to (x_max, y_max)
  S->node_master.resize(N);
  for (int64_t fy = 0; fy < Gy; ++fy)
    for (int64_t fx = 0; fx < G; ++fx)
      {
        const int64_t mx = fx == 0 ? G - 1 : fx, my = fy == 0 ? Gy - 1 : fy;
        S->node_master[id[fy * G + fx]] = id[my * G + mx];
      }
  for (int c = 0; c < 3; ++c)
    for (int64_t t = 0; t < N; ++t)
      if (S->node_master[t] != t)
        {
          S->c_slave.push_back((int32_t)(c * N + t));
          S->c_master.push_back((int32_t)(c * N + S->node_master[t]));
        }

  // cells of this rank: owned Morton chunk, then the cells needed to complete owned rows
  const int64_t c0 = chunk_begin(rank), c1 = chunk_begin(rank + 1);
  std::vector<int64_t> cells;
  for (int64_t c = c0; c < c1; ++c)
    cells.push_back(c);
  if (n_ranks > 1)
    {
      const int64_t o0 = S->node_offsets[rank], o1 = S->node_offsets[rank + 1];
      auto mine = [&](int32_t t) { return t >= o0 && t < o1; };
      // candidate ghost cells: the 8 neighbours (periodic wrap) of every owned cell, plus owned test
      std::vector<int8_t> mark(n_cells_g, 0);
      for (int64_t c = c0; c < c1; ++c)
        mark[c] = 1;
      std::vector<int64_t> extra;
      for (int64_t c = c0; c < c1; ++c)
        {
          int cx, cy;
          cell_xy(c, cx, cy);
          for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx)
              {
                const int ax = (cx + dx + nx) % nx, ay = (cy + dy + ny) % ny;
                const int64_t nb = cell_of(ax, ay);
                if (mark[nb])
                  continue;
                bool need = false;
                for (int a = 0; a < ns && !need; ++a)
                  {
                    const int32_t t = id[((int64_t)p * ay + oy[a]) * G + (int64_t)p * ax + ox[a]];
                    need = mine(t) || mine(S->node_master[t]);
                  }
                if (need)
                  {
                    mark[nb] = 1;
                    extra.push_back(nb);
                  }
              }
        }
      // a master owned here can also be reached only through periodic images of far cells
      // (e.g. the corner master): scan the boundary cells once
      for (int64_t c = 0; c < n_cells_g; ++c)
        {
          int cx, cy;
          cell_xy(c, cx, cy);
          if (mark[c] || (cx != 0 && cy != 0 && cx != nx - 1 && cy != ny - 1))
            continue;
          bool need = false;
          for (int a = 0; a < ns && !need; ++a)
            {
              const int32_t t = id[((int64_t)p * cy + oy[a]) * G + (int64_t)p * cx + ox[a]];
              need = mine(t) || mine(S->node_master[t]);
            }
          if (need)
            {
              mark[c] = 1;
              extra.push_back(c);
            }
        }
      std::sort(extra.begin(), extra.end());
      cells.insert(cells.end(), extra.begin(), extra.end());
    }

  const int64_t nc = (int64_t)cells.size();
  S->cell_dofs.resize((size_t)nc * dpc);
  S->cell_vertices.resize((size_t)nc * 8);
  for (int64_t k = 0; k < nc; ++k)
    {
      int cx, cy;
      cell_xy(cells[k], cx, cy);
      for (int v = 0; v < 4; ++v)
        {
          S->cell_vertices[k * 8 + v * 2 + 0] = x0 + (cx + (v & 1)) * hx;
          S->cell_vertices[k * 8 + v * 2 + 1] = y0 + (cy + (v >> 1)) * hy;
        }
      for (int i = 0; i < dpc; ++i)
        {
          const int a = i / 3, c = i % 3;
          const int32_t t = id[((int64_t)p * cy + oy[a]) * G + (int64_t)p * cx + ox[a]];
          S->cell_dofs[k * dpc + i] = (int32_t)(c * N + t);
        }
    }

  pfc_mesh &M = S->mesh;
  std::memset(&M, 0, sizeof(M));
  M.n_cells = nc;
  M.n_owned_cells = c1 - c0;
  M.n_dofs = 3 * N;
  M.dofs_per_cell = dpc;
  M.n_q = ns;
  M.cell_vertices = S->cell_vertices.data();
  M.cell_dofs = S->cell_dofs.data();
  M.local_component = S->local_comp.data();
  M.local_scalar_index = S->local_shape.data();
  M.shape_value = S->shape.data();
  M.shape_grad_ref = S->dshape.data();
  M.q_weight = S->qw.data();
  for (int c = 0; c < 3; ++c)
    {
      M.block_start[c] = c * N;
      M.block_size[c] = N;
    }
  M.n_constraints = (int64_t)S->c_slave.size();
  M.constraint_slave = S->c_slave.data();
  M.constraint_master = S->c_master.data();
  M.row_ptr = nullptr;
  M.col_idx = nullptr;
  M.rank = rank;
  M.n_ranks = n_ranks;
  M.node_offsets = S->node_offsets.data();
  *out = S;
  return PFC_OK;
}

extern "C" void pfc_standin_destroy(pfc_standin *s) { delete s; }
extern "C" const pfc_mesh *pfc_standin_mesh(const pfc_standin *s) { return s ? &s->mesh : nullptr; }
extern "C" int64_t pfc_standin_n_nodes(const pfc_standin *s) { return s ? s->N : 0; }
extern "C" const double *pfc_standin_support_points(const pfc_standin *s) { return s ? s->support.data() : nullptr; }

// HexagonalLattice<dim>::value, hexagonal_lattice.cpp:52-77 (q vectors :14-21), interpolated
// at the FE_Q support points (VectorTools::interpolate, mpi.cpp:285-287) followed by
// constraints.distribute (mpi.cpp:288).
extern "C" int pfc_standin_hexagonal_lattice(const pfc_standin *s, double A0, double psi0, int32_t n_disl,
                                             const double *pos, const double *burg, double *Psi)
{
  if (!s || !Psi || n_disl < 0 || (n_disl > 0 && (!pos || !burg)))
    return PFC_ERR_BAD_ARGUMENT;
  const int64_t N = s->N;
  const double q[3][2] = {{0.0, 1.0}, {0.5 * std::sqrt(3.0), -0.5}, {-0.5 * std::sqrt(3.0), -0.5}};
#pragma omp parallel for schedule(static)
  for (int64_t t = 0; t < N; ++t)
    {
      const double px = s->support[2 * t], py = s->support[2 * t + 1];
      double sum = 0;
      for (int n = 0; n < 3; ++n)
        {
          double phase = 0;
          for (int j = 0; j < n_disl; ++j)
            {
              const double snj = 1.0 / (2 * M_PI) * (q[n][0] * burg[2 * j] + q[n][1] * burg[2 * j + 1]);
              phase += snj * std::atan2(py - pos[2 * j + 1], px - pos[2 * j]);
            }
          sum += std::cos((q[n][0] * px + q[n][1] * py) - phase);
        }
      Psi[t] = psi0 + A0 * sum;
      Psi[N + t] = -A0 * sum;
      Psi[2 * N + t] = A0 * sum;
    }
  for (int c = 0; c < 3; ++c)
    for (int64_t t = 0; t < N; ++t)
      if (s->node_master[t] != t)
        Psi[c * N + t] = Psi[c * N + s->node_master[t]];
  return PFC_OK;
}

// Polycrystalline initial condition (harness extension for BASELINE.json config 4, not in the reference): periodic
// Voronoi grains (minimum-image distance to seeds[n][2], ties to the lowest index); grain g carries the cos-sum of
// HexagonalLattice<dim>::value (hexagonal_lattice.cpp:52-77, no dislocations) in its own frame
// p' = R(alpha_g)^T rho + o_g, rho = minimum-image displacement from the seed.  Same operations in the same order as
// the CPU checker's copy (test infrastructure): the two are compared bit for bit (tests/test_plan_cpu.py).
extern "C" int pfc_standin_polycrystal(const pfc_standin *s, double A0, double psi0, int32_t n_grains, const double *seeds,
                                       const double *alpha, const double *offset, double *Psi)
{
  if (!s || !Psi || n_grains < 1 || !seeds || !alpha || !offset)
    return PFC_ERR_BAD_ARGUMENT;
  const int64_t N = s->N;
  const double W = s->x1 - s->x0, H = s->y1 - s->y0;
  const double q[3][2] = {{0.0, 1.0}, {0.5 * std::sqrt(3.0), -0.5}, {-0.5 * std::sqrt(3.0), -0.5}};
#pragma omp parallel for schedule(static)
  for (int64_t t = 0; t < N; ++t)
    {
      const double px = s->support[2 * t], py = s->support[2 * t + 1];
      int best = 0;
      double bd = 0, bx = 0, by = 0;
      for (int g = 0; g < n_grains; ++g)
        {
          double dx = px - seeds[2 * g], dy = py - seeds[2 * g + 1];
          dx -= W * std::rint(dx / W);
          dy -= H * std::rint(dy / H);
          const double d2 = dx * dx + dy * dy;
          if (g == 0 || d2 < bd)
            {
              best = g;
              bd = d2;
              bx = dx;
              by = dy;
            }
        }
      const double ca = std::cos(alpha[best]), sa = std::sin(alpha[best]);
      const double qx = ca * bx + sa * by + offset[2 * best];
      const double qy = ca * by - sa * bx + offset[2 * best + 1];
      double sum = 0;
      for (int n = 0; n < 3; ++n)
        sum += std::cos((q[n][0] * qx + q[n][1] * qy) - 0.0);
      Psi[t] = psi0 + A0 * sum;
      Psi[N + t] = -A0 * sum;
      Psi[2 * N + t] = A0 * sum;
    }
  for (int c = 0; c < 3; ++c)
    for (int64_t t = 0; t < N; ++t)
      if (s->node_master[t] != t)
        Psi[c * N + t] = Psi[c * N + s->node_master[t]];
  return PFC_OK;
}
