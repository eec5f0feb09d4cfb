// pfc_lattice_plan.cpp — host detection of the uniform tensor lattice behind the arrays of pfc_mesh.
//
// The reference's mesh is always one hyper_rectangle refined globally and made periodic in x and y
// (make_grid, mpi.cpp:115-146), FE_Q(p)^3 with QGauss(p+1) (mpi.cpp:84-86,308).  Nothing in the C ABI says
// so: the structure is re-discovered here from cell vertices, dof maps, constraints and the shape tables,
// and every assumption is verified; any mismatch leaves the general (table-driven) path in charge.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>

#include "pfc_plan.h"

namespace pfc
{
namespace
{
const double kInvSqrt3 = 0.57735026918962576451, kSqrt06 = 0.77459666924148337704;

double gauss_point(int p, int q)
{
  if (p == 1)
    return q == 0 ? 0.5 - 0.5 * kInvSqrt3 : 0.5 + 0.5 * kInvSqrt3;
  return q == 0 ? 0.5 - 0.5 * kSqrt06 : (q == 1 ? 0.5 : 0.5 + 0.5 * kSqrt06);
}
} // namespace

double lattice_weight_1d(int p, int q)
{
  if (p == 1)
    return 0.5;
  return q == 1 ? 8.0 / 18.0 : 5.0 / 18.0;
}

double lattice_shape_1d(int p, int q, int j)
{
  const double x = gauss_point(p, q);
  if (p == 1)
    return j == 0 ? 1.0 - x : x;
  return j == 0 ? 2.0 * (x - 0.5) * (x - 1.0) : (j == 1 ? 4.0 * x * (1.0 - x) : 2.0 * x * (x - 0.5));
}

double lattice_dshape_1d(int p, int q, int j)
{
  const double x = gauss_point(p, q);
  if (p == 1)
    return j == 0 ? -1.0 : 1.0;
  return j == 0 ? 4.0 * x - 3.0 : (j == 1 ? 4.0 - 8.0 * x : 4.0 * x - 1.0);
}

bool detect_lattice(const pfc_mesh &m, Plan &P, const std::vector<int32_t> &cnode,
                    const std::vector<int32_t> &master, const std::vector<double> &h_of)
{
  Lattice &L = P.lat;
  L = Lattice();
  auto no = [&](const char *why) {
    L.why = why;
    L.active = false;
    return false;
  };
  const int p = P.degree, n1 = p + 1, ns = P.n_s;
  const int64_t N = P.N_global;
  L.p = p;

  // ---- 1. local scalar order -> (jx, jy): deal.II hierarchical or lexicographic; the caller's tables
  //         must be the tensor FE_Q(p) x QGauss(p+1) tables (q index x-fastest) in that order ----------
  int ax[9], ay[9];
  bool found = false;
  for (int cand = 0; cand < 2 && !found; ++cand)
    {
      if (cand == 0)
        { // vertices, lines (x=0, x=1, y=0, y=1), interior  [dealii-ext]
          if (p == 1)
            {
              const int X[4] = {0, 1, 0, 1}, Y[4] = {0, 0, 1, 1};
              for (int a = 0; a < 4; ++a)
                ax[a] = X[a], ay[a] = Y[a];
            }
          else
            {
              const int X[9] = {0, 2, 0, 2, 0, 2, 1, 1, 1}, Y[9] = {0, 0, 2, 2, 1, 1, 0, 2, 1};
              for (int a = 0; a < 9; ++a)
                ax[a] = X[a], ay[a] = Y[a];
            }
        }
      else
        for (int a = 0; a < ns; ++a)
          ax[a] = a % n1, ay[a] = a / n1;
      double err = 0;
      for (int qy = 0; qy < n1; ++qy)
        for (int qx = 0; qx < n1; ++qx)
          {
            const int q = qx + n1 * qy;
            err = std::max(err, std::fabs(m.q_weight[q] - lattice_weight_1d(p, qx) * lattice_weight_1d(p, qy)));
            for (int a = 0; a < ns; ++a)
              {
                const double sx = lattice_shape_1d(p, qx, ax[a]), sy = lattice_shape_1d(p, qy, ay[a]);
                const double dx = lattice_dshape_1d(p, qx, ax[a]), dy = lattice_dshape_1d(p, qy, ay[a]);
                err = std::max(err, std::fabs(m.shape_value[q * ns + a] - sx * sy));
                err = std::max(err, std::fabs(m.shape_grad_ref[(q * ns + a) * 2] - dx * sy));
                err = std::max(err, std::fabs(m.shape_grad_ref[(q * ns + a) * 2 + 1] - sx * dy));
              }
          }
      found = err <= 1e-12;
    }
  if (!found)
    return no("shape tables are not tensor FE_Q(p) x QGauss(p+1) in a known local order");

  // ---- 2. identical cells on an integer lattice ----------------------------------------------------
  const double hx = h_of[0], hy = h_of[1];
  const double X0 = m.cell_vertices[0], Y0 = m.cell_vertices[1];
  std::vector<int32_t> cix(m.n_cells), ciy(m.n_cells);
  {
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(+ : bad)
    for (int64_t c = 0; c < m.n_cells; ++c)
      {
        if (std::fabs(h_of[2 * c] - hx) > 1e-10 * hx || std::fabs(h_of[2 * c + 1] - hy) > 1e-10 * hy)
          bad++;
        const double fx = (m.cell_vertices[c * 8] - X0) / hx, fy = (m.cell_vertices[c * 8 + 1] - Y0) / hy;
        const double rx = std::nearbyint(fx), ry = std::nearbyint(fy);
        if (std::fabs(fx - rx) > 1e-6 || std::fabs(fy - ry) > 1e-6 || std::fabs(rx) > 1e8 || std::fabs(ry) > 1e8)
          bad++;
        cix[c] = (int32_t)rx;
        ciy[c] = (int32_t)ry;
      }
    if (bad)
      return no("cells are not identical rectangles on one lattice");
  }
  L.hx = hx;
  L.hy = hy;

  // ---- 3. raw lattice coordinates of the local nodes ---------------------------------------------------
  std::vector<int32_t> nx(N, INT_MIN), ny(N, INT_MIN);
  for (int64_t c = 0; c < m.n_cells; ++c)
    for (int a = 0; a < ns; ++a)
      {
        const int32_t t = cnode[c * ns + a];
        const int32_t x = p * cix[c] + ax[a], y = p * ciy[c] + ay[a];
        if (nx[t] == INT_MIN)
          nx[t] = x, ny[t] = y;
        else if (nx[t] != x || ny[t] != y)
          return no("a node has two lattice positions");
      }

  // ---- 4. periods (in nodes) from slave/master pairs whose raw positions are both known ----------------
  int64_t Px = 0, Py = 0;
  for (int64_t t = 0; t < N; ++t)
    if (master[t] != t && nx[t] != INT_MIN && nx[master[t]] != INT_MIN)
      {
        const int64_t dx = std::llabs((int64_t)nx[master[t]] - nx[t]), dy = std::llabs((int64_t)ny[master[t]] - ny[t]);
        if (dx != 0)
          {
            if (Px != 0 && Px != dx)
              return no("inconsistent x period");
            Px = dx;
          }
        if (dy != 0)
          {
            if (Py != 0 && Py != dy)
              return no("inconsistent y period");
            Py = dy;
          }
      }
  if (Px % p != 0 || Py % p != 0)
    return no("period is not a whole number of cells");

  // ---- 5. owned free nodes: a full rectangle [xa, xb] x [ya, yb] of raw positions ----------------------
  const int64_t own0 = P.own0, own1 = own0 + P.n_owned;
  int64_t xa = INT_MAX, xb = INT_MIN, ya = INT_MAX, yb = INT_MIN, n_free = 0;
  for (int64_t t = own0; t < own1; ++t)
    if (master[t] == t)
      {
        if (nx[t] == INT_MIN)
          return no("an owned free node has no cell of its own");
        xa = std::min<int64_t>(xa, nx[t]);
        xb = std::max<int64_t>(xb, nx[t]);
        ya = std::min<int64_t>(ya, ny[t]);
        yb = std::max<int64_t>(yb, ny[t]);
        ++n_free;
      }
  if (n_free == 0)
    return no("no owned free node");
  const int64_t w = xb - xa + 1, h = yb - ya + 1;
  if (w * h != n_free || w > INT32_MAX / 4 || h > INT32_MAX / 4 || w * h > INT32_MAX / 2)
    return no("owned free nodes do not fill a rectangle");
  if ((Px != 0 && w > Px) || (Py != 0 && h > Py))
    return no("owned rectangle exceeds the period");
  L.w = (int32_t)w;
  L.h = (int32_t)h;
  L.ox = X0 + (double)xa * (hx / p);
  L.oy = Y0 + (double)ya * (hy / p);
  L.Px = Px;
  L.Py = Py;
  L.order.assign((size_t)(w * h), -1);
  for (int64_t t = own0; t < own1; ++t)
    if (master[t] == t)
      {
        int32_t &slot = L.order[(size_t)((ny[t] - ya) * w + (nx[t] - xa))];
        if (slot != -1)
          return no("two owned nodes at one lattice position");
        slot = (int32_t)t;
      }

  // ---- 6. frame: all cells that touch an owned free node; periodic images placed by +- period ----------
  auto fdiv = [](int64_t a, int64_t b) { return a >= 0 ? a / b : -((-a + b - 1) / b); };
  const int64_t kx_min = fdiv(xa + p - 1, p) - 1, kx_max = fdiv(xb, p);
  const int64_t ky_min = fdiv(ya + p - 1, p) - 1, ky_max = fdiv(yb, p);
  const int64_t ncx = kx_max - kx_min + 1, ncy = ky_max - ky_min + 1;
  if (ncx * ncy > INT32_MAX / 16)
    return no("frame too large");
  L.ncx = (int32_t)ncx;
  L.ncy = (int32_t)ncy;
  L.fw = (int32_t)(p * ncx + 1);
  L.fh = (int32_t)(p * ncy + 1);
  L.ex0 = (int32_t)(xa - p * kx_min);
  L.ey0 = (int32_t)(ya - p * ky_min);
  L.lut_global.assign((size_t)L.fw * L.fh, -1);
  std::vector<int8_t> filled((size_t)(ncx * ncy), 0);
  const int64_t Pcx = Px / p, Pcy = Py / p;
  for (int64_t c = 0; c < m.n_cells; ++c)
    for (int my = -1; my <= 1; ++my)
      for (int mx = -1; mx <= 1; ++mx)
        {
          if ((mx != 0 && Pcx == 0) || (my != 0 && Pcy == 0))
            continue;
          const int64_t kx = cix[c] + mx * Pcx - kx_min, ky = ciy[c] + my * Pcy - ky_min;
          if (kx < 0 || kx >= ncx || ky < 0 || ky >= ncy)
            continue;
          filled[(size_t)(ky * ncx + kx)] = 1;
          for (int a = 0; a < ns; ++a)
            {
              const int32_t g = master[cnode[c * ns + a]];
              int32_t &slot = L.lut_global[(size_t)((p * ky + ay[a]) * L.fw + p * kx + ax[a])];
              if (slot != -1 && slot != g)
                return no("two resolved nodes at one frame position");
              slot = g;
            }
        }
  for (int8_t f : filled)
    if (!f)
      return no("a cell touching an owned node was not handed over");
  // the owned rectangle of the frame must be exactly the owned free nodes in lexicographic order
  for (int64_t ly = 0; ly < h; ++ly)
    for (int64_t lx = 0; lx < w; ++lx)
      if (L.lut_global[(size_t)((L.ey0 + ly) * L.fw + L.ex0 + lx)] != L.order[(size_t)(ly * w + lx)])
        return no("frame and owned rectangle disagree");
  // owned slaves (periodic images this rank numbers itself): their frame position, in internal-id order (global id)
  for (int64_t t = own0; t < own1; ++t)
    if (master[t] != t)
      {
        if (nx[t] == INT_MIN)
          return no("an owned slave has no cell of its own");
        const int64_t fx = nx[t] - xa + L.ex0, fy = ny[t] - ya + L.ey0;
        if (fx < 0 || fx >= L.fw || fy < 0 || fy >= L.fh)
          return no("an owned slave lies outside the frame");
        L.slave_fx.push_back((int32_t)fx);
        L.slave_fy.push_back((int32_t)fy);
      }
  L.active = true;
  return true;
}
} // namespace pfc
