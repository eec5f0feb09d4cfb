// pfc_matfree.cuh — K4': matrix-free application y = J(Psi_n) x on Q1 / Q2 quads (north_star item 2; the
// reference has no such path).  Same integrand as the assembly (mpi.cpp:421-444), applied to
// a vector instead of being stored:
//   y_psi_i = sum_q { N_i [ bm X0 + bv . grad X0 + cm X1 - 2 dt th X2 ] + dt th grad N_i . grad X2 } JxW
//   y_chi_i = sum_q { grad N_i . grad X0 + N_i X1 } JxW
//   y_phi_i = sum_q { grad N_i . grad X1 + N_i X2 } JxW
// with X_c = sum_j N_j x_c(m(j)) evaluated at the quadrature points (columns resolved through the periodic
// constraints, rows scattered to m(i)), bm = 1 - dt th (6 psi chi + 6 |grad psi|^2), bv = -12 dt th psi grad psi,
// cm = -dt th ((1 + eps) + 3 psi^2).  One warp handles 32/NS cells; lane = quadrature point in phase 1, test
// function in phase 2; colour-ordered read-modify-write scatter (no atomics, deterministic).
// Algorithmic bytes: x 24 N + y 24 N + (psi, chi) 16 N + 4 ints per cell node (gather, col, row, spos).
#pragma once
#include "pfc_kernels.cuh"

namespace pfc
{
struct MfArgs
{
  const int32_t *cell_gather, *cell_col, *cell_row, *cell_spos;
  const double *cell_h;
  const double *psi_n; // state vector (psi at c = 0, chi at c = 1)
  const double *tab;
  int64_t Np;
  double dtth, eps;
};

constexpr int kMfCoef = 12;

template <int NS>
static __global__ void __launch_bounds__(kAsmThreads) mf_apply_kernel(MfArgs A, const double *__restrict__ x,
                                                                       double *__restrict__ y, int64_t cell_begin,
                                                                       int64_t cell_end, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int CPW = 32 / NS;
  constexpr int WPB = kAsmThreads / 32;
  __shared__ double s_sh[NS * NS], s_dsh[NS * NS * 2], s_qw[NS];
  __shared__ double s_u[WPB][CPW][5][NS];
  __shared__ double s_coef[WPB][CPW][NS][kMfCoef];

  for (int i = threadIdx.x; i < NS * NS; i += kAsmThreads)
    s_sh[i] = A.tab[i];
  for (int i = threadIdx.x; i < NS * NS * 2; i += kAsmThreads)
    s_dsh[i] = A.tab[NS * NS + i];
  for (int i = threadIdx.x; i < NS; i += kAsmThreads)
    s_qw[i] = A.tab[NS * NS * 3 + i];
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cw = lane / NS, a = lane % NS;
  const int64_t cell = cell_begin + ((int64_t)blockIdx.x * WPB + warp) * CPW + cw;
  const bool valid = cw < CPW && cell < cell_end;
  double hx = 1, hy = 1;
  int32_t uraw = 0;
  if (valid)
    {
      hx = A.cell_h[cell * 2];
      hy = A.cell_h[cell * 2 + 1];
      uraw = A.cell_gather[cell * NS + a];
      const int32_t ucol = A.cell_col[cell * NS + a];
#pragma unroll
      for (int c = 0; c < 3; ++c)
        s_u[warp][cw][c][a] = x[vaddr(ucol, c, A.Np)];
      s_u[warp][cw][3][a] = A.psi_n[vaddr(uraw, 0, A.Np)];
      s_u[warp][cw][4][a] = A.psi_n[vaddr(uraw, 1, A.Np)];
    }
  const double ihx = 1.0 / hx, ihy = 1.0 / hy;
  __syncwarp();
  if (valid)
    {
      const int q = a;
      double v[5] = {0, 0, 0, 0, 0}, gx[5] = {0, 0, 0, 0, 0}, gy[5] = {0, 0, 0, 0, 0};
#pragma unroll
      for (int b = 0; b < NS; ++b)
        {
          const double n = s_sh[q * NS + b];
          const double dx = s_dsh[(q * NS + b) * 2] * ihx, dy = s_dsh[(q * NS + b) * 2 + 1] * ihy;
#pragma unroll
          for (int f = 0; f < 5; ++f)
            {
              const double uv = s_u[warp][cw][f][b];
              v[f] += uv * n;
              gx[f] += uv * dx;
              gy[f] += uv * dy;
            }
        }
      const double JxW = s_qw[q] * hx * hy;
      const double psi = v[3], chi = v[4];
      const double bm = (1.0 - A.dtth * (6 * psi * chi + 6 * (gx[3] * gx[3] + gy[3] * gy[3]))) * JxW;
      const double bx = -12.0 * A.dtth * psi * gx[3] * JxW, by = -12.0 * A.dtth * psi * gy[3] * JxW;
      const double cm = -A.dtth * ((1 + A.eps) + 3 * psi * psi) * JxW;
      double *cf = s_coef[warp][cw][q];
      cf[0] = bm * v[0] + bx * gx[0] + by * gy[0] + cm * v[1] - 2.0 * A.dtth * v[2] * JxW;
      cf[1] = A.dtth * gx[2] * JxW;
      cf[2] = A.dtth * gy[2] * JxW;
      cf[3] = v[1] * JxW;
      cf[4] = gx[0] * JxW;
      cf[5] = gy[0] * JxW;
      cf[6] = v[2] * JxW;
      cf[7] = gx[1] * JxW;
      cf[8] = gy[1] * JxW;
      cf[9] = bm;
      cf[10] = bx;
      cf[11] = by;
    }
  __syncwarp();
  if (!valid)
    return;
  const int i = a;
  double y0 = 0, y1 = 0, y2 = 0;
#pragma unroll
  for (int q = 0; q < NS; ++q)
    {
      const double ni = s_sh[q * NS + i];
      const double gix = s_dsh[(q * NS + i) * 2] * ihx, giy = s_dsh[(q * NS + i) * 2 + 1] * ihy;
      const double *cf = s_coef[warp][cw][q];
      y0 += ni * cf[0] + gix * cf[1] + giy * cf[2];
      y1 += ni * cf[3] + gix * cf[4] + giy * cf[5];
      y2 += ni * cf[6] + gix * cf[7] + giy * cf[8];
    }
  const int32_t row = A.cell_row[cell * NS + i];
  if (row >= 0)
    {
      y[row] += y0;
      y[A.Np + row] += y1;
      y[2 * A.Np + row] += y2;
    }
  if (A.cell_spos[cell * NS + i] >= 0)
    { // constrained local row: |local(i,i)| times the slave's own entry ([dealii-ext], as the assembly stores it)
      double bii = 0, mii = 0;
#pragma unroll
      for (int q = 0; q < NS; ++q)
        {
          const double ni = s_sh[q * NS + i];
          const double gix = s_dsh[(q * NS + i) * 2] * ihx, giy = s_dsh[(q * NS + i) * 2 + 1] * ihy;
          const double *cf = s_coef[warp][cw][q];
          bii += ni * (cf[9] * ni + cf[10] * gix + cf[11] * giy);
          mii += ni * ni * s_qw[q] * hx * hy;
        }
      y[uraw] += fabs(bii) * x[uraw];
      y[A.Np + uraw] += fabs(mii) * x[A.Np + uraw];
      y[2 * A.Np + uraw] += fabs(mii) * x[2 * A.Np + uraw];
    }
}
} // namespace pfc
