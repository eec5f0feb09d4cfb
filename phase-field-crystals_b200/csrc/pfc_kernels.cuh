// pfc_kernels.cuh — hand-written fp64 CUDA kernels (sm_100a) of the PFC Newton-Krylov hot path.
//
// All kernels are HBM-bound ("Kernels and rooflines"); tensor cores are not used.
// Vector layout (length Lv): [psi owned Np | chi owned Np | phi owned Np | ghosts interleaved
// [g][c]], i.e. component c of internal node u lives at  u < Np ? c*Np + u : 3*u + c.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace pfc
{
// device-resident control block of one GMRES solve (no host round trip per iteration)
struct Ctl
{
  int state;            // 0 iterate, 1 success, 2 failure (SolverControl::check)
  int reorth;           // sticky re-orthogonalisation flag (Kelley test tripped)
  int dim;              // Krylov dimension built in the current restart cycle
  int pass1;            // second Gram-Schmidt pass ran in this iteration
  long long iters;      // accumulated_iterations
  long long max_iters;
  double tol;
  double rho;           // current residual estimate
  double inv_norm;      // 1 / ||w|| of the newest basis vector
  double ww0;           // ||w||^2 before orthogonalisation
  long long kelley_trips;
};

__device__ __forceinline__ int64_t vaddr(int32_t u, int c, int64_t Np)
{
  return u < Np ? (int64_t)c * Np + u : (int64_t)3 * u + c;
}

// ------------------------------------------------------------------------------------------------
// K1-K3: quadrature-point assembly (mpi.cpp:350-462) of the iterate-dependent blocks B, C and
// of -R, or (CONST_PASS) of the constant blocks M (mass) and K (stiffness).
// One warp handles 32/NS cells; lane (cell, a) is quadrature point q = a in phase 1 and test
// function i = a in phase 2 (n_q == n_s for QGauss(p+1) x FE_Q(p)).  Shape/gradient tables are
// staged in shared memory; the scatter is a plain read-modify-write, conflict free inside one
// colour (no atomics, deterministic).
// ------------------------------------------------------------------------------------------------
struct AsmArgs
{
  const int32_t *cell_gather, *cell_row, *cell_pos, *cell_spos;
  const double *cell_h;
  const double *psi_n, *psi_n1;
  double *B, *C, *M, *K, *rhs;
  const double *tab; // shape[nq][ns], dshape[nq][ns][2], qw[nq]
  int64_t Np;
  double dt, theta, eps;
  int rhs_only; // matrix-free operator: only -R is needed, B and C are not stored
};

constexpr int kAsmThreads = 128;
constexpr int kNCoef = 13;

template <int NS, bool CONST_PASS>
static __global__ void __launch_bounds__(kAsmThreads) assemble_kernel(AsmArgs A, int64_t cell_begin, int64_t cell_end)
{
  constexpr int CPW = 32 / NS;               // cells per warp
  constexpr int WPB = kAsmThreads / 32;      // warps per block
  __shared__ double s_sh[NS * NS], s_dsh[NS * NS * 2], s_qw[NS];
  __shared__ double s_u[WPB][CPW][6][NS];
  __shared__ double s_coef[WPB][CPW][NS][kNCoef];

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
  if (valid)
    {
      hx = A.cell_h[cell * 2];
      hy = A.cell_h[cell * 2 + 1];
    }
  const double ihx = 1.0 / hx, ihy = 1.0 / hy;

  if (!CONST_PASS)
    {
      if (valid)
        {
          const int32_t u = A.cell_gather[cell * NS + a];
#pragma unroll
          for (int c = 0; c < 3; ++c)
            {
              const int64_t ad = vaddr(u, c, A.Np);
              s_u[warp][cw][c][a] = A.psi_n[ad];
              s_u[warp][cw][3 + c][a] = A.psi_n1[ad];
            }
        }
      __syncwarp();
      if (valid)
        {
          // phase 1: fields at quadrature point q = a (mpi.cpp:360-371)
          const int q = a;
          double v[6] = {0, 0, 0, 0, 0, 0}, gx[6] = {0, 0, 0, 0, 0, 0}, gy[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
          for (int b = 0; b < NS; ++b)
            {
              const double n = s_sh[q * NS + b];
              const double dx = s_dsh[(q * NS + b) * 2] * ihx, dy = s_dsh[(q * NS + b) * 2 + 1] * ihy;
#pragma unroll
              for (int f = 0; f < 6; ++f)
                {
                  const double uv = s_u[warp][cw][f][b];
                  v[f] += uv * n;
                  gx[f] += uv * dx;
                  gy[f] += uv * dy;
                }
            }
          const double JxW = s_qw[q] * hx * hy;
          const double psi = v[0], chi = v[1], phi = v[2], psi1 = v[3], chi1 = v[4], phi1 = v[5];
          const double g2 = gx[0] * gx[0] + gy[0] * gy[0], g21 = gx[3] * gx[3] + gy[3] * gy[3];
          const double dtth = A.dt * A.theta, dt1 = A.dt * (1.0 - A.theta);
          double *cf = s_coef[warp][cw][q];
          // residual integrand, mpi.cpp:388-417
          cf[0] = (-psi + psi1 + dtth * (2 * phi + (1 + A.eps) * chi + 3 * psi * psi * chi + 6 * psi * g2) +
                   dt1 * (2 * phi1 + (1 + A.eps) * chi1 + 3 * psi1 * psi1 * chi1 + 6 * psi1 * g21)) * JxW;
          cf[1] = A.dt * (A.theta * gx[2] + (1 - A.theta) * gx[5]) * JxW;
          cf[2] = A.dt * (A.theta * gy[2] + (1 - A.theta) * gy[5]) * JxW;
          cf[3] = -chi * JxW;
          cf[4] = gx[0] * JxW;
          cf[5] = gy[0] * JxW;
          cf[6] = -phi * JxW;
          cf[7] = gx[1] * JxW;
          cf[8] = gy[1] * JxW;
          // Jacobian coefficients of B and C, mpi.cpp:421-433
          cf[9] = (1.0 - dtth * (6 * psi * chi + 6 * g2)) * JxW;
          cf[10] = -12.0 * dtth * psi * gx[0] * JxW;
          cf[11] = -12.0 * dtth * psi * gy[0] * JxW;
          cf[12] = -dtth * ((1 + A.eps) + 3 * psi * psi) * JxW;
        }
      __syncwarp();
    }

  if (!valid)
    return;

  // phase 2: row i = a of the local blocks
  const int i = a;
  double m0[NS], m1[NS];
#pragma unroll
  for (int j = 0; j < NS; ++j)
    m0[j] = m1[j] = 0;
  double r0 = 0, r1 = 0, r2 = 0;
#pragma unroll
  for (int q = 0; q < NS; ++q)
    {
      const double ni = s_sh[q * NS + i];
      const double gix = s_dsh[(q * NS + i) * 2] * ihx, giy = s_dsh[(q * NS + i) * 2 + 1] * ihy;
      if (CONST_PASS)
        {
          const double JxW = s_qw[q] * hx * hy;
#pragma unroll
          for (int j = 0; j < NS; ++j)
            {
              const double nj = s_sh[q * NS + j];
              const double gjx = s_dsh[(q * NS + j) * 2] * ihx, gjy = s_dsh[(q * NS + j) * 2 + 1] * ihy;
              m0[j] += ni * nj * JxW;                    // M: eta_chi_i eta_chi_j, mpi.cpp:437,441
              m1[j] += (gix * gjx + giy * gjy) * JxW;    // K: grad eta_i . grad eta_j, mpi.cpp:435,439,443
            }
        }
      else
        {
          const double *cf = s_coef[warp][cw][q];
          r0 += ni * cf[0] - (gix * cf[1] + giy * cf[2]);
          r1 += ni * cf[3] - (gix * cf[4] + giy * cf[5]);
          r2 += ni * cf[6] - (gix * cf[7] + giy * cf[8]);
          const double bm = cf[9], bx = cf[10], by = cf[11], cm = cf[12];
          if (!A.rhs_only)
#pragma unroll
          for (int j = 0; j < NS; ++j)
            {
              const double nj = s_sh[q * NS + j];
              const double gjx = s_dsh[(q * NS + j) * 2] * ihx, gjy = s_dsh[(q * NS + j) * 2 + 1] * ihy;
              m0[j] += ni * (bm * nj + bx * gjx + by * gjy); // B
              m1[j] += ni * cm * nj;                         // C
            }
        }
    }

  // scatter through the periodic constraints (distribute_local_to_global, mpi.cpp:454-458)
  const int32_t row = A.cell_row[cell * NS + i];
  double *d0 = CONST_PASS ? A.M : A.B, *d1 = CONST_PASS ? A.K : A.C;
  if (row >= 0)
    {
      if (!CONST_PASS)
        {
          A.rhs[row] += r0;
          A.rhs[A.Np + row] += r1;
          A.rhs[2 * A.Np + row] += r2;
        }
      if (CONST_PASS || !A.rhs_only)
#pragma unroll
      for (int j = 0; j < NS; ++j)
        {
          const int32_t pos = A.cell_pos[(cell * NS + i) * NS + j];
          d0[pos] += m0[j];
          d1[pos] += m1[j];
        }
    }
  // constrained local row: |local(i,i)| on the slave's own diagonal ([dealii-ext])
  const int32_t sp = A.cell_spos[cell * NS + i];
  if (sp >= 0 && (CONST_PASS || !A.rhs_only))
    d0[sp] += fabs(m0[i]);
}

// ------------------------------------------------------------------------------------------------
// K4: y = J x on the shared-pattern block layout.  J = [[B, C, D], [K, M, 0], [0, K, M]] with
// D = dt*theta*(K - 2M) (mpi.cpp:421-443; tex:399-466).  SELL-32: one warp per slice, one lane per
// row, entries column-major inside the slice, so every load of col/B/C/M/K is a full coalesced line.
// Bytes per scalar entry: 4 + 4*8 = 36 (the monolithic CSR of the reference moves 7*12 = 84).
// ------------------------------------------------------------------------------------------------
struct SpmvArgs
{
  const int64_t *slice_ptr;
  const int32_t *col;
  const double *B, *C, *M, *K;
  int64_t Np, n_slices;
  double dtth;
};

constexpr int kSpmvThreads = 128;

static __global__ void __launch_bounds__(kSpmvThreads) spmv_sell_kernel(SpmvArgs S, const double *__restrict__ x,
                                                                  double *__restrict__ y, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  const int64_t w = ((int64_t)blockIdx.x * kSpmvThreads + threadIdx.x) >> 5;
  if (w >= S.n_slices)
    return;
  const int lane = threadIdx.x & 31;
  const int64_t base = S.slice_ptr[w];
  const int width = (int)((S.slice_ptr[w + 1] - base) >> 5);
  const int64_t Np = S.Np;
  double y0 = 0, y1 = 0, y2 = 0;
  const int64_t p0 = base + lane;
#pragma unroll 4
  for (int k = 0; k < width; ++k)
    {
      const int64_t p = p0 + (int64_t)k * 32;
      int32_t c = __ldg(S.col + p);
      const double vb = __ldg(S.B + p), vm = __ldg(S.M + p);
      if (c < 0)
        { // constrained row: diagonal only (B for psi, M for chi and phi)
          c = ~c;
          y0 += vb * x[c];
          y1 += vm * x[Np + c];
          y2 += vm * x[2 * Np + c];
        }
      else
        {
          const double vc = __ldg(S.C + p), vk = __ldg(S.K + p);
          const double x0 = x[vaddr(c, 0, Np)], x1 = x[vaddr(c, 1, Np)], x2 = x[vaddr(c, 2, Np)];
          y0 += vb * x0 + vc * x1 + S.dtth * (vk - 2.0 * vm) * x2;
          y1 += vk * x0 + vm * x1;
          y2 += vk * x1 + vm * x2;
        }
    }
  const int64_t r = w * 32 + lane;
  y[r] = y0;
  y[Np + r] = y1;
  y[2 * Np + r] = y2;
}
} // namespace pfc
