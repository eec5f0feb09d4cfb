// pfc_lattice_smooth.cuh — EXPERIMENTAL (off unless PFC_SMOOTHER=1, single rank): one Chebyshev sweep of the Schur solve
// with a constant-coefficient stand-in of the psi Schur complement, in ONE tiled kernel (DESIGN.md §9 item 2;
// iteration counts studied in tests/study_precond.py).
//   S~ x = bbar M x - cbar M t1 + dt theta (K - 2 M) t2,   t1 = M_L^-1 K x,  t2 = M_L^-1 K t1
// (S_L = B - C M_L^-1 K + D M_L^-1 K M_L^-1 K with B ~ bbar M, C ~ cbar M, D = dt theta (K - 2 M)), then
//   d = c0 d + c1 dinv (rhs - S~ x),  xout = x + d.
// All operators are tensor products of the assembled 1-D rows (Row1<P>), so a CTA stages x with a halo of 3p nodes in shared
// memory and applies the three stencil stages on shrinking tiles.  Not yet run on a GPU: pfc_selfcheck_smoother compares it
// with the composition of the K-stencil and mass-stencil kernels and is the first thing to run.
#ifndef PFC_LATTICE_SMOOTH_CUH
#define PFC_LATTICE_SMOOTH_CUH
#include "pfc_lattice.cuh"

namespace pfc
{
constexpr int kSmoothThreads = 256;

template <int P, int TX, int TY>
constexpr size_t lat_smooth_smem()
{
  return sizeof(double) * ((size_t)(TX + 6 * P) * (TY + 6 * P) + (size_t)(TX + 4 * P) * (TY + 4 * P) +
                           (size_t)(TX + 2 * P) * (TY + 2 * P) + 8 * (2 * P + 1));
}

// single rank: the owned free nodes are the whole periodic lattice, w and h are the periods
template <int P, int TX, int TY>
static __global__ void __launch_bounds__(kSmoothThreads) lat_smooth_kernel(LatGeom L, SmoothOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int NW = 2 * P + 1;
  constexpr int VX = TX + 6 * P, VY = TY + 6 * P; // x tile, halo 3p
  constexpr int AX = TX + 4 * P, AY = TY + 4 * P; // t1 tile, halo 2p
  constexpr int BX = TX + 2 * P, BY = TY + 2 * P; // t2 tile, halo p
  extern __shared__ double sm[];
  double *sv = sm, *s1 = sv + VX * VY, *s2 = s1 + AX * AY, *rows = s2 + BX * BY;
  // rows: [kx | mx | ky | my][type 0 = vertex, 1 = mid][NW], scaled to the physical 1-D operators
  double *kx = rows, *mx = rows + 2 * NW, *ky = rows + 4 * NW, *my = rows + 6 * NW;
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY; // first output node of the tile (owned coordinates)
  if (tid < 2 * NW)
    {
      const int type = tid / NW, dd = tid % NW - P;
      const double k1 = type ? Row1<P>::K(1, dd) : Row1<P>::K(0, dd), m1 = type ? Row1<P>::M(1, dd) : Row1<P>::M(0, dd);
      kx[tid] = k1 * L.ihx;
      mx[tid] = m1 * L.hx;
      ky[tid] = k1 * L.ihy;
      my[tid] = m1 * L.hy;
    }
  for (int idx = tid; idx < VX * VY; idx += kSmoothThreads)
    {
      int gx = (x0 - 3 * P + idx % VX) % L.w, gy = (y0 - 3 * P + idx / VX) % L.h;
      gx += gx < 0 ? L.w : 0;
      gy += gy < 0 ? L.h : 0;
      sv[idx] = O.x[(int64_t)gy * L.w + gx];
    }
  __syncthreads();

  // node type of tile-local position i of a tile whose first node has owned coordinate c0 (frame = owned + ex0 / ey0)
  auto type_of = [](int frame_coord) { return P == 2 ? (frame_coord & 1) : 0; };
  auto lumped = [](int type) { return P == 1 ? 1.0 : (type ? 2.0 / 3.0 : 1.0 / 3.0); };
  // (Ax (x) By + Cx (x) Dy) applied at src[(j0 + dy) * ldx + i0 + dx]
  auto stencil2 = [&](const double *src, int ldx, int i0, int j0, const double *ax, const double *by, const double *cx,
                      const double *dy_) {
    double acc = 0;
#pragma unroll
    for (int dy = 0; dy < NW; ++dy)
      {
        const double *row = src + (j0 + dy - P) * ldx + i0 - P;
        double ra = 0, rc = 0;
#pragma unroll
        for (int dx = 0; dx < NW; ++dx)
          {
            ra = fma(ax[dx], row[dx], ra);
            rc = fma(cx[dx], row[dx], rc);
          }
        acc = fma(by[dy], ra, acc);
        acc = fma(dy_[dy], rc, acc);
      }
    return acc;
  };
  const double ivol = L.ihx * L.ihy;

  // t1 = M_L^-1 K x on the 2p-halo tile
  for (int idx = tid; idx < AX * AY; idx += kSmoothThreads)
    {
      const int i = idx % AX, j = idx / AX;
      const int tx = type_of(L.ex0 + x0 - 2 * P + i), ty = type_of(L.ey0 + y0 - 2 * P + j);
      const double kv = stencil2(sv, VX, i + P, j + P, kx + tx * NW, my + ty * NW, mx + tx * NW, ky + ty * NW);
      s1[idx] = kv * ivol / (lumped(tx) * lumped(ty));
    }
  __syncthreads();
  // t2 = M_L^-1 K t1 on the p-halo tile
  for (int idx = tid; idx < BX * BY; idx += kSmoothThreads)
    {
      const int i = idx % BX, j = idx / BX;
      const int tx = type_of(L.ex0 + x0 - P + i), ty = type_of(L.ey0 + y0 - P + j);
      const double kv = stencil2(s1, AX, i + P, j + P, kx + tx * NW, my + ty * NW, mx + tx * NW, ky + ty * NW);
      s2[idx] = kv * ivol / (lumped(tx) * lumped(ty));
    }
  __syncthreads();
  // S~ x and the Chebyshev update on the output nodes
  const double c0 = O.cheb[0], c1 = O.cheb[1];
  for (int idx = tid; idx < TX * TY; idx += kSmoothThreads)
    {
      const int i = idx % TX, j = idx / TX;
      const int lx = x0 + i, ly = y0 + j;
      if (lx >= L.w || ly >= L.h)
        continue;
      const int tx = type_of(L.ex0 + lx), ty = type_of(L.ey0 + ly);
      const double *kxr = kx + tx * NW, *mxr = mx + tx * NW, *kyr = ky + ty * NW, *myr = my + ty * NW;
      // M x, M t1: (Mx (x) My) only -> second pair zero via the same helper with cx = mx, dy = 0 rows is wasteful; direct:
      double m_x = 0, m_t1 = 0, m_t2 = 0;
#pragma unroll
      for (int dy = 0; dy < NW; ++dy)
        {
          double rx = 0, r1 = 0, r2 = 0;
#pragma unroll
          for (int dx = 0; dx < NW; ++dx)
            {
              rx = fma(mxr[dx], sv[(j + 3 * P + dy - P) * VX + i + 3 * P + dx - P], rx);
              r1 = fma(mxr[dx], s1[(j + 2 * P + dy - P) * AX + i + 2 * P + dx - P], r1);
              r2 = fma(mxr[dx], s2[(j + P + dy - P) * BX + i + P + dx - P], r2);
            }
          m_x = fma(myr[dy], rx, m_x);
          m_t1 = fma(myr[dy], r1, m_t1);
          m_t2 = fma(myr[dy], r2, m_t2);
        }
      const double k_t2 = stencil2(s2, BX, i + P, j + P, kxr, myr, mxr, kyr);
      const double Sx = O.bbar * m_x - O.cbar * m_t1 + O.dtth * (k_t2 - 2.0 * m_t2);
      const int64_t r = (int64_t)ly * L.w + lx;
      const double dn = c0 * O.d[r] + c1 * O.dinv[r] * (O.rhs[r] - Sx);
      O.d[r] = dn;
      O.xout[r] = sv[(j + 3 * P) * VX + i + 3 * P] + dn;
    }
}
} // namespace pfc
#endif
