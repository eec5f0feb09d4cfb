// pfc_reduced.cuh — kernels of the psi-only reduced Newton solve (PFC_PRECOND_SCHUR_REDUCED, Q1 lattice path).
//
// The serial class of the reference solves  S dpsi = f,  S = B + (D M^-1 K - C) M^-1 K  with CG mass inverses inside every
// operator application (phase_field_crystal_system.cpp:319-369).  On the uniform periodic Q1 lattice K
// and M are circulant and commute, so with dpsi = M^2 y no inverse is left in the operator:
//
//     T y = (B M^2 + D K^2 - C K M) y = f          Krylov vectors of Np doubles instead of 3 Np
//     dpsi = M^2 y,  dchi = u - K M y,  dphi = K^2 y - w,     u = M^-1 G,  w = M^-1 (K u - H),  f = F - C u + D w
//
// (u, w: two mass solves per Newton iteration, skipped while ||(G, H)|| is far below the tolerance: the auxiliary equations
// are linear, so G = H = 0 to rounding after the first reduced solve.)  T y is row 0 of J applied to the three fields
// (M^2 y, -K M y, K^2 y), whose other two rows vanish: lat_pstage_kernel below makes the three fields, lat_j1_kernel<ROW0>
// (pfc_ring_q1.cuh) applies the row.  The same three fields of the solution ARE the Newton update.  Preconditioner:
// Jacobi-Chebyshev on the constant-coefficient S~ = M + dt theta (K - 2 M)(M_L^-1 K)^2 of the ring sweeps — the stiff part of
// T's symbol (dt theta (k - 2 m) k^2) equals S~'s, so GMRES needs about half the iterations of the 3-block solve
// (tests/study_reduced.py).
#pragma once
#include "pfc_ring_q1.cuh"

namespace pfc
{
// (M^2 z, -K M z, K^2 z) on the owned nodes and one node beyond, from the scalar field z (valid three nodes beyond the owned
// rectangle), as a 3-component vector.  A CTA makes TX x TY outputs: z on the tile + 2, (M z, K z) on the tile + 1 in
// shared memory, then the second stencil.  K, M: consistent 9-point stencils (q1::xsum / ysum).
template <int TX, int TY, int NT>
static __global__ void __launch_bounds__(NT) lat_pstage_kernel(LatGeom L, LatOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int IW = TX + 4, IH = TY + 4, AW = TX + 2, AH = TY + 2;
  __shared__ double s_z[IH][IW];
  __shared__ double s_a[AH][AW]; // M z
  __shared__ double s_b[AH][AW]; // K z
  const int tid = threadIdx.x;
  const int X0 = L.ex0 - 1 + blockIdx.x * TX, Y0 = L.ey0 - 1 + blockIdx.y * TY; // padded position of output (0, 0) of the tile
  const q1::Cst C(L);
  for (int idx = tid; idx < IW * IH; idx += NT)
    {
      const int X = X0 - 2 + idx % IW, Y = Y0 - 2 + idx / IW;
      s_z[idx / IW][idx % IW] = (X >= 0 && X < L.fw && Y >= 0 && Y < L.fh) ? O.s0[lat_id(L, X, Y)] : 0.0;
    }
  __syncthreads();
  for (int idx = tid; idx < AW * AH; idx += NT)
    {
      const int i = idx % AW, j = idx / AW; // centre at s_z[j + 1][i + 1]
      double wk[3], wm[3], Kv, Mv;
#pragma unroll
      for (int r = 0; r < 3; ++r)
        q1::xsum(&s_z[j + r][i], wk[r], wm[r]);
      q1::ysum<true>(C, 2, wk, wm, Kv, Mv);
      s_a[j][i] = Mv;
      s_b[j][i] = Kv;
    }
  __syncthreads();
  const int64_t Np = L.Np;
  for (int idx = tid; idx < TX * TY; idx += NT)
    {
      const int i = idx % TX, j = idx / TX; // centre at s_a[j + 1][i + 1]
      const int X = X0 + i, Y = Y0 + j;
      if (X > L.ex0 + L.w || Y > L.ey0 + L.h || X >= L.fw || Y >= L.fh)
        continue;
      double wk[3], wm[3], Ka, Ma, Kb, Mnone;
#pragma unroll
      for (int r = 0; r < 3; ++r)
        q1::xsum(&s_a[j + r][i], wk[r], wm[r]);
      q1::ysum<true>(C, 2, wk, wm, Ka, Ma);
#pragma unroll
      for (int r = 0; r < 3; ++r)
        q1::xsum(&s_b[j + r][i], wk[r], wm[r]);
      q1::ysum<false>(C, 2, wk, wm, Kb, Mnone);
      const int32_t u = lat_id(L, X, Y);
      O.out3[vaddr(u, 0, Np)] = Ma;
      O.out3[vaddr(u, 1, Np)] = -Ka;
      O.out3[vaddr(u, 2, Np)] = Kb;
    }
}

// first Chebyshev step on a whole scalar field [owned | ring]: d = x = c0 dinv r  (no stencil: x_0 = 0)
static __global__ void __launch_bounds__(256) cheb0_kernel(double *__restrict__ d, double *__restrict__ x, const double *__restrict__ r,
                                                           const double *__restrict__ dinv, const double *__restrict__ cheb,
                                                           int64_t n, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  const double c0 = cheb[0];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    {
      const double v = c0 * dinv[i] * r[i];
      d[i] = v;
      x[i] = v;
    }
}

static __global__ void fill_kernel(double *x, double v, int64_t n)
{
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    x[i] = v;
}

// y = a x + b y elementwise (b == 0: y is not read); optional control gate
static __global__ void __launch_bounds__(256) axpby_kernel(double *__restrict__ y, double a, const double *__restrict__ x, double b,
                                                           int64_t n, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = b == 0.0 ? a * x[i] : fma(a, x[i], b * y[i]);
}
} // namespace pfc
