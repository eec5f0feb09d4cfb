// pfc_ring.cuh — streaming stencil kernels of the ring path: the node-stencil parts of the block-Schur preconditioner
// (pfc_precond.cuh) fused into three kernels that run on the padded rectangle of the ring overlay with shrinking validity
// and NO halo exchange in between (pfc_plan.h, Ring):
//   t kernel:          m1 = M_L^-1 r1,  t = M_L^-1 (r2 - K m1)
//   Chebyshev sweep:   d = c0 d + c1 dinv (rhs - S~ x),  x += d,   S~ = bbar M + dt theta (K - 2 M) (M_L^-1 K)^2
//   back substitution: chi = M_L^-1 (r1 - K psi),  phi = M_L^-1 (r2 - K chi)
// S~ is the psi Schur complement S_L = B - C M_L^-1 K + D M_L^-1 K M_L^-1 K with its iterate-dependent blocks replaced
// by B ~ bbar M, C ~ 0 (studied in tests/study_precond.py: the GMRES iteration counts stay within one or two of the
// exact S_L).  It is used INSIDE the smoother only; the Schur right-hand side and the outer operator stay exact, and
// the stopping test of GMRES is the true residual, so the answer does not change.
//
// All operators are tensor products of the assembled 1-D rows of FE_Q(p) (Row1<P>): K = Kx (x) My + Mx (x) Ky,
// M = Mx (x) My.  A CTA owns a block of columns and marches down a strip of rows; a thread owns one column, keeps
// the x-direction sums of the last 2p+1 rows of every stage in registers and combines them in y.  Neighbouring
// columns meet through one shared-memory row per stage (double buffered: one __syncthreads per row for all stages,
// every stage lagging the one before by p + 1 rows).  Each stage loses p columns at the block edges, so a block of NT
// threads produces NT - 2 p stages columns.  Bytes per node: sweep x, rhs, dinv, d in + d, x out = 48.
#pragma once
#include "pfc_lattice.cuh"

namespace pfc
{
template <int P>
struct RingTab
{
  static constexpr int NW = 2 * P + 1;
  double v[4][2][NW]; // [kx | mx | ky | my][vertex | mid][offset + P], physical 1-D rows
};

template <int P>
__device__ __forceinline__ void ring_fill_tab(RingTab<P> &T, const LatGeom &L, int tid)
{
  constexpr int NW = 2 * P + 1;
  if (tid < 2 * NW)
    {
      const int type = tid / NW, dd = tid % NW - P;
      const double k1 = type ? Row1<P>::K(1, dd) : Row1<P>::K(0, dd), m1 = type ? Row1<P>::M(1, dd) : Row1<P>::M(0, dd);
      T.v[0][type][dd + P] = k1 * L.ihx;
      T.v[1][type][dd + P] = m1 * L.hx;
      T.v[2][type][dd + P] = k1 * L.ihy;
      T.v[3][type][dd + P] = m1 * L.hy;
    }
}

// 1 / lumped mass of the node at padded position (X, Y): 1 / (hx hy wx wy), w = 1 (Q1), 1/3 vertex, 2/3 mid (Q2)
template <int P>
__device__ __forceinline__ double ring_iml(const LatGeom &L, int X, int Y)
{
  if (P == 1)
    return L.ihx * L.ihy;
  const double wx = (X & 1) ? 1.5 : 3.0, wy = (Y & 1) ? 1.5 : 3.0;
  return L.ihx * L.ihy * wx * wy;
}

__device__ __forceinline__ int64_t ring_vaddr(const LatGeom &L, int X, int Y, int c)
{
  const int64_t u = lat_id(L, X, Y);
  return u < L.Np ? (int64_t)c * L.Np + u : 3 * u + c;
}

// x-direction sums of one exchanged row: (Kx v, Mx v) at this thread's column
template <int P>
__device__ __forceinline__ void ring_xsum(const double *row, const double (&kx)[2 * P + 1], const double (&mx)[2 * P + 1],
                                          double &k, double &m)
{
  k = 0, m = 0;
#pragma unroll
  for (int dd = 0; dd <= 2 * P; ++dd)
    {
      const double v = row[dd];
      k = fma(kx[dd], v, k);
      m = fma(mx[dd], v, m);
    }
}
template <int N>
__device__ __forceinline__ void ring_push(double (&w)[N], double v)
{
#pragma unroll
  for (int i = 0; i + 1 < N; ++i)
    w[i] = w[i + 1];
  w[N - 1] = v;
}
// (K v)(row) = sum my rK + ky rM and (M v)(row) = sum my rM from the windows of x sums centred on `row`
template <int P>
__device__ __forceinline__ void ring_ysum(const RingTab<P> &T, int row, const double (&rK)[2 * P + 1],
                                          const double (&rM)[2 * P + 1], double &Kv, double &Mv)
{
  const int ty = P == 2 ? (row & 1) : 0;
  Kv = 0, Mv = 0;
#pragma unroll
  for (int dd = 0; dd <= 2 * P; ++dd)
    {
      const double my = T.v[3][ty][dd], ky = T.v[2][ty][dd];
      Kv = fma(my, rK[dd], Kv);
      Kv = fma(ky, rM[dd], Kv);
      Mv = fma(my, rM[dd], Mv);
    }
}

// ---- t kernel: m1 = M_L^-1 r1, t = M_L^-1 (r2 - K m1) on the window of margin lo -------------------------------
template <int P, int NT, int PF>
static __global__ void __launch_bounds__(NT) ring_t_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int NW = 2 * P + 1, HX = P, OUTW = NT - 2 * HX;
  __shared__ double s_ex[2][NT + 2 * P];
  __shared__ RingTab<P> T;
  const int tid = threadIdx.x;
  ring_fill_tab<P>(T, L, tid);
  for (int i = tid; i < 2 * (NT + 2 * P); i += NT)
    (&s_ex[0][0])[i] = 0.0;
  const int X = O.lo + blockIdx.x * OUTW - HX + tid;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  const bool colok = X >= 0 && X < L.fw, outcol = tid >= HX && tid < NT - HX && X < L.fw - O.lo;
  __syncthreads();
  const int tx = P == 2 ? (X & 1) : 0;
  double kx[NW], mx[NW], rK[NW], rM[NW];
#pragma unroll
  for (int dd = 0; dd < NW; ++dd)
    kx[dd] = T.v[0][tx][dd], mx[dd] = T.v[1][tx][dd], rK[dd] = 0, rM[dd] = 0;
  const int n0 = Y0 - HX, n1 = Y1 - 1 + HX;
  // operands are fetched PF rows ahead into a register ring, so that PF rows of loads are in flight per thread
  struct Pre
  {
    double r1n, o_r1, o_r2;
    bool wr;
  };
  auto fetch = [&](int n, Pre &q) {
    q.r1n = (colok && n <= n1 && n < L.fh) ? O.in[ring_vaddr(L, X, n, 1)] : 0.0;
    const int m = n - P;
    q.wr = outcol && m >= Y0 && m < Y1;
    q.o_r1 = q.o_r2 = 0.0;
    if (q.wr)
      {
        q.o_r1 = O.in[ring_vaddr(L, X, m, 1)];
        q.o_r2 = O.in[ring_vaddr(L, X, m, 2)];
      }
  };
  Pre pre[PF];
#pragma unroll
  for (int j = 0; j < PF; ++j)
    fetch(n0 + j, pre[j]);
  for (int nb = n0; nb <= n1; nb += PF)
    {
#pragma unroll
      for (int j = 0; j < PF; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          const Pre cur = pre[j];
          fetch(n + PF, pre[j]);
          const int par = (n - n0) & 1, m = n - P;
          s_ex[par][tid + P] = cur.r1n * ring_iml<P>(L, X, n);
          __syncthreads();
          double k, mm;
          ring_xsum<P>(&s_ex[par][tid], kx, mx, k, mm);
          ring_push(rK, k);
          ring_push(rM, mm);
          if (cur.wr)
            {
              double Kv, Mv;
              ring_ysum<P>(T, m, rK, rM, Kv, Mv);
              const double il = ring_iml<P>(L, X, m);
              const int64_t r = lat_id(L, X, m);
              O.t[r] = il * (cur.o_r2 - Kv);
              O.m1[r] = il * cur.o_r1;
            }
        }
    }
}

// ---- Chebyshev sweep on S~ -------------------------------------------------------------------------------------
// epi 0: d = c0 d + c1 dinv (rhs - S~ x), out = x + d;  epi 2: out = dinv S~ x (power iteration);  epi 3: out = S~ x
template <int P, int NT, int PF>
static __global__ void __launch_bounds__(NT) ring_sweep_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int NW = 2 * P + 1, HX = 3 * P, OUTW = NT - 2 * HX, DEL = 2 * P + 3;
  __shared__ double s_ex[2][3][NT + 2 * P];
  __shared__ RingTab<P> T;
  const int tid = threadIdx.x;
  ring_fill_tab<P>(T, L, tid);
  for (int i = tid; i < 6 * (NT + 2 * P); i += NT)
    (&s_ex[0][0][0])[i] = 0.0;
  const int X = O.lo + blockIdx.x * OUTW - HX + tid;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  const bool colok = X >= 0 && X < L.fw, outcol = tid >= HX && tid < NT - HX && X < L.fw - O.lo;
  __syncthreads();
  const int tx = P == 2 ? (X & 1) : 0;
  double kx[NW], mx[NW], r1K[NW], r1M[NW], r2K[NW], r2M[NW], r3K[NW], r3M[NW], mxd[DEL];
#pragma unroll
  for (int dd = 0; dd < NW; ++dd)
    {
      kx[dd] = T.v[0][tx][dd], mx[dd] = T.v[1][tx][dd];
      r1K[dd] = r1M[dd] = r2K[dd] = r2M[dd] = r3K[dd] = r3M[dd] = 0;
    }
#pragma unroll
  for (int i = 0; i < DEL; ++i)
    mxd[i] = 0;
  double c0 = 0, c1 = 0;
  if (O.epi == 0)
    c0 = O.cheb[0], c1 = O.cheb[1];
  auto ldx = [&](int Y) { return O.x[O.x_vec ? ring_vaddr(L, X, Y, 0) : lat_id(L, X, Y)]; };
  const int n0 = Y0 - HX, nx1 = Y1 - 1 + HX, n1 = nx1 + 2; // last x row needed; last iteration (output row Y1 - 1)
  struct Pre
  {
    double x, rhs, dinv, d, xo;
    int32_t r;
    bool wr;
  };
  auto fetch = [&](int n, Pre &q) {
    q.x = (colok && n <= nx1 && n < L.fh) ? ldx(n) : 0.0;
    const int m = n - 2 - 3 * P; // output row of iteration n
    q.wr = outcol && m >= Y0 && m < Y1;
    q.rhs = q.dinv = q.d = q.xo = 0.0;
    q.r = 0;
    if (q.wr)
      {
        q.r = lat_id(L, X, m);
        if (O.epi == 0)
          q.rhs = O.rhs[q.r], q.d = O.d[q.r], q.xo = ldx(m);
        if (O.epi != 3)
          q.dinv = O.dinv[q.r];
      }
  };
  Pre pre[PF];
#pragma unroll
  for (int j = 0; j < PF; ++j)
    fetch(n0 + j, pre[j]);
  double a1 = 0, a2 = 0;
  for (int nb = n0; nb <= n1; nb += PF)
    {
#pragma unroll
      for (int j = 0; j < PF; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          const Pre cur = pre[j];
          fetch(n + PF, pre[j]);
          const int par = (n - n0) & 1, m = n - 2 - 3 * P;
          s_ex[par][0][tid + P] = cur.x;
          s_ex[par][1][tid + P] = a1;
          s_ex[par][2][tid + P] = a2;
          __syncthreads();
          double k, mm, Kv, Mv;
          ring_xsum<P>(&s_ex[par][0][tid], kx, mx, k, mm); // x, row n
          ring_push(r1K, k);
          ring_push(r1M, mm);
          ring_xsum<P>(&s_ex[par][1][tid], kx, mx, k, mm); // t1, row n - 1 - p
          ring_push(r2K, k);
          ring_push(r2M, mm);
          ring_xsum<P>(&s_ex[par][2][tid], kx, mx, k, mm); // t2, row n - 2 - 2p
          ring_push(r3K, k);
          ring_push(r3M, mm);
          ring_ysum<P>(T, n - P, r1K, r1M, Kv, Mv); // t1 = M_L^-1 K x at row n - p; (M x) kept for the output row
          a1 = Kv * ring_iml<P>(L, X, n - P);
          ring_push(mxd, Mv);
          ring_ysum<P>(T, n - 1 - 2 * P, r2K, r2M, Kv, Mv); // t2 = M_L^-1 K t1 at row n - 1 - 2p
          a2 = Kv * ring_iml<P>(L, X, n - 1 - 2 * P);
          if (cur.wr)
            {
              ring_ysum<P>(T, m, r3K, r3M, Kv, Mv); // K t2, M t2 at the output row
              const double Sx = O.bbar * mxd[0] + O.dtth * (Kv - 2.0 * Mv);
              const int64_t r = cur.r;
              double res;
              if (O.epi == 0)
                {
                  const double dn = c0 * cur.d + c1 * cur.dinv * (cur.rhs - Sx);
                  O.d[r] = dn;
                  res = cur.xo + dn;
                }
              else
                res = O.epi == 2 ? cur.dinv * Sx : Sx;
              O.out[O.out_vec ? (r < L.Np ? r : 3 * r) : r] = res;
            }
        }
    }
}

// ---- back substitution: chi = M_L^-1 (r1 - K psi), phi = M_L^-1 (r2 - K chi); psi = component 0 of z -------------
template <int P, int NT, int PF>
static __global__ void __launch_bounds__(NT) ring_back_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int NW = 2 * P + 1, HX = 2 * P, OUTW = NT - 2 * HX;
  __shared__ double s_ex[2][2][NT + 2 * P];
  __shared__ RingTab<P> T;
  const int tid = threadIdx.x;
  ring_fill_tab<P>(T, L, tid);
  for (int i = tid; i < 4 * (NT + 2 * P); i += NT)
    (&s_ex[0][0][0])[i] = 0.0;
  const int X = O.lo + blockIdx.x * OUTW - HX + tid;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  const bool colok = X >= 0 && X < L.fw, outcol = tid >= HX && tid < NT - HX && X < L.fw - O.lo;
  __syncthreads();
  const int tx = P == 2 ? (X & 1) : 0;
  double kx[NW], mx[NW], r1K[NW], r1M[NW], r2K[NW], r2M[NW];
#pragma unroll
  for (int dd = 0; dd < NW; ++dd)
    {
      kx[dd] = T.v[0][tx][dd], mx[dd] = T.v[1][tx][dd];
      r1K[dd] = r1M[dd] = r2K[dd] = r2M[dd] = 0;
    }
  const int n0 = Y0 - HX, nx1 = Y1 - 1 + HX, n1 = nx1 + 1;
  struct Pre
  {
    double x, o_r1, o_r2;
    bool chi_ok, wr;
  };
  auto fetch = [&](int n, Pre &q) {
    q.x = (colok && n <= nx1 && n < L.fh) ? O.z[ring_vaddr(L, X, n, 0)] : 0.0;
    const int mc = n - P, m = n - 1 - 2 * P; // rows of chi / phi produced in iteration n
    q.chi_ok = colok && mc >= 0 && mc < L.fh && n <= n1;
    q.wr = outcol && m >= Y0 && m < Y1;
    q.o_r1 = q.chi_ok ? O.in[ring_vaddr(L, X, mc, 1)] : 0.0;
    q.o_r2 = q.wr ? O.in[ring_vaddr(L, X, m, 2)] : 0.0;
  };
  Pre pre[PF];
#pragma unroll
  for (int j = 0; j < PF; ++j)
    fetch(n0 + j, pre[j]);
  double a1 = 0;
  for (int nb = n0; nb <= n1; nb += PF)
    {
#pragma unroll
      for (int j = 0; j < PF; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          const Pre cur = pre[j];
          fetch(n + PF, pre[j]);
          const int par = (n - n0) & 1, mc = n - P, m = n - 1 - 2 * P;
          s_ex[par][0][tid + P] = cur.x;
          s_ex[par][1][tid + P] = a1;
          __syncthreads();
          double k, mm, Kv, Mv;
          ring_xsum<P>(&s_ex[par][0][tid], kx, mx, k, mm); // psi, row n
          ring_push(r1K, k);
          ring_push(r1M, mm);
          ring_xsum<P>(&s_ex[par][1][tid], kx, mx, k, mm); // chi, row n - 1 - p
          ring_push(r2K, k);
          ring_push(r2M, mm);
          ring_ysum<P>(T, mc, r1K, r1M, Kv, Mv);
          a1 = cur.chi_ok ? ring_iml<P>(L, X, mc) * (cur.o_r1 - Kv) : 0.0;
          if (outcol && mc >= Y0 && mc < Y1)
            O.z[ring_vaddr(L, X, mc, 1)] = a1;
          if (cur.wr)
            {
              ring_ysum<P>(T, m, r2K, r2M, Kv, Mv);
              O.z[ring_vaddr(L, X, m, 2)] = ring_iml<P>(L, X, m) * (cur.o_r2 - Kv);
            }
        }
    }
}
} // namespace pfc
