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

// The windows of the last 2p+1 x sums of a stage stay in registers WITHOUT shifting: the row loop is unrolled by NW = 2p+1,
// row phase j (compile time after unrolling) writes slot j, and element dd (0 = oldest row) of the window is slot
// (j + 1 + dd) % NW.  (K v)(row) = sum my rK + ky rM and (M v)(row) = sum my rM, window centred on `row`.
template <int P>
__device__ __forceinline__ void ring_ysum(const RingTab<P> &T, int row, int j, const double (&rK)[2 * P + 1],
                                          const double (&rM)[2 * P + 1], double &Kv, double &Mv)
{
  constexpr int NW = 2 * P + 1;
  const int ty = P == 2 ? (row & 1) : 0;
  Kv = 0, Mv = 0;
#pragma unroll
  for (int dd = 0; dd < NW; ++dd)
    {
      const int sl = (j + 1 + dd) % NW;
      const double my = T.v[3][ty][dd], ky = T.v[2][ty][dd];
      Kv = fma(my, rK[sl], Kv);
      Kv = fma(ky, rM[sl], Kv);
      Mv = fma(my, rM[sl], Mv);
    }
}

// Addresses of one thread's column.  FAST: every row and column the CTA touches (warm-up rows included) lies in the
// owned rectangle, where the id is (Y - ey0) w + (X - ex0) and vectors are component-major; otherwise the ring index.
// Nearly all CTAs are FAST; the kernels are issue bound, and the general index costs more than the arithmetic.
template <bool FAST>
struct RingCol
{
  const LatGeom &L;
  const int X, base, w, Np;
  __device__ __forceinline__ RingCol(const LatGeom &L_, int X_)
    : L(L_), X(X_), base(X_ - L_.ex0 - L_.ey0 * L_.w), w(L_.w), Np((int)L_.Np)
  {}
  __device__ __forceinline__ int id(int Y) const { return FAST ? base + Y * w : lat_id(L, X, Y); }
  __device__ __forceinline__ int64_t v(int Y, int c) const { return FAST ? (int64_t)(c * Np + base + Y * w) : ring_vaddr(L, X, Y, c); }
  // component 0 of the node with id r in a 3-vector
  __device__ __forceinline__ int64_t v0(int r) const { return (FAST || r < Np) ? (int64_t)r : 3 * (int64_t)r; }
};

// CTA-uniform: columns [X0, X0 + NT) and rows [Ya, Yb] inside the owned rectangle
__device__ __forceinline__ bool ring_cta_fast(const LatGeom &L, int X0, int NT, int Ya, int Yb)
{
  return X0 >= L.ex0 && X0 + NT <= L.ex0 + L.w && Ya >= L.ey0 && Yb < L.ey0 + L.h;
}

// ---- t kernel: m1 = M_L^-1 r1, t = M_L^-1 (r2 - K m1) on the window of margin lo -------------------------------
template <int P, int NT, int PF, bool FAST>
__device__ __forceinline__ void ring_t_rows(const LatGeom &L, const RingOp &O, double (&s_ex)[2][NT + 2 * P],
                                            const RingTab<P> &T, int X, int Y0, int Y1)
{
  constexpr int NW = 2 * P + 1, HX = P;
  static_assert(NW % PF == 0, "the prefetch ring turns with the row phase");
  const int tid = threadIdx.x;
  const RingCol<FAST> A(L, X);
  const bool colok = FAST || (X >= 0 && X < L.fw);
  const bool outcol = tid >= HX && tid < NT - HX && (FAST || X < L.fw - O.lo);
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
    q.r1n = (colok && n <= n1 && (FAST || n < L.fh)) ? O.in[A.v(n, 1)] : 0.0;
    const int m = n - P;
    q.wr = outcol && m >= Y0 && m < Y1;
    q.o_r1 = q.o_r2 = 0.0;
    if (q.wr)
      {
        q.o_r1 = O.in[A.v(m, 1)];
        q.o_r2 = O.in[A.v(m, 2)];
      }
  };
  Pre pre[PF];
#pragma unroll
  for (int j = 0; j < PF; ++j)
    fetch(n0 + j, pre[j]);
  for (int nb = n0; nb <= n1; nb += NW)
    {
#pragma unroll
      for (int j = 0; j < NW; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          const Pre cur = pre[j % PF];
          fetch(n + PF, pre[j % PF]);
          const int par = (n - n0) & 1, m = n - P;
          s_ex[par][tid + P] = cur.r1n * ring_iml<P>(L, X, n);
          __syncthreads();
          ring_xsum<P>(&s_ex[par][tid], kx, mx, rK[j], rM[j]);
          if (cur.wr)
            {
              double Kv, Mv;
              ring_ysum<P>(T, m, j, rK, rM, Kv, Mv);
              const double il = ring_iml<P>(L, X, m);
              const int r = A.id(m);
              O.t[r] = il * (cur.o_r2 - Kv);
              O.m1[r] = il * cur.o_r1;
            }
        }
    }
}

template <int P, int NT, int PF>
static __global__ void __launch_bounds__(NT) ring_t_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int HX = P, OUTW = NT - 2 * HX;
  __shared__ double s_ex[2][NT + 2 * P];
  __shared__ RingTab<P> T;
  const int tid = threadIdx.x;
  ring_fill_tab<P>(T, L, tid);
  for (int i = tid; i < 2 * (NT + 2 * P); i += NT)
    (&s_ex[0][0])[i] = 0.0;
  const int X0 = O.lo + blockIdx.x * OUTW - HX;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  __syncthreads();
  if (ring_cta_fast(L, X0, NT, Y0 - HX, Y1 - 1 + HX))
    ring_t_rows<P, NT, PF, true>(L, O, s_ex, T, X0 + tid, Y0, Y1);
  else
    ring_t_rows<P, NT, PF, false>(L, O, s_ex, T, X0 + tid, Y0, Y1);
}

// ---- Chebyshev sweep on S~ -------------------------------------------------------------------------------------
// epi 0: d = c0 d + c1 dinv (rhs - S~ x), out = x + d;  epi 2: out = dinv S~ x (power iteration);  epi 3: out = S~ x
// Stage s lags stage s - 1 by p + 1 rows: in iteration n the thread exchanges x(n), t1(n - 1 - p), t2(n - 2 - 2p) and
// finishes the output row m = n - 2 - 3p.  (M x)(m) was formed 2p + 2 iterations earlier; it waits in a private
// shared-memory ring (s_del), which costs one load and one store instead of 2p + 2 register moves.
template <int P, int NT, int PF, bool FAST>
__device__ __forceinline__ void ring_sweep_rows(const LatGeom &L, const RingOp &O, double (&s_ex)[2][3][NT + 2 * P],
                                                double (&s_del)[2 * P + 2][NT], const RingTab<P> &T, int X, int Y0, int Y1)
{
  constexpr int NW = 2 * P + 1, HX = 3 * P, DEL = 2 * P + 2;
  static_assert(NW % PF == 0, "the prefetch ring turns with the row phase");
  const int tid = threadIdx.x;
  const RingCol<FAST> A(L, X);
  const bool colok = FAST || (X >= 0 && X < L.fw);
  const bool outcol = tid >= HX && tid < NT - HX && (FAST || X < L.fw - O.lo);
  const int tx = P == 2 ? (X & 1) : 0;
  double kx[NW], mx[NW], r1K[NW], r1M[NW], r2K[NW], r2M[NW], r3K[NW], r3M[NW];
#pragma unroll
  for (int dd = 0; dd < NW; ++dd)
    {
      kx[dd] = T.v[0][tx][dd], mx[dd] = T.v[1][tx][dd];
      r1K[dd] = r1M[dd] = r2K[dd] = r2M[dd] = r3K[dd] = r3M[dd] = 0;
    }
  const int epi = O.epi;
  double c0 = 0, c1 = 0;
  if (epi == 0)
    c0 = O.cheb[0], c1 = O.cheb[1];
  const bool xv = O.x_vec, ov = O.out_vec;
  auto ldx = [&](int Y) { return O.x[xv ? A.v(Y, 0) : (int64_t)A.id(Y)]; };
  const int n0 = Y0 - HX, nx1 = Y1 - 1 + HX, n1 = nx1 + 2; // last x row needed; last iteration (output row Y1 - 1)
  struct Pre
  {
    double x, rhs, dinv, d, xo;
    int r;
    bool wr;
  };
  auto fetch = [&](int n, Pre &q) {
    q.x = (colok && n <= nx1 && (FAST || n < L.fh)) ? ldx(n) : 0.0;
    const int m = n - 2 - 3 * P; // output row of iteration n
    q.wr = outcol && m >= Y0 && m < Y1;
    q.rhs = q.dinv = q.d = q.xo = 0.0;
    q.r = 0;
    if (q.wr)
      {
        q.r = A.id(m);
        if (epi == 0)
          q.rhs = O.rhs[q.r], q.d = O.d[q.r], q.xo = ldx(m);
        if (epi != 3)
          q.dinv = O.dinv[q.r];
      }
  };
  Pre pre[PF];
#pragma unroll
  for (int j = 0; j < PF; ++j)
    fetch(n0 + j, pre[j]);
  double a1 = 0, a2 = 0;
  int dl = 0;
  for (int nb = n0; nb <= n1; nb += NW)
    {
#pragma unroll
      for (int j = 0; j < NW; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          const Pre cur = pre[j % PF];
          fetch(n + PF, pre[j % PF]);
          const int par = (n - n0) & 1, m = n - 2 - 3 * P;
          s_ex[par][0][tid + P] = cur.x;
          s_ex[par][1][tid + P] = a1;
          s_ex[par][2][tid + P] = a2;
          __syncthreads();
          double Kv, Mv;
          ring_xsum<P>(&s_ex[par][0][tid], kx, mx, r1K[j], r1M[j]); // x, row n
          ring_xsum<P>(&s_ex[par][1][tid], kx, mx, r2K[j], r2M[j]); // t1, row n - 1 - p
          ring_xsum<P>(&s_ex[par][2][tid], kx, mx, r3K[j], r3M[j]); // t2, row n - 2 - 2p
          ring_ysum<P>(T, n - P, j, r1K, r1M, Kv, Mv); // t1 = M_L^-1 K x at row n - p; (M x) waits for its output row
          a1 = Kv * ring_iml<P>(L, X, n - P);
          const double Mx_m = s_del[dl][tid];
          s_del[dl][tid] = Mv;
          dl = dl + 1 == DEL ? 0 : dl + 1;
          ring_ysum<P>(T, n - 1 - 2 * P, j, r2K, r2M, Kv, Mv); // t2 = M_L^-1 K t1 at row n - 1 - 2p
          a2 = Kv * ring_iml<P>(L, X, n - 1 - 2 * P);
          if (cur.wr)
            {
              ring_ysum<P>(T, m, j, r3K, r3M, Kv, Mv); // K t2, M t2 at the output row
              const double Sx = O.bbar * Mx_m + O.dtth * (Kv - 2.0 * Mv);
              double res;
              if (epi == 0)
                {
                  const double dn = c0 * cur.d + c1 * cur.dinv * (cur.rhs - Sx);
                  O.d[cur.r] = dn;
                  res = cur.xo + dn;
                }
              else
                res = epi == 2 ? cur.dinv * Sx : Sx;
              O.out[ov ? A.v0(cur.r) : (int64_t)cur.r] = res;
            }
        }
    }
}

template <int P, int NT, int PF>
static __global__ void __launch_bounds__(NT) ring_sweep_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int HX = 3 * P, OUTW = NT - 2 * HX;
  __shared__ double s_ex[2][3][NT + 2 * P];
  __shared__ double s_del[2 * P + 2][NT];
  __shared__ RingTab<P> T;
  const int tid = threadIdx.x;
  ring_fill_tab<P>(T, L, tid);
  for (int i = tid; i < 6 * (NT + 2 * P); i += NT)
    (&s_ex[0][0][0])[i] = 0.0;
  for (int i = 0; i < 2 * P + 2; ++i)
    s_del[i][tid] = 0.0;
  const int X0 = O.lo + blockIdx.x * OUTW - HX;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  __syncthreads();
  if (ring_cta_fast(L, X0, NT, Y0 - HX, Y1 - 1 + HX))
    ring_sweep_rows<P, NT, PF, true>(L, O, s_ex, s_del, T, X0 + tid, Y0, Y1);
  else
    ring_sweep_rows<P, NT, PF, false>(L, O, s_ex, s_del, T, X0 + tid, Y0, Y1);
}

// ---- back substitution: chi = M_L^-1 (r1 - K psi), phi = M_L^-1 (r2 - K chi); psi = component 0 of z -------------
template <int P, int NT, int PF, bool FAST>
__device__ __forceinline__ void ring_back_rows(const LatGeom &L, const RingOp &O, double (&s_ex)[2][2][NT + 2 * P],
                                               const RingTab<P> &T, int X, int Y0, int Y1)
{
  constexpr int NW = 2 * P + 1, HX = 2 * P;
  static_assert(NW % PF == 0, "the prefetch ring turns with the row phase");
  const int tid = threadIdx.x;
  const RingCol<FAST> A(L, X);
  const bool colok = FAST || (X >= 0 && X < L.fw);
  const bool outcol = tid >= HX && tid < NT - HX && (FAST || X < L.fw - O.lo);
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
    q.x = (colok && n <= nx1 && (FAST || n < L.fh)) ? O.z[A.v(n, 0)] : 0.0;
    const int mc = n - P, m = n - 1 - 2 * P; // rows of chi / phi produced in iteration n
    q.chi_ok = colok && mc >= (FAST ? n0 : 0) && (FAST || mc < L.fh) && n <= n1;
    q.wr = outcol && m >= Y0 && m < Y1;
    q.o_r1 = q.chi_ok ? O.in[A.v(mc, 1)] : 0.0;
    q.o_r2 = q.wr ? O.in[A.v(m, 2)] : 0.0;
  };
  Pre pre[PF];
#pragma unroll
  for (int j = 0; j < PF; ++j)
    fetch(n0 + j, pre[j]);
  double a1 = 0;
  for (int nb = n0; nb <= n1; nb += NW)
    {
#pragma unroll
      for (int j = 0; j < NW; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          const Pre cur = pre[j % PF];
          fetch(n + PF, pre[j % PF]);
          const int par = (n - n0) & 1, mc = n - P, m = n - 1 - 2 * P;
          s_ex[par][0][tid + P] = cur.x;
          s_ex[par][1][tid + P] = a1;
          __syncthreads();
          double Kv, Mv;
          ring_xsum<P>(&s_ex[par][0][tid], kx, mx, r1K[j], r1M[j]); // psi, row n
          ring_xsum<P>(&s_ex[par][1][tid], kx, mx, r2K[j], r2M[j]); // chi, row n - 1 - p
          ring_ysum<P>(T, mc, j, r1K, r1M, Kv, Mv);
          a1 = cur.chi_ok ? ring_iml<P>(L, X, mc) * (cur.o_r1 - Kv) : 0.0;
          if (outcol && mc >= Y0 && mc < Y1)
            O.z[A.v(mc, 1)] = a1;
          if (cur.wr)
            {
              ring_ysum<P>(T, m, j, r2K, r2M, Kv, Mv);
              O.z[A.v(m, 2)] = ring_iml<P>(L, X, m) * (cur.o_r2 - Kv);
            }
        }
    }
}

template <int P, int NT, int PF>
static __global__ void __launch_bounds__(NT) ring_back_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int HX = 2 * P, OUTW = NT - 2 * HX;
  __shared__ double s_ex[2][2][NT + 2 * P];
  __shared__ RingTab<P> T;
  const int tid = threadIdx.x;
  ring_fill_tab<P>(T, L, tid);
  for (int i = tid; i < 4 * (NT + 2 * P); i += NT)
    (&s_ex[0][0][0])[i] = 0.0;
  const int X0 = O.lo + blockIdx.x * OUTW - HX;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  __syncthreads();
  if (ring_cta_fast(L, X0, NT, Y0 - HX, Y1 - 1 + HX + 1))
    ring_back_rows<P, NT, PF, true>(L, O, s_ex, T, X0 + tid, Y0, Y1);
  else
    ring_back_rows<P, NT, PF, false>(L, O, s_ex, T, X0 + tid, Y0, Y1);
}
} // namespace pfc
