// pfc_ring_q1.cuh — Q1 kernels in closed form.  The 1-D rows of FE_Q(1) and the 2 x 2 Gauss rule are written out, so
// that the iterate-dependent blocks B, C of the Jacobian (assemble_system, mpi.cpp:421-433) cost about 130 FP64
// instructions per cell and the constant blocks K, M, D = dt theta (K - 2 M) are 9-point stencils:
//   lat_j1_kernel:      y = J z on the owned nodes (J apply of GMRES), tiled like lat_kernel<1, LAT_J>
//   ring_rhs1_kernel:   m1 = M_L^-1 r1,  t = M_L^-1 (r2 - K m1),  rhs = r0 - C m1 - D t,  d = x = dinv rhs / theta
//                       (t kernel + Schur right-hand side of precond_apply_ring in one streaming pass; m1 and t never
//                       leave the SM: 64 B per node instead of 32 + 72)
//   ring_sweep1_kernel, ring_back1_kernel: the pipelines of ring_sweep_kernel / ring_back_kernel with the rows as
//                       immediates (fewer registers, more resident warps)
// Why: lat_kernel<1, LAT_J> integrates every block at the Gauss points, about 390 FP64 instructions per cell, and runs at
// 77 % of the FP64 pipe (profiles/r02i): on B200 (64 FP64 lanes per SM against 44 B per clock of HBM) that pipe, not
// HBM, bounds it.
#pragma once
#include "pfc_ring.cuh"

namespace pfc
{
namespace q1
{
constexpr double G0 = 0.21132486540518711775; // Gauss points of [0, 1]
constexpr double G1 = 0.78867513459481288225;

// x direction of a node row for the cell right of the thread's column: values at the two Gauss abscissae, difference
struct XS
{
  double a0, a1, b;
};
__device__ __forceinline__ XS xstage(double uL, double uR)
{
  XS s;
  s.b = uR - uL;
  s.a0 = fma(G0, s.b, uL);
  s.a1 = fma(G1, s.b, uL);
  return s;
}
// values at the Gauss points [qy][qx] from the x stages of the cell's top (t) and bottom (b) node rows;
// dy[qx] = reference-cell y derivative at Gauss column qx (independent of qy)
__device__ __forceinline__ void values(const XS &t, const XS &b, double (&v)[2][2], double (&dy)[2])
{
  dy[0] = b.a0 - t.a0;
  dy[1] = b.a1 - t.a1;
  v[0][0] = fma(G0, dy[0], t.a0);
  v[0][1] = fma(G0, dy[1], t.a1);
  v[1][0] = fma(G1, dy[0], t.a0);
  v[1][1] = fma(G1, dy[1], t.a1);
}
// reference-cell x derivative at Gauss row qy (independent of qx)
__device__ __forceinline__ void xderiv(const XS &t, const XS &b, double (&dx)[2])
{
  const double db = b.b - t.b;
  dx[0] = fma(G0, db, t.b);
  dx[1] = fma(G1, db, t.b);
}
// sum_q N_i(q) A(q) for the four corners: c[iy][ix]  (A without JxW: the caller scales once per node)
__device__ __forceinline__ void test(const double (&A)[2][2], double (&c)[2][2])
{
  double t0[2], t1[2]; // [qy] tested in x with N_0, N_1
#pragma unroll
  for (int qy = 0; qy < 2; ++qy)
    {
      const double dA = A[qy][0] - A[qy][1];
      t1[qy] = fma(G0, dA, A[qy][1]); // G0 A0 + G1 A1
      t0[qy] = fma(G1, dA, A[qy][1]); // G1 A0 + G0 A1
    }
  const double d0 = t0[0] - t0[1], d1 = t1[0] - t1[1];
  c[0][0] = fma(G1, d0, t0[1]);
  c[1][0] = fma(G0, d0, t0[1]);
  c[0][1] = fma(G1, d1, t1[1]);
  c[1][1] = fma(G0, d1, t1[1]);
}

// unscaled 1-D sums of a row at the thread's column: k = 2 v0 - vL - vR (Kx h), m = vL + 4 v0 + vR (6 Mx / h)
__device__ __forceinline__ void xsum(const double *row, double &k, double &m)
{
  const double s = row[0] + row[2], c = row[1];
  k = fma(2.0, c, -s);
  m = fma(4.0, c, s);
}
// constants of the 2-D operators built from those sums
struct Cst
{
  double kA, kB, mC; // K = kA (k- + 4 k0 + k+) + kB (2 m0 - m- - m+),  M = mC (m- + 4 m0 + m+)
  __device__ __forceinline__ explicit Cst(const LatGeom &L)
    : kA(L.ihx * L.hy * (1.0 / 6.0)), kB(L.hx * L.ihy * (1.0 / 6.0)), mC(L.hx * L.hy * (1.0 / 36.0))
  {}
};
// windows hold rows in slots (j + 1 + dd) % 3, dd = 0 oldest (as ring_ysum); the centre row is dd = 1
template <bool WANT_M>
__device__ __forceinline__ void ysum(const Cst &C, int j, const double (&wk)[3], const double (&wm)[3], double &Kv, double &Mv)
{
  const int a = (j + 1) % 3, b = (j + 2) % 3, c = j % 3; // oldest, centre, newest
  const double sk = wk[a] + wk[c], sm = wm[a] + wm[c];
  Kv = C.kA * fma(4.0, wk[b], sk) + C.kB * fma(2.0, wm[b], -sm);
  if (WANT_M)
    Mv = C.mC * fma(4.0, wm[b], sm);
}
} // namespace q1

// cp.async (LDGSTS), 8 bytes: global -> shared without a register in between
__device__ __forceinline__ void cp_async8(double *smem_dst, const double *gsrc)
{
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}


// ---- fused t kernel + Schur right-hand side + first Chebyshev step (Q1) -------------------------------------------
// A thread owns node column X and the cell column between X and X + 1.  Iteration n: publish m1(n) = r1(n) / ml,
// t(n - 2), psi_n(n) and the hand-over of the cell formed in iteration n - 1;  K m1 (n - 1) -> t(n - 1);  D t (n - 3);
// cell (rows n - 1, n) of C m1: its two left corners stay with the thread, the two right corners go to the right
// neighbour through the exchange row of the NEXT iteration (still one __syncthreads per row);  output row m = n - 3.
template <int NT, bool FAST>
__device__ __forceinline__ void ring_rhs1_rows(const LatGeom &L, const RingOp &O, double (&s_ex)[2][4][NT + 2],
                                               int X, int Y0, int Y1)
{
  constexpr int HX = 2;
  const int tid = threadIdx.x;
  const RingCol<FAST> A(L, X);
  const q1::Cst C(L);
  const bool colok = FAST || (X >= 0 && X < L.fw);
  const bool outcol = tid >= HX && tid < NT - HX && (FAST || X < L.fw - O.lo);
  const double iml = L.ihx * L.ihy, jxw = 0.25 * L.hx * L.hy;
  const double cm0 = -O.dtth * (1.0 + O.eps), cm2 = -3.0 * O.dtth;
  const double cheb0 = O.cheb[0];
  double w0k[3] = {0, 0, 0}, w0m[3] = {0, 0, 0}, w1k[3] = {0, 0, 0}, w1m[3] = {0, 0, 0};
  const int n0 = Y0 - HX, nl1 = Y1 + 1, n1 = Y1 + 2; // last row of r1 / psi needed; last iteration (output row Y1 - 1)
  struct Pre
  {
    double r1, ps, r2, r0, dinv;
    int r;
    bool wr, t_ok;
  };
  auto fetch = [&](int n, Pre &q) {
    const bool ld = colok && n <= nl1 && (FAST || n < L.fh);
    q.r1 = ld ? O.in[A.v(n, 1)] : 0.0;
    q.ps = ld ? O.psi[A.v(n, 0)] : 0.0;
    const int mt = n - 1, m = n - 3; // row of t formed in iteration n; output row
    q.t_ok = colok && mt >= n0 && mt <= nl1 - 1 && (FAST || (mt >= 0 && mt < L.fh));
    q.r2 = q.t_ok ? O.in[A.v(mt, 2)] : 0.0;
    q.wr = outcol && m >= Y0 && m < Y1;
    q.r0 = q.dinv = 0.0;
    q.r = 0;
    if (q.wr)
      {
        q.r = A.id(m);
        q.r0 = O.in[A.v(m, 0)];
        q.dinv = O.dinv[q.r];
      }
  };
  Pre pre;
  fetch(n0, pre);
  double tpub = 0, rpub = 0;                   // t(n - 2) and the hand-over, published in the next iteration
  q1::XS m_top = {0, 0, 0}, p_top = {0, 0, 0}; // x stages of row n - 1
  double c01p = 0, c11p = 0, Pprev = 0, Vhold = 0;
  for (int nb = n0; nb <= n1; nb += 3)
    {
#pragma unroll
      for (int j = 0; j < 3; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          const Pre cur = pre;
          fetch(n + 1, pre);
          const int par = (n - n0) & 1;
          const double m1n = cur.r1 * iml;
          s_ex[par][0][tid + 1] = m1n;
          s_ex[par][1][tid + 1] = tpub;
          s_ex[par][2][tid + 1] = cur.ps;
          s_ex[par][3][tid + 1] = rpub;
          __syncthreads();
          double Kv, Mv;
          q1::xsum(&s_ex[par][0][tid], w0k[j], w0m[j]); // m1, row n
          q1::ysum<false>(C, j, w0k, w0m, Kv, Mv);     // K m1 at row n - 1
          tpub = cur.t_ok ? iml * (cur.r2 - Kv) : 0.0;  // t(n - 1)
          q1::xsum(&s_ex[par][1][tid], w1k[j], w1m[j]); // t, row n - 2
          q1::ysum<true>(C, j, w1k, w1m, Kv, Mv);      // K t, M t at row n - 3
          const double Dt = O.dtth * fma(-2.0, Mv, Kv);
          // cell between columns X, X + 1 and rows n - 1, n:  C m1 = <N, cm m1>,  cm = -dt th ((1 + eps) + 3 psi^2)
          const q1::XS m_bot = q1::xstage(m1n, s_ex[par][0][tid + 2]);
          const q1::XS p_bot = q1::xstage(cur.ps, s_ex[par][2][tid + 2]);
          double mv[2][2], pv[2][2], dy[2], Aq[2][2], c[2][2];
          q1::values(m_top, m_bot, mv, dy);
          q1::values(p_top, p_bot, pv, dy);
#pragma unroll
          for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b)
              Aq[a][b] = fma(cm2 * pv[a][b], pv[a][b], cm0) * mv[a][b];
          q1::test(Aq, c);
          m_top = m_bot;
          p_top = p_bot;
          const double Pnew = c[0][0] + c01p; // node (X, n - 1): this cell and the one above it
          rpub = c[0][1] + c11p;              // node (X + 1, n - 1): handed to the right neighbour
          c01p = c[1][0];
          c11p = c[1][1];
          const double Vuse = Vhold;          // complete sum of row n - 3
          Vhold = Pprev + s_ex[par][3][tid];  // row n - 2: own part + the left neighbour's hand-over
          Pprev = Pnew;
          if (cur.wr)
            {
              const double rhs = cur.r0 - jxw * Vuse - Dt;
              const double x0 = cheb0 * cur.dinv * rhs;
              O.out[cur.r] = rhs;
              O.d[cur.r] = x0;
              O.out2[cur.r] = x0; // the first iterate
            }
        }
    }
}

template <int NT>
static __global__ void __launch_bounds__(NT) ring_rhs1_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int HX = 2, OUTW = NT - 2 * HX;
  __shared__ double s_ex[2][4][NT + 2];
  const int tid = threadIdx.x;
  for (int i = tid; i < 8 * (NT + 2); i += NT)
    (&s_ex[0][0][0])[i] = 0.0;
  const int X0 = O.lo + blockIdx.x * OUTW - HX;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  __syncthreads();
  if (ring_cta_fast(L, X0, NT, Y0 - HX, Y1 + 1))
    ring_rhs1_rows<NT, true>(L, O, s_ex, X0 + tid, Y0, Y1);
  else
    ring_rhs1_rows<NT, false>(L, O, s_ex, X0 + tid, Y0, Y1);
}

// ---- Chebyshev sweep on S~, Q1 closed form (pipeline of ring_sweep_kernel, pfc_ring.cuh) ---------------------------
template <int NT, int PF, bool FAST>
__device__ __forceinline__ void ring_sweep1_rows(const LatGeom &L, const RingOp &O, double (&s_ex)[2][3][NT + 2],
                                                 double (&s_del)[4][NT], int X, int Y0, int Y1)
{
  constexpr int HX = 3, DEL = 4;
  static_assert(3 % PF == 0, "the prefetch ring turns with the row phase");
  const int tid = threadIdx.x;
  const RingCol<FAST> A(L, X);
  const q1::Cst C(L);
  const bool colok = FAST || (X >= 0 && X < L.fw);
  const bool outcol = tid >= HX && tid < NT - HX && (FAST || X < L.fw - O.lo);
  const double iml = L.ihx * L.ihy;
  double w1k[3] = {0, 0, 0}, w1m[3] = {0, 0, 0}, w2k[3] = {0, 0, 0}, w2m[3] = {0, 0, 0}, w3k[3] = {0, 0, 0}, w3m[3] = {0, 0, 0};
  const int epi = O.epi;
  double c0 = 0, c1 = 0;
  if (epi == 0)
    c0 = O.cheb[0], c1 = O.cheb[1];
  const bool xv = O.x_vec, ov = O.out_vec;
  auto ldx = [&](int Y) { return O.x[xv ? A.v(Y, 0) : (int64_t)A.id(Y)]; };
  const int n0 = Y0 - HX, nx1 = Y1 - 1 + HX, n1 = nx1 + 2;
  struct Pre
  {
    double x, rhs, dinv, d, xo;
    int r;
    bool wr;
  };
  auto fetch = [&](int n, Pre &q) {
    q.x = (colok && n <= nx1 && (FAST || n < L.fh)) ? ldx(n) : 0.0;
    const int m = n - 5; // output row of iteration n
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
  for (int nb = n0; nb <= n1; nb += 3)
    {
#pragma unroll
      for (int j = 0; j < 3; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          const Pre cur = pre[j % PF];
          fetch(n + PF, pre[j % PF]);
          const int par = (n - n0) & 1;
          s_ex[par][0][tid + 1] = cur.x;
          s_ex[par][1][tid + 1] = a1;
          s_ex[par][2][tid + 1] = a2;
          __syncthreads();
          double Kv, Mv;
          q1::xsum(&s_ex[par][0][tid], w1k[j], w1m[j]); // x, row n
          q1::xsum(&s_ex[par][1][tid], w2k[j], w2m[j]); // t1, row n - 2
          q1::xsum(&s_ex[par][2][tid], w3k[j], w3m[j]); // t2, row n - 4
          q1::ysum<true>(C, j, w1k, w1m, Kv, Mv);       // t1 = M_L^-1 K x at row n - 1; (M x) waits for its output row
          a1 = Kv * iml;
          const double Mx_m = s_del[dl][tid];
          s_del[dl][tid] = Mv;
          dl = (dl + 1) & (DEL - 1);
          q1::ysum<false>(C, j, w2k, w2m, Kv, Mv);      // t2 = M_L^-1 K t1 at row n - 3
          a2 = Kv * iml;
          if (cur.wr)
            {
              q1::ysum<true>(C, j, w3k, w3m, Kv, Mv);   // K t2, M t2 at the output row n - 5
              const double Sx = O.bbar * Mx_m + O.dtth * fma(-2.0, Mv, Kv);
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

template <int NT, int PF>
static __global__ void __launch_bounds__(NT) ring_sweep1_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int HX = 3, OUTW = NT - 2 * HX;
  __shared__ double s_ex[2][3][NT + 2];
  __shared__ double s_del[4][NT];
  const int tid = threadIdx.x;
  for (int i = tid; i < 6 * (NT + 2); i += NT)
    (&s_ex[0][0][0])[i] = 0.0;
  for (int i = 0; i < 4; ++i)
    s_del[i][tid] = 0.0;
  const int X0 = O.lo + blockIdx.x * OUTW - HX;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  __syncthreads();
  if (ring_cta_fast(L, X0, NT, Y0 - HX, Y1 - 1 + HX))
    ring_sweep1_rows<NT, PF, true>(L, O, s_ex, s_del, X0 + tid, Y0, Y1);
  else
    ring_sweep1_rows<NT, PF, false>(L, O, s_ex, s_del, X0 + tid, Y0, Y1);
}

// ---- the same sweep with its operands staged by cp.async (epi 0, interior CTAs) -------------------------------------
// ring_sweep1_kernel prefetches rows into registers: 150 registers, 12 warps per SM, and a warp that waits for its
// oldest load also waits for the newest ones (6 scoreboards).  Here every thread copies its element of x(n + D) and of
// rhs, dinv, d of the output row of iteration n + D straight into shared-memory rings (LDGSTS, no register, no
// scoreboard), one commit group per row; wait_group<D> + the row's __syncthreads make row n visible.  x lands directly
// in the exchange row its neighbours read.  CTAs on the boundary of the owned rectangle run ring_sweep1_rows.
// D = prefetch distance in rows; ring depths: x D + 7 (> D + 5: the output row is read again as x_old), operands D + 1
// FIRST (reduced solve, pfc_reduced.cuh): the sweep that follows the stencil-free step 0, x_0 = d_0 = alpha r with the
// constant Jacobi scaling: only r is staged (as "x"; S~ x_0 = alpha S~ r), nothing else is read, d is written only if
// another sweep follows (O.d != 0).  Replaces a pass that wrote x_0, d_0 and a sweep that read them back with rhs and dinv.
template <int NT, int D, bool FIRST = false>
static __global__ void __launch_bounds__(NT) ring_sweep1a_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int HX = 3, OUTW = NT - 2 * HX, RX = D + 7, RO = D + 1;
  constexpr int N_X = RX * (NT + 2), N_OP = 3 * RO * NT, N_EX = 2 * 2 * (NT + 2), N_DEL = 4 * NT;
  __shared__ double raw[N_X + N_OP + N_EX + N_DEL];
  const int tid = threadIdx.x;
  for (int i = tid; i < N_X + N_OP + N_EX + N_DEL; i += NT)
    raw[i] = 0.0;
  const int X0 = O.lo + blockIdx.x * OUTW - HX;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  __syncthreads();
  // interior CTAs address by arithmetic; CTAs that touch the ring take the general index for the ADDRESSES only (same pipeline)
  const bool fast = ring_cta_fast(L, X0, NT, Y0 - HX, Y1 - 1 + HX);
  const int X = X0 + tid;
  const bool colok = fast || (X >= 0 && X < L.fw);
  double(&s_x)[RX][NT + 2] = *reinterpret_cast<double(*)[RX][NT + 2]>(raw);
  double(&s_op)[3][RO][NT] = *reinterpret_cast<double(*)[3][RO][NT]>(raw + N_X);
  double(&s_ex)[2][2][NT + 2] = *reinterpret_cast<double(*)[2][2][NT + 2]>(raw + N_X + N_OP);
  double(&s_del)[4][NT] = *reinterpret_cast<double(*)[4][NT]>(raw + N_X + N_OP + N_EX);
  const q1::Cst C(L);
  const bool outcol = tid >= HX && tid < NT - HX && (fast || X < L.fw - O.lo);
  const double iml = L.ihx * L.ihy, c0 = O.cheb[0], c1 = O.cheb[1];
  const int w = L.w;
  const int64_t base = (int64_t)(X - L.ex0) - (int64_t)L.ey0 * w; // owned node (X, Y): id = base + Y w
  const bool xv = O.x_vec, ov = O.out_vec;
  auto sid = [&](int Y) { return fast ? base + (int64_t)Y * w : (int64_t)lat_id(L, X, Y); }; // index in a scalar field
  auto xid = [&](int Y) { return (fast || !xv) ? sid(Y) : ring_vaddr(L, X, Y, 0); };         // x: scalar field or component 0
  const int n0 = Y0 - HX, nx1 = Y1 - 1 + HX, n1 = nx1 + 2;
  // one commit group per row: x(n) and the operands of the output row n - 5
  auto issue = [&](int n) {
    if (n <= nx1)
      {
        if (colok)
          cp_async8(&s_x[(n - n0) % RX][tid + 1], (FIRST ? O.rhs : O.x) + xid(n));
        else
          s_x[(n - n0) % RX][tid + 1] = 0.0;
      }
    const int m = n - 5;
    if (!FIRST && outcol && m >= Y0 && m < Y1)
      {
        const int so = (n - n0) % RO;
        const int64_t r = sid(m);
        cp_async8(&s_op[0][so][tid], O.rhs + r);
        cp_async8(&s_op[1][so][tid], O.dinv + r);
        cp_async8(&s_op[2][so][tid], O.d + r);
      }
    cp_async_commit();
  };
#pragma unroll
  for (int k = 0; k < D; ++k)
    issue(n0 + k);
  double w1k[3] = {0, 0, 0}, w1m[3] = {0, 0, 0}, w2k[3] = {0, 0, 0}, w2m[3] = {0, 0, 0}, w3k[3] = {0, 0, 0}, w3m[3] = {0, 0, 0};
  double a1 = 0, a2 = 0;
  int dl = 0;
  for (int nb = n0; nb <= n1; nb += 3)
    {
#pragma unroll
      for (int j = 0; j < 3; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          issue(n + D);
          const int par = (n - n0) & 1, sx = (n - n0) % RX;
          s_ex[par][0][tid + 1] = a1;
          s_ex[par][1][tid + 1] = a2;
          cp_async_wait<D>();
          __syncthreads();
          double Kv, Mv;
          if (n <= nx1)
            q1::xsum(&s_x[sx][tid], w1k[j], w1m[j]); // x, row n
          else
            w1k[j] = w1m[j] = 0.0;
          q1::xsum(&s_ex[par][0][tid], w2k[j], w2m[j]); // t1, row n - 2
          q1::xsum(&s_ex[par][1][tid], w3k[j], w3m[j]); // t2, row n - 4
          q1::ysum<true>(C, j, w1k, w1m, Kv, Mv);       // t1 = M_L^-1 K x at row n - 1; (M x) waits for its output row
          a1 = Kv * iml;
          const double Mx_m = s_del[dl][tid];
          s_del[dl][tid] = Mv;
          dl = (dl + 1) & 3;
          q1::ysum<false>(C, j, w2k, w2m, Kv, Mv);      // t2 = M_L^-1 K t1 at row n - 3
          a2 = Kv * iml;
          const int m = n - 5;
          if (outcol && m >= Y0 && m < Y1)
            {
              q1::ysum<true>(C, j, w3k, w3m, Kv, Mv);   // K t2, M t2 at the output row
              const double Sx = O.bbar * Mx_m + O.dtth * fma(-2.0, Mv, Kv);
              const int64_t r = sid(m);
              if (FIRST)
                {
                  const double rm = s_x[(m - n0) % RX][tid + 1], x0 = O.alpha * rm; // x_0 = d_0 = alpha r
                  const double dn = c0 * x0 + c1 * O.dinvc * (rm - O.alpha * Sx);
                  if (O.d)
                    O.d[r] = dn;
                  O.out[(ov && r >= L.Np) ? 3 * r : r] = x0 + dn;
                }
              else
                {
                  const int so = (n - n0) % RO;
                  const double dn = c0 * s_op[2][so][tid] + c1 * s_op[1][so][tid] * (s_op[0][so][tid] - Sx);
                  O.d[r] = dn;
                  O.out[(ov && r >= L.Np) ? 3 * r : r] = s_x[(m - n0) % RX][tid + 1] + dn;
                }
            }
        }
    }
}

// ---- back substitution, Q1 closed form (pipeline of ring_back_kernel) -----------------------------------------------
template <int NT, int PF, bool FAST>
__device__ __forceinline__ void ring_back1_rows(const LatGeom &L, const RingOp &O, double (&s_ex)[2][2][NT + 2], int X,
                                                int Y0, int Y1)
{
  constexpr int HX = 2;
  static_assert(3 % PF == 0, "the prefetch ring turns with the row phase");
  const int tid = threadIdx.x;
  const RingCol<FAST> A(L, X);
  const q1::Cst C(L);
  const bool colok = FAST || (X >= 0 && X < L.fw);
  const bool outcol = tid >= HX && tid < NT - HX && (FAST || X < L.fw - O.lo);
  const double iml = L.ihx * L.ihy;
  double w1k[3] = {0, 0, 0}, w1m[3] = {0, 0, 0}, w2k[3] = {0, 0, 0}, w2m[3] = {0, 0, 0};
  const int n0 = Y0 - HX, nx1 = Y1 - 1 + HX, n1 = nx1 + 1;
  struct Pre
  {
    double x, o_r1, o_r2;
    bool chi_ok, wr;
  };
  auto fetch = [&](int n, Pre &q) {
    q.x = (colok && n <= nx1 && (FAST || n < L.fh)) ? O.z[A.v(n, 0)] : 0.0;
    const int mc = n - 1, m = n - 3; // rows of chi / phi produced in iteration n
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
  for (int nb = n0; nb <= n1; nb += 3)
    {
#pragma unroll
      for (int j = 0; j < 3; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          const Pre cur = pre[j % PF];
          fetch(n + PF, pre[j % PF]);
          const int par = (n - n0) & 1, mc = n - 1, m = n - 3;
          s_ex[par][0][tid + 1] = cur.x;
          s_ex[par][1][tid + 1] = a1;
          __syncthreads();
          double Kv, Mv;
          q1::xsum(&s_ex[par][0][tid], w1k[j], w1m[j]); // psi, row n
          q1::xsum(&s_ex[par][1][tid], w2k[j], w2m[j]); // chi, row n - 2
          q1::ysum<false>(C, j, w1k, w1m, Kv, Mv);
          a1 = cur.chi_ok ? iml * (cur.o_r1 - Kv) : 0.0;
          if (outcol && mc >= Y0 && mc < Y1)
            O.z[A.v(mc, 1)] = a1;
          if (cur.wr)
            {
              q1::ysum<false>(C, j, w2k, w2m, Kv, Mv);
              O.z[A.v(m, 2)] = iml * (cur.o_r2 - Kv);
            }
        }
    }
}

template <int NT, int PF>
static __global__ void __launch_bounds__(NT) ring_back1_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int HX = 2, OUTW = NT - 2 * HX;
  __shared__ double s_ex[2][2][NT + 2];
  const int tid = threadIdx.x;
  for (int i = tid; i < 4 * (NT + 2); i += NT)
    (&s_ex[0][0][0])[i] = 0.0;
  const int X0 = O.lo + blockIdx.x * OUTW - HX;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  __syncthreads();
  if (ring_cta_fast(L, X0, NT, Y0 - HX, Y1 - 1 + HX + 1))
    ring_back1_rows<NT, PF, true>(L, O, s_ex, X0 + tid, Y0, Y1);
  else
    ring_back1_rows<NT, PF, false>(L, O, s_ex, X0 + tid, Y0, Y1);
}

// ---- fused right-hand side kernel and back substitution with cp.async staging (interior CTAs), as ring_sweep1a_kernel --
template <int NT, int D>
static __global__ void __launch_bounds__(NT) ring_rhs1a_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int HX = 2, OUTW = NT - 2 * HX, RX = D + 2, RO = D + 1;
  constexpr int N_X = 2 * RX * (NT + 2), N_OP = 3 * RO * NT, N_EX = 2 * 2 * (NT + 2), N_SLOW = 2 * 4 * (NT + 2);
  constexpr int N_RAW = N_X + N_OP + N_EX > N_SLOW ? N_X + N_OP + N_EX : N_SLOW;
  __shared__ double raw[N_RAW];
  const int tid = threadIdx.x;
  for (int i = tid; i < N_RAW; i += NT)
    raw[i] = 0.0;
  const int X0 = O.lo + blockIdx.x * OUTW - HX;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  __syncthreads();
  const bool fast = ring_cta_fast(L, X0, NT, Y0 - HX, Y1 + 1); // else: general index for the addresses, same pipeline
  const int X = X0 + tid;
  const bool colok = fast || (X >= 0 && X < L.fw);
  double(&s_x)[2][RX][NT + 2] = *reinterpret_cast<double(*)[2][RX][NT + 2]>(raw); // r1, psi_n: rows n .. n + D
  double(&s_op)[3][RO][NT] = *reinterpret_cast<double(*)[3][RO][NT]>(raw + N_X);  // r2(n - 1), r0(n - 3), dinv(n - 3)
  double(&s_ex)[2][2][NT + 2] = *reinterpret_cast<double(*)[2][2][NT + 2]>(raw + N_X + N_OP); // t(n - 2), hand-over
  const q1::Cst C(L);
  const bool outcol = tid >= HX && tid < NT - HX && (fast || X < L.fw - O.lo);
  const double iml = L.ihx * L.ihy, jxw = 0.25 * L.hx * L.hy;
  const double cm0 = -O.dtth * (1.0 + O.eps), cm2 = -3.0 * O.dtth, cheb0 = O.cheb[0];
  const int w = L.w;
  const int64_t Np = L.Np;
  const int64_t base = (int64_t)(X - L.ex0) - (int64_t)L.ey0 * w; // owned node (X, Y): id = base + Y w
  auto sid = [&](int Y) { return fast ? base + (int64_t)Y * w : (int64_t)lat_id(L, X, Y); };             // scalar field
  auto vid = [&](int Y, int c) { return fast ? c * Np + base + (int64_t)Y * w : ring_vaddr(L, X, Y, c); }; // 3-vector
  const int n0 = Y0 - HX, nl1 = Y1 + 1, n1 = Y1 + 2;
  auto issue = [&](int n) {
    const int sx = (n - n0) % RX, so = (n - n0) % RO;
    if (n <= nl1)
      {
        if (colok)
          {
            cp_async8(&s_x[0][sx][tid + 1], O.in + vid(n, 1));
            cp_async8(&s_x[1][sx][tid + 1], O.psi + vid(n, 0));
          }
        else
          s_x[0][sx][tid + 1] = s_x[1][sx][tid + 1] = 0.0;
      }
    if (n - 1 >= n0 && n - 1 <= nl1 - 1)
      {
        if (colok)
          cp_async8(&s_op[0][so][tid], O.in + vid(n - 1, 2));
        else
          s_op[0][so][tid] = 0.0;
      }
    const int m = n - 3;
    if (outcol && m >= Y0 && m < Y1)
      {
        cp_async8(&s_op[1][so][tid], O.in + vid(m, 0));
        cp_async8(&s_op[2][so][tid], O.dinv + sid(m));
      }
    cp_async_commit();
  };
#pragma unroll
  for (int k = 0; k < D; ++k)
    issue(n0 + k);
  double w0k[3] = {0, 0, 0}, w0m[3] = {0, 0, 0}, w1k[3] = {0, 0, 0}, w1m[3] = {0, 0, 0};
  double tpub = 0, rpub = 0;
  q1::XS m_top = {0, 0, 0}, p_top = {0, 0, 0};
  double c01p = 0, c11p = 0, Pprev = 0, Vhold = 0;
  for (int nb = n0; nb <= n1; nb += 3)
    {
#pragma unroll
      for (int j = 0; j < 3; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          issue(n + D);
          const int par = (n - n0) & 1, sx = (n - n0) % RX, so = (n - n0) % RO;
          s_ex[par][0][tid + 1] = tpub;
          s_ex[par][1][tid + 1] = rpub;
          cp_async_wait<D>();
          __syncthreads();
          const bool have = n <= nl1; // rows beyond the last one needed count as zero
          const double m1L = have ? s_x[0][sx][tid] * iml : 0.0, m1n = have ? s_x[0][sx][tid + 1] * iml : 0.0;
          const double m1R = have ? s_x[0][sx][tid + 2] * iml : 0.0;
          const double psn = have ? s_x[1][sx][tid + 1] : 0.0, psR = have ? s_x[1][sx][tid + 2] : 0.0;
          {
            const double sm = m1L + m1R;
            w0k[j] = fma(2.0, m1n, -sm);
            w0m[j] = fma(4.0, m1n, sm);
          }
          double Kv, Mv;
          q1::ysum<false>(C, j, w0k, w0m, Kv, Mv); // K m1 at row n - 1
          const bool t_ok = n - 1 >= n0 && n - 1 <= nl1 - 1;
          tpub = t_ok ? iml * (s_op[0][so][tid] - Kv) : 0.0; // t(n - 1)
          q1::xsum(&s_ex[par][0][tid], w1k[j], w1m[j]);     // t, row n - 2
          q1::ysum<true>(C, j, w1k, w1m, Kv, Mv);          // K t, M t at row n - 3
          const double Dt = O.dtth * fma(-2.0, Mv, Kv);
          const q1::XS m_bot = q1::xstage(m1n, m1R), p_bot = q1::xstage(psn, psR);
          double mv[2][2], pv[2][2], dy[2], Aq[2][2], c[2][2];
          q1::values(m_top, m_bot, mv, dy);
          q1::values(p_top, p_bot, pv, dy);
#pragma unroll
          for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b)
              Aq[a][b] = fma(cm2 * pv[a][b], pv[a][b], cm0) * mv[a][b];
          q1::test(Aq, c);
          m_top = m_bot;
          p_top = p_bot;
          const double Pnew = c[0][0] + c01p;
          rpub = c[0][1] + c11p;
          c01p = c[1][0];
          c11p = c[1][1];
          const double Vuse = Vhold;
          Vhold = Pprev + s_ex[par][1][tid];
          Pprev = Pnew;
          const int m = n - 3;
          if (outcol && m >= Y0 && m < Y1)
            {
              const double rhs = s_op[1][so][tid] - jxw * Vuse - Dt;
              const double x0 = cheb0 * s_op[2][so][tid] * rhs;
              const int64_t r = sid(m);
              O.out[r] = rhs;
              O.d[r] = x0;
              O.out2[r] = x0;
            }
        }
    }
}

template <int NT, int D>
static __global__ void __launch_bounds__(NT) ring_back1a_kernel(LatGeom L, RingOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int HX = 2, OUTW = NT - 2 * HX, RX = D + 2, RO = D + 1;
  constexpr int N_X = RX * (NT + 2), N_OP = 2 * RO * NT, N_EX = 2 * (NT + 2), N_SLOW = 2 * 2 * (NT + 2);
  constexpr int N_RAW = N_X + N_OP + N_EX > N_SLOW ? N_X + N_OP + N_EX : N_SLOW;
  __shared__ double raw[N_RAW];
  const int tid = threadIdx.x;
  for (int i = tid; i < N_RAW; i += NT)
    raw[i] = 0.0;
  const int X0 = O.lo + blockIdx.x * OUTW - HX;
  const int Y0 = O.lo + blockIdx.y * O.rs, Y1 = min(Y0 + O.rs, L.fh - O.lo);
  __syncthreads();
  const bool fast = ring_cta_fast(L, X0, NT, Y0 - HX, Y1 - 1 + HX + 1); // else: general index for the addresses only
  const int X = X0 + tid;
  const bool colok = fast || (X >= 0 && X < L.fw);
  double(&s_x)[RX][NT + 2] = *reinterpret_cast<double(*)[RX][NT + 2]>(raw);       // psi (component 0 of z): rows n .. n + D
  double(&s_op)[2][RO][NT] = *reinterpret_cast<double(*)[2][RO][NT]>(raw + N_X);  // r1(n - 1), r2(n - 3)
  double(&s_ex)[2][NT + 2] = *reinterpret_cast<double(*)[2][NT + 2]>(raw + N_X + N_OP); // chi(n - 2)
  const q1::Cst C(L);
  const bool outcol = tid >= HX && tid < NT - HX && (fast || X < L.fw - O.lo);
  const double iml = L.ihx * L.ihy;
  const int w = L.w;
  const int64_t Np = L.Np;
  const int64_t base = (int64_t)(X - L.ex0) - (int64_t)L.ey0 * w;
  auto vid = [&](int Y, int c) { return fast ? c * Np + base + (int64_t)Y * w : ring_vaddr(L, X, Y, c); }; // 3-vector
  const int n0 = Y0 - HX, nx1 = Y1 - 1 + HX, n1 = nx1 + 1;
  auto issue = [&](int n) {
    const int sx = (n - n0) % RX, so = (n - n0) % RO;
    if (n <= nx1)
      {
        if (colok)
          cp_async8(&s_x[sx][tid + 1], O.z + vid(n, 0));
        else
          s_x[sx][tid + 1] = 0.0;
      }
    if (n - 1 >= n0 && n <= n1)
      {
        if (colok)
          cp_async8(&s_op[0][so][tid], O.in + vid(n - 1, 1));
        else
          s_op[0][so][tid] = 0.0;
      }
    const int m = n - 3;
    if (outcol && m >= Y0 && m < Y1)
      cp_async8(&s_op[1][so][tid], O.in + vid(m, 2));
    cp_async_commit();
  };
#pragma unroll
  for (int k = 0; k < D; ++k)
    issue(n0 + k);
  double w1k[3] = {0, 0, 0}, w1m[3] = {0, 0, 0}, w2k[3] = {0, 0, 0}, w2m[3] = {0, 0, 0};
  double a1 = 0;
  for (int nb = n0; nb <= n1; nb += 3)
    {
#pragma unroll
      for (int j = 0; j < 3; ++j)
        {
          const int n = nb + j;
          if (n > n1)
            break;
          issue(n + D);
          const int par = (n - n0) & 1, sx = (n - n0) % RX, so = (n - n0) % RO, mc = n - 1, m = n - 3;
          s_ex[par][tid + 1] = a1;
          cp_async_wait<D>();
          __syncthreads();
          double Kv, Mv;
          if (n <= nx1)
            q1::xsum(&s_x[sx][tid], w1k[j], w1m[j]); // psi, row n
          else
            w1k[j] = w1m[j] = 0.0;
          q1::xsum(&s_ex[par][tid], w2k[j], w2m[j]); // chi, row n - 2
          q1::ysum<false>(C, j, w1k, w1m, Kv, Mv);
          a1 = (mc >= n0) ? iml * (s_op[0][so][tid] - Kv) : 0.0;
          if (outcol && mc >= Y0 && mc < Y1)
            O.z[vid(mc, 1)] = a1;
          if (outcol && m >= Y0 && m < Y1)
            {
              q1::ysum<false>(C, j, w2k, w2m, Kv, Mv);
              O.z[vid(m, 2)] = iml * (s_op[1][so][tid] - Kv);
            }
        }
    }
}

// ---- J apply on tiles of cells (Q1) ---------------------------------------------------------------------------------
// y = J z on the owned nodes.  Tiling as lat_kernel<1, LAT_J> (one thread per cell, the nodes of the inner
// (CX-1) x (CY-1) cells are the tile's own), but only B z_psi + C z_chi goes through the Gauss points (q1:: algebra);
// K z_psi + M z_chi, K z_chi + M z_phi and D z_phi are 9-point stencils on the staged nodal values.
// With half the arithmetic of the general kernel the load latency of a tile is no longer hidden by the arithmetic of
// the other resident CTAs (measured: 522 -> 510 us), so a CTA walks over NTILE tiles of a tile row and stages the
// nodal values of the next tile with cp.async (LDGSTS, 8 bytes: a tile row starts at any node) while it works on the
// current one; two shared-memory buffers.  NTILE = 1: plain version, no pipeline.
// ROWMAP: the node phase maps thread (cx, cy) of the cell phase to tile node (cx + 1, cy + 1) (the last thread column and
// row idle) instead of packing the (CX-1) x (CY-1) nodes into the first threads: a warp then reads one shared-memory row
// with consecutive lanes (no bank conflict where the packed rows of CX - 1 wrap inside a warp).
// ROW0: only the first block row, out3[r] = B z_psi + C z_chi + D z_phi (the operator of the reduced solve, pfc_reduced.cuh:
// its input makes the other two rows vanish); out3 is then a scalar field.
template <int CX, int CY, int NTILE, int MINB, bool ROWMAP = false, bool ROW0 = false>
static __global__ void __launch_bounds__(CX *CY, MINB) lat_j1_kernel(LatGeom L, LatOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int NT = CX * CY, NF = 5, TNX = CX + 1, TNY = CY + 1, TN = TNX * TNY, NBUF = NTILE > 1 ? 2 : 1;
  constexpr int ITER = (TN + NT - 1) / NT, ONX = CX - 1, ONY = CY - 1;
  __shared__ double s_in[NBUF][NF][TN]; // z_psi, z_chi, z_phi, psi_n, chi_n
  __shared__ double s_c[4][NT];         // corner contributions [iy * 2 + ix] of every cell
  const int tid = threadIdx.x;
  const int ntx = (L.ncx + CX - 2) / (CX - 1);
  const int bx0 = blockIdx.x * NTILE, nt = min(NTILE, ntx - bx0);
  const int ay0 = blockIdx.y * (CY - 1), ny0 = ay0 - 1; // first inner cell row of the tile row; padded row of tile node row 0
  const int64_t Np = L.Np;
  const double dtth = O.dt * O.theta;
  const bool rows_owned = ny0 >= L.ey0 && ny0 + TNY <= L.ey0 + L.h;

  // stage the nodal values of tile bx into buffer buf (asynchronously; zeros outside the padded rectangle)
  auto stage = [&](int buf, int bx) {
    const int nx0 = bx * (CX - 1) - 1;
    const bool fast = rows_owned && nx0 >= L.ex0 && nx0 + TNX <= L.ex0 + L.w; // all nodes owned: arithmetic ids
#pragma unroll
    for (int it = 0; it < ITER; ++it)
      {
        const int idx = tid + it * NT;
        if (idx >= TN)
          break;
        const int ex = nx0 + idx % TNX, ey = ny0 + idx / TNX;
        if (fast)
          {
            const int64_t u = (int64_t)(ey - L.ey0) * L.w + (ex - L.ex0);
            cp_async8(&s_in[buf][0][idx], O.in3a + u);
            cp_async8(&s_in[buf][1][idx], O.in3a + Np + u);
            cp_async8(&s_in[buf][2][idx], O.in3a + 2 * Np + u);
            cp_async8(&s_in[buf][3][idx], O.in3b + u);
            cp_async8(&s_in[buf][4][idx], O.in3b + Np + u);
          }
        else if (ex >= 0 && ex < L.fw && ey >= 0 && ey < L.fh)
          {
            const int32_t u = lat_id(L, ex, ey);
            cp_async8(&s_in[buf][0][idx], O.in3a + vaddr(u, 0, Np));
            cp_async8(&s_in[buf][1][idx], O.in3a + vaddr(u, 1, Np));
            cp_async8(&s_in[buf][2][idx], O.in3a + vaddr(u, 2, Np));
            cp_async8(&s_in[buf][3][idx], O.in3b + vaddr(u, 0, Np));
            cp_async8(&s_in[buf][4][idx], O.in3b + vaddr(u, 1, Np));
          }
        else
          {
#pragma unroll
            for (int k = 0; k < NF; ++k)
              s_in[buf][k][idx] = 0.0;
          }
      }
    cp_async_commit();
  };

  stage(0, bx0);
  for (int k = 0; k < nt; ++k)
    {
      const int buf = NBUF == 2 ? (k & 1) : 0;
      if (NBUF == 2 && k + 1 < nt)
        {
          stage(buf ^ 1, bx0 + k + 1);
          cp_async_wait<1>();
        }
      else
        cp_async_wait<0>();
      __syncthreads();
      const int ax0 = (bx0 + k) * (CX - 1), nx0 = ax0 - 1;
      const double(*S)[TN] = s_in[buf];
      { // ---- one cell per thread: B z_psi + C z_chi at the 2 x 2 Gauss points (mpi.cpp:421-433) ----
        const int cx = tid % CX, cy = tid / CX;
        const int kx = ax0 - 1 + cx, ky = ay0 - 1 + cy;
        const bool cell_ok = kx >= 0 && kx < L.ncx && ky >= 0 && ky < L.ncy;
        const int o = cy * TNX + cx;
        auto xs = [&](int fld, int row) { return q1::xstage(S[fld][o + row * TNX], S[fld][o + row * TNX + 1]); };
        double psi[2][2], chi[2][2], v0[2][2], v1[2][2], pdy[2], pdx[2], zdy[2], zdx[2], tmp[2];
        {
          const q1::XS t = xs(3, 0), b = xs(3, 1);
          q1::values(t, b, psi, pdy);
          q1::xderiv(t, b, pdx);
        }
        q1::values(xs(4, 0), xs(4, 1), chi, tmp);
        {
          const q1::XS t = xs(0, 0), b = xs(0, 1);
          q1::values(t, b, v0, zdy);
          q1::xderiv(t, b, zdx);
        }
        q1::values(xs(1, 0), xs(1, 1), v1, tmp);
        // physical gradients: d/dx at Gauss row qy = dx[qy] / hx, d/dy at Gauss column qx = dy[qx] / hy
        const double ihx2 = L.ihx * L.ihx, ihy2 = L.ihy * L.ihy;
        const double cm0 = -dtth * (1.0 + O.eps);
        double A[2][2], c[2][2];
#pragma unroll
        for (int qy = 0; qy < 2; ++qy)
#pragma unroll
          for (int qx = 0; qx < 2; ++qx)
            {
              const double p = psi[qy][qx];
              const double g2 = fma(pdx[qy] * pdx[qy], ihx2, pdy[qx] * pdy[qx] * ihy2); // |grad psi|^2
              const double gz = fma(pdx[qy] * zdx[qy], ihx2, pdy[qx] * zdy[qx] * ihy2); // grad psi . grad z_psi
              const double bm = fma(-6.0 * dtth, fma(p, chi[qy][qx], g2), 1.0);
              const double cm = fma(-3.0 * dtth * p, p, cm0);
              A[qy][qx] = fma(bm, v0[qy][qx], fma(-12.0 * dtth * p, gz, cm * v1[qy][qx]));
            }
        q1::test(A, c);
        s_c[0][tid] = cell_ok ? c[0][0] : 0.0;
        s_c[1][tid] = cell_ok ? c[0][1] : 0.0;
        s_c[2][tid] = cell_ok ? c[1][0] : 0.0;
        s_c[3][tid] = cell_ok ? c[1][1] : 0.0;
      }
      __syncthreads();
      // ---- nodes of the inner cells: the four corner contributions in a fixed order + the constant 9-point stencils ----
      if (ROWMAP ? (tid % CX < ONX && tid / CX < ONY) : tid < ONX * ONY)
        {
          const int i = 1 + (ROWMAP ? tid % CX : tid % ONX), j = 1 + (ROWMAP ? tid / CX : tid / ONX); // tile node
          if (lat_in_window(L, -1, nx0 + i, ny0 + j))
            {
              const double var =
                (s_c[3][(j - 1) * CX + i - 1] + s_c[2][(j - 1) * CX + i]) + (s_c[1][j * CX + i - 1] + s_c[0][j * CX + i]);
              const q1::Cst C(L);
              double K0 = 0, K1 = 0, K2, M1 = 0, M2, Mnone;
              double wk[3], wm[3];
              if (!ROW0)
                {
#pragma unroll
                  for (int r = 0; r < 3; ++r)
                    q1::xsum(&S[0][(j - 1 + r) * TNX + i - 1], wk[r], wm[r]);
                  q1::ysum<false>(C, 2, wk, wm, K0, Mnone); // phase 2: slots (0, 1, 2) = (oldest, centre, newest)
#pragma unroll
                  for (int r = 0; r < 3; ++r)
                    q1::xsum(&S[1][(j - 1 + r) * TNX + i - 1], wk[r], wm[r]);
                  q1::ysum<true>(C, 2, wk, wm, K1, M1);
                }
#pragma unroll
              for (int r = 0; r < 3; ++r)
                q1::xsum(&S[2][(j - 1 + r) * TNX + i - 1], wk[r], wm[r]);
              q1::ysum<true>(C, 2, wk, wm, K2, M2);
              const int64_t r = (int64_t)(ny0 + j - L.ey0) * L.w + (nx0 + i - L.ex0); // owned: ly w + lx
              O.out3[r] = fma(0.25 * L.hx * L.hy, var, dtth * fma(-2.0, M2, K2));
              if (!ROW0)
                {
                  O.out3[Np + r] = K0 + M1;
                  O.out3[2 * Np + r] = K1 + M2;
                }
            }
        }
      if (k + 1 < nt)
        __syncthreads(); // s_c and this buffer are rewritten by the next round
    }
}
} // namespace pfc
