// pfc_lattice.cuh — matrix-free, sum-factorised cell kernels on the uniform periodic tensor lattice
// (north_star item 2; the reference has no such path, it multiplies with assembled Trilinos blocks).
//
// The integrand is the one of assemble_system (mpi.cpp:388-444), applied to vectors instead of being stored:
//   J = [[B, C, D], [K, M, 0], [0, K, M]],  B = M - 6 dt th (<N, psi chi N> + <N, |grad psi|^2 N> + 2 <N, psi grad psi . grad N>),
//   C = -dt th ((1 + eps) M + 3 <N, psi^2 N>),  D = dt th (K - 2 M)                       (tex:399-466)
// Layout (pfc_plan.h, Lattice): the owned free nodes are a w x h rectangle numbered lexicographically, so a
// tile of cells reads its nodes as contiguous row segments; only the ring outside the rectangle (ghosts,
// periodic images) goes through a lookup table.  No per-cell index tables, no colouring, no atomics:
// one CTA evaluates CX x CY cells (one thread per cell, everything in registers, 1-D tables are compile-time
// constants, interpolation and testing are sum-factorised), parks the (p+1)^2 local results per cell in shared
// memory, and the nodes of the inner (CX-1) x (CY-1) cells gather their 1, 2 or 4 contributions in a fixed
// order (deterministic) and run the fused epilogue.  The outer row/column of cells is recomputed by the
// neighbouring CTA (overlap (CX CY) / ((CX-1)(CY-1))).
// Algorithmic bytes per application (N nodes): J apply: x 24 N + y 24 N + (psi, chi) 16 N = 64 N.
#pragma once
#include <utility>

#include "pfc_krylov.cuh"
#include "pfc_lattice_args.h"

namespace pfc
{
// ---- compile-time 1-D tables: FE_Q(p) on equidistant nodes at the QGauss(p+1) points of [0, 1] ----------
template <int P>
struct Tab1;
template <>
struct Tab1<1>
{
  static __host__ __device__ constexpr double g(int q) { return q == 0 ? 0.5 - 0.5 * 0.57735026918962576451 : 0.5 + 0.5 * 0.57735026918962576451; }
  static __host__ __device__ constexpr double S(int q, int j) { return j == 0 ? 1.0 - g(q) : g(q); }
  static __host__ __device__ constexpr double D(int, int j) { return j == 0 ? -1.0 : 1.0; }
  static __host__ __device__ constexpr double W(int) { return 0.5; }
};
template <>
struct Tab1<2>
{
  static __host__ __device__ constexpr double g(int q)
  {
    return q == 0 ? 0.5 - 0.5 * 0.77459666924148337704 : (q == 1 ? 0.5 : 0.5 + 0.5 * 0.77459666924148337704);
  }
  static __host__ __device__ constexpr double S(int q, int j)
  {
    return j == 0 ? 2.0 * (g(q) - 0.5) * (g(q) - 1.0) : (j == 1 ? 4.0 * g(q) * (1.0 - g(q)) : 2.0 * g(q) * (g(q) - 0.5));
  }
  static __host__ __device__ constexpr double D(int q, int j) { return j == 0 ? 4.0 * g(q) - 3.0 : (j == 1 ? 4.0 - 8.0 * g(q) : 4.0 * g(q) - 1.0); }
  static __host__ __device__ constexpr double W(int q) { return q == 1 ? 8.0 / 18.0 : 5.0 / 18.0; }
};

// compile-time loop: f(std::integral_constant<int, 0>) ... f(std::integral_constant<int, N - 1>)
template <class F, int... I>
__device__ __forceinline__ void sfor_impl(F &f, std::integer_sequence<int, I...>)
{
  (f(std::integral_constant<int, I>{}), ...);
}
template <int N, class F>
__device__ __forceinline__ void sfor(F &&f)
{
  sfor_impl(f, std::make_integer_sequence<int, N>{});
}
#define PFC_IC(name, X) constexpr int name = decltype(X)::value

// values (V) and/or physical gradients (G) of a cell's nodal values u[jy][jx] at the (p+1)^2 Gauss points,
// q arrays indexed [qy][qx]; sum factorisation: x direction first, then y
template <int P, bool V, bool G>
__device__ __forceinline__ void lat_interp(const double (&u)[P + 1][P + 1], double ihx, double ihy,
                                           double (&v)[P + 1][P + 1], double (&gx)[P + 1][P + 1],
                                           double (&gy)[P + 1][P + 1])
{
  constexpr int n = P + 1;
  double a[n][n], b[n][n]; // [jy][qx]: x-interpolated values / x-derivatives
  sfor<n>([&](auto JY) {
    PFC_IC(jy, JY);
    sfor<n>([&](auto QX) {
      PFC_IC(qx, QX);
      double sa = 0, sb = 0;
      sfor<n>([&](auto JX) {
        PFC_IC(jx, JX);
        constexpr double s = Tab1<P>::S(qx, jx), d = Tab1<P>::D(qx, jx);
        if constexpr (s != 0.0)
          sa = fma(s, u[jy][jx], sa);
        if constexpr (G && d != 0.0)
          sb = fma(d, u[jy][jx], sb);
      });
      a[jy][qx] = sa;
      b[jy][qx] = sb;
    });
  });
  sfor<n>([&](auto QY) {
    PFC_IC(qy, QY);
    sfor<n>([&](auto QX) {
      PFC_IC(qx, QX);
      double sv = 0, sx = 0, sy = 0;
      sfor<n>([&](auto JY) {
        PFC_IC(jy, JY);
        constexpr double s = Tab1<P>::S(qy, jy), d = Tab1<P>::D(qy, jy);
        if constexpr (s != 0.0)
          {
            if constexpr (V)
              sv = fma(s, a[jy][qx], sv);
            if constexpr (G)
              sx = fma(s, b[jy][qx], sx);
          }
        if constexpr (G && d != 0.0)
          sy = fma(d, a[jy][qx], sy);
      });
      if constexpr (V)
        v[qy][qx] = sv;
      if constexpr (G)
        {
          gx[qy][qx] = sx * ihx;
          gy[qy][qx] = sy * ihy;
        }
    });
  });
}

// y[iy][ix] += sum_q N_i(q) A(q) + dN_i/dx(q) Gx(q) + dN_i/dy(q) Gy(q)     (A, Gx, Gy already carry JxW)
template <int P, bool HA, bool HG>
__device__ __forceinline__ void lat_test(const double (&A)[P + 1][P + 1], const double (&Gx)[P + 1][P + 1],
                                         const double (&Gy)[P + 1][P + 1], double ihx, double ihy,
                                         double (&y)[P + 1][P + 1])
{
  constexpr int n = P + 1;
  double t1[n][n], t2[n][n]; // [qy][ix]: x direction tested
  sfor<n>([&](auto QY) {
    PFC_IC(qy, QY);
    sfor<n>([&](auto IX) {
      PFC_IC(ix, IX);
      double s1 = 0, s2 = 0;
      sfor<n>([&](auto QX) {
        PFC_IC(qx, QX);
        constexpr double s = Tab1<P>::S(qx, ix), d = Tab1<P>::D(qx, ix);
        if constexpr (s != 0.0)
          {
            if constexpr (HA)
              s1 = fma(s, A[qy][qx], s1);
            if constexpr (HG)
              s2 = fma(s, Gy[qy][qx], s2);
          }
        if constexpr (HG && d != 0.0)
          s1 = fma(d * ihx, Gx[qy][qx], s1);
      });
      t1[qy][ix] = s1;
      t2[qy][ix] = s2;
    });
  });
  sfor<n>([&](auto IY) {
    PFC_IC(iy, IY);
    sfor<n>([&](auto IX) {
      PFC_IC(ix, IX);
      double s0 = y[iy][ix];
      sfor<n>([&](auto QY) {
        PFC_IC(qy, QY);
        constexpr double s = Tab1<P>::S(qy, iy), d = Tab1<P>::D(qy, iy);
        if constexpr (s != 0.0)
          s0 = fma(s, t1[qy][ix], s0);
        if constexpr (HG && d != 0.0)
          s0 = fma(d * ihy, t2[qy][ix], s0);
      });
      y[iy][ix] = s0;
    });
  });
}

// internal id of padded position (ex, ey): owned free node, or ghost Np + ring index (ring_index, pfc_plan.h)
__device__ __forceinline__ int32_t lat_id(const LatGeom &L, int ex, int ey)
{
  const int lx = ex - L.ex0, ly = ey - L.ey0;
  if ((unsigned)lx < (unsigned)L.w && (unsigned)ly < (unsigned)L.h)
    return ly * L.w + lx;
  int r;
  if (ey < L.ey0)
    r = ey * L.fw + ex;
  else if (ey >= L.ey0 + L.h)
    r = (ey - L.h) * L.fw + ex;
  else
    r = (L.fh - L.h) * L.fw + ly * (L.fw - L.w) + (ex < L.ex0 ? ex : ex - L.w);
  return (int32_t)L.Np + r;
}

// output window of a cell kernel: the owned free nodes (lo < 0), or all padded positions at least lo nodes inside
__device__ __forceinline__ bool lat_in_window(const LatGeom &L, int lo, int X, int Y)
{
  if (lo < 0)
    return (unsigned)(X - L.ex0) < (unsigned)L.w && (unsigned)(Y - L.ey0) < (unsigned)L.h;
  return X >= lo && X < L.fw - lo && Y >= lo && Y < L.fh - lo;
}

template <int MODE>
struct LatShape
{
  static constexpr int NF = MODE == LAT_J ? 5 : (MODE == LAT_S ? 5 : (MODE == LAT_K ? 1 : (MODE == LAT_T ? 2 : 6)));
  static constexpr int NOUT = MODE == LAT_J ? 3 : ((MODE == LAT_R || MODE == LAT_T) ? 4 : 1);
};

template <int P, int MODE, int CX, int CY>
constexpr size_t lat_smem_bytes()
{
  return sizeof(double) * ((size_t)LatShape<MODE>::NF * (P * CX + 1) * (P * CY + 1) +
                           (size_t)LatShape<MODE>::NOUT * (P + 1) * (P + 1) * CX * CY);
}

template <int P, int MODE, int CX, int CY>
static __global__ void __launch_bounds__(CX *CY) lat_kernel(LatGeom L, LatOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int n = P + 1, NT = CX * CY, NF = LatShape<MODE>::NF, NOUT = LatShape<MODE>::NOUT;
  constexpr int TNX = P * CX + 1, TNY = P * CY + 1, TN = TNX * TNY;
  extern __shared__ double smem[];
  double *s_in = smem;             // [NF][TNY][TNX]
  double *s_out = smem + NF * TN;  // [NOUT][n*n][NT]
  const int tid = threadIdx.x;
  const int ax0 = blockIdx.x * (CX - 1), ay0 = blockIdx.y * (CY - 1); // first inner cell of the tile
  const int nx0 = P * (ax0 - 1), ny0 = P * (ay0 - 1);                 // frame position of tile node (0, 0)
  const int64_t Np = L.Np;

  // ---- stage the tile's nodal values: all node ids first, then all loads in flight, then the stores ----
  {
    constexpr int ITER = (TN + NT - 1) / NT;
    int32_t uid[ITER];
#pragma unroll
    for (int it = 0; it < ITER; ++it)
      {
        const int idx = tid + it * NT;
        const int i = idx % TNX, j = idx / TNX;
        const int ex = nx0 + i, ey = ny0 + j;
        uid[it] = (idx < TN && ex >= 0 && ex < L.fw && ey >= 0 && ey < L.fh) ? lat_id(L, ex, ey) : -1;
      }
    double f[ITER][NF];
#pragma unroll
    for (int it = 0; it < ITER; ++it)
      {
        const int32_t u = uid[it];
#pragma unroll
        for (int k = 0; k < NF; ++k)
          f[it][k] = 0;
        if (u >= 0)
          {
            if constexpr (MODE == LAT_J)
              {
                f[it][0] = O.in3a[vaddr(u, 0, Np)];
                f[it][1] = O.in3a[vaddr(u, 1, Np)];
                f[it][2] = O.in3a[vaddr(u, 2, Np)];
                f[it][3] = O.in3b[vaddr(u, 0, Np)];
                f[it][4] = O.in3b[vaddr(u, 1, Np)];
              }
            else if constexpr (MODE == LAT_S)
              {
                f[it][0] = O.sB != 0.0 ? O.s0[u] : 0.0;
                f[it][1] = O.s1[u];
                f[it][2] = O.s2[u];
                f[it][3] = O.in3b[vaddr(u, 0, Np)];
                f[it][4] = O.in3b[vaddr(u, 1, Np)];
              }
            else if constexpr (MODE == LAT_K)
              f[it][0] = O.s0[u];
            else if constexpr (MODE == LAT_T)
              {
                f[it][0] = O.in3a[vaddr(u, 0, Np)];
                f[it][1] = O.in3a[vaddr(u, 1, Np)];
              }
            else
              {
#pragma unroll
                for (int c = 0; c < 3; ++c)
                  {
                    f[it][c] = O.in3a[vaddr(u, c, Np)];
                    f[it][3 + c] = O.in3b[vaddr(u, c, Np)];
                  }
              }
          }
      }
#pragma unroll
    for (int it = 0; it < ITER; ++it)
      {
        const int idx = tid + it * NT;
        if (idx < TN)
          {
#pragma unroll
            for (int k = 0; k < NF; ++k)
              s_in[k * TN + idx] = f[it][k];
          }
      }
  }
  __syncthreads();

  // ---- one cell per thread ----
  const int cx = tid % CX, cy = tid / CX;
  const int kx = ax0 - 1 + cx, ky = ay0 - 1 + cy;
  const bool cell_ok = kx >= 0 && kx < L.ncx && ky >= 0 && ky < L.ncy;
  const double hx = L.hx, hy = L.hy, ihx = L.ihx, ihy = L.ihy, vol = hx * hy;
  const double *cell_in = s_in + (P * cy) * TNX + P * cx;
  auto load_u = [&](int f, double(&u)[n][n]) {
#pragma unroll
    for (int jy = 0; jy < n; ++jy)
#pragma unroll
      for (int jx = 0; jx < n; ++jx)
        u[jy][jx] = cell_in[f * TN + jy * TNX + jx];
  };
  auto store_y = [&](int o, const double(&y)[n][n]) {
#pragma unroll
    for (int jy = 0; jy < n; ++jy)
#pragma unroll
      for (int jx = 0; jx < n; ++jx)
        s_out[(o * n * n + jy * n + jx) * NT + tid] = cell_ok ? y[jy][jx] : 0.0;
  };
  auto jxw = [&](double(&q)[n][n], double s) { // q *= s * JxW
    sfor<n>([&](auto QY) {
      PFC_IC(qy, QY);
      sfor<n>([&](auto QX) {
        PFC_IC(qx, QX);
        q[qy][qx] *= s * (Tab1<P>::W(qx) * Tab1<P>::W(qy)) * vol;
      });
    });
  };
  double u[n][n], v[n][n], gx[n][n], gy[n][n];
  const double dtth = O.dt * O.theta;

  if constexpr (MODE == LAT_K)
    {
      double y[n][n] = {};
      load_u(0, u);
      lat_interp<P, false, true>(u, ihx, ihy, v, gx, gy);
      jxw(gx, 1.0);
      jxw(gy, 1.0);
      lat_test<P, false, true>(v, gx, gy, ihx, ihy, y);
      store_y(0, y);
    }
  else if constexpr (MODE == LAT_J || MODE == LAT_S)
    {
      // linearisation coefficients at the Gauss points (mpi.cpp:421-433)
      double bm[n][n], bx[n][n], by[n][n], cm[n][n], A0[n][n];
      {
        double chi[n][n];
        load_u(4, u);
        lat_interp<P, true, false>(u, ihx, ihy, chi, gx, gy);
        load_u(3, u);
        lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy);
#pragma unroll
        for (int a = 0; a < n; ++a)
#pragma unroll
          for (int b = 0; b < n; ++b)
            {
              const double psi = v[a][b];
              bm[a][b] = 1.0 - dtth * (6.0 * psi * chi[a][b] + 6.0 * (gx[a][b] * gx[a][b] + gy[a][b] * gy[a][b]));
              bx[a][b] = -12.0 * dtth * psi * gx[a][b];
              by[a][b] = -12.0 * dtth * psi * gy[a][b];
              cm[a][b] = -dtth * ((1.0 + O.eps) + 3.0 * psi * psi);
            }
      }
      if constexpr (MODE == LAT_J)
        {
          double y1[n][n] = {}, y2[n][n] = {}, y0[n][n] = {};
          // X0 = x_psi
          load_u(0, u);
          lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy);
#pragma unroll
          for (int a = 0; a < n; ++a)
#pragma unroll
            for (int b = 0; b < n; ++b)
              A0[a][b] = bm[a][b] * v[a][b] + bx[a][b] * gx[a][b] + by[a][b] * gy[a][b];
          jxw(gx, 1.0);
          jxw(gy, 1.0);
          lat_test<P, false, true>(v, gx, gy, ihx, ihy, y1); // K x_psi
          // X1 = x_chi
          load_u(1, u);
          lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy);
#pragma unroll
          for (int a = 0; a < n; ++a)
#pragma unroll
            for (int b = 0; b < n; ++b)
              A0[a][b] = fma(cm[a][b], v[a][b], A0[a][b]);
          jxw(v, 1.0);
          jxw(gx, 1.0);
          jxw(gy, 1.0);
          lat_test<P, true, false>(v, gx, gy, ihx, ihy, y1); // + M x_chi
          store_y(1, y1);
          lat_test<P, false, true>(v, gx, gy, ihx, ihy, y2); // K x_chi
          // X2 = x_phi
          load_u(2, u);
          lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy);
#pragma unroll
          for (int a = 0; a < n; ++a)
#pragma unroll
            for (int b = 0; b < n; ++b)
              A0[a][b] = fma(-2.0 * dtth, v[a][b], A0[a][b]);
          jxw(v, 1.0);
          lat_test<P, true, false>(v, gx, gy, ihx, ihy, y2); // + M x_phi
          store_y(2, y2);
          jxw(A0, 1.0);
          jxw(gx, dtth);
          jxw(gy, dtth);
          lat_test<P, true, true>(A0, gx, gy, ihx, ihy, y0); // B x_psi + C x_chi + D x_phi
          store_y(0, y0);
        }
      else
        {
          double y0[n][n] = {};
#pragma unroll
          for (int a = 0; a < n; ++a)
#pragma unroll
            for (int b = 0; b < n; ++b)
              A0[a][b] = 0.0;
          if (O.sB != 0.0)
            {
              load_u(0, u);
              lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy);
#pragma unroll
              for (int a = 0; a < n; ++a)
#pragma unroll
                for (int b = 0; b < n; ++b)
                  A0[a][b] = O.sB * (bm[a][b] * v[a][b] + bx[a][b] * gx[a][b] + by[a][b] * gy[a][b]);
            }
          load_u(1, u);
          lat_interp<P, true, false>(u, ihx, ihy, v, gx, gy);
#pragma unroll
          for (int a = 0; a < n; ++a)
#pragma unroll
            for (int b = 0; b < n; ++b)
              A0[a][b] = fma(O.sC * cm[a][b], v[a][b], A0[a][b]);
          load_u(2, u);
          lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy);
#pragma unroll
          for (int a = 0; a < n; ++a)
#pragma unroll
            for (int b = 0; b < n; ++b)
              A0[a][b] = fma(-2.0 * dtth * O.sD, v[a][b], A0[a][b]);
          jxw(A0, 1.0);
          jxw(gx, dtth * O.sD);
          jxw(gy, dtth * O.sD);
          lat_test<P, true, true>(A0, gx, gy, ihx, ihy, y0);
          store_y(0, y0);
        }
    }
  else if constexpr (MODE == LAT_T)
    { // right-hand side of the stress projection (stress_calculator_mpi.cpp:283-300).  The shape function of tensor
      // component (a,b) on scalar shape N is tau = N e_a (x) e_b: tau : X = N X_ab, (div tau)_a = d_b N,
      // tr tau = delta_ab N, grad tr tau = delta_ab grad N.  With P = grad psi, G = grad (psi + chi), S = psi + chi:
      //   rhs_ab = int N (G_a P_b + P_a G_b) + d_b N S P_a + delta_ab [ N (F - |P|^2) - psi grad N . P ]
      //   F = (chi^2 + (1 + eps) psi^2 + psi^4 / 2) / 2
      double psi[n][n], px[n][n], py[n][n], A[n][n], cx_[n][n], cy_[n][n];
      load_u(0, u);
      lat_interp<P, true, true>(u, ihx, ihy, psi, px, py);
      load_u(1, u);
      lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy); // chi, then G = grad psi + grad chi
      double D[n][n], S[n][n];                            // D = F - |P|^2
#pragma unroll
      for (int a = 0; a < n; ++a)
#pragma unroll
        for (int b = 0; b < n; ++b)
          {
            const double p_ = psi[a][b], c_ = v[a][b], p2 = p_ * p_;
            gx[a][b] += px[a][b];
            gy[a][b] += py[a][b];
            S[a][b] = p_ + c_;
            D[a][b] = 0.5 * (c_ * c_ + (1.0 + O.eps) * p2 + 0.5 * p2 * p2) - (px[a][b] * px[a][b] + py[a][b] * py[a][b]);
          }
      sfor<4>([&](auto C) {
        PFC_IC(c, C);
        constexpr int ta = c / 2, tb = c % 2;
        double y[n][n] = {};
#pragma unroll
        for (int a = 0; a < n; ++a)
#pragma unroll
          for (int b = 0; b < n; ++b)
            {
              const double Pa = ta == 0 ? px[a][b] : py[a][b], Pb = tb == 0 ? px[a][b] : py[a][b];
              const double Ga = ta == 0 ? gx[a][b] : gy[a][b], Gb = tb == 0 ? gx[a][b] : gy[a][b];
              double av = Ga * Pb + Pa * Gb, dx = 0.0, dy = 0.0;
              if (tb == 0)
                dx = S[a][b] * Pa;
              else
                dy = S[a][b] * Pa;
              if (ta == tb)
                {
                  av += D[a][b];
                  dx -= psi[a][b] * px[a][b];
                  dy -= psi[a][b] * py[a][b];
                }
              A[a][b] = av;
              cx_[a][b] = dx;
              cy_[a][b] = dy;
            }
        jxw(A, 1.0);
        jxw(cx_, 1.0);
        jxw(cy_, 1.0);
        lat_test<P, true, true>(A, cx_, cy_, ihx, ihy, y);
        store_y(c, y);
      });
    }
  else
    { // LAT_R: -R and diag(B); fields 0..2 = (psi, chi, phi)_n, 3..5 = (psi, chi, phi)_{n-1}
      const double dt1 = O.dt * (1.0 - O.theta);
      double psi[n][n], pgx[n][n], pgy[n][n], cf0[n][n], cfx[n][n], cfy[n][n], tx[n][n], ty[n][n];
      double y1[n][n] = {}, y2[n][n] = {}, y0[n][n] = {};
      load_u(0, u);
      lat_interp<P, true, true>(u, ihx, ihy, psi, pgx, pgy);
#pragma unroll
      for (int a = 0; a < n; ++a)
#pragma unroll
        for (int b = 0; b < n; ++b)
          tx[a][b] = pgx[a][b], ty[a][b] = pgy[a][b];
      jxw(tx, -1.0);
      jxw(ty, -1.0);
      lat_test<P, false, true>(v, tx, ty, ihx, ihy, y1); // - grad eta_chi . grad psi    (mpi.cpp:407-409)
      load_u(1, u);
      lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy); // chi
      {
        double dB[n][n];
        sfor<n>([&](auto IY) {
          PFC_IC(iy, IY);
          sfor<n>([&](auto IX) {
            PFC_IC(ix, IX);
            double s = 0;
            sfor<n>([&](auto QY) {
              PFC_IC(qy, QY);
              sfor<n>([&](auto QX) {
                PFC_IC(qx, QX);
                constexpr double Ni = Tab1<P>::S(qx, ix) * Tab1<P>::S(qy, iy);
                constexpr double Dx = Tab1<P>::D(qx, ix) * Tab1<P>::S(qy, iy), Dy = Tab1<P>::S(qx, ix) * Tab1<P>::D(qy, iy);
                constexpr double W = Tab1<P>::W(qx) * Tab1<P>::W(qy);
                if constexpr (Ni != 0.0)
                  {
                    const double p_ = psi[qy][qx];
                    const double bm = 1.0 - dtth * (6.0 * p_ * v[qy][qx] + 6.0 * (pgx[qy][qx] * pgx[qy][qx] + pgy[qy][qx] * pgy[qy][qx]));
                    const double bxq = -12.0 * dtth * p_ * pgx[qy][qx], byq = -12.0 * dtth * p_ * pgy[qy][qx];
                    s += (Ni * W) * (bm * Ni + bxq * (Dx * ihx) + byq * (Dy * ihy));
                  }
              });
            });
            dB[iy][ix] = s * vol;
          });
        });
        store_y(3, dB);
      }
#pragma unroll
      for (int a = 0; a < n; ++a)
#pragma unroll
        for (int b = 0; b < n; ++b)
          {
            const double p_ = psi[a][b], g2 = pgx[a][b] * pgx[a][b] + pgy[a][b] * pgy[a][b];
            cf0[a][b] = -p_ + dtth * ((1.0 + O.eps) * v[a][b] + 3.0 * p_ * p_ * v[a][b] + 6.0 * p_ * g2);
          }
      jxw(v, -1.0);
      jxw(gx, -1.0);
      jxw(gy, -1.0);
      lat_test<P, true, false>(v, gx, gy, ihx, ihy, y1); // - eta_chi chi
      store_y(1, y1);
      lat_test<P, false, true>(v, gx, gy, ihx, ihy, y2); // - grad eta_phi . grad chi
      load_u(2, u);
      lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy); // phi
#pragma unroll
      for (int a = 0; a < n; ++a)
#pragma unroll
        for (int b = 0; b < n; ++b)
          {
            cf0[a][b] = fma(2.0 * dtth, v[a][b], cf0[a][b]);
            cfx[a][b] = dtth * gx[a][b];
            cfy[a][b] = dtth * gy[a][b];
          }
      jxw(v, -1.0);
      lat_test<P, true, false>(v, gx, gy, ihx, ihy, y2); // - eta_phi phi
      store_y(2, y2);
      load_u(3, u);
      if (dt1 == 0.0)
        {
          lat_interp<P, true, false>(u, ihx, ihy, v, gx, gy);
#pragma unroll
          for (int a = 0; a < n; ++a)
#pragma unroll
            for (int b = 0; b < n; ++b)
              cf0[a][b] += v[a][b];
        }
      else
        { // (1 - theta) terms at the previous time level (mpi.cpp:396-402)
          lat_interp<P, true, true>(u, ihx, ihy, psi, pgx, pgy);
          load_u(4, u);
          lat_interp<P, true, false>(u, ihx, ihy, v, gx, gy);
#pragma unroll
          for (int a = 0; a < n; ++a)
#pragma unroll
            for (int b = 0; b < n; ++b)
              {
                const double p_ = psi[a][b], g2 = pgx[a][b] * pgx[a][b] + pgy[a][b] * pgy[a][b];
                cf0[a][b] += p_ + dt1 * ((1.0 + O.eps) * v[a][b] + 3.0 * p_ * p_ * v[a][b] + 6.0 * p_ * g2);
              }
          load_u(5, u);
          lat_interp<P, true, true>(u, ihx, ihy, v, gx, gy);
#pragma unroll
          for (int a = 0; a < n; ++a)
#pragma unroll
            for (int b = 0; b < n; ++b)
              {
                cf0[a][b] = fma(2.0 * dt1, v[a][b], cf0[a][b]);
                cfx[a][b] = fma(dt1, gx[a][b], cfx[a][b]);
                cfy[a][b] = fma(dt1, gy[a][b], cfy[a][b]);
              }
        }
      jxw(cf0, 1.0);
      jxw(cfx, -1.0);
      jxw(cfy, -1.0);
      lat_test<P, true, true>(cf0, cfx, cfy, ihx, ihy, y0);
      store_y(0, y0);
    }
  // ---- nodes of the inner cells: gather the local results in a fixed order, fused epilogue ----
  constexpr int ONX = P * (CX - 1), ONY = P * (CY - 1), EIT = (ONX * ONY + NT - 1) / NT;
  // epilogue operands of the Schur-row modes are fetched before the barrier, so their latency overlaps the wait
  double e_add[EIT], e_d[EIT], e_dinv[EIT], e_s0[EIT];
  if constexpr (MODE == LAT_S)
    {
#pragma unroll
      for (int it = 0; it < EIT; ++it)
        {
          const int idx = tid + it * NT;
          const int X = nx0 + P + idx % ONX, Y = ny0 + P + idx / ONX;
          e_add[it] = e_d[it] = e_dinv[it] = e_s0[it] = 0.0;
          if (idx < ONX * ONY && lat_in_window(L, O.lo, X, Y))
            {
              const int64_t r = lat_id(L, X, Y);
              if (O.epi != 2)
                e_add[it] = O.add[O.add_c < 0 ? r : vaddr(r, O.add_c, Np)];
              if (O.epi == 1)
                {
                  e_d[it] = O.d[r];
                  e_s0[it] = O.s0[r];
                }
              if (O.epi != 0 || O.d)
                e_dinv[it] = O.dinv[r];
            }
        }
    }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < EIT; ++it)
    {
      const int idx = tid + it * NT;
      if (idx >= ONX * ONY)
        break;
      const int i = P + idx % ONX, j = P + idx / ONX; // tile node
      if (!lat_in_window(L, MODE == LAT_S ? O.lo : -1, nx0 + i, ny0 + j))
        continue; // not an owned free node (Schur row: not in the output window)
      const int ccx = i / P, jx = i % P, ccy = j / P, jy = j % P;
      double val[NOUT];
#pragma unroll
      for (int o = 0; o < NOUT; ++o)
        {
          // cells (ccy, ccx) [local node (jy, jx)] and, on a vertex line, the cell before [local node P]
          double s = 0;
          if (jy == 0)
            {
              if (jx == 0)
                s += s_out[(o * n * n + P * n + P) * NT + (ccy - 1) * CX + ccx - 1];
              s += s_out[(o * n * n + P * n + jx) * NT + (ccy - 1) * CX + ccx];
            }
          if (jx == 0)
            s += s_out[(o * n * n + jy * n + P) * NT + ccy * CX + ccx - 1];
          s += s_out[(o * n * n + jy * n + jx) * NT + ccy * CX + ccx];
          val[o] = s;
        }
      const int64_t r = lat_id(L, nx0 + i, ny0 + j); // owned: ly w + lx
      if constexpr (MODE == LAT_J)
        {
          O.out3[r] = val[0];
          O.out3[Np + r] = val[1];
          O.out3[2 * Np + r] = val[2];
        }
      else if constexpr (MODE == LAT_K)
        {
          double res = O.k * val[0];
          if (O.add)
            res += O.a * O.add[r];
          res *= O.inv_ml[r];
          O.out[r] = res;
          if (O.out2)
            O.out2[r] = res;
        }
      else if constexpr (MODE == LAT_S)
        {
          if (O.epi == 0)
            {
              const double rhs = e_add[it] + val[0];
              O.out[r] = rhs;
              if (O.d)
                { // first Chebyshev step fused: d = x = dinv rhs / theta
                  const double x0 = O.cheb[0] * e_dinv[it] * rhs;
                  O.d[r] = x0;
                  O.out2[r] = x0;
                }
            }
          else if (O.epi == 1)
            {
              const double res = e_add[it] - val[0];
              const double dn = O.cheb[0] * e_d[it] + O.cheb[1] * e_dinv[it] * res;
              O.d[r] = dn;
              O.out[r] = e_s0[it] + dn;
            }
          else
            O.out[r] = e_dinv[it] * val[0];
        }
      else if constexpr (MODE == LAT_T)
        {
#pragma unroll
          for (int c = 0; c < 4; ++c)
            O.out[c * O.zstride + r] = val[c];
        }
      else
        {
          O.out3[r] = val[0];
          O.out3[Np + r] = val[1];
          O.out3[2 * Np + r] = val[2];
          if (O.out)
            O.out[r] = val[3] != 0.0 ? 1.0 / val[3] : 1.0;
        }
    }
}

// ---- K product as a node stencil -----------------------------------------------------------------------
// K = Kx (x) My + Mx (x) Ky with the assembled 1-D rows of FE_Q(p) on a uniform periodic line (exactly what the
// cell quadrature produces, QGauss(p+1) integrates these polynomials exactly).  Row of a vertex node: offsets
// -p..p; of a Q2 mid node: -1..1.  Unit-length rows; M scales with h, K with 1/h.
template <int P>
struct Row1;
template <>
struct Row1<1>
{
  static __host__ __device__ constexpr double M(int, int d) { return d == 0 ? 4.0 / 6.0 : 1.0 / 6.0; }
  static __host__ __device__ constexpr double K(int, int d) { return d == 0 ? 2.0 : -1.0; }
};
template <>
struct Row1<2>
{
  static __host__ __device__ constexpr double M(int mid, int d)
  {
    return mid ? (d == 0 ? 16.0 / 30.0 : ((d == 1 || d == -1) ? 2.0 / 30.0 : 0.0))
               : (d == 0 ? 8.0 / 30.0 : ((d == 1 || d == -1) ? 2.0 / 30.0 : -1.0 / 30.0));
  }
  static __host__ __device__ constexpr double K(int mid, int d)
  {
    return mid ? (d == 0 ? 16.0 / 3.0 : ((d == 1 || d == -1) ? -8.0 / 3.0 : 0.0))
               : (d == 0 ? 14.0 / 3.0 : ((d == 1 || d == -1) ? -8.0 / 3.0 : 1.0 / 3.0));
  }
};


// out = (a add + k K s0) inv_ml on the owned free nodes; one thread = one column x kKsY rows: the 1-D x sums of
// 2 p + kKsY rows are formed once and combined per output row.  Interior threads address by arithmetic,
// threads whose support leaves the owned rectangle go through the frame lookup.
template <int P, int kKsX, int kKsY> // threads along x, rows per thread
static __global__ void __launch_bounds__(kKsX) lat_kstencil_kernel(LatGeom L, LatOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int R = kKsY + 2 * P;
  const int lx = blockIdx.x * kKsX + threadIdx.x, ly0 = blockIdx.y * kKsY;
  if (lx >= L.w)
    return;
  const double *__restrict__ v = O.s0;
  const int midx = (P == 2) ? ((L.ex0 + lx) & 1) : 0;
  double kx[2 * P + 1], mx[2 * P + 1];
  sfor<2 * P + 1>([&](auto DD) {
    PFC_IC(dd, DD);
    kx[dd] = (midx ? Row1<P>::K(1, dd - P) : Row1<P>::K(0, dd - P)) * L.ihx;
    mx[dd] = (midx ? Row1<P>::M(1, dd - P) : Row1<P>::M(0, dd - P)) * L.hx;
  });
  double rK[R], rM[R];
  const bool interior = lx - P >= 0 && lx + P < L.w && ly0 - P >= 0 && ly0 + kKsY - 1 + P < L.h;
  if (interior)
    {
      const double *p0 = v + (int64_t)(ly0 - P) * L.w + lx;
#pragma unroll
      for (int r = 0; r < R; ++r)
        {
          double sk = 0, sm = 0;
#pragma unroll
          for (int dd = 0; dd <= 2 * P; ++dd)
            {
              const double val = __ldg(p0 + (int64_t)r * L.w + (dd - P));
              sk = fma(kx[dd], val, sk);
              sm = fma(mx[dd], val, sm);
            }
          rK[r] = sk;
          rM[r] = sm;
        }
    }
  else
    {
#pragma unroll
      for (int r = 0; r < R; ++r)
        {
          const int ey = min(max(L.ey0 + ly0 - P + r, 0), L.fh - 1);
          double sk = 0, sm = 0;
#pragma unroll
          for (int dd = 0; dd <= 2 * P; ++dd)
            {
              const int ex = min(max(L.ex0 + lx + dd - P, 0), L.fw - 1);
              const double val = v[lat_id(L, ex, ey)];
              sk = fma(kx[dd], val, sk);
              sm = fma(mx[dd], val, sm);
            }
          rK[r] = sk;
          rM[r] = sm;
        }
    }
#pragma unroll
  for (int j = 0; j < kKsY; ++j)
    {
      const int ly = ly0 + j;
      if (ly >= L.h)
        break;
      const int midy = (P == 2) ? ((L.ey0 + ly) & 1) : 0;
      double s = 0;
      sfor<2 * P + 1>([&](auto DD) {
        PFC_IC(dd, DD);
        const double my = (midy ? Row1<P>::M(1, dd - P) : Row1<P>::M(0, dd - P)) * L.hy;
        const double ky = (midy ? Row1<P>::K(1, dd - P) : Row1<P>::K(0, dd - P)) * L.ihy;
        s = fma(my, rK[j + dd], s);
        s = fma(ky, rM[j + dd], s);
      });
      const int64_t r = (int64_t)ly * L.w + lx;
      double res = O.k * s;
      if (O.add)
        res += O.a * O.add[r];
      res *= O.inv_ml[r];
      O.out[r] = res;
      if (O.out2)
        O.out2[r] = res;
    }
}

// out_c = M v_c for the batched scalar fields c of zmask (blockIdx.z): consistent mass matrix Mx (x) My, same
// thread layout as the K stencil above.
template <int P, int kKsX, int kKsY>
static __global__ void __launch_bounds__(kKsX) lat_mass_kernel(LatGeom L, LatOp O)
{
  const int zmask = O.zmask_dev ? (O.zmask & *O.zmask_dev) : O.zmask;
  if (!((zmask >> blockIdx.z) & 1))
    return;
  constexpr int R = kKsY + 2 * P;
  const int lx = blockIdx.x * kKsX + threadIdx.x, ly0 = blockIdx.y * kKsY;
  if (lx >= L.w)
    return;
  const double *__restrict__ v = O.s0 + blockIdx.z * O.zstride;
  double *__restrict__ out = O.out + blockIdx.z * O.zstride;
  const int midx = (P == 2) ? ((L.ex0 + lx) & 1) : 0;
  double mx[2 * P + 1];
  sfor<2 * P + 1>([&](auto DD) {
    PFC_IC(dd, DD);
    mx[dd] = (midx ? Row1<P>::M(1, dd - P) : Row1<P>::M(0, dd - P)) * L.hx;
  });
  double rM[R];
  const bool interior = lx - P >= 0 && lx + P < L.w && ly0 - P >= 0 && ly0 + kKsY - 1 + P < L.h;
#pragma unroll
  for (int r = 0; r < R; ++r)
    {
      double sm = 0;
      if (interior)
        {
          const double *p0 = v + (int64_t)(ly0 - P + r) * L.w + lx;
#pragma unroll
          for (int dd = 0; dd <= 2 * P; ++dd)
            sm = fma(mx[dd], __ldg(p0 + (dd - P)), sm);
        }
      else
        {
          const int ey = min(max(L.ey0 + ly0 - P + r, 0), L.fh - 1);
#pragma unroll
          for (int dd = 0; dd <= 2 * P; ++dd)
            {
              const int ex = min(max(L.ex0 + lx + dd - P, 0), L.fw - 1);
              sm = fma(mx[dd], v[lat_id(L, ex, ey)], sm);
            }
        }
      rM[r] = sm;
    }
#pragma unroll
  for (int j = 0; j < kKsY; ++j)
    {
      const int ly = ly0 + j;
      if (ly >= L.h)
        break;
      const int midy = (P == 2) ? ((L.ey0 + ly) & 1) : 0;
      double s = 0;
      sfor<2 * P + 1>([&](auto DD) {
        PFC_IC(dd, DD);
        const double my = (midy ? Row1<P>::M(1, dd - P) : Row1<P>::M(0, dd - P)) * L.hy;
        s = fma(my, rM[j + dd], s);
      });
      out[(int64_t)ly * L.w + lx] = s;
    }
}

// Two columns per thread: the 2 p + 1 wide row supports of columns lx and lx + 1 lie in x[lx - 2 .. lx + 3], i.e.
// three aligned 16-byte loads per row instead of 2 (2 p + 1) 8-byte loads (the one-column kernel is bound by
// load instructions, not bytes).  lx is even, w even on the fast path.
template <int P, int KX, int KY>
static __global__ void __launch_bounds__(KX) lat_kstencil2_kernel(LatGeom L, LatOp O, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  constexpr int R = KY + 2 * P, NW = 2 * P + 1;
  const int lx = 2 * (blockIdx.x * KX + threadIdx.x), ly0 = blockIdx.y * KY;
  if (lx >= L.w)
    return;
  const double *__restrict__ v = O.s0;
  // column A = lx, column B = lx + 1: vertex / mid alternate for Q2
  const int midA = (P == 2) ? ((L.ex0 + lx) & 1) : 0, midB = (P == 2) ? (midA ^ 1) : 0;
  double kA[NW], mA[NW], kB[NW], mB[NW];
  sfor<NW>([&](auto DD) {
    PFC_IC(dd, DD);
    kA[dd] = (midA ? Row1<P>::K(1, dd - P) : Row1<P>::K(0, dd - P)) * L.ihx;
    mA[dd] = (midA ? Row1<P>::M(1, dd - P) : Row1<P>::M(0, dd - P)) * L.hx;
    kB[dd] = (midB ? Row1<P>::K(1, dd - P) : Row1<P>::K(0, dd - P)) * L.ihx;
    mB[dd] = (midB ? Row1<P>::M(1, dd - P) : Row1<P>::M(0, dd - P)) * L.hx;
  });
  double rKA[R], rMA[R], rKB[R], rMB[R];
  const bool fast = (L.w & 1) == 0 && lx - 2 >= 0 && lx + 3 < L.w && ly0 - P >= 0 && ly0 + KY - 1 + P < L.h;
  if (fast)
    {
      const double2 *p0 = reinterpret_cast<const double2 *>(v + (int64_t)(ly0 - P) * L.w + lx - 2);
      const int64_t w2 = L.w >> 1;
#pragma unroll
      for (int r = 0; r < R; ++r)
        {
          const double2 a = __ldg(p0 + r * w2), b = __ldg(p0 + r * w2 + 1), c = __ldg(p0 + r * w2 + 2);
          const double val[6] = {a.x, a.y, b.x, b.y, c.x, c.y};
          double ska = 0, sma = 0, skb = 0, smb = 0;
#pragma unroll
          for (int dd = 0; dd < NW; ++dd)
            {
              ska = fma(kA[dd], val[2 + dd - P], ska);
              sma = fma(mA[dd], val[2 + dd - P], sma);
              skb = fma(kB[dd], val[3 + dd - P], skb);
              smb = fma(mB[dd], val[3 + dd - P], smb);
            }
          rKA[r] = ska, rMA[r] = sma, rKB[r] = skb, rMB[r] = smb;
        }
    }
  else
    {
#pragma unroll
      for (int r = 0; r < R; ++r)
        {
          const int ey = min(max(L.ey0 + ly0 - P + r, 0), L.fh - 1);
          double val[6];
#pragma unroll
          for (int k = 0; k < 6; ++k)
            val[k] = v[lat_id(L, min(max(L.ex0 + lx - 2 + k, 0), L.fw - 1), ey)];
          double ska = 0, sma = 0, skb = 0, smb = 0;
#pragma unroll
          for (int dd = 0; dd < NW; ++dd)
            {
              ska = fma(kA[dd], val[2 + dd - P], ska);
              sma = fma(mA[dd], val[2 + dd - P], sma);
              skb = fma(kB[dd], val[3 + dd - P], skb);
              smb = fma(mB[dd], val[3 + dd - P], smb);
            }
          rKA[r] = ska, rMA[r] = sma, rKB[r] = skb, rMB[r] = smb;
        }
    }
  const bool hasB = lx + 1 < L.w;
#pragma unroll
  for (int j = 0; j < KY; ++j)
    {
      const int ly = ly0 + j;
      if (ly >= L.h)
        break;
      const int midy = (P == 2) ? ((L.ey0 + ly) & 1) : 0;
      double sa = 0, sb = 0;
      sfor<NW>([&](auto DD) {
        PFC_IC(dd, DD);
        const double my = (midy ? Row1<P>::M(1, dd - P) : Row1<P>::M(0, dd - P)) * L.hy;
        const double ky = (midy ? Row1<P>::K(1, dd - P) : Row1<P>::K(0, dd - P)) * L.ihy;
        sa = fma(my, rKA[j + dd], sa);
        sa = fma(ky, rMA[j + dd], sa);
        sb = fma(my, rKB[j + dd], sb);
        sb = fma(ky, rMB[j + dd], sb);
      });
      const int64_t r = (int64_t)ly * L.w + lx;
      double ra = O.k * sa, rb = O.k * sb;
      if (fast)
        {
          if (O.add)
            {
              const double2 ad = *reinterpret_cast<const double2 *>(O.add + r);
              ra += O.a * ad.x;
              rb += O.a * ad.y;
            }
          const double2 im = *reinterpret_cast<const double2 *>(O.inv_ml + r);
          const double2 res = make_double2(ra * im.x, rb * im.y);
          *reinterpret_cast<double2 *>(O.out + r) = res;
          if (O.out2)
            *reinterpret_cast<double2 *>(O.out2 + r) = res;
        }
      else
        {
          if (O.add)
            {
              ra += O.a * O.add[r];
              if (hasB)
                rb += O.a * O.add[r + 1];
            }
          ra *= O.inv_ml[r];
          O.out[r] = ra;
          if (O.out2)
            O.out2[r] = ra;
          if (hasB)
            {
              rb *= O.inv_ml[r + 1];
              O.out[r + 1] = rb;
              if (O.out2)
                O.out2[r + 1] = rb;
            }
        }
    }
}

// inv_ml of the lattice: 1 / (row sum of M) = 1 / (hx hy wx wy), w = 1-D lumped weight of the node's column
// type (Q1: 1; Q2: vertex 1/3, mid 2/3); rows past the free nodes (slaves, padding) get 1
template <int P>
static __global__ void lat_inv_ml_kernel(LatGeom L, double *inv_ml, int64_t n_rows)
{
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rows)
    return;
  double v = 1.0;
  if (r < (int64_t)L.w * L.h)
    {
      const int ex = L.ex0 + (int)(r % L.w), ey = L.ey0 + (int)(r / L.w);
      const double wx = P == 1 ? 1.0 : (ex % P == 0 ? 1.0 / 3.0 : 2.0 / 3.0);
      const double wy = P == 1 ? 1.0 : (ey % P == 0 ? 1.0 / 3.0 : 2.0 / 3.0);
      v = 1.0 / (L.hx * L.hy * wx * wy);
    }
  inv_ml[r] = v;
}
} // namespace pfc
