// pfc_precond.cuh — kernels of the block-Schur right preconditioner (PFC_PRECOND_BLOCK_SCHUR).
//
// J = [[B, C, D], [K, M, 0], [0, K, M]] (mpi.cpp:421-443, tex:399-466).  With the row-sum lumped mass M_L
// (diagonal) in place of M the block elimination the reference sketches in solve_and_update
// (mpi.cpp:496-572; serial.cpp:319-369; tex:469-662) is explicit:
//     m1  = M_L^-1 r1,   t = M_L^-1 (r2 - K m1)
//     psi = S_L^-1 (r0 - C m1 - D t),   S_L = B - C M_L^-1 K + D M_L^-1 K M_L^-1 K
//     chi = M_L^-1 (r1 - K psi),   phi = M_L^-1 (r2 - K chi)
// and S_L^-1 is approximated by `sweeps` steps of Jacobi-Chebyshev (diag(B)^-1 S_L has a spectrum in
// about [1.2, 5] for Q1 and [1.3, 17] for Q2 at the shipped parameters, measured by power iteration).
// The preconditioner is a fixed linear operator, so plain right-preconditioned GMRES applies and the
// stopping test stays the true residual ||b - J x|| <= tol of the reference.
// All vectors here are scalar fields in internal node order: [owned Np | ghosts n_ghost].
#pragma once
#include "pfc_krylov.cuh"

namespace pfc
{
struct SellArgs
{
  const int64_t *slice_ptr;
  const int32_t *col;
  const double *B, *C, *M, *K;
  int64_t Np, n_slices;
  double dtth;
};

// out[r] = (a * add[r] + k * sum_j K[r,j] v[j]) * inv_ml[r]         (12 B per stored entry)
static __global__ void __launch_bounds__(kSpmvThreads) kmv_kernel(SellArgs S, const double *__restrict__ add, double a,
                                                                   const double *__restrict__ v, double k,
                                                                   const double *__restrict__ inv_ml,
                                                                   double *__restrict__ out, double *__restrict__ out2,
                                                                   const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  const int64_t w = ((int64_t)blockIdx.x * kSpmvThreads + threadIdx.x) >> 5;
  if (w >= S.n_slices)
    return;
  const int lane = threadIdx.x & 31;
  const int64_t base = S.slice_ptr[w];
  const int width = (int)((S.slice_ptr[w + 1] - base) >> 5);
  const int64_t p0 = base + lane;
  double s = 0;
#pragma unroll 4
  for (int j = 0; j < width; ++j)
    {
      const int64_t p = p0 + (int64_t)j * 32;
      const int32_t c = __ldg(S.col + p);
      const double vk = __ldg(S.K + p);
      if (c >= 0) // constrained rows have no K entry
        s += vk * v[c];
    }
  const int64_t r = w * 32 + lane;
  double res = k * s;
  if (add)
    res += a * add[r];
  res *= inv_ml[r];
  out[r] = res;
  if (out2)
    out2[r] = res;
}

// y = sB * B v0 + sC * C v1 + sD * D v2 on scalar fields, D = dtth (K - 2M)     (36 B per stored entry)
//   MODE 0: out[r] = add[r] + y                                   (Schur right-hand side: sB = 0, sC = sD = -1)
//   MODE 1: Chebyshev step:  res = add[r] - y;  d = c1 d + c2 dinv res;  out[r] = v0[r] + d   (sB = 1, sC = -1, sD = 1)
//   MODE 2: out[r] = dinv[r] * y                                  (power iteration on diag(B)^-1 S_L)
template <int MODE>
static __global__ void __launch_bounds__(kSpmvThreads) bcd_kernel(SellArgs S, const double *__restrict__ v0,
                                                                   const double *__restrict__ v1,
                                                                   const double *__restrict__ v2, double sB, double sC,
                                                                   double sD, const double *__restrict__ add,
                                                                   const double *__restrict__ dinv, double *__restrict__ d,
                                                                   const double *__restrict__ cheb, double *__restrict__ out,
                                                                   const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  const int64_t w = ((int64_t)blockIdx.x * kSpmvThreads + threadIdx.x) >> 5;
  if (w >= S.n_slices)
    return;
  const int lane = threadIdx.x & 31;
  const int64_t base = S.slice_ptr[w];
  const int width = (int)((S.slice_ptr[w + 1] - base) >> 5);
  const int64_t p0 = base + lane;
  double y = 0;
#pragma unroll 4
  for (int j = 0; j < width; ++j)
    {
      const int64_t p = p0 + (int64_t)j * 32;
      int32_t c = __ldg(S.col + p);
      const double vb = __ldg(S.B + p);
      if (c < 0)
        { // constrained row: S_L row is the B diagonal only
          if (MODE != 0)
            y += sB * vb * v0[~c];
        }
      else
        {
          const double vc = __ldg(S.C + p), vm = __ldg(S.M + p), vk = __ldg(S.K + p);
          double t = sC * vc * v1[c] + sD * S.dtth * (vk - 2.0 * vm) * v2[c];
          if (MODE != 0)
            t += sB * vb * v0[c];
          y += t;
        }
    }
  const int64_t r = w * 32 + lane;
  if (MODE == 0)
    out[r] = add[r] + y;
  else if (MODE == 1)
    {
      const double res = add[r] - y;
      const double dn = cheb[0] * d[r] + cheb[1] * dinv[r] * res;
      d[r] = dn;
      out[r] = v0[r] + dn;
    }
  else
    out[r] = dinv[r] * y;
}

// per row: inv_ml = 1 / sum_j M[r,j] (CONST) or dinvB = 1 / B[r,r]; padding rows get 1
template <bool ROWSUM_M>
static __global__ void __launch_bounds__(kSpmvThreads) rowdiag_kernel(SellArgs S, double *__restrict__ out)
{
  const int64_t w = ((int64_t)blockIdx.x * kSpmvThreads + threadIdx.x) >> 5;
  if (w >= S.n_slices)
    return;
  const int lane = threadIdx.x & 31;
  const int64_t base = S.slice_ptr[w];
  const int width = (int)((S.slice_ptr[w + 1] - base) >> 5);
  const int64_t r = w * 32 + lane;
  double s = 0;
  for (int j = 0; j < width; ++j)
    {
      const int64_t p = base + lane + (int64_t)j * 32;
      int32_t c = S.col[p];
      if (ROWSUM_M)
        s += S.M[p];
      else
        {
          if (c < 0)
            c = ~c;
          if (c == r && S.B[p] != 0.0 && s == 0.0)
            s = S.B[p];
        }
    }
  out[r] = (s != 0.0) ? 1.0 / s : 1.0;
}

// out = a * x * y   (y may be null: out = a * x), optionally out2 = out; scalar from the device if a_dev
static __global__ void vec_mul_kernel(double *out, double *out2, double a, const double *a_dev, const double *x,
                                      const double *y, int64_t n, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  const double s = a_dev ? a * (*a_dev) : a;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    {
      const double v = y ? s * x[i] * y[i] : s * x[i];
      out[i] = v;
      if (out2)
        out2[i] = v;
    }
}

// out = a x + b y
static __global__ void vec_axpby_kernel(double *out, double a, const double *x, double b, const double *y, int64_t n)
{
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = a * x[i] + b * y[i];
}

// scalar halo pack: buf[k] = v[rows[k]]
static __global__ void halo_pack1_kernel(double *buf, const double *v, const int32_t *rows, int64_t n)
{
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n)
    buf[k] = v[rows[k]];
}
} // namespace pfc
