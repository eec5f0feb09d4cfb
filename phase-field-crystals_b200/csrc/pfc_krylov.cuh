// pfc_krylov.cuh — fused GMRES vector kernels (K5): multi-dot, multi-axpy with norm, Givens
// update on the device.  Restates what deal.II's SolverGMRES does inside
// solver_gmres.solve(system_matrix, dPsi_n, system_rhs, PreconditionIdentity()) (mpi.cpp:480-482)
// [dealii-ext]: restarted GMRES, x0 = 0, Gram-Schmidt with the Kelley re-orthogonalisation
// test, Givens QR of the Hessenberg matrix, stop on |gamma_dim| <= tol.
#pragma once
#include "pfc_kernels.cuh"
#include "pfc_p2p.cuh"

namespace pfc
{
constexpr int kVecThreads = 256;
constexpr int kMaxBasis = 64;         // max n_tmp_vectors supported
constexpr int kPartialStride = 64;    // doubles per block in the partial-sum scratch

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Deterministic grid reduction of NV values per block: per-warp shuffle, per-block shared memory,
// per-grid "last block sums the partials in block order".  out[i] = sum over blocks.
// ar.R > 0: the last block also all-reduces the NV results over the ranks (peer-memory slots, rank order), so no
// separate collective kernel follows.
template <int NV>
__device__ __forceinline__ void grid_reduce(double (&acc)[NV], double *partial, double *out,
                                            unsigned int *counter, const ArDev &ar)
{
  __shared__ double s_red[kVecThreads / 32][NV];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < NV; ++i)
    {
      const double s = warp_sum(acc[i]);
      if (lane == 0)
        s_red[warp][i] = s;
    }
  __syncthreads();
  if (threadIdx.x < NV)
    {
      double s = 0;
#pragma unroll
      for (int w = 0; w < kVecThreads / 32; ++w)
        s += s_red[w][threadIdx.x];
      partial[(size_t)blockIdx.x * kPartialStride + threadIdx.x] = s;
    }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0)
    s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (s_last)
    {
      // warp w sums values w, w + 8, ...; its lanes stride over the blocks, then shuffle-reduce:
      // fixed order, so the result does not depend on which block happened to finish last
      for (int i = warp; i < NV; i += kVecThreads / 32)
        {
          double s = 0;
          for (unsigned int b = lane; b < gridDim.x; b += 32)
            s += __ldcg(partial + (size_t)b * kPartialStride + i);
          s = warp_sum(s);
          if (lane == 0)
            out[i] = s;
        }
      if (threadIdx.x == 0)
        *counter = 0;
      if (ar.R > 1)
        {
          __syncthreads();
          ar_block_allreduce(ar, out, NV);
        }
    }
}

// out[i] = V_i . w for i < DIM, out[DIM] = w . w        (one pass over DIM + 1 vectors)
// pass 1 (second Gram-Schmidt pass) runs only if ctl->reorth is set.
// Loads are issued in batches of kGsBatch vectors before their products: under register pressure (DIM >= 21) ptxas
// otherwise alternates one load with one product, and a thread then has two loads in flight instead of eight
// (measured 3.7 instead of 6.7 TB/s at DIM 24).  DIM >= 12 asks for two CTAs per SM (128 registers).
constexpr int kGsBatch = 8;
template <int DIM>
static __global__ void __launch_bounds__(kVecThreads, DIM >= 12 ? 2 : 3)
  multidot_kernel(const double *__restrict__ V, int64_t stride, const double *__restrict__ w, int64_t n2, double *partial,
                  double *out, unsigned int *counter, const Ctl *ctl, int pass, ArDev ar)
{
  (void)pass;
  if (ctl && ctl->state != 0)
    return;
  double acc[DIM + 1];
#pragma unroll
  for (int i = 0; i <= DIM; ++i)
    acc[i] = 0;
  const double2 *w2 = reinterpret_cast<const double2 *>(w);
  for (int64_t idx = (int64_t)blockIdx.x * kVecThreads + threadIdx.x; idx < n2; idx += (int64_t)gridDim.x * kVecThreads)
    {
      const double2 wv = w2[idx];
      acc[DIM] += wv.x * wv.x + wv.y * wv.y;
#pragma unroll
      for (int i0 = 0; i0 < DIM; i0 += kGsBatch)
        {
          double2 v[kGsBatch];
#pragma unroll
          for (int b = 0; b < kGsBatch; ++b)
            if (i0 + b < DIM)
              v[b] = __ldg(reinterpret_cast<const double2 *>(V + (int64_t)(i0 + b) * stride) + idx);
#pragma unroll
          for (int b = 0; b < kGsBatch; ++b)
            if (i0 + b < DIM)
              acc[i0 + b] += v[b].x * wv.x + v[b].y * wv.y;
        }
    }
  grid_reduce<DIM + 1>(acc, partial, out, counter, ar);
}

// w -= sum_i h_i V_i ; out[0] = ||w||^2 after the update.  The last block also applies the
// Kelley test (pass 0): if ||w'|| <= 10 sqrt(eps) ||w|| the sticky reorth flag is raised.
template <int DIM>
static __global__ void __launch_bounds__(kVecThreads, DIM >= 12 ? 2 : 3)
  multiaxpy_kernel(const double *__restrict__ V, int64_t stride, double *__restrict__ w, int64_t n2,
                   const double *__restrict__ h, const double *__restrict__ is2, double *partial, double *out,
                   unsigned int *counter, Ctl *ctl, int pass, ArDev ar)
{
  (void)pass;
  if (ctl->state != 0)
    return;
  double hh[DIM]; // is2: the basis vectors are stored unnormalised (s_i v_i) and h holds (s_i v_i) . w: coefficient h_i / s_i^2
#pragma unroll
  for (int i = 0; i < DIM; ++i)
    hh[i] = is2 ? h[i] * is2[i] : h[i];
  double acc[1] = {0};
  double2 *w2 = reinterpret_cast<double2 *>(w);
  for (int64_t idx = (int64_t)blockIdx.x * kVecThreads + threadIdx.x; idx < n2; idx += (int64_t)gridDim.x * kVecThreads)
    {
      double2 wv = w2[idx];
#pragma unroll
      for (int i0 = 0; i0 < DIM; i0 += kGsBatch)
        {
          double2 v[kGsBatch];
#pragma unroll
          for (int b = 0; b < kGsBatch; ++b)
            if (i0 + b < DIM)
              v[b] = __ldg(reinterpret_cast<const double2 *>(V + (int64_t)(i0 + b) * stride) + idx);
#pragma unroll
          for (int b = 0; b < kGsBatch; ++b)
            if (i0 + b < DIM)
              {
                wv.x -= hh[i0 + b] * v[b].x;
                wv.y -= hh[i0 + b] * v[b].y;
              }
        }
      w2[idx] = wv;
      acc[0] += wv.x * wv.x + wv.y * wv.y;
    }
  grid_reduce<1>(acc, partial, out, counter, ar);
}

// per-solve small state on the device
struct GmresSmall
{
  double H[kMaxBasis * kMaxBasis]; // H[i * ld + j], ld = kMaxBasis
  double gamma[kMaxBasis + 1], ci[kMaxBasis], si[kMaxBasis], y[kMaxBasis];
  double dots0[kMaxBasis + 1];     // pass 0: V_i . w, [dim] = w . w
  double dots1[kMaxBasis + 1];     // pass 1
  double nrm2[2];                  // ||w||^2 after pass 0 / pass 1
  // lazy normalisation (preconditioned solves): basis vector i is stored as s_i v_i, s_0 = 1, s_{j+1} = ||w'_j||
  double s2[kMaxBasis + 1], is2[kMaxBasis + 1]; // s_i^2 and 1 / s_i^2
};

// Hessenberg column, Givens rotation, residual estimate, SolverControl::check.
// two_pass: the iteration ran both Gram-Schmidt passes (PFC_ORTHO_CGS2, or after a Kelley trip).
// Otherwise the Kelley test [dealii-ext: norm_vv > 10 * norm_vv_start * sqrt(eps)] is applied to the single
// pass; if it trips, the iteration is discarded and the cycle ends with state 3: the host continues the
// solve with two-pass cycles (the re-orthogonalisation flag is sticky, as in deal.II).
static __global__ void givens_kernel(GmresSmall *g, Ctl *ctl, int inner, int two_pass, int lazy)
{
  if (ctl->state != 0)
    return;
  const int dim = inner + 1;
  if (!two_pass)
    {
      const double before = sqrt(g->dots0[dim]), after = sqrt(g->nrm2[0]);
      if (!(after > 10.0 * before * sqrt(2.220446049250313e-16)))
        {
          ctl->reorth = 1;
          ctl->kelley_trips++;
          ctl->state = 3;
          return;
        }
    }
  double h[kMaxBasis + 1];
  for (int i = 0; i < dim; ++i)
    h[i] = g->dots0[i] + (two_pass ? g->dots1[i] : 0.0);
  const double n2w = two_pass ? g->nrm2[1] : g->nrm2[0];
  double s = sqrt(n2w);
  if (lazy)
    { // Lazy normalisation: the stored vectors are s_i v_i and w is s_j (J P^-1 v_j), so the dots carry s_i s_j and the
      // norm carries s_j; the new vector stays as it is, s_{j+1} = ||w'|| (no pass over it to divide by its norm).
      const double isj = sqrt(g->is2[inner]);
      for (int i = 0; i < dim; ++i)
        h[i] *= sqrt(g->is2[i]) * isj;
      g->s2[dim] = n2w;
      g->is2[dim] = n2w != 0.0 ? 1.0 / n2w : 1.0;
      s *= isj;
    }
  h[dim] = s;
  ctl->inv_norm = (s != 0.0) ? 1.0 / s : 1.0;
  for (int i = 0; i < inner; ++i)
    {
      const double sn = g->si[i], cs = g->ci[i], d = h[i];
      h[i] = cs * d + sn * h[i + 1];
      h[i + 1] = -sn * d + cs * h[i + 1];
    }
  const double r = 1.0 / sqrt(h[inner] * h[inner] + h[inner + 1] * h[inner + 1]);
  g->si[inner] = h[inner + 1] * r;
  g->ci[inner] = h[inner] * r;
  h[inner] = g->ci[inner] * h[inner] + g->si[inner] * h[inner + 1];
  g->gamma[inner + 1] = -g->si[inner] * g->gamma[inner];
  g->gamma[inner] *= g->ci[inner];
  for (int i = 0; i < dim; ++i)
    g->H[i * kMaxBasis + inner] = h[i];
  ctl->iters += 1;
  ctl->dim = dim;
  const double rho = fabs(g->gamma[dim]);
  ctl->rho = rho;
  if (rho <= ctl->tol)
    ctl->state = 1;
  else if (ctl->iters >= ctl->max_iters || isnan(rho))
    ctl->state = 2;
}

// w *= ctl->inv_norm (normalisation of the newest basis vector)
static __global__ void __launch_bounds__(kVecThreads) scale_kernel(double *__restrict__ w, int64_t n2, const Ctl *ctl)
{
  if (ctl->state != 0)
    return;
  const double a = ctl->inv_norm;
  double2 *w2 = reinterpret_cast<double2 *>(w);
  for (int64_t idx = (int64_t)blockIdx.x * kVecThreads + threadIdx.x; idx < n2; idx += (int64_t)gridDim.x * kVecThreads)
    {
      double2 v = w2[idx];
      v.x *= a;
      v.y *= a;
      w2[idx] = v;
    }
}

// restart: v0 = b - A x (A x is in v0 on entry, or use_ax = 0 when x = 0), out[0] = ||v0||^2
static __global__ void __launch_bounds__(kVecThreads) residual_kernel(double *__restrict__ v0, const double *__restrict__ b,
                                                               int64_t n2, int use_ax, double *partial, double *out,
                                                               unsigned int *counter, const Ctl *ctl, ArDev ar)
{
  if (ctl->state != 0)
    return;
  double acc[1] = {0};
  double2 *v2 = reinterpret_cast<double2 *>(v0);
  const double2 *b2 = reinterpret_cast<const double2 *>(b);
  for (int64_t idx = (int64_t)blockIdx.x * kVecThreads + threadIdx.x; idx < n2; idx += (int64_t)gridDim.x * kVecThreads)
    {
      double2 r = b2[idx];
      if (use_ax)
        {
          const double2 p = v2[idx];
          r.x -= p.x;
          r.y -= p.y;
        }
      v2[idx] = r;
      acc[0] += r.x * r.x + r.y * r.y;
    }
  grid_reduce<1>(acc, partial, out, counter, ar);
}

// start of a restart cycle: rho = ||b - A x||, check, gamma_0 = rho
static __global__ void cycle_begin_kernel(GmresSmall *g, Ctl *ctl)
{
  if (ctl->state != 0)
    return;
  const double rho = sqrt(g->nrm2[0]);
  ctl->rho = rho;
  ctl->dim = 0;
  if (rho <= ctl->tol)
    ctl->state = 1;
  else if (ctl->iters >= ctl->max_iters || isnan(rho))
    ctl->state = 2;
  g->gamma[0] = rho;
  g->s2[0] = g->is2[0] = 1.0; // v_0 is normalised by the scale kernel that follows
  ctl->inv_norm = rho != 0.0 ? 1.0 / rho : 1.0;
}

// gate of the end part when a cycle is enqueued in two pieces: the end runs (gate->state = 0) once the solve has
// stopped or the restart length is reached, otherwise the continuation graph follows and runs it
static __global__ void end_gate_kernel(const Ctl *ctl, Ctl *gate, int restart)
{
  gate->state = (ctl->state != 0 || ctl->dim >= restart) ? 0 : 1;
}

// end of a cycle: back substitution H y = gamma (H1.backward), dim = ctl->dim
static __global__ void backsolve_kernel(GmresSmall *g, Ctl *ctl, const Ctl *gate, int lazy)
{
  if (gate->state != 0)
    return;
  const int dim = ctl->dim;
  for (int i = dim - 1; i >= 0; --i)
    {
      double s = g->gamma[i];
      for (int j = i + 1; j < dim; ++j)
        s -= g->H[i * kMaxBasis + j] * g->y[j];
      g->y[i] = s / g->H[i * kMaxBasis + i];
    }
  if (lazy) // x += sum y_i v_i = sum (y_i / s_i) (s_i v_i)
    for (int i = 0; i < dim; ++i)
      g->y[i] *= sqrt(g->is2[i]);
}

// x += sum_{i < dim} y_i V_i ; dim read from the device; consumes the cycle (dim := 0)
static __global__ void __launch_bounds__(kVecThreads) xupdate_kernel(const double *__restrict__ V, int64_t stride,
                                                              double *__restrict__ x, int64_t n2,
                                                              const GmresSmall *g, const Ctl *ctl, const Ctl *gate)
{
  const int dim = ctl->dim;
  if (dim <= 0 || gate->state != 0)
    return;
  __shared__ double s_y[kMaxBasis];
  if (threadIdx.x < dim)
    s_y[threadIdx.x] = g->y[threadIdx.x];
  __syncthreads();
  double2 *x2 = reinterpret_cast<double2 *>(x);
  for (int64_t idx = (int64_t)blockIdx.x * kVecThreads + threadIdx.x; idx < n2; idx += (int64_t)gridDim.x * kVecThreads)
    {
      double2 xv = x2[idx];
#pragma unroll 4
      for (int i = 0; i < dim; ++i)
        {
          const double2 v = __ldg(reinterpret_cast<const double2 *>(V + (int64_t)i * stride) + idx);
          xv.x += s_y[i] * v.x;
          xv.y += s_y[i] * v.y;
        }
      x2[idx] = xv;
    }
}

// ---- Newton-level vector kernels ------------------------------------------------------------------
// constraints.distribute (mpi.cpp:485): v[slave] = v[master], three components
static __global__ void distribute_kernel(double *v, const int32_t *slave, const int32_t *master, int64_t n, int64_t Np)
{
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n)
    return;
  const int32_t s = slave[k], m = master[k];
#pragma unroll
  for (int c = 0; c < 3; ++c)
    v[vaddr(s, c, Np)] = v[vaddr(m, c, Np)];
}

// Psi_n += dPsi_n (mpi.cpp:487-491) over the whole local vector
static __global__ void axpy1_kernel(double *__restrict__ y, const double *__restrict__ x, int64_t n, const Ctl *ctl = nullptr)
{
  if (ctl && ctl->state != 0)
    return;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] += x[i];
}

// API order <-> internal order.  map[k] = internal address of API dof k's value, or -1
// (gather takes owned entries only, address < limit = 3 Np, so rank-wise results can be summed)
static __global__ void gather_kernel(double *dst, const double *src, const int64_t *map, int64_t n, int64_t limit)
{
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n && map[k] >= 0 && map[k] < limit)
    dst[k] = src[map[k]];
}
static __global__ void scatter_kernel(double *dst, const double *src, const int64_t *map, int64_t n)
{
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n && map[k] >= 0)
    dst[map[k]] = src[k];
}

// halo pack: buf[3k + c] = v[c*Np + rows[k]]
static __global__ void halo_pack_kernel(double *buf, const double *v, const int32_t *rows, int64_t n, int64_t Np)
{
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n)
    return;
  const int32_t r = rows[k];
  buf[3 * k] = v[r];
  buf[3 * k + 1] = v[Np + r];
  buf[3 * k + 2] = v[2 * Np + r];
}
} // namespace pfc
