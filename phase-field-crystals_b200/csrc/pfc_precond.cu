// pfc_precond.cu — host drivers of the block-Schur right preconditioner (kernels: pfc_precond.cuh).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "pfc_comm.h"
#include "pfc_handle.h"

using namespace pfc;

#define PFC_TRY(expr)    \
  do                     \
    {                    \
      int rc_ = (expr);  \
      if (rc_ != PFC_OK) \
        return rc_;      \
    }                    \
  while (0)

static SellArgs sell_args(const pfc_handle *h)
{
  SellArgs S;
  S.slice_ptr = h->d_slice_ptr.p;
  S.col = h->d_col.p;
  S.B = h->d_B.p;
  S.C = h->d_C.p;
  S.M = h->d_M.p;
  S.K = h->d_K.p;
  S.Np = h->plan.Np;
  S.n_slices = h->plan.n_slices;
  S.dtth = h->prm.dt * h->prm.theta;
  return S;
}

static inline int sell_blocks(const pfc_handle *h)
{
  return (int)((h->plan.n_slices * 32 + kSpmvThreads - 1) / kSpmvThreads);
}
static inline int vgrid(const pfc_handle *h, int64_t n)
{
  return (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)h->n_sm * 8));
}

// K / B-C-D products of the assembled path (the lattice path runs precond_apply_ring below)
static void run_kmv(pfc_handle *h, const SellArgs &S, int nb, const double *add, double a, const double *v, double k,
                    double *out, double *out2, const Ctl *ctl)
{
  kmv_kernel<<<nb, kSpmvThreads, 0, h->stream>>>(S, add, a, v, k, h->d_inv_ml.p, out, out2, ctl);
}

template <int MODE>
static void run_bcd(pfc_handle *h, const SellArgs &S, int nb, const double *v0, const double *v1, const double *v2,
                    double sB, double sC, double sD, const double *add, const double *dinv, double *d,
                    const double *cheb, double *out, const Ctl *ctl)
{
  bcd_kernel<MODE><<<nb, kSpmvThreads, 0, h->stream>>>(S, v0, v1, v2, sB, sC, sD, add, dinv, d, cheb, out, ctl);
}

// work vectors inside d_pw
enum { W_M1 = 0, W_T, W_RHS, W_XA, W_XB, W_D, W_T1, W_T2, W_CHI };
static inline double *wv(const pfc_handle *h, int k) { return h->d_pw.p + (int64_t)k * h->sstride; }

// lumped mass: inv_ml = 1 / rowsum(M); M is iterate independent (assembled once in pfc_setup)
int precond_after_setup(pfc_handle *h)
{
  if (!h->lat) // the lattice path knows the lumped mass in closed form (lat_setup)
    {
      rowdiag_kernel<true><<<sell_blocks(h), kSpmvThreads, 0, h->stream>>>(sell_args(h), h->d_inv_ml.p);
      h->launches++;
    }
  PFC_CUDA(h, cudaGetLastError());
  // deterministic start vector of the power iterations: zero on constrained / padding rows
  const Plan &P = h->plan;
  std::vector<double> v((size_t)h->sstride * 3, 0.0);
  for (int64_t r = 0; r < P.Np; ++r)
    if (P.perm[r] >= 0 && P.row_master[r] == r)
      {
        const uint64_t g = (uint64_t)P.perm[r] * 0x9E3779B97F4A7C15ull;
        v[r] = 0.5 + (double)((g >> 40) & 0xFFFF) / 65536.0;            // for lambda_max
        v[h->sstride + r] = ((g >> 20) & 1) ? 1.0 : -1.0;               // rough start for the other end
      }
  PFC_CUDA(h, cudaMemcpyAsync(h->d_pv.p, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  h->eig_valid = false;
  return PFC_OK;
}

// y = diag(B)^-1 S_L v   (scalar fields), the operator whose spectrum the Chebyshev interval must cover
static int apply_dinv_SL(pfc_handle *h, double *v, double *y)
{
  if (h->lat)
    { // the smoother's operator on the lattice path is the constant-coefficient S~ (pfc_ring.cuh)
      PFC_TRY(halo_exchange_scalar(h, v));
      RingOp o;
      std::memset(&o, 0, sizeof(o));
      o.x = v;
      o.dinv = h->d_dinvB.p;
      o.out = y;
      o.bbar = 1.0;
      o.dtth = h->prm.dt * h->prm.theta;
      o.lo = 3 * h->plan.degree;
      o.epi = 2;
      return ring_launch_sweep(h, o, nullptr);
    }
  const SellArgs S = sell_args(h);
  const int nb = sell_blocks(h);
  cudaStream_t s = h->stream;
  double *t1 = wv(h, W_T1), *t2 = wv(h, W_T2);
  PFC_TRY(halo_exchange_scalar(h, v));
  run_kmv(h, S, nb, nullptr, 0.0, v, 1.0, t1, nullptr, nullptr);
  PFC_TRY(halo_exchange_scalar(h, t1));
  run_kmv(h, S, nb, nullptr, 0.0, t1, 1.0, t2, nullptr, nullptr);
  PFC_TRY(halo_exchange_scalar(h, t2));
  run_bcd<2>(h, S, nb, v, t1, t2, 1.0, -1.0, 1.0, nullptr, h->d_dinvB.p, nullptr, nullptr, y, nullptr);
  h->launches += 3;
  return PFC_OK;
}

static int scalar_norm(pfc_handle *h, const double *v, double *out)
{
  const int64_t n2 = h->plan.Np / 2;
  GmresSmall *g = h->d_small.p;
  multidot_kernel<0><<<vgrid(h, n2), kVecThreads, 0, h->stream>>>(nullptr, 0, v, n2, h->d_partial.p, &g->nrm2[0],
                                                                  h->d_counter.p, nullptr, 0, ArDev{});
  h->launches++;
  PFC_TRY(allreduce_sum(h, &g->nrm2[0], 1));
  PFC_CUDA(h, cudaMemcpyAsync(h->h_scalar, &g->nrm2[0], sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  *out = std::sqrt(h->h_scalar[0]);
  return PFC_OK;
}

// after every assembly of B: Jacobi scaling; every few solves: spectrum bounds by (shifted) power iteration,
// warm-started from the previous vectors; then the Chebyshev coefficients.
int precond_refresh(pfc_handle *h)
{
  const Plan &P = h->plan;
  cudaStream_t s = h->stream;
  if (!h->lat) // the lattice residual kernel already left 1 / diag(B) in d_dinvB
    {
      rowdiag_kernel<false><<<sell_blocks(h), kSpmvThreads, 0, s>>>(sell_args(h), h->d_dinvB.p);
      h->launches++;
    }
  else
    PFC_TRY(halo_exchange_scalar(h, h->d_dinvB.p)); // the sweeps run on the ring too
  const bool estimate = !h->eig_valid || h->solves_since_eig >= 16;
  if (estimate)
    {
      const int iters = h->eig_valid ? 8 : 30;
      double *v = h->d_pv.p, *v2 = h->d_pv.p + h->sstride, *y = h->d_pv.p + 2 * h->sstride;
      double lam = 0, nv = 0;
      const int g = vgrid(h, P.Np);
      for (int pass = 0; pass < 2; ++pass)
        {
          double *x = pass == 0 ? v : v2;
          PFC_TRY(scalar_norm(h, x, &nv));
          if (!(nv > 0))
            return h->fail(PFC_ERR_NONFINITE, "preconditioner: zero power-iteration vector");
          vec_mul_kernel<<<g, 256, 0, s>>>(x, nullptr, 1.0 / nv, nullptr, x, nullptr, P.Np, nullptr);
          double mu = 0;
          for (int it = 0; it < iters; ++it)
            {
              PFC_TRY(apply_dinv_SL(h, x, y));
              if (pass == 1) // shifted operator lam_max I - A: dominant eigenvalue lam_max - lam_min
                vec_axpby_kernel<<<g, 256, 0, s>>>(y, lam, x, -1.0, y, P.Np);
              PFC_TRY(scalar_norm(h, y, &mu));
              if (!std::isfinite(mu) || !(mu > 0))
                return h->fail(PFC_ERR_NONFINITE, "preconditioner: power iteration broke down");
              vec_mul_kernel<<<g, 256, 0, s>>>(x, nullptr, 1.0 / mu, nullptr, y, nullptr, P.Np, nullptr);
              h->launches += 2;
            }
          if (pass == 0)
            lam = 1.1 * mu; // power iteration converges from below: safety margin
          else
            {
              h->eig_max = lam;
              h->eig_min = std::max(0.85 * (lam - mu), lam / 200.0);
            }
        }
      h->eig_valid = true;
      h->solves_since_eig = 0;
      // Chebyshev coefficients for steps 1 .. sweeps-1 (step 0 is x = dinv r / theta)
      const double theta = 0.5 * (h->eig_max + h->eig_min), delta = 0.5 * (h->eig_max - h->eig_min);
      std::vector<double> c(64, 0.0);
      const double sigma = theta / delta;
      double rho = 1.0 / sigma;
      c[0] = 1.0 / theta;
      for (int i = 1; i < h->sweeps; ++i)
        {
          const double rho_new = 1.0 / (2.0 * sigma - rho);
          c[2 * i] = rho_new * rho;
          c[2 * i + 1] = 2.0 * rho_new / delta;
          rho = rho_new;
        }
      PFC_CUDA(h, cudaMemcpyAsync(h->d_cheb.p, c.data(), 64 * sizeof(double), cudaMemcpyHostToDevice, s));
      PFC_CUDA(h, cudaStreamSynchronize(s));
    }
  h->solves_since_eig++;
  PFC_CUDA(h, cudaGetLastError());
  return PFC_OK;
}

// scalar field [owned | ring] -> component 0 of a 3-component vector (only needed when there is a single sweep)
static __global__ void ring_to_vec0_kernel(double *__restrict__ out, const double *__restrict__ x, int64_t n, int64_t Np,
                                           const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  for (int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u < n; u += (int64_t)gridDim.x * blockDim.x)
    out[u < Np ? u : 3 * u] = x[u];
}

// Lattice path: the whole application on the ring overlay, ONE halo exchange (of `in`), then local kernels on windows that
// shrink by the reach of every stage (margins in nodes, p = degree, s = sweeps):
//   t kernel p | Schur right-hand side + first Chebyshev step 2p | sweep i 2p + 3p i | back substitution 3p s + p
// so that out = (psi, chi, phi) is valid one cell beyond the owned nodes and the operator needs no exchange of its own
// (ring depth (3 s + 2) p, pfc_api.cu).
static int precond_apply_ring(pfc_handle *h, double *in, double *out, const Ctl *ctl)
{
  const Plan &P = h->plan;
  const int p = P.degree, sw = h->sweeps;
  double *m1 = wv(h, W_M1), *t = wv(h, W_T), *rhs = wv(h, W_RHS), *xa = wv(h, W_XA), *xb = wv(h, W_XB), *d = wv(h, W_D);
  PFC_TRY(halo_exchange(h, in, ctl));
  if (p == 1 && h->q1_fast)
    { // Q1: t kernel, Schur right-hand side and first Chebyshev step in one pass (pfc_ring_q1.cuh)
      RingOp a;
      std::memset(&a, 0, sizeof(a));
      a.in = in;
      a.psi = h->d_psi_n.p;
      a.dinv = h->d_dinvB.p;
      a.cheb = h->d_cheb.p;
      a.d = d;
      a.out = rhs;
      a.out2 = xa;
      a.dtth = h->prm.dt * h->prm.theta;
      a.eps = h->prm.eps;
      a.lo = 2 * p;
      PFC_TRY(ring_launch_rhs1(h, a, ctl));
    }
  else
    {
      RingOp a;
      std::memset(&a, 0, sizeof(a));
      a.in = in;
      a.t = t;
      a.m1 = m1;
      a.lo = p;
      PFC_TRY(ring_launch_t(h, a, ctl));
      LatOp op = lat_op(h); // rhs = r0 - C m1 - D t (exact, iterate-dependent C), d = x = dinv rhs / theta
      op.in3b = h->d_psi_n.p;
      op.s1 = m1;
      op.s2 = t;
      op.sB = 0.0;
      op.sC = -1.0;
      op.sD = -1.0;
      op.add = in;
      op.add_c = 0;
      op.dinv = h->d_dinvB.p;
      op.cheb = h->d_cheb.p;
      op.d = d;
      op.out = rhs;
      op.out2 = xa;
      op.epi = 0;
      op.lo = 2 * p;
      PFC_TRY(lat_launch(h, LAT_S, op, ctl));
    }
  double *x = xa, *xn = xb;
  for (int i = 1; i < sw; ++i)
    {
      RingOp o;
      std::memset(&o, 0, sizeof(o));
      o.x = x;
      o.rhs = rhs;
      o.dinv = h->d_dinvB.p;
      o.cheb = h->d_cheb.p + 2 * i;
      o.d = d;
      o.bbar = 1.0;
      o.dtth = h->prm.dt * h->prm.theta;
      o.lo = 2 * p + 3 * p * i;
      o.epi = 0;
      if (i == sw - 1)
        o.out = out, o.out_vec = 1; // the last sweep writes psi straight into the result vector
      else
        o.out = xn;
      PFC_TRY(ring_launch_sweep(h, o, ctl));
      std::swap(x, xn);
    }
  if (sw == 1)
    {
      const int64_t n = P.Np + P.n_ghost;
      ring_to_vec0_kernel<<<vgrid(h, n), 256, 0, h->stream>>>(out, xa, n, P.Np, ctl);
      h->launches++;
    }
  RingOp b;
  std::memset(&b, 0, sizeof(b));
  b.in = in;
  b.z = out;
  b.lo = 3 * p * sw + p;
  PFC_TRY(ring_launch_back(h, b, ctl));
  PFC_CUDA(h, cudaGetLastError());
  return PFC_OK;
}

// out = P^-1 in   (in, out: solver vectors of 3 components; stream-ordered, graph-capturable).  Lattice path: the ring
// of `in` is refreshed here and out is valid one cell beyond the owned nodes.
int precond_apply(pfc_handle *h, double *in, double *out, const Ctl *ctl)
{
  if (h->lat)
    return precond_apply_ring(h, in, out, ctl);
  const Plan &P = h->plan;
  const SellArgs S = sell_args(h);
  const int nb = sell_blocks(h), g = vgrid(h, P.Np);
  cudaStream_t s = h->stream;
  const int64_t Np = P.Np;
  const double *r0 = in, *r1 = in + Np, *r2 = in + 2 * Np;
  double *m1 = wv(h, W_M1), *t = wv(h, W_T), *rhs = wv(h, W_RHS), *xa = wv(h, W_XA), *xb = wv(h, W_XB);
  double *d = wv(h, W_D), *t1 = wv(h, W_T1), *t2 = wv(h, W_T2), *chi = wv(h, W_CHI);
  const double *iml = h->d_inv_ml.p;

  vec_mul_kernel<<<g, 256, 0, s>>>(m1, nullptr, 1.0, nullptr, r1, iml, Np, ctl);                     // m1 = M_L^-1 r1
  PFC_TRY(halo_exchange_scalar(h, m1, ctl));
  run_kmv(h, S, nb, r2, 1.0, m1, -1.0, t, nullptr, ctl);                // t = M_L^-1 (r2 - K m1)
  PFC_TRY(halo_exchange_scalar(h, t, ctl));
  // Schur right-hand side rhs = r0 - C m1 - D t, then Chebyshev on S_L psi = rhs with Jacobi scaling diag(B)^-1:
  // first step d = x = dinv rhs / theta
  run_bcd<0>(h, S, nb, nullptr, m1, t, 0.0, -1.0, -1.0, r0, nullptr, nullptr, nullptr, rhs, ctl);
  vec_mul_kernel<<<g, 256, 0, s>>>(d, xa, 1.0, h->d_cheb.p, rhs, h->d_dinvB.p, Np, ctl);          // d = x = dinv rhs / theta
  h->launches += 4;
  double *x = xa, *xn = xb;
  for (int i = 1; i < h->sweeps; ++i)
    {
      PFC_TRY(halo_exchange_scalar(h, x, ctl));
      run_kmv(h, S, nb, nullptr, 0.0, x, 1.0, t1, nullptr, ctl);
      PFC_TRY(halo_exchange_scalar(h, t1, ctl));
      run_kmv(h, S, nb, nullptr, 0.0, t1, 1.0, t2, nullptr, ctl);
      PFC_TRY(halo_exchange_scalar(h, t2, ctl));
      run_bcd<1>(h, S, nb, x, t1, t2, 1.0, -1.0, 1.0, rhs, h->d_dinvB.p, d, h->d_cheb.p + 2 * i, xn, ctl);
      std::swap(x, xn);
      h->launches += 3;
    }
  // back substitution; psi -> out[0], chi -> out[1], phi -> out[2]
  PFC_TRY(halo_exchange_scalar(h, x, ctl));
  run_kmv(h, S, nb, r1, 1.0, x, -1.0, chi, out + Np, ctl);              // chi = M_L^-1 (r1 - K psi)
  PFC_TRY(halo_exchange_scalar(h, chi, ctl));
  run_kmv(h, S, nb, r2, 1.0, chi, -1.0, out + 2 * Np, nullptr, ctl);    // phi = M_L^-1 (r2 - K chi)
  vec_mul_kernel<<<g, 256, 0, s>>>(out, nullptr, 1.0, nullptr, x, nullptr, Np, ctl);
  h->launches += 3;
  PFC_CUDA(h, cudaGetLastError());
  return PFC_OK;
}

// bench.py roofline legs of the preconditioner (include/pfc_b200.h): ms5 / bytes4, see the header
extern "C" int pfc_time_precond(pfc_handle *h, int32_t reps, double *ms5, double *bytes4)
{
  if (!h || reps < 1 || !ms5)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->assembled || h->prm.preconditioner != PFC_PRECOND_BLOCK_SCHUR)
    return h->fail(PFC_ERR_STATE, "pfc_time_precond needs pfc_assemble and PFC_PRECOND_BLOCK_SCHUR");
  PFC_CUDA(h, cudaSetDevice(h->device));
  PFC_TRY(precond_refresh(h));
  const Plan &P = h->plan;
  const SellArgs S = sell_args(h);
  const int nb = sell_blocks(h), p = P.degree;
  cudaStream_t s = h->stream;
  double *t1 = wv(h, W_T1), *t2 = wv(h, W_T2), *x = wv(h, W_XA), *xn = wv(h, W_XB);
  const int n_legs = h->lat ? 5 : 3;
  for (int i = 0; i < 5; ++i)
    ms5[i] = 0;
  // each leg: `reps` launches captured in one CUDA graph (the solver runs its cycle as a graph too), so the
  // number is device time per launch, not host launch rate; 3-component operands rotate through the Krylov basis
  const int half = std::max(1, h->n_basis / 2);
  float ms;
  for (int leg = 0; leg < n_legs; ++leg)
    {
      if (leg == 3 && h->lat && p == 1 && h->q1_fast)
        continue; // Q1: the t kernel is fused into the right-hand side kernel (leg 2)
      cudaGraph_t graph = nullptr;
      cudaGraphExec_t exec = nullptr;
      const int64_t before = h->launches;
      PFC_CUDA(h, cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
      int rc = PFC_OK;
      for (int i = 0; i < reps && rc == PFC_OK; ++i)
        {
          double *in = h->d_V.p + (int64_t)(i % half) * h->stride, *out = h->d_V.p + (int64_t)(half + i % half) * h->stride;
          if (leg == 0)
            rc = precond_apply(h, in, h->d_z.p, nullptr);
          else if (h->lat)
            {
              RingOp o;
              std::memset(&o, 0, sizeof(o));
              o.in = in;
              o.bbar = 1.0;
              o.dtth = h->prm.dt * h->prm.theta;
              if (leg == 1)
                {
                  o.x = (i & 1) ? xn : x, o.out = (i & 1) ? x : xn, o.rhs = wv(h, W_RHS), o.dinv = h->d_dinvB.p;
                  o.cheb = h->d_cheb.p + 2, o.d = wv(h, W_D), o.lo = 5 * p;
                  rc = ring_launch_sweep(h, o, nullptr);
                }
              else if (leg == 2 && p == 1 && h->q1_fast)
                {
                  o.psi = h->d_psi_n.p, o.dinv = h->d_dinvB.p, o.cheb = h->d_cheb.p, o.d = wv(h, W_D), o.out = wv(h, W_RHS);
                  o.out2 = x, o.eps = h->prm.eps, o.lo = 2 * p;
                  rc = ring_launch_rhs1(h, o, nullptr);
                }
              else if (leg == 2)
                {
                  LatOp op = lat_op(h);
                  op.in3b = h->d_psi_n.p, op.s1 = wv(h, W_M1), op.s2 = wv(h, W_T), op.sC = -1.0, op.sD = -1.0;
                  op.add = in, op.add_c = 0, op.dinv = h->d_dinvB.p, op.cheb = h->d_cheb.p, op.d = wv(h, W_D);
                  op.out = wv(h, W_RHS), op.out2 = x, op.epi = 0, op.lo = 2 * p;
                  rc = lat_launch(h, LAT_S, op, nullptr);
                }
              else if (leg == 3)
                {
                  o.t = wv(h, W_T), o.m1 = wv(h, W_M1), o.lo = p;
                  rc = ring_launch_t(h, o, nullptr);
                }
              else
                {
                  o.z = out, o.lo = 3 * p * h->sweeps + p;
                  rc = ring_launch_back(h, o, nullptr);
                }
            }
          else if (leg == 1)
            run_kmv(h, S, nb, nullptr, 0.0, x, 1.0, t1, nullptr, nullptr);
          else
            run_bcd<1>(h, S, nb, x, t1, t2, 1.0, -1.0, 1.0, wv(h, W_RHS), h->d_dinvB.p, wv(h, W_D), h->d_cheb.p + 2, xn, nullptr);
        }
      cudaError_t e = cudaStreamEndCapture(s, &graph);
      h->launches = before;
      if (rc != PFC_OK)
        {
          if (graph)
            cudaGraphDestroy(graph);
          return rc;
        }
      PFC_CUDA(h, e);
      PFC_CUDA(h, cudaGraphInstantiate(&exec, graph, 0));
      cudaGraphDestroy(graph);
      cudaGraphLaunch(exec, s); // warm-up
      PFC_CUDA(h, cudaEventRecord(h->ev[0], s));
      cudaGraphLaunch(exec, s);
      PFC_CUDA(h, cudaEventRecord(h->ev[1], s));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      cudaGraphExecDestroy(exec);
      PFC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
      ms5[leg] = ms / reps;
    }
  PFC_CUDA(h, cudaGetLastError());
  if (bytes4)
    {
      bytes4[0] = h->lat ? lat_bytes(h, RING_SWEEP) : 12.0 * (double)P.n_ent + 32.0 * (double)P.Np;
      bytes4[1] = h->lat ? lat_bytes(h, (p == 1 && h->q1_fast) ? RING_RHS1 : RING_RHS) : 36.0 * (double)P.n_ent + 64.0 * (double)P.Np;
      bytes4[2] = (h->lat && !(p == 1 && h->q1_fast)) ? lat_bytes(h, RING_T) : 0.0;
      bytes4[3] = h->lat ? lat_bytes(h, RING_BACK) : 0.0;
    }
  return PFC_OK;
}

// Q1 closed-form kernels (pfc_ring_q1.cuh) against the kernels they replace, on a pseudo-random vector tied to the global
// node numbers: max_rel2[0] = J apply (lat_j1_kernel vs lat_kernel<1, LAT_J>), max_rel2[1] = preconditioner application
// (fused right-hand side kernel, closed-form sweeps and back substitution vs t kernel + lat_kernel<1, LAT_S> + the
// general ring kernels); max |a - b| / max |b| over the owned rows of all three components.  [collective]
extern "C" int pfc_selfcheck_q1(pfc_handle *h, double *max_rel2)
{
  if (!h || !max_rel2)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || !h->lat || h->plan.degree != 1 || !h->assembled || h->prm.preconditioner != PFC_PRECOND_BLOCK_SCHUR)
    return h->fail(PFC_ERR_UNSUPPORTED, "pfc_selfcheck_q1 needs the Q1 lattice path, pfc_assemble and PFC_PRECOND_BLOCK_SCHUR");
  PFC_CUDA(h, cudaSetDevice(h->device));
  PFC_TRY(precond_refresh(h));
  const Plan &P = h->plan;
  const LatGeom &G = h->lat_geom;
  const int64_t n = (int64_t)G.w * G.h, Np = P.Np;
  cudaStream_t s = h->stream;
  double *in = h->d_V.p, *ya = h->d_V.p + h->stride, *yb = h->d_V.p + 2 * h->stride;
  std::vector<double> v((size_t)h->stride, 0.0), a((size_t)3 * Np), b((size_t)3 * Np);
  for (int c = 0; c < 3; ++c)
    for (int64_t r = 0; r < n; ++r)
      {
        uint64_t st = ((uint64_t)P.perm[r] * 3 + c) * 0x9E3779B97F4A7C15ull + 88172645463325252ull;
        st ^= st << 13, st ^= st >> 7, st ^= st << 17;
        v[c * Np + r] = (double)(st >> 11) / 9007199254740992.0 - 0.5;
      }
  const bool keep = h->q1_fast;
  int rc = PFC_OK;
  for (int leg = 0; leg < 2 && rc == PFC_OK; ++leg)
    {
      PFC_CUDA(h, cudaMemcpyAsync(in, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice, s));
      PFC_CUDA(h, cudaMemsetAsync(ya, 0, h->stride * sizeof(double), s));
      PFC_CUDA(h, cudaMemsetAsync(yb, 0, h->stride * sizeof(double), s));
      for (int variant = 0; variant < 2 && rc == PFC_OK; ++variant)
        {
          h->q1_fast = variant == 0;
          double *y = variant == 0 ? ya : yb;
          if (leg == 0)
            {
              rc = halo_exchange(h, in, nullptr);
              LatOp op = lat_op(h);
              op.in3a = in;
              op.in3b = h->d_psi_n.p;
              op.out3 = y;
              if (rc == PFC_OK)
                rc = lat_launch(h, LAT_J, op, nullptr);
            }
          else
            rc = precond_apply(h, in, y, nullptr);
        }
      h->q1_fast = keep;
      if (rc != PFC_OK)
        return rc;
      PFC_CUDA(h, cudaGetLastError());
      PFC_CUDA(h, cudaMemcpyAsync(a.data(), ya, a.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
      PFC_CUDA(h, cudaMemcpyAsync(b.data(), yb, b.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      double worst = 0, scale = 0;
      for (int c = 0; c < 3; ++c)
        for (int64_t r = 0; r < n; ++r)
          {
            worst = std::max(worst, std::fabs(a[c * Np + r] - b[c * Np + r]));
            scale = std::max(scale, std::fabs(b[c * Np + r]));
          }
      double e = scale > 0 ? worst / scale : 1.0; // an all-zero reference is a failure, too
      if (!(e == e))
        e = 1.0;
      max_rel2[leg] = e;
    }
  return PFC_OK;
}
