// pfc_reduced.cu — host side of the psi-only reduced Newton solve (PFC_PRECOND_SCHUR_REDUCED; kernels and the derivation:
// pfc_reduced.cuh; GMRES driver: pfc_solver.cu, which calls reduced_prepare / reduced_apply / reduced_precond /
// reduced_finish).  Restates what solve_and_update of the serial class does (phase_field_crystal_system.cpp:319-369: psi_rhs
// :349-351, GMRES on the psi system :353-355, back substitution :357-361) with the substitution dpsi = M^2 y.
#include <cmath>
#include <cstring>
#include <vector>

#include "pfc_comm.h"
#include "pfc_handle.h"
#include "pfc_krylov.cuh"
#include "pfc_reduced.cuh"

using namespace pfc;

#define PFC_TRY(expr)            \
  do                             \
    {                            \
      int rc_ = (expr);          \
      if (rc_ != PFC_OK)         \
        return rc_;              \
    }                            \
  while (0)

namespace
{
inline int vgrid(const pfc_handle *h, int64_t n)
{
  return (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)h->n_sm * 8));
}

// Symbol of S~ = M + dt theta (K - 2 M)(M_L^-1 K)^2 on the torus (Q1, cell hx x hy): its mean is the diagonal entry,
// its range over the mean the spectrum of the Jacobi-scaled operator on any periodic lattice of this cell.
void stilde_symbol(double hx, double hy, double dtth, double &diag, double &lmin, double &lmax)
{
  constexpr int n = 96;
  const double two_pi = 6.283185307179586476925286766559;
  double sum = 0, lo = 1e300, hi = -1e300;
  for (int a = 0; a < n; ++a)
    for (int b = 0; b < n; ++b)
      {
        const double cx = std::cos(two_pi * a / n), cy = std::cos(two_pi * b / n);
        const double mx = (4.0 + 2.0 * cx) / 6.0, my = (4.0 + 2.0 * cy) / 6.0, kx = 2.0 - 2.0 * cx, ky = 2.0 - 2.0 * cy;
        const double m = hx * hy * mx * my, k = hy / hx * kx * my + hx / hy * mx * ky, ml = hx * hy;
        const double s = m + dtth * (k - 2.0 * m) * (k / ml) * (k / ml);
        sum += s;
        lo = std::min(lo, s);
        hi = std::max(hi, s);
      }
  diag = sum / (n * n);
  lmin = lo / diag;
  lmax = hi / diag;
}

// squared l2 norm over the ranks of n doubles on the device -> host
int norm2(pfc_handle *h, const double *v, int64_t n, double *out)
{
  GmresSmall *g = h->d_small.p;
  multidot_kernel<0><<<vgrid(h, n / 2), kVecThreads, 0, h->stream>>>(nullptr, 0, v, n / 2, h->d_partial.p, &g->nrm2[0], h->d_counter.p,
                                                                    nullptr, 0, ArDev{});
  h->launches++;
  PFC_TRY(allreduce_sum(h, &g->nrm2[0], 1));
  PFC_CUDA(h, cudaMemcpyAsync(h->h_scalar, &g->nrm2[0], sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  *out = h->h_scalar[0];
  return PFC_OK;
}

// x = M^-1 b by conjugate gradients on the consistent mass stencil (lat_mass_kernel), absolute tolerance on ||b - M x||.
// Host driven (two reductions per iteration): it runs only while the auxiliary equations are violated, i.e. in the first
// Newton iteration after pfc_set_state.  Work fields r, p, q: scalar fields of the arena (p takes part in the exchange).
int mass_solve(pfc_handle *h, const double *b, double *x, double *r, double *p, double *q, double tol, double *res2)
{
  const int64_t Np = h->plan.Np;
  cudaStream_t s = h->stream;
  GmresSmall *g = h->d_small.p;
  const int grid = vgrid(h, Np);
  PFC_CUDA(h, cudaMemsetAsync(x, 0, Np * sizeof(double), s));
  PFC_CUDA(h, cudaMemsetAsync(q, 0, h->sstride * sizeof(double), s));
  PFC_CUDA(h, cudaMemcpyAsync(r, b, Np * sizeof(double), cudaMemcpyDeviceToDevice, s));
  PFC_CUDA(h, cudaMemcpyAsync(p, b, Np * sizeof(double), cudaMemcpyDeviceToDevice, s));
  double rr = 0;
  PFC_TRY(norm2(h, r, Np, &rr));
  for (int it = 0; it < 500 && std::sqrt(rr) > tol; ++it)
    {
      PFC_TRY(halo_exchange_scalar(h, p));
      LatOp o = lat_op(h);
      o.s0 = p;
      o.out = q;
      o.zstride = h->sstride;
      o.zmask = 1;
      PFC_TRY(lat_launch(h, LAT_M, o, nullptr));
      multidot_kernel<1><<<vgrid(h, Np / 2), kVecThreads, 0, s>>>(p, 0, q, Np / 2, h->d_partial.p, g->dots0, h->d_counter.p, nullptr, 0, ArDev{});
      h->launches++;
      PFC_TRY(allreduce_sum(h, g->dots0, 2));
      PFC_CUDA(h, cudaMemcpyAsync(h->h_scalar, g->dots0, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      const double pq = h->h_scalar[0];
      if (!(pq > 0) || !std::isfinite(pq))
        return h->fail(PFC_ERR_NONFINITE, "reduced solve: mass CG broke down (p.Mp = %.3e)", pq);
      const double alpha = rr / pq;
      axpby_kernel<<<grid, 256, 0, s>>>(x, alpha, p, 1.0, Np, nullptr);
      axpby_kernel<<<grid, 256, 0, s>>>(r, -alpha, q, 1.0, Np, nullptr);
      double rr_new = 0;
      PFC_TRY(norm2(h, r, Np, &rr_new));
      axpby_kernel<<<grid, 256, 0, s>>>(p, 1.0, r, rr_new / rr, Np, nullptr);
      h->launches += 3;
      rr = rr_new;
      h->red_mass_iters++;
    }
  if (!(std::sqrt(rr) <= tol))
    return h->fail(PFC_ERR_GMRES_NOT_CONVERGED, "reduced solve: mass CG did not reach %.1e (residual %.3e)", tol, std::sqrt(rr));
  *res2 = rr;
  return PFC_OK;
}
} // namespace

// host-only view for the CPU tests (tests/test_reduced_cpu.py): out3 = {diag(S~), min, max of the Jacobi-scaled symbol}
extern "C" int pfc_reduced_symbol(double hx, double hy, double dt_theta, double *out3)
{
  if (!out3 || !(hx > 0) || !(hy > 0))
    return PFC_ERR_BAD_ARGUMENT;
  stilde_symbol(hx, hy, dt_theta, out3[0], out3[1], out3[2]);
  return PFC_OK;
}

// end of pfc_setup on the Q1 lattice path: work fields, the constant Jacobi scaling and the Chebyshev coefficients
int reduced_setup(pfc_handle *h)
{
  const Plan &P = h->plan;
  const LatGeom &G = h->lat_geom;
  cudaStream_t s = h->stream;
  if (P.Np % 2 != 0 || h->d_pw.n < (size_t)9 * h->sstride)
    return h->fail(PFC_ERR_STATE, "reduced solve: work fields missing");
  double *w = h->d_pw.p;
  const int64_t ss = h->sstride;
  h->red_xa = w, h->red_xb = w + ss, h->red_d = w + 2 * ss, h->red_z = w + 3 * ss, h->red_y = w + 4 * ss;
  h->red_u = w + 5 * ss, h->red_f = w + 6 * ss, h->red_um = w + 7 * ss, h->red_wm = w + 8 * ss;
  const double dtth = h->prm.dt * h->prm.theta;
  double diag, lmin, lmax;
  stilde_symbol(G.hx, G.hy, dtth, diag, lmin, lmax);
  if (!(diag > 0) || !(lmin > 0))
    return h->fail(PFC_ERR_UNSUPPORTED, "reduced solve: S~ is not positive definite for this cell size and dt theta (symbol minimum %.3e)", lmin * diag);
  const int64_t n = P.Np + P.n_ghost;
  PFC_CUDA(h, h->d_dinvS.alloc((size_t)n));
  fill_kernel<<<vgrid(h, n), 256, 0, s>>>(h->d_dinvS.p, 1.0 / diag, n);
  h->launches++;
  // interval [0.85 min, 1.1 max], not wider than 200 (as precond_refresh); coefficients for steps 1 .. sweeps-1, step 0 is
  // x = dinv r / theta
  const double hi = 1.1 * lmax, lo = std::max(0.85 * lmin, hi / 200.0);
  const double theta = 0.5 * (hi + lo), delta = 0.5 * (hi - lo), sigma = theta / delta;
  std::vector<double> c(64, 0.0);
  double rho = 1.0 / sigma;
  c[0] = 1.0 / theta;
  for (int i = 1; i < h->sweeps; ++i)
    {
      const double rho_new = 1.0 / (2.0 * sigma - rho);
      c[2 * i] = rho_new * rho;
      c[2 * i + 1] = 2.0 * rho_new / delta;
      rho = rho_new;
    }
  h->red_dinvc = 1.0 / diag;
  h->red_alpha = c[0] / diag;
  PFC_CUDA(h, h->d_cheb_red.upload(c, s));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  PFC_CUDA(h, cudaGetLastError());
  h->reduced = true;
  h->operator_note += std::string(h->operator_note.empty() ? "" : "; ") + "psi-only reduced solve active (PFC_PRECOND_SCHUR_REDUCED)";
  return PFC_OK;
}

// out = T y = row 0 of J (M^2 y, -K M y, K^2 y).  halo: y holds owned values only (else: valid three nodes beyond them)
int reduced_apply(pfc_handle *h, double *y, double *out, const Ctl *ctl, bool halo)
{
  if (halo)
    PFC_TRY(halo_exchange_scalar(h, y, ctl));
  LatOp a = lat_op(h);
  a.s0 = y;
  a.out3 = h->d_z.p;
  PFC_TRY(lat_launch(h, LAT_P, a, ctl));
  LatOp b = lat_op(h);
  b.in3a = h->d_z.p;
  b.in3b = h->d_psi_n.p;
  b.out3 = out;
  return lat_launch(h, LAT_J0, b, ctl);
}

// *out = Chebyshev_sweeps(S~) in, valid 3 (sweeps - 1) nodes inside the padded rectangle: one exchange (of `in`), then
// local kernels on shrinking windows (as precond_apply_ring).  Step 0 has no stencil (x_0 = d_0 = alpha in): the first
// sweep makes it on the fly from `in` (ring_sweep1a_kernel<FIRST>), so two steps are ONE pass: in -> out, 16 B per node.
int reduced_precond(pfc_handle *h, double *in, double **out, const Ctl *ctl)
{
  const Plan &P = h->plan;
  const int sw = h->sweeps;
  PFC_TRY(halo_exchange_scalar(h, in, ctl));
  if (sw == 1)
    {
      const int64_t n = P.Np + P.n_ghost;
      cheb0_kernel<<<vgrid(h, n), 256, 0, h->stream>>>(h->red_d, h->red_xa, in, h->d_dinvS.p, h->d_cheb_red.p, n, ctl);
      h->launches++;
      *out = h->red_xa;
      PFC_CUDA(h, cudaGetLastError());
      return PFC_OK;
    }
  double *x = h->red_xa, *xn = h->red_xb;
  for (int i = 1; i < sw; ++i)
    {
      RingOp o;
      std::memset(&o, 0, sizeof(o));
      o.rhs = in;
      o.cheb = h->d_cheb_red.p + 2 * i;
      o.bbar = 1.0;
      o.dtth = h->prm.dt * h->prm.theta;
      o.lo = 3 * i;
      o.epi = 0;
      if (i == 1)
        {
          o.first = 1;
          o.alpha = h->red_alpha;
          o.dinvc = h->red_dinvc;
          o.d = sw > 2 ? h->red_d : nullptr;
          o.out = sw == 2 ? h->red_z : x;
        }
      else
        {
          o.x = x;
          o.dinv = h->d_dinvS.p;
          o.d = h->red_d;
          o.out = i == sw - 1 ? h->red_z : xn;
          std::swap(x, xn);
        }
      PFC_TRY(ring_launch_sweep(h, o, ctl));
    }
  *out = h->red_z;
  PFC_CUDA(h, cudaGetLastError());
  return PFC_OK;
}

// Before GMRES: f, and the tolerance left for the psi row.  The 3-block residual of the update built by reduced_finish is
// (f - T y, G - M u, H - (K u - M w)); while ||(G, H)|| is below a tenth of the tolerance u = w = 0 and f = F.
int reduced_prepare(pfc_handle *h, double *tol)
{
  const Plan &P = h->plan;
  const int64_t Np = P.Np;
  cudaStream_t s = h->stream;
  const double tol0 = *tol;
  const int grid = vgrid(h, Np);
  double aux2 = 0;
  PFC_TRY(norm2(h, h->d_rhs.p + Np, 2 * Np, &aux2));
  if (!std::isfinite(aux2))
    return h->fail(PFC_ERR_NONFINITE, "reduced solve: non-finite right-hand side");
  // rows of the Krylov vectors past the owned nodes (Np is rounded up) are never written by the operator
  const int64_t n_own = (int64_t)h->lat_geom.w * h->lat_geom.h;
  if (Np > n_own)
    PFC_CUDA(h, cudaMemset2DAsync(h->d_V.p + n_own, h->sstride * sizeof(double), 0, (Np - n_own) * sizeof(double), h->n_basis, s));
  h->red_have_uw = std::sqrt(aux2) > 0.1 * tol0;
  if (!h->red_have_uw)
    {
      PFC_CUDA(h, cudaMemcpyAsync(h->red_f, h->d_rhs.p, Np * sizeof(double), cudaMemcpyDeviceToDevice, s));
      h->red_aux2 = aux2;
    }
  else
    {
      const double mtol = 0.05 * tol0;
      double *t3 = h->d_z.p, *y3 = h->d_u.p; // 3-component vectors
      double r2u = 0, r2w = 0;
      // u = M^-1 G
      PFC_TRY(mass_solve(h, h->d_rhs.p + Np, h->red_um, h->red_xa, h->red_xb, h->red_d, mtol, &r2u));
      // K u = row 2 of J (0, u, 0);  w = M^-1 (K u - H)
      PFC_CUDA(h, cudaMemsetAsync(t3, 0, h->stride * sizeof(double), s));
      PFC_CUDA(h, cudaMemcpyAsync(t3 + Np, h->red_um, Np * sizeof(double), cudaMemcpyDeviceToDevice, s));
      PFC_TRY(halo_exchange(h, t3));
      {
        LatOp o = lat_op(h);
        o.in3a = t3;
        o.in3b = h->d_psi_n.p;
        o.out3 = y3;
        PFC_TRY(lat_launch(h, LAT_J, o, nullptr));
      }
      axpby_kernel<<<grid, 256, 0, s>>>(h->red_y, 1.0, y3 + 2 * Np, 0.0, Np, nullptr);             // K u
      axpby_kernel<<<grid, 256, 0, s>>>(h->red_y, -1.0, h->d_rhs.p + 2 * Np, 1.0, Np, nullptr);   // K u - H
      h->launches += 2;
      PFC_TRY(mass_solve(h, h->red_y, h->red_wm, h->red_xa, h->red_xb, h->red_d, mtol, &r2w));
      // f = F - (C u - D w) = F - row 0 of J (0, u, -w); rows 1, 2 of the same product give the residuals left in G, H
      axpby_kernel<<<grid, 256, 0, s>>>(t3 + 2 * Np, -1.0, h->red_wm, 0.0, Np, nullptr);
      h->launches++;
      PFC_TRY(halo_exchange(h, t3));
      {
        LatOp o = lat_op(h);
        o.in3a = t3;
        o.in3b = h->d_psi_n.p;
        o.out3 = y3;
        PFC_TRY(lat_launch(h, LAT_J, o, nullptr));
      }
      axpby_kernel<<<grid, 256, 0, s>>>(h->red_f, 1.0, h->d_rhs.p, 0.0, Np, nullptr);
      axpby_kernel<<<grid, 256, 0, s>>>(h->red_f, -1.0, y3, 1.0, Np, nullptr);
      axpby_kernel<<<vgrid(h, 2 * Np), 256, 0, s>>>(y3 + Np, 1.0, h->d_rhs.p + Np, -1.0, 2 * Np, nullptr); // (G, H) - rows 1, 2
      h->launches += 3;
      PFC_TRY(norm2(h, y3 + Np, 2 * Np, &h->red_aux2));
    }
  PFC_CUDA(h, cudaGetLastError());
  if (!(h->red_aux2 < tol0 * tol0))
    return h->fail(PFC_ERR_GMRES_NOT_CONVERGED, "reduced solve: the auxiliary rows keep a residual of %.3e (tolerance %.1e)",
                   std::sqrt(h->red_aux2), tol0);
  *tol = std::sqrt(tol0 * tol0 - h->red_aux2);
  return PFC_OK;
}

// After GMRES: dPsi = (M^2 y, u - K M y, K^2 y - w) into the solver's update vector (owned rows; pfc_solve_update
// exchanges, distributes and adds it)
int reduced_finish(pfc_handle *h)
{
  const int64_t Np = h->plan.Np;
  cudaStream_t s = h->stream;
  PFC_TRY(halo_exchange_scalar(h, h->red_y));
  LatOp a = lat_op(h);
  a.s0 = h->red_y;
  a.out3 = h->d_x.p;
  PFC_TRY(lat_launch(h, LAT_P, a, nullptr));
  if (h->red_have_uw)
    {
      const int grid = vgrid(h, Np);
      axpby_kernel<<<grid, 256, 0, s>>>(h->d_x.p + Np, 1.0, h->red_um, 1.0, Np, nullptr);
      axpby_kernel<<<grid, 256, 0, s>>>(h->d_x.p + 2 * Np, -1.0, h->red_wm, 1.0, Np, nullptr);
      h->launches += 2;
    }
  PFC_CUDA(h, cudaGetLastError());
  return PFC_OK;
}

// bench.py roofline legs of the reduced solve (include/pfc_b200.h), device time per launch replayed from a CUDA graph,
// operands rotating through the Krylov basis (scalar fields, sstride apart) so that they do not stay in L2:
//   0 operator T y (both kernels)   1 lat_pstage_kernel   2 lat_j1_kernel<ROW0>
//   3 preconditioner (exchange + sweeps)   4 first sweep (ring_sweep1a_kernel<FIRST>)   5 a later sweep (0 if sweeps < 3)
extern "C" int pfc_time_reduced(pfc_handle *h, int32_t reps, double *ms6, double *bytes6)
{
  if (!h || reps < 1 || !ms6)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || !h->reduced || !h->has_state)
    return h->fail(PFC_ERR_STATE, "pfc_time_reduced needs an active reduced solve (PFC_PRECOND_SCHUR_REDUCED on the Q1 lattice path) and a state");
  PFC_CUDA(h, cudaSetDevice(h->device));
  cudaStream_t s = h->stream;
  const int half = std::max(1, h->n_basis / 2), sw = h->sweeps;
  const double N = (double)h->lat_geom.w * h->lat_geom.h;
  auto vin = [&](int i) { return h->d_V.p + (int64_t)(i % half) * h->sstride; };
  auto vout = [&](int i) { return h->d_V.p + (int64_t)(half + i % half) * h->sstride; };
  for (int leg = 0; leg < 6; ++leg)
    {
      ms6[leg] = 0;
      if (bytes6)
        bytes6[leg] = 0;
      if ((leg >= 4 && sw < 2) || (leg == 5 && sw < 3))
        continue;
      cudaGraph_t graph = nullptr;
      cudaGraphExec_t exec = nullptr;
      const int64_t before = h->launches;
      PFC_CUDA(h, cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
      int rc = PFC_OK;
      for (int i = 0; i < reps && rc == PFC_OK; ++i)
        {
          double *z = nullptr;
          RingOp o;
          std::memset(&o, 0, sizeof(o));
          o.bbar = 1.0, o.dtth = h->prm.dt * h->prm.theta, o.epi = 0;
          LatOp a = lat_op(h);
          switch (leg)
            {
            case 0: rc = reduced_apply(h, vin(i), vout(i), nullptr, false); break;
            case 1: a.s0 = vin(i), a.out3 = h->d_z.p, rc = lat_launch(h, LAT_P, a, nullptr); break;
            case 2: a.in3a = h->d_V.p + (int64_t)(i % 4) * h->stride, a.in3b = h->d_psi_n.p, a.out3 = vout(i);
              rc = lat_launch(h, LAT_J0, a, nullptr);
              break;
            case 3: rc = reduced_precond(h, vin(i), &z, nullptr); break;
            case 4:
              o.rhs = vin(i), o.cheb = h->d_cheb_red.p + 2, o.lo = 3, o.first = 1, o.alpha = h->red_alpha, o.dinvc = h->red_dinvc;
              o.d = sw > 2 ? h->red_d : nullptr, o.out = vout(i);
              rc = ring_launch_sweep(h, o, nullptr);
              break;
            default:
              o.x = vin(i), o.rhs = vin(i + 1), o.dinv = h->d_dinvS.p, o.d = h->red_d, o.cheb = h->d_cheb_red.p + 4, o.lo = 6, o.out = vout(i);
              rc = ring_launch_sweep(h, o, nullptr);
              break;
            }
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
      PFC_CUDA(h, cudaGraphLaunch(exec, s)); // warm-up
      PFC_CUDA(h, cudaEventRecord(h->ev[0], s));
      PFC_CUDA(h, cudaGraphLaunch(exec, s));
      PFC_CUDA(h, cudaEventRecord(h->ev[1], s));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      float ms = 0;
      PFC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
      cudaGraphExecDestroy(exec);
      ms6[leg] = ms / reps;
      if (bytes6)
        {
          // algorithmic bytes per node: z in + three fields out | three fields + psi, chi in, one out | r in, x out
          // (+ d out when a sweep follows) | x, rhs, dinv, d in, d, x out
          const double per_node[6] = {32.0 + 48.0, 32.0, 48.0, 0.0, sw > 2 ? 24.0 : 16.0, 48.0};
          bytes6[leg] = per_node[leg] * N;
        }
    }
  // the 3-component test operands of leg 2 and the rotation leave the basis dirty; the next solve rebuilds what it reads
  return PFC_OK;
}
