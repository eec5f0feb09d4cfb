// pfc_stress.cu — L2 projection of the PFC stress tensor on the lattice path.
// Replaces StressCalculatorMPI<dim>::calculate_stress (src/stress_tools/stress_calculator_mpi.cpp:345-353):
//   calculate_mass_matrix  :141-200   -> consistent mass matrix as a node stencil (LAT_M), never stored
//   calculate_righthand_side :204-314 -> one thread-per-cell kernel for all dim*dim components (LAT_T)
//   solve                  :317-341   -> the dim*dim unpreconditioned CG solves run as ONE batched iteration
//                                        (independent alpha/beta per component, converged components masked
//                                        out, so every component sees exactly its own CG iterates), then
//                                        constraints.distribute
// The state is read where the time stepper left it in HBM; only the 4 N results travel to the host.
#include <algorithm>
#include <cmath>
#include <vector>

#include "pfc_comm.h"
#include "pfc_handle.h"
#include "pfc_krylov.cuh"
#include "pfc_lattice_args.h"

using namespace pfc;

#define PFC_TRY(expr)    \
  do                     \
    {                    \
      int rc_ = (expr);  \
      if (rc_ != PFC_OK) \
        return rc_;      \
    }                    \
  while (0)

namespace
{
// device-resident state of the batched CG: the host only polls it every few iterations
struct StressCtl
{
  double gh[4];       // g_c . g_c of the previous iteration
  double beta[4];
  long long iters[4]; // SolverControl::last_step() per component
  int mask;           // bit c set = component c still iterates
  int it;
  int fail;           // 1: max iterations, 2: non-finite
  int fail_comp;
  double fail_res;
};

// g = -b, d = b (b arrives in d), out[c] = b_c . b_c
__global__ void __launch_bounds__(kVecThreads) stress_init_kernel(double *__restrict__ g, const double *__restrict__ d,
                                                                  int64_t zstride, int64_t n, double *partial,
                                                                  double *out, unsigned int *counter)
{
  double acc[4] = {0, 0, 0, 0};
  for (int64_t i = (int64_t)blockIdx.x * kVecThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kVecThreads)
#pragma unroll
    for (int c = 0; c < 4; ++c)
      {
        const double b = d[c * zstride + i];
        g[c * zstride + i] = -b;
        acc[c] = fma(b, b, acc[c]);
      }
  grid_reduce<4>(acc, partial, out, counter, ArDev{});
}

// out[c] = d_c . h_c for the components still iterating
__global__ void __launch_bounds__(kVecThreads) stress_dot_kernel(const double *__restrict__ d, const double *__restrict__ hh,
                                                                 int64_t zstride, int64_t n, const StressCtl *ctl,
                                                                 double *partial, double *out, unsigned int *counter)
{
  const int mask = ctl->mask;
  if (!mask)
    return;
  double acc[4] = {0, 0, 0, 0};
  for (int64_t i = (int64_t)blockIdx.x * kVecThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kVecThreads)
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if ((mask >> c) & 1)
        acc[c] = fma(d[c * zstride + i], hh[c * zstride + i], acc[c]);
  grid_reduce<4>(acc, partial, out, counter, ArDev{});
}

// alpha_c = gh_c / dh_c;  x_c += alpha_c d_c;  g_c += alpha_c h_c;  out[c] = g_c . g_c
__global__ void __launch_bounds__(kVecThreads) stress_update_kernel(double *__restrict__ x, double *__restrict__ g,
                                                                    const double *__restrict__ d,
                                                                    const double *__restrict__ hh, int64_t zstride,
                                                                    int64_t n, const StressCtl *ctl, const double *dh,
                                                                    double *partial, double *out, unsigned int *counter)
{
  const int mask = ctl->mask;
  if (!mask)
    return;
  double acc[4] = {0, 0, 0, 0}, alpha[4];
#pragma unroll
  for (int c = 0; c < 4; ++c)
    alpha[c] = ((mask >> c) & 1) ? ctl->gh[c] / dh[c] : 0.0;
  for (int64_t i = (int64_t)blockIdx.x * kVecThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kVecThreads)
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if ((mask >> c) & 1)
        {
          const int64_t k = c * zstride + i;
          x[k] = fma(alpha[c], d[k], x[k]);
          const double gv = fma(alpha[c], hh[k], g[k]);
          g[k] = gv;
          acc[c] = fma(gv, gv, acc[c]);
        }
  grid_reduce<4>(acc, partial, out, counter, ArDev{});
}

// SolverControl::check per component (stress_calculator_mpi.cpp:320, :329-334), then beta for the next direction
__global__ void stress_ctl_kernel(StressCtl *ctl, const double *rr, double tol, long long max_iters, int first)
{
  if (first)
    {
      ctl->it = 0;
      ctl->fail = 0;
      ctl->mask = 0;
    }
  else if (!ctl->mask)
    return;
  else
    ctl->it++;
  const int it = ctl->it;
  int mask = first ? 15 : ctl->mask;
  for (int c = 0; c < 4; ++c)
    {
      if (!((mask >> c) & 1))
        continue;
      const double r2 = rr[c], res = sqrt(r2);
      ctl->beta[c] = 0.0;
      if (res <= tol)
        {
          ctl->iters[c] = it;
          mask &= ~(1 << c);
        }
      else if (it >= max_iters || !isfinite(res))
        {
          ctl->fail = isfinite(res) ? 1 : 2;
          ctl->fail_comp = c;
          ctl->fail_res = res;
          mask = 0;
          break;
        }
      else
        {
          if (!first)
            ctl->beta[c] = r2 / ctl->gh[c];
          ctl->gh[c] = r2;
        }
    }
  ctl->mask = mask;
}

// d_c = beta_c d_c - g_c
__global__ void __launch_bounds__(kVecThreads) stress_dir_kernel(double *__restrict__ d, const double *__restrict__ g,
                                                                 int64_t zstride, int64_t n, const StressCtl *ctl)
{
  const int mask = ctl->mask;
  if (!mask)
    return;
  for (int64_t i = (int64_t)blockIdx.x * kVecThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kVecThreads)
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if ((mask >> c) & 1)
        {
          const int64_t k = c * zstride + i;
          d[k] = fma(ctl->beta[c], d[k], -g[k]);
        }
}

// constraints.distribute (stress_calculator_mpi.cpp:336): v_c[slave] = v_c[master]
__global__ void stress_distribute_kernel(double *v, const int32_t *slave, const int32_t *master, int64_t n, int64_t zstride)
{
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n)
    return;
#pragma unroll
  for (int c = 0; c < 4; ++c)
    v[c * zstride + slave[k]] = v[c * zstride + master[k]];
}

constexpr int kPollEvery = 8; // CG iterations between two looks at the control block
} // namespace

static int stress_impl(pfc_handle *h, double eps, double tol, double *stress, int64_t *cg_iters, bool owned)
{
  if (!h || !stress || !cg_iters)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || !h->has_state)
    return h->fail(PFC_ERR_STATE, "pfc_stress needs pfc_setup and pfc_set_state first");
  if (!h->lat)
    return h->fail(PFC_ERR_UNSUPPORTED,
                   "pfc_stress runs on the lattice path (operator_kind = PFC_OPERATOR_MATRIX_FREE on a uniform periodic mesh)");
  PFC_CUDA(h, cudaSetDevice(h->device));
  const Plan &P = h->plan;
  cudaStream_t s = h->stream;
  const int64_t ss = (P.Np + P.n_ghost + 15) / 16 * 16, Np = P.Np;
  const int64_t N = P.n_dofs / 3;
  constexpr int64_t kTail = 32; // doubles after the fields: dh[4], rr[4], StressCtl
  static_assert(sizeof(StressCtl) <= (kTail - 8) * sizeof(double), "control block does not fit");
  if (h->d_stress.n != (size_t)(16 * ss + kTail))
    {
      PFC_CUDA(h, h->d_stress.alloc((size_t)(16 * ss + kTail)));
      h->stress_map_owned = -1;
    }
  const int64_t n_out = owned ? 4 * P.n_owned : 4 * N;
  if (h->stress_map_owned != (owned ? 1 : 0))
    {
      // output position -> address in a batched scalar field; owned nodes only (like pfc_get_state):
      // API order [4][N] by global node, or [4][n_owned] by owned node
      std::vector<int64_t> map((size_t)n_out, -1);
      if (owned)
        {
          for (int c = 0; c < 4; ++c)
            for (int64_t j = 0; j < P.n_owned; ++j)
              map[c * P.n_owned + j] = c * ss + P.iperm[j];
        }
      else
        for (int64_t r = 0; r < Np; ++r)
          if (P.perm[r] >= 0)
            for (int c = 0; c < 4; ++c)
              map[c * N + P.perm[r]] = c * ss + r;
      PFC_CUDA(h, h->d_stress_map.upload(map, s));
      PFC_CUDA(h, h->d_stress_io.alloc((size_t)n_out));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      h->stress_map_owned = owned ? 1 : 0;
    }
  h->stress_valid = false;
  double *x = h->d_stress.p, *g = x + 4 * ss, *d = x + 8 * ss, *hh = x + 12 * ss, *small = x + 16 * ss;
  double *dh = small, *rr = small + 4;
  StressCtl *ctl = reinterpret_cast<StressCtl *>(small + 8);
  PFC_CUDA(h, cudaMemsetAsync(x, 0, (size_t)(16 * ss + kTail) * sizeof(double), s));
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((Np + kVecThreads - 1) / kVecThreads, 4 * (int64_t)h->n_sm));

  // right-hand side (all four components in one pass over the state), into d
  PFC_TRY(halo_exchange(h, h->d_psi_n.p));
  {
    LatOp o = lat_op(h);
    o.eps = eps;
    o.in3a = h->d_psi_n.p;
    o.out = d;
    o.zstride = ss;
    PFC_TRY(lat_launch(h, LAT_T, o, nullptr));
  }
  // SolverControl(dof_handler.n_dofs()) of the dim*dim-component handler, default tolerance (stress.cpp:320)
  const long long max_iters = 4 * (long long)N;
  stress_init_kernel<<<grid, kVecThreads, 0, s>>>(g, d, ss, Np, h->d_partial.p, rr, h->d_counter.p);
  PFC_TRY(allreduce_sum(h, rr, 4));
  stress_ctl_kernel<<<1, 1, 0, s>>>(ctl, rr, tol, max_iters, 1);
  h->launches += 2;
  StressCtl host{};
  auto poll = [&]() -> int {
    PFC_CUDA(h, cudaMemcpyAsync(&host, ctl, sizeof(StressCtl), cudaMemcpyDeviceToHost, s));
    PFC_CUDA(h, cudaStreamSynchronize(s));
    return PFC_OK;
  };
  PFC_TRY(poll());
  while (host.mask && !host.fail)
    {
      // the host's mask is a superset of the device's (components only ever leave): exchange and launch for it,
      // the kernels themselves follow the device mask
      for (int k = 0; k < kPollEvery; ++k)
        {
          for (int c = 0; c < 4; ++c) // ring values of the search directions (single rank: periodic wrap-around copies)
            if ((host.mask >> c) & 1)
              PFC_TRY(halo_exchange_scalar(h, d + c * ss));
          LatOp o = lat_op(h);
          o.s0 = d;
          o.out = hh;
          o.zstride = ss;
          o.zmask = host.mask;
          o.zmask_dev = &ctl->mask;
          PFC_TRY(lat_launch(h, LAT_M, o, nullptr));
          stress_dot_kernel<<<grid, kVecThreads, 0, s>>>(d, hh, ss, Np, ctl, h->d_partial.p, dh, h->d_counter.p);
          PFC_TRY(allreduce_sum(h, dh, 4));
          stress_update_kernel<<<grid, kVecThreads, 0, s>>>(x, g, d, hh, ss, Np, ctl, dh, h->d_partial.p, rr, h->d_counter.p);
          PFC_TRY(allreduce_sum(h, rr, 4));
          stress_ctl_kernel<<<1, 1, 0, s>>>(ctl, rr, tol, max_iters, 0);
          stress_dir_kernel<<<grid, kVecThreads, 0, s>>>(d, g, ss, Np, ctl);
          h->launches += 4;
        }
      PFC_TRY(poll());
    }
  if (host.fail == 2)
    return h->fail(PFC_ERR_NONFINITE, "pfc_stress: non-finite residual (component %d)", host.fail_comp);
  if (host.fail)
    return h->fail(PFC_ERR_GMRES_NOT_CONVERGED, "pfc_stress: CG did not converge in %d iterations (component %d, residual %.3e)",
                   host.it, host.fail_comp, host.fail_res);
  for (int c = 0; c < 4; ++c)
    cg_iters[c] = host.iters[c];
  // constraints.distribute, then the API order
  for (int c = 0; c < 4; ++c)
    PFC_TRY(halo_exchange_scalar(h, x + c * ss));
  const int64_t nsl = (int64_t)P.slave_id.size();
  if (nsl > 0)
    {
      stress_distribute_kernel<<<(int)((nsl + 255) / 256), 256, 0, s>>>(x, h->d_slave.p, h->d_slave_master.p, nsl, ss);
      h->launches++;
    }
  PFC_CUDA(h, cudaMemsetAsync(h->d_stress_io.p, 0, (size_t)n_out * sizeof(double), s));
  gather_kernel<<<(int)((n_out + 255) / 256), 256, 0, s>>>(h->d_stress_io.p, x, h->d_stress_map.p, n_out, 4 * ss);
  h->launches++;
  PFC_CUDA(h, cudaGetLastError());
  PFC_CUDA(h, cudaMemcpyAsync(stress, h->d_stress_io.p, (size_t)n_out * sizeof(double), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  h->stress_valid = true; // x (owned + ghost, distributed) stays in d_stress for pfc_output
  return PFC_OK;
}

extern "C" int pfc_stress(pfc_handle *h, double eps, double tol, double *stress, int64_t *cg_iters)
{
  return stress_impl(h, eps, tol, stress, cg_iters, false);
}
extern "C" int pfc_stress_owned(pfc_handle *h, double eps, double tol, double *stress, int64_t *cg_iters)
{
  return stress_impl(h, eps, tol, stress, cg_iters, true);
}

// ---- output: download owned + ghost values, hand them to the host VTU writer ----
// kind 0: Psi_n as "psi", "chi", "phi" (output_configuration, mpi.cpp:597-619); 1: the assembled right-hand side under
// the same names (output_rhs, mpi.cpp:624-645); 2: the result of the last pfc_stress as tensor "stress"
// (StressCalculatorMPI::output_stress, stress.cpp:358-384).
extern "C" int pfc_output(pfc_handle *h, const pfc_mesh *mesh, int32_t kind, const char *directory, const char *basename,
                          uint32_t counter)
{
  if (!h || !mesh || !directory || !basename || kind < 0 || kind > 2)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || (kind == 0 && !h->has_state))
    return h->fail(PFC_ERR_STATE, "pfc_output needs pfc_setup and pfc_set_state first");
  if (kind == 1 && !h->rhs_valid)
    return h->fail(PFC_ERR_STATE, "pfc_output(rhs) before the first assembly");
  if (kind == 2 && !h->stress_valid)
    return h->fail(PFC_ERR_STATE, "pfc_output(stress) before pfc_stress");
  PFC_CUDA(h, cudaSetDevice(h->device));
  const Plan &P = h->plan;
  const int64_t N = P.n_dofs / 3, Np = P.Np, ss = (P.Np + P.n_ghost + 15) / 16 * 16;
  if (mesh->block_size[0] != N || mesh->rank != P.rank || mesh->n_ranks != P.n_ranks)
    return h->fail(PFC_ERR_BAD_ARGUMENT, "pfc_output: not the mesh pfc_setup was given");
  const int nf = kind == 2 ? 4 : 3;
  std::vector<double> values((size_t)nf * N, 0.0), dev;
  if (kind == 2)
    {
      dev.resize((size_t)4 * ss);
      PFC_CUDA(h, cudaMemcpyAsync(dev.data(), h->d_stress.p, dev.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
      PFC_CUDA(h, cudaStreamSynchronize(h->stream));
      for (int c = 0; c < 4; ++c)
        {
          for (int64_t r = 0; r < Np; ++r)
            if (P.perm[r] >= 0)
              values[c * N + P.perm[r]] = dev[c * ss + r];
          for (int64_t g = 0; g < P.n_ghost; ++g)
            if (P.ghost_global[g] >= 0) // ring overlay: ring positions beyond the handed-over cells have no global id
              values[c * N + P.ghost_global[g]] = dev[c * ss + Np + g];
        }
    }
  else
    {
      double *src = kind == 0 ? h->d_psi_n.p : h->d_rhs.p;
      if (kind == 1) // ghost rows of the right-hand side live on their owners
        PFC_TRY(halo_exchange(h, src));
      dev.resize((size_t)P.Lv);
      PFC_CUDA(h, cudaMemcpyAsync(dev.data(), src, dev.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
      PFC_CUDA(h, cudaStreamSynchronize(h->stream));
      for (int c = 0; c < 3; ++c)
        {
          for (int64_t r = 0; r < Np; ++r)
            if (P.perm[r] >= 0)
              values[c * N + P.perm[r]] = dev[c * Np + r];
          for (int64_t g = 0; g < P.n_ghost; ++g)
            if (P.ghost_global[g] >= 0)
              values[c * N + P.ghost_global[g]] = dev[3 * (Np + g) + c];
        }
    }
  // slaves of other ranks that sit on this rank's cells (ring overlay: a ring position carries the id of the node it
  // resolves to, not the slave's): constraints.distribute on the host, value of the master (mpi.cpp:609, stress.cpp:366)
  if (kind != 1)
    {
      std::vector<char> known((size_t)N, 0);
      for (int64_t r = 0; r < Np; ++r)
        if (P.perm[r] >= 0)
          known[P.perm[r]] = 1;
      for (int64_t g = 0; g < P.n_ghost; ++g)
        if (P.ghost_global[g] >= 0)
          known[P.ghost_global[g]] = 1;
      for (int64_t k = 0; k < mesh->n_constraints; ++k)
        {
          const int64_t sl = mesh->constraint_slave[k], ms = mesh->constraint_master[k];
          if (sl < N && ms < N && !known[sl] && known[ms])
            for (int c = 0; c < nf; ++c)
              values[c * N + sl] = values[c * N + ms];
        }
    }
  static const char *const state_names[3] = {"psi", "chi", "phi"};
  static const char *const stress_names[4] = {"stress", "stress", "stress", "stress"};
  static const int32_t scalar3[3] = {0, 0, 0}, tensor4[4] = {1, 1, 1, 1};
  char err[400] = {0};
  const int rc = pfc_vtu_write(mesh, directory, basename, counter, nf, kind == 2 ? stress_names : state_names,
                               kind == 2 ? tensor4 : scalar3, values.data(), err, sizeof(err));
  if (rc != PFC_OK)
    return h->fail(rc, "%s", err);
  return PFC_OK;
}
