// pfc_solver.cu — drivers of the hot path: assemble_system (mpi.cpp:300-467), the GMRES solve
// of solve_and_update_no_precondition (mpi.cpp:471-492) and the Newton loop of
// iterate_timestep (mpi.cpp:576-593), all stream-ordered on the handle's stream.
#include <cmath>
#include <cstring>

#include "pfc_comm.h"
#include "pfc_handle.h"
#include "pfc_matfree.cuh"

using namespace pfc;

#define PFC_TRY(expr)            \
  do                             \
    {                            \
      int rc_ = (expr);          \
      if (rc_ != PFC_OK)         \
        return rc_;              \
    }                            \
  while (0)

static inline int vec_grid(const pfc_handle *h, int64_t n2)
{
  const int64_t need = (n2 + kVecThreads - 1) / kVecThreads;
  const int64_t cap = (int64_t)h->n_sm * 4; // persistent grid-stride: 4 CTAs of 256 threads per SM
  return (int)std::max<int64_t>(1, std::min<int64_t>(need, std::min<int64_t>(cap, 4096)));
}

// ---- assembly --------------------------------------------------------------------------------------
template <int NS, bool CONST_PASS>
static void launch_asm_colors(pfc_handle *h, const AsmArgs &A)
{
  const Plan &P = h->plan;
  constexpr int cells_per_block = (32 / NS) * (kAsmThreads / 32);
  for (int c = 0; c < P.n_colors; ++c)
    {
      const int64_t b = P.color_ptr[c], e = P.color_ptr[c + 1];
      if (e <= b)
        continue;
      const int blocks = (int)((e - b + cells_per_block - 1) / cells_per_block);
      assemble_kernel<NS, CONST_PASS><<<blocks, kAsmThreads, 0, h->stream>>>(A, b, e);
      h->launches++;
    }
}

int precond_refresh(pfc_handle *h);                                                    // pfc_precond.cu
int precond_apply(pfc_handle *h, double *in, double *out, const pfc::Ctl *ctl); // pfc_precond.cu

static inline bool is_general_mf(const pfc_handle *h)
{
  return !h->lat && (h->prm.operator_kind == PFC_OPERATOR_MATRIX_FREE || h->prm.operator_kind == PFC_OPERATOR_MATRIX_FREE_GENERAL);
}

int launch_assembly(pfc_handle *h, bool const_pass)
{
  const Plan &P = h->plan;
  if (h->lat)
    { // -R by the lattice cell kernel; 1 / diag(B) for the Jacobi scaling of the preconditioner on the way
      if (const_pass)
        return PFC_OK;
      LatOp op = lat_op(h);
      op.in3a = h->d_psi_n.p;
      op.in3b = h->d_psi_n1.p;
      op.out3 = h->d_rhs.p;
      op.out = h->prm.preconditioner == PFC_PRECOND_BLOCK_SCHUR ? h->d_dinvB.p : nullptr;
      return lat_launch(h, LAT_R, op, nullptr);
    }
  AsmArgs A;
  A.cell_gather = h->d_cell_gather.p;
  A.cell_row = h->d_cell_row.p;
  A.cell_pos = h->d_cell_pos.p;
  A.cell_spos = h->d_cell_spos.p;
  A.cell_h = h->d_cell_h.p;
  A.psi_n = h->d_psi_n.p;
  A.psi_n1 = h->d_psi_n1.p;
  A.B = h->d_B.p;
  A.C = h->d_C.p;
  A.M = h->d_M.p;
  A.K = h->d_K.p;
  A.rhs = h->d_rhs.p;
  A.tab = h->d_tab.p;
  A.Np = P.Np;
  A.dt = h->prm.dt;
  A.theta = h->prm.theta;
  A.eps = h->prm.eps;
  // matrix-free operator without preconditioner: B and C are never read
  A.rhs_only = (!const_pass && is_general_mf(h) && h->prm.preconditioner == PFC_PRECOND_NONE) ? 1 : 0;
  cudaStream_t s = h->stream;
  if (const_pass)
    {
      PFC_CUDA(h, cudaMemsetAsync(h->d_M.p, 0, h->d_M.n * sizeof(double), s));
      PFC_CUDA(h, cudaMemsetAsync(h->d_K.p, 0, h->d_K.n * sizeof(double), s));
      if (P.n_s == 4)
        launch_asm_colors<4, true>(h, A);
      else
        launch_asm_colors<9, true>(h, A);
    }
  else
    {
      // system_matrix = 0; system_rhs = 0 (mpi.cpp:305-306)
      PFC_CUDA(h, cudaMemsetAsync(h->d_B.p, 0, h->d_B.n * sizeof(double), s));
      PFC_CUDA(h, cudaMemsetAsync(h->d_C.p, 0, h->d_C.n * sizeof(double), s));
      PFC_CUDA(h, cudaMemsetAsync(h->d_rhs.p, 0, h->d_rhs.n * sizeof(double), s));
      if (P.n_s == 4)
        launch_asm_colors<4, false>(h, A);
      else
        launch_asm_colors<9, false>(h, A);
    }
  PFC_CUDA(h, cudaGetLastError());
  return PFC_OK;
}

// ---- operator application ----------------------------------------------------------------------------
static int launch_spmv(pfc_handle *h, const double *x, double *y, const Ctl *ctl)
{
  const Plan &P = h->plan;
  SpmvArgs S;
  S.slice_ptr = h->d_slice_ptr.p;
  S.col = h->d_col.p;
  S.B = h->d_B.p;
  S.C = h->d_C.p;
  S.M = h->d_M.p;
  S.K = h->d_K.p;
  S.Np = P.Np;
  S.n_slices = P.n_slices;
  S.dtth = h->prm.dt * h->prm.theta;
  const int blocks = (int)((P.n_slices * 32 + kSpmvThreads - 1) / kSpmvThreads);
  spmv_sell_kernel<<<blocks, kSpmvThreads, 0, h->stream>>>(S, x, y, ctl);
  h->launches++;
  return PFC_OK;
}

// matrix-free y = J x: zero y, then one launch per colour (pfc_matfree.cuh)
template <int NS>
static int launch_mf(pfc_handle *h, const double *x, double *y, const Ctl *ctl)
{
  const Plan &P = h->plan;
  MfArgs A;
  A.cell_gather = h->d_cell_gather.p;
  A.cell_col = h->d_cell_col.p;
  A.cell_row = h->d_cell_row.p;
  A.cell_spos = h->d_cell_spos.p;
  A.cell_h = h->d_cell_h.p;
  A.psi_n = h->d_psi_n.p;
  A.tab = h->d_tab.p;
  A.Np = P.Np;
  A.dtth = h->prm.dt * h->prm.theta;
  A.eps = h->prm.eps;
  PFC_CUDA(h, cudaMemsetAsync(y, 0, 3 * P.Np * sizeof(double), h->stream));
  constexpr int cells_per_block = (32 / NS) * (kAsmThreads / 32);
  for (int c = 0; c < P.n_colors; ++c)
    {
      const int64_t b = P.color_ptr[c], e = P.color_ptr[c + 1];
      if (e <= b)
        continue;
      const int blocks = (int)((e - b + cells_per_block - 1) / cells_per_block);
      mf_apply_kernel<NS><<<blocks, kAsmThreads, 0, h->stream>>>(A, x, y, b, e, ctl);
      h->launches++;
    }
  return PFC_OK;
}

// y = J x with the halo of x refreshed first (multi-rank)
static int apply_operator(pfc_handle *h, double *x, double *y, const Ctl *ctl, bool halo = true)
{
  if (halo)
    PFC_TRY(halo_exchange(h, x, ctl));
  if (h->lat)
    {
      LatOp op = lat_op(h);
      op.in3a = x;
      op.in3b = h->d_psi_n.p;
      op.out3 = y;
      return lat_launch(h, LAT_J, op, ctl);
    }
  if (is_general_mf(h))
    return h->plan.n_s == 4 ? launch_mf<4>(h, x, y, ctl) : launch_mf<9>(h, x, y, ctl);
  return launch_spmv(h, x, y, ctl);
}

// ---- what a GMRES solve works on ---------------------------------------------------------------------
// The 3-block Newton system (vectors of 3 Np doubles, operator J, block-Schur right preconditioner), or the psi-only
// reduced system T y = f of PFC_PRECOND_SCHUR_REDUCED (pfc_reduced.cu: scalar fields of Np doubles, always preconditioned).
struct KrylovSys
{
  bool reduced, precond;
  int64_t n;  // doubles of the owned part of a vector (even)
  int64_t st; // doubles between basis vectors
  double *V, *x, *z, *u;
  const double *rhs;
};
int reduced_apply(pfc_handle *h, double *y, double *out, const pfc::Ctl *ctl, bool halo); // pfc_reduced.cu
int reduced_precond(pfc_handle *h, double *in, double **out, const pfc::Ctl *ctl);        // pfc_reduced.cu
int reduced_prepare(pfc_handle *h, double *tol);                                          // pfc_reduced.cu
int reduced_finish(pfc_handle *h);                                                        // pfc_reduced.cu

static KrylovSys krylov_sys(pfc_handle *h)
{
  KrylovSys K;
  K.reduced = h->reduced;
  if (h->reduced)
    {
      K.precond = true;
      K.n = h->plan.Np;
      K.st = h->sstride;
      K.V = h->d_V.p;
      K.x = h->red_y;
      K.z = nullptr; // the preconditioner names its output field
      K.u = h->red_u;
      K.rhs = h->red_f;
    }
  else
    {
      K.precond = h->prm.preconditioner == PFC_PRECOND_BLOCK_SCHUR;
      K.n = 3 * h->plan.Np;
      K.st = h->stride;
      K.V = h->d_V.p;
      K.x = h->d_x.p;
      K.z = h->d_z.p;
      K.u = h->d_u.p;
      K.rhs = h->d_rhs.p;
    }
  return K;
}
// out = A in (halo: refresh the ghosts of `in` first)
static int sys_apply(pfc_handle *h, const KrylovSys &K, double *in, double *out, const Ctl *ctl, bool halo)
{
  return K.reduced ? reduced_apply(h, in, out, ctl, halo) : apply_operator(h, in, out, ctl, halo);
}
// *out = field holding P^-1 in
static int sys_precond(pfc_handle *h, const KrylovSys &K, double *in, double **out, const Ctl *ctl)
{
  if (K.reduced)
    return reduced_precond(h, in, out, ctl);
  *out = K.z;
  return precond_apply(h, in, K.z, ctl);
}

// ---- template dispatch over the Krylov dimension ---------------------------------------------------
using DotFn = void (*)(const double *, int64_t, const double *, int64_t, double *, double *, unsigned int *, const Ctl *, int, ArDev);
using AxpyFn = void (*)(const double *, int64_t, double *, int64_t, const double *, const double *, double *, double *, unsigned int *, Ctl *,
                        int, ArDev);
constexpr int kMaxDim = 32;
template <int D>
struct Tab
{
  static void fill(DotFn *d, AxpyFn *a)
  {
    d[D] = multidot_kernel<D>;
    a[D] = multiaxpy_kernel<D>;
    Tab<D - 1>::fill(d, a);
  }
};
template <>
struct Tab<0>
{
  static void fill(DotFn *d, AxpyFn *a)
  {
    d[0] = multidot_kernel<0>;
    a[0] = nullptr;
  }
};
static DotFn g_dot[kMaxDim + 1];
static AxpyFn g_axpy[kMaxDim + 1];
static int g_dot_occ[kMaxDim + 1], g_axpy_occ[kMaxDim + 1]; // resident CTAs per SM of every instantiation
static bool g_tab_ready = false;
static void ensure_tab()
{
  if (!g_tab_ready)
    {
      Tab<kMaxDim>::fill(g_dot, g_axpy);
      for (int d = 0; d <= kMaxDim; ++d)
        {
          int nd = 0, na = 0;
          if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nd, g_dot[d], kVecThreads, 0) != cudaSuccess || nd < 1)
            nd = 1;
          if (!g_axpy[d] || cudaOccupancyMaxActiveBlocksPerMultiprocessor(&na, g_axpy[d], kVecThreads, 0) != cudaSuccess || na < 1)
            na = 1;
          g_dot_occ[d] = nd;
          g_axpy_occ[d] = na;
        }
      cudaGetLastError();
      g_tab_ready = true;
    }
}
// The Gram-Schmidt kernels are persistent grid-stride loops: exactly one wave of CTAs.  Their register count, and with
// it the CTAs an SM holds, changes with the Krylov dimension (2 to 8); a grid of 4 per SM left a quarter of the vector
// to a second wave of one CTA per SM at the dimensions that fit only 3 (measured: 2.7 instead of 6.7 TB/s at dimension 24,
// profiles/r02o_krylov_dims.log).
static inline int gs_grid(const pfc_handle *h, int64_t n2, int occ)
{
  const int64_t need = (n2 + kVecThreads - 1) / kVecThreads;
  return (int)std::max<int64_t>(1, std::min<int64_t>(need, std::min<int64_t>((int64_t)h->n_sm * occ, 4096)));
}

// ---- one restart cycle, or the piece [i0, i1) of its iterations, enqueued on the stream (captured into a CUDA
//      graph when enabled).  i0 == 0 starts the cycle; every piece ends with the gated end part -----------------
static int enqueue_cycle(pfc_handle *h, bool first_cycle, bool two_pass, int i0, int i1)
{
  ensure_tab();
  cudaStream_t s = h->stream;
  const KrylovSys K = krylov_sys(h);
  const int64_t n2 = K.n / 2;
  const int grid = vec_grid(h, n2);
  Ctl *ctl = h->d_ctl.p;
  Ctl *gate = h->d_ctl_end.p;
  GmresSmall *g = h->d_small.p;
  double *V = K.V;
  const int64_t st = K.st;
  double *partial = h->d_partial.p;
  unsigned int *cnt = h->d_counter.p;
  const bool precond = K.precond;
  // peer-memory path: the reduction kernels all-reduce their results in their last block (no collective kernel)
  const ArDev ar = comm_ar_endpoint(h);
  const bool fused = ar.R > 1;
  // Lazy normalisation of the basis (GmresSmall::s2): no scale pass per iteration.  Preconditioned solves only; the
  // unpreconditioned path keeps the arithmetic of deal.II's SolverGMRES (the tests compare its iteration counts with the CPU restatement's).
  const bool lazy = precond;
  const double *is2 = lazy ? g->is2 : nullptr;

  if (i0 == 0)
    {
      // v0 = b - A x ; rho = ||v0|| ; check ; v0 /= rho          ([dealii-ext] start of the outer loop)
      if (!first_cycle)
        PFC_TRY(sys_apply(h, K, K.x, V, ctl, true));
      residual_kernel<<<grid, kVecThreads, 0, s>>>(V, K.rhs, n2, first_cycle ? 0 : 1, partial, &g->nrm2[0], cnt, ctl, ar);
      h->launches++;
      if (!fused)
        PFC_TRY(allreduce_sum(h, &g->nrm2[0], 1, ctl));
      cycle_begin_kernel<<<1, 1, 0, s>>>(g, ctl);
      scale_kernel<<<grid, kVecThreads, 0, s>>>(V, n2, ctl);
      h->launches += 2;
    }

  for (int inner = i0; inner < i1; ++inner)
    {
      const int dim = inner + 1;
      double *w = V + (int64_t)dim * st;
      if (precond)
        { // right preconditioning: w = A P^-1 v_inner
          double *z = nullptr;
          PFC_TRY(sys_precond(h, K, V + (int64_t)inner * st, &z, ctl));
          // lattice path: P^-1 v is already valid beyond the owned nodes (ring overlay), no exchange
          PFC_TRY(sys_apply(h, K, z, w, ctl, !h->lat));
        }
      else
        PFC_TRY(sys_apply(h, K, V + (int64_t)inner * st, w, ctl, true)); // A.vmult
      g_dot[dim]<<<gs_grid(h, n2, g_dot_occ[dim]), kVecThreads, 0, s>>>(V, st, w, n2, partial, g->dots0, cnt, ctl, 0, ar);
      if (!fused)
        PFC_TRY(allreduce_sum(h, g->dots0, dim + 1, ctl));
      g_axpy[dim]<<<gs_grid(h, n2, g_axpy_occ[dim]), kVecThreads, 0, s>>>(V, st, w, n2, g->dots0, is2, partial, &g->nrm2[0], cnt, ctl, 0, ar);
      if (!fused)
        PFC_TRY(allreduce_sum(h, &g->nrm2[0], 1, ctl));
      if (two_pass)
        {
          g_dot[dim]<<<gs_grid(h, n2, g_dot_occ[dim]), kVecThreads, 0, s>>>(V, st, w, n2, partial, g->dots1, cnt, ctl, 1, ar);
          if (!fused)
            PFC_TRY(allreduce_sum(h, g->dots1, dim + 1, ctl));
          g_axpy[dim]<<<gs_grid(h, n2, g_axpy_occ[dim]), kVecThreads, 0, s>>>(V, st, w, n2, g->dots1, is2, partial, &g->nrm2[1], cnt, ctl, 1, ar);
          if (!fused)
            PFC_TRY(allreduce_sum(h, &g->nrm2[1], 1, ctl));
          h->launches += 2;
        }
      givens_kernel<<<1, 1, 0, s>>>(g, ctl, inner, two_pass ? 1 : 0, lazy ? 1 : 0);
      if (!lazy)
        scale_kernel<<<grid, kVecThreads, 0, s>>>(w, n2, ctl);
      h->launches += lazy ? 4 : 5;
    }
  // end part: runs when the solve has stopped or the restart length is reached, else the continuation piece follows
  end_gate_kernel<<<1, 1, 0, s>>>(ctl, gate, h->restart);
  backsolve_kernel<<<1, 1, 0, s>>>(g, ctl, gate, lazy ? 1 : 0);
  if (precond)
    { // x += P^-1 (V y)
      PFC_CUDA(h, cudaMemsetAsync(K.u, 0, K.n * sizeof(double), s));
      xupdate_kernel<<<grid, kVecThreads, 0, s>>>(V, st, K.u, n2, g, ctl, gate);
      double *z = nullptr;
      PFC_TRY(sys_precond(h, K, K.u, &z, gate));
      axpy1_kernel<<<grid, kVecThreads, 0, s>>>(K.x, z, K.n, gate);
      h->launches++;
    }
  else
    xupdate_kernel<<<grid, kVecThreads, 0, s>>>(V, st, K.x, n2, g, ctl, gate);
  h->launches += 3;
  PFC_CUDA(h, cudaMemcpyAsync(h->h_ctl, ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaGetLastError());
  return PFC_OK;
}

static int run_cycle(pfc_handle *h, bool first_cycle, bool two_pass, int i0, int i1)
{
  if (!h->prm.use_cuda_graph)
    return enqueue_cycle(h, first_cycle, two_pass, i0, i1);
  const int key = (((i0 * 64 + i1) * 2 + (first_cycle ? 0 : 1)) * 2 + (two_pass ? 1 : 0)) * 2 + (h->reduced ? 1 : 0);
  pfc_handle::CycleGraph &cg = h->cycle_graphs[key];
  if (!cg.exec)
    {
      const int64_t before = h->launches;
      cudaGraph_t graph = nullptr;
      PFC_CUDA(h, cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
      const int rc = enqueue_cycle(h, first_cycle, two_pass, i0, i1);
      cudaError_t e = cudaStreamEndCapture(h->stream, &graph);
      if (rc != PFC_OK)
        {
          if (graph)
            cudaGraphDestroy(graph);
          return rc;
        }
      PFC_CUDA(h, e);
      PFC_CUDA(h, cudaGraphInstantiate(&cg.exec, graph, 0));
      cudaGraphDestroy(graph);
      cg.launches = h->launches - before; // kernel nodes recorded in the graph
      h->launches = before;               // capture launched nothing
    }
  PFC_CUDA(h, cudaGraphLaunch(cg.exec, h->stream));
  h->launches += cg.launches;
  return PFC_OK;
}

// GMRES as SolverGMRES::solve is used at mpi.cpp:478-482.  x := 0 (mpi.cpp:481).
static int gmres_solve(pfc_handle *h, double tol, int64_t *iters, double *final_res)
{
  const Plan &P = h->plan;
  cudaStream_t s = h->stream;
  const KrylovSys K = krylov_sys(h);
  Ctl c0;
  std::memset(&c0, 0, sizeof(c0));
  c0.tol = tol;
  c0.max_iters = h->prm.gmres_max_iters > 0 ? h->prm.gmres_max_iters : P.n_dofs;
  *h->h_ctl = c0;
  PFC_CUDA(h, cudaMemcpyAsync(h->d_ctl.p, h->h_ctl, sizeof(Ctl), cudaMemcpyHostToDevice, s));
  // owned part only: the ghost tail is refreshed before every use
  PFC_CUDA(h, cudaMemsetAsync(K.x, 0, K.n * sizeof(double), s));
  PFC_CUDA(h, cudaStreamSynchronize(s)); // h_ctl is reused as the read-back buffer
  bool first = true, two_pass = h->prm.orthogonalization == PFC_ORTHO_CGS2;
  // first piece of a cycle: as many iterations as the same solve of the previous time step needed (+ slack, multiple
  // of 4 so that few graphs exist); the kernels of a longer graph would all be launched just to skip themselves
  const int slot = std::min(h->solve_index, 7);
  const int pred = h->pred_iters[slot];
  for (;;)
    {
      int L = h->restart;
      if (first && pred > 0 && h->prm.use_cuda_graph)
        L = std::min(h->restart, (pred + 1 + 3) / 4 * 4);
      PFC_TRY(run_cycle(h, first, two_pass, 0, L));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      if (h->h_ctl->state == 0 && h->h_ctl->dim == L && L < h->restart)
        { // not there yet: the rest of the cycle
          PFC_TRY(run_cycle(h, first, two_pass, L, h->restart));
          PFC_CUDA(h, cudaStreamSynchronize(s));
        }
      first = false;
      if (h->h_ctl->state == 3)
        { // Kelley test tripped: go on with two Gram-Schmidt passes per iteration (sticky)
          two_pass = true;
          h->h_ctl->state = 0;
          PFC_CUDA(h, cudaMemcpyAsync(h->d_ctl.p, h->h_ctl, sizeof(Ctl), cudaMemcpyHostToDevice, s));
          PFC_CUDA(h, cudaStreamSynchronize(s));
          continue;
        }
      if (h->h_ctl->state != 0)
        break;
    }
  h->pred_iters[slot] = (int)std::min<long long>(h->h_ctl->iters, h->restart);
  h->solve_index++;
  if (iters)
    *iters = h->h_ctl->iters;
  if (final_res)
    *final_res = h->h_ctl->rho;
  if (h->h_ctl->state == 2)
    {
      if (std::isnan(h->h_ctl->rho))
        return h->fail(PFC_ERR_NONFINITE, "GMRES residual is NaN after %lld iterations", h->h_ctl->iters);
      return h->fail(PFC_ERR_GMRES_NOT_CONVERGED, "GMRES did not converge in %lld iterations (residual %.3e)",
                     h->h_ctl->iters, h->h_ctl->rho);
    }
  return PFC_OK;
}

// ---- C ABI: hot path ----------------------------------------------------------------------------------
extern "C" int pfc_assemble(pfc_handle *h)
{
  if (!h)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || !h->has_state)
    return h->fail(PFC_ERR_STATE, "pfc_assemble needs pfc_setup and pfc_set_state first");
  PFC_CUDA(h, cudaSetDevice(h->device));
  PFC_TRY(launch_assembly(h, false));
  h->assembled = h->rhs_valid = true;
  return PFC_OK;
}

static int residual_norm_async(pfc_handle *h)
{
  ensure_tab();
  const int64_t n2 = 3 * h->plan.Np / 2;
  GmresSmall *g = h->d_small.p;
  g_dot[0]<<<vec_grid(h, n2), kVecThreads, 0, h->stream>>>(nullptr, 0, h->d_rhs.p, n2, h->d_partial.p, &g->nrm2[0],
                                                            h->d_counter.p, nullptr, 0, ArDev{});
  h->launches++;
  PFC_TRY(allreduce_sum(h, &g->nrm2[0], 1));
  PFC_CUDA(h, cudaMemcpyAsync(h->h_scalar, &g->nrm2[0], sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  return PFC_OK;
}

extern "C" int pfc_residual_norm(pfc_handle *h, double *norm)
{
  if (!h || !norm)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->assembled)
    return h->fail(PFC_ERR_STATE, "pfc_residual_norm before pfc_assemble");
  PFC_CUDA(h, cudaSetDevice(h->device));
  PFC_TRY(residual_norm_async(h));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  *norm = std::sqrt(h->h_scalar[0]);
  return PFC_OK;
}

extern "C" int pfc_solve_update(pfc_handle *h, int64_t *gmres_iters, double *final_residual)
{
  if (!h)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->assembled)
    return h->fail(PFC_ERR_STATE, "pfc_solve_update before pfc_assemble");
  PFC_CUDA(h, cudaSetDevice(h->device));
  const Plan &P = h->plan;
  if (h->reduced)
    { // psi-only reduced system: right-hand side (and the two mass solves, if the auxiliary residuals are not negligible),
      // GMRES on scalar fields, then the Newton update from its solution (pfc_reduced.cu)
      double tol = h->prm.gmres_tol;
      PFC_TRY(reduced_prepare(h, &tol));
      double res = 0;
      PFC_TRY(gmres_solve(h, tol, gmres_iters, &res));
      if (final_residual)
        *final_residual = std::sqrt(res * res + h->red_aux2); // of the 3-block system: rows 1, 2 keep their (tiny) residual
      PFC_TRY(reduced_finish(h));
    }
  else
    {
      if (h->prm.preconditioner == PFC_PRECOND_BLOCK_SCHUR)
        PFC_TRY(precond_refresh(h));
      PFC_TRY(gmres_solve(h, h->prm.gmres_tol, gmres_iters, final_residual));
    }
  // constraints.distribute(dPsi_n) (mpi.cpp:485); Psi_n += dPsi_n (mpi.cpp:487-491)
  PFC_TRY(halo_exchange(h, h->d_x.p));
  const int64_t nsl = (int64_t)P.slave_id.size();
  if (nsl > 0)
    {
      distribute_kernel<<<(int)((nsl + 255) / 256), 256, 0, h->stream>>>(h->d_x.p, h->d_slave.p, h->d_slave_master.p, nsl, P.Np);
      h->launches++;
    }
  axpy1_kernel<<<vec_grid(h, P.Lv), kVecThreads, 0, h->stream>>>(h->d_psi_n.p, h->d_x.p, P.Lv);
  h->launches++;
  PFC_CUDA(h, cudaGetLastError());
  h->assembled = false;
  return PFC_OK;
}

extern "C" int pfc_step(pfc_handle *h, pfc_step_stats *st)
{
  if (!h)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || !h->has_state)
    return h->fail(PFC_ERR_STATE, "pfc_step needs pfc_setup and pfc_set_state first");
  PFC_CUDA(h, cudaSetDevice(h->device));
  pfc_step_stats local;
  if (!st)
    st = &local;
  std::memset(st, 0, sizeof(*st));
  const int64_t launches0 = h->launches;
  h->solve_index = 0;
  cudaStream_t s = h->stream;
  PFC_CUDA(h, cudaEventRecord(h->ev[0], s));
  int status = PFC_WARN_NEWTON_MAX_ITERS;
  float ms;
  for (int it = 0; it < h->prm.newton_max_iters; ++it) // mpi.cpp:579
    {
      PFC_CUDA(h, cudaEventRecord(h->ev[1], s));
      PFC_TRY(launch_assembly(h, false)); // mpi.cpp:582
      h->assembled = h->rhs_valid = true;
      PFC_TRY(residual_norm_async(h)); // mpi.cpp:584
      PFC_CUDA(h, cudaEventRecord(h->ev[2], s));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      PFC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]));
      st->ms_assembly += ms;
      const double residual = std::sqrt(h->h_scalar[0]);
      if (st->n_assemblies < PFC_MAX_NEWTON_RECORD)
        st->residual[st->n_assemblies] = residual;
      st->n_assemblies++;
      if (!std::isfinite(residual))
        {
          st->status = PFC_ERR_NONFINITE;
          return h->fail(PFC_ERR_NONFINITE, "non-finite Newton residual");
        }
      if (residual < h->prm.newton_tol) // mpi.cpp:586-587
        {
          status = PFC_OK;
          break;
        }
      int64_t its = 0;
      double fr = 0;
      PFC_CUDA(h, cudaEventRecord(h->ev[1], s));
      const int rc = pfc_solve_update(h, &its, &fr); // mpi.cpp:590
      PFC_CUDA(h, cudaEventRecord(h->ev[2], s));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      PFC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]));
      st->ms_solver += ms;
      if (st->n_solves < PFC_MAX_NEWTON_RECORD)
        {
          st->gmres_iters[st->n_solves] = its;
          st->gmres_final_residual[st->n_solves] = fr;
        }
      st->n_solves++;
      st->gmres_iters_total += its;
      if (rc != PFC_OK)
        {
          st->status = rc;
          return rc;
        }
    }
  // Psi_n_1 = Psi_n (mpi.cpp:592), also when Newton did not converge
  PFC_CUDA(h, cudaMemcpyAsync(h->d_psi_n1.p, h->d_psi_n.p, h->d_psi_n.n * sizeof(double), cudaMemcpyDeviceToDevice, s));
  PFC_CUDA(h, cudaEventRecord(h->ev[3], s));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  PFC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[0], h->ev[3]));
  st->ms_total = ms;
  st->status = status;
  st->kernel_launches = h->launches - launches0;
  h->assembled = false;
  return status == PFC_OK ? PFC_OK : PFC_WARN_NEWTON_MAX_ITERS;
}

extern "C" int pfc_run(pfc_handle *h, int32_t n_timesteps, pfc_step_stats *last, double *ms_total)
{
  if (!h || n_timesteps < 0)
    return PFC_ERR_BAD_ARGUMENT;
  double total = 0;
  pfc_step_stats st;
  int worst = PFC_OK;
  for (int t = 0; t < n_timesteps; ++t) // mpi.cpp:685-688
    {
      const int rc = pfc_step(h, &st);
      total += st.ms_total;
      if (rc != PFC_OK && rc != PFC_WARN_NEWTON_MAX_ITERS)
        {
          if (last)
            *last = st;
          return rc;
        }
      if (rc != PFC_OK)
        worst = rc;
    }
  if (last)
    *last = st;
  if (ms_total)
    *ms_total = total;
  return worst;
}

// ---- inspection / timing ---------------------------------------------------------------------------
extern "C" int pfc_apply_operator(pfc_handle *h, const double *x, double *y)
{
  if (!h || !x || !y)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->assembled)
    return h->fail(PFC_ERR_STATE, "pfc_apply_operator before pfc_assemble");
  PFC_CUDA(h, cudaSetDevice(h->device));
  const Plan &P = h->plan;
  const int64_t n = P.n_dofs;
  const int blocks = (int)((n + 255) / 256);
  double *vx = h->d_V.p, *vy = h->d_V.p + h->stride;
  PFC_CUDA(h, cudaMemsetAsync(vx, 0, 3 * P.Np * sizeof(double), h->stream));
  PFC_CUDA(h, cudaMemcpyAsync(h->d_io.p, x, n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  scatter_kernel<<<blocks, 256, 0, h->stream>>>(vx, h->d_io.p, h->d_api_map.p, n);
  h->launches++;
  PFC_TRY(apply_operator(h, vx, vy, nullptr));
  PFC_CUDA(h, cudaMemsetAsync(h->d_io.p, 0, n * sizeof(double), h->stream));
  gather_kernel<<<blocks, 256, 0, h->stream>>>(h->d_io.p, vy, h->d_api_map.p, n, 3 * P.Np);
  h->launches++;
  PFC_CUDA(h, cudaGetLastError());
  PFC_CUDA(h, cudaMemcpyAsync(y, h->d_io.p, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  return PFC_OK;
}

extern "C" int pfc_time_operator(pfc_handle *h, int32_t reps, double *ms_mean, double *bytes_per_launch)
{
  if (!h || reps < 1 || !ms_mean)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->assembled)
    return h->fail(PFC_ERR_STATE, "pfc_time_operator before pfc_assemble");
  PFC_CUDA(h, cudaSetDevice(h->device));
  const Plan &P = h->plan;
  // operands rotate through the Krylov basis (what the solver does), so that x and y do not stay in L2 between launches
  const int half = std::max(1, h->n_basis / 2);
  auto vx = [&](int i) { return h->d_V.p + (int64_t)(i % half) * h->stride; };
  auto vy = [&](int i) { return h->d_V.p + (int64_t)(half + i % half) * h->stride; };
  for (int i = 0; i < 3; ++i)
    PFC_TRY(apply_operator(h, vx(i), vy(i), nullptr));
  PFC_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
  for (int i = 0; i < reps; ++i)
    PFC_TRY(apply_operator(h, vx(i), vy(i), nullptr));
  PFC_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  float ms;
  PFC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
  *ms_mean = ms / reps;
  if (bytes_per_launch)
    // layout-true algorithmic bytes: per stored entry col 4 + B,C,M,K 32; per row x read 24 + y write 24;
    // slice pointers 8 per 32 rows
    *bytes_per_launch = h->lat ? lat_bytes(h, LAT_J) : is_general_mf(h)
                          ? 64.0 * (double)P.Np + (double)P.n_cells * (16.0 * P.n_s + 16.0) // x, y, psi, chi + cell tables
                          : 36.0 * (double)P.n_ent + 48.0 * (double)P.Np + 8.0 * (double)P.n_slices;
  return PFC_OK;
}

extern "C" int pfc_time_assembly(pfc_handle *h, int32_t reps, double *ms_mean, double *bytes_per_launch)
{
  if (!h || reps < 1 || !ms_mean)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || !h->has_state)
    return h->fail(PFC_ERR_STATE, "pfc_time_assembly needs pfc_setup and pfc_set_state first");
  PFC_CUDA(h, cudaSetDevice(h->device));
  const Plan &P = h->plan;
  PFC_TRY(launch_assembly(h, false));
  PFC_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
  for (int i = 0; i < reps; ++i)
    PFC_TRY(launch_assembly(h, false));
  PFC_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  float ms;
  PFC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
  *ms_mean = ms / reps;
  h->assembled = h->rhs_valid = true;
  if (bytes_per_launch)
    // B, C zeroed + written (read-modify-write counted once each way): 2 * 2 * 8 per entry;
    // state read 48 / node, rhs write 24 / node; per cell tables (gather, row, spos, pos, h)
    *bytes_per_launch = h->lat ? lat_bytes(h, LAT_R)
                               : 32.0 * (double)P.n_ent + 72.0 * (double)P.Np +
                                   (double)P.n_cells * (4.0 * (3 * P.n_s + P.n_s * P.n_s) + 16.0);
  return PFC_OK;
}

extern "C" int pfc_time_krylov(pfc_handle *h, int32_t dim, int32_t reps, double *ms_dot, double *bytes_dot,
                               double *ms_axpy, double *bytes_axpy)
{
  if (!h || reps < 1 || !ms_dot || !ms_axpy)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup)
    return h->fail(PFC_ERR_STATE, "pfc_time_krylov before pfc_setup");
  if (dim < 1 || dim > h->restart)
    return h->fail(PFC_ERR_BAD_ARGUMENT, "dim must be in [1, restart]");
  PFC_CUDA(h, cudaSetDevice(h->device));
  ensure_tab();
  cudaStream_t s = h->stream;
  const KrylovSys K = krylov_sys(h); // the vectors of the active solve: 3 Np doubles, or Np (reduced solve)
  const int64_t n2 = K.n / 2, stride = K.st;
  GmresSmall *g = h->d_small.p;
  double *V = K.V, *w = V + (int64_t)dim * stride;
  // neutral control block: state 0, no second pass; coefficients 0 so that w is unchanged
  PFC_CUDA(h, cudaMemsetAsync(h->d_ctl.p, 0, sizeof(Ctl), s));
  PFC_CUDA(h, cudaMemsetAsync(g->dots1, 0, sizeof(g->dots1), s));
  float ms;
  for (int k = 0; k < 2; ++k)
    {
      for (int i = 0; i < 2; ++i) // warm-up
        if (k == 0)
          g_dot[dim]<<<gs_grid(h, n2, g_dot_occ[dim]), kVecThreads, 0, s>>>(V, stride, w, n2, h->d_partial.p, g->dots0, h->d_counter.p, h->d_ctl.p, 0, ArDev{});
        else
          g_axpy[dim]<<<gs_grid(h, n2, g_axpy_occ[dim]), kVecThreads, 0, s>>>(V, stride, w, n2, g->dots1, nullptr, h->d_partial.p, &g->nrm2[1], h->d_counter.p, h->d_ctl.p, 0, ArDev{});
      PFC_CUDA(h, cudaEventRecord(h->ev[0], s));
      for (int i = 0; i < reps; ++i)
        if (k == 0)
          g_dot[dim]<<<gs_grid(h, n2, g_dot_occ[dim]), kVecThreads, 0, s>>>(V, stride, w, n2, h->d_partial.p, g->dots0, h->d_counter.p, h->d_ctl.p, 0, ArDev{});
        else
          g_axpy[dim]<<<gs_grid(h, n2, g_axpy_occ[dim]), kVecThreads, 0, s>>>(V, stride, w, n2, g->dots1, nullptr, h->d_partial.p, &g->nrm2[1], h->d_counter.p, h->d_ctl.p, 0, ArDev{});
      h->launches += reps + 2;
      PFC_CUDA(h, cudaEventRecord(h->ev[1], s));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      PFC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
      (k == 0 ? *ms_dot : *ms_axpy) = ms / reps;
    }
  PFC_CUDA(h, cudaGetLastError());
  const double vec = 8.0 * (double)K.n; // one vector
  if (bytes_dot)
    *bytes_dot = (dim + 1) * vec;
  if (bytes_axpy)
    *bytes_axpy = (dim + 2) * vec;
  return PFC_OK;
}
