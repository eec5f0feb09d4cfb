// pfc_api.cu — C ABI (include/pfc_b200.h) over the CUDA kernels: lifecycle, setup, state.
// The solver drivers live in pfc_solver.cu.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>

#include "pfc_comm.h"
#include "pfc_handle.h"

using namespace pfc;

static thread_local std::string g_create_error;

extern "C" const char *pfc_version(void) { return "pfc_b200 0.1 (sm_100a)"; }

extern "C" void pfc_default_params(pfc_params *p)
{
  if (!p)
    return;
  std::memset(p, 0, sizeof(*p));
  // parameter-files/hexagonal_lattice.toml:3-18 + deal.II defaults used at mpi.cpp:478-480
  p->dim = 2;
  p->degree = 1;
  p->dt = 0.1;
  p->theta = 1.0;
  p->eps = -0.8;
  p->newton_tol = 1e-8;
  p->newton_max_iters = 200;
  p->gmres_tol = 1e-10;
  p->gmres_max_iters = 0;
  p->gmres_n_tmp_vectors = 30;
  p->operator_kind = PFC_OPERATOR_ASSEMBLED;
  p->orthogonalization = PFC_ORTHO_CGS;
  p->preconditioner = PFC_PRECOND_NONE;
  p->precond_sweeps = 0;
  p->device = -1;
  p->use_cuda_graph = 1;
}

extern "C" const char *pfc_last_error(const pfc_handle *h) { return h ? h->err.c_str() : g_create_error.c_str(); }

extern "C" int pfc_create(const pfc_params *p, pfc_handle **out)
{
  if (!p || !out)
    {
      g_create_error = "null argument";
      return PFC_ERR_BAD_ARGUMENT;
    }
  *out = nullptr;
  // argument checks mirror the std::invalid_argument paths of app_mpi.cpp:25-110
  if (p->dim != 2)
    {
      g_create_error = "only dim = 2 is supported (dim 3 is instantiated but never run upstream)";
      return PFC_ERR_UNSUPPORTED;
    }
  if (p->degree < 1 || p->degree > 2)
    {
      g_create_error = "degree must be 1 or 2";
      return PFC_ERR_UNSUPPORTED;
    }
  if (!(p->dt > 0) || !(p->theta >= 0 && p->theta <= 1) || !(p->newton_tol > 0) || p->newton_max_iters < 1 ||
      !(p->gmres_tol > 0) || !std::isfinite(p->eps))
    {
      g_create_error = "invalid time-stepping / tolerance parameter";
      return PFC_ERR_BAD_ARGUMENT;
    }
  const int ntmp = p->gmres_n_tmp_vectors > 0 ? p->gmres_n_tmp_vectors : 30;
  if (ntmp < 3 || ntmp > 34)
    {
      g_create_error = "gmres_n_tmp_vectors must be in [3, 34]";
      return PFC_ERR_UNSUPPORTED;
    }
  if (p->orthogonalization != PFC_ORTHO_CGS && p->orthogonalization != PFC_ORTHO_CGS2)
    {
      g_create_error = "orthogonalization must be PFC_ORTHO_CGS or PFC_ORTHO_CGS2";
      return PFC_ERR_UNSUPPORTED;
    }
  if ((p->operator_kind != PFC_OPERATOR_ASSEMBLED && p->operator_kind != PFC_OPERATOR_MATRIX_FREE &&
       p->operator_kind != PFC_OPERATOR_MATRIX_FREE_GENERAL) ||
      (p->preconditioner != PFC_PRECOND_NONE && p->preconditioner != PFC_PRECOND_BLOCK_SCHUR &&
       p->preconditioner != PFC_PRECOND_SCHUR_REDUCED) ||
      p->precond_sweeps > 16)
    {
      g_create_error = "operator_kind / preconditioner not available in this build";
      return PFC_ERR_UNSUPPORTED;
    }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    {
      g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (there is no CPU fallback)";
      return PFC_ERR_CUDA;
    }
  pfc_handle *h = new pfc_handle();
  h->prm = *p;
  h->prm.gmres_n_tmp_vectors = ntmp;
  h->restart = ntmp - 2;
  h->n_basis = ntmp - 1;
  // Chebyshev degree: measured optimum of the lattice path on M2-sized meshes (profiles/README.md): Q2 5, Q1 3
  // default Chebyshev degree: measured best per path.  Matrix-free lattice path with its cheap streaming sweeps: Q1 4 (M4: 226
  // against 237 ms per step at 3 or 5, profiles/r02r), Q2 6 (M2: 24.7 against 25.5 ms at 5, profiles/r02x); assembled: 3 / 5
  h->sweeps = p->precond_sweeps > 0 ? p->precond_sweeps
                                    : (p->operator_kind == PFC_OPERATOR_MATRIX_FREE ? (p->degree == 2 ? 6 : 4) : (p->degree == 2 ? 5 : 3));
  // The reduced solve shares buffers, ring and fall-back with the block-Schur preconditioner: from here on it is that
  // preconditioner plus a flag (pfc_setup decides whether the mesh allows it).  Its S~ is well conditioned after the
  // Jacobi scaling (spectrum about [0.6, 2.5]): two Chebyshev steps do what four do (tests/study_reduced.py)
  h->want_reduced = p->preconditioner == PFC_PRECOND_SCHUR_REDUCED;
  if (h->want_reduced)
    {
      h->prm.preconditioner = PFC_PRECOND_BLOCK_SCHUR;
      if (p->precond_sweeps <= 0 && p->operator_kind == PFC_OPERATOR_MATRIX_FREE && p->degree == 1)
        h->sweeps = 2;
    }
  int dev = p->device;
  if (dev < 0)
    cudaGetDevice(&dev);
  if (dev >= ndev || (e = cudaSetDevice(dev)) != cudaSuccess)
    {
      g_create_error = "invalid device ordinal";
      delete h;
      return PFC_ERR_CUDA;
    }
  h->device = dev;
  cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, dev);
  if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess)
    {
      g_create_error = std::string("cudaStreamCreate: ") + cudaGetErrorString(e);
      delete h;
      return PFC_ERR_CUDA;
    }
  for (auto &ev : h->ev)
    cudaEventCreate(&ev);
  cudaMallocHost((void **)&h->h_ctl, sizeof(Ctl));
  cudaMallocHost((void **)&h->h_scalar, 8 * sizeof(double));
  *out = h;
  return PFC_OK;
}

extern "C" void pfc_destroy(pfc_handle *h)
{
  if (!h)
    return;
  cudaSetDevice(h->device);
  if (h->stream)
    cudaStreamSynchronize(h->stream);
  for (auto &kv : h->cycle_graphs)
    if (kv.second.exec)
      cudaGraphExecDestroy(kv.second.exec);
  h->cycle_graphs.clear();
  comm_destroy(h);
  for (auto &ev : h->ev)
    if (ev)
      cudaEventDestroy(ev);
  if (h->h_ctl)
    cudaFreeHost(h->h_ctl);
  if (h->h_scalar)
    cudaFreeHost(h->h_scalar);
  if (h->stream)
    cudaStreamDestroy(h->stream);
  delete h;
}

int launch_assembly(pfc_handle *h, bool const_pass); // pfc_solver.cu
int reduced_setup(pfc_handle *h);                     // pfc_reduced.cu
int precond_after_setup(pfc_handle *h);               // pfc_solver.cu

extern "C" int pfc_setup(pfc_handle *h, const pfc_mesh *mesh)
{
  if (!h || !mesh)
    return PFC_ERR_BAD_ARGUMENT;
  PFC_CUDA(h, cudaSetDevice(h->device));
  h->is_setup = h->has_state = h->assembled = h->rhs_valid = h->stress_valid = false;
  h->stress_map_owned = -1; // maps derived from the previous plan
  h->d_owned_map.release();
  h->d_stress.release();
  comm_p2p_destroy(h); // peers hold mappings of the arena that is about to be replaced
  if (mesh->n_ranks > 1 && (!h->comm || h->comm_size != mesh->n_ranks || h->comm_rank != mesh->rank))
    return h->fail(PFC_ERR_STATE, "pfc_comm_init must precede pfc_setup for n_ranks > 1 (same rank / n_ranks)");
  Plan &P = h->plan;
  const int plan_flags = h->prm.operator_kind == PFC_OPERATOR_MATRIX_FREE ? (PLAN_LATTICE | PLAN_SKIP_TABLES) : 0;
  int rc = build_plan(*mesh, h->prm.degree, P, plan_flags);
  if (rc != PFC_OK)
    return h->fail(rc, "pfc_setup: %s", P.error.c_str());
  h->lat = false;
  h->operator_note.clear();
  if (plan_flags & PLAN_LATTICE)
    { // the lattice kernels need every rank on the lattice: the ranks exchange their places on it (collective) and
      // replace deal.II's one-cell ghost layer by the deep ring; otherwise all ranks take the general cell kernel
      std::vector<RingCard> cards;
      rc = comm_gather_cards(h, ring_card(P), cards);
      if (rc != PFC_OK)
        return rc;
      std::string why;
      for (size_t r = 0; r < cards.size(); ++r)
        if (!cards[r].active && why.empty())
          why = "rank " + std::to_string(r) + (r == (size_t)P.rank ? ": " + P.lat.why : std::string(": no uniform lattice with a rectangular owned region"));
      if (why.empty())
        {
          // depth: what one preconditioned GMRES iteration consumes (t, Schur right-hand side, Chebyshev sweeps,
          // back substitution, operator), DESIGN.md "ring"; unpreconditioned: the operator's one cell
          const int p = h->prm.degree;
          const int depth = h->prm.preconditioner == PFC_PRECOND_BLOCK_SCHUR ? (3 * h->sweeps + 2) * p : p;
          rc = ring_overlay(P, cards, depth);
          if (rc != PFC_OK)
            why = P.error;
        }
      if (why.empty())
        h->lat = true;
      else
        {
          h->operator_note = "lattice kernels declined (" + why + "): using the general colour-ordered matrix-free kernel";
          if (P.lat.active)
            { // this rank's plan skipped the tables the general kernel needs
              rc = build_plan(*mesh, h->prm.degree, P, 0);
              if (rc != PFC_OK)
                return h->fail(rc, "pfc_setup: %s", P.error.c_str());
            }
        }
    }

  // host pattern check (mpi.cpp:232-244): bit-identical or refused
  if (mesh->row_ptr && mesh->col_idx && P.n_ranks == 1 && P.has_tables)
    {
      std::vector<int64_t> rp;
      std::vector<int32_t> ci;
      monolithic_pattern(P, rp, ci, nullptr, nullptr);
      if (std::memcmp(rp.data(), mesh->row_ptr, rp.size() * sizeof(int64_t)) != 0 ||
          std::memcmp(ci.data(), mesh->col_idx, ci.size() * sizeof(int32_t)) != 0)
        return h->fail(PFC_ERR_BAD_ARGUMENT, "pfc_setup: host sparsity pattern differs from the derived pattern");
    }
  if (h->lat)
    {
      rc = comm_ring_upload(h);
      if (rc != PFC_OK)
        return rc;
    }
  else if (P.n_ranks > 1)
    {
      rc = comm_exchange_send_lists(h);
      if (rc != PFC_OK)
        return rc;
    }

  cudaStream_t s = h->stream;
  PFC_CUDA(h, h->d_cell_gather.upload(P.cell_gather, s));
  PFC_CUDA(h, h->d_cell_row.upload(P.cell_row, s));
  PFC_CUDA(h, h->d_cell_pos.upload(P.cell_pos, s));
  PFC_CUDA(h, h->d_cell_spos.upload(P.cell_spos, s));
  PFC_CUDA(h, h->d_cell_col.upload(P.cell_col, s));
  PFC_CUDA(h, h->d_cell_h.upload(P.cell_h, s));
  PFC_CUDA(h, h->d_col.upload(P.col, s));
  PFC_CUDA(h, h->d_slice_ptr.upload(P.slice_ptr, s));
  PFC_CUDA(h, h->d_slave.upload(P.slave_id, s));
  PFC_CUDA(h, h->d_slave_master.upload(P.slave_master, s));
  {
    std::vector<double> tab;
    tab.insert(tab.end(), P.shape.begin(), P.shape.end());
    tab.insert(tab.end(), P.dshape.begin(), P.dshape.end());
    tab.insert(tab.end(), P.qw.begin(), P.qw.end());
    PFC_CUDA(h, h->d_tab.upload(tab, s));
    PFC_CUDA(h, cudaStreamSynchronize(s)); // tab is a local
  }
  {
    // API dof k -> internal address (owned and ghost), -1 if not local
    std::vector<int64_t> map(P.n_dofs, -1);
    for (int64_t r = 0; r < P.Np; ++r)
      if (P.perm[r] >= 0)
        for (int c = 0; c < 3; ++c)
          map[P.block_start[c] + P.perm[r]] = c * P.Np + r;
    for (int64_t g = 0; g < P.n_ghost; ++g)
      if (P.ghost_global[g] >= 0) // ring overlay: only the positions of the handed-over frame have a global id
        for (int c = 0; c < 3; ++c)
          map[P.block_start[c] + P.ghost_global[g]] = 3 * (P.Np + g) + c;
    PFC_CUDA(h, h->d_api_map.upload(map, s));
    PFC_CUDA(h, cudaStreamSynchronize(s));
  }
  h->stride = (P.Lv + 15) / 16 * 16;
  const size_t ne = (size_t)P.n_ent;
  PFC_CUDA(h, h->d_B.alloc(ne));
  PFC_CUDA(h, h->d_C.alloc(ne));
  PFC_CUDA(h, h->d_M.alloc(ne));
  PFC_CUDA(h, h->d_K.alloc(ne));
  PFC_CUDA(h, h->d_rhs.alloc(h->stride));
  PFC_CUDA(h, h->d_psi_n.alloc(h->stride));
  PFC_CUDA(h, h->d_psi_n1.alloc(h->stride));
  h->sstride = (P.Np + P.n_ghost + 15) / 16 * 16;
  {
    const size_t n3 = (size_t)h->n_basis + 3;
    h->arena_off_scalar = (int64_t)(n3 * h->stride);
    const bool pc = h->prm.preconditioner == PFC_PRECOND_BLOCK_SCHUR;
    PFC_CUDA(h, h->d_arena.alloc(n3 * h->stride + (pc ? 12 * (size_t)h->sstride : 0)));
    PFC_CUDA(h, cudaMemsetAsync(h->d_arena.p, 0, h->d_arena.n * sizeof(double), s));
    double *a = h->d_arena.p;
    h->d_V.view(a, (size_t)h->stride * h->n_basis);
    h->d_x.view(a + (size_t)h->stride * h->n_basis, h->stride);
    h->d_z.view(a + (size_t)h->stride * (h->n_basis + 1), h->stride);
    h->d_u.view(a + (size_t)h->stride * (h->n_basis + 2), h->stride);
    if (pc)
      {
        h->d_pw.view(a + h->arena_off_scalar, (size_t)h->sstride * 9);
        h->d_pv.view(a + h->arena_off_scalar + 9 * h->sstride, (size_t)h->sstride * 3);
      }
  }
  PFC_CUDA(h, h->d_io.alloc(P.n_dofs));
  PFC_CUDA(h, h->d_ctl.alloc(1));
  PFC_CUDA(h, h->d_ctl_end.alloc(1));
  PFC_CUDA(h, cudaMemsetAsync(h->d_ctl_end.p, 0, sizeof(Ctl), s));
  PFC_CUDA(h, h->d_small.alloc(1));
  PFC_CUDA(h, h->d_partial.alloc((size_t)kPartialStride * 4096));
  PFC_CUDA(h, h->d_counter.alloc(4));
  h->eig_valid = false;
  if (h->prm.preconditioner == PFC_PRECOND_BLOCK_SCHUR)
    {
      PFC_CUDA(h, h->d_inv_ml.alloc(P.Np));
      PFC_CUDA(h, h->d_dinvB.alloc(P.Np));
      PFC_CUDA(h, h->d_cheb.alloc(64));
      PFC_CUDA(h, cudaMemsetAsync(h->d_cheb.p, 0, 64 * sizeof(double), s));
    }
  PFC_CUDA(h, cudaMemsetAsync(h->d_counter.p, 0, 4 * sizeof(unsigned int), s));
  PFC_CUDA(h, cudaMemsetAsync(h->d_ctl.p, 0, sizeof(Ctl), s));
  PFC_CUDA(h, cudaMemsetAsync(h->d_small.p, 0, sizeof(GmresSmall), s));
  for (DevBuf<double> *b : {&h->d_B, &h->d_C, &h->d_M, &h->d_K})
    PFC_CUDA(h, cudaMemsetAsync(b->p, 0, b->n * sizeof(double), s));
  for (DevBuf<double> *b : {&h->d_rhs, &h->d_psi_n, &h->d_psi_n1})
    PFC_CUDA(h, cudaMemsetAsync(b->p, 0, b->n * sizeof(double), s));
  if (P.n_ranks > 1)
    {
      rc = comm_p2p_setup(h); // falls back to NCCL (h->p2p = false) when peer memory cannot be mapped
      if (rc != PFC_OK)
        return rc;
    }
  for (auto &kv : h->cycle_graphs)
    if (kv.second.exec)
      cudaGraphExecDestroy(kv.second.exec);
  h->cycle_graphs.clear();
  for (int &v : h->pred_iters)
    v = 0;
  h->is_setup = true;
  // iterate-independent blocks M and K, once (they are re-assembled every Newton iteration
  // upstream, mpi.cpp:435-443, with identical values); the lattice path stores no matrix at all
  rc = h->lat ? lat_setup(h) : launch_assembly(h, true);
  if (rc != PFC_OK)
    {
      h->is_setup = false;
      return rc;
    }
  if (h->prm.preconditioner == PFC_PRECOND_BLOCK_SCHUR)
    {
      rc = precond_after_setup(h);
      if (rc != PFC_OK)
        {
          h->is_setup = false;
          return rc;
        }
    }
  h->reduced = false;
  if (h->want_reduced)
    {
      if (h->lat && P.degree == 1 && h->q1_fast)
        {
          rc = reduced_setup(h);
          if (rc != PFC_OK)
            {
              h->is_setup = false;
              return rc;
            }
        }
      else
        h->operator_note += std::string(h->operator_note.empty() ? "" : "; ") +
                            "PFC_PRECOND_SCHUR_REDUCED needs the Q1 lattice path: using PFC_PRECOND_BLOCK_SCHUR";
    }
  PFC_CUDA(h, cudaStreamSynchronize(s));
  return PFC_OK;
}

extern "C" int pfc_set_state(pfc_handle *h, const double *Psi_n, const double *Psi_n_1)
{
  if (!h || !Psi_n || !Psi_n_1)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup)
    return h->fail(PFC_ERR_STATE, "pfc_set_state before pfc_setup");
  PFC_CUDA(h, cudaSetDevice(h->device));
  const int64_t n = h->plan.n_dofs;
  const int blocks = (int)((n + 255) / 256);
  const double *src[2] = {Psi_n, Psi_n_1};
  double *dst[2] = {h->d_psi_n.p, h->d_psi_n1.p};
  for (int k = 0; k < 2; ++k)
    {
      if (k == 0 || src[1] != src[0]) // same host array for both time levels: one upload, two scatters
        PFC_CUDA(h, cudaMemcpyAsync(h->d_io.p, src[k], n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
      scatter_kernel<<<blocks, 256, 0, h->stream>>>(dst[k], h->d_io.p, h->d_api_map.p, n);
      h->launches++;
      if (h->lat)
        { // the ring beyond the handed-over ghost layer comes from its owners [collective for n_ranks > 1]
          const int rc = halo_exchange(h, dst[k]);
          if (rc != PFC_OK)
            return rc;
        }
    }
  PFC_CUDA(h, cudaGetLastError());
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  h->has_state = true;
  h->assembled = false;
  return PFC_OK;
}

// ---- owned-range transfers (multi-rank hosts hold only owned_partitioning, mpi.cpp:167-176) ----
static int owned_map(pfc_handle *h)
{
  const Plan &P = h->plan;
  if (h->d_owned_map.n == (size_t)(3 * P.n_owned))
    return PFC_OK;
  std::vector<int64_t> map((size_t)(3 * P.n_owned));
  for (int c = 0; c < 3; ++c)
    for (int64_t j = 0; j < P.n_owned; ++j)
      map[c * P.n_owned + j] = c * P.Np + P.iperm[j];
  PFC_CUDA(h, h->d_owned_map.upload(map, h->stream));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  return PFC_OK;
}

extern "C" int pfc_operator_info(const pfc_handle *h, char *buf, int32_t buflen)
{
  if (!h || !h->is_setup)
    return -1;
  const int kind = h->lat ? PFC_OPERATOR_MATRIX_FREE
                          : (h->prm.operator_kind == PFC_OPERATOR_ASSEMBLED ? PFC_OPERATOR_ASSEMBLED : PFC_OPERATOR_MATRIX_FREE_GENERAL);
  if (buf && buflen > 0)
    std::snprintf(buf, (size_t)buflen, "%s", h->operator_note.c_str());
  return kind;
}

extern "C" int64_t pfc_n_owned_nodes(const pfc_handle *h) { return (h && h->is_setup) ? h->plan.n_owned : -1; }

extern "C" int pfc_set_state_owned(pfc_handle *h, const double *Psi_n, const double *Psi_n_1)
{
  if (!h || !Psi_n || !Psi_n_1)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup)
    return h->fail(PFC_ERR_STATE, "pfc_set_state_owned before pfc_setup");
  PFC_CUDA(h, cudaSetDevice(h->device));
  int rc = owned_map(h);
  if (rc != PFC_OK)
    return rc;
  const int64_t n = 3 * h->plan.n_owned;
  const int blocks = (int)((n + 255) / 256);
  const double *src[2] = {Psi_n, Psi_n_1};
  double *dst[2] = {h->d_psi_n.p, h->d_psi_n1.p};
  for (int k = 0; k < 2; ++k)
    {
      if (k == 0 || src[1] != src[0])
        PFC_CUDA(h, cudaMemcpyAsync(h->d_io.p, src[k], n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
      scatter_kernel<<<blocks, 256, 0, h->stream>>>(dst[k], h->d_io.p, h->d_owned_map.p, n);
      h->launches++;
      rc = halo_exchange(h, dst[k]); // ghost values from their owners (update_ghost_values)
      if (rc != PFC_OK)
        return rc;
    }
  PFC_CUDA(h, cudaGetLastError());
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  h->has_state = true;
  h->assembled = false;
  return PFC_OK;
}

extern "C" int pfc_get_state_owned(pfc_handle *h, double *Psi_n, double *Psi_n_1)
{
  if (!h)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || !h->has_state)
    return h->fail(PFC_ERR_STATE, "pfc_get_state_owned before pfc_set_state");
  PFC_CUDA(h, cudaSetDevice(h->device));
  int rc = owned_map(h);
  if (rc != PFC_OK)
    return rc;
  const int64_t n = 3 * h->plan.n_owned;
  const int blocks = (int)((n + 255) / 256);
  double *host[2] = {Psi_n, Psi_n_1};
  const double *src[2] = {h->d_psi_n.p, h->d_psi_n1.p};
  for (int k = 0; k < 2; ++k)
    if (host[k])
      {
        gather_kernel<<<blocks, 256, 0, h->stream>>>(h->d_io.p, src[k], h->d_owned_map.p, n, 3 * h->plan.Np);
        h->launches++;
        PFC_CUDA(h, cudaMemcpyAsync(host[k], h->d_io.p, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        PFC_CUDA(h, cudaStreamSynchronize(h->stream));
      }
  PFC_CUDA(h, cudaGetLastError());
  return PFC_OK;
}

static int download_vector(pfc_handle *h, const double *dvec, double *host)
{
  const int64_t n = h->plan.n_dofs;
  const int blocks = (int)((n + 255) / 256);
  PFC_CUDA(h, cudaMemsetAsync(h->d_io.p, 0, n * sizeof(double), h->stream));
  gather_kernel<<<blocks, 256, 0, h->stream>>>(h->d_io.p, dvec, h->d_api_map.p, n,
                                                   3 * h->plan.Np);
  h->launches++;
  PFC_CUDA(h, cudaGetLastError());
  PFC_CUDA(h, cudaMemcpyAsync(host, h->d_io.p, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  return PFC_OK;
}

extern "C" int pfc_get_state(pfc_handle *h, double *Psi_n, double *Psi_n_1)
{
  if (!h)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || !h->has_state)
    return h->fail(PFC_ERR_STATE, "pfc_get_state before pfc_set_state");
  PFC_CUDA(h, cudaSetDevice(h->device));
  int rc = PFC_OK;
  if (Psi_n)
    rc = download_vector(h, h->d_psi_n.p, Psi_n);
  if (rc == PFC_OK && Psi_n_1)
    rc = download_vector(h, h->d_psi_n1.p, Psi_n_1);
  return rc;
}

extern "C" int pfc_get_rhs(pfc_handle *h, double *rhs)
{
  if (!h || !rhs)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->assembled)
    return h->fail(PFC_ERR_STATE, "pfc_get_rhs before pfc_assemble");
  PFC_CUDA(h, cudaSetDevice(h->device));
  return download_vector(h, h->d_rhs.p, rhs);
}

extern "C" int64_t pfc_get_nnz(pfc_handle *h) { return (h && h->is_setup && h->plan.has_tables) ? h->plan.nnz_mono : -1; }

extern "C" int pfc_get_pattern(pfc_handle *h, int64_t *row_ptr, int32_t *col_idx)
{
  if (!h || !row_ptr || !col_idx)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup)
    return h->fail(PFC_ERR_STATE, "pfc_get_pattern before pfc_setup");
  if (h->plan.n_ranks != 1)
    return h->fail(PFC_ERR_UNSUPPORTED, "monolithic inspection is single-rank only");
  if (!h->plan.has_tables)
    return h->fail(PFC_ERR_UNSUPPORTED, "the lattice matrix-free path stores no sparsity pattern (use PFC_OPERATOR_ASSEMBLED)");
  std::vector<int64_t> rp;
  std::vector<int32_t> ci;
  monolithic_pattern(h->plan, rp, ci, nullptr, nullptr);
  std::memcpy(row_ptr, rp.data(), rp.size() * sizeof(int64_t));
  std::memcpy(col_idx, ci.data(), ci.size() * sizeof(int32_t));
  return PFC_OK;
}

extern "C" int pfc_get_matrix_values(pfc_handle *h, double *values)
{
  if (!h || !values)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->assembled)
    return h->fail(PFC_ERR_STATE, "pfc_get_matrix_values before pfc_assemble");
  if (h->plan.n_ranks != 1)
    return h->fail(PFC_ERR_UNSUPPORTED, "monolithic inspection is single-rank only");
  if (!h->plan.has_tables)
    return h->fail(PFC_ERR_UNSUPPORTED, "the lattice matrix-free path stores no matrix (use PFC_OPERATOR_ASSEMBLED)");
  PFC_CUDA(h, cudaSetDevice(h->device));
  const Plan &P = h->plan;
  std::vector<double> B(P.n_ent), C(P.n_ent), M(P.n_ent), K(P.n_ent);
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  PFC_CUDA(h, cudaMemcpy(B.data(), h->d_B.p, P.n_ent * sizeof(double), cudaMemcpyDeviceToHost));
  PFC_CUDA(h, cudaMemcpy(C.data(), h->d_C.p, P.n_ent * sizeof(double), cudaMemcpyDeviceToHost));
  PFC_CUDA(h, cudaMemcpy(M.data(), h->d_M.p, P.n_ent * sizeof(double), cudaMemcpyDeviceToHost));
  PFC_CUDA(h, cudaMemcpy(K.data(), h->d_K.p, P.n_ent * sizeof(double), cudaMemcpyDeviceToHost));
  std::vector<int64_t> rp;
  std::vector<int32_t> ci, pos;
  std::vector<int8_t> blk;
  monolithic_pattern(P, rp, ci, &blk, &pos);
  const double dtth = h->prm.dt * h->prm.theta;
  for (size_t k = 0; k < ci.size(); ++k)
    {
      const int32_t p = pos[k];
      switch (blk[k])
        {
        case 0: values[k] = B[p]; break;
        case 1: values[k] = C[p]; break;
        case 2: values[k] = dtth * (K[p] - 2.0 * M[p]); break; // D, as the SpMV kernel forms it
        case 3: values[k] = K[p]; break;
        default: values[k] = M[p]; break;
        }
    }
  return PFC_OK;
}

extern "C" int pfc_get_dof_permutation(pfc_handle *h, int32_t *internal_of_node)
{
  if (!h || !internal_of_node)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup)
    return h->fail(PFC_ERR_STATE, "pfc_get_dof_permutation before pfc_setup");
  const Plan &P = h->plan;
  for (int64_t t = 0; t < P.N_global; ++t)
    internal_of_node[t] = -1;
  for (int64_t i = 0; i < P.n_owned; ++i)
    internal_of_node[P.own0 + i] = P.iperm[i];
  for (int64_t g = 0; g < P.n_ghost; ++g)
    if (P.ghost_global[g] >= 0)
      internal_of_node[P.ghost_global[g]] = (int32_t)(P.Np + g);
  return PFC_OK;
}
