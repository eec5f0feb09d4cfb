// pfc_handle.h — the opaque handle behind the C ABI: device buffers, stream, NCCL communicator.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

#include "pfc_lattice_args.h"
#include "pfc_precond.cuh"
#include "pfc_plan.h"

struct ncclComm;
namespace pfc
{
struct P2PState;
}

namespace pfc
{
template <class T>
struct DevBuf
{
  T *p = nullptr;
  size_t n = 0;
  cudaError_t alloc(size_t count)
  {
    release();
    n = count;
    if (count == 0)
      return cudaSuccess;
    return cudaMalloc((void **)&p, count * sizeof(T));
  }
  cudaError_t upload(const std::vector<T> &h, cudaStream_t s)
  {
    cudaError_t e = alloc(h.size());
    if (e != cudaSuccess || h.empty())
      return e;
    return cudaMemcpyAsync(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, s);
  }
  bool owns = true;
  // non-owning window into another allocation (the peer-visible arena)
  void view(T *ptr, size_t count)
  {
    release();
    p = ptr;
    n = count;
    owns = false;
  }
  void release()
  {
    if (p && owns)
      cudaFree(p);
    p = nullptr;
    n = 0;
    owns = true;
  }
  ~DevBuf() { release(); }
  DevBuf() = default;
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
};
} // namespace pfc

struct pfc_handle
{
  pfc_params prm;
  std::string err;
  int device = 0;
  int n_sm = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  bool is_setup = false, has_state = false, assembled = false;
  pfc::Plan plan;
  int64_t stride = 0; // doubles between Krylov basis vectors (Lv rounded up to 16)
  int n_basis = 0;    // n_tmp_vectors - 1 basis vectors (v_0 .. v_restart)
  int restart = 0;    // n_tmp_vectors - 2

  // mesh tables
  pfc::DevBuf<int32_t> d_cell_gather, d_cell_row, d_cell_pos, d_cell_spos, d_cell_col, d_col, d_slave, d_slave_master;
  pfc::DevBuf<double> d_cell_h, d_tab;
  pfc::DevBuf<int64_t> d_slice_ptr, d_api_map, d_stress_map, d_owned_map;
  pfc::DevBuf<double> d_stress, d_stress_io; // pfc_stress work fields (allocated on first use)
  std::string operator_note;   // why the lattice kernels were declined (pfc_operator_info)
  bool stress_valid = false;
  bool rhs_valid = false; // d_rhs holds the right-hand side of the last assembly (system_rhs as output_rhs sees it)
  int stress_map_owned = -1; // which output map d_stress_map holds (-1 none, 0 [4][N], 1 [4][n_owned])
  // matrix blocks (SELL entry order) and vectors
  pfc::DevBuf<double> d_B, d_C, d_M, d_K, d_rhs, d_psi_n, d_psi_n1, d_x, d_V, d_io;
  // Krylov control
  pfc::DevBuf<pfc::Ctl> d_ctl;
  pfc::DevBuf<pfc::GmresSmall> d_small;
  pfc::DevBuf<double> d_partial;
  pfc::DevBuf<unsigned int> d_counter;
  pfc::Ctl *h_ctl = nullptr; // pinned
  double *h_scalar = nullptr; // pinned [4]
  // restart-cycle graphs, keyed by (first cycle (x = 0) | later, one | two GS passes, iterations [i0, i1)): a cycle is
  // launched as [0, L) with L predicted from the previous time step and, if it has not converged by then, [L, restart)
  struct CycleGraph
  {
    cudaGraphExec_t exec = nullptr;
    int64_t launches = 0;
  };
  std::map<int, CycleGraph> cycle_graphs;
  pfc::DevBuf<pfc::Ctl> d_ctl_end;  // gate of the cycle's end part: state 0 = run (converged, failed or restart length reached)
  int pred_iters[8] = {0, 0, 0, 0, 0, 0, 0, 0}; // iterations of the k-th solve of the previous time step
  int solve_index = 0;

  // multi-GPU
  ncclComm *comm = nullptr;
  int comm_rank = 0, comm_size = 1;
  pfc::DevBuf<int32_t> d_send_rows, d_self_src, d_self_dst, d_recv_pos; // ring overlay lists (pfc_ring_plan.cpp)
  pfc::DevBuf<double> d_send_buf, d_recv_buf;
  // every vector that takes part in a halo exchange lives in one allocation, so that peers can address it:
  //   [n3 three-component vectors of `stride` doubles: Krylov basis, x, z, u | 12 scalar fields of `sstride`]
  pfc::DevBuf<double> d_arena;
  int64_t arena_off_scalar = 0; // doubles before the scalar fields
  // peer-to-peer path over NVLink (pfc_comm.cu): ghost values are stored straight into the neighbours' vectors and
  // small all-reduces go through per-rank slots, both synchronised by epoch flags in peer-visible memory
  bool p2p = false;
  pfc::P2PState *p2p_state = nullptr;

  // block-Schur preconditioner (pfc_precond.cuh)
  int64_t sstride = 0;                    // doubles between scalar work vectors (Np + n_ghost, rounded to 16)
  pfc::DevBuf<double> d_inv_ml, d_dinvB, d_pw, d_z, d_u, d_cheb, d_pv;
  int sweeps = 4;
  double eig_min = 0, eig_max = 0;
  bool eig_valid = false;
  int solves_since_eig = 0;

  // psi-only reduced solve (PFC_PRECOND_SCHUR_REDUCED, pfc_reduced.cu): active on the Q1 lattice path
  bool want_reduced = false, reduced = false;
  pfc::DevBuf<double> d_dinvS, d_cheb_red; // 1 / diag(S~) as a field [owned | ring]; Chebyshev coefficients
  double *red_y = nullptr, *red_u = nullptr, *red_f = nullptr, *red_z = nullptr; // scalar work fields inside d_pw
  double *red_xa = nullptr, *red_xb = nullptr, *red_d = nullptr, *red_um = nullptr, *red_wm = nullptr;
  double red_dinvc = 0, red_alpha = 0; // 1 / diag(S~); coefficient of Chebyshev step 0 (x_0 = alpha r)
  bool red_have_uw = false; // this solve carries u = M^-1 G, w = M^-1 (K u - H)
  double red_aux2 = 0;      // squared residual the reduced solve leaves in rows 1, 2 of the 3-block system
  int64_t red_mass_iters = 0;

  // lattice matrix-free path (pfc_lattice.cu): active when operator_kind = PFC_OPERATOR_MATRIX_FREE and the plan found
  // the uniform periodic tensor lattice
  bool lat = false;
  bool q1_fast = true;          // Q1: closed-form kernels of pfc_ring_q1.cuh; off: PFC_Q1_CLOSED_FORM=0 / self-check
  int j1_variant = 35;          // Q1 J apply kernel and tiling (lat_launch, pfc_lattice.cu); PFC_J1_VARIANT at pfc_setup
  bool lat_k_by_cells = false; // K product by the cell-quadrature kernel instead of the node stencil (self-check)
  pfc::LatGeom lat_geom;

  int64_t launches = 0;

  int fail(int code, const char *fmt, ...)
  {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    err = buf;
    return code;
  }
};

#define PFC_CUDA(h, call)                                                                                      \
  do                                                                                                           \
    {                                                                                                          \
      cudaError_t e_ = (call);                                                                                 \
      if (e_ != cudaSuccess)                                                                                   \
        return (h)->fail(PFC_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));     \
    }                                                                                                          \
  while (0)
