// pfc_lattice.cu — host launcher of the lattice kernels (pfc_lattice.cuh).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "pfc_comm.h"
#include "pfc_handle.h"
#include "pfc_lattice.cuh"
#include "pfc_ring.cuh"
#include "pfc_ring_q1.cuh"
#include "pfc_reduced.cuh"

namespace pfc
{
namespace
{
// cells per CTA (one thread per cell); the inner (CX-1) x (CY-1) cells are the CTA's own
constexpr int kCX = 16, kCY = 16;

template <int P, int MODE, int CX, int CY>
int launch_tile(pfc_handle *h, const LatOp &op, const Ctl *ctl)
{
  static bool opted_in = false; // per instantiation; one device per process (one rank per GPU)
  if (!opted_in)
    {
      PFC_CUDA(h, cudaFuncSetAttribute(lat_kernel<P, MODE, CX, CY>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)lat_smem_bytes<P, MODE, CX, CY>()));
      opted_in = true;
    }
  const LatGeom &G = h->lat_geom;
  const dim3 grid((G.ncx + CX - 2) / (CX - 1), (G.ncy + CY - 2) / (CY - 1));
  lat_kernel<P, MODE, CX, CY><<<grid, CX * CY, lat_smem_bytes<P, MODE, CX, CY>(), h->stream>>>(G, op, ctl);
  h->launches++;
  return PFC_OK;
}

template <int P, int MODE>
int launch_mode(pfc_handle *h, const LatOp &op, const Ctl *ctl)
{
  if constexpr (MODE == LAT_J || MODE == LAT_S)
    { // the two hot cell kernels: 16 x 8 cells per CTA for Q2 (two CTAs per SM instead of one: J 49 -> 35 us, Schur row
      // 30 -> 27 us on M2), 16 x 16 for Q1; PFC_LAT_TILE overrides (tuning aid, profiles/tile_sweep.sh)
      static const int tile = getenv("PFC_LAT_TILE") ? atoi(getenv("PFC_LAT_TILE")) : -1;
      switch (tile >= 0 ? tile : (P == 2 ? 1 : 0))
        {
        case 1: return launch_tile<P, MODE, 16, 8>(h, op, ctl);
        case 2: return launch_tile<P, MODE, 32, 8>(h, op, ctl);
        case 3: return launch_tile<P, MODE, 8, 8>(h, op, ctl);
        default: break;
        }
    }
  return launch_tile<P, MODE, kCX, kCY>(h, op, ctl);
}
} // namespace

LatOp lat_op(const pfc_handle *h)
{
  LatOp o;
  std::memset(&o, 0, sizeof(o));
  o.dt = h->prm.dt;
  o.theta = h->prm.theta;
  o.eps = h->prm.eps;
  o.inv_ml = h->d_inv_ml.p;
  o.lo = -1;
  o.add_c = -1;
  return o;
}

int lat_launch(pfc_handle *h, int mode, const LatOp &op, const Ctl *ctl)
{
  const bool q1 = h->plan.degree == 1;
  if (mode == LAT_K && !h->lat_k_by_cells)
    { // node-stencil form of the K product
      static const int variant = getenv("PFC_KS_VARIANT") ? atoi(getenv("PFC_KS_VARIANT")) : 0;
      const LatGeom &G = h->lat_geom;
#define PFC_KS(KX, KY)                                                                   \
  {                                                                                      \
    const dim3 grid((G.w + KX - 1) / KX, (G.h + KY - 1) / KY);                           \
    if (q1)                                                                              \
      lat_kstencil_kernel<1, KX, KY><<<grid, KX, 0, h->stream>>>(G, op, ctl);            \
    else                                                                                 \
      lat_kstencil_kernel<2, KX, KY><<<grid, KX, 0, h->stream>>>(G, op, ctl);            \
  }
#define PFC_KS2(KX, KY)                                                                  \
  {                                                                                      \
    const dim3 grid((G.w + 2 * KX - 1) / (2 * KX), (G.h + KY - 1) / KY);                 \
    if (q1)                                                                              \
      lat_kstencil2_kernel<1, KX, KY><<<grid, KX, 0, h->stream>>>(G, op, ctl);           \
    else                                                                                 \
      lat_kstencil2_kernel<2, KX, KY><<<grid, KX, 0, h->stream>>>(G, op, ctl);           \
  }
      switch (variant)
        {
        case 10: PFC_KS2(64, 4) break;
        case 11: PFC_KS2(128, 4) break;
        case 12: PFC_KS2(64, 8) break;
        case 13: PFC_KS2(64, 2) break;
        case 14: PFC_KS2(32, 4) break;
        case 2: PFC_KS(128, 8) break;
        case 6: PFC_KS(64, 8) break;
        default: PFC_KS(128, 4) break;
        }
#undef PFC_KS
#undef PFC_KS2
      h->launches++;
      return PFC_OK;
    }
  switch (mode)
    {
    case LAT_J:
      if (q1 && h->q1_fast)
        { // closed-form Q1 algebra, constant blocks as stencils, cp.async pipeline over a tile row (pfc_ring_q1.cuh)
          const int variant = h->j1_variant; // PFC_J1_VARIANT at pfc_setup (tuning aid)
          const LatGeom &G = h->lat_geom;
#define PFC_J1(CX, CY, NTILE, MINB, ...)                                                                          \
  {                                                                                                           \
    const int ntx = (G.ncx + CX - 2) / (CX - 1), nty = (G.ncy + CY - 2) / (CY - 1);                           \
    lat_j1_kernel<CX, CY, NTILE, MINB, ##__VA_ARGS__><<<dim3((ntx + NTILE - 1) / NTILE, nty), CX * CY, 0, h->stream>>>(G, op, ctl); \
  }
          switch (variant)
            {
            case 36: PFC_J1(32, 8, 4, 3, true) break;
            case 37: PFC_J1(32, 8, 16, 3, true) break;
            case 38: PFC_J1(32, 6, 8, 4, true) break;
            case 30: PFC_J1(32, 4, 8, 6, true) break;
            case 33: PFC_J1(32, 4, 8, 7, true) break;
            case 34: PFC_J1(32, 8, 8, 4, true) break;
            case 35: PFC_J1(32, 8, 8, 3, true) break;
            case 0: PFC_J1(16, 16, 1, 4) break;
            case 2: PFC_J1(16, 16, 8, 2) break;
            case 3: PFC_J1(32, 8, 4, 3) break;
            case 4: PFC_J1(16, 8, 1, 8) break;
            case 5: PFC_J1(16, 8, 8, 6) break;
            case 7: PFC_J1(16, 16, 4, 4) break;
            case 8: PFC_J1(16, 16, 4, 2) break;
            case 1: PFC_J1(16, 16, 4, 3) break;
            case 9: PFC_J1(64, 4, 4, 3) break;
            case 10: PFC_J1(32, 4, 16, 6) break;
            case 11: PFC_J1(32, 4, 4, 6) break;
            case 12: PFC_J1(32, 8, 8, 3) break;
            case 13: PFC_J1(64, 4, 8, 3) break;
            case 6: PFC_J1(32, 4, 8, 6) break; // best with the packed node phase (profiles/r02k_j1_tuning.log): 413 us on M4
            default: PFC_J1(32, 8, 8, 3, true) break; // 35: measured best on B200 (profiles/r04c_j1_rowmap.log, r04e_*): 378 us
            }
#undef PFC_J1
          h->launches++;
          return PFC_OK;
        }
      return q1 ? launch_mode<1, LAT_J>(h, op, ctl) : launch_mode<2, LAT_J>(h, op, ctl);
    case LAT_J0:
      {
        if (!(q1 && h->q1_fast))
          return h->fail(PFC_ERR_UNSUPPORTED, "LAT_J0 is a Q1 closed-form kernel");
        const LatGeom &G = h->lat_geom;
        constexpr int CX = 32, CY = 8, NTILE = 8;
        const int ntx = (G.ncx + CX - 2) / (CX - 1), nty = (G.ncy + CY - 2) / (CY - 1);
        lat_j1_kernel<CX, CY, NTILE, 3, true, true><<<dim3((ntx + NTILE - 1) / NTILE, nty), CX * CY, 0, h->stream>>>(G, op, ctl);
        h->launches++;
        return PFC_OK;
      }
    case LAT_P:
      {
        if (!q1)
          return h->fail(PFC_ERR_UNSUPPORTED, "LAT_P is a Q1 kernel");
        const LatGeom &G = h->lat_geom;
        constexpr int TX = 64, TY = 16;
        lat_pstage_kernel<TX, TY, 256><<<dim3((G.w + 2 + TX - 1) / TX, (G.h + 2 + TY - 1) / TY), 256, 0, h->stream>>>(G, op, ctl);
        h->launches++;
        return PFC_OK;
      }
    case LAT_S: return q1 ? launch_mode<1, LAT_S>(h, op, ctl) : launch_mode<2, LAT_S>(h, op, ctl);
    case LAT_K: return q1 ? launch_mode<1, LAT_K>(h, op, ctl) : launch_mode<2, LAT_K>(h, op, ctl);
    case LAT_R: return q1 ? launch_mode<1, LAT_R>(h, op, ctl) : launch_mode<2, LAT_R>(h, op, ctl);
    case LAT_T: return q1 ? launch_mode<1, LAT_T>(h, op, ctl) : launch_tile<2, LAT_T, 16, 8>(h, op, ctl);
    case LAT_M:
      {
        const LatGeom &G = h->lat_geom;
        const dim3 grid((G.w + 127) / 128, (G.h + 3) / 4, 4);
        if (q1)
          lat_mass_kernel<1, 128, 4><<<grid, 128, 0, h->stream>>>(G, op);
        else
          lat_mass_kernel<2, 128, 4><<<grid, 128, 0, h->stream>>>(G, op);
        h->launches++;
        return PFC_OK;
      }
    }
  return h->fail(PFC_ERR_BAD_ARGUMENT, "lat_launch: unknown mode");
}

// ---- streaming stencil kernels of the ring path (pfc_ring.cuh) ----------------------------------------------------
namespace
{
constexpr int kRingThreads = 128;

// Rows per strip.  A strip of rs rows costs rs + warm row iterations (warm = the pipeline's warm-up rows, repeated by every
// strip), and the CTAs run in waves of n_sm x (resident CTAs per SM): the choice minimises waves x (rs + warm) over
// 16 <= rs <= 64.  Longer strips lose even when they make exactly one wave (the CTAs on the boundary of the owned
// rectangle take the general index path and finish last: 260 instead of 226 us per sweep on M4 with 317-row strips);
// shorter ones pay the warm-up too often (profiles/r02u_rs_sweep.log).
int ring_rows_per_strip(const pfc_handle *h, int col_blocks, int rows, int warm, int resident)
{
  static const int forced = getenv("PFC_RING_RS") ? atoi(getenv("PFC_RING_RS")) : 0; // tuning aid
  if (forced > 0)
    return forced;
  const int64_t slots = (int64_t)h->n_sm * std::max(1, resident);
  int best = 16;
  int64_t best_cost = INT64_MAX;
  for (int rs = 16; rs <= std::max(16, std::min(rows, 64)); ++rs)
    {
      const int64_t ctas = (int64_t)col_blocks * ((rows + rs - 1) / rs);
      const int64_t cost = ((ctas + slots - 1) / slots) * (rs + warm);
      if (cost <= best_cost) // ties: the longer strip
        best_cost = cost, best = rs;
    }
  return best;
}

// resident CTAs per SM of a kernel (cached per function)
template <class K>
int ring_resident(K kernel)
{
  static K seen[16];
  static int occ[16], n_seen = 0;
  for (int i = 0; i < n_seen; ++i)
    if (seen[i] == kernel)
      return occ[i];
  int nb = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, kRingThreads, 0) != cudaSuccess || nb < 1)
    nb = 1, cudaGetLastError();
  if (n_seen < 16)
    seen[n_seen] = kernel, occ[n_seen++] = nb;
  return nb;
}

template <int STAGES, class K1, class K2>
int ring_launch(pfc_handle *h, RingOp op, const Ctl *ctl, K1 q1, K2 q2)
{
  const LatGeom &G = h->lat_geom;
  const int p = h->plan.degree, outw = kRingThreads - 2 * STAGES * p;
  const int cols = G.fw - 2 * op.lo, rows = G.fh - 2 * op.lo;
  if (cols <= 0 || rows <= 0 || op.lo < STAGES * p)
    return h->fail(PFC_ERR_STATE, "ring kernel: window margin %d outside the ring", op.lo);
  const int cb = (cols + outw - 1) / outw;
  const int warm = 2 * STAGES * p + STAGES - 1; // rows before the first and after the last output row of a strip
  op.rs = ring_rows_per_strip(h, cb, rows, warm, p == 1 ? ring_resident(q1) : ring_resident(q2));
  const dim3 grid(cb, (rows + op.rs - 1) / op.rs);
  if (p == 1)
    q1<<<grid, kRingThreads, 0, h->stream>>>(G, op, ctl);
  else
    q2<<<grid, kRingThreads, 0, h->stream>>>(G, op, ctl);
  h->launches++;
  return PFC_OK;
}
} // namespace

// prefetch depth of the register rings: 1 row, or (deep) one turn of the row phase = 2p + 1 rows; PFC_RING_PF=1|2 forces
// shallow | deep (tuning aid)
static bool ring_deep(bool dflt)
{
  static const int forced = getenv("PFC_RING_PF") ? atoi(getenv("PFC_RING_PF")) : 0;
  return forced > 0 ? forced > 1 : dflt;
}
#define PFC_RING_DISPATCH(STAGES, KERNEL, DEEP1, DEEP2)                                                         \
  if (ring_deep(h->plan.degree == 1 ? DEEP1 : DEEP2))                                                           \
    return ring_launch<STAGES>(h, op, ctl, KERNEL<1, kRingThreads, 3>, KERNEL<2, kRingThreads, 5>);             \
  return ring_launch<STAGES>(h, op, ctl, KERNEL<1, kRingThreads, 1>, KERNEL<2, kRingThreads, 1>);
// Q1: closed-form variants (pfc_ring_q1.cuh) unless switched off (self-check)
#define PFC_RING_DISPATCH1(STAGES, KERNEL1, DEEP1)                                                              \
  if (h->plan.degree == 1 && h->q1_fast)                                                                        \
    {                                                                                                           \
      if (ring_deep(DEEP1))                                                                                     \
        return ring_launch<STAGES>(h, op, ctl, KERNEL1<kRingThreads, 3>, KERNEL1<kRingThreads, 3>);             \
      return ring_launch<STAGES>(h, op, ctl, KERNEL1<kRingThreads, 1>, KERNEL1<kRingThreads, 1>);               \
    }
int ring_launch_t(pfc_handle *h, const RingOp &op, const Ctl *ctl) { PFC_RING_DISPATCH(1, ring_t_kernel, true, false) }
int ring_launch_sweep(pfc_handle *h, const RingOp &op, const Ctl *ctl)
{
  if (op.first)
    {
      if (!(h->plan.degree == 1 && h->q1_fast && op.epi == 0))
        return h->fail(PFC_ERR_UNSUPPORTED, "ring sweep: the first-step form is a Q1 closed-form kernel");
      return ring_launch<3>(h, op, ctl, ring_sweep1a_kernel<kRingThreads, 4, true>, ring_sweep1a_kernel<kRingThreads, 4, true>);
    }
  static const int staged = getenv("PFC_SWEEP_STAGED") ? atoi(getenv("PFC_SWEEP_STAGED")) : 4; // 0: off; 3, 4, 6: distance
  if (h->plan.degree == 1 && h->q1_fast && op.epi == 0 && staged > 0) // Chebyshev step: operands staged by cp.async
    {
      if (staged == 3)
        return ring_launch<3>(h, op, ctl, ring_sweep1a_kernel<kRingThreads, 3>, ring_sweep1a_kernel<kRingThreads, 3>);
      if (staged == 6)
        return ring_launch<3>(h, op, ctl, ring_sweep1a_kernel<kRingThreads, 6>, ring_sweep1a_kernel<kRingThreads, 6>);
      return ring_launch<3>(h, op, ctl, ring_sweep1a_kernel<kRingThreads, 4>, ring_sweep1a_kernel<kRingThreads, 4>);
    }
  PFC_RING_DISPATCH1(3, ring_sweep1_kernel, true)
  PFC_RING_DISPATCH(3, ring_sweep_kernel, true, false)
}
int ring_launch_back(pfc_handle *h, const RingOp &op, const Ctl *ctl)
{
  static const int staged = getenv("PFC_BACK_STAGED") ? atoi(getenv("PFC_BACK_STAGED")) : 4; // 0: off (tuning / A-B aid)
  if (h->plan.degree == 1 && h->q1_fast && staged > 0)
    return ring_launch<2>(h, op, ctl, ring_back1a_kernel<kRingThreads, 4>, ring_back1a_kernel<kRingThreads, 4>);
  PFC_RING_DISPATCH1(2, ring_back1_kernel, false)
  PFC_RING_DISPATCH(2, ring_back_kernel, false, false)
}
int ring_launch_rhs1(pfc_handle *h, const RingOp &op, const Ctl *ctl)
{
  if (h->plan.degree != 1)
    return h->fail(PFC_ERR_UNSUPPORTED, "ring_launch_rhs1 is a Q1 kernel");
  static const int staged = getenv("PFC_RHS_STAGED") ? atoi(getenv("PFC_RHS_STAGED")) : 4; // 0: off (tuning / A-B aid)
  if (staged > 0)
    return ring_launch<2>(h, op, ctl, ring_rhs1a_kernel<kRingThreads, 4>, ring_rhs1a_kernel<kRingThreads, 4>);
  return ring_launch<2>(h, op, ctl, ring_rhs1_kernel<kRingThreads>, ring_rhs1_kernel<kRingThreads>);
}
#undef PFC_RING_DISPATCH1
#undef PFC_RING_DISPATCH

// nodes N = w h; every field is read once (the halo ring through L2) and every output written once
double lat_bytes(const pfc_handle *h, int mode)
{
  const double N = (double)h->lat_geom.w * h->lat_geom.h;
  switch (mode)
    {
    case LAT_J: return 64.0 * N;  // x 24 + y 24 + psi, chi 16
    case LAT_S: return 80.0 * N;  // v0, v1, v2, psi, chi 40 + add, dinv, d 24 + d, out 16   (Chebyshev step)
    case LAT_K: return 32.0 * N;  // v 8 + add 8 + inv_ml 8 + out 8
    case LAT_T: return 48.0 * N;  // psi, chi 16 + four right-hand sides 32
    case LAT_M: return 64.0 * N;  // four fields in, four out
    case RING_T: return 32.0 * N;     // r1, r2 in, t, m1 out
    case RING_SWEEP: return 48.0 * N; // x, rhs, dinv, d in, d, x out
    case RING_BACK: return 40.0 * N;  // psi, r1, r2 in, chi, phi out
    case RING_RHS1: return 64.0 * N;  // r0, r1, r2, psi, dinv in, rhs, d, x out (t kernel fused, Q1)
    case LAT_J0: return 48.0 * N;     // three fields 24 + psi, chi 16 + out 8
    case LAT_P: return 32.0 * N;      // z in, three fields out
    case RING_RHS: return 72.0 * N;   // m1, t, psi, chi, r0, dinv in, rhs, d, x out
    default: return (h->prm.theta == 1.0 ? 32.0 : 48.0) * N + 32.0 * N; // state in, rhs + 1/diag(B) out
    }
}

} // namespace pfc

extern "C" int pfc_selfcheck(pfc_handle *h, double *max_rel)
{
  using namespace pfc;
  if (!h || !max_rel)
    return PFC_ERR_BAD_ARGUMENT;
  *max_rel = 0.0;
  if (!h->is_setup)
    return h->fail(PFC_ERR_STATE, "pfc_selfcheck before pfc_setup");
  if (!h->lat)
    return PFC_OK;
  PFC_CUDA(h, cudaSetDevice(h->device));
  const Plan &P = h->plan;
  const int64_t n = P.Np + P.n_ghost;
  if (h->stride < n)
    return h->fail(PFC_ERR_STATE, "pfc_selfcheck: work space too small");
  double *x = h->d_V.p, *ya = h->d_V.p + h->stride, *yb = h->d_V.p + 2 * h->stride;
  std::vector<double> hx((size_t)n, 0.0), a((size_t)P.Np), b((size_t)P.Np);
  for (int64_t r = 0; r < P.Np; ++r)
    if (P.perm[r] >= 0)
      hx[r] = 0.5 + (double)(((uint64_t)P.perm[r] * 0x9E3779B97F4A7C15ull >> 40) & 0xFFFF) / 65536.0;
  cudaStream_t s = h->stream;
  PFC_CUDA(h, cudaMemcpyAsync(x, hx.data(), n * sizeof(double), cudaMemcpyHostToDevice, s));
  PFC_CUDA(h, cudaMemsetAsync(ya, 0, 2 * h->stride * sizeof(double), s));
  {
    const int rc = halo_exchange_scalar(h, x); // ring values from their owners
    if (rc != PFC_OK)
      return rc;
  }
  LatOp op = lat_op(h);
  op.s0 = x;
  op.k = 1.0;
  for (int pass = 0; pass < 2; ++pass)
    {
      h->lat_k_by_cells = pass == 1;
      op.out = pass == 0 ? ya : yb;
      const int rc = lat_launch(h, LAT_K, op, nullptr);
      h->lat_k_by_cells = false;
      if (rc != PFC_OK)
        return rc;
    }
  PFC_CUDA(h, cudaGetLastError());
  PFC_CUDA(h, cudaMemcpyAsync(a.data(), ya, P.Np * sizeof(double), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaMemcpyAsync(b.data(), yb, P.Np * sizeof(double), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  double dmax = 0, amax = 0;
  for (int64_t r = 0; r < P.Np; ++r)
    {
      dmax = std::max(dmax, std::fabs(a[r] - b[r]));
      amax = std::max(amax, std::fabs(b[r]));
    }
  // leave the Krylov work space as pfc_setup left it
  PFC_CUDA(h, cudaMemsetAsync(x, 0, 3 * h->stride * sizeof(double), s));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  *max_rel = amax > 0 ? dmax / amax : (dmax > 0 ? 1.0 : 0.0);
  return PFC_OK;
}

namespace pfc
{
int lat_setup(pfc_handle *h)
{
  const Lattice &L = h->plan.lat;
  const Ring &R = h->plan.ring; // padded rectangle of the ring overlay: every position has an arithmetic id
  cudaStream_t s = h->stream;
  LatGeom &G = h->lat_geom;
  h->q1_fast = !(getenv("PFC_Q1_CLOSED_FORM") && atoi(getenv("PFC_Q1_CLOSED_FORM")) == 0); // tuning / A-B aid
  h->j1_variant = getenv("PFC_J1_VARIANT") ? atoi(getenv("PFC_J1_VARIANT")) : 35;              // tuning / A-B aid
  G.ncx = R.ncx;
  G.ncy = R.ncy;
  G.ex0 = R.ex0;
  G.ey0 = R.ey0;
  G.w = L.w;
  G.h = L.h;
  G.fw = R.fw;
  G.fh = R.fh;
  G.Np = h->plan.Np;
  G.hx = L.hx;
  G.hy = L.hy;
  G.ihx = 1.0 / L.hx;
  G.ihy = 1.0 / L.hy;
  if (h->d_inv_ml.n != (size_t)h->plan.Np)
    PFC_CUDA(h, h->d_inv_ml.alloc(h->plan.Np));
  // 1 / diag(B) lives on the ring too (the Chebyshev sweeps run there): [owned | ghosts]
  PFC_CUDA(h, h->d_dinvB.alloc((size_t)(h->plan.Np + h->plan.n_ghost)));
  PFC_CUDA(h, cudaMemsetAsync(h->d_dinvB.p, 0, h->d_dinvB.n * sizeof(double), s));
  const int64_t n = h->plan.Np;
  if (h->plan.degree == 1)
    lat_inv_ml_kernel<1><<<(int)((n + 255) / 256), 256, 0, s>>>(G, h->d_inv_ml.p, n);
  else
    lat_inv_ml_kernel<2><<<(int)((n + 255) / 256), 256, 0, s>>>(G, h->d_inv_ml.p, n);
  lat_inv_ml_kernel<1><<<(int)((n + 255) / 256), 256, 0, s>>>(G, h->d_dinvB.p, n); // placeholder values (1 on slaves/padding)
  h->launches += 2;
  PFC_CUDA(h, cudaGetLastError());
  PFC_CUDA(h, cudaStreamSynchronize(s));
  return PFC_OK;
}
} // namespace pfc

// S~ x by the streaming sweep kernel (pfc_ring.cuh) against the composition of the K-stencil and mass-stencil kernels with
// a halo exchange before every product: max |a - b| / max |b| over the owned free nodes.  [collective]
extern "C" int pfc_selfcheck_smoother(pfc_handle *h, double *max_rel)
{
  using namespace pfc;
  if (!h || !max_rel)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup || !h->lat)
    return h->fail(PFC_ERR_UNSUPPORTED, "pfc_selfcheck_smoother needs the lattice path");
  PFC_CUDA(h, cudaSetDevice(h->device));
  const LatGeom &G = h->lat_geom;
  const Plan &P = h->plan;
  const int p = P.degree;
  if (h->plan.ring.D < 3 * p)
    return h->fail(PFC_ERR_UNSUPPORTED, "pfc_selfcheck_smoother needs a ring of depth >= 3 degree (preconditioner on)");
  const int64_t n = (int64_t)G.w * G.h, ss = (P.Np + P.n_ghost + 15) / 16 * 16;
  const double bbar = 0.9, dtth = h->prm.dt * h->prm.theta;
  DevBuf<double> F, Gm, Kt2, out;
  PFC_CUDA(h, F.alloc((size_t)3 * ss));   // x, t1, t2
  PFC_CUDA(h, Gm.alloc((size_t)4 * ss));  // M x, M t1, M t2
  PFC_CUDA(h, Kt2.alloc((size_t)ss));
  PFC_CUDA(h, out.alloc((size_t)ss));
  std::vector<double> x((size_t)ss, 0.0);
  for (int64_t r = 0; r < n; ++r)
    { // tied to the global node: the same field on every partition
      uint64_t st = (uint64_t)P.perm[r] * 0x9E3779B97F4A7C15ull + 88172645463325252ull;
      st ^= st << 13, st ^= st >> 7, st ^= st << 17;
      x[r] = (double)(st >> 11) / 9007199254740992.0 - 0.5;
    }
  cudaStream_t s = h->stream;
  PFC_CUDA(h, cudaMemsetAsync(F.p, 0, F.n * sizeof(double), s));
  PFC_CUDA(h, cudaMemsetAsync(Gm.p, 0, Gm.n * sizeof(double), s));
  PFC_CUDA(h, cudaMemsetAsync(Kt2.p, 0, Kt2.n * sizeof(double), s));
  PFC_CUDA(h, cudaMemsetAsync(out.p, 0, out.n * sizeof(double), s));
  PFC_CUDA(h, cudaMemcpyAsync(F.p, x.data(), ss * sizeof(double), cudaMemcpyHostToDevice, s));
  auto kprod = [&](double *in, double *o) { // o = (K in) M_L^-1
    int rc = halo_exchange_scalar(h, in);
    if (rc != PFC_OK)
      return rc;
    LatOp q = lat_op(h);
    q.s0 = in;
    q.k = 1.0;
    q.out = o;
    return lat_launch(h, LAT_K, q, nullptr);
  };
  int rc = kprod(F.p, F.p + ss);
  if (rc == PFC_OK)
    rc = kprod(F.p + ss, F.p + 2 * ss);
  if (rc == PFC_OK)
    rc = kprod(F.p + 2 * ss, Kt2.p);
  if (rc != PFC_OK)
    return rc;
  {
    LatOp q = lat_op(h);
    q.s0 = F.p;
    q.out = Gm.p;
    q.zstride = ss;
    q.zmask = 7;
    rc = lat_launch(h, LAT_M, q, nullptr);
    if (rc != PFC_OK)
      return rc;
  }
  RingOp so;
  std::memset(&so, 0, sizeof(so));
  so.x = F.p;
  so.out = out.p;
  so.bbar = bbar;
  so.dtth = dtth;
  so.lo = 3 * p;
  so.epi = 3;
  rc = ring_launch_sweep(h, so, nullptr);
  if (rc != PFC_OK)
    return rc;
  std::vector<double> gm((size_t)3 * ss), kt2((size_t)ss), iml((size_t)n), xo((size_t)ss);
  PFC_CUDA(h, cudaMemcpyAsync(gm.data(), Gm.p, gm.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaMemcpyAsync(kt2.data(), Kt2.p, kt2.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaMemcpyAsync(iml.data(), h->d_inv_ml.p, iml.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaMemcpyAsync(xo.data(), out.p, xo.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  PFC_CUDA(h, cudaGetLastError());
  double worst = 0, scale = 0;
  for (int64_t r = 0; r < n; ++r)
    {
      const double ref = bbar * gm[r] + dtth * (kt2[r] / iml[r] - 2.0 * gm[2 * ss + r]);
      worst = std::max(worst, std::fabs(xo[r] - ref));
      scale = std::max(scale, std::fabs(ref));
    }
  *max_rel = scale > 0 ? worst / scale : worst;
  return PFC_OK;
}
