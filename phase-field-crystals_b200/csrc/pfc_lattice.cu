// pfc_lattice.cu — host launcher of the lattice kernels (pfc_lattice.cuh).
#include <cstring>

#include "pfc_handle.h"
#include "pfc_lattice.cuh"

namespace pfc
{
namespace
{
// cells per CTA (one thread per cell); the inner (CX-1) x (CY-1) cells are the CTA's own
constexpr int kCX = 16, kCY = 16;

template <int P, int MODE>
int launch_mode(pfc_handle *h, const LatOp &op, const Ctl *ctl)
{
  const LatGeom &G = h->lat_geom;
  const dim3 grid((G.ncx + kCX - 2) / (kCX - 1), (G.ncy + kCY - 2) / (kCY - 1));
  lat_kernel<P, MODE, kCX, kCY><<<grid, kCX * kCY, lat_smem_bytes<P, MODE, kCX, kCY>(), h->stream>>>(G, op, ctl);
  h->launches++;
  return PFC_OK;
}

template <int P, int MODE>
cudaError_t opt_in_smem()
{
  return cudaFuncSetAttribute(lat_kernel<P, MODE, kCX, kCY>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)lat_smem_bytes<P, MODE, kCX, kCY>());
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
  return o;
}

int lat_launch(pfc_handle *h, int mode, const LatOp &op, const Ctl *ctl)
{
  const bool q1 = h->plan.degree == 1;
  switch (mode)
    {
    case LAT_J: return q1 ? launch_mode<1, LAT_J>(h, op, ctl) : launch_mode<2, LAT_J>(h, op, ctl);
    case LAT_S: return q1 ? launch_mode<1, LAT_S>(h, op, ctl) : launch_mode<2, LAT_S>(h, op, ctl);
    case LAT_K: return q1 ? launch_mode<1, LAT_K>(h, op, ctl) : launch_mode<2, LAT_K>(h, op, ctl);
    case LAT_R: return q1 ? launch_mode<1, LAT_R>(h, op, ctl) : launch_mode<2, LAT_R>(h, op, ctl);
    }
  return h->fail(PFC_ERR_BAD_ARGUMENT, "lat_launch: unknown mode");
}

// nodes N = w h; every field is read once (the halo ring through L2) and every output written once
double lat_bytes(const pfc_handle *h, int mode)
{
  const double N = (double)h->lat_geom.w * h->lat_geom.h;
  switch (mode)
    {
    case LAT_J: return 64.0 * N;  // x 24 + y 24 + psi, chi 16
    case LAT_S: return 80.0 * N;  // v0, v1, v2, psi, chi 40 + add, dinv, d 24 + d, out 16   (Chebyshev step)
    case LAT_K: return 32.0 * N;  // v 8 + add 8 + inv_ml 8 + out 8
    default: return (h->prm.theta == 1.0 ? 32.0 : 48.0) * N + 32.0 * N; // state in, rhs + 1/diag(B) out
    }
}

int lat_setup(pfc_handle *h)
{
  const Lattice &L = h->plan.lat;
  cudaStream_t s = h->stream;
  PFC_CUDA(h, h->d_lut.upload(L.lut, s));
  LatGeom &G = h->lat_geom;
  G.ncx = L.ncx;
  G.ncy = L.ncy;
  G.ex0 = L.ex0;
  G.ey0 = L.ey0;
  G.w = L.w;
  G.h = L.h;
  G.fw = L.fw;
  G.fh = L.fh;
  G.lut = h->d_lut.p;
  G.Np = h->plan.Np;
  G.hx = L.hx;
  G.hy = L.hy;
  PFC_CUDA(h, (opt_in_smem<1, LAT_J>()));
  PFC_CUDA(h, (opt_in_smem<1, LAT_S>()));
  PFC_CUDA(h, (opt_in_smem<1, LAT_K>()));
  PFC_CUDA(h, (opt_in_smem<1, LAT_R>()));
  PFC_CUDA(h, (opt_in_smem<2, LAT_J>()));
  PFC_CUDA(h, (opt_in_smem<2, LAT_S>()));
  PFC_CUDA(h, (opt_in_smem<2, LAT_K>()));
  PFC_CUDA(h, (opt_in_smem<2, LAT_R>()));
  if (h->d_inv_ml.n != (size_t)h->plan.Np)
    PFC_CUDA(h, h->d_inv_ml.alloc(h->plan.Np));
  if (h->d_dinvB.n != (size_t)h->plan.Np)
    PFC_CUDA(h, h->d_dinvB.alloc(h->plan.Np));
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
