// pfc_lattice_args.h — argument blocks of the lattice kernels (pfc_lattice.cuh) and their host launcher
// (pfc_lattice.cu).
#pragma once
#include <cstdint>

struct pfc_handle;

namespace pfc
{
struct Ctl;
// ---- kernel arguments ---------------------------------------------------------------------------------
struct LatGeom
{
  int32_t ncx, ncy, ex0, ey0, w, h, fw, fh;
  const int32_t *lut;
  int64_t Np;
  double hx, hy, ihx, ihy; // cell size and its reciprocals
};

enum
{
  LAT_J = 0, // y = J x                       (vec3 in3a = x, state in3b; out3 = y)
  LAT_S = 1, // y0 = sB B v0 + sC C v1 + sD D v2 on scalar fields, epilogue `epi` (see bcd_kernel, pfc_precond.cuh)
  LAT_K = 2, // out = (a add + k K s0) inv_ml  (kmv_kernel)
  LAT_R = 3, // out3 = -R(Psi_n = in3a, Psi_n-1 = in3b), out = 1 / diag(B)   (assemble_system, mpi.cpp:388-417)
  LAT_T = 4, // out[c zstride + r] = right-hand side of the stress projection, c = a*2 + b, from the state in3a
             // (stress_calculator_mpi.cpp:204-314)
  LAT_M = 5  // out[c zstride + r] = (M s0[c zstride + .])_r for the fields of zmask: consistent mass matrix as a
             // node stencil (stress_calculator_mpi.cpp:141-200)
};

struct LatOp
{
  const double *in3a, *in3b;   // 3-component vectors (vaddr layout)
  const double *s0, *s1, *s2;  // scalar fields [owned Np | ghosts]
  const double *add, *dinv, *cheb, *inv_ml;
  double *d, *out, *out2, *out3;
  double a, k, sB, sC, sD;
  double dt, theta, eps;
  int epi;
  int64_t zstride; // LAT_T / LAT_M: doubles between the batched scalar fields (tensor components)
  int zmask;       // LAT_M: bit c set = field c is processed
  const int *zmask_dev; // LAT_M: if set, the mask is read from the device (and-ed with zmask)
};


// EXPERIMENTAL constant-coefficient Chebyshev sweep (pfc_lattice_smooth.cuh; PFC_SMOOTHER=1, single rank)
struct SmoothOp
{
  const double *x, *rhs, *dinv, *cheb; // cheb[0] = coefficient of d, cheb[1] = coefficient of dinv * residual
  double *d, *xout;
  double bbar, cbar, dtth;
};
int lat_smooth(pfc_handle *h, const SmoothOp &op, const Ctl *ctl);

// host side (pfc_lattice.cu)
int lat_setup(pfc_handle *h);                                              // after build_plan: lut upload, inv_ml
int lat_launch(pfc_handle *h, int mode, const LatOp &op, const Ctl *ctl);  // stream-ordered, graph-capturable
LatOp lat_op(const pfc_handle *h);                                          // zeroed LatOp with dt, theta, eps
double lat_bytes(const pfc_handle *h, int mode);                            // algorithmic bytes of one launch
} // namespace pfc
