// pfc_lattice_args.h — argument blocks of the lattice kernels (pfc_lattice.cuh) and their host launcher
// (pfc_lattice.cu).
#pragma once
#include <cstdint>

struct pfc_handle;

namespace pfc
{
struct Ctl;
// ---- kernel arguments ---------------------------------------------------------------------------------
// Padded rectangle of the ring overlay (Ring in pfc_plan.h): positions [0, fw) x [0, fh), owned free nodes at
// [ex0, ex0 + w) x [ey0, ey0 + h) with id ly w + lx, everything else a ghost with id Np + ring index (lat_id below);
// (0, 0) is a cell corner, ncx x ncy cells.
struct LatGeom
{
  int32_t ncx, ncy, ex0, ey0, w, h, fw, fh;
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
  LAT_M = 5, // out[c zstride + r] = (M s0[c zstride + .])_r for the fields of zmask: consistent mass matrix as a
             // node stencil (stress_calculator_mpi.cpp:141-200)
  RING_T = 6, RING_SWEEP = 7, RING_BACK = 8, RING_RHS = 9, RING_RHS1 = 10, // legs of the ring path (lat_bytes)
  LAT_J0 = 11, // out3[r] = (J x)_psi only, out3 a scalar field (Q1 closed form; operator of the reduced solve, pfc_reduced.cuh)
  LAT_P = 12   // out3 = (M^2 s0, -K M s0, K^2 s0) on the owned nodes and one node beyond (pfc_reduced.cuh)
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
  int lo;          // LAT_S: output window = padded positions at least lo nodes inside (< 0: the owned free nodes)
  int add_c;       // LAT_S: `add` is component add_c of a 3-component vector (< 0: scalar field)
  int64_t zstride; // LAT_T / LAT_M: doubles between the batched scalar fields (tensor components)
  int zmask;       // LAT_M: bit c set = field c is processed
  const int *zmask_dev; // LAT_M: if set, the mask is read from the device (and-ed with zmask)
};


// streaming stencil kernels of the ring path (pfc_ring.cuh): t kernel, Chebyshev sweep, back substitution
struct RingOp
{
  const double *in;          // 3-component input vector r = (r0, r1, r2) of the preconditioner
  const double *x;           // sweep: scalar field x (or component 0 of xv when x_vec)
  const double *rhs, *dinv, *cheb;
  double *d, *t, *m1;        // Chebyshev direction; t kernel outputs
  double *out;               // sweep output: scalar field, or 3-component vector when out_vec (component 0)
  double *z;                 // back substitution: 3-component vector, x in component 0, writes chi, phi
  double bbar, dtth;
  int lo, rs, epi, out_vec, x_vec;
  // Q1 fused right-hand side kernel (pfc_ring_q1.cuh)
  const double *psi;         // state Psi_n (3-component vector): coefficient of C from its psi
  double *out2;              // first Chebyshev iterate
  double eps;
  // reduced solve (pfc_reduced.cuh): first != 0: x_0 = d_0 = alpha rhs is not read but made from rhs, dinv = dinvc
  int first;
  double alpha, dinvc;
};
int ring_launch_rhs1(pfc_handle *h, const RingOp &op, const Ctl *ctl); // Q1 only
int ring_launch_t(pfc_handle *h, const RingOp &op, const Ctl *ctl);
int ring_launch_sweep(pfc_handle *h, const RingOp &op, const Ctl *ctl);
int ring_launch_back(pfc_handle *h, const RingOp &op, const Ctl *ctl);

// host side (pfc_lattice.cu)
int lat_setup(pfc_handle *h);                                              // after build_plan: lut upload, inv_ml
int lat_launch(pfc_handle *h, int mode, const LatOp &op, const Ctl *ctl);  // stream-ordered, graph-capturable
LatOp lat_op(const pfc_handle *h);                                          // zeroed LatOp with dt, theta, eps
double lat_bytes(const pfc_handle *h, int mode);                            // algorithmic bytes of one launch
} // namespace pfc
