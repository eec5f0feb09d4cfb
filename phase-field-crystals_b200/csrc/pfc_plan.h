// pfc_plan.h — host-side products of pfc_setup: everything the kernels index with.
//
// The reference builds these through deal.II in setup_dofs (mpi.cpp:150-274): DoF maps,
// AffineConstraints, BlockSparsityPattern.  Here the flattened host arrays of pfc_mesh are
// turned into the device layout described in DESIGN.md ("Data layout in HBM").
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/pfc_b200.h"

namespace pfc
{
constexpr int kSlice = 32;      // SELL slice height = one warp
constexpr int kSigma = 4096;    // sorting window (rows) of SELL-C-sigma

// Uniform tensor lattice view of the local mesh (pfc_lattice_plan.cpp).  The reference only ever builds one
// hyper_rectangle refined globally (mpi.cpp:119-122,145): identical axis-aligned cells, periodic in x and y.
// When the arrays handed to pfc_setup have that structure the matrix-free path numbers the owned free
// nodes lexicographically and addresses everything by lattice arithmetic (no per-cell index tables).
// Local frame: ncx x ncy cells = every cell that touches an owned free node; frame node (ex, ey),
// 0 <= ex <= p ncx, holds the resolved (periodic master) node at that position; the owned free nodes are
// the rectangle [ex0, ex0 + w) x [ey0, ey0 + h) with internal id (ey - ey0) w + (ex - ex0); every other
// frame position goes through lut (ghost ids, or owned ids of periodic images).
struct Lattice
{
  bool active = false;
  int p = 0;
  int32_t ncx = 0, ncy = 0, ex0 = 0, ey0 = 0, w = 0, h = 0, fw = 0, fh = 0;
  double hx = 0, hy = 0;
  std::vector<int32_t> lut;        // [fh][fw] internal id (after numbering)
  std::vector<int32_t> lut_global; // [fh][fw] global resolved node (detection product)
  std::vector<int32_t> order;      // [w h] global node of owned free node (ly, lx), lexicographic
  std::string why;                 // why the lattice path is not available
  // what the ring overlay (pfc_ring_plan.cpp) needs to place this rank on the global periodic lattice
  double ox = 0, oy = 0;           // physical position of the owned free node (0, 0)
  int64_t Px = 0, Py = 0;          // period in nodes as far as this rank can see it (0: no periodic pair among its nodes)
  std::vector<int32_t> slave_fx, slave_fy; // frame position of the owned slaves, in the order of their internal ids
};

// One rank's place on the global lattice: what the ranks all-gather once in pfc_setup (lattice path).
struct RingCard
{
  double ox, oy, hx, hy;     // physical position of owned free node (0, 0); cell size
  int64_t Px, Py;            // period in nodes (0 if unknown to the rank)
  int32_t w, h, ex0, ey0;    // owned rectangle; its position in the rank's one-cell frame (ex0 mod p = node type)
  int32_t p, active;
};

// Deep ghost ring of the lattice path (pfc_ring_plan.cpp, DESIGN.md "ring").  The rank's local index space becomes the
// padded rectangle [0, fw) x [0, fh) of lattice positions with the owned free nodes at [ex0, ex0 + w) x [ey0, ey0 + h)
// and padded (0, 0) on a cell corner; every position outside the owned rectangle is a ghost with id Np + ring index
// (arithmetic, see ring_index), filled by ONE halo exchange from whichever rank owns that position modulo the period
// (possibly this rank itself).  A ring of depth D lets the whole preconditioner + operator chain of a GMRES iteration run
// on local data with shrinking validity: 14 -> 1 exchanges per iteration.
struct Ring
{
  bool active = false;
  int32_t D = 0;                      // requested depth (every margin is >= D)
  int32_t ex0 = 0, ey0 = 0, fw = 0, fh = 0, ncx = 0, ncy = 0; // padded geometry (replaces the frame of Lattice)
  int64_t n_ring = 0;                 // fw fh - w h ring positions (+ 1 scratch slot at index n_ring)
  std::vector<int32_t> self_src, self_dst; // ring positions owned by this rank itself: tail[self_dst] = owned[self_src]
  std::vector<int32_t> recv_pos;      // [recv_ptr.back()] ring index of the k-th received value (peer-major order)
};

inline int64_t ring_index(int32_t X, int32_t Y, int32_t ex0, int32_t ey0, int32_t w, int32_t h, int32_t fw, int32_t fh)
{
  if (Y < ey0)
    return (int64_t)Y * fw + X;
  if (Y >= ey0 + h)
    return (int64_t)(Y - h) * fw + X;
  return (int64_t)(fh - h) * fw + (int64_t)(Y - ey0) * (fw - w) + (X < ex0 ? X : X - w);
}

enum
{
  PLAN_LATTICE = 1,      // try the lattice numbering (matrix-free path)
  PLAN_SKIP_TABLES = 2   // with an active lattice: skip SELL pattern, colouring and per-cell tables
};

struct Plan
{
  // sizes
  int degree = 0, n_s = 0, n_q = 0, dpc = 0;
  int64_t n_cells = 0;          // cells this rank integrates
  int64_t N_global = 0;         // scalar nodes per component, global
  int64_t n_dofs = 0;           // 3 N_global
  int64_t block_start[3] = {0, 0, 0};
  int rank = 0, n_ranks = 1;
  std::vector<int64_t> node_offsets; // [n_ranks+1]
  int64_t own0 = 0;             // first owned global node
  int64_t n_owned = 0;          // owned scalar nodes
  int64_t Np = 0;               // n_owned rounded up to kSlice: rows of the SELL matrix
  int64_t n_ghost = 0;          // ghost scalar nodes
  int64_t Lv = 0;               // vector length: 3*Np + 3*n_ghost (ghost tail interleaved [g][c])
  // numbering.  internal id u: u < Np owned row, u >= Np ghost (Np + ghost index)
  std::vector<int32_t> perm;        // [Np] internal row -> global node, -1 padding
  std::vector<int32_t> iperm;       // [n_owned] (global node - own0) -> internal row
  std::vector<int32_t> ghost_global;// [n_ghost] global node, sorted by (owner, id)
  std::vector<int32_t> row_len;     // [Np] real entries of the row (0 padding)
  std::vector<int32_t> row_master;  // [Np] internal id of the resolved master of the row's node
                                    //      (== row if free, -1 padding); may be a ghost id
  // SELL-32-sigma scalar pattern shared by the coupled blocks
  int64_t n_slices = 0;
  std::vector<int64_t> slice_ptr;   // [n_slices+1] offsets (in entries) into col/val arrays
  std::vector<int32_t> col;         // [n_ent] internal column id; ~row on the single diagonal
                                    //         entry of a constrained (slave) row; padding = row, 0.0
  int64_t n_ent = 0;
  // cells, sorted by colour
  int n_colors = 0;
  std::vector<int64_t> color_ptr;   // [n_colors+1]
  std::vector<int32_t> cell_src;    // [n_cells] index of the cell in the caller's arrays
  std::vector<int32_t> cell_gather; // [n_cells][n_s] internal id of the raw node (state gather)
  std::vector<int32_t> cell_row;    // [n_cells][n_s] internal row of the resolved node, -1 if not owned
  std::vector<int32_t> cell_pos;    // [n_cells][n_s][n_s] entry position of (m(a), m(b)), -1 if row not owned
  std::vector<int32_t> cell_col;    // [n_cells][n_s] internal id of the resolved node (matrix-free column gather)
  std::vector<int32_t> cell_spos;   // [n_cells][n_s] entry position of the slave diagonal, -1 if not slave/owned
  std::vector<double>  cell_h;      // [n_cells][2] hx, hy
  // reference tables
  std::vector<double> shape, dshape, qw;
  // periodic slaves among local nodes: v[slave] = v[master]
  std::vector<int32_t> slave_id, slave_master; // internal ids (slave is owned or ghost)
  // halo (n_ranks > 1)
  std::vector<int32_t> peers;           // ranks exchanged with
  std::vector<int64_t> send_ptr;        // [n_peers+1] into send_rows
  std::vector<int32_t> send_rows;       // internal owned rows to pack for each peer
  std::vector<int64_t> recv_ptr;        // [n_peers+1] ghost index ranges (ghosts sorted by owner)
  // monolithic view (single rank only): pattern the reference would build (mpi.cpp:232-244)
  int64_t nnz_mono = 0;
  bool has_tables = true;               // SELL pattern + per-cell tables were built
  Lattice lat;
  Ring ring;
  std::string error;
};

// builds the plan; returns PFC_OK or an error code with plan.error set
int build_plan(const pfc_mesh &mesh, int degree, Plan &plan, int flags = 0);

// pfc_lattice_plan.cpp: fills plan.lat (lut_global, order, frame sizes); false + lat.why when the mesh is not a
// uniform periodic tensor lattice with a rectangular owned region
bool detect_lattice(const pfc_mesh &mesh, Plan &plan, const std::vector<int32_t> &cnode,
                    const std::vector<int32_t> &master, const std::vector<double> &h_of);
// pfc_ring_plan.cpp: this rank's card, and the overlay that replaces the plan's one-cell ghost layer by the deep ring
// (ghost ids, halo lists incl. self copies, slave masters, ghost_global for the handed-over frame nodes).
// `cards` holds all ranks' cards in rank order.  Pure host code (CPU tests run it for every rank of a partition).
RingCard ring_card(const Plan &plan);
int ring_overlay(Plan &plan, const std::vector<RingCard> &cards, int depth);
// built-in FE_Q(p) x QGauss(p+1) 1-D tables the lattice kernels are compiled with
double lattice_shape_1d(int p, int q, int j);
double lattice_dshape_1d(int p, int q, int j);
double lattice_weight_1d(int p, int q);

// monolithic CSR (mpi.cpp:220-244) derived from the plan (single rank): row_ptr [n_dofs+1],
// col_idx [nnz], and for every monolithic entry where it lives: block id (0 B,1 C,2 D,3 K,4 M)
// and the SELL entry position
void monolithic_pattern(const Plan &plan, std::vector<int64_t> &row_ptr, std::vector<int32_t> &col_idx,
                        std::vector<int8_t> *block_of, std::vector<int32_t> *pos_of);

// requests the peers' send lists: needs_from[r] = global nodes this rank needs from rank r.
// (the exchange of these lists happens in pfc_comm_init)
} // namespace pfc
