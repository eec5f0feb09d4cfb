// pfc_plan.cpp — host preprocessing of pfc_setup (see pfc_plan.h).
//
// Reference semantics restated here (never the reference's code; deal.II does this upstream):
//   * coupling table: all `always` except (phi,psi) and (chi,phi)            mpi.cpp:220-229
//   * pattern = {(m(i), m(j)) : i, j on a common cell, coupled} U {(s,s) : s constrained}
//     with keep_constrained_dofs = false                                      mpi.cpp:232-237
//   * constraints are identity periodic lines slave = master                   mpi.cpp:188-217
#include "pfc_plan.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <numeric>

namespace pfc
{
namespace
{
int fail(Plan &p, int code, const std::string &msg)
{
  p.error = msg;
  return code;
}
} // namespace

int build_plan(const pfc_mesh &m, int degree, Plan &P, int flags)
{
  P = Plan();
  if (degree < 1 || degree > 2)
    return fail(P, PFC_ERR_UNSUPPORTED, "degree must be 1 or 2");
  const int ns = (degree + 1) * (degree + 1);
  if (m.dofs_per_cell != 3 * ns || m.n_q != ns)
    return fail(P, PFC_ERR_BAD_ARGUMENT, "dofs_per_cell / n_q do not match FESystem(FE_Q(degree)^3) with QGauss(degree+1)");
  if (!m.cell_vertices || !m.cell_dofs || !m.local_component || !m.local_scalar_index ||
      !m.shape_value || !m.shape_grad_ref || !m.q_weight)
    return fail(P, PFC_ERR_BAD_ARGUMENT, "null mesh array");
  if (m.n_cells <= 0 || m.n_cells > INT32_MAX / (ns * ns))
    return fail(P, PFC_ERR_UNSUPPORTED, "n_cells out of range for 32-bit entry positions");
  const int64_t N = m.block_size[0];
  if (m.block_size[1] != N || m.block_size[2] != N || m.n_dofs != 3 * N || N <= 0 || 3 * N > INT32_MAX)
    return fail(P, PFC_ERR_UNSUPPORTED, "the three blocks must have equal size (component_wise numbering of identical FE_Q)");
  P.degree = degree;
  P.n_s = ns;
  P.n_q = ns;
  P.dpc = 3 * ns;
  P.n_cells = m.n_cells;
  P.N_global = N;
  P.n_dofs = m.n_dofs;
  for (int c = 0; c < 3; ++c)
    P.block_start[c] = m.block_start[c];
  P.rank = m.n_ranks > 1 ? m.rank : 0;
  P.n_ranks = m.n_ranks > 1 ? m.n_ranks : 1;
  if (P.n_ranks > 1)
    {
      if (!m.node_offsets)
        return fail(P, PFC_ERR_BAD_ARGUMENT, "node_offsets required for n_ranks > 1");
      P.node_offsets.assign(m.node_offsets, m.node_offsets + P.n_ranks + 1);
      if (P.node_offsets.front() != 0 || P.node_offsets.back() != N)
        return fail(P, PFC_ERR_BAD_ARGUMENT, "node_offsets must span [0, N]");
    }
  else
    P.node_offsets = {0, N};
  P.own0 = P.node_offsets[P.rank];
  P.n_owned = P.node_offsets[P.rank + 1] - P.own0;
  if (P.n_owned <= 0)
    return fail(P, PFC_ERR_BAD_ARGUMENT, "rank owns no dofs");
  const int dpc = P.dpc;

  // ---- local index tables: (component, scalar shape) -> local dof ---------------------------
  std::vector<int> loc_of(3 * ns, -1);
  for (int i = 0; i < dpc; ++i)
    {
      const int c = m.local_component[i], a = m.local_scalar_index[i];
      if (c < 0 || c > 2 || a < 0 || a >= ns || loc_of[c * ns + a] != -1)
        return fail(P, PFC_ERR_BAD_ARGUMENT, "local_component / local_scalar_index is not a bijection");
      loc_of[c * ns + a] = i;
    }
  P.shape.assign(m.shape_value, m.shape_value + ns * ns);
  P.dshape.assign(m.shape_grad_ref, m.shape_grad_ref + ns * ns * 2);
  P.qw.assign(m.q_weight, m.q_weight + ns);

  // ---- scalar nodes of every cell; block congruence ------------------------------------------
  std::vector<int32_t> cnode((size_t)m.n_cells * ns);
  {
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(+ : bad)
    for (int64_t cell = 0; cell < m.n_cells; ++cell)
      for (int a = 0; a < ns; ++a)
        {
          const int64_t t = (int64_t)m.cell_dofs[cell * dpc + loc_of[a]] - m.block_start[0];
          if (t < 0 || t >= N)
            bad++;
          for (int c = 1; c < 3; ++c)
            if ((int64_t)m.cell_dofs[cell * dpc + loc_of[c * ns + a]] - m.block_start[c] != t)
              bad++;
          cnode[cell * ns + a] = (int32_t)t;
        }
    if (bad)
      return fail(P, PFC_ERR_UNSUPPORTED, "blocks are not index-congruent (dof(c, node) != block_start[c] + node)");
  }

  // ---- geometry: axis-aligned rectangles only (hyper_rectangle, mpi.cpp:119-122) --------------
  std::vector<double> h_of((size_t)m.n_cells * 2);
  {
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(+ : bad)
    for (int64_t cell = 0; cell < m.n_cells; ++cell)
      {
        const double *v = m.cell_vertices + cell * 8;
        const double hx = v[2] - v[0], hy = v[5] - v[1];
        const double tol = 1e-12 * (std::fabs(hx) + std::fabs(hy));
        if (!(hx > 0) || !(hy > 0) || std::fabs(v[3] - v[1]) > tol || std::fabs(v[4] - v[0]) > tol ||
            std::fabs(v[6] - v[2]) > tol || std::fabs(v[7] - v[5]) > tol)
          bad++;
        h_of[cell * 2] = hx;
        h_of[cell * 2 + 1] = hy;
      }
    if (bad)
      return fail(P, PFC_ERR_UNSUPPORTED, "cells must be axis-aligned rectangles with lexicographic vertices");
  }

  // ---- constraints -> resolved master node (global), identical for the three components -------
  std::vector<int32_t> master(N);
  std::iota(master.begin(), master.end(), 0);
  {
    if (m.n_constraints % 3 != 0 || (m.n_constraints > 0 && (!m.constraint_slave || !m.constraint_master)))
      return fail(P, PFC_ERR_UNSUPPORTED, "constraints must come in component triples");
    std::vector<int8_t> seen(N, 0);
    for (int64_t k = 0; k < m.n_constraints; ++k)
      {
        const int64_t s = m.constraint_slave[k], ms = m.constraint_master[k];
        int c = -1;
        for (int b = 0; b < 3; ++b)
          if (s >= m.block_start[b] && s < m.block_start[b] + N)
            c = b;
        if (c < 0 || ms < m.block_start[c] || ms >= m.block_start[c] + N)
          return fail(P, PFC_ERR_UNSUPPORTED, "constraint couples different components");
        const int32_t ts = (int32_t)(s - m.block_start[c]), tm = (int32_t)(ms - m.block_start[c]);
        if (ts == tm)
          return fail(P, PFC_ERR_BAD_ARGUMENT, "dof constrained to itself");
        if (seen[ts] & (1 << c))
          return fail(P, PFC_ERR_UNSUPPORTED, "dof constrained twice (only identity periodic lines are supported)");
        if (seen[ts] && master[ts] != tm)
          return fail(P, PFC_ERR_UNSUPPORTED, "constraints differ between components");
        seen[ts] |= (int8_t)(1 << c);
        master[ts] = tm;
      }
    for (int64_t t = 0; t < N; ++t)
      if (seen[t] != 0 && seen[t] != 7)
        return fail(P, PFC_ERR_UNSUPPORTED, "constraints differ between components");
    // close chains
    for (int64_t t = 0; t < N; ++t)
      {
        int32_t r = master[t];
        int guard = 0;
        while (master[r] != r && guard++ < 64)
          r = master[r];
        if (master[r] != r)
          return fail(P, PFC_ERR_BAD_ARGUMENT, "constraint cycle");
        master[t] = r;
      }
  }

  // ---- local nodes: owned rows + ghosts --------------------------------------------------------
  const int64_t own0 = P.own0, own1 = own0 + P.n_owned;
  auto owned = [&](int64_t t) { return t >= own0 && t < own1; };
  auto owner_of = [&](int64_t t) {
    return (int)(std::upper_bound(P.node_offsets.begin(), P.node_offsets.end(), t) - P.node_offsets.begin()) - 1;
  };
  std::vector<int32_t> g2l(N, -1); // global node -> internal id (filled below)
  {
    std::vector<int8_t> touched(N, 0);
    for (int64_t k = 0; k < m.n_cells * ns; ++k)
      {
        touched[cnode[k]] = 1;
        touched[master[cnode[k]]] = 1;
      }
    for (int64_t t = own0; t < own1; ++t)
      if (master[t] != t)
        touched[master[t]] = 1; // owned slave -> its master must be addressable (distribute)
    for (int64_t t = 0; t < N; ++t)
      if (touched[t] && !owned(t))
        P.ghost_global.push_back((int32_t)t);
    // global ids ascend and node_offsets are monotone, so this is already sorted by (owner, id)
    P.n_ghost = (int64_t)P.ghost_global.size();
    if (P.n_ranks == 1 && P.n_ghost != 0)
      return fail(P, PFC_ERR_BAD_ARGUMENT, "internal: ghosts on a single rank");
  }

  const int64_t n_owned = P.n_owned;
  if (flags & PLAN_LATTICE)
    detect_lattice(m, P, cnode, master, h_of);
  const bool skip_tables = P.lat.active && (flags & PLAN_SKIP_TABLES);
  P.has_tables = !skip_tables;

  // ---- rows of the scalar pattern (owned nodes), columns as ascending GLOBAL node ids ----------
  // adjacency: resolved owned node -> cells
  const int64_t n_cells_tab = skip_tables ? 0 : m.n_cells, n_rows_tab = skip_tables ? 0 : n_owned;
  std::vector<int64_t> adj_ptr(n_owned + 1, 0);
  for (int64_t k = 0; k < n_cells_tab * ns; ++k)
    {
      const int32_t r = master[cnode[k]];
      if (owned(r))
        adj_ptr[r - own0 + 1]++;
    }
  for (int64_t i = 0; i < n_rows_tab; ++i)
    adj_ptr[i + 1] += adj_ptr[i];
  std::vector<int32_t> adj(adj_ptr[n_owned]);
  {
    std::vector<int64_t> fill(adj_ptr.begin(), adj_ptr.end() - 1);
    for (int64_t cell = 0; cell < n_cells_tab; ++cell)
      for (int a = 0; a < ns; ++a)
        {
          const int32_t r = master[cnode[cell * ns + a]];
          if (owned(r))
            adj[fill[r - own0]++] = (int32_t)cell;
        }
  }
  std::vector<int32_t> row_len(n_owned);
  std::vector<int64_t> grow_ptr(n_owned + 1, 0);
  std::vector<int32_t> gcol; // global column ids per row, ascending
  {
    const int maxrow = 4 * ns * 4; // generous: a node is shared by <= 4 cells (+ periodic images)
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n_rows_tab; ++i)
      {
        const int64_t t = own0 + i;
        if (master[t] != t)
          {
            row_len[i] = 1;
            continue;
          }
        std::vector<int32_t> tmp;
        tmp.reserve(maxrow);
        for (int64_t k = adj_ptr[i]; k < adj_ptr[i + 1]; ++k)
          for (int a = 0; a < ns; ++a)
            tmp.push_back(master[cnode[(int64_t)adj[k] * ns + a]]);
        std::sort(tmp.begin(), tmp.end());
        row_len[i] = (int32_t)(std::unique(tmp.begin(), tmp.end()) - tmp.begin());
      }
    for (int64_t i = 0; i < n_rows_tab; ++i)
      grow_ptr[i + 1] = grow_ptr[i] + row_len[i];
    gcol.resize(grow_ptr[n_owned]);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n_rows_tab; ++i)
      {
        const int64_t t = own0 + i;
        if (master[t] != t)
          {
            gcol[grow_ptr[i]] = (int32_t)t;
            continue;
          }
        std::vector<int32_t> tmp;
        for (int64_t k = adj_ptr[i]; k < adj_ptr[i + 1]; ++k)
          for (int a = 0; a < ns; ++a)
            tmp.push_back(master[cnode[(int64_t)adj[k] * ns + a]]);
        std::sort(tmp.begin(), tmp.end());
        tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
        std::copy(tmp.begin(), tmp.end(), gcol.begin() + grow_ptr[i]);
      }
  }
  for (int64_t i = 0; i < n_rows_tab; ++i)
    if (master[own0 + i] == own0 + i && row_len[i] == 0)
      return fail(P, PFC_ERR_BAD_ARGUMENT, "an owned free dof is not touched by any cell handed over");

  // ---- internal numbering: stable sort by descending row length inside windows of kSigma rows --
  P.Np = (n_owned + kSlice - 1) / kSlice * kSlice;
  P.perm.assign(P.Np, -1);
  P.iperm.assign(n_owned, -1);
  {
    std::vector<int32_t> idx(n_owned);
    std::iota(idx.begin(), idx.end(), 0);
    if (P.lat.active)
      { // lattice numbering: owned free nodes lexicographic (y, x), then the owned slaves by global id
        int64_t r = 0;
        for (int32_t g : P.lat.order)
          idx[r++] = (int32_t)(g - own0);
        for (int64_t t = own0; t < own1; ++t)
          if (master[t] != t)
            idx[r++] = (int32_t)(t - own0);
        if (r != n_owned)
          return fail(P, PFC_ERR_BAD_ARGUMENT, "internal: lattice numbering does not cover the owned nodes");
      }
    else
      for (int64_t w0 = 0; w0 < n_owned; w0 += kSigma)
        {
          const int64_t w1 = std::min<int64_t>(n_owned, w0 + kSigma);
          std::stable_sort(idx.begin() + w0, idx.begin() + w1,
                           [&](int32_t a, int32_t b) { return row_len[a] > row_len[b]; });
        }
    for (int64_t r = 0; r < n_owned; ++r)
      {
        P.perm[r] = (int32_t)(own0 + idx[r]);
        P.iperm[idx[r]] = (int32_t)r;
      }
  }
  for (int64_t t = own0; t < own1; ++t)
    g2l[t] = P.iperm[t - own0];
  for (int64_t g = 0; g < P.n_ghost; ++g)
    g2l[P.ghost_global[g]] = (int32_t)(P.Np + g);
  P.Lv = 3 * P.Np + 3 * P.n_ghost;
  if (P.Lv > INT32_MAX)
    return fail(P, PFC_ERR_UNSUPPORTED, "local vector too long for 32-bit indices");

  P.row_master.assign(P.Np, -1);
  P.row_len.assign(P.Np, 0);
  for (int64_t r = 0; r < n_owned; ++r)
    {
      P.row_master[r] = g2l[master[P.perm[r]]];
      P.row_len[r] = row_len[P.perm[r] - own0];
    }

  // ---- lattice: frame position -> internal id ----------------------------------------------------
  if (P.lat.active)
    {
      Lattice &L = P.lat;
      L.lut.resize(L.lut_global.size());
      for (size_t k = 0; k < L.lut_global.size(); ++k)
        {
          const int32_t g = L.lut_global[k];
          if (g < 0 || g2l[g] < 0)
            return fail(P, PFC_ERR_BAD_ARGUMENT, "internal: lattice frame position without a local node");
          L.lut[k] = g2l[g];
        }
      for (int64_t ly = 0; ly < L.h; ++ly)
        for (int64_t lx = 0; lx < L.w; ++lx)
          if (L.lut[(size_t)((L.ey0 + ly) * L.fw + L.ex0 + lx)] != (int32_t)(ly * L.w + lx))
            return fail(P, PFC_ERR_BAD_ARGUMENT, "internal: lattice numbering mismatch");
      std::vector<int32_t>().swap(L.lut_global);
    }

  P.n_slices = P.Np / kSlice;
  P.slice_ptr.assign(P.n_slices + 1, 0);
  if (!skip_tables)
    {
  // ---- SELL-32 arrays ---------------------------------------------------------------------------
  P.n_slices = P.Np / kSlice;
  P.slice_ptr.assign(P.n_slices + 1, 0);
  for (int64_t s = 0; s < P.n_slices; ++s)
    {
      int32_t w = 0;
      for (int l = 0; l < kSlice; ++l)
        {
          const int32_t t = P.perm[s * kSlice + l];
          if (t >= 0)
            w = std::max(w, row_len[t - own0]);
        }
      P.slice_ptr[s + 1] = P.slice_ptr[s] + (int64_t)w * kSlice;
    }
  P.n_ent = P.slice_ptr[P.n_slices];
  if (P.n_ent > INT32_MAX)
    return fail(P, PFC_ERR_UNSUPPORTED, "scalar pattern too large for 32-bit entry positions");
  P.col.resize(P.n_ent);
#pragma omp parallel for schedule(static)
  for (int64_t s = 0; s < P.n_slices; ++s)
    {
      const int64_t w = (P.slice_ptr[s + 1] - P.slice_ptr[s]) / kSlice;
      for (int l = 0; l < kSlice; ++l)
        {
          const int64_t r = s * kSlice + l;
          const int32_t t = P.perm[r];
          const int64_t i = t >= 0 ? t - own0 : -1;
          const int32_t len = t >= 0 ? row_len[i] : 0;
          const bool slave = t >= 0 && master[t] != t;
          for (int64_t k = 0; k < w; ++k)
            {
              int32_t c;
              if (k < len)
                c = slave ? ~(int32_t)r : g2l[gcol[grow_ptr[i] + k]];
              else
                c = (int32_t)r; // padding: value 0, harmless column
              P.col[P.slice_ptr[s] + k * kSlice + l] = c;
            }
        }
    }

  // entry position of (row r, global column g): binary search in the row's ascending global list
  auto entry_pos = [&](int64_t r, int32_t gc) -> int32_t {
    const int64_t i = P.perm[r] - own0;
    const int32_t *b = gcol.data() + grow_ptr[i], *e = b + row_len[i];
    const int32_t *f = std::lower_bound(b, e, gc);
    if (f == e || *f != gc)
      return -1;
    return (int32_t)(P.slice_ptr[r / kSlice] + (int64_t)(f - b) * kSlice + (r % kSlice));
  };

  // ---- colouring of the cells: no two cells of one colour touch the same raw or resolved node --
  std::vector<int32_t> color(m.n_cells, -1);
  {
    std::vector<uint32_t> used(N, 0); // colours already present at a node (bitmask, <= 32 colours)
    int ncol = 0;
    for (int64_t cell = 0; cell < m.n_cells; ++cell)
      {
        uint32_t forbid = 0;
        for (int a = 0; a < ns; ++a)
          {
            const int32_t t = cnode[cell * ns + a];
            forbid |= used[t] | used[master[t]];
          }
        int c = 0;
        while (c < 32 && (forbid >> c) & 1u)
          ++c;
        if (c >= 32)
          return fail(P, PFC_ERR_UNSUPPORTED, "colouring needs more than 32 colours");
        color[cell] = c;
        ncol = std::max(ncol, c + 1);
        for (int a = 0; a < ns; ++a)
          {
            const int32_t t = cnode[cell * ns + a];
            used[t] |= 1u << c;
            used[master[t]] |= 1u << c;
          }
      }
    P.n_colors = ncol;
  }
  P.color_ptr.assign(P.n_colors + 1, 0);
  for (int64_t cell = 0; cell < m.n_cells; ++cell)
    P.color_ptr[color[cell] + 1]++;
  for (int c = 0; c < P.n_colors; ++c)
    P.color_ptr[c + 1] += P.color_ptr[c];
  P.cell_src.resize(m.n_cells);
  {
    std::vector<int64_t> fill(P.color_ptr.begin(), P.color_ptr.end() - 1);
    for (int64_t cell = 0; cell < m.n_cells; ++cell)
      P.cell_src[fill[color[cell]]++] = (int32_t)cell;
  }

  // ---- per-cell tables in colour order ---------------------------------------------------------
  P.cell_gather.resize((size_t)m.n_cells * ns);
  P.cell_row.resize((size_t)m.n_cells * ns);
  P.cell_spos.resize((size_t)m.n_cells * ns);
  P.cell_col.resize((size_t)m.n_cells * ns);
  P.cell_pos.resize((size_t)m.n_cells * ns * ns);
  P.cell_h.resize((size_t)m.n_cells * 2);
  int bad_pos = 0;
#pragma omp parallel for schedule(static) reduction(+ : bad_pos)
  for (int64_t k = 0; k < m.n_cells; ++k)
    {
      const int64_t cell = P.cell_src[k];
      P.cell_h[k * 2] = h_of[cell * 2];
      P.cell_h[k * 2 + 1] = h_of[cell * 2 + 1];
      for (int a = 0; a < ns; ++a)
        for (int b = a + 1; b < ns; ++b)
          if (master[cnode[cell * ns + a]] == master[cnode[cell * ns + b]])
            bad_pos++; // two local nodes of one cell identified by periodicity (1-cell-wide mesh)
      for (int a = 0; a < ns; ++a)
        {
          const int32_t t = cnode[cell * ns + a], mt = master[t];
          P.cell_gather[k * ns + a] = g2l[t];
          if (g2l[t] < 0)
            bad_pos++;
          const bool row_owned = owned(mt);
          const int32_t r = row_owned ? g2l[mt] : -1;
          P.cell_row[k * ns + a] = r;
          P.cell_col[k * ns + a] = g2l[mt];
          if (g2l[mt] < 0)
            bad_pos++;
          P.cell_spos[k * ns + a] = (mt != t && owned(t)) ? entry_pos(g2l[t], t) : -1;
          if (mt != t && owned(t) && P.cell_spos[k * ns + a] < 0)
            bad_pos++;
          for (int b = 0; b < ns; ++b)
            {
              int32_t pos = -1;
              if (row_owned)
                {
                  pos = entry_pos(r, master[cnode[cell * ns + b]]);
                  if (pos < 0)
                    bad_pos++;
                }
              P.cell_pos[(k * ns + a) * ns + b] = pos;
            }
        }
    }
  if (bad_pos)
    return fail(P, PFC_ERR_BAD_ARGUMENT, "local entry missing from the derived pattern, or a cell whose own nodes are identified by periodicity");

    }
  // ---- periodic slaves among the local nodes (constraints.distribute, mpi.cpp:485) -------------
  for (int64_t t = 0; t < N; ++t)
    if (master[t] != t && g2l[t] >= 0)
      {
        if (g2l[master[t]] < 0)
          return fail(P, PFC_ERR_BAD_ARGUMENT, "internal: master of a local slave is not local");
        P.slave_id.push_back(g2l[t]);
        P.slave_master.push_back(g2l[master[t]]);
      }

  // ---- halo receive ranges (ghosts are sorted by owner) -----------------------------------------
  if (P.n_ranks > 1)
    {
      P.recv_ptr.push_back(0);
      for (int64_t g = 0; g < P.n_ghost;)
        {
          const int o = owner_of(P.ghost_global[g]);
          int64_t e = g;
          while (e < P.n_ghost && owner_of(P.ghost_global[e]) == o)
            ++e;
          P.peers.push_back(o);
          P.recv_ptr.push_back(e);
          g = e;
        }
    }

  // ---- monolithic nnz (single rank) ------------------------------------------------------------
  if (P.n_ranks == 1 && !skip_tables)
    {
      int64_t nnz = 0;
      for (int64_t i = 0; i < n_owned; ++i)
        nnz += (master[own0 + i] != own0 + i) ? 3 : 7 * (int64_t)row_len[i];
      P.nnz_mono = nnz;
    }
  return PFC_OK;
}

void monolithic_pattern(const Plan &P, std::vector<int64_t> &row_ptr, std::vector<int32_t> &col_idx,
                        std::vector<int8_t> *block_of, std::vector<int32_t> *pos_of)
{
  // block ids: 0 B(0,0) 1 C(0,1) 2 D(0,2) 3 K(1,0),(2,1) 4 M(1,1),(2,2); slave diagonals: comp 0 -> B, 1,2 -> M
  static const int coupled[3][3] = {{1, 1, 1}, {1, 1, 0}, {0, 1, 1}};
  static const int8_t blk[3][3] = {{0, 1, 2}, {3, 4, -1}, {-1, 3, 4}};
  const int64_t N = P.N_global;
  row_ptr.assign(3 * N + 1, 0);
  // entries of the internal row r in ascending GLOBAL column order: recover by sorting (small rows)
  struct Ent
  {
    int32_t gcol, pos;
  };
  auto row_entries = [&](int64_t r, std::vector<Ent> &out, bool &slave) {
    out.clear();
    slave = false;
    const int64_t s = r / kSlice, l = r % kSlice;
    const int64_t w = (P.slice_ptr[s + 1] - P.slice_ptr[s]) / kSlice;
    (void)w;
    for (int64_t k = 0; k < P.row_len[r]; ++k)
      {
        const int64_t pos = P.slice_ptr[s] + k * kSlice + l;
        const int32_t c = P.col[pos];
        if (c < 0)
          {
            slave = true;
            out.push_back({P.perm[r], (int32_t)pos});
            return;
          }
        const int32_t g = c < P.Np ? P.perm[c] : P.ghost_global[c - P.Np];
        out.push_back({g, (int32_t)pos});
      }
    std::sort(out.begin(), out.end(), [](const Ent &a, const Ent &b) { return a.gcol < b.gcol; });
  };
  std::vector<Ent> ent;
  std::vector<int32_t> len(N, 0);
  std::vector<int8_t> is_slave(N, 0);
  for (int64_t t = 0; t < N; ++t)
    {
      bool slave;
      row_entries(P.iperm[t], ent, slave);
      len[t] = (int32_t)ent.size();
      is_slave[t] = slave;
    }
  for (int c = 0; c < 3; ++c)
    for (int64_t t = 0; t < N; ++t)
      {
        const int nb = coupled[c][0] + coupled[c][1] + coupled[c][2];
        row_ptr[c * N + t + 1] = row_ptr[c * N + t] + (is_slave[t] ? 1 : (int64_t)nb * len[t]);
      }
  const int64_t nnz = row_ptr[3 * N];
  col_idx.resize(nnz);
  if (block_of)
    block_of->resize(nnz);
  if (pos_of)
    pos_of->resize(nnz);
  for (int64_t t = 0; t < N; ++t)
    {
      bool slave;
      row_entries(P.iperm[t], ent, slave);
      for (int c = 0; c < 3; ++c)
        {
          int64_t o = row_ptr[c * N + t];
          if (slave)
            {
              col_idx[o] = (int32_t)(c * N + t);
              if (block_of)
                (*block_of)[o] = c == 0 ? 0 : 4;
              if (pos_of)
                (*pos_of)[o] = ent[0].pos;
              continue;
            }
          for (int d = 0; d < 3; ++d)
            if (coupled[c][d])
              for (const Ent &e : ent)
                {
                  col_idx[o] = (int32_t)(d * N + e.gcol);
                  if (block_of)
                    (*block_of)[o] = blk[c][d];
                  if (pos_of)
                    (*pos_of)[o] = e.pos;
                  ++o;
                }
        }
    }
}
} // namespace pfc
