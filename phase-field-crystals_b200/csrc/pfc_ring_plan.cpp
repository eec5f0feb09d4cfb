// pfc_ring_plan.cpp — the deep ghost ring of the lattice path (host side; see Ring in pfc_plan.h).
//
// The reference exchanges ghost values inside every Epetra vmult (mpi.cpp:482) through deal.II's one-cell ghost layer
// (mpi.cpp:178-186).  Here a rank's ghosts are instead ALL lattice positions within `depth` nodes of its owned
// rectangle, owners found by arithmetic on the global periodic lattice (the ranks all-gather their RingCards once):
// nothing beyond the handed-over mesh is needed, because on the uniform lattice position (X, Y) mod period lies in
// exactly one rank's rectangle and its index there is (Y - y_r) w_r + (X - x_r).  Single rank: the ring is the periodic
// wrap-around of the rank's own nodes (self copies).
#include <algorithm>
#include <cmath>
#include <cstdio>

#include "pfc_plan.h"

namespace pfc
{
RingCard ring_card(const Plan &P)
{
  RingCard c;
  const Lattice &L = P.lat;
  c.ox = L.ox;
  c.oy = L.oy;
  c.hx = L.hx;
  c.hy = L.hy;
  c.Px = L.Px;
  c.Py = L.Py;
  c.w = L.w;
  c.h = L.h;
  c.ex0 = L.ex0;
  c.ey0 = L.ey0;
  c.p = L.p;
  c.active = L.active ? 1 : 0;
  return c;
}

namespace
{
struct Geo
{
  int32_t ex0, ey0, w, h, fw, fh;
  int64_t gx, gy; // global lattice position of the owned free node (0, 0)
};

inline int64_t pmod(int64_t a, int64_t m)
{
  const int64_t r = a % m;
  return r < 0 ? r + m : r;
}

// margins >= D with padded (0, 0) and (fw - 1, fh - 1) on cell corners
Geo geometry(const RingCard &c, int D, double ox_min, double oy_min)
{
  Geo g;
  const int p = c.p;
  g.w = c.w;
  g.h = c.h;
  g.ex0 = D + (int32_t)pmod(c.ex0 - D, p);
  g.ey0 = D + (int32_t)pmod(c.ey0 - D, p);
  int dr = D, dt = D;
  while ((g.ex0 + g.w + dr - 1) % p)
    ++dr;
  while ((g.ey0 + g.h + dt - 1) % p)
    ++dt;
  g.fw = g.ex0 + g.w + dr;
  g.fh = g.ey0 + g.h + dt;
  g.gx = std::llround((c.ox - ox_min) / (c.hx / p));
  g.gy = std::llround((c.oy - oy_min) / (c.hy / p));
  return g;
}
} // namespace

int ring_overlay(Plan &P, const std::vector<RingCard> &cards, int depth)
{
  auto fail = [&](const char *why) {
    P.error = std::string("ring overlay: ") + why;
    return PFC_ERR_UNSUPPORTED;
  };
  const int R = P.n_ranks, me = P.rank;
  if ((int)cards.size() != R || !P.lat.active || depth < P.degree)
    return fail("needs an active lattice on this rank, one card per rank and depth >= degree");
  int64_t PX = 0, PY = 0, area = 0;
  double ox_min = cards[0].ox, oy_min = cards[0].oy;
  for (const RingCard &c : cards)
    {
      if (!c.active)
        return fail("a rank has no lattice");
      if (c.p != cards[0].p || std::fabs(c.hx - cards[0].hx) > 1e-10 * c.hx || std::fabs(c.hy - cards[0].hy) > 1e-10 * c.hy)
        return fail("ranks disagree on degree or cell size");
      if ((c.Px && PX && c.Px != PX) || (c.Py && PY && c.Py != PY))
        return fail("ranks disagree on the period");
      PX = std::max(PX, c.Px);
      PY = std::max(PY, c.Py);
      ox_min = std::min(ox_min, c.ox);
      oy_min = std::min(oy_min, c.oy);
      area += (int64_t)c.w * c.h;
    }
  if (PX <= 0 || PY <= 0 || area != PX * PY)
    return fail("the owned rectangles do not tile the periodic lattice");
  std::vector<Geo> G(R);
  for (int r = 0; r < R; ++r)
    G[r] = geometry(cards[r], depth, ox_min, oy_min);
  // owner of a global position: (rank, index in its owned rectangle)
  auto owner_of = [&](int64_t gX, int64_t gY, int &q, int32_t &idx) {
    for (int r = 0; r < R; ++r)
      {
        const int64_t lx = pmod(gX - G[r].gx, PX), ly = pmod(gY - G[r].gy, PY);
        if (lx < G[r].w && ly < G[r].h)
          {
            q = r;
            idx = (int32_t)(ly * G[r].w + lx);
            return true;
          }
      }
    return false;
  };
  // ring of rank r in ring-index order: owner rank and owner index of every position
  auto ring_of = [&](int r, std::vector<int32_t> &own_rank, std::vector<int32_t> &own_idx) {
    const Geo &g = G[r];
    const int64_t n = (int64_t)g.fw * g.fh - (int64_t)g.w * g.h;
    own_rank.assign((size_t)n, -1);
    own_idx.assign((size_t)n, -1);
    bool ok = true;
#pragma omp parallel for schedule(static) reduction(&& : ok)
    for (int32_t Y = 0; Y < g.fh; ++Y)
      for (int32_t X = 0; X < g.fw; ++X)
        {
          if (X >= g.ex0 && X < g.ex0 + g.w && Y >= g.ey0 && Y < g.ey0 + g.h)
            continue;
          int q = -1;
          int32_t idx = -1;
          if (!owner_of(g.gx - g.ex0 + X, g.gy - g.ey0 + Y, q, idx))
            ok = false;
          const int64_t k = ring_index(X, Y, g.ex0, g.ey0, g.w, g.h, g.fw, g.fh);
          own_rank[(size_t)k] = q;
          own_idx[(size_t)k] = idx;
        }
    return ok;
  };

  const Geo &g = G[me];
  Ring &Rg = P.ring;
  Rg = Ring();
  Rg.D = depth;
  Rg.ex0 = g.ex0;
  Rg.ey0 = g.ey0;
  Rg.fw = g.fw;
  Rg.fh = g.fh;
  Rg.ncx = (g.fw - 1) / P.degree;
  Rg.ncy = (g.fh - 1) / P.degree;
  Rg.n_ring = (int64_t)g.fw * g.fh - (int64_t)g.w * g.h;
  if (P.Np + Rg.n_ring + 1 > INT32_MAX / 3)
    return fail("local vector too long for 32-bit indices");

  std::vector<int32_t> orank, oidx;
  if (!ring_of(me, orank, oidx))
    return fail("a ring position has no owner");
  // receive side: per peer the ring positions it fills, ascending; self copies
  std::vector<std::vector<int32_t>> recv_of(R), send_of(R);
  for (int64_t k = 0; k < Rg.n_ring; ++k)
    {
      const int q = orank[(size_t)k];
      if (q == me)
        {
          Rg.self_src.push_back(oidx[(size_t)k]);
          Rg.self_dst.push_back((int32_t)k);
        }
      else
        recv_of[q].push_back((int32_t)k);
    }
  // send side: what every other rank's ring takes from this rank, in that rank's ring order
  for (int q = 0; q < R; ++q)
    {
      if (q == me)
        continue;
      std::vector<int32_t> qr, qi;
      if (!ring_of(q, qr, qi))
        return fail("a ring position of a peer has no owner");
      for (size_t k = 0; k < qr.size(); ++k)
        if (qr[k] == me)
          send_of[q].push_back(qi[k]);
    }
  // the peer-memory protocol paces a pair by its two-way traffic (pfc_comm.cu): make every pair two-way with one
  // dummy value into the scratch slot n_ring where the geometry is one-directional
  for (int q = 0; q < R; ++q)
    {
      if (q == me)
        continue;
      if (recv_of[q].empty() && !send_of[q].empty())
        recv_of[q].push_back((int32_t)Rg.n_ring);
      else if (!recv_of[q].empty() && send_of[q].empty())
        send_of[q].push_back(0);
    }
  P.peers.clear();
  P.recv_ptr.assign(1, 0);
  P.send_ptr.assign(1, 0);
  P.send_rows.clear();
  for (int q = 0; q < R; ++q)
    if (q != me && !recv_of[q].empty())
      {
        P.peers.push_back(q);
        Rg.recv_pos.insert(Rg.recv_pos.end(), recv_of[q].begin(), recv_of[q].end());
        P.send_rows.insert(P.send_rows.end(), send_of[q].begin(), send_of[q].end());
        P.recv_ptr.push_back((int64_t)Rg.recv_pos.size());
        P.send_ptr.push_back((int64_t)P.send_rows.size());
      }

  // ghost ids := ring positions; global ids are known for the positions of the handed-over one-cell frame only
  const Lattice &L = P.lat;
  std::vector<int32_t> gg((size_t)Rg.n_ring + 1, -1);
  for (int32_t fy = 0; fy < L.fh; ++fy)
    for (int32_t fx = 0; fx < L.fw; ++fx)
      {
        const int32_t lx = fx - L.ex0, ly = fy - L.ey0;
        if (lx >= 0 && lx < L.w && ly >= 0 && ly < L.h)
          continue;
        const int32_t id = L.lut[(size_t)fy * L.fw + fx];
        if (id >= P.Np) // periodic images of the rank's own nodes stay anonymous copies: their global id maps to the owned row
          gg[(size_t)ring_index(lx + g.ex0, ly + g.ey0, g.ex0, g.ey0, g.w, g.h, g.fw, g.fh)] = P.ghost_global[id - P.Np];
      }
  P.ghost_global.swap(gg);
  P.n_ghost = Rg.n_ring + 1;
  P.Lv = 3 * P.Np + 3 * P.n_ghost;
  // constraints.distribute (mpi.cpp:485) for the owned slaves: the master's value sits at the slave's own position
  P.slave_id.clear();
  P.slave_master.clear();
  for (size_t k = 0; k < L.slave_fx.size(); ++k)
    {
      const int32_t X = L.slave_fx[k] - L.ex0 + g.ex0, Y = L.slave_fy[k] - L.ey0 + g.ey0;
      const int32_t id = (int32_t)((int64_t)L.w * L.h + (int64_t)k);
      const int32_t m = (int32_t)(P.Np + ring_index(X, Y, g.ex0, g.ey0, g.w, g.h, g.fw, g.fh));
      P.slave_id.push_back(id);
      P.slave_master.push_back(m);
      P.row_master[id] = m;
    }
  Rg.active = true;
  return PFC_OK;
}
} // namespace pfc
