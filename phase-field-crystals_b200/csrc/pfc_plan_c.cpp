// pfc_plan_c.cpp — C view of the host plan (include/pfc_b200.h, "host-only view").
#include <cstdio>
#include <cstring>

#include "pfc_plan.h"

struct pfc_plan
{
  pfc::Plan P;
};

extern "C" int pfc_plan_build(const pfc_mesh *mesh, int32_t degree, pfc_plan **out, char *err, int32_t err_len)
{
  return pfc_plan_build_ex(mesh, degree, 0, out, err, err_len);
}

extern "C" void pfc_plan_get_lattice(const pfc_plan *p, pfc_lattice_view *v)
{
  if (!p || !v)
    return;
  const pfc::Lattice &L = p->P.lat;
  std::memset(v, 0, sizeof(*v));
  v->active = L.active ? 1 : 0;
  v->degree = L.p;
  v->ncx = L.ncx;
  v->ncy = L.ncy;
  v->ex0 = L.ex0;
  v->ey0 = L.ey0;
  v->w = L.w;
  v->h = L.h;
  v->fw = L.fw;
  v->fh = L.fh;
  v->hx = L.hx;
  v->hy = L.hy;
  v->lut = L.lut.data();
  std::snprintf(v->why, sizeof(v->why), "%s", L.why.c_str());
}

extern "C" int pfc_plan_build_ex(const pfc_mesh *mesh, int32_t degree, int32_t flags, pfc_plan **out, char *err,
                                 int32_t err_len)
{
  if (!mesh || !out)
    return PFC_ERR_BAD_ARGUMENT;
  *out = nullptr;
  pfc_plan *p = new pfc_plan();
  const int rc = pfc::build_plan(*mesh, degree, p->P, flags);
  if (rc != PFC_OK)
    {
      if (err && err_len > 0)
        std::snprintf(err, (size_t)err_len, "%s", p->P.error.c_str());
      delete p;
      return rc;
    }
  *out = p;
  return PFC_OK;
}

extern "C" void pfc_plan_get_view(const pfc_plan *p, pfc_plan_view *v)
{
  if (!p || !v)
    return;
  const pfc::Plan &P = p->P;
  std::memset(v, 0, sizeof(*v));
  v->n_s = P.n_s;
  v->n_colors = P.n_colors;
  v->rank = P.rank;
  v->n_ranks = P.n_ranks;
  v->n_cells = P.n_cells;
  v->N_global = P.N_global;
  v->own0 = P.own0;
  v->n_owned = P.n_owned;
  v->Np = P.Np;
  v->n_ghost = P.n_ghost;
  v->Lv = P.Lv;
  v->n_slices = P.n_slices;
  v->n_ent = P.n_ent;
  v->nnz_mono = P.nnz_mono;
  v->n_slaves = (int64_t)P.slave_id.size();
  v->n_peers = (int64_t)P.peers.size();
  v->perm = P.perm.data();
  v->iperm = P.iperm.data();
  v->ghost_global = P.ghost_global.data();
  v->row_len = P.row_len.data();
  v->row_master = P.row_master.data();
  v->col = P.col.data();
  v->slice_ptr = P.slice_ptr.data();
  v->color_ptr = P.color_ptr.data();
  v->cell_src = P.cell_src.data();
  v->cell_gather = P.cell_gather.data();
  v->cell_row = P.cell_row.data();
  v->cell_pos = P.cell_pos.data();
  v->cell_spos = P.cell_spos.data();
  v->cell_h = P.cell_h.data();
  v->slave_id = P.slave_id.data();
  v->slave_master = P.slave_master.data();
  v->peers = P.peers.data();
  v->recv_ptr = P.recv_ptr.data();
}

extern "C" int pfc_plan_monolithic(const pfc_plan *p, int64_t *row_ptr, int32_t *col_idx, int8_t *block_of, int32_t *pos_of)
{
  if (!p || !row_ptr || !col_idx)
    return PFC_ERR_BAD_ARGUMENT;
  if (p->P.n_ranks != 1)
    return PFC_ERR_UNSUPPORTED;
  std::vector<int64_t> rp;
  std::vector<int32_t> ci, pos;
  std::vector<int8_t> blk;
  pfc::monolithic_pattern(p->P, rp, ci, &blk, &pos);
  std::memcpy(row_ptr, rp.data(), rp.size() * sizeof(int64_t));
  std::memcpy(col_idx, ci.data(), ci.size() * sizeof(int32_t));
  if (block_of)
    std::memcpy(block_of, blk.data(), blk.size());
  if (pos_of)
    std::memcpy(pos_of, pos.data(), pos.size() * sizeof(int32_t));
  return PFC_OK;
}

// all ranks of a partition in one process (CPU tests): gather the cards, overlay every plan
extern "C" int pfc_plan_ring_overlay(pfc_plan *const *plans, int32_t n_ranks, int32_t depth, char *err, int32_t err_len)
{
  if (!plans || n_ranks < 1)
    return PFC_ERR_BAD_ARGUMENT;
  std::vector<pfc::RingCard> cards;
  for (int r = 0; r < n_ranks; ++r)
    {
      if (!plans[r] || plans[r]->P.rank != r || plans[r]->P.n_ranks != n_ranks)
        return PFC_ERR_BAD_ARGUMENT;
      cards.push_back(pfc::ring_card(plans[r]->P));
    }
  for (int r = 0; r < n_ranks; ++r)
    {
      const int rc = pfc::ring_overlay(plans[r]->P, cards, depth);
      if (rc != PFC_OK)
        {
          if (err && err_len > 0)
            std::snprintf(err, (size_t)err_len, "rank %d: %s", r, plans[r]->P.error.c_str());
          return rc;
        }
    }
  return PFC_OK;
}

extern "C" void pfc_plan_get_ring(const pfc_plan *p, pfc_ring_view *v)
{
  if (!p || !v)
    return;
  const pfc::Plan &P = p->P;
  const pfc::Ring &R = P.ring;
  std::memset(v, 0, sizeof(*v));
  v->active = R.active ? 1 : 0;
  v->depth = R.D;
  v->ex0 = R.ex0;
  v->ey0 = R.ey0;
  v->fw = R.fw;
  v->fh = R.fh;
  v->ncx = R.ncx;
  v->ncy = R.ncy;
  v->n_ring = R.n_ring;
  v->n_self = (int64_t)R.self_src.size();
  v->n_recv = (int64_t)R.recv_pos.size();
  v->n_send = (int64_t)P.send_rows.size();
  v->n_peers = (int64_t)P.peers.size();
  v->self_src = R.self_src.data();
  v->self_dst = R.self_dst.data();
  v->recv_pos = R.recv_pos.data();
  v->send_rows = P.send_rows.data();
  v->peers = P.peers.data();
  v->recv_ptr = P.recv_ptr.data();
  v->send_ptr = P.send_ptr.data();
}

extern "C" void pfc_plan_destroy(pfc_plan *p) { delete p; }
