// pfc_dealii_adapter.cc — see pfc_dealii_adapter.h.  NOT compiled in this repository's image (no deal.II there).
#include "pfc_dealii_adapter.h"

#include <deal.II/base/quadrature_lib.h>
#include <deal.II/dofs/dof_tools.h>
#include <deal.II/fe/fe_q.h>
#include <deal.II/lac/solver_control.h>

#include <stdexcept>

namespace pfc_b200
{
using namespace dealii;

Adapter::~Adapter()
{
  if (gpu)
    pfc_destroy(gpu);
}

void Adapter::check(int rc) const
{
  if (rc == PFC_OK || rc == PFC_WARN_NEWTON_MAX_ITERS) // the reference leaves its Newton loop silently (mpi.cpp:579-592)
    return;
  if (rc == PFC_ERR_GMRES_NOT_CONVERGED) // what SolverGMRES throws at mpi.cpp:482
    throw SolverControl::NoConvergence(0, 0.0);
  throw std::runtime_error(pfc_last_error(gpu));
}

void Adapter::setup(MPI_Comm comm_, const DoFHandler<2> &dof_handler, const FESystem<2> &fe_system,
                    const AffineConstraints<double> &constraints, const std::vector<IndexSet> &owned_partitioning,
                    const Options &opt)
{
  comm                     = comm_;
  const unsigned int rank  = Utilities::MPI::this_mpi_process(comm);
  const unsigned int ranks = Utilities::MPI::n_mpi_processes(comm);

  pfc_params prm;
  pfc_default_params(&prm);
  prm.degree           = fe_system.degree;
  prm.dt               = opt.dt;
  prm.theta            = opt.theta;
  prm.eps              = opt.eps;
  prm.newton_tol       = opt.newton_tol;
  prm.newton_max_iters = opt.newton_max_iters;
  prm.operator_kind    = opt.operator_kind;
  prm.preconditioner   = opt.preconditioner;
  prm.device           = rank % opt.gpus_per_node;
  if (pfc_create(&prm, &gpu) != PFC_OK)
    throw std::runtime_error(pfc_last_error(nullptr));
  if (ranks > 1)
    { // NCCL communicator over the MPI communicator of mpi.cpp:79
      char id[128];
      if (rank == 0)
        check(pfc_nccl_unique_id(id));
      MPI_Bcast(id, 128, MPI_BYTE, 0, comm);
      check(pfc_comm_init(gpu, id, rank, ranks));
    }

  // ---- tables of the scalar element at the Gauss points of mpi.cpp:308 ----
  const FiniteElement<2> &fe_q = fe_system.base_element(0);
  const QGauss<2>         quad(fe_system.degree + 1);
  const unsigned int      n_s = fe_q.n_dofs_per_cell(), n_q = quad.size(), dpc = fe_system.n_dofs_per_cell();
  std::vector<double>     shape(n_q * n_s), grad(n_q * n_s * 2), weight(n_q);
  for (unsigned int q = 0; q < n_q; ++q)
    {
      weight[q] = quad.weight(q);
      for (unsigned int a = 0; a < n_s; ++a)
        {
          shape[q * n_s + a]          = fe_q.shape_value(a, quad.point(q));
          const Tensor<1, 2> g        = fe_q.shape_grad(a, quad.point(q));
          grad[(q * n_s + a) * 2 + 0] = g[0];
          grad[(q * n_s + a) * 2 + 1] = g[1];
        }
    }
  std::vector<std::int32_t> local_component(dpc), local_scalar(dpc);
  for (unsigned int i = 0; i < dpc; ++i)
    {
      const auto ci      = fe_system.system_to_component_index(i);
      local_component[i] = ci.first;
      local_scalar[i]    = ci.second;
    }

  // ---- cells: locally owned first, then the ghost layer (incl. periodic neighbours, mpi.cpp:143-144) ----
  std::vector<std::int32_t>            cell_dofs;
  std::vector<double>                  cell_vertices;
  std::vector<types::global_dof_index> dofs(dpc);
  std::int64_t                         n_owned_cells = 0, n_ghost_cells = 0;
  for (int pass = 0; pass < 2; ++pass)
    for (const auto &cell : dof_handler.active_cell_iterators())
      if (pass == 0 ? cell->is_locally_owned() : cell->is_ghost())
        {
          cell->get_dof_indices(dofs);
          for (const auto d : dofs)
            cell_dofs.push_back(static_cast<std::int32_t>(d));
          for (unsigned int v = 0; v < 4; ++v)
            {
              cell_vertices.push_back(cell->vertex(v)[0]);
              cell_vertices.push_back(cell->vertex(v)[1]);
            }
          ++(pass == 0 ? n_owned_cells : n_ghost_cells);
        }

  // ---- blocks, ownership (mpi.cpp:160-176), constraints (mpi.cpp:188-217) ----
  const std::vector<types::global_dof_index> per_block = DoFTools::count_dofs_per_fe_block(dof_handler);
  n_owned                                              = owned_partitioning[0].n_elements();
  const std::vector<std::int64_t> counts               = Utilities::MPI::all_gather(comm, n_owned);
  std::vector<std::int64_t>       node_offsets(ranks + 1, 0);
  for (unsigned int r = 0; r < ranks; ++r)
    node_offsets[r + 1] = node_offsets[r] + counts[r];
  std::vector<std::int32_t> slave, master;
  for (const auto &line : constraints.get_lines())
    {
      if (line.entries.size() != 1 || line.entries[0].second != 1.0)
        throw std::runtime_error("pfc_b200: only identity periodic constraints are supported");
      slave.push_back(static_cast<std::int32_t>(line.index));
      master.push_back(static_cast<std::int32_t>(line.entries[0].first));
    }

  pfc_mesh m{};
  m.rank               = rank;
  m.n_ranks            = ranks;
  m.n_cells            = n_owned_cells + n_ghost_cells;
  m.n_owned_cells      = n_owned_cells;
  m.n_dofs             = dof_handler.n_dofs();
  m.dofs_per_cell      = dpc;
  m.n_q                = n_q;
  m.cell_dofs          = cell_dofs.data();
  m.cell_vertices      = cell_vertices.data();
  m.local_component    = local_component.data();
  m.local_scalar_index = local_scalar.data();
  m.shape_value        = shape.data();
  m.shape_grad_ref     = grad.data();
  m.q_weight           = weight.data();
  std::int64_t start   = 0;
  for (int b = 0; b < 3; ++b)
    {
      m.block_start[b] = start;
      m.block_size[b]  = per_block[b];
      start += per_block[b];
    }
  m.node_offsets      = node_offsets.data();
  m.n_constraints     = static_cast<std::int64_t>(slave.size());
  m.constraint_slave  = slave.data();
  m.constraint_master = master.data();
  m.row_ptr           = nullptr; // or the host pattern of mpi.cpp:232-244: pfc_setup then checks it bit for bit
  m.col_idx           = nullptr;
  check(pfc_setup(gpu, &m));
  up.resize(3 * n_owned);
  up1.resize(3 * n_owned);
}

void Adapter::set_state(const LA::MPI::BlockVector &Psi_n, const LA::MPI::BlockVector &Psi_n_1)
{
  for (unsigned int b = 0; b < 3; ++b)
    { // the owned range of a block is contiguous (component-wise renumbering, mpi.cpp:157-158)
      const auto range = Psi_n.block(b).local_range();
      for (auto i = range.first; i < range.second; ++i)
        {
          up[b * n_owned + (i - range.first)]  = Psi_n.block(b)[i];
          up1[b * n_owned + (i - range.first)] = Psi_n_1.block(b)[i];
        }
    }
  check(pfc_set_state_owned(gpu, up.data(), up1.data()));
}

void Adapter::get_state(LA::MPI::BlockVector &Psi_n) const
{
  std::vector<double> own(3 * n_owned);
  check(pfc_get_state_owned(gpu, own.data(), nullptr));
  for (unsigned int b = 0; b < 3; ++b)
    {
      const auto range = Psi_n.block(b).local_range();
      for (auto i = range.first; i < range.second; ++i)
        Psi_n.block(b)[i] = own[b * n_owned + (i - range.first)];
    }
  Psi_n.compress(VectorOperation::insert);
}

pfc_step_stats Adapter::step()
{
  pfc_step_stats st;
  check(pfc_step(gpu, &st));
  return st; // the caller prints "Residual is:" (mpi.cpp:585) and "Number of iterations:" (mpi.cpp:483) from st
}

std::array<std::int64_t, 4> Adapter::stress(double eps, std::vector<double> &owned_stress)
{
  std::array<std::int64_t, 4> its{};
  owned_stress.resize(4 * n_owned);
  check(pfc_stress_owned(gpu, eps, 1e-10, owned_stress.data(), its.data()));
  return its;
}
} // namespace pfc_b200
