// pfc_dealii_adapter.h — flattens what a deal.II host owns into the pfc_mesh of include/pfc_b200.h and drives the handle.
// Meant to be called from PhaseFieldCrystalSystemMPI<dim> (reference: src/phase_field_crystal_systems/
// phase_field_crystal_system_mpi.{hpp,cpp}); the call sites are listed in INTEGRATION.md.  Needs deal.II >= 9.3 with
// Trilinos, p4est and MPI, none of which exist in this repository's image: this file is NOT compiled or tested there.
#pragma once
#include <deal.II/base/index_set.h>
#include <deal.II/base/mpi.h>
#include <deal.II/dofs/dof_handler.h>
#include <deal.II/fe/fe_system.h>
#include <deal.II/lac/affine_constraints.h>
#include <deal.II/lac/generic_linear_algebra.h>

#include <array>
#include <vector>

#include <pfc_b200.h>

namespace pfc_b200
{
namespace LA = dealii::LinearAlgebraTrilinos;

struct Options
{
  double dt, theta, eps, newton_tol;
  unsigned int newton_max_iters;
  int operator_kind  = PFC_OPERATOR_MATRIX_FREE;  // lattice kernels on make_grid's meshes; 0 = assembled, bit-exact pattern
  int preconditioner = PFC_PRECOND_BLOCK_SCHUR;   // PFC_PRECOND_NONE = solve_and_update_no_precondition as upstream;
                                                  // PFC_PRECOND_SCHUR_REDUCED = the serial class's psi-only Schur solve
                                                  // (fastest on FE_Q(1) + matrix-free; elsewhere it acts as BLOCK_SCHUR)
  int gpus_per_node  = 8;
};

// One instance per MPI rank = one GPU.  dim = 2 (the C ABI refuses anything else).
class Adapter
{
public:
  Adapter() = default;
  ~Adapter();
  Adapter(const Adapter &) = delete;

  // end of setup_dofs() (mpi.cpp:274): mesh, DoF maps, tables, constraints, ownership
  void setup(MPI_Comm comm, const dealii::DoFHandler<2> &dof_handler, const dealii::FESystem<2> &fe_system,
             const dealii::AffineConstraints<double> &constraints, const std::vector<dealii::IndexSet> &owned_partitioning,
             const Options &opt);
  // end of initialize_fe_field() (mpi.cpp:296) / before output (mpi.cpp:597): owned ranges only, ghosts travel GPU to GPU
  void set_state(const LA::MPI::BlockVector &Psi_n, const LA::MPI::BlockVector &Psi_n_1);
  void get_state(LA::MPI::BlockVector &Psi_n) const; // caller: Psi_n.update_ghost_values() on its ghosted copy
  // iterate_timestep() (mpi.cpp:576-593): the whole Newton loop; throws what the reference would throw
  pfc_step_stats step();
  // stress_calculator->calculate_stress (mpi.cpp:673-676, 690-693); [4][n_owned]
  std::array<std::int64_t, 4> stress(double eps, std::vector<double> &owned_stress);

  pfc_handle *handle() { return gpu; }

private:
  void check(int rc) const;
  pfc_handle               *gpu = nullptr;
  MPI_Comm                  comm = MPI_COMM_NULL;
  std::int64_t              n_owned = 0;
  std::vector<double>       up, up1; // [3][n_owned] staging
};
} // namespace pfc_b200
