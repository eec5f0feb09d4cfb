// pfc_comm.h — multi-GPU plumbing: NCCL (loaded lazily with dlopen, so a single-GPU host needs
// no NCCL at all) for the two exchanges the path has:
//   halo of ghost entries before each operator application  (Epetra Import inside vmult, mpi.cpp:482)
//   all-reduce of the Gram-Schmidt coefficients and norms    (MPI_Allreduce inside dots, mpi.cpp:482,584)
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

struct pfc_handle;

namespace pfc
{
struct Ctl;
// ctl (optional): the peer-memory kernels skip themselves once the solve has ended (Ctl.state != 0 is identical on all
// ranks: it derives from all-reduced values); NCCL operations always run
void comm_destroy(pfc_handle *h);
int comm_exchange_send_lists(pfc_handle *h);           // after build_plan: fills plan.send_ptr / send_rows
struct RingCard;
int comm_gather_cards(pfc_handle *h, const RingCard &mine, std::vector<RingCard> &all); // lattice path: collective
int comm_ring_upload(pfc_handle *h);                   // after ring_overlay: device copies of the ring lists
int halo_exchange(pfc_handle *h, double *v, const Ctl *ctl = nullptr);           // v's ghost tail := owners' values (stream-ordered)
int halo_exchange_scalar(pfc_handle *h, double *v, const Ctl *ctl = nullptr);    // scalar field [owned Np | ghosts]
int allreduce_sum(pfc_handle *h, double *dev, size_t n, const Ctl *ctl = nullptr); // in place, stream-ordered; no-op on one rank
} // namespace pfc

// ---- peer-to-peer path (one node, NVLink): see pfc_comm.cu ------------------------------------------------
namespace pfc
{
struct Ctl;
int comm_p2p_setup(pfc_handle *h);     // after the arena exists; collective; h->p2p = false -> NCCL stays in charge
void comm_p2p_destroy(pfc_handle *h);
struct ArDev;
ArDev comm_ar_endpoint(const pfc_handle *h); // R == 0 unless the peer-memory path is active (pfc_p2p.cuh)
void comm_p2p_unknown_predecessor(pfc_handle *h); // before a graph capture / launch: force a flag-only round first
} // namespace pfc
