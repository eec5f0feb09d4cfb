// pfc_comm.cu — NCCL plumbing (see pfc_comm.h).  NCCL is resolved at run time with dlopen so
// that libpfc_b200.so has no link-time dependency on it.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>
#include <mutex>

#include "pfc_comm.h"
#include "pfc_handle.h"

namespace
{
struct NcclApi
{
  void *lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

NcclApi &nccl()
{
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    // RTLD_NOLOAD first: reuse the copy a host process (e.g. torch) already mapped
    void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!lib)
      lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib)
      lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib)
      {
        api.error = std::string("dlopen(libnccl.so.2): ") + dlerror();
        return;
      }
    api.lib = lib;
#define PFC_SYM(name)                                                          \
  api.name = reinterpret_cast<decltype(api.name)>(dlsym(lib, "nccl" #name));  \
  if (!api.name)                                                               \
    api.error = "missing symbol nccl" #name;
    PFC_SYM(GetUniqueId)
    PFC_SYM(CommInitRank)
    PFC_SYM(CommDestroy)
    PFC_SYM(AllReduce)
    PFC_SYM(AllGather)
    PFC_SYM(Send)
    PFC_SYM(Recv)
    PFC_SYM(GroupStart)
    PFC_SYM(GroupEnd)
    PFC_SYM(GetErrorString)
#undef PFC_SYM
  });
  return api;
}
} // namespace

#define PFC_NCCL(h, call)                                                                                \
  do                                                                                                     \
    {                                                                                                    \
      ncclResult_t r_ = (call);                                                                          \
      if (r_ != ncclSuccess)                                                                             \
        return (h)->fail(PFC_ERR_NCCL, "%s:%d %s: %s", __FILE__, __LINE__, #call, nccl().GetErrorString(r_)); \
    }                                                                                                    \
  while (0)

extern "C" int pfc_nccl_unique_id(void *id_bytes128)
{
  if (!id_bytes128)
    return PFC_ERR_BAD_ARGUMENT;
  NcclApi &n = nccl();
  if (!n.lib || !n.error.empty())
    return PFC_ERR_NCCL;
  ncclUniqueId id;
  if (n.GetUniqueId(&id) != ncclSuccess)
    return PFC_ERR_NCCL;
  static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
  std::memcpy(id_bytes128, &id, sizeof(id));
  return PFC_OK;
}

extern "C" int pfc_comm_init(pfc_handle *h, const void *id_bytes128, int rank, int n_ranks)
{
  if (!h || !id_bytes128 || n_ranks < 1 || rank < 0 || rank >= n_ranks)
    return PFC_ERR_BAD_ARGUMENT;
  NcclApi &n = nccl();
  if (!n.lib || !n.error.empty())
    return h->fail(PFC_ERR_NCCL, "NCCL unavailable: %s", n.error.c_str());
  PFC_CUDA(h, cudaSetDevice(h->device));
  pfc::comm_destroy(h);
  ncclUniqueId id;
  std::memcpy(&id, id_bytes128, sizeof(id));
  ncclComm_t c = nullptr;
  PFC_NCCL(h, n.CommInitRank(&c, n_ranks, id, rank));
  h->comm = c;
  h->comm_rank = rank;
  h->comm_size = n_ranks;
  return PFC_OK;
}

namespace pfc
{
void comm_destroy(pfc_handle *h)
{
  if (h->comm)
    {
      nccl().CommDestroy((ncclComm_t)h->comm);
      h->comm = nullptr;
    }
  h->comm_rank = 0;
  h->comm_size = 1;
}

int allreduce_sum(pfc_handle *h, double *dev, size_t n)
{
  if (h->comm_size <= 1)
    return PFC_OK;
  PFC_NCCL(h, nccl().AllReduce(dev, dev, n, ncclDouble, ncclSum, (ncclComm_t)h->comm, h->stream));
  return PFC_OK;
}

// Every rank tells the owners which of their nodes it needs; owners turn the lists into internal
// row numbers to pack.  (deal.II/Epetra build the same Import plan inside compress()/vmult.)
int comm_exchange_send_lists(pfc_handle *h)
{
  Plan &P = h->plan;
  NcclApi &n = nccl();
  const int R = P.n_ranks;
  ncclComm_t comm = (ncclComm_t)h->comm;
  cudaStream_t s = h->stream;
  // counts[r][o] = number of nodes rank r needs from owner o
  std::vector<long long> mine(R, 0), all((size_t)R * R, 0);
  for (size_t p = 0; p < P.peers.size(); ++p)
    mine[P.peers[p]] = P.recv_ptr[p + 1] - P.recv_ptr[p];
  DevBuf<long long> d_mine, d_all;
  PFC_CUDA(h, d_mine.upload(mine, s));
  PFC_CUDA(h, d_all.alloc((size_t)R * R));
  PFC_NCCL(h, n.AllGather(d_mine.p, d_all.p, R, ncclInt64, comm, s));
  PFC_CUDA(h, cudaMemcpyAsync(all.data(), d_all.p, all.size() * sizeof(long long), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaStreamSynchronize(s));

  // the peer set must be symmetric for the grouped send/recv: union of who I need and who needs me
  std::vector<int32_t> peers;
  for (int r = 0; r < R; ++r)
    if (r != P.rank && (all[(size_t)P.rank * R + r] > 0 || all[(size_t)r * R + P.rank] > 0))
      peers.push_back(r);
  std::vector<int64_t> recv_ptr(1, 0), send_ptr(1, 0);
  for (int r : peers)
    {
      recv_ptr.push_back(recv_ptr.back() + all[(size_t)P.rank * R + r]);
      send_ptr.push_back(send_ptr.back() + all[(size_t)r * R + P.rank]);
    }
  if (recv_ptr.back() != P.n_ghost)
    return h->fail(PFC_ERR_NCCL, "halo plan: ghost count mismatch");
  DevBuf<int32_t> d_need, d_asked;
  PFC_CUDA(h, d_need.upload(P.ghost_global, s)); // sorted by owner == peer order
  PFC_CUDA(h, d_asked.alloc((size_t)std::max<int64_t>(send_ptr.back(), 1)));
  PFC_NCCL(h, n.GroupStart());
  for (size_t p = 0; p < peers.size(); ++p)
    {
      const int64_t nr = recv_ptr[p + 1] - recv_ptr[p], nsd = send_ptr[p + 1] - send_ptr[p];
      if (nr > 0)
        PFC_NCCL(h, n.Send(d_need.p + recv_ptr[p], (size_t)nr, ncclInt32, peers[p], comm, s));
      if (nsd > 0)
        PFC_NCCL(h, n.Recv(d_asked.p + send_ptr[p], (size_t)nsd, ncclInt32, peers[p], comm, s));
    }
  PFC_NCCL(h, n.GroupEnd());
  std::vector<int32_t> asked((size_t)send_ptr.back());
  if (!asked.empty())
    PFC_CUDA(h, cudaMemcpyAsync(asked.data(), d_asked.p, asked.size() * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  P.send_rows.resize(asked.size());
  for (size_t k = 0; k < asked.size(); ++k)
    {
      const int64_t t = asked[k];
      if (t < P.own0 || t >= P.own0 + P.n_owned)
        return h->fail(PFC_ERR_NCCL, "halo plan: peer asked for a node this rank does not own");
      P.send_rows[k] = P.iperm[t - P.own0];
    }
  P.peers = peers;
  P.recv_ptr = recv_ptr;
  P.send_ptr = send_ptr;
  PFC_CUDA(h, h->d_send_rows.upload(P.send_rows, s));
  PFC_CUDA(h, h->d_send_buf.alloc(3 * std::max<size_t>(P.send_rows.size(), 1)));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  return PFC_OK;
}

int halo_exchange(pfc_handle *h, double *v)
{
  const Plan &P = h->plan;
  if (P.n_ranks <= 1)
    return PFC_OK;
  NcclApi &n = nccl();
  ncclComm_t comm = (ncclComm_t)h->comm;
  const int64_t ns = (int64_t)P.send_rows.size();
  if (ns > 0)
    {
      halo_pack_kernel<<<(int)((ns + 255) / 256), 256, 0, h->stream>>>(h->d_send_buf.p, v, h->d_send_rows.p, ns, P.Np);
      h->launches++;
    }
  PFC_NCCL(h, n.GroupStart());
  for (size_t p = 0; p < P.peers.size(); ++p)
    {
      const int64_t nr = P.recv_ptr[p + 1] - P.recv_ptr[p], nsd = P.send_ptr[p + 1] - P.send_ptr[p];
      if (nsd > 0)
        PFC_NCCL(h, n.Send(h->d_send_buf.p + 3 * P.send_ptr[p], (size_t)(3 * nsd), ncclDouble, P.peers[p], comm, h->stream));
      if (nr > 0)
        PFC_NCCL(h, n.Recv(v + 3 * P.Np + 3 * P.recv_ptr[p], (size_t)(3 * nr), ncclDouble, P.peers[p], comm, h->stream));
    }
  PFC_NCCL(h, n.GroupEnd());
  return PFC_OK;
}

int halo_exchange_scalar(pfc_handle *h, double *v)
{
  const Plan &P = h->plan;
  if (P.n_ranks <= 1)
    return PFC_OK;
  NcclApi &n = nccl();
  ncclComm_t comm = (ncclComm_t)h->comm;
  const int64_t ns = (int64_t)P.send_rows.size();
  if (ns > 0)
    {
      halo_pack1_kernel<<<(int)((ns + 255) / 256), 256, 0, h->stream>>>(h->d_send_buf.p, v, h->d_send_rows.p, ns);
      h->launches++;
    }
  PFC_NCCL(h, n.GroupStart());
  for (size_t p = 0; p < P.peers.size(); ++p)
    {
      const int64_t nr = P.recv_ptr[p + 1] - P.recv_ptr[p], nsd = P.send_ptr[p + 1] - P.send_ptr[p];
      if (nsd > 0)
        PFC_NCCL(h, n.Send(h->d_send_buf.p + P.send_ptr[p], (size_t)nsd, ncclDouble, P.peers[p], comm, h->stream));
      if (nr > 0)
        PFC_NCCL(h, n.Recv(v + P.Np + P.recv_ptr[p], (size_t)nr, ncclDouble, P.peers[p], comm, h->stream));
    }
  PFC_NCCL(h, n.GroupEnd());
  return PFC_OK;
}
} // namespace pfc
