// pfc_comm.cu — NCCL plumbing (see pfc_comm.h).  NCCL is resolved at run time with dlopen so
// that libpfc_b200.so has no link-time dependency on it.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "pfc_comm.h"
#include "pfc_handle.h"
#include "pfc_p2p.cuh"

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

extern "C" int pfc_time_comm(pfc_handle *h, int32_t reps, double *ms3)
{
  using namespace pfc;
  if (!h || reps < 1 || !ms3)
    return PFC_ERR_BAD_ARGUMENT;
  if (!h->is_setup)
    return h->fail(PFC_ERR_STATE, "pfc_time_comm before pfc_setup");
  ms3[0] = ms3[1] = ms3[2] = 0;
  if (h->comm_size <= 1)
    return PFC_OK;
  PFC_CUDA(h, cudaSetDevice(h->device));
  cudaStream_t s = h->stream;
  const bool pc = h->prm.preconditioner == PFC_PRECOND_BLOCK_SCHUR;
  for (int leg = 0; leg < 3; ++leg)
    {
      if (leg == 0 && !pc)
        continue;
      cudaGraph_t graph = nullptr;
      cudaGraphExec_t exec = nullptr;
      const int64_t before = h->launches;
      PFC_CUDA(h, cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
      comm_p2p_unknown_predecessor(h);
      int rc = PFC_OK;
      for (int i = 0; i < reps && rc == PFC_OK; ++i)
        { // alternate between two vectors, as the solver does
          if (leg == 0)
            rc = halo_exchange_scalar(h, h->d_pw.p + (i & 1) * h->sstride);
          else if (leg == 1)
            rc = halo_exchange(h, h->d_V.p + (i & 1) * h->stride);
          else
            rc = allreduce_sum(h, h->d_small.p->dots1, 16);
        }
      cudaError_t e = cudaStreamEndCapture(s, &graph);
      h->launches = before;
      if (rc != PFC_OK)
        {
          if (graph)
            cudaGraphDestroy(graph);
          return rc;
        }
      PFC_CUDA(h, e);
      PFC_CUDA(h, cudaGraphInstantiate(&exec, graph, 0));
      cudaGraphDestroy(graph);
      cudaGraphLaunch(exec, s);
      PFC_CUDA(h, cudaEventRecord(h->ev[0], s));
      cudaGraphLaunch(exec, s);
      PFC_CUDA(h, cudaEventRecord(h->ev[1], s));
      PFC_CUDA(h, cudaStreamSynchronize(s));
      comm_p2p_unknown_predecessor(h);
      cudaGraphExecDestroy(exec);
      float ms;
      PFC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
      ms3[leg] = ms / reps;
    }
  PFC_CUDA(h, cudaMemsetAsync(h->d_small.p, 0, sizeof(GmresSmall), s));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  return PFC_OK;
}

extern "C" int pfc_comm_mode(const pfc_handle *h) { return !h || h->comm_size <= 1 ? 0 : (h->p2p ? 2 : 1); }

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
static int p2p_allreduce(pfc_handle *h, double *dev, size_t n, const Ctl *ctl);
static int p2p_halo(pfc_handle *h, double *v, int nc, const Ctl *ctl);

void comm_destroy(pfc_handle *h)
{
  comm_p2p_destroy(h);
  if (h->comm)
    {
      nccl().CommDestroy((ncclComm_t)h->comm);
      h->comm = nullptr;
    }
  h->comm_rank = 0;
  h->comm_size = 1;
}

int allreduce_sum(pfc_handle *h, double *dev, size_t n, const Ctl *ctl)
{
  if (h->comm_size <= 1)
    return PFC_OK;
  if (h->p2p)
    return p2p_allreduce(h, dev, n, ctl);
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

// ring overlay (lattice path): tail[self_dst] = owned[self_src] (positions this rank owns itself: periodic wrap) and,
// on the NCCL transport, received values from the staging buffer to their ring positions
template <int NC>
static __global__ void ring_unpack_kernel(double *__restrict__ v, int64_t Np, const int32_t *__restrict__ self_src,
                                          const int32_t *__restrict__ self_dst, int64_t n_self,
                                          const double *__restrict__ stage, const int32_t *__restrict__ recv_pos,
                                          int64_t n_recv, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n_self; i += nth)
    {
      const int64_t s = self_src[i], t = Np + self_dst[i];
      if (NC == 3)
        {
#pragma unroll
          for (int c = 0; c < 3; ++c)
            v[3 * t + c] = v[(int64_t)c * Np + s];
        }
      else
        v[t] = v[s];
    }
  for (int64_t g = tid; g < n_recv; g += nth)
    {
      const int64_t t = Np + recv_pos[g];
      if (NC == 3)
        {
#pragma unroll
          for (int c = 0; c < 3; ++c)
            v[3 * t + c] = stage[3 * g + c];
        }
      else
        v[t] = stage[g];
    }
}

static int ring_unpack(pfc_handle *h, double *v, int nc, bool staged, const Ctl *ctl)
{
  const Plan &P = h->plan;
  const int64_t n_self = (int64_t)P.ring.self_src.size(), n_recv = staged ? (int64_t)P.ring.recv_pos.size() : 0;
  const int64_t work = std::max(n_self, n_recv);
  if (work == 0)
    return PFC_OK;
  const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((work + 255) / 256, (int64_t)h->n_sm * 4));
  if (nc == 3)
    ring_unpack_kernel<3><<<blocks, 256, 0, h->stream>>>(v, P.Np, h->d_self_src.p, h->d_self_dst.p, n_self, h->d_recv_buf.p,
                                                          h->d_recv_pos.p, n_recv, ctl);
  else
    ring_unpack_kernel<1><<<blocks, 256, 0, h->stream>>>(v, P.Np, h->d_self_src.p, h->d_self_dst.p, n_self, h->d_recv_buf.p,
                                                          h->d_recv_pos.p, n_recv, ctl);
  h->launches++;
  return PFC_OK;
}

static int nccl_halo(pfc_handle *h, double *v, int nc, const Ctl *ctl)
{
  const Plan &P = h->plan;
  NcclApi &n = nccl();
  ncclComm_t comm = (ncclComm_t)h->comm;
  const bool ring = P.ring.active;
  const int64_t ns = (int64_t)P.send_rows.size();
  if (ns > 0)
    {
      if (nc == 3)
        halo_pack_kernel<<<(int)((ns + 255) / 256), 256, 0, h->stream>>>(h->d_send_buf.p, v, h->d_send_rows.p, ns, P.Np);
      else
        halo_pack1_kernel<<<(int)((ns + 255) / 256), 256, 0, h->stream>>>(h->d_send_buf.p, v, h->d_send_rows.p, ns);
      h->launches++;
    }
  // ghosts sorted by owner: received straight into the tail; ring: into the staging buffer, then to their positions
  double *dst = ring ? h->d_recv_buf.p : v + (int64_t)nc * P.Np;
  PFC_NCCL(h, n.GroupStart());
  for (size_t p = 0; p < P.peers.size(); ++p)
    {
      const int64_t nr = P.recv_ptr[p + 1] - P.recv_ptr[p], nsd = P.send_ptr[p + 1] - P.send_ptr[p];
      if (nsd > 0)
        PFC_NCCL(h, n.Send(h->d_send_buf.p + nc * P.send_ptr[p], (size_t)(nc * nsd), ncclDouble, P.peers[p], comm, h->stream));
      if (nr > 0)
        PFC_NCCL(h, n.Recv(dst + nc * P.recv_ptr[p], (size_t)(nc * nr), ncclDouble, P.peers[p], comm, h->stream));
    }
  PFC_NCCL(h, n.GroupEnd());
  if (ring)
    return ring_unpack(h, v, nc, true, ctl);
  return PFC_OK;
}

static int halo_any(pfc_handle *h, double *v, int nc, const Ctl *ctl)
{
  const Plan &P = h->plan;
  if (P.n_ranks <= 1)
    return P.ring.active ? ring_unpack(h, v, nc, false, ctl) : PFC_OK; // single rank: periodic wrap-around copies
  if (h->p2p)
    return p2p_halo(h, v, nc, ctl);
  return nccl_halo(h, v, nc, ctl);
}

// device copies of the ring lists (after ring_overlay); the NCCL transport also gets its staging buffers
int comm_ring_upload(pfc_handle *h)
{
  const Plan &P = h->plan;
  cudaStream_t s = h->stream;
  PFC_CUDA(h, h->d_self_src.upload(P.ring.self_src, s));
  PFC_CUDA(h, h->d_self_dst.upload(P.ring.self_dst, s));
  PFC_CUDA(h, h->d_recv_pos.upload(P.ring.recv_pos, s));
  PFC_CUDA(h, h->d_send_rows.upload(P.send_rows, s));
  PFC_CUDA(h, h->d_send_buf.alloc(3 * std::max<size_t>(P.send_rows.size(), 1)));
  PFC_CUDA(h, h->d_recv_buf.alloc(3 * std::max<size_t>(P.ring.recv_pos.size(), 1)));
  PFC_CUDA(h, cudaStreamSynchronize(s));
  return PFC_OK;
}

// all ranks' RingCards (collective); single rank: the own card
int comm_gather_cards(pfc_handle *h, const RingCard &mine, std::vector<RingCard> &all)
{
  const int R = h->plan.n_ranks;
  all.assign((size_t)R, mine);
  if (R <= 1)
    return PFC_OK;
  static_assert(sizeof(RingCard) % 8 == 0, "RingCard is exchanged as int64 words");
  NcclApi &n = nccl();
  const size_t words = sizeof(RingCard) / 8;
  DevBuf<long long> d_mine, d_all;
  PFC_CUDA(h, d_mine.alloc(words));
  PFC_CUDA(h, d_all.alloc(words * R));
  PFC_CUDA(h, cudaMemcpyAsync(d_mine.p, &mine, sizeof(RingCard), cudaMemcpyHostToDevice, h->stream));
  PFC_NCCL(h, n.AllGather(d_mine.p, d_all.p, words, ncclInt64, (ncclComm_t)h->comm, h->stream));
  PFC_CUDA(h, cudaMemcpyAsync(all.data(), d_all.p, sizeof(RingCard) * R, cudaMemcpyDeviceToHost, h->stream));
  PFC_CUDA(h, cudaStreamSynchronize(h->stream));
  return PFC_OK;
}

int halo_exchange(pfc_handle *h, double *v, const Ctl *ctl) { return halo_any(h, v, 3, ctl); }
int halo_exchange_scalar(pfc_handle *h, double *v, const Ctl *ctl) { return halo_any(h, v, 1, ctl); }

// =====================================================================================================
// Peer-to-peer path over NVLink / NVSwitch (single node).  A preconditioned GMRES iteration has ~14 halo
// exchanges and 2 all-reduces of a few doubles; what they cost is latency, not bandwidth.  Here both are ONE
// kernel each, no fence and no separate flag: every double travels as one 16-byte store {lo, tag, hi, tag}
// into a peer-mapped (CUDA IPC) buffer of the receiver, 8-byte halves are written atomically, and the
// receiver polls the slot until both tags carry the current epoch (the scheme of NCCL's LL protocol).
//   * halo exchange: store the values the neighbours need into their buffers, then poll the own buffer and
//     unpack into the ghost tail of the vector (the consumer kernel follows in stream order);
//   * all-reduce: store the <= 72 values into the own slot row on every rank, poll the rows of all ranks and
//     sum them in rank order -> bitwise identical results on all ranks (fixed order).
// Epochs are counted on the device (the kernels sit in CUDA graphs that are replayed) and all ranks run the
// same sequence of collective kernels.  A rank cannot send exchange n+2 before it has received n+1, which
// its neighbour sends after it has consumed n, so two buffer halves (epoch parity) are enough.
// =====================================================================================================
struct P2PState
{
  int R = 0, me = 0;
  P2PSync *sync[kMaxRanksP2P] = {}; // [rank]: mapped address (own: local allocation)
  uint4 *ll[kMaxRanksP2P] = {};     // [rank]: halo buffer [2][3 n_ghost(rank)]
  int64_t n_ghost[kMaxRanksP2P] = {};
  int64_t recv_off[kMaxRanksP2P] = {}; // [rank q]: where my values start in q's ghost order (nodes), -1 none
  bool mapped_ll[kMaxRanksP2P] = {}, mapped_sync[kMaxRanksP2P] = {};
};

struct P2PHaloArgs
{
  uint4 *dst[kMaxRanksP2P];       // peer buffer, already offset to my first slot (parity 0)
  int64_t half[kMaxRanksP2P];     // slots per parity half of that peer: 3 n_ghost(peer)
  int64_t send_ptr[kMaxRanksP2P + 1];
  int n_peers;
  const uint4 *mine;              // own buffer
  int64_t my_half, n_ghost;       // n_ghost: values received (slots of the own buffer in use)
  // ring overlay: the k-th received value goes to ring position recv_pos[k] (NULL: ghost k), and the positions this
  // rank owns itself are copied locally
  const int32_t *recv_pos, *self_src, *self_dst;
  int64_t n_self;
};


// grid is capped (all CTAs resident): every thread first stores all its values, then polls
template <int NC>
static __global__ void __launch_bounds__(256) p2p_halo_kernel(P2PHaloArgs H, double *__restrict__ v,
                                                               const int32_t *__restrict__ rows, int64_t Np,
                                                               P2PSync *sl, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  const unsigned long long e = sl->halo_epoch + 1;
  const unsigned int tag = ll_tag(e);
  const int64_t par = (int64_t)(e & 1);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  const int64_t n_send = H.send_ptr[H.n_peers];
  for (int64_t k = tid; k < n_send; k += nth)
    {
      int p = 0;
      while (k >= H.send_ptr[p + 1])
        ++p;
      const int64_t j = k - H.send_ptr[p];
      const int32_t r = rows[k];
      uint4 *d = H.dst[p] + par * H.half[p] + NC * j;
#pragma unroll
      for (int c = 0; c < NC; ++c)
        ll_store(d + c, v[(int64_t)c * Np + r], tag);
    }
  for (int64_t i = tid; i < H.n_self; i += nth)
    {
      const int64_t s = H.self_src[i], t = Np + H.self_dst[i];
      if (NC == 3)
        {
#pragma unroll
          for (int c = 0; c < 3; ++c)
            v[3 * t + c] = v[(int64_t)c * Np + s];
        }
      else
        v[t] = v[s];
    }
  const uint4 *m = H.mine + par * H.my_half;
  for (int64_t g = tid; g < H.n_ghost; g += nth)
    {
      const int64_t t = Np + (H.recv_pos ? (int64_t)H.recv_pos[g] : g);
      if (NC == 3)
        {
#pragma unroll
          for (int c = 0; c < 3; ++c)
            v[3 * t + c] = ll_wait(m + 3 * g + c, tag);
        }
      else
        v[t] = ll_wait(m + g, tag);
    }
  // the last CTA to finish advances the epoch (the next collective kernel starts after this one has ended)
  __syncthreads();
  if (threadIdx.x == 0 && atomicAdd(&sl->block_counter, 1u) == gridDim.x - 1)
    {
      sl->halo_epoch = e;
      sl->block_counter = 0;
    }
}

static __global__ void __launch_bounds__(128) p2p_allreduce_kernel(ArDev A, double *dev, int n, const Ctl *ctl)
{
  if (ctl && ctl->state != 0)
    return;
  ar_block_allreduce(A, dev, n);
}

ArDev comm_ar_endpoint(const pfc_handle *h)
{
  ArDev A;
  std::memset(&A, 0, sizeof(A));
  if (h->p2p && h->p2p_state)
    {
      for (int q = 0; q < kMaxRanksP2P; ++q)
        A.sync[q] = h->p2p_state->sync[q];
      A.R = h->p2p_state->R;
      A.me = h->p2p_state->me;
    }
  return A;
}

static int p2p_allreduce(pfc_handle *h, double *dev, size_t n, const Ctl *ctl)
{
  if (n > (size_t)kArSlot)
    return h->fail(PFC_ERR_NCCL, "p2p all-reduce of %zu values (max %d)", n, kArSlot);
  p2p_allreduce_kernel<<<1, 128, 0, h->stream>>>(comm_ar_endpoint(h), dev, (int)n, ctl);
  h->launches++;
  return PFC_OK;
}

static int p2p_halo(pfc_handle *h, double *v, int nc, const Ctl *ctl)
{
  const Plan &P = h->plan;
  P2PState &S = *h->p2p_state;
  P2PHaloArgs H;
  std::memset(&H, 0, sizeof(H));
  H.n_peers = (int)P.peers.size();
  for (int p = 0; p < H.n_peers; ++p)
    {
      const int q = P.peers[p];
      H.half[p] = 3 * S.n_ghost[q];
      H.dst[p] = S.ll[q] + (int64_t)nc * std::max<int64_t>(S.recv_off[q], 0);
      H.send_ptr[p] = P.send_ptr[p];
    }
  H.send_ptr[H.n_peers] = P.send_ptr[H.n_peers];
  H.mine = S.ll[S.me];
  H.my_half = 3 * P.n_ghost;
  H.n_ghost = P.ring.active ? (int64_t)P.ring.recv_pos.size() : P.n_ghost;
  if (P.ring.active)
    {
      H.recv_pos = h->d_recv_pos.p;
      H.self_src = h->d_self_src.p;
      H.self_dst = h->d_self_dst.p;
      H.n_self = (int64_t)P.ring.self_src.size();
    }
  const int64_t work = std::max<int64_t>(std::max<int64_t>(H.send_ptr[H.n_peers], H.n_ghost), H.n_self);
  const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((work + 255) / 256, (int64_t)h->n_sm * 2));
  if (nc == 3)
    p2p_halo_kernel<3><<<blocks, 256, 0, h->stream>>>(H, v, h->d_send_rows.p, P.Np, S.sync[S.me], ctl);
  else
    p2p_halo_kernel<1><<<blocks, 256, 0, h->stream>>>(H, v, h->d_send_rows.p, P.Np, S.sync[S.me], ctl);
  h->launches++;
  return PFC_OK;
}

void comm_p2p_destroy(pfc_handle *h)
{
  P2PState *S = h->p2p_state;
  if (!S)
    return;
  cudaSetDevice(h->device);
  if (h->stream)
    cudaStreamSynchronize(h->stream);
  for (int q = 0; q < kMaxRanksP2P; ++q)
    {
      if (S->mapped_ll[q])
        cudaIpcCloseMemHandle(S->ll[q]);
      if (S->mapped_sync[q])
        cudaIpcCloseMemHandle(S->sync[q]);
    }
  if (S->sync[S->me])
    cudaFree(S->sync[S->me]);
  if (S->ll[S->me])
    cudaFree(S->ll[S->me]);
  delete S;
  h->p2p_state = nullptr;
  h->p2p = false;
}

void comm_p2p_unknown_predecessor(pfc_handle *) {} // the flag-in-data protocol needs no ordering hint

// collective: every rank publishes IPC handles of its halo buffer and sync block plus its ghost layout; peers are
// mapped with cudaIpcOpenMemHandle.  Any failure on any rank (no peer access, PFC_NO_P2P=1, more than 8 ranks)
// leaves all ranks on the NCCL path.
int comm_p2p_setup(pfc_handle *h)
{
  comm_p2p_destroy(h);
  const Plan &P = h->plan;
  NcclApi &n = nccl();
  ncclComm_t comm = (ncclComm_t)h->comm;
  cudaStream_t s = h->stream;
  const int R = P.n_ranks, me = P.rank;
  struct Card
  {
    cudaIpcMemHandle_t ll, sync;
    int64_t n_ghost, ok;
    int64_t recv_off[kMaxRanksP2P]; // [source rank] start (nodes) of that rank's values in my ghost order
  };
  static_assert(sizeof(Card) % 8 == 0, "Card is exchanged as int64 words");
  Card mine;
  std::memset(&mine, 0, sizeof(mine));
  const char *env = getenv("PFC_NO_P2P");
  bool ok = R <= kMaxRanksP2P && !(env && atoi(env) != 0);
  // The two-parity LL buffer is paced by the two-way traffic of a pair: a rank cannot send exchange n + 2 to a peer before
  // it has received n + 1 from it.  A pair with data in one direction only has no such pacing, so any such pair on any
  // rank keeps ALL ranks on NCCL (the ring overlay makes every pair two-way with a dummy value, pfc_ring_plan.cpp; the
  // general ghost lists of the assembled path are symmetric on every mesh make_grid produces, but nothing enforces it).
  for (size_t p = 0; p < P.peers.size() && ok; ++p)
    if ((P.recv_ptr[p + 1] - P.recv_ptr[p] > 0) != (P.send_ptr[p + 1] - P.send_ptr[p] > 0))
      ok = false;
  P2PState *S = new P2PState();
  S->R = R;
  S->me = me;
  P2PSync *sl = nullptr;
  uint4 *ll = nullptr;
  const size_t ll_bytes = sizeof(uint4) * 2 * 3 * (size_t)std::max<int64_t>(P.n_ghost, 1);
  if (ok && (cudaMalloc((void **)&sl, sizeof(P2PSync)) != cudaSuccess || cudaMalloc((void **)&ll, ll_bytes) != cudaSuccess))
    ok = false;
  if (ok)
    {
      cudaMemsetAsync(sl, 0, sizeof(P2PSync), s);
      cudaMemsetAsync(ll, 0, ll_bytes, s);
      cudaStreamSynchronize(s);
      ok = cudaIpcGetMemHandle(&mine.ll, ll) == cudaSuccess && cudaIpcGetMemHandle(&mine.sync, sl) == cudaSuccess;
    }
  cudaGetLastError();
  S->sync[me] = sl;
  S->ll[me] = ll;
  mine.n_ghost = P.n_ghost;
  mine.ok = ok ? 1 : 0;
  for (int q = 0; q < kMaxRanksP2P; ++q)
    mine.recv_off[q] = -1;
  for (size_t p = 0; p < P.peers.size(); ++p)
    if (P.peers[p] < kMaxRanksP2P)
      mine.recv_off[P.peers[p]] = P.recv_ptr[p];
  std::vector<Card> all((size_t)R);
  {
    DevBuf<long long> d_mine, d_all;
    const size_t words = sizeof(Card) / 8;
    PFC_CUDA(h, d_mine.alloc(words));
    PFC_CUDA(h, d_all.alloc(words * R));
    PFC_CUDA(h, cudaMemcpyAsync(d_mine.p, &mine, sizeof(Card), cudaMemcpyHostToDevice, s));
    PFC_NCCL(h, n.AllGather(d_mine.p, d_all.p, words, ncclInt64, comm, s));
    PFC_CUDA(h, cudaMemcpyAsync(all.data(), d_all.p, sizeof(Card) * R, cudaMemcpyDeviceToHost, s));
    PFC_CUDA(h, cudaStreamSynchronize(s));
  }
  for (int q = 0; q < R; ++q)
    ok = ok && all[q].ok == 1;
  if (ok)
    for (int q = 0; q < R && ok; ++q)
      {
        S->n_ghost[q] = all[q].n_ghost;
        S->recv_off[q] = all[q].recv_off[me];
        if (q == me)
          continue;
        void *pl = nullptr, *ps = nullptr;
        if (cudaIpcOpenMemHandle(&ps, all[q].sync, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
          ok = false;
        else
          {
            S->sync[q] = (P2PSync *)ps;
            S->mapped_sync[q] = true;
          }
        if (ok && cudaIpcOpenMemHandle(&pl, all[q].ll, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
          ok = false;
        else if (ok)
          {
            S->ll[q] = (uint4 *)pl;
            S->mapped_ll[q] = true;
          }
      }
  cudaGetLastError();
  // agree on the outcome: one rank without peer access sends everybody to NCCL
  {
    DevBuf<long long> d_flag;
    long long f = ok ? 1 : 0, fsum = 0;
    PFC_CUDA(h, d_flag.alloc(1));
    PFC_CUDA(h, cudaMemcpyAsync(d_flag.p, &f, sizeof(f), cudaMemcpyHostToDevice, s));
    PFC_NCCL(h, n.AllReduce(d_flag.p, d_flag.p, 1, ncclInt64, ncclSum, comm, s));
    PFC_CUDA(h, cudaMemcpyAsync(&fsum, d_flag.p, sizeof(fsum), cudaMemcpyDeviceToHost, s));
    PFC_CUDA(h, cudaStreamSynchronize(s));
    ok = fsum == R;
  }
  h->p2p_state = S;
  h->p2p = ok;
  if (!ok)
    comm_p2p_destroy(h);
  return PFC_OK;
}
} // namespace pfc
