// pfc_p2p.cuh — device side of the peer-memory protocol (pfc_comm.cu): flag-in-data stores and the slot all-reduce
// that the Gram-Schmidt kernels run in their last block (pfc_krylov.cuh).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace pfc
{
constexpr int kMaxRanksP2P = 8;
constexpr int kArSlot = 72;

struct P2PSync
{ // one per rank, peer-visible
  uint4 ar_ll[2][kMaxRanksP2P][kArSlot]; // [epoch parity][source rank][value]
  unsigned long long halo_epoch, ar_epoch; // local
  unsigned int block_counter;
};


// all-reduce endpoint handed to kernels by value; R == 0: no fused all-reduce (single rank, or NCCL does it)
struct ArDev
{
  P2PSync *sync[kMaxRanksP2P];
  int R, me;
};

__device__ __forceinline__ unsigned int ll_tag(unsigned long long e) { return (unsigned int)(e & 0x7fffffffu) | 0x80000000u; }
__device__ __forceinline__ void ll_store(uint4 *p, double v, unsigned int tag)
{
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"((unsigned int)b), "r"(tag),
               "r"((unsigned int)(b >> 32)), "r"(tag)
               : "memory");
}
// bounded wait: a peer that never arrives must not hang the GPU (trap -> CUDA error on the host)
__device__ __forceinline__ double ll_wait(const uint4 *p, unsigned int tag)
{
  unsigned int lo, t0, hi, t1;
  const long long c0 = clock64();
  for (;;)
    {
      asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(lo), "=r"(t0), "=r"(hi), "=r"(t1) : "l"(p) : "memory");
      if (t0 == tag && t1 == tag)
        break;
      if (clock64() - c0 > 20000000000ll)
        __trap();
    }
  return __longlong_as_double((long long)(((unsigned long long)hi << 32) | lo));
}


// sum of vals[0..n) over all ranks, in rank order, result back in vals[] (identical bits on every rank).
// Called by all threads of ONE block of the grid (blockDim.x >= n), every rank calling in the same sequence.
__device__ __forceinline__ void ar_block_allreduce(const ArDev &A, double *vals, int n)
{
  P2PSync *sl = A.sync[A.me];
  const unsigned long long e = sl->ar_epoch + 1;
  const unsigned int tag = ll_tag(e);
  const int b = (int)(e & 1), t = threadIdx.x;
  if (t < n)
    {
      const double val = vals[t];
      for (int q = 0; q < A.R; ++q)
        ll_store(&A.sync[q]->ar_ll[b][A.me][t], val, tag);
      double s = 0;
      for (int src = 0; src < A.R; ++src)
        s += ll_wait(&sl->ar_ll[b][src][t], tag);
      vals[t] = s;
    }
  __syncthreads();
  if (t == 0)
    sl->ar_epoch = e;
}
} // namespace pfc
