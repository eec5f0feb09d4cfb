// pfc_noalloc.cu — pfc_dot_noalloc: a dot product / squared norm of two device vectors that allocates nothing.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the norm kernel of the solver
// (multidot_kernel<0> + grid_reduce of pfc_krylov.cuh, behind pfc_residual_norm: system_rhs.l2_norm(), mpi.cpp:584) with its
// scratch — the per-block partial sums and the "last block" counter — in a file-scope static __device__ array instead of
// the handle's cudaMalloc'ed d_partial / d_counter.  The existing entry points and kernels are untouched.
//
// Scratch: kNoallocMaxBlocks partial sums + one counter word, 64 KiB + 256 B per GPU that loads the library, held until
// the process exits.  Block b owns the fixed chunk [b, b + 1) * kNoallocChunk of the vectors, so the partial sums, their
// summation order and therefore the result do not depend on the SM count or on the launch; the array bounds the size:
// n <= PFC_NOALLOC_MAX_N = kNoallocMaxBlocks * kNoallocChunk = 134 217 728 doubles (1 GiB per vector; M4's 3 Np = 50 M).
// Every call shares the array: two calls that overlap on the GPU (different streams, or graphs replayed concurrently)
// race on it; calls on one stream are ordered and safe.  The counter is cleared on the stream before every launch, the
// partial sums are fully rewritten by every launch.
#include <algorithm>
#include <cstdint>

#include "pfc_handle.h"

namespace
{
constexpr int kNoallocThreads = 256;
constexpr int kNoallocChunk = 16384;     // doubles per block
constexpr int kNoallocMaxBlocks = 8192;  // => PFC_NOALLOC_MAX_N
constexpr int kNoallocCounterAt = kNoallocMaxBlocks; // the counter word lives in the 256-byte tail of the array
static_assert((int64_t)kNoallocMaxBlocks * kNoallocChunk == PFC_NOALLOC_MAX_N, "include/pfc_b200.h states the limit");

// In host code this name is a host-side stand-in, NOT a device address: the address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) double g_noalloc_scratch[kNoallocMaxBlocks + 32];

constexpr int kMaxDevices = 64;
double *g_scratch_ptr[kMaxDevices] = {}; // device address of g_noalloc_scratch on device d, resolved once per device

// once per device, with that device current
int noalloc_setup(pfc_handle *h, int dev, double **out)
{
  if (dev < 0 || dev >= kMaxDevices)
    return h->fail(PFC_ERR_BAD_ARGUMENT, "pfc_dot_noalloc: device %d out of range", dev);
  if (!g_scratch_ptr[dev])
    {
      double *p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
      if (e != cudaSuccess || !p)
        return h->fail(PFC_ERR_CUDA, "pfc_dot_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      g_scratch_ptr[dev] = p;
    }
  *out = g_scratch_ptr[dev];
  return PFC_OK;
}

__device__ __forceinline__ double warp_sum_na(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

// out[0] = sum_i x[i] y[i]; block b sums its chunk in a fixed order, the last block to finish sums the partials in block
// order (fixed order as well: the result does not depend on which block that is)
__global__ void __launch_bounds__(kNoallocThreads) noalloc_dot_kernel(const double *__restrict__ x, const double *__restrict__ y,
                                                                     int64_t n, double *partial, unsigned int *counter,
                                                                     double *out)
{
  __shared__ double s_red[kNoallocThreads / 32];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i0 = (int64_t)blockIdx.x * kNoallocChunk, i1 = min(n, i0 + kNoallocChunk);
  double acc = 0;
  for (int64_t i = i0 + threadIdx.x; i < i1; i += kNoallocThreads)
    acc = fma(x[i], y[i], acc);
  acc = warp_sum_na(acc);
  if (lane == 0)
    s_red[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0)
    {
      double s = 0;
#pragma unroll
      for (int w = 0; w < kNoallocThreads / 32; ++w)
        s += s_red[w];
      partial[blockIdx.x] = s;
      __threadfence();
      s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
    }
  __syncthreads();
  if (!s_last)
    return;
  __threadfence();
  double s = 0;
  for (unsigned int b = threadIdx.x; b < gridDim.x; b += kNoallocThreads)
    s += __ldcg(partial + b);
  s = warp_sum_na(s);
  __syncthreads(); // s_red is reused
  if (lane == 0)
    s_red[warp] = s;
  __syncthreads();
  if (threadIdx.x == 0)
    {
      double t = 0;
#pragma unroll
      for (int w = 0; w < kNoallocThreads / 32; ++w)
        t += s_red[w];
      out[0] = t;
      *counter = 0;
    }
}
} // namespace

// include/pfc_b200.h.  x_dev, y_dev: n doubles each in device memory (may be the same pointer: squared norm); out_dev: one
// double in device memory; stream: a cudaStream_t, or NULL for the handle's stream.  Allocates nothing on the device; the
// launch and the 4-byte memset before it are stream-ordered and capturable.  n = 0 writes 0 (and resolves the scratch
// address: a warm-up call before a capture).
extern "C" int pfc_dot_noalloc(pfc_handle *h, const double *x_dev, const double *y_dev, int64_t n, double *out_dev, void *stream)
{
  if (!h)
    return PFC_ERR_BAD_ARGUMENT;
  if (!x_dev || !y_dev || !out_dev || n < 0)
    return h->fail(PFC_ERR_BAD_ARGUMENT, "pfc_dot_noalloc: null pointer or negative length");
  if (n > PFC_NOALLOC_MAX_N)
    return h->fail(PFC_ERR_BAD_ARGUMENT,
                   "pfc_dot_noalloc: n = %lld exceeds PFC_NOALLOC_MAX_N = %lld (the static scratch holds one partial sum per "
                   "%d elements, %d of them)",
                   (long long)n, (long long)PFC_NOALLOC_MAX_N, kNoallocChunk, kNoallocMaxBlocks);
  PFC_CUDA(h, cudaSetDevice(h->device));
  double *scratch = nullptr;
  if (const int rc = noalloc_setup(h, h->device, &scratch); rc != PFC_OK)
    return rc;
  cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
  unsigned int *counter = reinterpret_cast<unsigned int *>(scratch + kNoallocCounterAt);
  PFC_CUDA(h, cudaMemsetAsync(counter, 0, sizeof(unsigned int), s)); // must start at zero, whatever the last call left
  const int blocks = (int)std::max<int64_t>(1, (n + kNoallocChunk - 1) / kNoallocChunk);
  noalloc_dot_kernel<<<blocks, kNoallocThreads, 0, s>>>(x_dev, y_dev, n, scratch, counter, out_dev);
  PFC_CUDA(h, cudaGetLastError());
  return PFC_OK;
}
