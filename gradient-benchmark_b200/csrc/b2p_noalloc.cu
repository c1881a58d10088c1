// b2p_coalescent_constant_noalloc — an entry point that allocates NOTHING (include/b200phylo.h).
//
// Some callers cannot call cudaMalloc: code under CUDA-graph capture, tight serving loops,
// harnesses that forbid it.  The constant-coalescent kernel (b2p_timetree.cuh) needs scratch
// memory — per-CTA partial sums and a completion ticket — and a set-up object (b2p_timetree)
// that allocates it.  This variant takes only the caller's inputs and outputs; its scratch is
// a file-scope `static __device__` array of a fixed, modest size that every GPU which loads
// this library holds until the process exits.
//
// No counterpart in the reference.
//
// Sharing rules.  Every call on a device uses the same array and finds the last call's
// contents in it: the ticket is cleared on the caller's stream before each launch and the
// partial sums are fully rewritten before they are read.  Two calls that OVERLAP on the GPU
// (different streams, no ordering between them) race on the array — the caller serialises.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <mutex>
#include <string>

#include "../../include/b200phylo.h"

namespace {

constexpr int NA_NT = 256, NA_LANES = 8, NA_EV = NA_NT / NA_LANES;  // events per CTA and pass
constexpr int NA_MAX_GRID = 1024;
constexpr int NA_MAX_DEVICES = 64;
// layout of the scratch array: [NA_MAX_GRID] doubles (per-CTA partial sums) | ticket (unsigned)
constexpr size_t NA_TICKET_OFF = sizeof(double) * NA_MAX_GRID;
constexpr size_t NA_SCRATCH_BYTES = NA_TICKET_OFF + 256;

// The scratch itself.  In HOST code this name is a host-side stand-in, not a device address:
// only cudaGetSymbolAddress (below, once per device) yields the pointer a kernel may be given.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_SCRATCH_BYTES];

thread_local std::string g_noalloc_error;

int na_fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_noalloc_error = buf;
  return code;
}

__device__ __forceinline__ double na_block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < NA_NT / 32; ++w) s += red[w];
  __syncthreads();
  return s;  // valid in thread 0
}

// Constant-population coalescent density of the node heights and its gradient; same sum by
// parts as tt_coalescent_kernel (b2p_timetree.cuh): S = sum over events of t_e x (lineage-pair
// count before - after), where an event's lineage count follows from its RANK in the order
// (height, tips before internal nodes, id).  Without a set-up object there are no sorted tip
// heights, so the tips ordered before an event are counted by scanning too: 8 lanes per event
// scan the N-1 internal heights and the N tip heights (O(N^2 / SMs) compares in all).
//   logp = -S/theta - (N-1) log theta, grad_heights[i] = -(k_i - 1)/theta, grad_theta = S/theta^2 - (N-1)/theta
__global__ void __launch_bounds__(NA_NT) coalescent_noalloc_kernel(int N, const double* __restrict__ tip_h,
                                                                    const double* __restrict__ heights,
                                                                    const double* __restrict__ theta_p, double* __restrict__ logp,
                                                                    double* __restrict__ grad_heights, double* __restrict__ grad_theta,
                                                                    unsigned char* __restrict__ scratch) {
  __shared__ double red[NA_NT / 32];
  __shared__ unsigned last;
  double* parts = reinterpret_cast<double*>(scratch);
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch + NA_TICKET_OFF);
  const int M = N - 1, E = 2 * N - 1;
  const int sub = threadIdx.x & (NA_LANES - 1);
  const double theta = *theta_p;
  double part = 0.0;
  for (int e0 = blockIdx.x * NA_EV; e0 < E; e0 += gridDim.x * NA_EV) {  // whole CTA iterates together
    const int e = e0 + threadIdx.x / NA_LANES;
    const bool live = e < E, internal = e >= N;
    const double te = !live ? 0.0 : internal ? heights[e - N] : tip_h[e];
    int int_before = 0, tip_before = 0;
    if (live) {
      // internal nodes ordered before e: strictly lower; for an internal node also equal height and lower id
      for (int j = sub; j < M; j += NA_LANES) {
        const double hj = heights[j];
        int_before += (hj < te) || (internal && hj == te && j < e - N);
      }
      // tips ordered before e: an internal node comes after every tip of its height; a tip after equal tips of lower id
      for (int j = sub; j < N; j += NA_LANES) {
        const double hj = tip_h[j];
        tip_before += (hj < te) || (hj == te && (internal || j < e));
      }
    }
#pragma unroll
    for (int o = NA_LANES / 2; o > 0; o >>= 1) {
      int_before += __shfl_xor_sync(0xffffffffu, int_before, o);
      tip_before += __shfl_xor_sync(0xffffffffu, tip_before, o);
    }
    if (live && sub == 0) {
      const int k = tip_before - int_before;  // lineages just before the event
      if (internal) {
        part += te * (double)(k - 1);
        if (grad_heights) grad_heights[e - N] = -(double)(k - 1) / theta;
      } else {
        part -= te * (double)k;
      }
    }
  }
  part = na_block_sum(part, red);
  if (threadIdx.x == 0) {
    parts[blockIdx.x] = part;
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  double s = 0.0;
  for (int j = threadIdx.x; j < (int)gridDim.x; j += NA_NT) s += __ldcg(parts + j);  // fixed order: deterministic
  s = na_block_sum(s, red);
  if (threadIdx.x == 0) {
    if (logp) *logp = -s / theta - (double)(N - 1) * log(theta);
    if (grad_theta) *grad_theta = s / (theta * theta) - (double)(N - 1) / theta;
    *ticket = 0u;
  }
}

// Device address of the scratch array, resolved once per device with that device current.
std::mutex g_na_lock;
unsigned char* g_na_ptr[NA_MAX_DEVICES] = {};
int g_na_sms[NA_MAX_DEVICES] = {};

int na_setup(int device, unsigned char** ptr, int* n_sms) {
  std::lock_guard<std::mutex> guard(g_na_lock);
  if (!g_na_ptr[device]) {
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return na_fail(B2P_ERR_CUDA, "b2p_coalescent_constant_noalloc: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    unsigned char* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return na_fail(B2P_ERR_CUDA, "b2p_coalescent_constant_noalloc: cudaGetSymbolAddress: %s (was the library built for this GPU?)",
                     cudaGetErrorString(e));
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) return na_fail(B2P_ERR_CUDA, "b2p_coalescent_constant_noalloc: cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
    g_na_sms[device] = sms;
    g_na_ptr[device] = p;
  }
  *ptr = g_na_ptr[device];
  *n_sms = g_na_sms[device];
  return B2P_OK;
}

}  // namespace

extern "C" {

int32_t b2p_noalloc_max_tips(void) { return B2P_NOALLOC_MAX_TIPS; }

const char* b2p_noalloc_last_error(void) { return g_noalloc_error.c_str(); }

int b2p_coalescent_constant_noalloc(int32_t device, int32_t n_tips, const double* d_tip_heights, const double* d_heights,
                                    const double* d_theta, double* d_logp, double* d_grad_heights, double* d_grad_theta,
                                    void* cuda_stream) {
  if (!d_tip_heights || !d_heights || !d_theta || (!d_logp && !d_grad_heights && !d_grad_theta))
    return na_fail(B2P_ERR_ARG, "b2p_coalescent_constant_noalloc: null pointer");
  if (n_tips < 2) return na_fail(B2P_ERR_ARG, "b2p_coalescent_constant_noalloc: n_tips must be at least 2 (got %d)", n_tips);
  if (n_tips > B2P_NOALLOC_MAX_TIPS)
    return na_fail(B2P_ERR_UNSUPPORTED,
                   "b2p_coalescent_constant_noalloc: n_tips = %d exceeds the fixed limit of %d tips of the no-allocation entry "
                   "point; use b2p_timetree_create + b2p_coalescent_constant for larger trees",
                   n_tips, B2P_NOALLOC_MAX_TIPS);
  if (device < 0 || device >= NA_MAX_DEVICES) return na_fail(B2P_ERR_ARG, "b2p_coalescent_constant_noalloc: device %d out of range", device);
  unsigned char* scratch = nullptr;
  int n_sms = 0;
  int rc = na_setup(device, &scratch, &n_sms);
  if (rc) return rc;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return na_fail(B2P_ERR_CUDA, "b2p_coalescent_constant_noalloc: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  cudaStream_t st = (cudaStream_t)cuda_stream;
  // the array keeps the last call's contents: the ticket must start at zero (the partial sums
  // are rewritten by every CTA before the last one reads them)
  e = cudaMemsetAsync(scratch + NA_TICKET_OFF, 0, sizeof(unsigned), st);
  if (e != cudaSuccess) return na_fail(B2P_ERR_CUDA, "b2p_coalescent_constant_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(e));
  const int E = 2 * n_tips - 1;
  int grid = (E + NA_EV - 1) / NA_EV;
  if (grid > 4 * n_sms) grid = 4 * n_sms;
  if (grid > NA_MAX_GRID) grid = NA_MAX_GRID;
  coalescent_noalloc_kernel<<<grid, NA_NT, 0, st>>>(n_tips, d_tip_heights, d_heights, d_theta, d_logp, d_grad_heights, d_grad_theta, scratch);
  e = cudaGetLastError();
  if (e != cudaSuccess) return na_fail(B2P_ERR_CUDA, "b2p_coalescent_constant_noalloc: launch: %s", cudaGetErrorString(e));
  return B2P_OK;
}

}  // extern "C"
