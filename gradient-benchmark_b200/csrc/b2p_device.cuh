// Device-side helpers for the 4-state traversal kernels (sm_100a only).
//
// Rescaling constants and warp-level reductions shared by the walk kernel (b2p_walk.cuh),
// the dense-state kernels (b2p_dense.cuh) and the small kernels (b2p_aux.cuh).
// Data layout in HBM: DESIGN.md §3.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2p {

// Lazy power-of-two rescaling thresholds (exact in binary floating point).
template <typename T> struct Scale;
template <> struct Scale<double> {
  static __device__ __forceinline__ double threshold() { return 6.223015277861142e-61; }  // 2^-200
  static __device__ __forceinline__ double factor() { return 1.6069380442589903e+60; }    // 2^+200
  static __device__ __forceinline__ double log_factor() { return 138.62943611198907; }    // 200 ln 2
};
template <> struct Scale<float> {
  static __device__ __forceinline__ float threshold() { return 9.094947017729282e-13f; }  // 2^-40
  static __device__ __forceinline__ float factor() { return 1099511627776.0f; }           // 2^+40
  static __device__ __forceinline__ double log_factor() { return 27.725887222397812; }    // 40 ln 2
};

// ---- reductions over the K category lanes of one pattern -----------------------
// Lanes of a pattern are adjacent and K-aligned (tile origin is a multiple of K).  Every
// shuffle is a FULL-WARP shuffle: all 32 lanes of a warp must reach it (the traversal kernel
// keeps its K != 1 control flow warp-uniform around these calls; lanes without work take part
// with neutral values).  A per-group member mask would let idle lanes skip, but a shuffle whose
// mask differs from lane to lane is serialised by the compiler over the distinct mask values
// (a MATCH.ANY loop: eight trips per shuffle at K = 4) — measured 15 % of the K = 4 kernels.
template <int K, typename T>
__device__ __forceinline__ T group_sum(T v, int lane, int kk) {
  if (K == 1) return v;
  if (K == 2 || K == 4 || K == 8 || K == 16 || K == 32) {
#pragma unroll
    for (int o = 1; o < K; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
  }
  // generic K (runtime kk): gather from the group's lanes
  const int base = lane - lane % kk;
  T s = (T)0;
  for (int j = 0; j < kk; ++j) s += __shfl_sync(0xffffffffu, v, (base + j) & 31);
  return s;
}
template <int K, typename T>
__device__ __forceinline__ T group_max(T v, int lane, int kk) {
  if (K == 1) return v;
  if (K == 2 || K == 4 || K == 8 || K == 16 || K == 32) {
#pragma unroll
    for (int o = 1; o < K; o <<= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
  }
  const int base = lane - lane % kk;
  T s = v;
  for (int j = 0; j < kk; ++j) s = fmax(s, __shfl_sync(0xffffffffu, v, (base + j) & 31));
  return s;
}

template <int K>
__device__ __forceinline__ int group_max_int(int v, int lane, int kk) {
  if (K == 1) return v;
  if (K == 2 || K == 4 || K == 8 || K == 16 || K == 32) {
#pragma unroll
    for (int o = 1; o < K; o <<= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
  }
  const int base = lane - lane % kk;
  int s = v;
  for (int j = 0; j < kk; ++j) s = max(s, __shfl_sync(0xffffffffu, v, (base + j) & 31));
  return s;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace b2p
