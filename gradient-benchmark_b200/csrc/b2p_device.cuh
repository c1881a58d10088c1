// Device-side helpers for the 4-state traversal kernels (sm_100a only).
//
// Data layout in HBM:
//   partials  [internal node][column][state]   column = pattern * K + category
//   one column of one node = 4 states = 32 B (fp64) / 16 B (fp32) = one 256-bit or
//   128-bit access per thread, so a warp moves 1 KB / 512 B contiguous per node.
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

// ---- one column (4 states) global load / store --------------------------------
// Partials are written and re-read inside one launch, so these are coherent
// (non-.nc) accesses.  fp64 uses the sm_100 256-bit LDG/STG.
__device__ __forceinline__ void ld_col(const double* p, double (&x)[4]) {
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(x[0]), "=d"(x[1]), "=d"(x[2]), "=d"(x[3])
               : "l"(p)
               : "memory");
}
__device__ __forceinline__ void st_col(double* p, const double (&x)[4]) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(x[0]), "d"(x[1]),
               "d"(x[2]), "d"(x[3])
               : "memory");
}
__device__ __forceinline__ void ld_col(const float* p, float (&x)[4]) {
  asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(x[0]), "=f"(x[1]), "=f"(x[2]), "=f"(x[3])
               : "l"(p)
               : "memory");
}
__device__ __forceinline__ void st_col(float* p, const float (&x)[4]) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(x[0]), "f"(x[1]),
               "f"(x[2]), "f"(x[3])
               : "memory");
}


// ---- Ampere-style async copies (LDGSTS): global -> shared without registers -------
// .cg = cache at L2 only, so a column written by another warp of this CTA earlier in
// the launch is always read coherently.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// One column staged in shared memory.  fp64: two 16 B halves in separate 512 B planes
// (conflict-free LDS.128); fp32: one 16 B unit.
__device__ __forceinline__ void lds_col(const unsigned char* op, int lane, double (&x)[4]) {
  const double2 a = *reinterpret_cast<const double2*>(op + lane * 16);
  const double2 b = *reinterpret_cast<const double2*>(op + 512 + lane * 16);
  x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y;
}
__device__ __forceinline__ void lds_col(const unsigned char* op, int lane, float (&x)[4]) {
  const float4 a = *reinterpret_cast<const float4*>(op + lane * 16);
  x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
}
__device__ __forceinline__ void cp_col(unsigned char* op, int lane, const double* g) {
  cp_async16(op + lane * 16, g);
  cp_async16(op + 512 + lane * 16, g + 2);
}
__device__ __forceinline__ void cp_col(unsigned char* op, int lane, const float* g) {
  cp_async16(op + lane * 16, g);
}

template <typename T>
__device__ __forceinline__ void tip_col(uint8_t m, T (&x)[4]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) x[j] = (T)((m >> j) & 1);
}

// ---- reductions over the K category lanes of one pattern -----------------------
// Lanes of a pattern are adjacent and K-aligned (tile origin is a multiple of K).
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

// 16-byte aligned row of a staged 4x4 matrix
__device__ __forceinline__ void load_row(const double* m, double (&r)[4]) {
  const double2 a = *reinterpret_cast<const double2*>(m);
  const double2 b = *reinterpret_cast<const double2*>(m + 2);
  r[0] = a.x; r[1] = a.y; r[2] = b.x; r[3] = b.y;
}
__device__ __forceinline__ void load_row(const float* m, float (&r)[4]) {
  const float4 a = *reinterpret_cast<const float4*>(m);
  r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
}

// y = M x with M row-major 4x4 in shared memory (broadcast / conflict-free reads)
template <typename T>
__device__ __forceinline__ void matvec(const T* __restrict__ m, const T (&x)[4], T (&y)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
    y[i] = m[4 * i] * x[0] + m[4 * i + 1] * x[1] + m[4 * i + 2] * x[2] + m[4 * i + 3] * x[3];
}
// y = M^T x
template <typename T>
__device__ __forceinline__ void matvec_t(const T* __restrict__ m, const T (&x)[4], T (&y)[4]) {
#pragma unroll
  for (int j = 0; j < 4; ++j)
    y[j] = m[j] * x[0] + m[4 + j] * x[1] + m[8 + j] * x[2] + m[12 + j] * x[3];
}

}  // namespace b2p
