// Dense-state path (sm_100a): 5..64 states, e.g. the 61-state codon models of BASELINE.json
// configs[4].  Here the state contraction is a real 64x64 matrix product per node, the path
// is fp64-compute-bound, and the work is organised as tiled
// matrix products: one CTA = one node x one tile of 64 patterns, 256 threads, each thread a
// 4 states x 4 patterns register tile; operands staged in shared memory ([state][pattern]
// layout in HBM, so a tile row is one contiguous 512-byte read).  Nodes of one tree level
// are launched together (level-synchronous: at 61 states a level is ~10^2 us of arithmetic,
// launch latency is noise).  Same message formulation as the 4-state walk kernel:
//   post-order  x_v = m_l o m_r ;  m_v = P_v x_v
//   pre-order   q_l = u_v o m_r, q_r = u_v o m_l ; L = q_l.m_l ; dL/dt_l = r q_l.(Q m_l) ; u_l = P_l^T q_l
// A tip child's message is column `code` of P (Q m = column `code` of Q P), a missing state
// is the all-ones vector.  K = 1 rate category, fp64 only.
#pragma once
#include "b2p_device.cuh"

namespace b2p {

constexpr int SP = 64;   // padded state count
constexpr int DT = 64;   // patterns per tile
constexpr int DNT = 256; // threads per CTA

struct DenseNode { int v, l, r, pad; };

struct DenseParams {
  int n_tips, n_states, n_cols, cols_pad;  // cols_pad: multiple of DT
  int n_tiles, n_branches;
  double rate;               // the single category's rate
  const double* P;           // [2N-2][SP][SP] row-major, zero padded
  const double* PT;          // transposed
  const double* D;           // Q P (tip branches use its columns)
  const double* Q;           // [SP][SP] row-major
  const double* QT;
  const double* freqs;       // [SP]
  const uint8_t* codes;      // [N][cols_pad], >= n_states: missing
  const double* weights;     // [cols_pad], 0 for padding
  double* msg;               // [N-1][SP][cols_pad]
  double* pre;               // [N-1][SP][cols_pad]
  int* scount;               // [cols_pad]
  double* lpart;             // [n_tiles]
  double* gpart;             // [n_tiles][2N-2]
};

// ------------------------------------------------------------------ transition matrices
// One CTA per branch: P = V diag(exp(lambda r t)) V^-1, its transpose, and D = Q P.
__global__ void __launch_bounds__(DNT) dense_transition_kernel(const double* __restrict__ evec, const double* __restrict__ ivec,
                                                              const double* __restrict__ evals, const double* __restrict__ q,
                                                              const double* __restrict__ bl, double rate, int S,
                                                              double* __restrict__ P, double* __restrict__ PT, double* __restrict__ D) {
  __shared__ double ex[SP];
  __shared__ double sp[SP][SP + 1];
  const int b = blockIdx.x;
  if (threadIdx.x < SP) ex[threadIdx.x] = threadIdx.x < S ? exp(evals[threadIdx.x] * bl[b] * rate) : 0.0;
  __syncthreads();
  // four entries per thread at a time: independent accumulators hide the latency of the 61-long chains
  for (int e0 = threadIdx.x; e0 < SP * SP; e0 += 4 * DNT) {
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    const int l = e0 % SP;  // e0 + u * DNT keeps the column (DNT is a multiple of SP)
    if (l < S)
      for (int j = 0; j < S; ++j) {
        const double w = ex[j] * ivec[j * S + l];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = (e0 + u * DNT) / SP;
          if (i < S) acc[u] = fma(evec[i * S + j], w, acc[u]);
        }
      }
#pragma unroll
    for (int u = 0; u < 4; ++u) sp[(e0 + u * DNT) / SP][l] = acc[u];
  }
  __syncthreads();
  double* p = P + (size_t)b * SP * SP;
  double* pt = PT + (size_t)b * SP * SP;
  double* d = D + (size_t)b * SP * SP;
  for (int e0 = threadIdx.x; e0 < SP * SP; e0 += 4 * DNT) {
    const int l = e0 % SP;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (int j = 0; j < S; ++j) {
      const double w = sp[j][l];
#pragma unroll
      for (int u = 0; u < 4; ++u) acc[u] = fma(q[((e0 + u * DNT) / SP) * SP + j], w, acc[u]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int e = e0 + u * DNT, i = e / SP;
      p[e] = sp[i][l];
      pt[e] = sp[l][i];
      d[e] = acc[u];
    }
  }
}

// ---------------------------------------------------------------------------- helpers
// Message of child c for (state k, column col): internal -> msg row, tip -> P[k][code] (or 1).
__device__ __forceinline__ double child_message(const DenseParams& p, int c, int k, int col) {
  if (c >= p.n_tips) return p.msg[((size_t)(c - p.n_tips) * SP + k) * p.cols_pad + col];
  if (k >= p.n_states) return 0.0;
  const unsigned code = p.codes[(size_t)c * p.cols_pad + col];
  return code < (unsigned)p.n_states ? p.P[((size_t)c * SP + k) * SP + code] : 1.0;
}

// acc[a][b] = sum_k A[k][4*ti + a] * X[k][4*tc + b]   (both operands in shared memory)
__device__ __forceinline__ void tile_product(const double (*A)[SP], const double (*X)[DT], int ti, int tc, double (&acc)[4][4]) {
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
#pragma unroll 4
  for (int k = 0; k < SP; ++k) {
    const double2 a0 = *reinterpret_cast<const double2*>(&A[k][4 * ti]);
    const double2 a1 = *reinterpret_cast<const double2*>(&A[k][4 * ti + 2]);
    const double2 x0 = *reinterpret_cast<const double2*>(&X[k][4 * tc]);
    const double2 x1 = *reinterpret_cast<const double2*>(&X[k][4 * tc + 2]);
    const double av[4] = {a0.x, a0.y, a1.x, a1.y}, xv[4] = {x0.x, x0.y, x1.x, x1.y};
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) acc[a][b] = fma(av[a], xv[b], acc[a][b]);
  }
}

__device__ __forceinline__ void load_matrix(double (*A)[SP], const double* __restrict__ g) {
  for (int e = threadIdx.x * 2; e < SP * SP; e += DNT * 2)
    *reinterpret_cast<double2*>(&A[0][0] + e) = *reinterpret_cast<const double2*>(g + e);
}

// Lazy power-of-two rescale of a register tile: the column maximum over all 64 states is
// gathered over the 16 threads (ti = lane & 15) that share the tile's columns.
__device__ __forceinline__ void rescale_tile(double (&acc)[4][4], int (&bumped)[4]) {
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    int h = 0;
#pragma unroll
    for (int a = 0; a < 4; ++a) h = max(h, __double2hiint(acc[a][b]));
#pragma unroll
    for (int o = 1; o < 16; o <<= 1) h = max(h, __shfl_xor_sync(0xffffffffu, h, o));
    bumped[b] = (h < 0x33700000 && h > 0) ? 1 : 0;  // < 2^-200
    if (bumped[b]) {
#pragma unroll
      for (int a = 0; a < 4; ++a) acc[a][b] *= Scale<double>::factor();
    }
  }
}

// sum over the CTA (DNT threads), result valid in thread 0
__device__ __forceinline__ double cta_sum(double v, double* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < DNT / 32; ++w) s += red[w];
  return s;
}

// ---------------------------------------------------------------------- post-order level
// grid (nodes of the level, tiles).  Dynamic shared memory: A [SP][SP], X [SP][DT].
__global__ void __launch_bounds__(DNT, 1) dense_post_kernel(const DenseParams p, const DenseNode* __restrict__ nodes, int root) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*A)[SP] = reinterpret_cast<double (*)[SP]>(dsm);
  double (*X)[DT] = reinterpret_cast<double (*)[DT]>(dsm + SP * SP * 8);
  __shared__ double red[DNT / 32];
  const DenseNode nd = nodes[blockIdx.x];
  const int col0 = blockIdx.y * DT, t = threadIdx.x;
  const int ti = t & 15, tc = t >> 4;
  if (nd.v != root) load_matrix(A, p.PT + (size_t)nd.v * SP * SP);
  for (int e = t; e < SP * DT; e += DNT) {
    const int k = e / DT, c = e % DT;
    X[k][c] = child_message(p, nd.l, k, col0 + c) * child_message(p, nd.r, k, col0 + c);
  }
  __syncthreads();
  if (nd.v != root) {
    // x_v is rescaled BEFORE the product (same rule as the 4-state kernel): the column
    // maximum of X decides, the product carries the factor through.
    double acc[4][4];
    tile_product(A, X, ti, tc, acc);
    // column maxima of x (threads 0..DT-1, one column each)
    __shared__ int bump[DT];
    if (t < DT) {
      int h = 0;
      for (int k = 0; k < SP; ++k) h = max(h, __double2hiint(X[k][t]));
      const int b = (h < 0x33700000 && h > 0) ? 1 : 0;
      bump[t] = b;
      if (b) atomicAdd(p.scount + col0 + t, 1);
    }
    __syncthreads();
    double* out = p.msg + (size_t)(nd.v - p.n_tips) * SP * p.cols_pad + col0 + 4 * tc;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      double y[4];
#pragma unroll
      for (int b = 0; b < 4; ++b) y[b] = bump[4 * tc + b] ? acc[a][b] * Scale<double>::factor() : acc[a][b];
      double* o = out + (size_t)(4 * ti + a) * p.cols_pad;
      *reinterpret_cast<double2*>(o) = make_double2(y[0], y[1]);
      *reinterpret_cast<double2*>(o + 2) = make_double2(y[2], y[3]);
    }
  } else {
    double term = 0.0;
    if (t < DT) {
      double site = 0.0;
      int h = 0;
      for (int k = 0; k < p.n_states; ++k) { site += p.freqs[k] * X[k][t]; h = max(h, __double2hiint(X[k][t])); }
      const int b = (h < 0x33700000 && h > 0) ? 1 : 0;
      if (b) site *= Scale<double>::factor();
      const double w = p.weights[col0 + t];
      if (w != 0.0) term = w * (log(site) - (double)(p.scount[col0 + t] + b) * Scale<double>::log_factor());
    }
    const double s = cta_sum(term, red);
    if (t == 0) p.lpart[blockIdx.y] = s;
  }
}

// ----------------------------------------------------------------------- pre-order level
// grid (nodes of the depth level, tiles, 2): blockIdx.z picks the child the CTA works for, so a
// CTA holds one matrix and that child's message / outside vector (3 x 32 KB): two CTAs fit an
// SM and a narrow level near the root spreads over twice as many SMs.  Both CTAs of a node read
// the same three operand tiles (the second read is an L2 hit).
__global__ void __launch_bounds__(DNT, 2) dense_pre_kernel(const DenseParams p, const DenseNode* __restrict__ nodes, int root) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*A)[SP] = reinterpret_cast<double (*)[SP]>(dsm);
  double (*M)[DT] = reinterpret_cast<double (*)[DT]>(dsm + 1 * SP * SP * 8);   // the child's message
  double (*Qc)[DT] = reinterpret_cast<double (*)[DT]>(dsm + 2 * SP * SP * 8);  // everything outside its branch
  __shared__ double red[DNT / 32];
  __shared__ double s_inv[DT], s_num[DT];
  const DenseNode nd = nodes[blockIdx.x];
  const int side = blockIdx.z;
  const int c = side ? nd.r : nd.l;
  const int col0 = blockIdx.y * DT, t = threadIdx.x;
  const int ti = t & 15, tc = t >> 4;
  for (int e = t; e < SP * DT; e += DNT) {
    const int k = e / DT, cc = e % DT;
    const double ml = child_message(p, nd.l, k, col0 + cc), mr = child_message(p, nd.r, k, col0 + cc);
    const double u = nd.v == root ? (k < p.n_states ? p.freqs[k] : 0.0)
                                  : p.pre[((size_t)(nd.v - p.n_tips) * SP + k) * p.cols_pad + col0 + cc];
    M[k][cc] = side ? mr : ml;
    Qc[k][cc] = u * (side ? ml : mr);
  }
  if (t < DT) s_num[t] = 0.0;
  __syncthreads();
  // site likelihood (denominator) and, for a tip child, the numerator: one column per thread
  if (t < DT) {
    double den = 0.0;
    for (int k = 0; k < SP; ++k) den += Qc[k][t] * M[k][t];
    s_inv[t] = p.weights[col0 + t] * p.rate / den;
    if (c < p.n_tips) {
      const unsigned code = p.codes[(size_t)c * p.cols_pad + col0 + t];
      double num = 0.0;  // missing state: Q 1 = 0
      if (code < (unsigned)p.n_states) {
        const double* d = p.D + (size_t)c * SP * SP + code;
        for (int i = 0; i < p.n_states; ++i) num += Qc[i][t] * d[(size_t)i * SP];
      }
      s_num[t] = num;
    }
  }
  // internal child: numerator q.(Q m) and the child's pre-order partial P^T q
  if (c >= p.n_tips) {
    double acc[4][4];
    __syncthreads();
    load_matrix(A, p.QT);  // (Q m)[i] = sum_k Q[i][k] m[k]: A[k][i] = Q^T
    __syncthreads();
    tile_product(A, M, ti, tc, acc);
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      double part = 0.0;
#pragma unroll
      for (int a = 0; a < 4; ++a) part += Qc[4 * ti + a][4 * tc + b] * acc[a][b];
#pragma unroll
      for (int o = 1; o < 16; o <<= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
      if (ti == 0) s_num[4 * tc + b] = part;
    }
    __syncthreads();
    load_matrix(A, p.P + (size_t)c * SP * SP);  // (P^T q)[j] = sum_i P[i][j] q[i]: A[i][j] = P
    __syncthreads();
    tile_product(A, Qc, ti, tc, acc);
    int bumped[4];
    rescale_tile(acc, bumped);
    double* out = p.pre + (size_t)(c - p.n_tips) * SP * p.cols_pad + col0 + 4 * tc;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      double* o = out + (size_t)(4 * ti + a) * p.cols_pad;
      *reinterpret_cast<double2*>(o) = make_double2(acc[a][0], acc[a][1]);
      *reinterpret_cast<double2*>(o + 2) = make_double2(acc[a][2], acc[a][3]);
    }
  }
  __syncthreads();
  double g = 0.0;
  if (t < DT && p.weights[col0 + t] != 0.0) g = s_num[t] * s_inv[t];
  const double sg = cta_sum(g, red);
  if (t == 0) p.gpart[(size_t)blockIdx.y * p.n_branches + c] = sg;
}

}  // namespace b2p
