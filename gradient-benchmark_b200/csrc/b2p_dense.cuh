// Dense-state path (sm_100a): 5..64 states, e.g. the 61-state codon models of BASELINE.json
// configs[4].  Here the state contraction is a real 64x64 matrix product per node, the path
// is fp64-compute-bound, and the work is organised as tiled
// matrix products: one CTA = one node x one tile of 64 patterns, 256 threads, the 64 x 64 x 64
// products on the fp64 tensor cores (mma.sync m8n8k4, each warp 64 states x 8 patterns);
// operands staged in shared memory ([state][pattern] layout in HBM, so a tile row is one
// contiguous 512-byte read).  Nodes of one tree level are launched together (level-synchronous:
// at 61 states a level is ~10^2 us of arithmetic, launch latency is noise); tip branches, which
// need no product, run in one launch of their own.  Same message formulation as the 4-state
// walk kernel:
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
constexpr int LD = 68;   // row stride (doubles) of every shared-memory tile: 544 B = 32 B mod 128 B, so the four
                         // k-rows of an mma fragment (lane & 3) land in four different 8-bank groups and a
                         // half-warp's 8-byte fragment read is one conflict-free wavefront
constexpr int DENSE_TILE_BYTES = SP * LD * 8;

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
  int* done_post;            // [N-1][n_tiles] 1 once the node's message tile is in HBM (cleared per evaluation)
  int* done_pre;             // [N-1][n_tiles] 1 once the node's pre-order partial tile is in HBM
  int* err;                  // set when a dependency did not arrive within the time limit
  double* lpart;             // [n_tiles]
  double* gpart;             // [n_tiles][2N-2]
  unsigned long long* prof;  // debug build (-DB2P_DBG, B2P_DENSE_PROF=1): [2][8] summed clocks per phase of the post / pre CTAs
};
#ifdef B2P_DBG
#define DENSE_STAMP(i) do { if (p.prof && threadIdx.x == 0) stamp[i] = clock64(); } while (0)
#define DENSE_STAMPS_FLUSH(k, n) do { if (p.prof && threadIdx.x == 0) { for (int i_ = 1; i_ < (n); ++i_) atomicAdd(p.prof + 8 * (k) + i_, (unsigned long long)(stamp[i_] - stamp[i_ - 1])); atomicAdd(p.prof + 8 * (k), 1ull); } } while (0)
#else
#define DENSE_STAMP(i) do { } while (0)
#define DENSE_STAMPS_FLUSH(k, n) do { } while (0)
#endif

// ---------------------------------------------------------------------------- helpers
// Thread t of the CTA owns, in every 64x64 operand tile, the column pair cp = 2 (t & 31) of the
// rows kr + 8 j (kr = t >> 5, j = 0..7).  All fetches of a tile half (j = 4 h .. 4 h + 3) are issued
// into registers before the first shared-memory store, so a thread keeps up to 16 independent
// 16-byte loads in flight: the load phase runs at memory throughput, not at one latency per row.
constexpr int DH = 4;  // rows per thread and tile half

// Rows k0 + 8 u (u < DH) of child c's message at columns col, col + 1: an internal child's row of
// msg, a tip's P[k][code] (1 for a missing state, 0 in the padding rows).
__device__ __forceinline__ void fetch_message(const DenseParams& p, int c, int col, int k0, unsigned code0, unsigned code1,
                                              double2 (&v)[DH]) {
  if (c >= p.n_tips) {
    const double* g = p.msg + ((size_t)(c - p.n_tips) * SP + k0) * p.cols_pad + col;
#pragma unroll
    for (int u = 0; u < DH; ++u) v[u] = *reinterpret_cast<const double2*>(g + (size_t)u * 8 * p.cols_pad);
  } else {
    const double* g = p.P + ((size_t)c * SP + k0) * SP;
    const bool known0 = code0 < (unsigned)p.n_states, known1 = code1 < (unsigned)p.n_states;
#pragma unroll
    for (int u = 0; u < DH; ++u) {
      const double fill = k0 + 8 * u < p.n_states ? 1.0 : 0.0;
      v[u].x = known0 ? g[u * 8 * SP + code0] : fill;
      v[u].y = known1 ? g[u * 8 * SP + code1] : fill;
    }
  }
}

__device__ __forceinline__ void fetch_codes(const DenseParams& p, int c, int col, unsigned& code0, unsigned& code1) {
  code0 = code1 = 0;
  if (c < p.n_tips) {
    const uchar2 cd = *reinterpret_cast<const uchar2*>(p.codes + (size_t)c * p.cols_pad + col);
    code0 = cd.x; code1 = cd.y;
  }
}

__device__ __forceinline__ void fetch_matrix(const double* __restrict__ g, int k0, int cp, double2 (&v)[DH]) {
#pragma unroll
  for (int u = 0; u < DH; ++u) v[u] = *reinterpret_cast<const double2*>(g + (k0 + 8 * u) * SP + cp);
}

// Programmatic dependent launch: the level kernels are launched with the stream-serialization
// attribute, so a level's CTAs become resident while the previous level drains and run their
// prologue (task, tip codes, matrices that older launches wrote) before level_wait() — which
// returns once the previous launch has completed and its writes are visible.  A CTA releases its
// own dependents only AFTER its wait, so by induction everything older than the previous launch
// is complete whenever a prologue runs.
__device__ __forceinline__ void level_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// One launch per pass, dependencies by flag.  A pass is ONE grid whose blocks are laid out in level
// order (block = task * n_tiles + tile): a task's tile depends only on the same tile of its
// children (post-order) or of its parent (pre-order), so instead of a launch per tree level — with
// every level waiting for the slowest CTA of the one before, and 1-4-node levels near the root
// that cost a launch latency each — a CTA waits for exactly the tiles it reads.  Blocks are
// dispatched in index order, so a producer is always resident or finished before its consumer
// spins: the waits cannot deadlock.  They are bounded all the same (~2 s, then the error word).
__device__ __forceinline__ void tile_wait(const int* flag, int* err) {
  const long long t0 = clock64();
  int v;
  do {
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if (!v && clock64() - t0 > 4000000000ll) { *err = 1; break; }
  } while (!v);
}
// after the CTA's stores of the tile and a __syncthreads(): one thread publishes
__device__ __forceinline__ void tile_publish(int* flag) {
  __threadfence();
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(flag), "r"(1) : "memory");
}

// Operands that need no arithmetic on the way in (matrices, an internal child's own message) go
// straight from global to shared memory (LDGSTS): no registers, so the register-staged operands
// of BOTH tile halves can be in flight with them in one latency round.
__device__ __forceinline__ void copy_rows_async(double (*T)[LD], const double* __restrict__ g, size_t row_stride, int kr, int cp) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(&T[kr + 8 * j][cp]);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(g + (size_t)(kr + 8 * j) * row_stride + cp) : "memory");
  }
}
__device__ __forceinline__ void copy_rows_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ void store_rows(double (*T)[LD], int k0, int cp, const double2 (&v)[DH]) {
#pragma unroll
  for (int u = 0; u < DH; ++u) *reinterpret_cast<double2*>(&T[k0 + 8 * u][cp]) = v[u];
}

// fp64 tensor-core product (mma.sync m8n8k4, the DMMA path of sm_100a: the 4 x 4 register-tile
// version of this loop was bound by shared-memory wavefronts, 8 doubles read per 16 FMAs).
// Warp w owns the 8 columns 8 w .. 8 w + 7 of the tile and all 64 rows:
//   acc[mi][j] = sum_k A[k][8 mi + g] * X[k][8 w + 2 tg + j],  g = lane >> 2, tg = lane & 3.
// Fragments: a = A[k0 + tg][8 mi + g], b = X[k0 + tg][8 w + g]; nine 8-byte reads feed 8 mmas
// (2048 FMAs per warp).
__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
__device__ __forceinline__ void tile_product(const double (*A)[LD], const double (*X)[LD], int warp, int lane, double (&acc)[8][2]) {
  const int g = lane >> 2, tg = lane & 3;
#pragma unroll
  for (int mi = 0; mi < 8; ++mi) acc[mi][0] = acc[mi][1] = 0.0;
#pragma unroll 4
  for (int k0 = 0; k0 < SP; k0 += 4) {
    const double b = X[k0 + tg][8 * warp + g];
    const double* ar = &A[k0 + tg][g];
#pragma unroll
    for (int mi = 0; mi < 8; ++mi) dmma(acc[mi], ar[8 * mi], b);
  }
}

// Two products that share the right-hand operand: acc1 = A1^T-contraction with X, acc2 = A2 with X
// (one fragment of X feeds 16 mmas).
__device__ __forceinline__ void tile_product2(const double (*A1)[LD], const double (*A2)[LD], const double (*X)[LD], int warp, int lane,
                                              double (&acc1)[8][2], double (&acc2)[8][2]) {
  const int g = lane >> 2, tg = lane & 3;
#pragma unroll
  for (int mi = 0; mi < 8; ++mi) acc1[mi][0] = acc1[mi][1] = acc2[mi][0] = acc2[mi][1] = 0.0;
#pragma unroll 2
  for (int k0 = 0; k0 < SP; k0 += 4) {
    const double b = X[k0 + tg][8 * warp + g];
    const double* a1 = &A1[k0 + tg][g];
    const double* a2 = &A2[k0 + tg][g];
#pragma unroll
    for (int mi = 0; mi < 8; ++mi) dmma(acc1[mi], a1[8 * mi], b);
#pragma unroll
    for (int mi = 0; mi < 8; ++mi) dmma(acc2[mi], a2[8 * mi], b);
  }
}

// The thread's part of a shared-memory tile in the accumulator layout: rows 8 mi + g, columns
// 8 w + 2 tg + {0, 1}.
__device__ __forceinline__ void read_block(const double (*T)[LD], int warp, int lane, double (&v)[8][2]) {
  const int g = lane >> 2, tg = lane & 3;
#pragma unroll
  for (int mi = 0; mi < 8; ++mi) {
    const double2 x = *reinterpret_cast<const double2*>(&T[8 * mi + g][8 * warp + 2 * tg]);
    v[mi][0] = x.x; v[mi][1] = x.y;
  }
}

// Sum / maximum over the 8 lanes (same tg) that share the thread's two columns; every one of them
// receives the result.
__device__ __forceinline__ double column_sum(double v) {
#pragma unroll
  for (int o = 4; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int column_max(int v) {
#pragma unroll
  for (int o = 4; o < 32; o <<= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// 1 if the column whose entries (this thread's share) are v[.][j] lies below 2^-200
__device__ __forceinline__ int column_small(const double (&v)[8][2], int j) {
  int h = 0;
#pragma unroll
  for (int mi = 0; mi < 8; ++mi) h = max(h, __double2hiint(v[mi][j]));
  h = column_max(h);
  return (h < 0x33700000 && h > 0) ? 1 : 0;
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

// ------------------------------------------------------------------ transition matrices
// One CTA per branch: P = V diag(exp(lambda r t)) V^-1, its transpose, and D = Q P, as two of the
// tensor-core tile products above.  evecT[j][i] = V[i][j], ivec[j][l] = V^-1[j][l], qT[j][i] = Q[i][j],
// all zero padded to SP x SP (so P, P^T and D come out zero padded).  Dynamic shared memory:
// three tiles (A, X, S).
__global__ void __launch_bounds__(DNT, 2) dense_transition_kernel(const double* __restrict__ evecT, const double* __restrict__ ivec,
                                                                 const double* __restrict__ evals, const double* __restrict__ qT,
                                                                 const double* __restrict__ bl, double rate, int S,
                                                                 double* __restrict__ P, double* __restrict__ PT, double* __restrict__ D) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*A)[LD] = reinterpret_cast<double (*)[LD]>(dsm);
  double (*X)[LD] = reinterpret_cast<double (*)[LD]>(dsm + 1 * DENSE_TILE_BYTES);
  double (*Sm)[LD] = reinterpret_cast<double (*)[LD]>(dsm + 2 * DENSE_TILE_BYTES);
  const int b = blockIdx.x, t = threadIdx.x;
  const int warp = t >> 5, lane = t & 31, g = lane >> 2, tg = lane & 3;
  const int kr = t >> 5, cp = 2 * (t & 31);
  const double tb = bl[b] * rate;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int k0 = kr + 32 * h;
    double2 am[DH], xm[DH];
    fetch_matrix(evecT, k0, cp, am);
    fetch_matrix(ivec, k0, cp, xm);
#pragma unroll
    for (int u = 0; u < DH; ++u) {
      const int k = k0 + 8 * u;
      const double e = k < S ? exp(evals[k] * tb) : 0.0;
      xm[u].x *= e; xm[u].y *= e;
    }
    store_rows(A, k0, cp, am);
    store_rows(X, k0, cp, xm);
  }
  __syncthreads();
  double2 qm[2][DH];
  fetch_matrix(qT, kr, cp, qm[0]);
  fetch_matrix(qT, kr + 32, cp, qm[1]);
  double acc[8][2];
  tile_product(A, X, warp, lane, acc);  // acc = P[8 mi + g][8 w + 2 tg + j]
  __syncthreads();
#pragma unroll
  for (int mi = 0; mi < 8; ++mi) {
    *reinterpret_cast<double2*>(&Sm[8 * mi + g][8 * warp + 2 * tg]) = make_double2(acc[mi][0], acc[mi][1]);
    X[8 * warp + 2 * tg][8 * mi + g] = acc[mi][0];      // X becomes P^T
    X[8 * warp + 2 * tg + 1][8 * mi + g] = acc[mi][1];
  }
  store_rows(A, kr, cp, qm[0]);
  store_rows(A, kr + 32, cp, qm[1]);
  __syncthreads();
  double* p = P + (size_t)b * SP * SP;
  double* pt = PT + (size_t)b * SP * SP;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int k = kr + 8 * j;
    *reinterpret_cast<double2*>(p + k * SP + cp) = *reinterpret_cast<const double2*>(&Sm[k][cp]);
    *reinterpret_cast<double2*>(pt + k * SP + cp) = *reinterpret_cast<const double2*>(&X[k][cp]);
  }
  tile_product(A, Sm, warp, lane, acc);  // D[i][l] = sum_j Q[i][j] P[j][l]
  double* d = D + (size_t)b * SP * SP + g * SP + 8 * warp + 2 * tg;
#pragma unroll
  for (int mi = 0; mi < 8; ++mi) *reinterpret_cast<double2*>(d + 8 * mi * SP) = make_double2(acc[mi][0], acc[mi][1]);
}

// ---------------------------------------------------------------------- post-order pass
// grid: internal nodes in level order x tiles (1-D, tile fastest).  Dynamic shared memory: A [SP][SP], X [SP][DT].
__global__ void __launch_bounds__(DNT, 3) dense_post_kernel(const DenseParams p, const DenseNode* __restrict__ nodes, int root) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*A)[LD] = reinterpret_cast<double (*)[LD]>(dsm);
  double (*X)[LD] = reinterpret_cast<double (*)[LD]>(dsm + DENSE_TILE_BYTES);
  __shared__ double red[DNT / 32];
  const int tile = blockIdx.x % p.n_tiles;
  const DenseNode nd = nodes[blockIdx.x / p.n_tiles];
  const int col0 = tile * DT, t = threadIdx.x;
  const int warp = t >> 5, lane = t & 31, g = lane >> 2, tg = lane & 3;
  const int kr = t >> 5, cp = 2 * (t & 31);
  const bool is_root = nd.v == root;
#ifdef B2P_DBG
  long long stamp[6];
#endif
  DENSE_STAMP(0);
  unsigned lc0, lc1, rc0, rc1;
  fetch_codes(p, nd.l, col0 + cp, lc0, lc1);
  fetch_codes(p, nd.r, col0 + cp, rc0, rc1);
  level_wait();  // P^T comes from the transition launch, the previous one
  if (!is_root) copy_rows_async(A, p.PT + (size_t)nd.v * SP * SP, SP, kr, cp);
  // the children's message tiles (internal children only: a tip's message is a column of P)
  // a tip child's message (columns of P: the transition launch wrote them) does not depend on any
  // flag: requested before the wait for the internal child's tile
  double2 ml[2][DH], mr[2][DH];
  const bool tip_l = nd.l < p.n_tips, tip_r = nd.r < p.n_tips;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    if (tip_l) fetch_message(p, nd.l, col0 + cp, kr + 32 * h, lc0, lc1, ml[h]);
    if (tip_r) fetch_message(p, nd.r, col0 + cp, kr + 32 * h, rc0, rc1, mr[h]);
  }
  if (t == 0 && !tip_l) tile_wait(p.done_post + (size_t)(nd.l - p.n_tips) * p.n_tiles + tile, p.err);
  if (t == 32 && !tip_r) tile_wait(p.done_post + (size_t)(nd.r - p.n_tips) * p.n_tiles + tile, p.err);
  __syncthreads();
  DENSE_STAMP(1);  // 0-1: codes, launch wait, dependency flags
  {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (!tip_l) fetch_message(p, nd.l, col0 + cp, kr + 32 * h, lc0, lc1, ml[h]);
      if (!tip_r) fetch_message(p, nd.r, col0 + cp, kr + 32 * h, rc0, rc1, mr[h]);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
      for (int u = 0; u < DH; ++u) { ml[h][u].x *= mr[h][u].x; ml[h][u].y *= mr[h][u].y; }
      store_rows(X, kr + 32 * h, cp, ml[h]);
    }
  }
  copy_rows_wait();
  __syncthreads();
  DENSE_STAMP(2);  // 1-2: operands from global memory into shared memory
  if (!is_root) {
    // x_v is rescaled BEFORE the product (same rule as the 4-state kernel): the column
    // maximum of X decides, the product carries the factor through.
    int bump[2];
    {
      double x[8][2];
      read_block(X, warp, lane, x);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        bump[j] = column_small(x, j);
        if (bump[j] && g == 0) atomicAdd(p.scount + col0 + 8 * warp + 2 * tg + j, 1);
      }
    }
    double acc[8][2];
    tile_product(A, X, warp, lane, acc);
    DENSE_STAMP(3);  // 2-3: rescale decision + product (thread 0's warp)
    double* out = p.msg + ((size_t)(nd.v - p.n_tips) * SP + g) * p.cols_pad + col0 + 8 * warp + 2 * tg;
    const double f0 = bump[0] ? Scale<double>::factor() : 1.0, f1 = bump[1] ? Scale<double>::factor() : 1.0;
#pragma unroll
    for (int mi = 0; mi < 8; ++mi)
      *reinterpret_cast<double2*>(out + (size_t)8 * mi * p.cols_pad) = make_double2(acc[mi][0] * f0, acc[mi][1] * f1);
    __syncthreads();
    DENSE_STAMP(4);  // 3-4: stores + the slowest warp
    if (t == 0) tile_publish(p.done_post + (size_t)(nd.v - p.n_tips) * p.n_tiles + tile);
    DENSE_STAMP(5);  // 4-5: fence + flag
    DENSE_STAMPS_FLUSH(0, 6);
  } else {
    double term = 0.0;
    if (t < DT) {
      double site = 0.0;
      int h = 0;
      for (int k = 0; k < p.n_states; ++k) { site += p.freqs[k] * X[k][t]; h = max(h, __double2hiint(X[k][t])); }
      const int b = (h < 0x33700000 && h > 0) ? 1 : 0;
      if (b) site *= Scale<double>::factor();
      const double w = p.weights[col0 + t];
      if (w != 0.0) term = w * (log(site) - (double)(__ldcg(p.scount + col0 + t) + b) * Scale<double>::log_factor());
    }
    const double s = cta_sum(term, red);
    if (t == 0) p.lpart[tile] = s;
  }
}

// ----------------------------------------------------------------------- pre-order pass
// grid: internal children by the depth of their parent x tiles (1-D, tile fastest).  A task is {v = parent, l = the
// child c the CTA works for, r = its sibling}.  With q_c = u_v o m_sibling everything outside the
// branch, both products of the task contract q_c:
//   numerator   q.(Q m_c) = (Q^T q).m_c        -> acc1 = Q^T q   (A1 = Q)
//   pre-order   u_c       = P_c^T q            -> acc2           (A2 = P_c)
// so one pass over the fragments of q feeds 16 mmas, both matrices arrive by cp.async while q is
// staged, and m_c is needed only elementwise: it is read from HBM straight into the accumulator
// layout (never staged).  Three tiles (Q, P_c, q) = 102 KB: two CTAs per SM.  Tip children need
// no product and no successor: dense_tip_kernel below handles all of them in one launch.
__global__ void __launch_bounds__(DNT, 2) dense_pre_kernel(const DenseParams p, const DenseNode* __restrict__ tasks, int root) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*A1)[LD] = reinterpret_cast<double (*)[LD]>(dsm);
  double (*A2)[LD] = reinterpret_cast<double (*)[LD]>(dsm + 1 * DENSE_TILE_BYTES);
  double (*Qc)[LD] = reinterpret_cast<double (*)[LD]>(dsm + 2 * DENSE_TILE_BYTES);  // everything outside the branch
  __shared__ double red[DNT / 32];
  const int tile = blockIdx.x % p.n_tiles;
  const DenseNode nd = tasks[blockIdx.x / p.n_tiles];
  const int c = nd.l, sib = nd.r;
  const int col0 = tile * DT, t = threadIdx.x;
  const int warp = t >> 5, lane = t & 31, g = lane >> 2, tg = lane & 3;
  const int kr = t >> 5, cp = 2 * (t & 31);
  const bool is_root = nd.v == root;
#ifdef B2P_DBG
  long long stamp[6];
#endif
  DENSE_STAMP(0);
  unsigned sc0, sc1;
  fetch_codes(p, sib, col0 + cp, sc0, sc1);
  copy_rows_async(A1, p.Q, SP, kr, cp);
  copy_rows_async(A2, p.P + (size_t)c * SP * SP, SP, kr, cp);
  level_wait();  // the messages come from the post-order launch, the previous one
  // What does not depend on the parent's flag is requested BEFORE waiting for it: the sibling's
  // message (into registers) and, towards L2, the child's own message, which is read after the
  // products in the accumulator layout (a 16-byte piece of every row per thread).
  double2 ms[2][DH], uu[2][DH];
#pragma unroll
  for (int h = 0; h < 2; ++h) fetch_message(p, sib, col0 + cp, kr + 32 * h, sc0, sc1, ms[h]);
  {
    const double* gm = p.msg + ((size_t)(c - p.n_tips) * SP + g) * p.cols_pad + col0 + 8 * warp + 2 * tg;
#pragma unroll
    for (int mi = 0; mi < 8; ++mi) asm volatile("prefetch.global.L2 [%0];" ::"l"(gm + (size_t)8 * mi * p.cols_pad));
  }
  // the parent's pre-order partial tile comes from an earlier block of this launch
  if (!is_root) {
    if (t == 0) tile_wait(p.done_pre + (size_t)(nd.v - p.n_tips) * p.n_tiles + tile, p.err);
    __syncthreads();
  }
  DENSE_STAMP(1);
  {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k0 = kr + 32 * h;
      if (is_root) {
#pragma unroll
        for (int u = 0; u < DH; ++u) {
          const double f = k0 + 8 * u < p.n_states ? p.freqs[k0 + 8 * u] : 0.0;
          uu[h][u] = make_double2(f, f);
        }
      } else {
        const double* gp = p.pre + ((size_t)(nd.v - p.n_tips) * SP + k0) * p.cols_pad + col0 + cp;
#pragma unroll
        for (int u = 0; u < DH; ++u) uu[h][u] = *reinterpret_cast<const double2*>(gp + (size_t)u * 8 * p.cols_pad);
      }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
      for (int u = 0; u < DH; ++u) { uu[h][u].x *= ms[h][u].x; uu[h][u].y *= ms[h][u].y; }
      store_rows(Qc, kr + 32 * h, cp, uu[h]);
    }
  }
  copy_rows_wait();
  __syncthreads();
  DENSE_STAMP(2);
  double acc1[8][2], acc2[8][2];
  tile_product2(A1, A2, Qc, warp, lane, acc1, acc2);
  DENSE_STAMP(3);
  // the child's own message in the accumulator layout (rows 8 mi + g, columns 8 w + 2 tg + {0, 1})
  double den[2] = {0.0, 0.0}, num[2] = {0.0, 0.0};
  {
    const double* gm = p.msg + ((size_t)(c - p.n_tips) * SP + g) * p.cols_pad + col0 + 8 * warp + 2 * tg;
    double2 m[8];
#pragma unroll
    for (int mi = 0; mi < 8; ++mi) m[mi] = *reinterpret_cast<const double2*>(gm + (size_t)8 * mi * p.cols_pad);
    double qb[8][2];
    read_block(Qc, warp, lane, qb);
#pragma unroll
    for (int mi = 0; mi < 8; ++mi) {
      den[0] = fma(qb[mi][0], m[mi].x, den[0]); den[1] = fma(qb[mi][1], m[mi].y, den[1]);
      num[0] = fma(acc1[mi][0], m[mi].x, num[0]); num[1] = fma(acc1[mi][1], m[mi].y, num[1]);
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) { den[j] = column_sum(den[j]); num[j] = column_sum(num[j]); }
  }
  const double f0 = column_small(acc2, 0) ? Scale<double>::factor() : 1.0;
  const double f1 = column_small(acc2, 1) ? Scale<double>::factor() : 1.0;
  double* out = p.pre + ((size_t)(c - p.n_tips) * SP + g) * p.cols_pad + col0 + 8 * warp + 2 * tg;
#pragma unroll
  for (int mi = 0; mi < 8; ++mi)
    *reinterpret_cast<double2*>(out + (size_t)8 * mi * p.cols_pad) = make_double2(acc2[mi][0] * f0, acc2[mi][1] * f1);
  __syncthreads();
  DENSE_STAMP(4);
  if (t == 0) tile_publish(p.done_pre + (size_t)(c - p.n_tips) * p.n_tiles + tile);
  DENSE_STAMP(5);
  DENSE_STAMPS_FLUSH(1, 6);
  double gsum = 0.0;
  if (g == 0) {
    const double2 w = *reinterpret_cast<const double2*>(p.weights + col0 + 8 * warp + 2 * tg);
    if (w.x != 0.0) gsum += w.x * p.rate * num[0] / den[0];
    if (w.y != 0.0) gsum += w.y * p.rate * num[1] / den[1];
  }
  const double sg = cta_sum(gsum, red);
  if (t == 0) p.gpart[(size_t)tile * p.n_branches + c] = sg;
}

// ------------------------------------------------------------------------- tip branches
// grid (tips, tiles), after the last pre-order level.  Task {v = parent, l = the tip c, r = its
// sibling}.  For a tip, m_c is column `code` of P_c and Q m_c column `code` of D_c = Q P_c (a
// missing state has m = 1, Q 1 = 0), so its gradient needs no product: warp w sums rows w + 8 j of
// q o m and q o (Q m) for two columns per lane, entirely in registers; eight partial sums per
// column meet in shared memory.  No tile staging: several CTAs share an SM and the launch runs at
// memory speed.
__global__ void __launch_bounds__(DNT, 3) dense_tip_kernel(const DenseParams p, const DenseNode* __restrict__ tasks, int root) {
  __shared__ double sden[DNT / 32][DT], snum[DNT / 32][DT];
  __shared__ double red[DNT / 32];
  const DenseNode nd = tasks[blockIdx.x];
  const int c = nd.l, sib = nd.r;
  const int col0 = blockIdx.y * DT, t = threadIdx.x;
  const int kr = t >> 5, cp = 2 * (t & 31);
  const bool is_root = nd.v == root;
  unsigned sc0, sc1, cc0, cc1;
  fetch_codes(p, sib, col0 + cp, sc0, sc1);
  fetch_codes(p, c, col0 + cp, cc0, cc1);
  const bool known0 = cc0 < (unsigned)p.n_states, known1 = cc1 < (unsigned)p.n_states;
  level_wait();
  double den0 = 0.0, den1 = 0.0, num0 = 0.0, num1 = 0.0;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int k0 = kr + 32 * h;
    double2 mc[DH], ms[DH], uu[DH], dc[DH];
    fetch_message(p, c, col0 + cp, k0, cc0, cc1, mc);
    fetch_message(p, sib, col0 + cp, k0, sc0, sc1, ms);
    const double* dg = p.D + ((size_t)c * SP + k0) * SP;
#pragma unroll
    for (int u = 0; u < DH; ++u) {
      dc[u].x = known0 ? dg[u * 8 * SP + cc0] : 0.0;
      dc[u].y = known1 ? dg[u * 8 * SP + cc1] : 0.0;
    }
    if (is_root) {
#pragma unroll
      for (int u = 0; u < DH; ++u) {
        const double f = k0 + 8 * u < p.n_states ? p.freqs[k0 + 8 * u] : 0.0;
        uu[u] = make_double2(f, f);
      }
    } else {
      const double* gp = p.pre + ((size_t)(nd.v - p.n_tips) * SP + k0) * p.cols_pad + col0 + cp;
#pragma unroll
      for (int u = 0; u < DH; ++u) uu[u] = *reinterpret_cast<const double2*>(gp + (size_t)u * 8 * p.cols_pad);
    }
#pragma unroll
    for (int u = 0; u < DH; ++u) {
      const double q0 = uu[u].x * ms[u].x, q1 = uu[u].y * ms[u].y;
      den0 = fma(q0, mc[u].x, den0); den1 = fma(q1, mc[u].y, den1);
      num0 = fma(q0, dc[u].x, num0); num1 = fma(q1, dc[u].y, num1);
    }
  }
  *reinterpret_cast<double2*>(&sden[kr][cp]) = make_double2(den0, den1);
  *reinterpret_cast<double2*>(&snum[kr][cp]) = make_double2(num0, num1);
  __syncthreads();
  double gsum = 0.0;
  if (t < DT) {
    double den = 0.0, num = 0.0;
#pragma unroll
    for (int w = 0; w < DNT / 32; ++w) { den += sden[w][t]; num += snum[w][t]; }
    const double wt = p.weights[col0 + t];
    if (wt != 0.0) gsum = wt * p.rate * num / den;
  }
  const double sg = cta_sum(gsum, red);
  if (t == 0) p.gpart[(size_t)blockIdx.y * p.n_branches + c] = sg;
}

}  // namespace b2p
