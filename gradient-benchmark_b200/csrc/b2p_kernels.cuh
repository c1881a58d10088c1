// Hand-written sm_100a kernels of the hot path:
//   K1 transition_kernel   P(t), dP/dt for every branch x category           (row a2)
//   K2/K3 traverse_kernel  tile-persistent post-order + pre-order + gradient (rows a3-a6)
//   K4 reduce_kernel       deterministic sum of per-CTA partial results      (rows a4, a6)
//
// Traversal design.  A *column* is one (pattern, category) pair; columns
// never interact except through the K lanes of one pattern (rescaling, the category
// sum).  A *tile* is <= 32 adjacent columns = the lanes of one warp.  A CTA owns one
// tile at a time and walks the WHOLE tree for it: warp w executes slot w of a static
// list schedule (Hu's highest-level-first, built on the host side of the C ABI), W
// independent nodes per round, rounds separated by __syncthreads().  All dependencies
// are therefore CTA-local: no grid sync, no flags, no atomics, one launch for both
// passes, and a child written in round r is re-read from L1/L2 in round r+1.
#pragma once
#include "b2p_device.cuh"

namespace b2p {

constexpr int kMaxCat = 32;

struct ModelParams {  // passed by value to K1
  double evec[16], ivec[16], evals[4];
  double rates[kMaxCat];
};

template <typename T>
struct TraverseParams {
  int n_tips, n_patterns, n_cat, n_cols;  // n_cols = P*K
  int cols_per_tile, n_tiles, n_rounds, n_branches;
  const int4* sched;        // [n_rounds][W] (node, left, right, _); node < 0 = idle slot
  const uint8_t* tips;      // [N][P] state masks
  const double* weights;    // [P]
  const T* pmat;            // [2N-2][K][16]
  const T* dpmat;           // [2N-2][K][16]
  T* post;                  // [N-1][n_cols][4]
  T* pre;                   // [N-1][n_cols][4]
  double* lpart;            // [grid]
  double* gpart;            // [grid][2N-2]
  double freqs[4];
  double cat_w[kMaxCat];
};

// ---------------------------------------------------------------------------- K1
// One thread per (branch, category, matrix entry).  P = V e^{lambda r t} V^-1,
// dP/dt = V (lambda r) e^{lambda r t} V^-1  (= r Q P).  Computed in fp64, stored as T.
template <typename T>
__global__ void transition_kernel(ModelParams m, const double* __restrict__ bl, int n_branches,
                                  int n_cat, T* __restrict__ pmat, T* __restrict__ dpmat) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int total = n_branches * n_cat * 16;
  if (idx >= total) return;
  const int e = idx & 15, bk = idx >> 4;
  const int k = bk % n_cat, b = bk / n_cat;
  const int i = e >> 2, l = e & 3;
  const double rt = m.rates[k], t = bl[b] * rt;
  double p = 0.0, dp = 0.0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const double c = m.evec[i * 4 + j] * m.ivec[j * 4 + l];
    const double ex = exp(m.evals[j] * t);
    p += c * ex;
    dp += c * ex * (m.evals[j] * rt);
  }
  pmat[idx] = (T)p;
  dpmat[idx] = (T)dp;
}

// ------------------------------------------------------------------------- K2/K3
template <typename T> struct MatStride;
template <> struct MatStride<double> { static constexpr int v = 18; };  // 144 B: 16B-aligned, bank shift 4
template <> struct MatStride<float> { static constexpr int v = 20; };   //  80 B: 16B-aligned, bank shift 20

// Shared memory of one warp: D stages, each = 3 operand planes (32 columns) followed by
// 4 x K transition matrices (row stride MatStride).  Sized on the host with the same
// formula (stage_bytes()).
template <typename T>
__host__ __device__ constexpr int op_bytes() { return 32 * 4 * (int)sizeof(T); }
template <typename T>
__host__ __device__ constexpr int stage_bytes(int kk) {
  return 3 * op_bytes<T>() + 4 * kk * MatStride<T>::v * (int)sizeof(T);
}

struct Task {
  int v, l, r, fl;     // fl: bits 0-1 post-order prefetch distance, bits 2-3 pre-order
  unsigned tl, tr;     // tip masks of this lane's pattern (when l / r are tips)
};

template <typename T, int K, int W, int D>
struct Ctx {
  const TraverseParams<T>& p;
  int lane, warp, kk, k, col, pat;
  bool active;
  double w_p;
  T cw;
  unsigned char* wsm;  // this warp's staged region
  int* scount;
  double ll_acc;
  double* gp;
  bool first_tile;
  __device__ __forceinline__ Ctx(const TraverseParams<T>& p_) : p(p_) {}
};

// Issue the asynchronous copies (LDGSTS) of one task's operands and matrices into a
// stage.  PASS 0 = post-order (children partials, P_l, P_r); PASS 1 = pre-order (adds
// the parent's pre-partial and dP_l, dP_r).  Tip children are 1-byte masks carried in
// registers instead.
template <int PASS, typename T, int K, int W, int D>
__device__ __forceinline__ void issue_task(Ctx<T, K, W, D>& c, Task& t, int stage) {
  const TraverseParams<T>& p = c.p;
  const int N = p.n_tips;
  unsigned char* st = c.wsm + stage * stage_bytes<T>(c.kk);
  if (c.active) {
    if (t.l >= N) cp_col(st, c.lane, p.post + ((size_t)(t.l - N) * p.n_cols + c.col) * 4);
    else t.tl = __ldg(p.tips + (size_t)t.l * p.n_patterns + c.pat);
    if (t.r >= N) cp_col(st + op_bytes<T>(), c.lane, p.post + ((size_t)(t.r - N) * p.n_cols + c.col) * 4);
    else t.tr = __ldg(p.tips + (size_t)t.r * p.n_patterns + c.pat);
    if (PASS == 1 && t.v != 2 * N - 2)
      cp_col(st + 2 * op_bytes<T>(), c.lane, p.pre + ((size_t)(t.v - N) * p.n_cols + c.col) * 4);
  }
  constexpr int GPM = 16 * (int)sizeof(T) / 16;  // 16-byte granules per 4x4 matrix
  constexpr int EPG = 16 / (int)sizeof(T);        // elements per granule
  const int per_mat = c.kk * GPM;
  const int total = (PASS ? 4 : 2) * per_mat;
  unsigned char* mats = st + 3 * op_bytes<T>();
  for (int g = c.lane; g < total; g += 32) {
    const int which = g / per_mat, rem = g - which * per_mat;
    const int k_ = rem / GPM, gi = rem - k_ * GPM;
    const T* src = ((which & 2) ? p.dpmat : p.pmat) +
                   ((size_t)((which & 1) ? t.r : t.l) * c.kk + k_) * 16 + gi * EPG;
    cp_async16(mats + ((which * c.kk + k_) * MatStride<T>::v + gi * EPG) * sizeof(T), src);
  }
}

template <typename T, int K, int W, int D>
__device__ __forceinline__ void compute_post(Ctx<T, K, W, D>& c, const Task& t, int stage) {
  const TraverseParams<T>& p = c.p;
  constexpr int MS = MatStride<T>::v;
  const int N = p.n_tips, root = 2 * N - 2;
  const unsigned char* st = c.wsm + stage * stage_bytes<T>(c.kk);
  const T* smk = reinterpret_cast<const T*>(st + 3 * op_bytes<T>()) + c.k * MS;
  const int blk = c.kk * MS;
  T xl[4], xr[4];
  if (t.l < N) tip_col<T>((uint8_t)t.tl, xl); else lds_col(st, c.lane, xl);
  if (t.r < N) tip_col<T>((uint8_t)t.tr, xr); else lds_col(st + op_bytes<T>(), c.lane, xr);
  if (!c.active) {
#pragma unroll
    for (int i = 0; i < 4; ++i) xl[i] = xr[i] = (T)0;
  }
  T a[4], b[4], y[4];
  matvec(smk, xl, a);
  matvec(smk + blk, xr, b);
#pragma unroll
  for (int i = 0; i < 4; ++i) y[i] = a[i] * b[i];
  T mx = fmax(fmax(y[0], y[1]), fmax(y[2], y[3]));
  mx = group_max<K>(mx, c.lane, c.kk);
  int bumped = 0;
  if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= Scale<T>::factor();
    bumped = 1;
  }
  if (c.active) st_col(p.post + ((size_t)(t.v - N) * p.n_cols + c.col) * 4, y);
  if (t.v != root) {
    if (bumped && c.k == 0) atomicAdd(&c.scount[c.lane], 1);
  } else {
    // the root is alone in the last round: every other node's rescale count is final
    T s = (T)p.freqs[0] * y[0] + (T)p.freqs[1] * y[1] + (T)p.freqs[2] * y[2] + (T)p.freqs[3] * y[3];
    s = group_sum<K>(s * c.cw, c.lane, c.kk);
    double term = 0.0;
    if (c.active && c.k == 0)
      term = c.w_p * (log((double)s) - (double)(c.scount[c.lane] + bumped) * Scale<T>::log_factor());
    c.ll_acc += warp_sum(term);
  }
}

template <typename T, int K, int W, int D>
__device__ __forceinline__ void compute_pre(Ctx<T, K, W, D>& c, const Task& t, int stage) {
  const TraverseParams<T>& p = c.p;
  constexpr int MS = MatStride<T>::v;
  const int N = p.n_tips, root = 2 * N - 2;
  const unsigned char* st = c.wsm + stage * stage_bytes<T>(c.kk);
  const T* smk = reinterpret_cast<const T*>(st + 3 * op_bytes<T>()) + c.k * MS;
  const int blk = c.kk * MS;
  T u[4], xl[4], xr[4];
  if (t.l < N) tip_col<T>((uint8_t)t.tl, xl); else lds_col(st, c.lane, xl);
  if (t.r < N) tip_col<T>((uint8_t)t.tr, xr); else lds_col(st + op_bytes<T>(), c.lane, xr);
  if (t.v != root) lds_col(st + 2 * op_bytes<T>(), c.lane, u);
  else {
#pragma unroll
    for (int i = 0; i < 4; ++i) u[i] = (T)p.freqs[i];
  }
  if (!c.active) {
#pragma unroll
    for (int i = 0; i < 4; ++i) xl[i] = xr[i] = u[i] = (T)0;
  }
  T ml[4], mr[4], dl[4], dr[4];
  matvec(smk, xl, ml);
  matvec(smk + blk, xr, mr);
  matvec(smk + 2 * blk, xl, dl);
  matvec(smk + 3 * blk, xr, dr);
  T ql[4], qr[4];
  T den = 0, numl = 0, numr = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    ql[i] = u[i] * mr[i];
    qr[i] = u[i] * ml[i];
    den += ql[i] * ml[i];
    numl += ql[i] * dl[i];
    numr += qr[i] * dr[i];
  }
  den = group_sum<K>(den * c.cw, c.lane, c.kk);
  numl = group_sum<K>(numl * c.cw, c.lane, c.kk);
  numr = group_sum<K>(numr * c.cw, c.lane, c.kk);
  double gl = 0.0, gr = 0.0;
  if (c.active && c.k == 0) {
    const double inv = c.w_p / (double)den;
    gl = (double)numl * inv;
    gr = (double)numr * inv;
  }
  gl = warp_sum(gl);
  gr = warp_sum(gr);
  if (c.lane == 0) {
    if (c.first_tile) { c.gp[t.l] = gl; c.gp[t.r] = gr; }
    else { c.gp[t.l] += gl; c.gp[t.r] += gr; }
  }
  if (t.l >= N) {
    T y[4];
    matvec_t(smk, ql, y);
    const T mx = group_max<K>(fmax(fmax(y[0], y[1]), fmax(y[2], y[3])), c.lane, c.kk);
    if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
      for (int i = 0; i < 4; ++i) y[i] *= Scale<T>::factor();
    }
    if (c.active) st_col(p.pre + ((size_t)(t.l - N) * p.n_cols + c.col) * 4, y);
  }
  if (t.r >= N) {
    T y[4];
    matvec_t(smk + blk, qr, y);
    const T mx = group_max<K>(fmax(fmax(y[0], y[1]), fmax(y[2], y[3])), c.lane, c.kk);
    if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
      for (int i = 0; i < 4; ++i) y[i] *= Scale<T>::factor();
    }
    if (c.active) st_col(p.pre + ((size_t)(t.r - N) * p.n_cols + c.col) * 4, y);
  }
}

// One pass over the schedule with a D-stage asynchronous operand pipeline.  In round rd a
// warp (1) issues its own task's copies if nothing could be prefetched (distance 0),
// (2) issues the copies of its tasks j rounds ahead whose prefetch distance is exactly
// j, (3) waits for the current stage only, computes, stores, (4) joins the CTA barrier.
// Groups are committed in the fixed order A, B_1..B_{D-1} every round so the
// cp.async.wait_group immediates are static.
template <int PASS, typename T, int K, int W, int D>
__device__ __forceinline__ void run_pass(Ctx<T, K, W, D>& c) {
  const TraverseParams<T>& p = c.p;
  const int R = p.n_rounds;
  auto fetch = [&](int rd) -> Task {
    Task t;
    t.v = -1; t.l = t.r = t.fl = 0; t.tl = t.tr = 0;
    if (rd < R) {
      const int4 e = __ldg(p.sched + (PASS ? (R - 1 - rd) : rd) * W + c.warp);
      t.v = e.x; t.l = e.y; t.r = e.z; t.fl = PASS ? ((e.w >> 2) & 3) : (e.w & 3);
    }
    return t;
  };
  Task t[D];
#pragma unroll
  for (int j = 0; j < D; ++j) t[j] = fetch(j);
  int stage = 0;
  for (int rd = 0; rd < R; ++rd) {
    Task tnew = fetch(rd + D);
    if (t[0].v >= 0 && t[0].fl == 0) issue_task<PASS>(c, t[0], stage);
    cp_async_commit();
#pragma unroll
    for (int j = 1; j < D; ++j) {
      int sj = stage + j;
      if (sj >= D) sj -= D;
      if (t[j].v >= 0 && t[j].fl == j) issue_task<PASS>(c, t[j], sj);
      cp_async_commit();
    }
    if (t[0].v >= 0) {
      if (t[0].fl == 0) cp_async_wait<D - 1>();
      else cp_async_wait<(D == 2) ? 1 : 4>();
      __syncwarp();
      if (PASS) compute_pre(c, t[0], stage);
      else compute_post(c, t[0], stage);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j + 1 < D; ++j) t[j] = t[j + 1];
    t[D - 1] = tnew;
    if (++stage == D) stage = 0;
  }
  cp_async_wait<0>();
}

template <typename T, int K, int W, bool GRAD, int MINB, int D>
__global__ void __launch_bounds__(W * 32, MINB) traverse_kernel(const TraverseParams<T> p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int scount[32];
  __shared__ double warp_ll[W];

  Ctx<T, K, W, D> c(p);
  c.lane = threadIdx.x & 31;
  c.warp = threadIdx.x >> 5;
  c.kk = (K > 0) ? K : p.n_cat;
  c.k = c.lane % c.kk;
  c.wsm = smem_raw + (size_t)c.warp * D * stage_bytes<T>(c.kk);
  c.scount = scount;
  c.ll_acc = 0.0;
  c.gp = p.gpart + (size_t)blockIdx.x * p.n_branches;
  c.cw = (T)p.cat_w[c.k];
  c.first_tile = true;

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, c.first_tile = false) {
    c.col = tile * p.cols_per_tile + c.lane;
    c.active = c.lane < p.cols_per_tile && c.col < p.n_cols;
    c.pat = c.active ? c.col / c.kk : 0;
    c.w_p = c.active ? __ldg(p.weights + c.pat) : 0.0;
    if (threadIdx.x < 32) scount[threadIdx.x] = 0;
    __syncthreads();
    run_pass<0>(c);            // post-order (row a3) + root log-likelihood (row a4)
    if (GRAD) run_pass<1>(c);  // pre-order (row a5) + branch gradient (row a6)
  }

  if (c.lane == 0) warp_ll[c.warp] = c.ll_acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < W; ++w) s += warp_ll[w];
    p.lpart[blockIdx.x] = s;
  }
}

// ---------------------------------------------------------------------------- K4
// grad[b] = sum_g gpart[g][b];  logl = sum_g lpart[g].  Fixed summation order.
__global__ void __launch_bounds__(512) reduce_kernel(const double* __restrict__ gpart,
                                                     const double* __restrict__ lpart, int n_parts,
                                                     int n_branches, double* __restrict__ grad,
                                                     double* __restrict__ logl) {
  __shared__ double sh[16][33];
  const int x = threadIdx.x & 31, y = threadIdx.x >> 5;
  if (blockIdx.x == gridDim.x - 1) {  // last block: log-likelihood
    double s = 0.0;
    for (int g = threadIdx.x; g < n_parts; g += 512) s += lpart[g];
    s = warp_sum(s);
    if (x == 0) sh[y][0] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int i = 0; i < 16; ++i) t += sh[i][0];
      *logl = t;
    }
    return;
  }
  const int b = blockIdx.x * 32 + x;
  double s = 0.0;
  if (b < n_branches)
    for (int g = y; g < n_parts; g += 16) s += gpart[(size_t)g * n_branches + b];
  sh[y][x] = s;
  __syncthreads();
  if (y == 0 && b < n_branches) {
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) t += sh[i][x];
    grad[b] = t;
  }
}

}  // namespace b2p
