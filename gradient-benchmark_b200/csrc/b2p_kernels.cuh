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

// One schedule slot: 32 bytes, read as two int4.  Offsets are pre-multiplied on the host
// so the kernel forms an operand address with one IMAD.WIDE:
//   internal node id: off = (id - N) * n_cols   (column units, same for post and pre)
//   tip id:           off = id * n_patterns     (bytes into the tip-mask array)
struct SchedEntry {
  int v, l, r, fl;  // v < 0 = idle slot; fl bits 0-1 / 2-3 = post / pre prefetch distance
  unsigned voff, loff, roff, pad;
};

template <typename T>
struct TraverseParams {
  int n_tips, n_patterns, n_cat, n_cols;  // n_cols = P*K
  int cols_per_tile, n_tiles, n_rounds, n_branches;
  const SchedEntry* sched;  // [n_rounds][W]
  const uint8_t* tips;      // [N][P] state masks
  const double* weights;    // [P]
  const T* pmat;            // [2N-2][K][16]
  T* post;                  // [N-1][n_cols][4]
  T* pre;                   // [N-1][n_cols][4]
  double* lpart;            // [grid]
  double* gpart;            // [grid][2N-2]
  T q[16];                  // rate matrix Q (unit rate), row-major
  double freqs[4];
  double cat_w[kMaxCat];
  double rates[kMaxCat];
};

// ---------------------------------------------------------------------------- K1
// One thread per (branch, category, matrix entry).  P = V e^{lambda r t} V^-1, computed
// in fp64, stored as T.  (dP/dt = r Q P is never formed: the pre-order kernel applies Q
// from the constant bank to P x instead.)
template <typename T>
__global__ void transition_kernel(ModelParams m, const double* __restrict__ bl, int n_branches,
                                  int n_cat, T* __restrict__ pmat) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int total = n_branches * n_cat * 16;
  if (idx >= total) return;
  const int e = idx & 15, bk = idx >> 4;
  const int k = bk % n_cat, b = bk / n_cat;
  const int i = e >> 2, l = e & 3;
  const double t = bl[b] * m.rates[k];
  double p = 0.0;
#pragma unroll
  for (int j = 0; j < 4; ++j) p += m.evec[i * 4 + j] * m.ivec[j * 4 + l] * exp(m.evals[j] * t);
  pmat[idx] = (T)p;
}

// ------------------------------------------------------------------------- K2/K3
template <typename T> struct MatStride;
template <> struct MatStride<double> { static constexpr int v = 18; };  // 144 B: 16B-aligned, bank shift 4
template <> struct MatStride<float> { static constexpr int v = 20; };   //  80 B: 16B-aligned, bank shift 20

// Shared memory of one warp: D stages, each = 3 operands x NC column planes (32 lanes
// each) followed by 2 x K transition matrices (row stride MatStride); then a 2-slot ring
// of schedule entries.  Sized on the host with the same formulas.
template <typename T>
__host__ __device__ constexpr int op_bytes() { return 32 * 4 * (int)sizeof(T); }
template <typename T>
__host__ __device__ constexpr int stage_bytes(int kk, int nc) {
  return 3 * nc * op_bytes<T>() + 2 * kk * MatStride<T>::v * (int)sizeof(T);
}
template <typename T>
__host__ __device__ constexpr int warp_smem_bytes(int kk, int nc, int d) {
  return d * stage_bytes<T>(kk, nc) + 64;
}

struct Task {
  int v, l, r, fl;
  unsigned voff, loff, roff;
  unsigned tl, tr;  // tip masks, 8 bits per sub-column
};

template <typename T, int K, int W, int NC, int D>
struct Ctx {
  const TraverseParams<T>& p;
  int lane, warp, kk, k, cpw;
  bool act[NC];
  double w_p[NC];
  T cw, cwr;            // category weight, weight * rate
  T* post_c;            // this lane's column 0 in node 0 of post / pre
  T* pre_c;
  const uint8_t* tips_c;
  unsigned char* wsm;   // this warp's staged region
  int* scount;          // [NC][32]
  double ll_acc;
  double* gp;
  bool first_tile;
  __device__ __forceinline__ Ctx(const TraverseParams<T>& p_) : p(p_) {}
};

// Issue the asynchronous copies (LDGSTS) of one task's operands and matrices into a
// stage.  PASS 0 = post-order (children partials); PASS 1 = pre-order (adds the parent's
// pre-partial).  Tip children are 1-byte masks carried in registers instead.
template <int PASS, typename T, int K, int W, int NC, int D>
__device__ __forceinline__ void issue_task(Ctx<T, K, W, NC, D>& c, Task& t, int stage) {
  const TraverseParams<T>& p = c.p;
  const int N = p.n_tips;
  unsigned char* st = c.wsm + stage * stage_bytes<T>(c.kk, NC);
  const int patstep = c.cpw / c.kk;
  if (t.l < N) t.tl = 0;
  if (t.r < N) t.tr = 0;
#pragma unroll
  for (int j = 0; j < NC; ++j) {
    if (c.act[j]) {
      const size_t cj = (size_t)j * c.cpw * 4;
      if (t.l >= N) cp_col(st + (0 * NC + j) * op_bytes<T>(), c.lane, c.post_c + cj + (size_t)t.loff * 4);
      else t.tl |= (unsigned)__ldg(c.tips_c + j * patstep + t.loff) << (8 * j);
      if (t.r >= N) cp_col(st + (1 * NC + j) * op_bytes<T>(), c.lane, c.post_c + cj + (size_t)t.roff * 4);
      else t.tr |= (unsigned)__ldg(c.tips_c + j * patstep + t.roff) << (8 * j);
      if (PASS == 1 && t.v != 2 * N - 2)
        cp_col(st + (2 * NC + j) * op_bytes<T>(), c.lane, c.pre_c + cj + (size_t)t.voff * 4);
    }
  }
  constexpr int GPM = (int)sizeof(T);       // 16-byte granules per 4x4 matrix (8 fp64 / 4 fp32)
  constexpr int EPG = 16 / (int)sizeof(T);  // elements per granule
  const int per_mat = c.kk * GPM;
  unsigned char* mats = st + 3 * NC * op_bytes<T>();
  for (int g = c.lane; g < 2 * per_mat; g += 32) {
    const int which = g >= per_mat, rem = g - which * per_mat;
    const int k_ = rem / GPM, gi = rem - k_ * GPM;
    const T* src = p.pmat + ((size_t)(which ? t.r : t.l) * c.kk + k_) * 16 + gi * EPG;
    cp_async16(mats + ((which * c.kk + k_) * MatStride<T>::v + gi * EPG) * sizeof(T), src);
  }
}

template <typename T, int K, int W, int NC, int D>
__device__ __forceinline__ void compute_post(Ctx<T, K, W, NC, D>& c, const Task& t, int stage) {
  const TraverseParams<T>& p = c.p;
  constexpr int MS = MatStride<T>::v;
  const int N = p.n_tips, root = 2 * N - 2;
  const unsigned char* st = c.wsm + stage * stage_bytes<T>(c.kk, NC);
  const T* ml_ = reinterpret_cast<const T*>(st + 3 * NC * op_bytes<T>()) + c.k * MS;
  const T* mr_ = ml_ + c.kk * MS;
  T xl[NC][4], xr[NC][4], y[NC][4];
#pragma unroll
  for (int j = 0; j < NC; ++j) {
    if (t.l < N) tip_col<T>((uint8_t)(t.tl >> (8 * j)), xl[j]);
    else lds_col(st + (0 * NC + j) * op_bytes<T>(), c.lane, xl[j]);
    if (t.r < N) tip_col<T>((uint8_t)(t.tr >> (8 * j)), xr[j]);
    else lds_col(st + (1 * NC + j) * op_bytes<T>(), c.lane, xr[j]);
    if (!c.act[j]) {
#pragma unroll
      for (int i = 0; i < 4; ++i) xl[j][i] = xr[j][i] = (T)0;
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    T rl[4], rr[4];
    load_row(ml_ + 4 * i, rl);
    load_row(mr_ + 4 * i, rr);
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      const T a = rl[0] * xl[j][0] + rl[1] * xl[j][1] + rl[2] * xl[j][2] + rl[3] * xl[j][3];
      const T b = rr[0] * xr[j][0] + rr[1] * xr[j][1] + rr[2] * xr[j][2] + rr[3] * xr[j][3];
      y[j][i] = a * b;
    }
  }
  int bumped[NC];
#pragma unroll
  for (int j = 0; j < NC; ++j) {
    T mx = fmax(fmax(y[j][0], y[j][1]), fmax(y[j][2], y[j][3]));
    mx = group_max<K>(mx, c.lane, c.kk);
    bumped[j] = 0;
    if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
      for (int i = 0; i < 4; ++i) y[j][i] *= Scale<T>::factor();
      bumped[j] = 1;
    }
    if (c.act[j]) st_col(c.post_c + (size_t)j * c.cpw * 4 + (size_t)t.voff * 4, y[j]);
  }
  if (t.v != root) {
#pragma unroll
    for (int j = 0; j < NC; ++j)
      if (bumped[j] && c.k == 0) atomicAdd(&c.scount[j * 32 + c.lane], 1);
  } else {
    // the root is alone in the last round: every other node's rescale count is final
    double term = 0.0;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      T s = (T)p.freqs[0] * y[j][0] + (T)p.freqs[1] * y[j][1] + (T)p.freqs[2] * y[j][2] +
            (T)p.freqs[3] * y[j][3];
      s = group_sum<K>(s * c.cw, c.lane, c.kk);
      if (c.act[j] && c.k == 0)
        term += c.w_p[j] * (log((double)s) -
                            (double)(c.scount[j * 32 + c.lane] + bumped[j]) * Scale<T>::log_factor());
    }
    c.ll_acc += warp_sum(term);
  }
}

template <typename T, int K, int W, int NC, int D>
__device__ __forceinline__ void compute_pre(Ctx<T, K, W, NC, D>& c, const Task& t, int stage) {
  const TraverseParams<T>& p = c.p;
  constexpr int MS = MatStride<T>::v;
  const int N = p.n_tips, root = 2 * N - 2;
  const unsigned char* st = c.wsm + stage * stage_bytes<T>(c.kk, NC);
  const T* ml_ = reinterpret_cast<const T*>(st + 3 * NC * op_bytes<T>()) + c.k * MS;
  const T* mr_ = ml_ + c.kk * MS;
  T ql[NC][4], qr[NC][4];  // first hold x_l, x_r, then q_l, q_r
  T ml[NC][4], mr[NC][4];
#pragma unroll
  for (int j = 0; j < NC; ++j) {
    if (t.l < N) tip_col<T>((uint8_t)(t.tl >> (8 * j)), ql[j]);
    else lds_col(st + (0 * NC + j) * op_bytes<T>(), c.lane, ql[j]);
    if (t.r < N) tip_col<T>((uint8_t)(t.tr >> (8 * j)), qr[j]);
    else lds_col(st + (1 * NC + j) * op_bytes<T>(), c.lane, qr[j]);
    if (!c.act[j]) {
#pragma unroll
      for (int i = 0; i < 4; ++i) ql[j][i] = qr[j][i] = (T)0;
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    T rl[4], rr[4];
    load_row(ml_ + 4 * i, rl);
    load_row(mr_ + 4 * i, rr);
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      ml[j][i] = rl[0] * ql[j][0] + rl[1] * ql[j][1] + rl[2] * ql[j][2] + rl[3] * ql[j][3];
      mr[j][i] = rr[0] * qr[j][0] + rr[1] * qr[j][1] + rr[2] * qr[j][2] + rr[3] * qr[j][3];
    }
  }
  double gl = 0.0, gr = 0.0;
#pragma unroll
  for (int j = 0; j < NC; ++j) {
    T u[4];
    if (t.v != root) lds_col(st + (2 * NC + j) * op_bytes<T>(), c.lane, u);
    else {
#pragma unroll
      for (int i = 0; i < 4; ++i) u[i] = (T)p.freqs[i];
    }
    if (!c.act[j]) {
#pragma unroll
      for (int i = 0; i < 4; ++i) u[i] = (T)0;
    }
    T den = 0, numl = 0, numr = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      ql[j][i] = u[i] * mr[j][i];
      qr[j][i] = u[i] * ml[j][i];
      den += ql[j][i] * ml[j][i];
    }
    // d/dt (P x) = r Q (P x): Q comes from the constant bank, no staged dP matrices
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const T dl = p.q[4 * i] * ml[j][0] + p.q[4 * i + 1] * ml[j][1] + p.q[4 * i + 2] * ml[j][2] +
                   p.q[4 * i + 3] * ml[j][3];
      const T dr = p.q[4 * i] * mr[j][0] + p.q[4 * i + 1] * mr[j][1] + p.q[4 * i + 2] * mr[j][2] +
                   p.q[4 * i + 3] * mr[j][3];
      numl += ql[j][i] * dl;
      numr += qr[j][i] * dr;
    }
    den = group_sum<K>(den * c.cw, c.lane, c.kk);
    numl = group_sum<K>(numl * c.cwr, c.lane, c.kk);
    numr = group_sum<K>(numr * c.cwr, c.lane, c.kk);
    if (c.act[j] && c.k == 0) {
      const double inv = c.w_p[j] / (double)den;
      gl += (double)numl * inv;
      gr += (double)numr * inv;
    }
  }
  gl = warp_sum(gl);
  gr = warp_sum(gr);
  if (c.lane == 0) {
    if (c.first_tile) { c.gp[t.l] = gl; c.gp[t.r] = gr; }
    else { c.gp[t.l] += gl; c.gp[t.r] += gr; }
  }
  // pre_c = P_c^T q_c for internal children (ml / mr are dead: reuse as outputs)
  if (t.l >= N || t.r >= N) {
#pragma unroll
    for (int j = 0; j < NC; ++j)
#pragma unroll
      for (int i = 0; i < 4; ++i) ml[j][i] = mr[j][i] = (T)0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      T rl[4], rr[4];
      load_row(ml_ + 4 * i, rl);
      load_row(mr_ + 4 * i, rr);
#pragma unroll
      for (int j = 0; j < NC; ++j)
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          ml[j][jj] += rl[jj] * ql[j][i];
          mr[j][jj] += rr[jj] * qr[j][i];
        }
    }
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      if (t.l >= N) {
        const T mx = group_max<K>(fmax(fmax(ml[j][0], ml[j][1]), fmax(ml[j][2], ml[j][3])), c.lane, c.kk);
        if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
          for (int i = 0; i < 4; ++i) ml[j][i] *= Scale<T>::factor();
        }
        if (c.act[j]) st_col(c.pre_c + (size_t)j * c.cpw * 4 + (size_t)t.loff * 4, ml[j]);
      }
      if (t.r >= N) {
        const T mx = group_max<K>(fmax(fmax(mr[j][0], mr[j][1]), fmax(mr[j][2], mr[j][3])), c.lane, c.kk);
        if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
          for (int i = 0; i < 4; ++i) mr[j][i] *= Scale<T>::factor();
        }
        if (c.act[j]) st_col(c.pre_c + (size_t)j * c.cpw * 4 + (size_t)t.roff * 4, mr[j]);
      }
    }
  }
}

// One pass over the schedule with a D-stage asynchronous operand pipeline.  In round rd a
// warp (1) issues its own task's copies if nothing could be prefetched (distance 0),
// (2) issues the copies of its tasks j rounds ahead whose prefetch distance is exactly
// j, (3) waits for the current stage only, computes, stores, (4) joins the CTA barrier.
// Groups are committed in the fixed order A, B_1..B_{D-1} every round so the
// cp.async.wait_group immediates are static.  Schedule entries travel through a 2-slot
// shared-memory ring, fetched D+1 rounds ahead inside group A.
template <int PASS, typename T, int K, int W, int NC, int D>
__device__ __forceinline__ void run_pass(Ctx<T, K, W, NC, D>& c) {
  const TraverseParams<T>& p = c.p;
  const int R = p.n_rounds;
  unsigned char* ring = c.wsm + D * stage_bytes<T>(c.kk, NC);
  auto gaddr = [&](int rd) -> const int4* {
    return reinterpret_cast<const int4*>(p.sched + (size_t)(PASS ? (R - 1 - rd) : rd) * W + c.warp);
  };
  auto unpack = [&](const int4& a, const int4& b) -> Task {
    Task t;
    t.v = a.x; t.l = a.y; t.r = a.z; t.fl = PASS ? ((a.w >> 2) & 3) : (a.w & 3);
    t.voff = (unsigned)b.x; t.loff = (unsigned)b.y; t.roff = (unsigned)b.z;
    t.tl = t.tr = 0;
    if (t.v < 0) t.fl = 0;
    return t;
  };
  auto idle = [&]() -> Task {
    Task t;
    t.v = -1; t.l = t.r = t.fl = 0; t.voff = t.loff = t.roff = 0; t.tl = t.tr = 0;
    return t;
  };
  auto fetch_direct = [&](int rd) -> Task {
    if (rd >= R) return idle();
    const int4* g = gaddr(rd);
    return unpack(__ldg(g), __ldg(g + 1));
  };
  auto issue_entry = [&](int rd) {
    if (rd < R && c.lane < 2) cp_async16(ring + (rd & 1) * 32 + c.lane * 16, gaddr(rd) + c.lane);
  };
  auto decode = [&](int rd) -> Task {
    if (rd >= R) return idle();
    const int4* s = reinterpret_cast<const int4*>(ring + (rd & 1) * 32);
    return unpack(s[0], s[1]);
  };

  Task t[D];
#pragma unroll
  for (int j = 0; j < D; ++j) t[j] = fetch_direct(j);
  issue_entry(D);
  cp_async_commit();
  int stage = 0;
  for (int rd = 0; rd < R; ++rd) {
    issue_entry(rd + D + 1);
    if (t[0].v >= 0 && t[0].fl == 0) issue_task<PASS>(c, t[0], stage);
    cp_async_commit();
#pragma unroll
    for (int j = 1; j < D; ++j) {
      int sj = stage + j;
      if (sj >= D) sj -= D;
      if (t[j].v >= 0 && t[j].fl == j) issue_task<PASS>(c, t[j], sj);
      cp_async_commit();
    }
    if (t[0].fl == 0) cp_async_wait<D - 1>();
    else cp_async_wait<(D == 2) ? 1 : 4>();
    __syncwarp();
    if (t[0].v >= 0) {
      if (PASS) compute_pre(c, t[0], stage);
      else compute_post(c, t[0], stage);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j + 1 < D; ++j) t[j] = t[j + 1];
    t[D - 1] = decode(rd + D);
    if (++stage == D) stage = 0;
  }
  cp_async_wait<0>();
  __syncwarp();
}

template <typename T, int K, int W, int NC, bool GRAD, int MINB, int D>
__global__ void __launch_bounds__(W * 32, MINB) traverse_kernel(const TraverseParams<T> p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int scount[NC * 32];
  __shared__ double warp_ll[W];

  Ctx<T, K, W, NC, D> c(p);
  c.lane = threadIdx.x & 31;
  c.warp = threadIdx.x >> 5;
  c.kk = (K > 0) ? K : p.n_cat;
  c.cpw = (32 / c.kk) * c.kk;
  c.k = c.lane % c.kk;
  c.wsm = smem_raw + (size_t)c.warp * warp_smem_bytes<T>(c.kk, NC, D);
  c.scount = scount;
  c.ll_acc = 0.0;
  c.gp = p.gpart + (size_t)blockIdx.x * p.n_branches;
  c.cw = (T)p.cat_w[c.k];
  c.cwr = (T)(p.cat_w[c.k] * p.rates[c.k]);
  c.first_tile = true;

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, c.first_tile = false) {
    const int col0 = tile * p.cols_per_tile + c.lane;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      const int col = col0 + j * c.cpw;
      c.act[j] = c.lane < c.cpw && j * c.cpw + c.lane < p.cols_per_tile && col < p.n_cols;
      c.w_p[j] = c.act[j] ? __ldg(p.weights + col / c.kk) : 0.0;
    }
    c.post_c = p.post + (size_t)col0 * 4;
    c.pre_c = p.pre + (size_t)col0 * 4;
    c.tips_c = p.tips + col0 / c.kk;
    if (threadIdx.x < NC * 32) scount[threadIdx.x] = 0;
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
