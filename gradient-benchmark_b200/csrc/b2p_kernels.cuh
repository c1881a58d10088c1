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

// Stage `n_mats` (2 = P_l,P_r ; 4 = + dP_l,dP_r) K x 16 matrices of one task into the
// warp's shared-memory region, layout [which][k][stride].
template <typename T>
__device__ __forceinline__ void stage_matrices(T* sm, const TraverseParams<T>& p, int l, int r,
                                               int n_mats, int lane) {
  constexpr int MS = MatStride<T>::v;
  const int kk16 = p.n_cat * 16;
  const int total = n_mats * kk16;
  for (int idx = lane; idx < total; idx += 32) {
    const int which = idx / kk16, rem = idx - which * kk16;
    const T* base = (which & 2) ? p.dpmat : p.pmat;
    const int br = (which & 1) ? r : l;
    const T v = __ldg(base + (size_t)br * kk16 + rem);
    sm[(which * p.n_cat + (rem >> 4)) * MS + (rem & 15)] = v;
  }
  __syncwarp();
}

template <typename T, int K, int W, bool GRAD, int MINB>
__global__ void __launch_bounds__(W * 32, MINB) traverse_kernel(const TraverseParams<T> p) {
  constexpr int MS = MatStride<T>::v;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int scount[32];
  __shared__ double warp_ll[W];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int kk = (K > 0) ? K : p.n_cat;
  const int n_mats = GRAD ? 4 : 2;
  T* sm = reinterpret_cast<T*>(smem_raw) + (size_t)warp * n_mats * kk * MS;
  const int N = p.n_tips, root = 2 * N - 2;
  const int k = lane % kk;
  const T* smk = sm + k * MS;  // this lane's category inside each staged matrix block
  const int blk = kk * MS;     // stride between staged matrices
  double ll_acc = 0.0;
  double* gp = p.gpart + (size_t)blockIdx.x * p.n_branches;
  bool first_tile = true;

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, first_tile = false) {
    const int col = tile * p.cols_per_tile + lane;
    const bool active = lane < p.cols_per_tile && col < p.n_cols;
    const int pat = active ? col / kk : 0;
    const double w_p = active ? __ldg(p.weights + pat) : 0.0;
    const T cw = (T)p.cat_w[k];
    if (threadIdx.x < 32) scount[threadIdx.x] = 0;
    __syncthreads();

    // ================= post-order (row a3) + root log-likelihood (row a4) ==========
    for (int rd = 0; rd < p.n_rounds; ++rd) {
      const int4 e = __ldg(p.sched + rd * W + warp);
      if (e.x >= 0) {
        stage_matrices(sm, p, e.y, e.z, 2, lane);
        T xl[4] = {0, 0, 0, 0}, xr[4] = {0, 0, 0, 0};
        if (active) {
          if (e.y < N) tip_col<T>(__ldg(p.tips + (size_t)e.y * p.n_patterns + pat), xl);
          else ld_col(p.post + ((size_t)(e.y - N) * p.n_cols + col) * 4, xl);
          if (e.z < N) tip_col<T>(__ldg(p.tips + (size_t)e.z * p.n_patterns + pat), xr);
          else ld_col(p.post + ((size_t)(e.z - N) * p.n_cols + col) * 4, xr);
        }
        T a[4], b[4], y[4];
        matvec(smk, xl, a);
        matvec(smk + blk, xr, b);
#pragma unroll
        for (int i = 0; i < 4; ++i) y[i] = a[i] * b[i];
        T mx = fmax(fmax(y[0], y[1]), fmax(y[2], y[3]));
        mx = group_max<K>(mx, lane, kk);
        int bumped = 0;
        if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
          for (int i = 0; i < 4; ++i) y[i] *= Scale<T>::factor();
          bumped = 1;
        }
        if (e.x != root) {
          if (bumped && k == 0) atomicAdd(&scount[lane], 1);
          if (active) st_col(p.post + ((size_t)(e.x - N) * p.n_cols + col) * 4, y);
        } else {
          // the root is alone in the last round: every other node's count is final
          if (active) st_col(p.post + ((size_t)(e.x - N) * p.n_cols + col) * 4, y);
          T s = (T)p.freqs[0] * y[0] + (T)p.freqs[1] * y[1] + (T)p.freqs[2] * y[2] +
                (T)p.freqs[3] * y[3];
          s = group_sum<K>(s * cw, lane, kk);
          double term = 0.0;
          if (active && k == 0)
            term = w_p * (log((double)s) - (double)(scount[lane] + bumped) * Scale<T>::log_factor());
          term = warp_sum(term);
          ll_acc += term;
        }
      }
      __syncthreads();
    }

    // ================= pre-order (row a5) + branch gradient (row a6) ================
    if (GRAD) {
      for (int rd = p.n_rounds - 1; rd >= 0; --rd) {
        const int4 e = __ldg(p.sched + rd * W + warp);
        if (e.x >= 0) {
          const int v = e.x, l = e.y, r = e.z;
          stage_matrices(sm, p, l, r, 4, lane);
          T u[4], xl[4] = {0, 0, 0, 0}, xr[4] = {0, 0, 0, 0};
#pragma unroll
          for (int i = 0; i < 4; ++i) u[i] = (T)p.freqs[i];
          if (active) {
            if (v != root) ld_col(p.pre + ((size_t)(v - N) * p.n_cols + col) * 4, u);
            if (l < N) tip_col<T>(__ldg(p.tips + (size_t)l * p.n_patterns + pat), xl);
            else ld_col(p.post + ((size_t)(l - N) * p.n_cols + col) * 4, xl);
            if (r < N) tip_col<T>(__ldg(p.tips + (size_t)r * p.n_patterns + pat), xr);
            else ld_col(p.post + ((size_t)(r - N) * p.n_cols + col) * 4, xr);
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
          den = group_sum<K>(den * cw, lane, kk);
          numl = group_sum<K>(numl * cw, lane, kk);
          numr = group_sum<K>(numr * cw, lane, kk);
          double gl = 0.0, gr = 0.0;
          if (active && k == 0) {
            const double inv = w_p / (double)den;
            gl = (double)numl * inv;
            gr = (double)numr * inv;
          }
          gl = warp_sum(gl);
          gr = warp_sum(gr);
          if (lane == 0) {
            if (first_tile) { gp[l] = gl; gp[r] = gr; }
            else { gp[l] += gl; gp[r] += gr; }
          }
          if (l >= N) {
            T y[4];
            matvec_t(smk, ql, y);
            T mx = group_max<K>(fmax(fmax(y[0], y[1]), fmax(y[2], y[3])), lane, kk);
            if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
              for (int i = 0; i < 4; ++i) y[i] *= Scale<T>::factor();
            }
            if (active) st_col(p.pre + ((size_t)(l - N) * p.n_cols + col) * 4, y);
          }
          if (r >= N) {
            T y[4];
            matvec_t(smk + blk, qr, y);
            T mx = group_max<K>(fmax(fmax(y[0], y[1]), fmax(y[2], y[3])), lane, kk);
            if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
              for (int i = 0; i < 4; ++i) y[i] *= Scale<T>::factor();
            }
            if (active) st_col(p.pre + ((size_t)(r - N) * p.n_cols + col) * 4, y);
          }
        }
        __syncthreads();
      }
    }
  }

  if (lane == 0) warp_ll[warp] = ll_acc;
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
