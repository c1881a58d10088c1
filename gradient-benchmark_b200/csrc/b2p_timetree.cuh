// Time-tree kernels (sm_100a): node heights from ratios / root height, the log-Jacobian of
// that map, branch lengths from heights, the constant-size coalescent — values and vector-
// Jacobian products.  SURVEY.md §8(f) rows 1-2; the functions the reference's micro benchmark
// times next to the tree likelihood (/root/reference/plot.Rmd:49-60).
//
// These are O(N) functions of a few thousand doubles: latency-bound, not bandwidth-bound.
// What matters is (i) one launch per function, no host round trip, so they chain with the
// likelihood kernels on one stream, and (ii) a short dependent chain.  The height recursion
// h_i = b_i + r_i (h_parent - b_i) depends on the parent, so nodes are laid out by depth
// ("level order", parents before children) and a single CTA walks the levels with all of a
// node's operands staged in shared memory.  Real time trees are ladder-like: most levels
// hold a handful of nodes.  Runs of levels no wider than a warp are walked by warp 0 alone
// with __syncwarp between levels (one __syncthreads per run instead of one per level).
//
// Arithmetic uses the non-contracting intrinsics so results match the CPU restatement
// (oracle/timetree_oracle.py) bit for bit where the summation order is the same.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2p {

constexpr int TT_NT = 512;  // threads of the single-CTA kernels

// One run of levels.  narrow: every level in [lvl_begin, lvl_end) has <= 32 nodes.
struct TTSegment { int lvl_begin, lvl_end, narrow, pad; };

struct TTDev {
  int n_tips, n_int;        // N, M = N-1
  int n_levels, n_segments;
  int in_smem;              // level arrays fit shared memory
  // level order (position idx = 0..M-1, root at 0)
  const int* node;          // [M] internal index i = node id - N
  const int* ppos;          // [M] position of the parent (-1: root)
  const int* kpos;          // [M][2] positions of the internal children, lower node id first (-1: tip)
  const double* bound_l;    // [M] b_i in level order
  const int* lvl_off;       // [n_levels + 1]
  const TTSegment* seg;     // [n_segments]
  // node order
  const int* parent_all;    // [2N-2] node id of the parent
  const int* kids;          // [M][2] node ids of the children, lower id first
  const double* bound;      // [M]
  const double* tip_h;      // [N]
  const double* tip_sorted; // [N] ascending
  const int* tip_rank;      // [N] rank of tip in (height, id) order
  // scratch
  double* work;             // [6 M + 8] fallback for the level arrays when they do not fit shared memory
  double* ev_t;             // [2N-1] coalescent: per-CTA partial sums
  unsigned* ticket;         // coalescent: CTA completion counter
};

__device__ __forceinline__ double tt_block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  const int nw = (blockDim.x + 31) >> 5;
  for (int w = 0; w < nw; ++w) s += red[w];  // every thread, same order
  return s;
}

// ------------------------------------------------------------------------- node heights
// params [M]: ratios of nodes N..2N-3, root height last.  heights [M] in node order.
// Extra inputs of the adjoint kernel when it closes a joint-density evaluation
// (b2p_joint_density_grad): more upstream gradients w.r.t. the heights, and the scalar terms to add up.
struct TTJoin {
  const double* g1;
  const double* g2;
  const double* term[3];
  double* total;
};

// SMEM: the level arrays live in shared memory (a compile-time fact, so the walk uses
// LDS/STS with 32-bit addresses rather than generic loads); otherwise in global scratch.
//
// A run of narrow levels is one warp's serial chain, and with a single warp issuing, its
// speed is the instruction count per level.  So the loop body is kept to the chain itself:
// every array has one spare slot M (r = b = 0, its own parent, no children) that idle lanes
// are pointed at instead of being predicated off, the root (always alone on level 0) is
// handled before the loop, and the next level's operands are fetched ahead of the chain.
template <bool SMEM>
__global__ void __launch_bounds__(TT_NT) tt_heights_kernel(const TTDev t, const double* __restrict__ params,
                                                          double* __restrict__ heights) {
  extern __shared__ __align__(16) unsigned char tt_sm[];
  const int M = t.n_int, tid = threadIdx.x;
  double2* rb = SMEM ? reinterpret_cast<double2*>(tt_sm) : reinterpret_cast<double2*>(t.work);  // {r, b}
  double* h = reinterpret_cast<double*>(rb + M + 1);
  int2* lv = reinterpret_cast<int2*>(h + M + 1);  // {first position, node count} per level
  int* pp = reinterpret_cast<int*>(lv + t.n_levels);
  for (int i = tid; i < M; i += TT_NT) {
    rb[i] = make_double2(params[t.node[i]], t.bound_l[i]);
    pp[i] = t.ppos[i];
  }
  for (int i = tid; i < t.n_levels; i += TT_NT) { const int o = t.lvl_off[i]; lv[i] = make_int2(o, t.lvl_off[i + 1] - o); }
  if (tid == 0) { rb[M] = make_double2(0.0, 0.0); pp[M] = M; h[M] = 0.0; }
  __syncthreads();
  if (tid == 0) h[0] = rb[0].x;  // the root's "ratio" slot carries the root height
  __syncthreads();
  TTSegment nxt = t.seg[0];
  for (int s = 0; s < t.n_segments; ++s) {
    const TTSegment sg = nxt;
    if (s + 1 < t.n_segments) nxt = t.seg[s + 1];  // in flight while this run is walked
    const int lb = sg.lvl_begin == 0 ? 1 : sg.lvl_begin;
    if (sg.narrow) {
      if (tid < 32 && lb < sg.lvl_end) {
        int2 v = lv[lb];
        int idx = tid < v.y ? v.x + tid : M;
        double2 q = rb[idx];
        int pc = pp[idx];
        for (int l = lb; l < sg.lvl_end; ++l) {
          const int cur = idx;
          const double2 c = q;
          const int pcur = pc;
          if (l + 1 < sg.lvl_end) {  // operands of the next level: independent of this level's result
            v = lv[l + 1];
            idx = tid < v.y ? v.x + tid : M;
            q = rb[idx];
            pc = pp[idx];
          }
          h[cur] = __dadd_rn(c.y, __dmul_rn(c.x, __dsub_rn(h[pcur], c.y)));
          __syncwarp();
        }
      }
    } else {
      const int2 v = lv[lb];
      for (int idx = v.x + tid; idx < v.x + v.y; idx += TT_NT) {
        const double2 c = rb[idx];
        h[idx] = __dadd_rn(c.y, __dmul_rn(c.x, __dsub_rn(h[pp[idx]], c.y)));
      }
    }
    __syncthreads();
  }
  for (int i = tid; i < M; i += TT_NT) heights[t.node[i]] = h[i];
}

// J^T g: grad_params[i] = a_i (h_parent - b_i), a_i = g_i + sum over internal children r_c a_c.
template <bool SMEM>
__global__ void __launch_bounds__(TT_NT) tt_heights_vjp_kernel(const TTDev t, const double* __restrict__ params,
                                                              const double* __restrict__ heights,
                                                              const double* __restrict__ grad_heights,
                                                              double* __restrict__ grad_params, const TTJoin join) {
  extern __shared__ __align__(16) unsigned char tt_sm[];
  const int M = t.n_int, tid = threadIdx.x;
  double* a = SMEM ? reinterpret_cast<double*>(tt_sm) : t.work;
  double* w = a + M + 1;  // r_i a_i; slot M stays 0 and stands in for a tip child
  double* r = w + M + 1;
  int2* kp = reinterpret_cast<int2*>(r + M + 1);
  int2* lv = kp + M + 1;
  for (int i = tid; i < M; i += TT_NT) {
    const int nd = t.node[i];
    double g = grad_heights[nd];
    if (join.g1) g += join.g1[nd];
    if (join.g2) g += join.g2[nd];
    a[i] = g;
    r[i] = t.ppos[i] < 0 ? 0.0 : params[nd];
    int2 k = reinterpret_cast<const int2*>(t.kpos)[i];
    if (k.x < 0) k.x = M;
    if (k.y < 0) k.y = M;
    kp[i] = k;
  }
  for (int i = tid; i < t.n_levels; i += TT_NT) { const int o = t.lvl_off[i]; lv[i] = make_int2(o, t.lvl_off[i + 1] - o); }
  if (tid == 0) {
    a[M] = 0.0; w[M] = 0.0; r[M] = 0.0; kp[M] = make_int2(M, M);
    if (join.total) {  // joint density: sum of the terms the earlier launches on this stream left behind
      double s = 0.0;
      for (int k = 0; k < 3; ++k) if (join.term[k]) s += *join.term[k];
      *join.total = s;
    }
  }
  __syncthreads();
  TTSegment nxt = t.seg[t.n_segments - 1];
  for (int s = t.n_segments - 1; s >= 0; --s) {
    const TTSegment sg = nxt;
    if (s > 0) nxt = t.seg[s - 1];
    if (sg.narrow) {
      if (tid < 32) {
        int2 v = lv[sg.lvl_end - 1];
        int idx = tid < v.y ? v.x + tid : M;
        int2 k = kp[idx];
        double ai = a[idx], ri = r[idx];
        for (int l = sg.lvl_end - 1; l >= sg.lvl_begin; --l) {
          const int cur = idx;
          const int2 kc = k;
          const double ac = ai, rc = ri;
          if (l > sg.lvl_begin) {
            v = lv[l - 1];
            idx = tid < v.y ? v.x + tid : M;
            k = kp[idx]; ai = a[idx]; ri = r[idx];
          }
          const double sum = __dadd_rn(__dadd_rn(ac, w[kc.x]), w[kc.y]);
          a[cur] = sum;
          w[cur] = __dmul_rn(rc, sum);
          __syncwarp();
        }
      }
    } else {
      const int2 v = lv[sg.lvl_begin];
      for (int idx = v.x + tid; idx < v.x + v.y; idx += TT_NT) {
        const int2 k = kp[idx];
        const double sum = __dadd_rn(__dadd_rn(a[idx], w[k.x]), w[k.y]);
        a[idx] = sum;
        w[idx] = __dmul_rn(r[idx], sum);
      }
    }
    __syncthreads();
  }
  for (int i = tid; i < M; i += TT_NT) {
    const int nd = t.node[i], pp = t.ppos[i];
    grad_params[nd] = pp < 0 ? a[i] : __dmul_rn(a[i], __dsub_rn(heights[t.node[pp]], t.bound_l[i]));
  }
}

// ------------------------------------------------------------------------ log-Jacobian
// logj = sum_{i != root} log(h_parent(i) - b_i);  grad[j] = sum_{internal children c} 1/(h_j - b_c)
__global__ void __launch_bounds__(TT_NT) tt_log_jacobian_kernel(const TTDev t, const double* __restrict__ heights,
                                                               double* __restrict__ logj, double* __restrict__ grad) {
  __shared__ double red[TT_NT / 32];
  const int M = t.n_int, N = t.n_tips;
  double s = 0.0;
  for (int i = threadIdx.x; i < M; i += TT_NT) {
    if (i < M - 1) s += log(heights[t.parent_all[N + i] - N] - t.bound[i]);
    if (grad) {
      const int2 k = reinterpret_cast<const int2*>(t.kids)[i];
      const double hj = heights[i];
      double g = 0.0;
      if (k.x >= N) g = __dadd_rn(g, __drcp_rn(__dsub_rn(hj, t.bound[k.x - N])));
      if (k.y >= N) g = __dadd_rn(g, __drcp_rn(__dsub_rn(hj, t.bound[k.y - N])));
      grad[i] = g;
    }
  }
  s = tt_block_sum(s, red);
  if (threadIdx.x == 0 && logj) *logj = s;
}

// ---------------------------------------------------------------------- branch lengths
// bl[c] = rate (h_parent(c) - h_c); rate may be null (1).
__global__ void tt_branch_lengths_kernel(const TTDev t, const double* __restrict__ heights,
                                         const double* __restrict__ rate, double* __restrict__ bl) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, N = t.n_tips;
  if (c >= 2 * N - 2) return;
  const double rt = rate ? *rate : 1.0;
  const double hc = c < N ? t.tip_h[c] : heights[c - N];
  bl[c] = __dmul_rn(rt, __dsub_rn(heights[t.parent_all[c] - N], hc));
}

// grad_heights[j] = rate (g[k0] + g[k1] - g[j]);  grad_rate = sum_c g[c] (h_parent - h_c)
__global__ void __launch_bounds__(TT_NT) tt_branch_lengths_vjp_kernel(const TTDev t, const double* __restrict__ heights,
                                                                     const double* __restrict__ rate,
                                                                     const double* __restrict__ grad_bl,
                                                                     double* __restrict__ grad_heights,
                                                                     double* __restrict__ grad_rate) {
  __shared__ double red[TT_NT / 32];
  const int M = t.n_int, N = t.n_tips;
  const double rt = rate ? *rate : 1.0;
  double s = 0.0;
  for (int i = threadIdx.x; i < M; i += TT_NT) {
    const int2 k = reinterpret_cast<const int2*>(t.kids)[i];
    const double hj = heights[i];
    const double g0 = grad_bl[k.x], g1 = grad_bl[k.y];
    double g = __dadd_rn(__dmul_rn(rt, g0), __dmul_rn(rt, g1));
    if (i < M - 1) g = __dsub_rn(g, __dmul_rn(rt, grad_bl[N + i]));
    if (grad_heights) grad_heights[i] = g;
    const double h0 = k.x < N ? t.tip_h[k.x] : heights[k.x - N];
    const double h1 = k.y < N ? t.tip_h[k.y] : heights[k.y - N];
    s += g0 * (hj - h0) + g1 * (hj - h1);
  }
  s = tt_block_sum(s, red);
  if (threadIdx.x == 0 && grad_rate) *grad_rate = s;
}

// ------------------------------------------------------------------ constant coalescent
// logp = -(1/theta) S - (N-1) log theta,  S = sum over intervals of C(k,2) x length.
// Summed by parts, S = sum over events of t_e (C(k before e) - C(k after e)): a merge of k
// lineages contributes t_e (k-1), a tip that joins k lineages -t_e k.  So every event only
// needs k_e = (tips ordered before it) - (internal nodes ordered before it), its rank in the
// order (height, tips before internal nodes, id) — found by counting, no sort, no neighbours:
// 8 lanes per event share the scan over the internal heights (staged in shared memory once
// per CTA: all events of a CTA scan the same array, so reading it from L2 in lockstep would
// leave one load in flight per CTA), the tips-before count is a binary search in the tip
// heights sorted at set-up.  CTAs leave partial sums; the last one to finish adds them in
// index order.  grad_heights[e] = -(k_e - 1)/theta, grad_theta = S/theta^2 - (N-1)/theta.
constexpr int TTC_NT = 256, TTC_LANES = 8, TTC_EV = TTC_NT / TTC_LANES;

template <bool SMEM>
__global__ void __launch_bounds__(TTC_NT) tt_coalescent_kernel(const TTDev t, const double* __restrict__ heights,
                                                              const double* __restrict__ theta_p,
                                                              double* __restrict__ logp, double* __restrict__ grad_heights,
                                                              double* __restrict__ grad_theta) {
  extern __shared__ __align__(16) unsigned char tt_sm[];
  __shared__ double red[TTC_NT / 32];
  __shared__ unsigned last;
  const int N = t.n_tips, M = t.n_int, E = 2 * N - 1;
  const int Mp = (M + 1) & ~1;
  const double* hs = heights;
  const double* ts = t.tip_sorted;
  if (SMEM) {
    double* sh = reinterpret_cast<double*>(tt_sm);
    for (int i = threadIdx.x; i < M; i += TTC_NT) sh[i] = heights[i];
    for (int i = threadIdx.x; i < N; i += TTC_NT) sh[Mp + i] = t.tip_sorted[i];
    hs = sh;
    ts = sh + Mp;
    __syncthreads();
  }
  const int sub = threadIdx.x & (TTC_LANES - 1);
  const double theta = *theta_p;
  const double2* h2 = reinterpret_cast<const double2*>(hs);
  const int half = M >> 1;
  double part = 0.0;
  for (int e0 = blockIdx.x * TTC_EV; e0 < E; e0 += gridDim.x * TTC_EV) {  // whole CTA iterates together
    const int e = e0 + threadIdx.x / TTC_LANES;
    const bool live = e < E, internal = e >= N;
    const double te = !live ? 0.0 : internal ? hs[e - N] : t.tip_h[e];
    const int self = e - N;
    // internal nodes strictly below, and at or below: a tip sorts before every internal node of its
    // height, so for a tip "below" is the strict count; an internal node counts itself in `le`
    // and, in the rare tie with another internal node, falls back to the id order.
    int lt = 0, le = 0;
    if (live) {
#pragma unroll 8
      for (int j2 = sub; j2 < half; j2 += TTC_LANES) {
        const double2 hv = h2[j2];
        lt += (hv.x < te) + (hv.y < te);
        le += (hv.x <= te) + (hv.y <= te);
      }
      if ((M & 1) && sub == 0) {
        const double hj = hs[M - 1];
        lt += hj < te;
        le += hj <= te;
      }
    }
#pragma unroll
    for (int o = TTC_LANES / 2; o > 0; o >>= 1) {
      lt += __shfl_xor_sync(0xffffffffu, lt, o);
      le += __shfl_xor_sync(0xffffffffu, le, o);
    }
    // other internal nodes at exactly this height (rare): lower id first.  Warp-uniform branch
    // so the shuffles stay converged; groups without a tie add nothing.
    const bool tied = live && internal && le - lt > 1;
    if (__any_sync(0xffffffffu, tied)) {
      int extra = 0;
      if (tied)
        for (int j = sub; j < self; j += TTC_LANES) extra += hs[j] == te;
#pragma unroll
      for (int o = TTC_LANES / 2; o > 0; o >>= 1) extra += __shfl_xor_sync(0xffffffffu, extra, o);
      lt += extra;
    }
    if (live && sub == 0) {
      if (internal) {
        const int below = lt;
        int lo = 0, hi = N;  // tips with height <= te
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (ts[mid] <= te) lo = mid + 1; else hi = mid; }
        const int k = lo - below;  // lineages just before the merge
        part += te * (double)(k - 1);
        if (grad_heights) grad_heights[self] = -(double)(k - 1) / theta;
      } else {
        part -= te * (double)(t.tip_rank[e] - lt);
      }
    }
  }
  part = tt_block_sum(part, red);
  if (threadIdx.x == 0) {
    t.ev_t[blockIdx.x] = part;
    __threadfence();
    last = atomicAdd(t.ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  double s = 0.0;
  for (int j = threadIdx.x; j < (int)gridDim.x; j += TTC_NT) s += __ldcg(t.ev_t + j);
  s = tt_block_sum(s, red);
  if (threadIdx.x == 0) {
    if (logp) *logp = -s / theta - (double)(N - 1) * log(theta);
    if (grad_theta) *grad_theta = s / (theta * theta) - (double)(N - 1) / theta;
    *t.ticket = 0u;  // ready for the next launch on this stream
  }
}

}  // namespace b2p
