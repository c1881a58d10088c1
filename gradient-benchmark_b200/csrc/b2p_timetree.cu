// C ABI of the time-tree functions (include/b200phylo.h, "Time trees").  Host side: validate
// the topology, lay the internal nodes out by depth, upload; each evaluation entry point is
// one kernel launch on the caller's stream, no synchronisation.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <numeric>
#include <string>
#include <vector>

#include "../../include/b200phylo.h"
#include "b2p_timetree.cuh"

using namespace b2p;

static thread_local std::string g_tt_create_error;

struct b2p_timetree {
  int n_tips = 0, device = 0, n_sms = 0;
  double* joint = nullptr;  // scratch of b2p_joint_density_grad: heights, bl, grad_bl, 3 x grad_heights, terms
  TTDev dev{};
  size_t smem_fwd = 0, smem_bwd = 0, smem_coal = 0;
  std::vector<void*> allocs;
  mutable std::string err;
};

static int tt_fail(const b2p_timetree* t, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (t) t->err = buf; else g_tt_create_error = buf;
  return code;
}

#define TCU(call)                                                                                      \
  do {                                                                                                 \
    cudaError_t e_ = (call);                                                                           \
    if (e_ != cudaSuccess)                                                                             \
      return tt_fail(t, B2P_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

template <typename T>
static int tt_upload(b2p_timetree* t, const std::vector<T>& v, const T** out) {
  void* p = nullptr;
  const size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
  if (cudaMalloc(&p, bytes) != cudaSuccess) return tt_fail(t, B2P_ERR_ALLOC, "cudaMalloc(%zu bytes) failed", bytes);
  t->allocs.push_back(p);
  TCU(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  *out = static_cast<const T*>(p);
  return B2P_OK;
}

extern "C" {

const char* b2p_timetree_last_error(const b2p_timetree* t) { return t ? t->err.c_str() : g_tt_create_error.c_str(); }

int b2p_timetree_destroy(b2p_timetree* t) {
  if (!t) return B2P_OK;
  cudaSetDevice(t->device);
  for (void* p : t->allocs) cudaFree(p);
  delete t;
  return B2P_OK;
}

int b2p_timetree_create(int32_t n_tips, const int32_t* h_triples, const double* h_tip_heights, int32_t device,
                        b2p_timetree** out) {
  b2p_timetree* t = nullptr;
  if (!out || !h_triples || !h_tip_heights) return tt_fail(t, B2P_ERR_ARG, "b2p_timetree_create: null argument");
  *out = nullptr;
  if (n_tips < 2) return tt_fail(t, B2P_ERR_ARG, "b2p_timetree_create: need n_tips >= 2");
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
    return tt_fail(t, B2P_ERR_CUDA, "b2p_timetree_create: no CUDA device; there is no CPU fallback");
  if (device < 0 || device >= n_dev) return tt_fail(t, B2P_ERR_ARG, "b2p_timetree_create: device %d out of range", device);
  const int N = n_tips, M = N - 1, n_nodes = 2 * N - 1, root = n_nodes - 1;
  std::vector<int> parent(n_nodes, -1), kids(2 * (size_t)M), depth(n_nodes, 0);
  std::vector<double> bound_all(n_nodes);
  for (int i = 0; i < N; ++i) {
    if (!(h_tip_heights[i] >= 0.0) || h_tip_heights[i] > 1e300)
      return tt_fail(t, B2P_ERR_ARG, "b2p_timetree_create: tip height %d is negative or not finite", i);
    bound_all[i] = h_tip_heights[i];
  }
  for (int e = 0; e < M; ++e) {
    const int v = h_triples[3 * e], l = h_triples[3 * e + 1], r = h_triples[3 * e + 2];
    if (v != N + e) return tt_fail(t, B2P_ERR_ARG, "b2p_timetree_create: triples must list internal nodes N..2N-2 in order (post-order ids)");
    if (l < 0 || r < 0 || l >= v || r >= v || l == r) return tt_fail(t, B2P_ERR_ARG, "b2p_timetree_create: child id out of range / not in post-order");
    if (parent[l] >= 0 || parent[r] >= 0) return tt_fail(t, B2P_ERR_ARG, "b2p_timetree_create: a node has two parents");
    parent[l] = parent[r] = v;
    kids[2 * e] = std::min(l, r);
    kids[2 * e + 1] = std::max(l, r);
    bound_all[v] = std::max(bound_all[l], bound_all[r]);
  }
  for (int v = 0; v < root; ++v)
    if (parent[v] < 0) return tt_fail(t, B2P_ERR_ARG, "b2p_timetree_create: tree is not connected (node without parent)");
  int n_levels = 1;
  for (int v = root - 1; v >= N; --v) { depth[v] = depth[parent[v]] + 1; n_levels = std::max(n_levels, depth[v] + 1); }
  // level order
  std::vector<int> lvl_off(n_levels + 1, 0), node(M), pos(M), ppos(M), kpos(2 * (size_t)M);
  for (int v = N; v < n_nodes; ++v) lvl_off[depth[v] + 1]++;
  for (int l = 0; l < n_levels; ++l) lvl_off[l + 1] += lvl_off[l];
  {
    std::vector<int> fill(lvl_off.begin(), lvl_off.end() - 1);
    for (int v = N; v < n_nodes; ++v) { const int p = fill[depth[v]]++; node[p] = v - N; pos[v - N] = p; }
  }
  std::vector<double> bound(M), bound_l(M);
  for (int i = 0; i < M; ++i) bound[i] = bound_all[N + i];
  for (int p = 0; p < M; ++p) {
    const int i = node[p], v = N + i;
    ppos[p] = v == root ? -1 : pos[parent[v] - N];
    bound_l[p] = bound[i];
    for (int s = 0; s < 2; ++s) { const int c = kids[2 * i + s]; kpos[2 * p + s] = c >= N ? pos[c - N] : -1; }
  }
  std::vector<TTSegment> seg;
  for (int l = 0; l < n_levels; ++l) {
    const bool narrow = lvl_off[l + 1] - lvl_off[l] <= 32;
    if (narrow && !seg.empty() && seg.back().narrow) seg.back().lvl_end = l + 1;
    else seg.push_back({l, l + 1, narrow ? 1 : 0, 0});
  }
  std::vector<int> parent_all(parent.begin(), parent.begin() + (n_nodes - 1));
  std::vector<double> tip_h(h_tip_heights, h_tip_heights + N), tip_sorted(tip_h);
  std::vector<int> by(N), tip_rank(N);
  std::iota(by.begin(), by.end(), 0);
  std::stable_sort(by.begin(), by.end(), [&](int a, int b) { return tip_h[a] < tip_h[b]; });
  for (int r = 0; r < N; ++r) { tip_rank[by[r]] = r; tip_sorted[r] = tip_h[by[r]]; }

  t = new (std::nothrow) b2p_timetree;
  if (!t) return tt_fail(nullptr, B2P_ERR_ALLOC, "b2p_timetree_create: out of host memory");
  t->n_tips = N; t->device = device;
  auto bail = [&](int code) { std::string m = t->err; b2p_timetree_destroy(t); g_tt_create_error = m; return code; };
  cudaDeviceProp prop;
  if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
    tt_fail(t, B2P_ERR_CUDA, "b2p_timetree_create: cannot select device %d", device);
    return bail(B2P_ERR_CUDA);
  }
  if (prop.major != 10) {
    tt_fail(t, B2P_ERR_UNSUPPORTED, "b2p_timetree_create: built for sm_100a (B200), device is sm_%d%d", prop.major, prop.minor);
    return bail(B2P_ERR_UNSUPPORTED);
  }
  t->n_sms = prop.multiProcessorCount;
  TTDev& d = t->dev;
  d.n_tips = N; d.n_int = M; d.n_levels = n_levels; d.n_segments = (int)seg.size();
  int rc = B2P_OK;
  if ((rc = tt_upload(t, node, &d.node)) || (rc = tt_upload(t, ppos, &d.ppos)) || (rc = tt_upload(t, kpos, &d.kpos)) ||
      (rc = tt_upload(t, bound_l, &d.bound_l)) || (rc = tt_upload(t, lvl_off, &d.lvl_off)) || (rc = tt_upload(t, seg, &d.seg)) ||
      (rc = tt_upload(t, parent_all, &d.parent_all)) || (rc = tt_upload(t, kids, &d.kids)) || (rc = tt_upload(t, bound, &d.bound)) ||
      (rc = tt_upload(t, tip_h, &d.tip_h)) || (rc = tt_upload(t, tip_sorted, &d.tip_sorted)) || (rc = tt_upload(t, tip_rank, &d.tip_rank)))
    return bail(rc);
  std::vector<double> zeros(6 * (size_t)M + 8, 0.0), zev(n_nodes, 0.0);
  std::vector<unsigned> zt(1, 0u);
  const double *work = nullptr, *ev_t = nullptr;
  const unsigned* ticket = nullptr;
  if ((rc = tt_upload(t, zeros, &work)) || (rc = tt_upload(t, zev, &ev_t)) ||
      (rc = tt_upload(t, zt, &ticket)))
    return bail(rc);
  {
    std::vector<double> zj(4 * (size_t)M + 2 * (size_t)(2 * N - 2) + 8, 0.0);
    const double* jp = nullptr;
    if ((rc = tt_upload(t, zj, &jp))) return bail(rc);
    t->joint = const_cast<double*>(jp);
  }
  d.work = const_cast<double*>(work); d.ev_t = const_cast<double*>(ev_t);
  d.ticket = const_cast<unsigned*>(ticket);
  // shared-memory footprints: heights: {r, b}, h (double) + parent position + {first, count} per level;
  // vjp: a, w, r (double) + two child positions + {first, count} per level
  // (each array has one spare slot, see b2p_timetree.cuh)
  t->smem_fwd = 28 * (size_t)(M + 1) + 8 * (size_t)n_levels + 16;
  t->smem_bwd = 32 * (size_t)(M + 1) + 8 * (size_t)n_levels + 16;
  d.in_smem = t->smem_bwd <= (size_t)prop.sharedMemPerBlockOptin ? 1 : 0;
  if (d.in_smem) {
    cudaError_t e1 = cudaFuncSetAttribute(tt_heights_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)prop.sharedMemPerBlockOptin);
    cudaError_t e2 = cudaFuncSetAttribute(tt_heights_vjp_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)prop.sharedMemPerBlockOptin);
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
      tt_fail(t, B2P_ERR_CUDA, "cudaFuncSetAttribute(smem=%zu): %s (was the library built for sm_100a?)", t->smem_bwd,
              cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
      return bail(B2P_ERR_CUDA);
    }
  } else {
    t->smem_fwd = t->smem_bwd = 0;
  }
  // coalescent: internal heights + sorted tip heights per CTA, when two CTAs still fit an SM
  t->smem_coal = 8 * ((size_t)((M + 1) & ~1) + N);
  if (t->smem_coal > 100 * 1024) t->smem_coal = 0;
  // (function attributes are process-wide: always ask for the cap, never for this tree's size)
  if (cudaFuncSetAttribute(tt_coalescent_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024) != cudaSuccess) {
    tt_fail(t, B2P_ERR_CUDA, "cudaFuncSetAttribute(coalescent smem=%zu) failed", t->smem_coal);
    return bail(B2P_ERR_CUDA);
  }
  *out = t;
  return B2P_OK;
}

int b2p_timetree_get_info(const b2p_timetree* t, int32_t* n_levels, int32_t* n_segments, int32_t* in_smem) {
  if (!t) return B2P_ERR_ARG;
  if (n_levels) *n_levels = t->dev.n_levels;
  if (n_segments) *n_segments = t->dev.n_segments;
  if (in_smem) *in_smem = t->dev.in_smem;
  return B2P_OK;
}

#define TT_LAUNCHED(name)                                                                         \
  do {                                                                                            \
    cudaError_t e_ = cudaGetLastError();                                                          \
    if (e_ != cudaSuccess) return tt_fail(t, B2P_ERR_CUDA, name ": launch failed: %s", cudaGetErrorString(e_)); \
  } while (0)

int b2p_timetree_heights(b2p_timetree* t, const double* d_params, double* d_heights, void* cuda_stream) {
  if (!t || !d_params || !d_heights) return t ? tt_fail(t, B2P_ERR_ARG, "b2p_timetree_heights: null pointer") : B2P_ERR_ARG;
  TCU(cudaSetDevice(t->device));
  if (t->dev.in_smem) tt_heights_kernel<true><<<1, TT_NT, t->smem_fwd, (cudaStream_t)cuda_stream>>>(t->dev, d_params, d_heights);
  else tt_heights_kernel<false><<<1, TT_NT, 0, (cudaStream_t)cuda_stream>>>(t->dev, d_params, d_heights);
  TT_LAUNCHED("b2p_timetree_heights");
  return B2P_OK;
}

int b2p_timetree_heights_vjp(b2p_timetree* t, const double* d_params, const double* d_heights, const double* d_grad_heights,
                             double* d_grad_params, void* cuda_stream) {
  if (!t || !d_params || !d_heights || !d_grad_heights || !d_grad_params)
    return t ? tt_fail(t, B2P_ERR_ARG, "b2p_timetree_heights_vjp: null pointer") : B2P_ERR_ARG;
  TCU(cudaSetDevice(t->device));
  if (t->dev.in_smem)
    tt_heights_vjp_kernel<true><<<1, TT_NT, t->smem_bwd, (cudaStream_t)cuda_stream>>>(t->dev, d_params, d_heights, d_grad_heights, d_grad_params, TTJoin{});
  else
    tt_heights_vjp_kernel<false><<<1, TT_NT, 0, (cudaStream_t)cuda_stream>>>(t->dev, d_params, d_heights, d_grad_heights, d_grad_params, TTJoin{});
  TT_LAUNCHED("b2p_timetree_heights_vjp");
  return B2P_OK;
}

int b2p_timetree_log_jacobian(b2p_timetree* t, const double* d_heights, double* d_logj, double* d_grad_heights, void* cuda_stream) {
  if (!t || !d_heights || (!d_logj && !d_grad_heights))
    return t ? tt_fail(t, B2P_ERR_ARG, "b2p_timetree_log_jacobian: null pointer") : B2P_ERR_ARG;
  TCU(cudaSetDevice(t->device));
  tt_log_jacobian_kernel<<<1, TT_NT, 0, (cudaStream_t)cuda_stream>>>(t->dev, d_heights, d_logj, d_grad_heights);
  TT_LAUNCHED("b2p_timetree_log_jacobian");
  return B2P_OK;
}

int b2p_timetree_branch_lengths(b2p_timetree* t, const double* d_heights, const double* d_rate, double* d_bl, void* cuda_stream) {
  if (!t || !d_heights || !d_bl) return t ? tt_fail(t, B2P_ERR_ARG, "b2p_timetree_branch_lengths: null pointer") : B2P_ERR_ARG;
  TCU(cudaSetDevice(t->device));
  const int nb = 2 * t->n_tips - 2;
  tt_branch_lengths_kernel<<<(nb + 255) / 256, 256, 0, (cudaStream_t)cuda_stream>>>(t->dev, d_heights, d_rate, d_bl);
  TT_LAUNCHED("b2p_timetree_branch_lengths");
  return B2P_OK;
}

int b2p_timetree_branch_lengths_vjp(b2p_timetree* t, const double* d_heights, const double* d_rate, const double* d_grad_bl,
                                    double* d_grad_heights, double* d_grad_rate, void* cuda_stream) {
  if (!t || !d_heights || !d_grad_bl || (!d_grad_heights && !d_grad_rate))
    return t ? tt_fail(t, B2P_ERR_ARG, "b2p_timetree_branch_lengths_vjp: null pointer") : B2P_ERR_ARG;
  TCU(cudaSetDevice(t->device));
  tt_branch_lengths_vjp_kernel<<<1, TT_NT, 0, (cudaStream_t)cuda_stream>>>(t->dev, d_heights, d_rate, d_grad_bl, d_grad_heights, d_grad_rate);
  TT_LAUNCHED("b2p_timetree_branch_lengths_vjp");
  return B2P_OK;
}

int b2p_coalescent_constant(b2p_timetree* t, const double* d_heights, const double* d_theta, double* d_logp,
                            double* d_grad_heights, double* d_grad_theta, void* cuda_stream) {
  if (!t || !d_heights || !d_theta || (!d_logp && !d_grad_heights && !d_grad_theta))
    return t ? tt_fail(t, B2P_ERR_ARG, "b2p_coalescent_constant: null pointer") : B2P_ERR_ARG;
  if (reinterpret_cast<uintptr_t>(d_heights) & 15) return tt_fail(t, B2P_ERR_ARG, "b2p_coalescent_constant: d_heights must be 16-byte aligned");
  TCU(cudaSetDevice(t->device));
  const int E = 2 * t->n_tips - 1;
  const int grid = std::min((E + TTC_EV - 1) / TTC_EV, t->n_sms * (t->smem_coal > 56 * 1024 ? 2 : 4));
  if (t->smem_coal)
    tt_coalescent_kernel<true><<<grid, TTC_NT, t->smem_coal, (cudaStream_t)cuda_stream>>>(t->dev, d_heights, d_theta, d_logp,
                                                                                       d_grad_heights, d_grad_theta);
  else
    tt_coalescent_kernel<false><<<grid, TTC_NT, 0, (cudaStream_t)cuda_stream>>>(t->dev, d_heights, d_theta, d_logp,
                                                                               d_grad_heights, d_grad_theta);
  TT_LAUNCHED("b2p_coalescent_constant");
  return B2P_OK;
}

int b2p_joint_density_grad(b2p_instance* like, b2p_timetree* t, const double* d_params, const double* d_rate,
                           const double* d_theta, int32_t include_jacobian, double* d_out, double* d_grad_params,
                           double* d_grad_rate, double* d_grad_theta, void* cuda_stream) {
  if (!t || !like || !d_params || !d_out || !d_grad_params)
    return t ? tt_fail(t, B2P_ERR_ARG, "b2p_joint_density_grad: null pointer") : B2P_ERR_ARG;
  if (d_theta && !d_grad_theta) return tt_fail(t, B2P_ERR_ARG, "b2p_joint_density_grad: d_theta without d_grad_theta");
  const int N = t->n_tips, M = N - 1, B = 2 * N - 2;
  b2p_info info;
  if (b2p_get_info(like, &info) != B2P_OK || info.n_tips != N)
    return tt_fail(t, B2P_ERR_ARG, "b2p_joint_density_grad: likelihood instance and time tree differ in tip count");
  TCU(cudaSetDevice(t->device));
  cudaStream_t st = (cudaStream_t)cuda_stream;
  double* heights = t->joint;          // [M]
  double* bl = heights + M + (M & 1);  // [B]  (kept 16-byte aligned)
  double* gbl = bl + B;                // [B]
  double* gh_bl = gbl + B;             // [M]
  double* gh_j = gh_bl + M;            // [M]
  double* gh_c = gh_j + M;             // [M]
  double* scratch_rate = gh_c + M;     // [1] when the caller does not want d/d rate
  int rc;
  if ((rc = b2p_timetree_heights(t, d_params, heights, st))) return rc;
  if ((rc = b2p_timetree_branch_lengths(t, heights, d_rate, bl, st))) return rc;
  if ((rc = b2p_loglik_grad(like, bl, d_out + 1, gbl, st)))
    return tt_fail(t, rc, "b2p_joint_density_grad: likelihood: %s", b2p_last_error(like));
  if ((rc = b2p_timetree_branch_lengths_vjp(t, heights, d_rate, gbl, gh_bl, d_grad_rate ? d_grad_rate : scratch_rate, st))) return rc;
  TTJoin join{};
  join.term[0] = d_out + 1;
  if (d_theta) {
    if ((rc = b2p_coalescent_constant(t, heights, d_theta, d_out + 2, gh_c, d_grad_theta, st))) return rc;
    join.g1 = gh_c;
    join.term[1] = d_out + 2;
  }
  if (include_jacobian) {
    if ((rc = b2p_timetree_log_jacobian(t, heights, d_out + 3, gh_j, st))) return rc;
    join.g2 = gh_j;
    join.term[2] = d_out + 3;
  }
  join.total = d_out;
  if (t->dev.in_smem)
    tt_heights_vjp_kernel<true><<<1, TT_NT, t->smem_bwd, st>>>(t->dev, d_params, heights, gh_bl, d_grad_params, join);
  else
    tt_heights_vjp_kernel<false><<<1, TT_NT, 0, st>>>(t->dev, d_params, heights, gh_bl, d_grad_params, join);
  TT_LAUNCHED("b2p_joint_density_grad");
  return B2P_OK;
}

}  // extern "C"
