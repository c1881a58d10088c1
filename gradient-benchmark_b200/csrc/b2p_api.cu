// libb200phylo.so — C ABI (include/b200phylo.h) over the sm_100a kernels.
// Host side: instance state, the list schedule of the tree, launch sizing.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <queue>
#include <string>
#include <vector>

#include "../../include/b200phylo.h"
#include "b2p_kernels.cuh"
#include "b2p_plan.h"

using namespace b2p;

static thread_local std::string g_create_error;

struct b2p_instance {
  b2p_config cfg{};
  int n_sms = 0;
  int W = 0, D = 2, NC = 1, cols_per_tile = 0, n_tiles = 0, grid = 0, blocks_per_sm = 0;
  int n_rounds = 0, tree_depth = 0, regs = 0;
  size_t smem_bytes = 0;
  bool have_topology = false, have_tips = false, have_weights = false, have_eigen = false,
       have_cats = false;
  // device scratch
  SchedEntry* d_sched = nullptr;
  uint8_t* d_tips = nullptr;
  double* d_weights = nullptr;
  void* d_pmat = nullptr;
  void* d_post = nullptr;
  void* d_pre = nullptr;
  double* d_lpart = nullptr;
  double* d_gpart = nullptr;
  double* d_bl_stage = nullptr;   // for the host-buffer entry point
  double* d_out_stage = nullptr;  // [1 + 2N-2]
  double* h_pin_in = nullptr;
  double* h_pin_out = nullptr;
  cudaStream_t own_stream = nullptr;
  static constexpr int kProfRing = 1024;  // event pairs kept while profiling is on
  std::vector<cudaEvent_t> ev0, ev1;
  int64_t prof_count = 0;
  bool profiling = false;
  int64_t device_bytes = 0;
  ModelParams model{};
  double freqs[4]{};
  double qmat[16]{};
  double cat_w[kMaxCat]{};
  const void* k_grad = nullptr;  // selected traversal kernels
  const void* k_eval = nullptr;
  // wave engine (b2p_wave.cuh)
  bool wave = true;
  WaveGeometry wg;
  WaveEntry* d_wpost = nullptr;
  WaveEntry* d_wpre = nullptr;
  int w_rounds_post = 0, w_rounds_pre = 0;
  long long w_ring_post = 0, w_ring_pre = 0, w_pre_global = 0;
  const void* wk_grad = nullptr;
  const void* wk_eval = nullptr;
  mutable std::string err;
};

static int fail(const b2p_instance* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (h) h->err = buf; else g_create_error = buf;
  return code;
}

#define CU(call)                                                                       \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess)                                                             \
      return fail(h, B2P_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_),    \
                  __FILE__, __LINE__);                                                 \
  } while (0)

template <typename T>
static int dev_alloc(b2p_instance* h, T** p, size_t n_bytes) {
  cudaError_t e = cudaMalloc((void**)p, n_bytes ? n_bytes : 1);
  if (e != cudaSuccess)
    return fail(h, B2P_ERR_ALLOC, "cudaMalloc(%zu bytes): %s", n_bytes, cudaGetErrorString(e));
  h->device_bytes += (int64_t)n_bytes;
  return B2P_OK;
}

// ---------------------------------------------------------------- kernel selection
// (W warps, NC columns per lane, min CTAs/SM) variants; D = 2 pipeline stages.
template <typename T, int K, int W, int NC, int MINB>
static void pick(const void** grad, const void** eval) {
  *grad = (const void*)traverse_kernel<T, K, W, NC, true, MINB, 2>;
  *eval = (const void*)traverse_kernel<T, K, W, NC, false, MINB, 2>;
}
template <typename T, int K>
static bool select_k(int W, int NC, const void** grad, const void** eval) {
  if (W == 8 && NC == 1) pick<T, K, 8, 1, 3>(grad, eval);
  else if (W == 8 && NC == 2) pick<T, K, 8, 2, 2>(grad, eval);
  else if (W == 16 && NC == 1) pick<T, K, 16, 1, 2>(grad, eval);
  else if (W == 16 && NC == 2) pick<T, K, 16, 2, 1>(grad, eval);
  else return false;
  return true;
}
template <typename T>
static bool select_kernels(int K, int W, int NC, const void** grad, const void** eval) {
  if (K == 1) return select_k<T, 1>(W, NC, grad, eval);
  if (K == 4) return select_k<T, 4>(W, NC, grad, eval);
  if (W == 8 && NC == 1) { pick<T, 0, 8, 1, 2>(grad, eval); return true; }
  return false;
}

template <typename T, int K>
static bool wave_select_nt(int nt, const void** grad, const void** eval) {
  if (nt == 1024) { *grad = (const void*)wave_kernel<T, K, 1024, true>; *eval = (const void*)wave_kernel<T, K, 1024, false>; }
  else if (nt == 768) { *grad = (const void*)wave_kernel<T, K, 768, true>; *eval = (const void*)wave_kernel<T, K, 768, false>; }
  else if (nt == 512) { *grad = (const void*)wave_kernel<T, K, 512, true>; *eval = (const void*)wave_kernel<T, K, 512, false>; }
  else if (nt == 256) { *grad = (const void*)wave_kernel<T, K, 256, true>; *eval = (const void*)wave_kernel<T, K, 256, false>; }
  else return false;
  return true;
}
template <typename T>
static bool wave_select(int K, int nt, const void** grad, const void** eval) {
  if (K == 1) return wave_select_nt<T, 1>(nt, grad, eval);
  if (K == 4) return wave_select_nt<T, 4>(nt, grad, eval);
  return wave_select_nt<T, 0>(nt, grad, eval);
}

static size_t elem_size(const b2p_instance* h) {
  return h->cfg.precision == B2P_PRECISION_F32 ? 4 : 8;
}

extern "C" {

int b2p_version(void) { return 1; }

const char* b2p_last_error(const b2p_instance* h) {
  return h ? h->err.c_str() : g_create_error.c_str();
}

int b2p_create(const b2p_config* cfg, b2p_instance** out) {
  b2p_instance* h = nullptr;
  if (!cfg || !out) return fail(h, B2P_ERR_ARG, "b2p_create: null argument");
  *out = nullptr;
  if (cfg->n_tips < 2 || cfg->n_patterns < 1 || cfg->n_categories < 1)
    return fail(h, B2P_ERR_ARG, "b2p_create: need n_tips >= 2, n_patterns >= 1, n_categories >= 1");
  if (cfg->n_categories > kMaxCat)
    return fail(h, B2P_ERR_UNSUPPORTED, "b2p_create: n_categories > %d", kMaxCat);
  if (cfg->n_states != 4)
    return fail(h, B2P_ERR_UNSUPPORTED, "b2p_create: only 4-state models are built (got %d)",
                cfg->n_states);
  if (cfg->precision != B2P_PRECISION_F64 && cfg->precision != B2P_PRECISION_F32)
    return fail(h, B2P_ERR_ARG, "b2p_create: precision must be 64 or 32");
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    return fail(h, B2P_ERR_CUDA, "b2p_create: no CUDA device (%s); there is no CPU fallback",
                cudaGetErrorString(e));
  if (cfg->device < 0 || cfg->device >= n_dev)
    return fail(h, B2P_ERR_ARG, "b2p_create: device %d out of range (0..%d)", cfg->device, n_dev - 1);
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, cfg->device)) != cudaSuccess)
    return fail(h, B2P_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
  if (prop.major != 10)
    return fail(h, B2P_ERR_UNSUPPORTED,
                "b2p_create: device is sm_%d%d; this library is built for sm_100a only", prop.major,
                prop.minor);
  if ((e = cudaSetDevice(cfg->device)) != cudaSuccess)
    return fail(h, B2P_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));

  h = new (std::nothrow) b2p_instance();
  if (!h) return fail(nullptr, B2P_ERR_ALLOC, "b2p_create: out of host memory");
  h->cfg = *cfg;
  h->n_sms = prop.multiProcessorCount;
  const int K = cfg->n_categories, N = cfg->n_tips, P = cfg->n_patterns;
  const size_t es = elem_size(h);
  const size_t n_cols = (size_t)P * K;

  // launch shape: W warps (= schedule width), NC columns per lane.  Defaults: two
  // columns per lane once that still yields at least one tile per SM.
  const bool generic_k = !(K == 1 || K == 4);
  const int cpw = (32 / K) * K;
  int W = cfg->warps_per_block, NC = 0;
  if (const char* s = getenv("B2P_WARPS")) W = atoi(s);
  if (const char* s = getenv("B2P_NC")) NC = atoi(s);
  if (W == 0) W = 8;
  if (NC == 0) NC = ((n_cols + 2 * cpw - 1) / (2 * cpw) >= (size_t)h->n_sms) ? 2 : 1;
  if (generic_k) { W = 8; NC = 1; }
  h->W = W; h->NC = NC; h->D = 2;
  h->cols_per_tile = NC * cpw;
  if (const char* s = getenv("B2P_COLS_PER_TILE")) {
    int c = atoi(s);
    if (c >= K && c <= NC * cpw && c % K == 0) h->cols_per_tile = c;
  }
  h->n_tiles = (int)((n_cols + h->cols_per_tile - 1) / h->cols_per_tile);
  bool ok = (cfg->precision == B2P_PRECISION_F32) ? select_kernels<float>(K, W, NC, &h->k_grad, &h->k_eval)
                                                  : select_kernels<double>(K, W, NC, &h->k_grad, &h->k_eval);
  if (!ok) { delete h; return fail(nullptr, B2P_ERR_ARG, "b2p_create: no kernel for W=%d NC=%d K=%d", W, NC, K); }
  h->smem_bytes = (size_t)W * (cfg->precision == B2P_PRECISION_F32 ? warp_smem_bytes<float>(K, NC, 2)
                                                                    : warp_smem_bytes<double>(K, NC, 2));
  if (h->smem_bytes > 227 * 1024) {
    delete h;
    return fail(nullptr, B2P_ERR_UNSUPPORTED, "b2p_create: K=%d needs %zu bytes of shared memory per CTA", K, h->smem_bytes);
  }

  int rc = B2P_OK;
  auto bail = [&](int code) { std::string m = h->err; b2p_destroy(h); g_create_error = m; return code; };
  for (const void* k : {h->k_grad, h->k_eval}) {
    if ((e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)h->smem_bytes)) != cudaSuccess) {
      fail(h, B2P_ERR_CUDA, "cudaFuncSetAttribute(smem=%zu): %s", h->smem_bytes, cudaGetErrorString(e));
      return bail(B2P_ERR_CUDA);
    }
  }
  cudaFuncAttributes fa;
  if ((e = cudaFuncGetAttributes(&fa, h->k_grad)) != cudaSuccess) {
    fail(h, B2P_ERR_CUDA, "cudaFuncGetAttributes: %s (was the library built for sm_100a?)",
         cudaGetErrorString(e));
    return bail(B2P_ERR_CUDA);
  }
  h->regs = fa.numRegs;
  int bps = 0;
  if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, h->k_grad, W * 32, h->smem_bytes)) !=
          cudaSuccess || bps < 1) {
    fail(h, B2P_ERR_CUDA, "occupancy query failed: %s", cudaGetErrorString(e));
    return bail(B2P_ERR_CUDA);
  }
  if (const char* s = getenv("B2P_BLOCKS_PER_SM")) { int c = atoi(s); if (c >= 1 && c < bps) bps = c; }
  h->blocks_per_sm = bps;
  h->grid = std::min(h->n_tiles, bps * h->n_sms);

  // wave engine geometry (default); B2P_ENGINE=v3 selects the older per-warp pipeline
  h->wave = true;
  if (const char* s = getenv("B2P_ENGINE")) h->wave = strcmp(s, "v3") != 0;
  if (h->wave) {
    int nt_force = 0, ct_force = 0;
    if (const char* s = getenv("B2P_NT")) nt_force = atoi(s);
    if (const char* s = getenv("B2P_CT")) ct_force = atoi(s);
    if (!plan_geometry(N, (int)n_cols, K, (int)es, h->n_sms, (size_t)prop.sharedMemPerBlockOptin - 1024, nt_force,
                       ct_force, h->wg)) {
      fail(h, B2P_ERR_UNSUPPORTED, "b2p_create: no wave geometry fits shared memory (K=%d)", K);
      return bail(B2P_ERR_UNSUPPORTED);
    }
    ok = (cfg->precision == B2P_PRECISION_F32) ? wave_select<float>(K, h->wg.nt, &h->wk_grad, &h->wk_eval)
                                               : wave_select<double>(K, h->wg.nt, &h->wk_grad, &h->wk_eval);
    if (!ok) { fail(h, B2P_ERR_ARG, "b2p_create: no wave kernel for NT=%d", h->wg.nt); return bail(B2P_ERR_ARG); }
    for (const void* k : {h->wk_grad, h->wk_eval}) {
      if ((e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->wg.smem)) != cudaSuccess) {
        fail(h, B2P_ERR_CUDA, "cudaFuncSetAttribute(wave smem=%zu): %s", h->wg.smem, cudaGetErrorString(e));
        return bail(B2P_ERR_CUDA);
      }
    }
    if ((e = cudaFuncGetAttributes(&fa, h->wk_grad)) != cudaSuccess) {
      fail(h, B2P_ERR_CUDA, "cudaFuncGetAttributes(wave): %s", cudaGetErrorString(e));
      return bail(B2P_ERR_CUDA);
    }
    h->regs = fa.numRegs;
    h->grid = std::max(h->grid, h->wg.grid);  // sizes the per-CTA partial arrays for either engine
  }

  const size_t n_br = 2 * (size_t)N - 2;
  if ((rc = dev_alloc(h, &h->d_sched, sizeof(SchedEntry) * (size_t)(N - 1) * W))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_tips, (size_t)N * P))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_weights, sizeof(double) * P))) return bail(rc);
  if ((rc = dev_alloc(h, (char**)&h->d_pmat, n_br * K * 16 * es))) return bail(rc);
  if ((rc = dev_alloc(h, (char**)&h->d_post, (size_t)(N - 1) * n_cols * 4 * es))) return bail(rc);
  if ((rc = dev_alloc(h, (char**)&h->d_pre, (size_t)(N - 1) * n_cols * 4 * es))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_lpart, sizeof(double) * h->grid))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_gpart, sizeof(double) * (size_t)h->grid * n_br))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_bl_stage, sizeof(double) * n_br))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_out_stage, sizeof(double) * (n_br + 1)))) return bail(rc);
  if ((e = cudaMallocHost((void**)&h->h_pin_in, sizeof(double) * n_br)) != cudaSuccess ||
      (e = cudaMallocHost((void**)&h->h_pin_out, sizeof(double) * (n_br + 1))) != cudaSuccess) {
    fail(h, B2P_ERR_ALLOC, "cudaMallocHost: %s", cudaGetErrorString(e));
    return bail(B2P_ERR_ALLOC);
  }
  if ((e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking)) != cudaSuccess) {
    fail(h, B2P_ERR_CUDA, "stream creation: %s", cudaGetErrorString(e));
    return bail(B2P_ERR_CUDA);
  }
  *out = h;
  return B2P_OK;
}

int b2p_destroy(b2p_instance* h) {
  if (!h) return B2P_OK;
  cudaSetDevice(h->cfg.device);
  cudaFree(h->d_wpost); cudaFree(h->d_wpre);
  cudaFree(h->d_sched); cudaFree(h->d_tips); cudaFree(h->d_weights); cudaFree(h->d_pmat); cudaFree(h->d_post); cudaFree(h->d_pre); cudaFree(h->d_lpart);
  cudaFree(h->d_gpart); cudaFree(h->d_bl_stage); cudaFree(h->d_out_stage);
  if (h->h_pin_in) cudaFreeHost(h->h_pin_in);
  if (h->h_pin_out) cudaFreeHost(h->h_pin_out);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  for (cudaEvent_t ev : h->ev0) cudaEventDestroy(ev);
  for (cudaEvent_t ev : h->ev1) cudaEventDestroy(ev);
  delete h;
  return B2P_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ list schedule
// Unit tasks (one per internal node) with in-tree precedence on W identical slots:
// Hu's highest-level-first rule is makespan-optimal.  Level = distance from the root.
// The pre-order pass runs the same rounds in reverse (optimal for the out-tree).
static int build_schedule(int N, int P, int K, const int32_t* triples, int W, int D, std::vector<SchedEntry>& sched,
                          int* depth_out, std::string& why) {
  const int n_nodes = 2 * N - 1, root = 2 * N - 2;
  std::vector<int> left(n_nodes, -1), right(n_nodes, -1), parent(n_nodes, -1);
  std::vector<char> seen(n_nodes, 0);
  for (int e = 0; e < N - 1; ++e) {
    const int v = triples[3 * e], l = triples[3 * e + 1], r = triples[3 * e + 2];
    if (v != N + e) { why = "triples must list internal nodes N..2N-2 in order (post-order ids)"; return -1; }
    if (l < 0 || r < 0 || l >= v || r >= v || l == r) { why = "child id out of range / not in post-order"; return -1; }
    if (seen[l] || seen[r]) { why = "a node has two parents"; return -1; }
    seen[l] = seen[r] = 1;
    left[v] = l; right[v] = r; parent[l] = v; parent[r] = v;
  }
  for (int v = 0; v < root; ++v)
    if (!seen[v]) { why = "tree is not connected (node without parent)"; return -1; }
  std::vector<int> depth(n_nodes, 0);
  int max_depth = 0;
  for (int v = root - 1; v >= 0; --v) {
    depth[v] = depth[parent[v]] + 1;
    if (v >= N) max_depth = std::max(max_depth, depth[v]);
  }
  *depth_out = max_depth + 1;
  std::vector<int> pending(n_nodes, 0);
  using Item = std::pair<int, int>;  // (depth, -node): deepest first, ties by lower id
  std::priority_queue<Item> ready;
  for (int v = N; v < n_nodes; ++v) {
    pending[v] = (left[v] >= N) + (right[v] >= N);
    if (pending[v] == 0) ready.push({depth[v], -v});
  }
  sched.clear();
  std::vector<int> round_nodes;
  int done = 0;
  while (done < N - 1) {
    round_nodes.clear();
    while ((int)round_nodes.size() < W && !ready.empty()) {
      round_nodes.push_back(-ready.top().second);
      ready.pop();
    }
    if (round_nodes.empty()) { why = "schedule stalled (cycle?)"; return -1; }
    for (int s = 0; s < W; ++s) {
      if (s < (int)round_nodes.size()) {
        const int v = round_nodes[s];
        sched.push_back(SchedEntry{v, left[v], right[v], 0, 0u, 0u, 0u, 0u});
      } else {
        sched.push_back(SchedEntry{-1, -1, -1, 0, 0u, 0u, 0u, 0u});
      }
    }
    for (int v : round_nodes) {
      ++done;
      if (v != root && --pending[parent[v]] == 0) ready.push({depth[parent[v]], -parent[v]});
    }
  }
  // Prefetch distances (in rounds) for the asynchronous operand pipeline: how many rounds
  // before its own a task's inputs are final.  Post-order: after both internal children;
  // pre-order (rounds run in reverse): after the parent.
  const int R = (int)(sched.size() / W);
  std::vector<int> round_of(n_nodes, -1);
  for (int i = 0; i < (int)sched.size(); ++i)
    if (sched[i].v >= 0) round_of[sched[i].v] = i / W;
  const uint64_t n_cols = (uint64_t)P * K;
  auto offset = [&](int id) -> unsigned {
    return id >= N ? (unsigned)((uint64_t)(id - N) * n_cols) : (unsigned)((uint64_t)id * P);
  };
  for (SchedEntry& e : sched) {
    if (e.v < 0) continue;
    const int r = round_of[e.v];
    int last_child = -1;
    if (e.l >= N) last_child = std::max(last_child, round_of[e.l]);
    if (e.r >= N) last_child = std::max(last_child, round_of[e.r]);
    const int d_post = std::max(0, std::min(D - 1, r - 1 - last_child));
    int d_pre = 0;
    if (e.v != root) d_pre = std::max(0, std::min(D - 1, round_of[parent[e.v]] - r - 1));
    e.fl = d_post | (d_pre << 2);
    e.voff = offset(e.v); e.loff = offset(e.l); e.roff = offset(e.r);
  }
  return R;
}

template <typename T>
static void fill_params(const b2p_instance* h, TraverseParams<T>& p) {
  const int N = h->cfg.n_tips, K = h->cfg.n_categories, P = h->cfg.n_patterns;
  p.n_tips = N; p.n_patterns = P; p.n_cat = K; p.n_cols = P * K;
  p.cols_per_tile = h->cols_per_tile; p.n_tiles = h->n_tiles; p.n_rounds = h->n_rounds;
  p.n_branches = 2 * N - 2;
  p.sched = h->d_sched; p.tips = h->d_tips; p.weights = h->d_weights;
  p.pmat = (const T*)h->d_pmat;
  p.post = (T*)h->d_post; p.pre = (T*)h->d_pre;
  p.lpart = h->d_lpart; p.gpart = h->d_gpart;
  for (int i = 0; i < 4; ++i) p.freqs[i] = h->freqs[i];
  for (int i = 0; i < 16; ++i) p.q[i] = (T)h->qmat[i];
  for (int k = 0; k < kMaxCat; ++k) {
    p.cat_w[k] = k < K ? h->cat_w[k] : 0.0;
    p.rates[k] = k < K ? h->model.rates[k] : 0.0;
  }
}

template <typename T>
static void fill_wave_params(const b2p_instance* h, WaveParams<T>& p) {
  const int N = h->cfg.n_tips, K = h->cfg.n_categories, P = h->cfg.n_patterns;
  const WaveGeometry& g = h->wg;
  p.n_tips = N; p.n_patterns = P; p.n_cat = K; p.n_cols = P * K;
  p.ct = g.ct; p.ts = g.ts; p.cpw = g.cpw; p.wn = g.wn;
  p.n_tiles = g.n_tiles; p.rounds_post = h->w_rounds_post; p.rounds_pre = h->w_rounds_pre;
  p.n_branches = 2 * N - 2;
  p.ring_cap = g.ring_cap; p.pj = g.pj;
  p.off_ent = g.off_ent; p.off_mats = g.off_mats; p.off_stage = g.off_stage; p.off_ring = g.off_ring;
  p.off_part = g.off_part; p.off_scount = g.off_scount;
  p.post = h->d_wpost; p.pre = h->d_wpre;
  p.tips = h->d_tips; p.weights = h->d_weights; p.pmat = (const T*)h->d_pmat;
  p.msg = (T*)h->d_post; p.pre_g = (T*)h->d_pre;
  p.lpart = h->d_lpart; p.gpart = h->d_gpart;
  for (int i = 0; i < 4; ++i) p.freqs[i] = h->freqs[i];
  for (int i = 0; i < 16; ++i) p.q[i] = (T)h->qmat[i];
  for (int k = 0; k < kMaxCatW; ++k) {
    p.cat_w[k] = k < K ? h->cat_w[k] : 0.0;
    p.rates[k] = k < K ? h->model.rates[k] : 0.0;
  }
}

template <typename T>
static cudaError_t launch_eval(b2p_instance* h, const double* d_bl, double* d_logl, double* d_grad,
                               cudaStream_t st) {
  const int N = h->cfg.n_tips, K = h->cfg.n_categories;
  const int n_br = 2 * N - 2;
  const int total = n_br * K * 16;
  transition_kernel<T><<<(total + 255) / 256, 256, 0, st>>>(h->model, d_bl, n_br, K, (T*)h->d_pmat);
  const int slot = (int)(h->prof_count % b2p_instance::kProfRing);
  if (h->profiling) cudaEventRecord(h->ev0[slot], st);
  cudaError_t e;
  int n_parts;
  if (h->wave) {
    WaveParams<T> wp;
    fill_wave_params(h, wp);
    void* args[] = {&wp};
    e = cudaLaunchKernel(d_grad ? h->wk_grad : h->wk_eval, dim3(h->wg.grid), dim3(h->wg.nt), args, h->wg.smem, st);
    n_parts = h->wg.grid;
  } else {
    TraverseParams<T> p;
    fill_params(h, p);
    void* args[] = {&p};
    n_parts = std::min(h->n_tiles, h->blocks_per_sm * h->n_sms);
    e = cudaLaunchKernel(d_grad ? h->k_grad : h->k_eval, dim3(n_parts), dim3(h->W * 32), args, h->smem_bytes, st);
  }
  if (e != cudaSuccess) return e;
  if (h->profiling) { cudaEventRecord(h->ev1[slot], st); ++h->prof_count; }
  const int nb = d_grad ? n_br : 0;
  reduce_kernel<<<(nb + 31) / 32 + 1, 512, 0, st>>>(h->d_gpart, h->d_lpart, n_parts, nb, d_grad, d_logl);
  return cudaGetLastError();
}

static int check_ready(b2p_instance* h) {
  if (!h) return B2P_ERR_ARG;
  if (!(h->have_topology && h->have_tips && h->have_weights && h->have_eigen && h->have_cats))
    return fail(h, B2P_ERR_ARG,
                "evaluation before setup is complete (topology=%d tips=%d weights=%d eigen=%d categories=%d)",
                h->have_topology, h->have_tips, h->have_weights, h->have_eigen, h->have_cats);
  return B2P_OK;
}

extern "C" {

int b2p_get_info(const b2p_instance* h, b2p_info* o) {
  if (!h || !o) return B2P_ERR_ARG;
  const int64_t N = h->cfg.n_tips, P = h->cfg.n_patterns, K = h->cfg.n_categories;
  const int64_t U = P * K * 4 * (int64_t)elem_size(h);
  o->n_tips = (int)N; o->n_patterns = (int)P; o->n_categories = (int)K; o->n_states = 4;
  o->precision = h->cfg.precision;
  o->warps_per_block = h->W; o->cols_per_tile = h->cols_per_tile; o->n_tiles = h->n_tiles;
  o->grid_blocks = h->grid; o->blocks_per_sm = h->blocks_per_sm; o->n_sms = h->n_sms;
  if (h->wave) {
    o->warps_per_block = h->wg.nt / 32; o->cols_per_tile = h->wg.ct; o->n_tiles = h->wg.n_tiles;
    o->grid_blocks = h->wg.grid; o->blocks_per_sm = 1;
  }
  o->post_rounds = h->n_rounds; o->tree_depth = h->tree_depth; o->regs_per_thread = h->regs;
  o->device_bytes = h->device_bytes;
  o->alg_bytes_loglik = (2 * N - 3) * U + N * P;
  o->alg_bytes_grad = (5 * N - 9) * U + 2 * N * P;
  return B2P_OK;
}

int b2p_set_topology(b2p_instance* h, const int32_t* h_triples) {
  if (!h || !h_triples) return h ? fail(h, B2P_ERR_ARG, "b2p_set_topology: null triples") : B2P_ERR_ARG;
  std::vector<SchedEntry> sched;
  std::string why;
  int depth = 0;
  const int rounds = build_schedule(h->cfg.n_tips, h->cfg.n_patterns, h->cfg.n_categories, h_triples, h->W, h->D, sched, &depth, why);
  if (rounds < 0) return fail(h, B2P_ERR_ARG, "b2p_set_topology: %s", why.c_str());
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaMemcpy(h->d_sched, sched.data(), sched.size() * sizeof(SchedEntry), cudaMemcpyHostToDevice));
  h->n_rounds = rounds;
  h->tree_depth = depth;
  if (h->wave) {
    WaveSchedule ws;
    if (!plan_schedule(h->cfg.n_tips, h_triples, h->wg.wn, h->wg.ring_cap, ws, why))
      return fail(h, B2P_ERR_ARG, "b2p_set_topology: %s", why.c_str());
    cudaFree(h->d_wpost); cudaFree(h->d_wpre);
    h->d_wpost = h->d_wpre = nullptr;
    CU(cudaMalloc((void**)&h->d_wpost, ws.post.size() * sizeof(WaveEntry)));
    CU(cudaMalloc((void**)&h->d_wpre, ws.pre.size() * sizeof(WaveEntry)));
    CU(cudaMemcpy(h->d_wpost, ws.post.data(), ws.post.size() * sizeof(WaveEntry), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->d_wpre, ws.pre.data(), ws.pre.size() * sizeof(WaveEntry), cudaMemcpyHostToDevice));
    h->w_rounds_post = ws.rounds_post; h->w_rounds_pre = ws.rounds_pre;
    h->w_ring_post = ws.ring_reads_post; h->w_ring_pre = ws.ring_reads_pre; h->w_pre_global = ws.pre_global_blocks;
    h->n_rounds = ws.rounds_post;
  }
  h->have_topology = true;
  return B2P_OK;
}

int b2p_set_tip_states(b2p_instance* h, const uint8_t* h_masks) {
  if (!h || !h_masks) return h ? fail(h, B2P_ERR_ARG, "b2p_set_tip_states: null masks") : B2P_ERR_ARG;
  const size_t n = (size_t)h->cfg.n_tips * h->cfg.n_patterns;
  for (size_t i = 0; i < n; ++i)
    if (h_masks[i] > 15) return fail(h, B2P_ERR_ARG, "b2p_set_tip_states: mask %u > 15 at %zu", h_masks[i], i);
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaMemcpy(h->d_tips, h_masks, n, cudaMemcpyHostToDevice));
  h->have_tips = true;
  return B2P_OK;
}

int b2p_set_pattern_weights(b2p_instance* h, const double* h_weights) {
  if (!h || !h_weights) return h ? fail(h, B2P_ERR_ARG, "b2p_set_pattern_weights: null") : B2P_ERR_ARG;
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaMemcpy(h->d_weights, h_weights, sizeof(double) * h->cfg.n_patterns, cudaMemcpyHostToDevice));
  h->have_weights = true;
  return B2P_OK;
}

int b2p_set_eigen(b2p_instance* h, const double* evec, const double* ivec, const double* evals,
                  const double* freqs) {
  if (!h || !evec || !ivec || !evals || !freqs)
    return h ? fail(h, B2P_ERR_ARG, "b2p_set_eigen: null argument") : B2P_ERR_ARG;
  memcpy(h->model.evec, evec, sizeof(double) * 16);
  memcpy(h->model.ivec, ivec, sizeof(double) * 16);
  memcpy(h->model.evals, evals, sizeof(double) * 4);
  memcpy(h->freqs, freqs, sizeof(double) * 4);
  for (int i = 0; i < 4; ++i)
    for (int l = 0; l < 4; ++l) {
      double acc = 0.0;
      for (int j = 0; j < 4; ++j) acc += evec[i * 4 + j] * evals[j] * ivec[j * 4 + l];
      h->qmat[i * 4 + l] = acc;
    }
  h->have_eigen = true;
  return B2P_OK;
}

int b2p_set_categories(b2p_instance* h, const double* rates, const double* weights) {
  if (!h || !rates || !weights)
    return h ? fail(h, B2P_ERR_ARG, "b2p_set_categories: null argument") : B2P_ERR_ARG;
  for (int k = 0; k < h->cfg.n_categories; ++k) {
    h->model.rates[k] = rates[k];
    h->cat_w[k] = weights[k];
  }
  h->have_cats = true;
  return B2P_OK;
}

static int eval_common(b2p_instance* h, const double* d_bl, double* d_logl, double* d_grad,
                       void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!d_bl || !d_logl) return fail(h, B2P_ERR_ARG, "evaluation: null device pointer");
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = (h->cfg.precision == B2P_PRECISION_F32)
                      ? launch_eval<float>(h, d_bl, d_logl, d_grad, st)
                      : launch_eval<double>(h, d_bl, d_logl, d_grad, st);
  if (e != cudaSuccess) return fail(h, B2P_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
  return B2P_OK;
}

int b2p_loglik(b2p_instance* h, const double* d_bl, double* d_logl, void* stream) {
  return eval_common(h, d_bl, d_logl, nullptr, stream);
}

int b2p_loglik_grad(b2p_instance* h, const double* d_bl, double* d_logl, double* d_grad, void* stream) {
  if (h && !d_grad) return fail(h, B2P_ERR_ARG, "b2p_loglik_grad: null d_grad");
  return eval_common(h, d_bl, d_logl, d_grad, stream);
}

int b2p_loglik_grad_host(b2p_instance* h, const double* h_bl, double* h_logl, double* h_grad) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!h_bl || !h_logl) return fail(h, B2P_ERR_ARG, "b2p_loglik_grad_host: null pointer");
  const size_t n_br = 2 * (size_t)h->cfg.n_tips - 2;
  CU(cudaSetDevice(h->cfg.device));
  memcpy(h->h_pin_in, h_bl, sizeof(double) * n_br);
  CU(cudaMemcpyAsync(h->d_bl_stage, h->h_pin_in, sizeof(double) * n_br, cudaMemcpyHostToDevice,
                     h->own_stream));
  rc = eval_common(h, h->d_bl_stage, h->d_out_stage, h_grad ? h->d_out_stage + 1 : nullptr,
                   h->own_stream);
  if (rc) return rc;
  const size_t n_out = h_grad ? n_br + 1 : 1;
  CU(cudaMemcpyAsync(h->h_pin_out, h->d_out_stage, sizeof(double) * n_out, cudaMemcpyDeviceToHost,
                     h->own_stream));
  CU(cudaStreamSynchronize(h->own_stream));
  *h_logl = h->h_pin_out[0];
  if (h_grad) memcpy(h_grad, h->h_pin_out + 1, sizeof(double) * n_br);
  return B2P_OK;
}

int b2p_set_profiling(b2p_instance* h, int enabled) {
  if (!h) return B2P_ERR_ARG;
  CU(cudaSetDevice(h->cfg.device));
  if (enabled && h->ev0.empty()) {
    h->ev0.resize(b2p_instance::kProfRing, nullptr);
    h->ev1.resize(b2p_instance::kProfRing, nullptr);
    for (int i = 0; i < b2p_instance::kProfRing; ++i) {
      CU(cudaEventCreate(&h->ev0[i]));
      CU(cudaEventCreate(&h->ev1[i]));
    }
  }
  h->profiling = enabled != 0;
  h->prof_count = 0;
  return B2P_OK;
}

int b2p_last_kernel_ms(b2p_instance* h, float* ms_out) {
  if (!h || !ms_out) return B2P_ERR_ARG;
  if (h->prof_count == 0) return fail(h, B2P_ERR_ARG, "b2p_last_kernel_ms: no profiled evaluation yet");
  const int slot = (int)((h->prof_count - 1) % b2p_instance::kProfRing);
  CU(cudaEventSynchronize(h->ev1[slot]));
  CU(cudaEventElapsedTime(ms_out, h->ev0[slot], h->ev1[slot]));
  return B2P_OK;
}

int b2p_kernel_ms_stats(b2p_instance* h, double* total_ms, int32_t* n_launches) {
  if (!h || !total_ms || !n_launches) return B2P_ERR_ARG;
  const int64_t n = std::min<int64_t>(h->prof_count, b2p_instance::kProfRing);
  double tot = 0.0;
  for (int64_t i = 0; i < n; ++i) {
    const int slot = (int)((h->prof_count - 1 - i) % b2p_instance::kProfRing);
    float ms = 0.f;
    CU(cudaEventSynchronize(h->ev1[slot]));
    CU(cudaEventElapsedTime(&ms, h->ev0[slot], h->ev1[slot]));
    tot += ms;
  }
  *total_ms = tot;
  *n_launches = (int32_t)n;
  return B2P_OK;
}

int b2p_get_post_partials(b2p_instance* h, int32_t node, double* h_out) {
  if (!h || !h_out) return B2P_ERR_ARG;
  const int N = h->cfg.n_tips;
  if (node < N || node > 2 * N - 2) return fail(h, B2P_ERR_ARG, "b2p_get_post_partials: node %d is not internal", node);
  const size_t n = (size_t)h->cfg.n_patterns * h->cfg.n_categories * 4;
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaDeviceSynchronize());
  if (h->cfg.precision == B2P_PRECISION_F64) {
    CU(cudaMemcpy(h_out, (const double*)h->d_post + (size_t)(node - N) * n, n * 8, cudaMemcpyDeviceToHost));
  } else {
    std::vector<float> tmp(n);
    CU(cudaMemcpy(tmp.data(), (const float*)h->d_post + (size_t)(node - N) * n, n * 4, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; ++i) h_out[i] = tmp[i];
  }
  return B2P_OK;
}

}  // extern "C"
