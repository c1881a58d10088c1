// libb200phylo.so — C ABI (include/b200phylo.h) over the sm_100a kernels.
// Host side: instance state, the wave planner (b2p_plan.h), launch sizing.
#include <algorithm>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200phylo.h"
#include "b2p_aux.cuh"
#include "b2p_plan.h"
#include "b2p_walk.cuh"
#include "b2p_dense.cuh"

using namespace b2p;

static thread_local std::string g_create_error;

struct b2p_instance {
  b2p_config cfg{};
  int n_sms = 0, regs = 0, ctas_per_sm = 1;
  bool have_topology = false, have_tips = false, have_weights = false, have_eigen = false,
       have_cats = false;
  // geometry + plan of the traversal kernel
  WalkGeometry wg;
  WalkPlan plan;  // records are released after upload; the statistics stay
  const void* kernel = nullptr;
  const void* kernel_outer = nullptr;  // B2P_FLAG_OUTER_PRODUCTS: the same walk, also summing the branches' outer products
  double* d_opart = nullptr;           // [grid][2N-2][K][16]
  double* d_mgpart = nullptr;          // b2p_model_gradients: [K][blocks][17] partial sums
  // device scratch
  unsigned char* d_records = nullptr;  // round records: descriptors | issue record | matrix rows (K1 writes these)
  int* d_mat_dst = nullptr;
  uint8_t* d_tips = nullptr;      // [tile][2N][ct_pad] masks per column, tip rows in schedule order
  uint8_t* d_tips_raw = nullptr;  // [N][P] as given by the caller
  int* d_tip_order = nullptr;     // [2N] tip id of each scheduled row
  double* d_weights = nullptr;
  void* d_msg = nullptr;   // [tile][N-1] blocks: messages m_v = P_v x_v
  void* d_pre = nullptr;   // [tile][N-1] blocks: pre-order partials that left shared memory
  double* d_lpart = nullptr;
  double* d_gpart = nullptr;
  long long* d_trace = nullptr;  // B2P_TRACE=<file>: per-round clock stamps of CTA 0 (debug)
  double* d_bl_stage = nullptr;   // for the host-buffer entry point
  double* d_out_stage = nullptr;  // [1 + 2N-2]
  // multi-GPU: peer-mapped receive buffers of the fused reduce + all-reduce (b2p_peer_*)
  struct Peer {
    int rank = -1, world = 0;
    unsigned char* own = nullptr;            // [2][world][2N-1] doubles, then [2][world][slices] flags, ticket, err, epoch
    void* mapped[kMaxPeers] = {};            // peers' buffers opened through CUDA IPC (own entry: own)
    bool attached = false;
  } peer;
  double* h_pin_in = nullptr;     // pinned + mapped: the kernels of the host entry point read / write these directly
  double* h_pin_out = nullptr;
  // the host entry point's launches, captured once per (with / without gradient) and replayed;
  // dropped whenever anything a launch argument depends on changes (graph_drop)
  cudaGraphExec_t graph_exec[2] = {nullptr, nullptr};
  bool use_graph = true;
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
  // b2p_set_gradient_operator: what the pre-order pass contracts instead of (Q, rates)
  bool grad_op_set = false;
  double grad_op[16]{};
  double grad_scale[kMaxCat]{};
  // dense-state path (5..64 states, b2p_dense.cuh): null for 4-state instances
  struct Dense {
    int S = 0, cols_pad = 0, n_tiles = 0;
    double *evec = nullptr, *ivec = nullptr, *evals = nullptr, *q = nullptr, *qt = nullptr, *freqs = nullptr;
    double *P = nullptr, *PT = nullptr, *D = nullptr, *weights = nullptr, *msg = nullptr, *pre = nullptr;
    uint8_t* codes = nullptr;
    int* scount = nullptr;                      // [cols_pad] rescale counts, then the dependency flags and the error word
    size_t sync_ints = 0;                       // ints cleared at the start of every evaluation (scount + flags + err)
    DenseNode* nodes = nullptr;                 // post-order levels, then pre-order levels
    std::vector<int> post_off, pre_off;         // level boundaries into `nodes`
    int root = 0, tip_off = 0;                  // tip tasks start at nodes[tip_off] (N of them)
  }* dense = nullptr;
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

static void graph_drop(b2p_instance* h) {
  for (cudaGraphExec_t& g : h->graph_exec)
    if (g) { cudaGraphExecDestroy(g); g = nullptr; }
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
template <typename T, int K>
static const void* walk_select_nt(int nt, int cpt) {
  if (cpt == 2) {
    switch (nt) {
      case 768: return (const void*)walk_kernel<T, K, 768, 2>;
      case 640: return (const void*)walk_kernel<T, K, 640, 2>;
      case 512: return (const void*)walk_kernel<T, K, 512, 2>;
      case 256: return (const void*)walk_kernel<T, K, 256, 2>;
    }
    return nullptr;
  }
  switch (nt) {
    case 1024: return (const void*)walk_kernel<T, K, 1024, 1>;
    case 768: return (const void*)walk_kernel<T, K, 768, 1>;
    case 512: return (const void*)walk_kernel<T, K, 512, 1>;
    case 256: return (const void*)walk_kernel<T, K, 256, 1>;
  }
  return nullptr;
}
static const void* walk_select_outer(int K, bool outer) {
  if (K == 1) return outer ? (const void*)walk_kernel<double, 1, 512, 1, true> : (const void*)walk_kernel<double, 1, 512, 1>;
  if (K == 4) return outer ? (const void*)walk_kernel<double, 4, 512, 1, true> : (const void*)walk_kernel<double, 4, 512, 1>;
  return nullptr;
}
template <typename T>
static const void* walk_select(int K, int nt, int cpt) {
  if (K == 1) return walk_select_nt<T, 1>(nt, cpt);
  if (K == 4) return walk_select_nt<T, 4>(nt, cpt);
  return walk_select_nt<T, 0>(nt, cpt);
}

static size_t elem_size(const b2p_instance* h) {
  return h->cfg.precision == B2P_PRECISION_F32 ? 4 : 8;
}


// ------------------------------------------------------------------ dense-state path (host)
static int dense_create(b2p_instance* h) {
  const int N = h->cfg.n_tips, P = h->cfg.n_patterns, S = h->cfg.n_states;
  b2p_instance::Dense* d = new (std::nothrow) b2p_instance::Dense();
  if (!d) return fail(h, B2P_ERR_ALLOC, "out of host memory");
  h->dense = d;
  d->S = S;
  d->n_tiles = (P + DT - 1) / DT;
  d->cols_pad = d->n_tiles * DT;
  const size_t n_br = 2 * (size_t)N - 2, mat = (size_t)SP * SP * sizeof(double);
  const size_t part = (size_t)(N - 1) * SP * d->cols_pad * sizeof(double);
  int rc;
  if ((rc = dev_alloc(h, &d->evec, mat)) || (rc = dev_alloc(h, &d->ivec, mat)) ||
      (rc = dev_alloc(h, &d->evals, sizeof(double) * SP)) || (rc = dev_alloc(h, &d->q, mat)) || (rc = dev_alloc(h, &d->qt, mat)) ||
      (rc = dev_alloc(h, &d->freqs, sizeof(double) * SP)) || (rc = dev_alloc(h, &d->P, n_br * mat)) ||
      (rc = dev_alloc(h, &d->PT, n_br * mat)) || (rc = dev_alloc(h, &d->D, n_br * mat)) ||
      (rc = dev_alloc(h, &d->weights, sizeof(double) * d->cols_pad)) || (rc = dev_alloc(h, (char**)&d->msg, part)) ||
      (rc = dev_alloc(h, (char**)&d->pre, part)) || (rc = dev_alloc(h, &d->codes, (size_t)N * d->cols_pad)) ||
      (rc = dev_alloc(h, &d->scount, sizeof(int) * (d->sync_ints = (size_t)d->cols_pad + 2 * (size_t)(N - 1) * d->n_tiles + 1))) || (rc = dev_alloc(h, &d->nodes, sizeof(DenseNode) * (3 * (size_t)N - 3))) ||
      (rc = dev_alloc(h, &h->d_lpart, sizeof(double) * d->n_tiles)) ||
      (rc = dev_alloc(h, &h->d_gpart, sizeof(double) * (size_t)d->n_tiles * n_br)) ||
      (rc = dev_alloc(h, &h->d_bl_stage, sizeof(double) * n_br)) || (rc = dev_alloc(h, &h->d_out_stage, sizeof(double) * (n_br + 1))))
    return rc;
  CU(cudaMemset(d->msg, 0, part));
  CU(cudaMemset(d->pre, 0, part));
  CU(cudaMallocHost((void**)&h->h_pin_in, sizeof(double) * n_br));
  CU(cudaMallocHost((void**)&h->h_pin_out, sizeof(double) * (n_br + 1)));
  CU(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
  CU(cudaFuncSetAttribute((const void*)dense_post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * DENSE_TILE_BYTES));
  CU(cudaFuncSetAttribute((const void*)dense_pre_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * DENSE_TILE_BYTES));
  CU(cudaFuncSetAttribute((const void*)dense_transition_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * DENSE_TILE_BYTES));
  cudaFuncAttributes fa;
  CU(cudaFuncGetAttributes(&fa, (const void*)dense_pre_kernel));
  h->regs = fa.numRegs;
  return B2P_OK;
}

// Level lists: post-order levels bottom-up (a node after both children), pre-order levels by depth.
static int dense_set_topology(b2p_instance* h, const int32_t* tr) {
  const int N = h->cfg.n_tips, n_nodes = 2 * N - 1, root = 2 * N - 2;
  b2p_instance::Dense* d = h->dense;
  std::vector<int> left(n_nodes, -1), right(n_nodes, -1), parent(n_nodes, -1), lvl(n_nodes, 0), dep(n_nodes, 0);
  std::vector<char> seen(n_nodes, 0);
  for (int e = 0; e < N - 1; ++e) {
    const int v = tr[3 * e], l = tr[3 * e + 1], r = tr[3 * e + 2];
    if (v != N + e) return fail(h, B2P_ERR_ARG, "b2p_set_topology: triples must list internal nodes N..2N-2 in order (post-order ids)");
    if (l < 0 || r < 0 || l >= v || r >= v || l == r) return fail(h, B2P_ERR_ARG, "b2p_set_topology: child id out of range / not in post-order");
    if (seen[l] || seen[r]) return fail(h, B2P_ERR_ARG, "b2p_set_topology: a node has two parents");
    seen[l] = seen[r] = 1;
    left[v] = l; right[v] = r; parent[l] = parent[r] = v;
    lvl[v] = std::max(l >= N ? lvl[l] + 1 : 0, r >= N ? lvl[r] + 1 : 0);
  }
  for (int v = 0; v < root; ++v)
    if (!seen[v]) return fail(h, B2P_ERR_ARG, "b2p_set_topology: tree is not connected (node without parent)");
  int max_dep = 0;
  for (int v = root - 1; v >= N; --v) { dep[v] = dep[parent[v]] + 1; max_dep = std::max(max_dep, dep[v]); }
  std::vector<DenseNode> nodes;
  d->post_off.assign(1, 0);
  for (int L = 0; L <= lvl[root]; ++L) {
    for (int v = N; v < n_nodes; ++v)
      if (lvl[v] == L) nodes.push_back({v, left[v], right[v], 0});
    d->post_off.push_back((int)nodes.size());
  }
  // pre-order tasks {parent, child, sibling}: internal children by the depth of their parent
  // (dense_pre_kernel, one launch per depth), then every tip (dense_tip_kernel, one launch)
  d->pre_off.assign(1, (int)nodes.size());
  for (int L = 0; L <= max_dep; ++L) {
    for (int v = N; v < n_nodes; ++v)
      if (dep[v] == L) {
        if (left[v] >= N) nodes.push_back({v, left[v], right[v], 0});
        if (right[v] >= N) nodes.push_back({v, right[v], left[v], 0});
      }
    d->pre_off.push_back((int)nodes.size());
  }
  d->tip_off = (int)nodes.size();
  for (int v = N; v < n_nodes; ++v) {
    if (left[v] < N) nodes.push_back({v, left[v], right[v], 0});
    if (right[v] < N) nodes.push_back({v, right[v], left[v], 0});
  }
  d->root = root;
  int pre_launches = 0;  // depths that have an internal child
  for (size_t L = 0; L + 1 < d->pre_off.size(); ++L) pre_launches += d->pre_off[L + 1] > d->pre_off[L];
  (void)pre_launches;
  // one launch per pass: the level lists only fix the ORDER of the blocks (dependencies go by flag)
  h->plan.rounds_post = 1; h->plan.rounds_pre = 1; h->plan.rounds = 2;
  h->plan.tree_depth = max_dep + 1;
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(d->nodes, nodes.data(), nodes.size() * sizeof(DenseNode), cudaMemcpyHostToDevice));
  h->have_topology = true;
  return B2P_OK;
}

static int dense_set_tips(b2p_instance* h, const uint8_t* codes) {
  const int N = h->cfg.n_tips, P = h->cfg.n_patterns;
  b2p_instance::Dense* d = h->dense;
  std::vector<uint8_t> pad((size_t)N * d->cols_pad, 255);
  for (int i = 0; i < N; ++i) memcpy(pad.data() + (size_t)i * d->cols_pad, codes + (size_t)i * P, P);
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaMemcpy(d->codes, pad.data(), pad.size(), cudaMemcpyHostToDevice));
  h->have_tips = true;
  return B2P_OK;
}

static int dense_set_weights(b2p_instance* h, const double* w) {
  b2p_instance::Dense* d = h->dense;
  std::vector<double> pad(d->cols_pad, 0.0);
  memcpy(pad.data(), w, sizeof(double) * h->cfg.n_patterns);
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaMemcpy(d->weights, pad.data(), pad.size() * sizeof(double), cudaMemcpyHostToDevice));
  h->have_weights = true;
  return B2P_OK;
}

static int dense_set_eigen(b2p_instance* h, const double* evec, const double* ivec, const double* evals, const double* freqs) {
  b2p_instance::Dense* d = h->dense;
  const int S = d->S;
  // zero-padded SP x SP operands of the tile products: V^T, V^-1, Q, Q^T
  std::vector<double> q((size_t)SP * SP, 0.0), qt((size_t)SP * SP, 0.0), ev(SP, 0.0), fr(SP, 0.0);
  std::vector<double> vt((size_t)SP * SP, 0.0), vi((size_t)SP * SP, 0.0);
  for (int i = 0; i < S; ++i)
    for (int j = 0; j < S; ++j) {
      vt[(size_t)j * SP + i] = evec[i * S + j];
      vi[(size_t)i * SP + j] = ivec[i * S + j];
    }
  for (int i = 0; i < S; ++i)
    for (int l = 0; l < S; ++l) {
      double acc = 0.0;
      for (int j = 0; j < S; ++j) acc += evec[i * S + j] * evals[j] * ivec[j * S + l];
      q[(size_t)i * SP + l] = acc;
      qt[(size_t)l * SP + i] = acc;
    }
  for (int i = 0; i < S; ++i) { ev[i] = evals[i]; fr[i] = freqs[i]; }
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaMemcpy(d->evec, vt.data(), vt.size() * sizeof(double), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d->ivec, vi.data(), vi.size() * sizeof(double), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d->evals, ev.data(), sizeof(double) * SP, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d->freqs, fr.data(), sizeof(double) * SP, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d->q, q.data(), q.size() * sizeof(double), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d->qt, qt.data(), qt.size() * sizeof(double), cudaMemcpyHostToDevice));
  h->have_eigen = true;
  return B2P_OK;
}

static PeerArgs peer_args(b2p_instance* h);
static cudaError_t dense_launch(b2p_instance* h, const double* d_bl, double* d_logl, double* d_grad, cudaStream_t st) {
  b2p_instance::Dense* d = h->dense;
  const int N = h->cfg.n_tips, n_br = 2 * N - 2;
  DenseParams p;
  p.n_tips = N; p.n_states = d->S; p.n_cols = h->cfg.n_patterns; p.cols_pad = d->cols_pad;
  p.n_tiles = d->n_tiles; p.n_branches = n_br; p.rate = h->model.rates[0];
  p.P = d->P; p.PT = d->PT; p.D = d->D; p.Q = d->q; p.QT = d->qt; p.freqs = d->freqs; p.codes = d->codes;
  p.weights = d->weights; p.msg = d->msg; p.pre = d->pre; p.scount = d->scount; p.lpart = h->d_lpart; p.gpart = h->d_gpart;
  p.done_post = d->scount + d->cols_pad;
  p.done_pre = p.done_post + (size_t)(N - 1) * d->n_tiles;
  p.err = p.done_pre + (size_t)(N - 1) * d->n_tiles;
  p.prof = nullptr;
#ifdef B2P_DBG
  static unsigned long long* g_dense_prof = nullptr;
  if (getenv("B2P_DENSE_PROF")) {
    if (!g_dense_prof) { cudaMalloc(&g_dense_prof, 16 * sizeof(unsigned long long)); cudaMemset(g_dense_prof, 0, 16 * sizeof(unsigned long long)); }
    else {  // print what the previous evaluation left, then clear
      unsigned long long hp[16];
      cudaMemcpy(hp, g_dense_prof, sizeof hp, cudaMemcpyDeviceToHost);
      for (int k = 0; k < 2; ++k)
        if (hp[8 * k]) {
          fprintf(stderr, "dense %s CTAs %llu, mean clocks per phase: wait %.0f, load %.0f, product %.0f, store %.0f, publish %.0f\n", k ? "pre " : "post",
                  hp[8 * k], (double)hp[8 * k + 1] / hp[8 * k], (double)hp[8 * k + 2] / hp[8 * k], (double)hp[8 * k + 3] / hp[8 * k],
                  (double)hp[8 * k + 4] / hp[8 * k], (double)hp[8 * k + 5] / hp[8 * k]);
        }
      cudaMemset(g_dense_prof, 0, sizeof hp);
    }
    p.prof = g_dense_prof;
  }
#endif
  cudaError_t e = cudaMemsetAsync(d->scount, 0, sizeof(int) * d->sync_ints, st);
  if (e != cudaSuccess) return e;
  const int slot = (int)(h->prof_count % b2p_instance::kProfRing);
  if (h->profiling) cudaEventRecord(h->ev0[slot], st);
  dense_transition_kernel<<<n_br, DNT, 3 * DENSE_TILE_BYTES, st>>>(d->evec, d->ivec, d->evals, d->qt, d_bl, p.rate, d->S, d->P, d->PT, d->D);
  // Programmatic dependent launch (b2p_dense.cuh, level_wait): a pass's CTAs become resident while
  // the previous launch drains and run their prologue before griddepcontrol.wait.  Inside a pass
  // the blocks are in level order and wait for the tiles they read (tile_wait / tile_publish).
  cudaLaunchAttribute pdl[1];
  pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  pdl[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(DNT); cfg.stream = st; cfg.attrs = pdl; cfg.numAttrs = 1;
  {
    const int n = d->post_off.back() - d->post_off.front();
    cfg.gridDim = dim3((unsigned)n * d->n_tiles); cfg.dynamicSmemBytes = 2 * DENSE_TILE_BYTES;
    if ((e = cudaLaunchKernelEx(&cfg, dense_post_kernel, p, (const DenseNode*)(d->nodes + d->post_off.front()), d->root)) != cudaSuccess) return e;
  }
  if (d_grad) {
    const int n = d->pre_off.back() - d->pre_off.front();
    if (n > 0) {
      cfg.gridDim = dim3((unsigned)n * d->n_tiles); cfg.dynamicSmemBytes = 3 * DENSE_TILE_BYTES;
      if ((e = cudaLaunchKernelEx(&cfg, dense_pre_kernel, p, (const DenseNode*)(d->nodes + d->pre_off.front()), d->root)) != cudaSuccess) return e;
    }
    cfg.gridDim = dim3(N, d->n_tiles); cfg.dynamicSmemBytes = 0;
    if ((e = cudaLaunchKernelEx(&cfg, dense_tip_kernel, p, (const DenseNode*)(d->nodes + d->tip_off), d->root)) != cudaSuccess) return e;
  }
  if (h->profiling) { cudaEventRecord(h->ev1[slot], st); ++h->prof_count; }
  const int nb = d_grad ? n_br : 0;
  if (h->peer.attached) {  // the reduction over tiles doubles as the all-reduce over ranks, as on the 4-state path
    PeerArgs pa = peer_args(h);
    reduce_exchange_kernel<<<(nb + 31) / 32 + 1, 512, 0, st>>>(h->d_gpart, h->d_lpart, d->n_tiles, nb, d_grad, d_logl, pa);
  } else {
    reduce_kernel<<<(nb + 31) / 32 + 1, 512, 0, st>>>(h->d_gpart, h->d_lpart, d->n_tiles, nb, d_grad, d_logl);
  }
  return cudaGetLastError();
}

extern "C" {

int b2p_version(void) { return 2; }

const char* b2p_last_error(const b2p_instance* h) {
  return h ? h->err.c_str() : g_create_error.c_str();
}

static void peer_release(b2p_instance* h);

int b2p_destroy(b2p_instance* h) {
  if (!h) return B2P_OK;
  cudaSetDevice(h->cfg.device);
  graph_drop(h);
  peer_release(h);
  cudaFree(h->d_records); cudaFree(h->d_mat_dst);
  cudaFree(h->d_tips); cudaFree(h->d_tips_raw); cudaFree(h->d_tip_order); cudaFree(h->d_weights); cudaFree(h->d_msg); cudaFree(h->d_pre);
  cudaFree(h->d_lpart); cudaFree(h->d_trace); cudaFree(h->d_gpart); cudaFree(h->d_opart); cudaFree(h->d_mgpart); cudaFree(h->d_bl_stage); cudaFree(h->d_out_stage);
  if (h->dense) {
    b2p_instance::Dense* d = h->dense;
    for (void* p : {(void*)d->evec, (void*)d->ivec, (void*)d->evals, (void*)d->q, (void*)d->qt, (void*)d->freqs, (void*)d->P,
                    (void*)d->PT, (void*)d->D, (void*)d->weights, (void*)d->msg, (void*)d->pre, (void*)d->codes,
                    (void*)d->scount, (void*)d->nodes})
      cudaFree(p);
    delete d;
  }
  if (h->h_pin_in) cudaFreeHost(h->h_pin_in);
  if (h->h_pin_out) cudaFreeHost(h->h_pin_out);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  for (cudaEvent_t ev : h->ev0) cudaEventDestroy(ev);
  for (cudaEvent_t ev : h->ev1) cudaEventDestroy(ev);
  delete h;
  return B2P_OK;
}

int b2p_create(const b2p_config* cfg, b2p_instance** out) {
  b2p_instance* h = nullptr;
  if (!cfg || !out) return fail(h, B2P_ERR_ARG, "b2p_create: null argument");
  *out = nullptr;
  if (cfg->n_tips < 2 || cfg->n_patterns < 1 || cfg->n_categories < 1)
    return fail(h, B2P_ERR_ARG, "b2p_create: need n_tips >= 2, n_patterns >= 1, n_categories >= 1");
  if (cfg->n_categories > kMaxCat)
    return fail(h, B2P_ERR_UNSUPPORTED, "b2p_create: n_categories > %d", kMaxCat);
  if (cfg->n_states != 4 && (cfg->n_states < 5 || cfg->n_states > SP))
    return fail(h, B2P_ERR_UNSUPPORTED, "b2p_create: n_states must be 4 or 5..%d (got %d)", SP, cfg->n_states);
  if (cfg->n_states != 4 && (cfg->n_categories != 1 || cfg->precision != B2P_PRECISION_F64))
    return fail(h, B2P_ERR_UNSUPPORTED, "b2p_create: the dense-state path (n_states = %d) is built for one rate category in fp64",
                cfg->n_states);
  if (cfg->precision != B2P_PRECISION_F64 && cfg->precision != B2P_PRECISION_F32)
    return fail(h, B2P_ERR_ARG, "b2p_create: precision must be 64 or 32");
  if (cfg->flags & ~B2P_FLAG_OUTER_PRODUCTS) return fail(h, B2P_ERR_ARG, "b2p_create: unknown flag bits 0x%x", cfg->flags);
  const bool outer = (cfg->flags & B2P_FLAG_OUTER_PRODUCTS) != 0;
  if (outer && (cfg->n_states != 4 || cfg->precision != B2P_PRECISION_F64 || (cfg->n_categories != 1 && cfg->n_categories != 4)))
    return fail(h, B2P_ERR_UNSUPPORTED,
                "b2p_create: B2P_FLAG_OUTER_PRODUCTS is built for 4 states, fp64 and 1 or 4 rate categories "
                "(otherwise: b2p_set_gradient_operator, one evaluation per entry)");
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
  h->use_graph = !(getenv("B2P_NO_GRAPH") && atoi(getenv("B2P_NO_GRAPH")));
  const int K = cfg->n_categories, N = cfg->n_tips, P = cfg->n_patterns;
  const size_t es = elem_size(h);
  const size_t n_cols = (size_t)P * K;
  int rc = B2P_OK;
  auto bail = [&](int code) { std::string m = h->err; b2p_destroy(h); g_create_error = m; return code; };

  if (cfg->n_states != 4) {
    if ((rc = dense_create(h))) return bail(rc);
    *out = h;
    return B2P_OK;
  }
  // geometry: threads per CTA (warps_per_block * 32 if given), tile width, copy depth
  int nt_force = cfg->warps_per_block > 0 ? cfg->warps_per_block * 32 : 0, ct_force = 0, depth = 1;
  if (const char* s = getenv("B2P_NT")) nt_force = atoi(s);
  if (const char* s = getenv("B2P_CT")) ct_force = atoi(s);
  if (const char* s = getenv("B2P_DEPTH")) depth = atoi(s);
  int n_copy = 3, cpt = 2;  // copy warps: every bulk copy and the fold (sweep: profiles/README.md)
  if (const char* s = getenv("B2P_NP")) n_copy = std::max(1, std::min(4, atoi(s)));
  if (const char* s = getenv("B2P_CPT")) cpt = atoi(s) == 1 ? 1 : 2;
  // B2P_CTAS=2: two 256-thread CTAs per SM, each with half the shared memory.  They run out of
  // phase (separate barriers), so one's operand reads overlap the other's arithmetic.
  h->ctas_per_sm = 1;
  if (const char* s = getenv("B2P_CTAS")) h->ctas_per_sm = atoi(s) == 2 ? 2 : 1;
  size_t smem_budget = (size_t)prop.sharedMemPerBlockOptin - 1024;  // static shared + reservation
  if (h->ctas_per_sm == 2) {
    smem_budget = ((size_t)prop.sharedMemPerMultiprocessor - 2 * 1024) / 2 - 1024;
    nt_force = 256;
    if (!getenv("B2P_NP")) n_copy = 2;
  }
  if (outer) {  // one column per thread (the 4x4 products live in registers), 512 threads, warp-aligned slots
    nt_force = 512; cpt = 1; h->ctas_per_sm = 1;
    if (!getenv("B2P_NP")) n_copy = 3;
    smem_budget = (size_t)prop.sharedMemPerBlockOptin - 1024;
  }
  bool planned = plan_geometry(N, (int)n_cols, K, (int)es, h->n_sms * h->ctas_per_sm, smem_budget, nt_force, ct_force, depth, n_copy, cpt, h->wg, outer);
  // fp32 needs 80-96 registers per thread, so 768 threads fit (more tasks per round), with four
  // copy warps for the larger rounds (sweep: profiles/README.md, round 2: 0.227 vs 0.270 ms at the
  // flu shape).
  if (planned && !outer && cfg->precision == B2P_PRECISION_F32 && !nt_force && !getenv("B2P_NP") && h->ctas_per_sm == 1) {
    WalkGeometry g2;
    if (plan_geometry(N, (int)n_cols, K, (int)es, h->n_sms, smem_budget, 768, ct_force, depth, 4, cpt, g2) && g2.cpt == h->wg.cpt) {
      h->wg = g2;
      nt_force = 768;
      n_copy = 4;
    }
  }
  // Copy warps: a round has about one load and one store per slot, and a warp issues a bulk copy
  // every ~100 cycles, so few slots need fewer copy warps (and a freed warp is compute threads):
  // two when that leaves at most ten slots (sweeps: profiles/README.md, round 2).
  if (planned && !outer && !getenv("B2P_NP") && n_copy == 3 && h->wg.wn <= 10) {
    WalkGeometry g2;
    if (plan_geometry(N, (int)n_cols, K, (int)es, h->n_sms * h->ctas_per_sm, smem_budget, nt_force, ct_force, depth, 2, cpt, g2) && g2.wn <= 10)
      h->wg = g2;
  }
  // Copy depth 2 (loads issued two rounds ahead) when three frames are a small part of shared
  // memory: the round time no longer contains the HBM latency.  With many slots the frames (matrix
  // rows of every task) would take the block pool's room instead.
  if (planned && !getenv("B2P_DEPTH") && 3 * h->wg.frame_bytes <= 48 * 1024) {
    WalkGeometry g2;
    if (plan_geometry(N, (int)n_cols, K, (int)es, h->n_sms * h->ctas_per_sm, smem_budget, nt_force, h->wg.ct, 2, h->wg.n_copy_warps, cpt, g2, outer) &&
        g2.wn == h->wg.wn &&
        // outer-product mode gives ~100 KB to staging and sum rows: with fewer than ~5.5 pool blocks per
        // slot the third frame costs more rounds than the hidden latency saves (flu shape: 0.700 -> 0.686 ms)
        !(outer && 2 * g2.n_pool < 11 * g2.wn))
      h->wg = g2;
  }
  if (!planned) {
    fail(h, B2P_ERR_UNSUPPORTED, "b2p_create: no kernel geometry fits shared memory (K=%d, threads=%d)", K, nt_force);
    return bail(B2P_ERR_UNSUPPORTED);
  }
  h->kernel = (cfg->precision == B2P_PRECISION_F32) ? walk_select<float>(K, h->wg.nt, h->wg.cpt) : walk_select<double>(K, h->wg.nt, h->wg.cpt);
  if (outer) { h->kernel = walk_select_outer(K, false); h->kernel_outer = walk_select_outer(K, true); }
  if (!h->kernel || (outer && (!h->kernel_outer || h->wg.nt != 512 || h->wg.cpt != 1))) {
    fail(h, B2P_ERR_ARG, "b2p_create: no kernel for %d threads", h->wg.nt);
    return bail(B2P_ERR_ARG);
  }
  // function attributes are process-wide: never lower what another live instance asked for
  static std::map<const void*, size_t> smem_asked;
  static std::mutex smem_asked_lock;
  {
    std::lock_guard<std::mutex> guard(smem_asked_lock);
    size_t& asked = smem_asked[h->kernel];
    asked = std::max(asked, h->wg.smem);
    e = cudaFuncSetAttribute(h->kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)asked);
    if (e == cudaSuccess && h->kernel_outer) {
      size_t& asked2 = smem_asked[h->kernel_outer];
      asked2 = std::max(asked2, h->wg.smem);
      e = cudaFuncSetAttribute(h->kernel_outer, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)asked2);
    }
  }
  if (e != cudaSuccess) {
    fail(h, B2P_ERR_CUDA, "cudaFuncSetAttribute(smem=%zu): %s", h->wg.smem, cudaGetErrorString(e));
    return bail(B2P_ERR_CUDA);
  }
  cudaFuncAttributes fa;
  if ((e = cudaFuncGetAttributes(&fa, h->kernel)) != cudaSuccess) {
    fail(h, B2P_ERR_CUDA, "cudaFuncGetAttributes: %s (was the library built for sm_100a?)", cudaGetErrorString(e));
    return bail(B2P_ERR_CUDA);
  }
  h->regs = fa.numRegs;

  const size_t n_br = 2 * (size_t)N - 2;
  const size_t part_bytes = (size_t)h->wg.n_tiles * (size_t)(N - 1) * (size_t)h->wg.blk;
  if ((rc = dev_alloc(h, &h->d_tips_raw, (size_t)N * P))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_weights, sizeof(double) * P))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_mat_dst, sizeof(int) * 2 * n_br))) return bail(rc);
  if ((rc = dev_alloc(h, (char**)&h->d_msg, part_bytes))) return bail(rc);
  if ((rc = dev_alloc(h, (char**)&h->d_pre, part_bytes))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_lpart, sizeof(double) * h->wg.grid))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_gpart, sizeof(double) * (size_t)h->wg.grid * n_br))) return bail(rc);
  if (outer && (rc = dev_alloc(h, &h->d_opart, sizeof(double) * (size_t)h->wg.grid * n_br * K * 16))) return bail(rc);
  if (outer && (rc = dev_alloc(h, &h->d_mgpart, sizeof(double) * 17 * K * ((n_br + kModelGradThreads - 1) / kModelGradThreads)))) return bail(rc);
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

}  // extern "C"

template <typename T>
static void fill_walk_params(const b2p_instance* h, bool grad, WalkParams<T>& p) {
  const int N = h->cfg.n_tips, K = h->cfg.n_categories, P = h->cfg.n_patterns;
  const WalkGeometry& g = h->wg;
  p.n_tiles = g.n_tiles; p.rounds_run = grad ? h->plan.rounds : h->plan.rounds_post; p.rounds_post = h->plan.rounds_post;
  p.n_branches = 2 * N - 2; p.n_cat = K; p.n_cols = P * K; p.n_patterns = P; p.n_copy_warps = g.n_copy_warps;
  p.ct = g.ct; p.ts = g.ts; p.cpw = g.cpw; p.wn = g.wn; p.stages = g.stages; p.depth = g.depth;
  p.blk = g.blk; p.plane = g.ct * 16; p.ct_pad = g.ct_pad; p.row_stride = g.row_stride;
  p.frame_bytes = g.frame_bytes; p.off_frame = g.off_frame; p.rel_cr = g.rel_cr;
  p.rel_ir = g.rel_ir; p.rel_tips = g.rel_tips; p.ir_jobs = 16 + g.ir_ids;
  p.off_bar = g.off_bar; p.off_gbuf = g.off_gbuf; p.off_scount = g.off_scount; p.off_freq = g.off_freq;
  p.boot_base = h->plan.boot_base; p.tail_base = h->plan.tail_base + (grad ? 0u : (unsigned)g.ir_ids);
  for (int i = 0; i < 4; ++i) { p.boot_n[i] = h->plan.boot_n[i]; p.boot_tx[i] = h->plan.boot_tx[i]; }
  p.tile_stride_part = (unsigned long long)(N - 1) * g.blk;
  p.tile_stride_tips = (unsigned long long)h->plan.rounds * 2ull * g.wn * g.ct_pad;
  p.records = h->d_records;
  p.msg = (unsigned char*)h->d_msg; p.pre = (unsigned char*)h->d_pre; p.tips = h->d_tips;
  p.weights = h->d_weights; p.lpart = h->d_lpart; p.gpart = h->d_gpart; p.trace = h->d_trace;
  p.opart = h->d_opart; p.off_outer = g.off_outer; p.outer_row = g.outer_row; p.off_ostage = g.off_ostage;
  p.dbg = 0;
#ifdef B2P_DBG
  if (const char* e = getenv("B2P_DBG")) p.dbg = atoi(e);
#endif
  for (int i = 0; i < 4; ++i) p.freqs[i] = h->freqs[i];
  for (int i = 0; i < 16; ++i) p.q[i] = (T)(h->grad_op_set ? h->grad_op[i] : h->qmat[i]);
  for (int k = 0; k < kMaxCatW; ++k) {
    p.cat_w[k] = k < K ? h->cat_w[k] : 0.0;
    p.rates[k] = k < K ? (h->grad_op_set ? h->grad_scale[k] : h->model.rates[k]) : 0.0;
  }
}

static int peer_slices(const b2p_instance* h) { return 1 + (2 * h->cfg.n_tips - 2 + 31) / 32; }
static size_t peer_flag_bytes(const b2p_instance* h, int world) { return sizeof(unsigned long long) * 2 * (size_t)world * peer_slices(h); }
static size_t peer_recv_bytes(const b2p_instance* h, int world) {
  return sizeof(double) * 2 * (size_t)world * (2 * (size_t)h->cfg.n_tips - 1);
}

// Arguments of the fused reduction + all-reduce (the epoch counter lives on the device: graph-safe).
static PeerArgs peer_args(b2p_instance* h) {
  PeerArgs pa{};
  b2p_instance::Peer& pr = h->peer;
  pa.rank = pr.rank; pa.world = pr.world; pa.n_out = 2 * h->cfg.n_tips - 1;
  const size_t recv = peer_recv_bytes(h, pr.world);
  for (int r = 0; r < pr.world; ++r) {
    unsigned char* base = (unsigned char*)pr.mapped[r];
    pa.recv[r] = (double*)base;
    pa.flags[r] = (unsigned long long*)(base + recv);
  }
  unsigned char* tail = pr.own + recv + peer_flag_bytes(h, pr.world);
  pa.n_slices = peer_slices(h);
  pa.ticket = (unsigned*)tail;
  pa.err = (unsigned*)(tail + 8);
  pa.epoch = (unsigned long long*)(tail + 16);
  return pa;
}

template <typename T>
static cudaError_t launch_eval(b2p_instance* h, const double* d_bl, double* d_logl, double* d_grad,
                               cudaStream_t st, double* d_outer = nullptr) {
  const int N = h->cfg.n_tips, K = h->cfg.n_categories;
  const int n_br = 2 * N - 2;
  const int total = n_br * K * 16;
  transition_kernel<T><<<(total + 255) / 256, 256, 0, st>>>(h->model, d_bl, n_br, N, K, h->d_mat_dst, h->wg.row_stride,
                                                            h->d_records);
  const int slot = (int)(h->prof_count % b2p_instance::kProfRing);
  if (h->profiling) cudaEventRecord(h->ev0[slot], st);
  WalkParams<T> wp;
  fill_walk_params(h, d_grad != nullptr || d_outer != nullptr, wp);
  void* args[] = {&wp};
  cudaError_t e = cudaLaunchKernel(d_outer ? h->kernel_outer : h->kernel, dim3(h->wg.grid), dim3(h->wg.nt), args, h->wg.smem, st);
  if (e != cudaSuccess) return e;
  if (h->profiling) { cudaEventRecord(h->ev1[slot], st); ++h->prof_count; }
  const int nb = (d_grad && !d_outer) ? n_br : 0;  // outer-product mode: the gradient follows from the outer products
  if (h->peer.attached) {
    PeerArgs pa = peer_args(h);
    reduce_exchange_kernel<<<(nb + 31) / 32 + 1, 512, 0, st>>>(h->d_gpart, h->d_lpart, h->wg.grid, nb, d_grad, d_logl, pa);
  } else {
    reduce_kernel<<<(nb + 31) / 32 + 1, 512, 0, st>>>(h->d_gpart, h->d_lpart, h->wg.grid, nb, d_grad, d_logl);
  }
  if (d_outer) {
    const int n_pairs = n_br * K * 8;
    reduce_outer_kernel<<<(n_pairs + 63) / 64, dim3(64, 8), 0, st>>>(h->d_opart, h->wg.grid, n_br, K, d_outer);
    if (d_grad) {
      OuterGradArgs ga;
      for (int i = 0; i < 16; ++i) ga.q[i] = h->qmat[i];
      for (int k = 0; k < kMaxCat; ++k) ga.rates[k] = k < K ? h->model.rates[k] : 0.0;
      grad_from_outer_kernel<<<(n_br + 127) / 128, 128, 0, st>>>(d_outer, n_br, K, ga, d_grad);
    }
  }
  return cudaGetLastError();
}

// Tip masks per tile in schedule order: one bulk copy per round brings the rows it needs.
static int retile_tips(b2p_instance* h) {
  const WalkGeometry& g = h->wg;
  const int N = h->cfg.n_tips;
  const int n_rows = h->plan.rounds * 2 * g.wn;
  const long long total = (long long)g.n_tiles * n_rows * g.ct_pad;
  (void)N;
  CU(cudaSetDevice(h->cfg.device));
  tile_tips_kernel<<<(unsigned)((total + 255) / 256), 256>>>(h->d_tips_raw, h->d_tip_order, g.n_tiles, n_rows, g.ct, g.ct_pad,
                                                              h->cfg.n_patterns, h->cfg.n_categories, h->d_tips);
  CU(cudaGetLastError());
  CU(cudaDeviceSynchronize());
  return B2P_OK;
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
  const int64_t U = P * K * h->cfg.n_states * (int64_t)elem_size(h);
  o->n_tips = (int)N; o->n_patterns = (int)P; o->n_categories = (int)K; o->n_states = h->cfg.n_states;
  if (h->dense) {
    memset(&o->precision, 0, sizeof(*o) - offsetof(b2p_info, precision));
    o->precision = h->cfg.precision; o->warps_per_block = DNT / 32; o->cols_per_tile = DT; o->n_tiles = h->dense->n_tiles;
    o->grid_blocks = h->dense->n_tiles; o->blocks_per_sm = 2; o->n_sms = h->n_sms;
    o->post_rounds = h->plan.rounds_post; o->pre_rounds = h->plan.rounds_pre; o->tree_depth = h->plan.tree_depth;
    o->regs_per_thread = h->regs; o->smem_bytes = 3 * DENSE_TILE_BYTES; o->device_bytes = h->device_bytes;
    o->alg_bytes_loglik = (2 * N - 3) * U + N * P;
    o->alg_bytes_grad = (5 * N - 9) * U + 2 * N * P;
    return B2P_OK;
  }
  o->precision = h->cfg.precision;
  o->warps_per_block = h->wg.nt / 32; o->cols_per_tile = h->wg.ct; o->n_tiles = h->wg.n_tiles;
  o->grid_blocks = h->wg.grid; o->blocks_per_sm = h->ctas_per_sm; o->n_sms = h->n_sms;
  o->post_rounds = h->plan.rounds_post; o->tree_depth = h->plan.tree_depth; o->regs_per_thread = h->regs;
  o->tasks_per_round = h->plan.wn_eff; o->pre_rounds = h->plan.rounds_pre; o->pool_blocks = h->wg.n_pool;
  o->smem_bytes = (int32_t)h->wg.smem;
  o->copy_depth = h->wg.depth;
  o->device_bytes = h->device_bytes;
  o->alg_bytes_loglik = (2 * N - 3) * U + N * P;
  o->alg_bytes_grad = (5 * N - 9) * U + 2 * N * P;
  o->resident_reads = h->plan.resident_reads;
  o->block_loads = h->plan.loads_msg + h->plan.loads_pre;
  o->block_stores = h->plan.stores_msg + h->plan.stores_pre;
  return B2P_OK;
}

int b2p_set_topology(b2p_instance* h, const int32_t* h_triples) {
  if (h) graph_drop(h);
  if (!h || !h_triples) return h ? fail(h, B2P_ERR_ARG, "b2p_set_topology: null triples") : B2P_ERR_ARG;
  if (h->dense) return dense_set_topology(h, h_triples);
  WalkPlan plan;
  std::string why;
  if (!plan_walk_auto(h->cfg.n_tips, h_triples, h->wg, plan, why))
    return fail(h, B2P_ERR_ARG, "b2p_set_topology: %s", why.c_str());
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaDeviceSynchronize());
  cudaFree(h->d_records); cudaFree(h->d_tips); cudaFree(h->d_tip_order);
  h->d_records = nullptr; h->d_tips = nullptr; h->d_tip_order = nullptr;
  h->have_topology = false;
  const size_t tips_bytes = (size_t)h->wg.n_tiles * plan.tip_order.size() * h->wg.ct_pad;
  CU(cudaMalloc((void**)&h->d_records, plan.records.size()));  // the matrix rows inside are zero until K1 runs
  CU(cudaMemcpy(h->d_records, plan.records.data(), plan.records.size(), cudaMemcpyHostToDevice));
  CU(cudaMalloc((void**)&h->d_tips, tips_bytes));
  CU(cudaMalloc((void**)&h->d_tip_order, plan.tip_order.size() * sizeof(int32_t)));
  CU(cudaMemcpy(h->d_mat_dst, plan.mat_dst.data(), plan.mat_dst.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  if ((int)plan.tip_order.size() != plan.rounds * 2 * h->wg.wn) return fail(h, B2P_ERR_ARG, "b2p_set_topology: internal error (tip rows)");
  CU(cudaMemcpy(h->d_tip_order, plan.tip_order.data(), plan.tip_order.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  plan.rec_off.clear(); plan.rec_off.shrink_to_fit();
  plan.rec_mats.clear(); plan.rec_mats.shrink_to_fit();
  plan.tip_order.clear(); plan.tip_order.shrink_to_fit();
  plan.records.clear(); plan.records.shrink_to_fit();
  plan.mat_dst.clear(); plan.mat_dst.shrink_to_fit();
  h->plan = std::move(plan);
  if (getenv("B2P_TRACE")) {
    cudaFree(h->d_trace);
    CU(cudaMalloc((void**)&h->d_trace, sizeof(long long) * 16 * h->plan.rounds));
    CU(cudaMemset(h->d_trace, 0, sizeof(long long) * 16 * h->plan.rounds));
  }
  h->have_topology = true;
  if (h->have_tips) { int rc = retile_tips(h); if (rc) return rc; }
  return B2P_OK;
}

int b2p_set_tip_states(b2p_instance* h, const uint8_t* h_masks) {
  if (h) graph_drop(h);
  if (!h || !h_masks) return h ? fail(h, B2P_ERR_ARG, "b2p_set_tip_states: null masks") : B2P_ERR_ARG;
  if (h->dense) return dense_set_tips(h, h_masks);  // bytes are state codes (>= n_states: missing)
  const int N = h->cfg.n_tips, P = h->cfg.n_patterns;
  const size_t n = (size_t)N * P;
  for (size_t i = 0; i < n; ++i)
    if (h_masks[i] > 15) return fail(h, B2P_ERR_ARG, "b2p_set_tip_states: mask %u > 15 at %zu", h_masks[i], i);
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaMemcpy(h->d_tips_raw, h_masks, n, cudaMemcpyHostToDevice));
  if (h->have_topology) { int rc = retile_tips(h); if (rc) return rc; }
  h->have_tips = true;
  return B2P_OK;
}

int b2p_set_pattern_weights(b2p_instance* h, const double* h_weights) {
  if (h) graph_drop(h);
  if (!h || !h_weights) return h ? fail(h, B2P_ERR_ARG, "b2p_set_pattern_weights: null") : B2P_ERR_ARG;
  if (h->dense) return dense_set_weights(h, h_weights);
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaMemcpy(h->d_weights, h_weights, sizeof(double) * h->cfg.n_patterns, cudaMemcpyHostToDevice));
  h->have_weights = true;
  return B2P_OK;
}

int b2p_set_eigen(b2p_instance* h, const double* evec, const double* ivec, const double* evals,
                  const double* freqs) {
  if (h) graph_drop(h);
  if (!h || !evec || !ivec || !evals || !freqs)
    return h ? fail(h, B2P_ERR_ARG, "b2p_set_eigen: null argument") : B2P_ERR_ARG;
  if (h->dense) return dense_set_eigen(h, evec, ivec, evals, freqs);
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
  if (h) graph_drop(h);
  if (!h || !rates || !weights)
    return h ? fail(h, B2P_ERR_ARG, "b2p_set_categories: null argument") : B2P_ERR_ARG;
  for (int k = 0; k < h->cfg.n_categories; ++k) {
    h->model.rates[k] = rates[k];
    h->cat_w[k] = weights[k];
  }
  h->have_cats = true;
  return B2P_OK;
}

int b2p_set_gradient_operator(b2p_instance* h, const double* h_op, const double* h_scale) {
  if (h) graph_drop(h);
  if (!h) return B2P_ERR_ARG;
  if (h->dense) return fail(h, B2P_ERR_UNSUPPORTED, "b2p_set_gradient_operator: 4-state instances only");
  if (!h_op) { h->grad_op_set = false; return B2P_OK; }
  for (int i = 0; i < 16; ++i) h->grad_op[i] = h_op[i];
  for (int k = 0; k < h->cfg.n_categories; ++k) h->grad_scale[k] = h_scale ? h_scale[k] : 1.0;
  h->grad_op_set = true;
  return B2P_OK;
}

int b2p_peer_status(b2p_instance* h, int32_t* attached, int32_t* error_word);

static int eval_common(b2p_instance* h, const double* d_bl, double* d_logl, double* d_grad,
                       void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!d_bl || !d_logl) return fail(h, B2P_ERR_ARG, "evaluation: null device pointer");
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = h->dense ? dense_launch(h, d_bl, d_logl, d_grad, st)
                  : (h->cfg.precision == B2P_PRECISION_F32) ? launch_eval<float>(h, d_bl, d_logl, d_grad, st)
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

int b2p_branch_outer_products(b2p_instance* h, const double* d_bl, double* d_logl, double* d_grad, double* d_outer,
                              void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!h->kernel_outer) return fail(h, B2P_ERR_ARG, "b2p_branch_outer_products: the instance was created without B2P_FLAG_OUTER_PRODUCTS");
  if (!d_bl || !d_logl || !d_outer) return fail(h, B2P_ERR_ARG, "b2p_branch_outer_products: null device pointer");
  if (h->grad_op_set) return fail(h, B2P_ERR_ARG, "b2p_branch_outer_products: a gradient operator is set (b2p_set_gradient_operator)");
  if (h->peer.attached) return fail(h, B2P_ERR_UNSUPPORTED, "b2p_branch_outer_products: not with attached peers (sum the K x (2N-2) x 16 values with the caller's all-reduce)");
  cudaError_t e = launch_eval<double>(h, d_bl, d_logl, d_grad, (cudaStream_t)stream, d_outer);
  if (e != cudaSuccess) return fail(h, B2P_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
  return B2P_OK;
}

int b2p_model_gradients(b2p_instance* h, const double* d_bl, const double* d_outer, int32_t root_child, double* d_out,
                        void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!h->d_mgpart) return fail(h, B2P_ERR_ARG, "b2p_model_gradients: the instance was created without B2P_FLAG_OUTER_PRODUCTS");
  if (!d_bl || !d_outer || !d_out) return fail(h, B2P_ERR_ARG, "b2p_model_gradients: null device pointer");
  const int N = h->cfg.n_tips, K = h->cfg.n_categories, n_br = 2 * N - 2;
  if (root_child < 0 || root_child >= n_br) return fail(h, B2P_ERR_ARG, "b2p_model_gradients: root_child %d out of range", root_child);
  ModelGradArgs a;
  memcpy(a.evec, h->model.evec, sizeof a.evec);
  memcpy(a.ivec, h->model.ivec, sizeof a.ivec);
  memcpy(a.evals, h->model.evals, sizeof a.evals);
  memcpy(a.q, h->qmat, sizeof a.q);
  memcpy(a.freqs, h->freqs, sizeof a.freqs);
  for (int k = 0; k < kMaxCat; ++k) a.rates[k] = k < K ? h->model.rates[k] : 0.0;
  cudaStream_t st = (cudaStream_t)stream;
  const int blocks = (n_br + kModelGradThreads - 1) / kModelGradThreads;
  model_grad_kernel<<<dim3(blocks, K), kModelGradThreads, 0, st>>>(d_outer, d_bl, n_br, a, h->d_mgpart);
  model_grad_finish_kernel<<<1, 64, 0, st>>>(h->d_mgpart, blocks, K, d_outer, n_br, root_child, a, d_out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(h, B2P_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
  return B2P_OK;
}

int b2p_loglik_grad_host(b2p_instance* h, const double* h_bl, double* h_logl, double* h_grad) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!h_bl || !h_logl) return fail(h, B2P_ERR_ARG, "b2p_loglik_grad_host: null pointer");
  const size_t n_br = 2 * (size_t)h->cfg.n_tips - 2;
  CU(cudaSetDevice(h->cfg.device));
  memcpy(h->h_pin_in, h_bl, sizeof(double) * n_br);
  // No copy engine on this path: the pinned staging buffers are mapped into the device's address
  // space (unified addressing), K1 reads the 2N-2 branch lengths from host memory and the
  // reduction kernel writes logL and the gradient straight back — two PCIe round trips of 16 KB
  // instead of two copy-engine launches with their own latency.  The three launches are captured
  // once into a CUDA graph and replayed with one call (4-state path; not while profiling or tracing).
  const int gi = h_grad ? 1 : 0;
  double* out = h->h_pin_out;
  const bool graph = h->use_graph && !h->dense && !h->profiling && !h->d_trace;
  if (graph) {
    if (!h->graph_exec[gi]) {
      cudaGraph_t g = nullptr;
      CU(cudaStreamBeginCapture(h->own_stream, cudaStreamCaptureModeThreadLocal));
      rc = eval_common(h, h->h_pin_in, out, h_grad ? out + 1 : nullptr, h->own_stream);
      cudaError_t e = cudaStreamEndCapture(h->own_stream, &g);
      if (rc) { if (g) cudaGraphDestroy(g); return rc; }
      if (e != cudaSuccess) return fail(h, B2P_ERR_CUDA, "graph capture: %s", cudaGetErrorString(e));
      e = cudaGraphInstantiate(&h->graph_exec[gi], g, 0);
      cudaGraphDestroy(g);
      if (e != cudaSuccess) { h->graph_exec[gi] = nullptr; return fail(h, B2P_ERR_CUDA, "graph instantiate: %s", cudaGetErrorString(e)); }
    }
    CU(cudaGraphLaunch(h->graph_exec[gi], h->own_stream));
  } else if (h->dense) {
    CU(cudaMemcpyAsync(h->d_bl_stage, h->h_pin_in, sizeof(double) * n_br, cudaMemcpyHostToDevice, h->own_stream));
    rc = eval_common(h, h->d_bl_stage, h->d_out_stage, h_grad ? h->d_out_stage + 1 : nullptr, h->own_stream);
    if (rc) return rc;
    CU(cudaMemcpyAsync(out, h->d_out_stage, sizeof(double) * (h_grad ? n_br + 1 : 1), cudaMemcpyDeviceToHost, h->own_stream));
  } else {
    rc = eval_common(h, h->h_pin_in, out, h_grad ? out + 1 : nullptr, h->own_stream);
    if (rc) return rc;
  }
  CU(cudaStreamSynchronize(h->own_stream));
  *h_logl = out[0];
  if (h_grad) memcpy(h_grad, out + 1, sizeof(double) * n_br);
  if (h->peer.attached && *h_logl != *h_logl) {  // NaN: the exchange kernel poisons its results when a peer timed out
    int32_t err = 0;
    b2p_peer_status(h, nullptr, &err);
    if (err) return fail(h, B2P_ERR_CUDA, "b2p_loglik_grad_host: rank %d did not deliver its sums within the time limit (fused all-reduce)", err - 1);
  }
  if (h->d_trace && h_grad) {
    const int R = h->plan.rounds;
    std::vector<long long> tr(16 * (size_t)R);
    CU(cudaMemcpy(tr.data(), h->d_trace, tr.size() * sizeof(long long), cudaMemcpyDeviceToHost));
    if (FILE* f = fopen(getenv("B2P_TRACE"), "w")) {
      for (int w = 0; w < 2; ++w)
        for (int q = 0; q < R; ++q) {
          const long long* r = tr.data() + ((size_t)w * R + q) * 8;
          fprintf(f, "%d %d %lld %lld %lld %lld %lld %lld %lld %lld\n", w, q, r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7]);
        }
      fclose(f);
    }
  }
  return B2P_OK;
}

// ------------------------------------------------------------------ multi-GPU peer exchange
int b2p_peer_export(b2p_instance* h, int32_t world, void* h_handle_out) {
  if (!h || !h_handle_out) return h ? fail(h, B2P_ERR_ARG, "b2p_peer_export: null pointer") : B2P_ERR_ARG;
  if (world < 2 || world > kMaxPeers) return fail(h, B2P_ERR_ARG, "b2p_peer_export: world must be 2..%d", kMaxPeers);
  if (h->peer.own) return fail(h, B2P_ERR_ARG, "b2p_peer_export: already exported");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  CU(cudaSetDevice(h->cfg.device));
  const size_t bytes = peer_recv_bytes(h, world) + peer_flag_bytes(h, world) + 32;  // + ticket, err, epoch
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) return fail(h, B2P_ERR_ALLOC, "b2p_peer_export: cudaMalloc(%zu) failed", bytes);
  h->peer.own = (unsigned char*)p;
  h->peer.world = world;
  CU(cudaMemset(p, 0, bytes));
  CU(cudaDeviceSynchronize());
  cudaIpcMemHandle_t hd;
  CU(cudaIpcGetMemHandle(&hd, p));
  memcpy(h_handle_out, &hd, sizeof hd);
  return B2P_OK;
}

int b2p_peer_attach(b2p_instance* h, int32_t rank, int32_t world, const void* h_handles) {
  if (!h || !h_handles) return h ? fail(h, B2P_ERR_ARG, "b2p_peer_attach: null pointer") : B2P_ERR_ARG;
  b2p_instance::Peer& pr = h->peer;
  if (!pr.own || pr.world != world) return fail(h, B2P_ERR_ARG, "b2p_peer_attach: call b2p_peer_export(world) first");
  if (rank < 0 || rank >= world) return fail(h, B2P_ERR_ARG, "b2p_peer_attach: rank out of range");
  if (pr.attached) return fail(h, B2P_ERR_ARG, "b2p_peer_attach: already attached");
  CU(cudaSetDevice(h->cfg.device));
  for (int r = 0; r < world; ++r) {
    if (r == rank) { pr.mapped[r] = pr.own; continue; }
    cudaIpcMemHandle_t hd;
    memcpy(&hd, (const unsigned char*)h_handles + 64 * (size_t)r, sizeof hd);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      for (int q = 0; q < r; ++q)
        if (q != rank && pr.mapped[q]) { cudaIpcCloseMemHandle(pr.mapped[q]); pr.mapped[q] = nullptr; }
      cudaGetLastError();
      return fail(h, B2P_ERR_CUDA, "b2p_peer_attach: cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
    }
    pr.mapped[r] = p;
  }
  pr.rank = rank;
  pr.attached = true;
  graph_drop(h);
  return B2P_OK;
}

int b2p_peer_status(b2p_instance* h, int32_t* attached, int32_t* error_word) {
  if (!h) return B2P_ERR_ARG;
  if (attached) *attached = h->peer.attached ? 1 : 0;
  if (error_word) {
    *error_word = 0;
    if (h->peer.own) {
      CU(cudaSetDevice(h->cfg.device));
      unsigned e = 0;
      const size_t off = peer_recv_bytes(h, h->peer.world) + peer_flag_bytes(h, h->peer.world) + 8;
      CU(cudaMemcpy(&e, h->peer.own + off, sizeof e, cudaMemcpyDeviceToHost));
      *error_word = (int32_t)e;
    }
  }
  return B2P_OK;
}

static void peer_release(b2p_instance* h) {
  b2p_instance::Peer& pr = h->peer;
  for (int r = 0; r < pr.world; ++r)
    if (pr.attached && r != pr.rank && pr.mapped[r]) cudaIpcCloseMemHandle(pr.mapped[r]);
  cudaFree(pr.own);
  pr = b2p_instance::Peer();
}

int b2p_set_profiling(b2p_instance* h, int enabled) {
  if (h) graph_drop(h);
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

int b2p_get_branch_partials(b2p_instance* h, int32_t node, double* h_out) {
  if (!h || !h_out) return B2P_ERR_ARG;
  const int N = h->cfg.n_tips;
  if (h->dense) return fail(h, B2P_ERR_UNSUPPORTED, "b2p_get_branch_partials: 4-state instances only");
  if (node < N || node >= 2 * N - 2)
    return fail(h, B2P_ERR_ARG, "b2p_get_branch_partials: node %d is not an internal non-root node", node);
  const WalkGeometry& g = h->wg;
  const long long n_cols = (long long)h->cfg.n_patterns * h->cfg.n_categories;
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaDeviceSynchronize());
  std::vector<unsigned char> blk(g.blk);
  for (int tile = 0; tile < g.n_tiles; ++tile) {
    const unsigned char* src = (const unsigned char*)h->d_msg + ((size_t)tile * (N - 1) + (node - N)) * g.blk;
    CU(cudaMemcpy(blk.data(), src, g.blk, cudaMemcpyDeviceToHost));
    for (int c = 0; c < g.ct; ++c) {
      const long long cg = (long long)tile * g.ct + c;
      if (cg >= n_cols) break;
      double* o = h_out + cg * 4;
      if (h->cfg.precision == B2P_PRECISION_F64) {  // two planes of [ct][2]
        const double* p0 = (const double*)blk.data() + 2 * c;
        const double* p1 = (const double*)(blk.data() + (size_t)g.ct * 16) + 2 * c;
        o[0] = p0[0]; o[1] = p0[1]; o[2] = p1[0]; o[3] = p1[1];
      } else {
        const float* p0 = (const float*)blk.data() + 4 * c;
        for (int i = 0; i < 4; ++i) o[i] = p0[i];
      }
    }
  }
  return B2P_OK;
}

}  // extern "C"
