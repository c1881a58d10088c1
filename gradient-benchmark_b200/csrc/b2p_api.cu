// libb200phylo.so — C ABI (include/b200phylo.h) over the sm_100a kernels.
// Host side: instance state, the wave planner (b2p_plan.h), launch sizing.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/b200phylo.h"
#include "b2p_aux.cuh"
#include "b2p_plan.h"

using namespace b2p;

static thread_local std::string g_create_error;

struct b2p_instance {
  b2p_config cfg{};
  int n_sms = 0, regs = 0, ctas_per_sm = 1;
  bool have_topology = false, have_tips = false, have_weights = false, have_eigen = false,
       have_cats = false;
  // geometry + schedules of the wave kernel
  WaveGeometry wg;
  int rounds_post = 0, rounds_pre = 0, tree_depth = 0;
  long long ring_post = 0, ring_pre = 0, pre_global = 0;
  const void* k_grad = nullptr;
  const void* k_eval = nullptr;
  // device scratch
  WaveTask* d_post_task = nullptr;
  WaveAhead* d_post_ahead = nullptr;
  WaveTask* d_pre_task = nullptr;
  WaveAhead* d_pre_ahead = nullptr;
  uint8_t* d_tips = nullptr;
  double* d_weights = nullptr;
  void* d_pmat = nullptr;
  void* d_msg = nullptr;   // [N-1][P*K][4] messages m_v = P_v x_v
  void* d_pre = nullptr;   // [N-1][P*K][4] pre-order partials that are read >= 2 rounds later
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
template <typename T, int K, int NT, int MINB>
static void wave_pick(const void** grad, const void** eval) {
  *grad = (const void*)wave_kernel<T, K, NT, MINB, true>;
  *eval = (const void*)wave_kernel<T, K, NT, MINB, false>;
}
template <typename T, int K>
static bool wave_select_nt(int nt, int minb, const void** grad, const void** eval) {
  if (minb == 1 && nt == 1024) wave_pick<T, K, 1024, 1>(grad, eval);
  else if (minb == 1 && nt == 768) wave_pick<T, K, 768, 1>(grad, eval);
  else if (minb == 1 && nt == 512) wave_pick<T, K, 512, 1>(grad, eval);
  else if (minb == 1 && nt == 256) wave_pick<T, K, 256, 1>(grad, eval);
  else if (minb == 2 && nt == 512) wave_pick<T, K, 512, 2>(grad, eval);
  else if (minb == 2 && nt == 256) wave_pick<T, K, 256, 2>(grad, eval);
  else return false;
  return true;
}
template <typename T>
static bool wave_select(int K, int nt, int minb, const void** grad, const void** eval) {
  if (K == 1) return wave_select_nt<T, 1>(nt, minb, grad, eval);
  if (K == 4) return wave_select_nt<T, 4>(nt, minb, grad, eval);
  return wave_select_nt<T, 0>(nt, minb, grad, eval);
}

static size_t elem_size(const b2p_instance* h) {
  return h->cfg.precision == B2P_PRECISION_F32 ? 4 : 8;
}

extern "C" {

int b2p_version(void) { return 2; }

const char* b2p_last_error(const b2p_instance* h) {
  return h ? h->err.c_str() : g_create_error.c_str();
}

int b2p_destroy(b2p_instance* h) {
  if (!h) return B2P_OK;
  cudaSetDevice(h->cfg.device);
  cudaFree(h->d_post_task); cudaFree(h->d_post_ahead); cudaFree(h->d_pre_task); cudaFree(h->d_pre_ahead);
  cudaFree(h->d_tips); cudaFree(h->d_weights); cudaFree(h->d_pmat); cudaFree(h->d_msg); cudaFree(h->d_pre);
  cudaFree(h->d_lpart); cudaFree(h->d_gpart); cudaFree(h->d_bl_stage); cudaFree(h->d_out_stage);
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
  int rc = B2P_OK;
  auto bail = [&](int code) { std::string m = h->err; b2p_destroy(h); g_create_error = m; return code; };

  // geometry: threads per CTA (warps_per_block * 32 if given), CTAs per SM, tile width
  int nt_force = cfg->warps_per_block > 0 ? cfg->warps_per_block * 32 : 0, ct_force = 0, cps = 1;
  if (const char* s = getenv("B2P_NT")) nt_force = atoi(s);
  if (const char* s = getenv("B2P_CT")) ct_force = atoi(s);
  if (const char* s = getenv("B2P_CTAS_PER_SM")) cps = std::max(1, std::min(2, atoi(s)));
  h->ctas_per_sm = cps;
  // per-CTA budget: the opt-in maximum, or half the SM's shared memory minus the 1 KB system reservation per CTA
  const size_t smem_budget = cps == 1 ? (size_t)prop.sharedMemPerBlockOptin - 1024
                                      : (size_t)prop.sharedMemPerMultiprocessor / 2 - 2048;
  if (cps == 2 && nt_force == 0) nt_force = 512;
  if (!plan_geometry(N, (int)n_cols, K, (int)es, h->n_sms, smem_budget, nt_force, ct_force, cps, h->wg)) {
    fail(h, B2P_ERR_UNSUPPORTED, "b2p_create: no kernel geometry fits shared memory (K=%d, threads=%d)", K, nt_force);
    return bail(B2P_ERR_UNSUPPORTED);
  }
  const bool ok = (cfg->precision == B2P_PRECISION_F32) ? wave_select<float>(K, h->wg.nt, cps, &h->k_grad, &h->k_eval)
                                                        : wave_select<double>(K, h->wg.nt, cps, &h->k_grad, &h->k_eval);
  if (!ok) { fail(h, B2P_ERR_ARG, "b2p_create: no kernel for %d threads x %d CTAs/SM", h->wg.nt, cps); return bail(B2P_ERR_ARG); }
  for (const void* k : {h->k_grad, h->k_eval}) {
    if ((e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->wg.smem)) != cudaSuccess) {
      fail(h, B2P_ERR_CUDA, "cudaFuncSetAttribute(smem=%zu): %s", h->wg.smem, cudaGetErrorString(e));
      return bail(B2P_ERR_CUDA);
    }
  }
  cudaFuncAttributes fa;
  if ((e = cudaFuncGetAttributes(&fa, h->k_grad)) != cudaSuccess) {
    fail(h, B2P_ERR_CUDA, "cudaFuncGetAttributes: %s (was the library built for sm_100a?)", cudaGetErrorString(e));
    return bail(B2P_ERR_CUDA);
  }
  h->regs = fa.numRegs;

  const size_t n_br = 2 * (size_t)N - 2;
  if ((rc = dev_alloc(h, &h->d_tips, (size_t)N * P))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_weights, sizeof(double) * P))) return bail(rc);
  if ((rc = dev_alloc(h, (char**)&h->d_pmat, n_br * K * 16 * es))) return bail(rc);
  if ((rc = dev_alloc(h, (char**)&h->d_msg, (size_t)(N - 1) * n_cols * 4 * es))) return bail(rc);
  if ((rc = dev_alloc(h, (char**)&h->d_pre, (size_t)(N - 1) * n_cols * 4 * es))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_lpart, sizeof(double) * h->wg.grid))) return bail(rc);
  if ((rc = dev_alloc(h, &h->d_gpart, sizeof(double) * (size_t)h->wg.grid * n_br))) return bail(rc);
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
static void fill_wave_params(const b2p_instance* h, WaveParams<T>& p) {
  const int N = h->cfg.n_tips, K = h->cfg.n_categories, P = h->cfg.n_patterns;
  const WaveGeometry& g = h->wg;
  p.n_tips = N; p.n_patterns = P; p.n_cat = K; p.n_cols = P * K;
  p.ct = g.ct; p.ts = g.ts; p.cpw = g.cpw; p.wn = g.wn;
  p.n_tiles = g.n_tiles; p.rounds_post = h->rounds_post; p.rounds_pre = h->rounds_pre;
  p.n_branches = 2 * N - 2;
  p.pj = g.pj; p.plane = g.ct * 16; p.stage_op_stride = g.wn * g.blk; p.pad0 = 0;
  p.off_task = g.off_task; p.off_ahead = g.off_ahead; p.off_mats = g.off_mats; p.off_stage = g.off_stage;
  p.off_part = g.off_part; p.off_scount = g.off_scount;
  p.post_task = h->d_post_task; p.post_ahead = h->d_post_ahead;
  p.pre_task = h->d_pre_task; p.pre_ahead = h->d_pre_ahead;
  p.tips = h->d_tips; p.weights = h->d_weights; p.pmat = (const T*)h->d_pmat;
  p.msg = (T*)h->d_msg; p.pre_g = (T*)h->d_pre;
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
  WaveParams<T> wp;
  fill_wave_params(h, wp);
  void* args[] = {&wp};
  cudaError_t e = cudaLaunchKernel(d_grad ? h->k_grad : h->k_eval, dim3(h->wg.grid), dim3(h->wg.nt), args,
                                   h->wg.smem, st);
  if (e != cudaSuccess) return e;
  if (h->profiling) { cudaEventRecord(h->ev1[slot], st); ++h->prof_count; }
  const int nb = d_grad ? n_br : 0;
  reduce_kernel<<<(nb + 31) / 32 + 1, 512, 0, st>>>(h->d_gpart, h->d_lpart, h->wg.grid, nb, d_grad, d_logl);
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
  o->warps_per_block = h->wg.nt / 32; o->cols_per_tile = h->wg.ct; o->n_tiles = h->wg.n_tiles;
  o->grid_blocks = h->wg.grid; o->blocks_per_sm = h->ctas_per_sm; o->n_sms = h->n_sms;
  o->post_rounds = h->rounds_post; o->tree_depth = h->tree_depth; o->regs_per_thread = h->regs;
  o->tasks_per_round = h->wg.wn; o->pre_rounds = h->rounds_pre; o->ring_blocks = h->wg.ring_cap;
  o->smem_bytes = (int32_t)h->wg.smem;
  o->device_bytes = h->device_bytes;
  o->alg_bytes_loglik = (2 * N - 3) * U + N * P;
  o->alg_bytes_grad = (5 * N - 9) * U + 2 * N * P;
  o->ring_reads_post = h->ring_post; o->ring_reads_pre = h->ring_pre; o->pre_global_blocks = h->pre_global;
  return B2P_OK;
}

int b2p_set_topology(b2p_instance* h, const int32_t* h_triples) {
  if (!h || !h_triples) return h ? fail(h, B2P_ERR_ARG, "b2p_set_topology: null triples") : B2P_ERR_ARG;
  WaveSchedule ws;
  std::string why;
  if (!plan_schedule(h->cfg.n_tips, h->cfg.n_patterns, h->cfg.n_categories, h_triples, h->wg, ws, why))
    return fail(h, B2P_ERR_ARG, "b2p_set_topology: %s", why.c_str());
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaDeviceSynchronize());
  cudaFree(h->d_post_task); cudaFree(h->d_post_ahead); cudaFree(h->d_pre_task); cudaFree(h->d_pre_ahead);
  h->d_post_task = h->d_pre_task = nullptr;
  h->d_post_ahead = h->d_pre_ahead = nullptr;
  h->have_topology = false;
  auto up = [&](auto** dst, const auto& v) -> cudaError_t {
    cudaError_t e = cudaMalloc((void**)dst, v.size() * sizeof(v[0]));
    if (e != cudaSuccess) return e;
    return cudaMemcpy(*dst, v.data(), v.size() * sizeof(v[0]), cudaMemcpyHostToDevice);
  };
  CU(up(&h->d_post_task, ws.post_task));
  CU(up(&h->d_post_ahead, ws.post_ahead));
  CU(up(&h->d_pre_task, ws.pre_task));
  CU(up(&h->d_pre_ahead, ws.pre_ahead));
  h->rounds_post = ws.rounds_post; h->rounds_pre = ws.rounds_pre; h->tree_depth = ws.depth;
  h->ring_post = ws.ring_reads_post; h->ring_pre = ws.ring_reads_pre; h->pre_global = ws.pre_global_blocks;
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

int b2p_get_branch_partials(b2p_instance* h, int32_t node, double* h_out) {
  if (!h || !h_out) return B2P_ERR_ARG;
  const int N = h->cfg.n_tips;
  if (node < N || node >= 2 * N - 2)
    return fail(h, B2P_ERR_ARG, "b2p_get_branch_partials: node %d is not an internal non-root node", node);
  const size_t n = (size_t)h->cfg.n_patterns * h->cfg.n_categories * 4;
  CU(cudaSetDevice(h->cfg.device));
  CU(cudaDeviceSynchronize());
  if (h->cfg.precision == B2P_PRECISION_F64) {
    CU(cudaMemcpy(h_out, (const double*)h->d_msg + (size_t)(node - N) * n, n * 8, cudaMemcpyDeviceToHost));
  } else {
    std::vector<float> tmp(n);
    CU(cudaMemcpy(tmp.data(), (const float*)h->d_msg + (size_t)(node - N) * n, n * 4, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; ++i) h_out[i] = tmp[i];
  }
  return B2P_OK;
}

}  // extern "C"
