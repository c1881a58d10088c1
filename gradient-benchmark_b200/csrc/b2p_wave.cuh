// Wave-scheduled traversal kernel (sm_100a) — SURVEY.md §8(a) rows a3-a6.
//
// One persistent CTA per SM owns a tile of CT adjacent columns (column = pattern x
// category) and walks the WHOLE tree for it, so every dependency is CTA-local: no grid
// sync, no flags, one launch for both passes.  Per round the CTA executes WN independent
// node tasks side by side: thread t works on (slot t / TS, column t % TS), i.e. slots are
// packed at thread granularity, not warp granularity, so an awkward tile width (68
// columns at 1000 taxa x 10k patterns on 148 SMs) still fills the lanes.
//
// Operands never wait on HBM/L2 latency inside a round:
//   * a block produced in round p and consumed in round p+1 is read from a shared-memory
//     RING (two parities of ring_cap blocks);
//   * a block consumed later (>= p+2) is read back from global through a per-thread
//     STAGE filled with cp.async one round ahead (own column => no cross-thread hazard,
//     single-buffered: the registers are loaded first, the slot is re-armed at once);
//   * tip states (1 byte per pattern) travel one round ahead in a register;
//   * the 4x4 transition matrices of a round are staged cooperatively one round ahead.
// The host side of the C ABI builds the static schedule (Hu's level rule per pass, ring
// budget per round) and routes every operand / output (b2p_api.cu).
//
// Message formulation: what a node stores is not its partial x_v but the message to its
// parent  m_v = P_v x_v.  The post-order task of v is then  x_v = m_l o m_r ; m_v = P_v x_v
// and the pre-order task needs no matrix products for its inputs:
//   q_l = u_v o m_r,  q_r = u_v o m_l,  L = sum q_l.m_l,
//   dL/dt_l = r_k q_l.(Q m_l),  u_l = P_l^T q_l  (same for r).
#pragma once
#include "b2p_device.cuh"

namespace b2p {

constexpr int kMaxCatW = 32;

enum : unsigned char { KIND_TIP = 0, KIND_RING = 1, KIND_GLOBAL = 2, KIND_ROOT = 3 };
enum : unsigned char { F_O0_RING = 1, F_O0_GLOBAL = 2, F_O1_RING = 4, F_O1_GLOBAL = 8 };

struct WaveEntry {  // 32 bytes, copied as two 16-byte granules
  int v, l, r;      // node ids (tips < N); v < 0 = idle slot
  unsigned char kl, kr, kv, flags;
  unsigned short ring_l, ring_r, ring_v, ring_o0, ring_o1, pad0;
  int pad1;
};
static_assert(sizeof(WaveEntry) == 32, "WaveEntry must be 32 bytes");

template <typename T>
struct WaveParams {
  int n_tips, n_patterns, n_cat, n_cols;
  int ct, ts, cpw, wn;  // tile columns, threads per slot, columns per warp, tasks per round
  int n_tiles, rounds_post, rounds_pre, n_branches;
  int ring_cap, pj;     // ring blocks per parity, partial sums per slot
  int off_ent, off_mats, off_stage, off_ring, off_part, off_scount;  // smem byte offsets
  const WaveEntry* post;  // [rounds_post][wn]
  const WaveEntry* pre;   // [rounds_pre][wn]
  const uint8_t* tips;    // [N][P]
  const double* weights;  // [P]
  const T* pmat;          // [2N-2][K][16]
  T* msg;                 // [N-1][n_cols][4]  m_v = P_v x_v (root row unused)
  T* pre_g;               // [N-1][n_cols][4]  pre-order partials consumed >= 2 rounds later
  double* lpart;          // [grid]
  double* gpart;          // [grid][2N-2]
  T q[16];
  double freqs[4];
  double cat_w[kMaxCatW];
  double rates[kMaxCatW];
};

// padded shared-memory stride of one 4x4 matrix, in bytes (bank shift between categories)
template <typename T> struct MatBytes;
template <> struct MatBytes<double> { static constexpr int v = 144; };
template <> struct MatBytes<float> { static constexpr int v = 80; };

// one column (4 states) in shared memory.  fp64: two 16-byte halves in separate planes
// (consecutive threads -> consecutive 16 B => conflict-free LDS.128); fp32: one 16 B unit.
__device__ __forceinline__ void s_ld(const unsigned char* p, int plane, double (&x)[4]) {
  const double2 a = *reinterpret_cast<const double2*>(p);
  const double2 b = *reinterpret_cast<const double2*>(p + plane);
  x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y;
}
__device__ __forceinline__ void s_ld(const unsigned char* p, int, float (&x)[4]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
}
__device__ __forceinline__ void s_st(unsigned char* p, int plane, const double (&x)[4]) {
  *reinterpret_cast<double2*>(p) = make_double2(x[0], x[1]);
  *reinterpret_cast<double2*>(p + plane) = make_double2(x[2], x[3]);
}
__device__ __forceinline__ void s_st(unsigned char* p, int, const float (&x)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(x[0], x[1], x[2], x[3]);
}
__device__ __forceinline__ void s_cp(unsigned char* p, int plane, const double* g) {
  cp_async16(p, g);
  cp_async16(p + plane, g + 2);
}
__device__ __forceinline__ void s_cp(unsigned char* p, int, const float* g) { cp_async16(p, g); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <typename T> __host__ __device__ constexpr int col_planes() { return sizeof(T) == 8 ? 2 : 1; }

// row i of a staged matrix times x
template <typename T>
__device__ __forceinline__ T row_dot(const T* m, int i, const T (&x)[4]) {
  T r[4];
  load_row(m + 4 * i, r);
  return r[0] * x[0] + r[1] * x[1] + r[2] * x[2] + r[3] * x[3];
}


// Per-thread state of the wave kernel.  The thread -> (slot, column) map is static.
template <typename T, int K, int NT>
struct WaveCtx {
  const WaveParams<T>& p;
  unsigned char* smem;
  int t, lane, warp;
  int s, c;             // slot, column within the tile
  int kk, k;            // categories, this thread's category
  bool in_slot;         // s < wn and the lane maps to a column of the tile
  bool act;             // in_slot and the column exists (col_g < n_cols)
  int col_g;            // global column
  int pat;              // global pattern
  double w_p;           // pattern weight (used by k == 0 lanes)
  T cw, cwr;            // category weight, weight * rate
  unsigned same;        // bit i: lane + (1 << i) is in the same slot (segmented reductions)
  bool head;            // first lane of this slot's segment in the warp
  int part_j;           // index of this warp among the warps that touch slot s
  int tip_l, tip_r;     // tip masks of the NEXT round's task (register prefetch)
  double ll_acc;        // log-likelihood contributions finalised by this thread
  double* gp;           // this CTA's row of per-branch partial gradients
  bool first_tile;
  __device__ __forceinline__ WaveCtx(const WaveParams<T>& p_) : p(p_) {}

  __device__ __forceinline__ int blk_bytes() const { return p.ct * 4 * (int)sizeof(T); }
  __device__ __forceinline__ int ring_plane() const { return p.ct * 16; }
  __device__ __forceinline__ WaveEntry* ent(int round) const {
    return reinterpret_cast<WaveEntry*>(smem + p.off_ent) + (round % 3) * p.wn;
  }
  __device__ __forceinline__ unsigned char* mats(int parity) const {
    return smem + p.off_mats + parity * (p.wn * 3 * kk * MatBytes<T>::v);
  }
  __device__ __forceinline__ const T* mat(int parity, int which) const {
    return reinterpret_cast<const T*>(mats(parity) + ((s * 3 + which) * kk + k) * MatBytes<T>::v);
  }
  __device__ __forceinline__ unsigned char* stage(int op) const {
    return smem + p.off_stage + op * (NT * 16 * col_planes<T>()) + t * 16;
  }
  __device__ __forceinline__ unsigned char* ring(int parity, int idx) const {
    return smem + p.off_ring + (parity * p.ring_cap + idx) * blk_bytes() + c * 16;
  }
  // partial sums of one slot: {int l, r; double a[pj], b[pj]}
  __device__ __forceinline__ int part_stride() const { return 16 + 16 * p.pj; }
  __device__ __forceinline__ unsigned char* part(int parity, int slot) const {
    return smem + p.off_part + (parity * p.wn + slot) * part_stride();
  }
  __device__ __forceinline__ int* scount() const { return reinterpret_cast<int*>(smem + p.off_scount); }
  __device__ __forceinline__ size_t goff(int node) const {
    return ((size_t)(node - p.n_tips) * p.n_cols + col_g) * 4;
  }
};

// Sum a value over the lanes of this thread's slot inside the warp; valid in `head` lanes.
template <typename C>
__device__ __forceinline__ double seg_sum(const C& c, double v) {
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const double o = __shfl_down_sync(0xffffffffu, v, 1 << i);
    if ((c.same >> i) & 1) v += o;
  }
  return v;
}

// Cooperative copies for round `rn` (issued one round ahead): the slot matrices.
template <int PASS, typename T, int K, int NT>
__device__ __forceinline__ void prefetch_mats(const WaveCtx<T, K, NT>& c, int rn) {
  const WaveParams<T>& p = c.p;
  constexpr int G = (int)sizeof(T) * 16 / 16;  // 16-byte granules per 4x4 matrix: 8 fp64, 4 fp32
  constexpr int EPG = 16 / (int)sizeof(T);
  const int kk = K > 0 ? K : c.kk;
  const int per_slot = 3 * kk * G;
  const int total = p.wn * per_slot;
  const WaveEntry* en = c.ent(rn);
  unsigned char* dst = c.mats(rn & 1);
  const int root = 2 * p.n_tips - 2;
  for (int g = c.t; g < total; g += NT) {
    const int sl = g / per_slot, rem = g - sl * per_slot;
    const int which = rem / (kk * G), rem2 = rem - which * (kk * G);
    const int k_ = rem2 / G, gi = rem2 - k_ * G;
    const WaveEntry& e = en[sl];
    if (e.v < 0) continue;
    const int node = which == 0 ? e.v : (which == 1 ? e.l : e.r);
    bool need;
    if (PASS == 0) need = which == 0 ? (e.v != root) : (node < p.n_tips);
    else need = which != 0;
    if (need)
      cp_async16(dst + ((sl * 3 + which) * kk + k_) * MatBytes<T>::v + gi * 16,
                 p.pmat + ((size_t)node * kk + k_) * 16 + gi * EPG);
  }
}

// Copy the schedule entries of round `rn` into the 3-deep entry ring.
template <typename T, int K, int NT>
__device__ __forceinline__ void prefetch_entries(const WaveCtx<T, K, NT>& c, const WaveEntry* sched, int rn,
                                                 int n_rounds) {
  if (rn >= n_rounds) return;
  const int n16 = c.p.wn * 2;
  const int4* src = reinterpret_cast<const int4*>(sched + (size_t)rn * c.p.wn);
  int4* dst = reinterpret_cast<int4*>(c.ent(rn));
  for (int g = c.t; g < n16; g += NT) cp_async16(dst + g, src + g);
}

// This thread's own operands of round `rn`: global blocks into the stage, tip masks into
// registers.  Must run after the thread has consumed the stage for the current round.
template <int PASS, typename T, int K, int NT>
__device__ __forceinline__ void prefetch_operands(WaveCtx<T, K, NT>& c, int rn) {
  const WaveParams<T>& p = c.p;
  c.tip_l = c.tip_r = 15;
  if (!c.act) return;
  const WaveEntry e = c.ent(rn)[c.s];
  if (e.v < 0) return;
  constexpr int PL = NT * 16;
  if (e.kl == KIND_GLOBAL) s_cp(c.stage(0), PL, p.msg + c.goff(e.l));
  else if (e.kl == KIND_TIP) c.tip_l = __ldg(p.tips + (size_t)e.l * p.n_patterns + c.pat);
  if (e.kr == KIND_GLOBAL) s_cp(c.stage(1), PL, p.msg + c.goff(e.r));
  else if (e.kr == KIND_TIP) c.tip_r = __ldg(p.tips + (size_t)e.r * p.n_patterns + c.pat);
  if (PASS == 1 && e.kv == KIND_GLOBAL) s_cp(c.stage(2), PL, p.pre_g + c.goff(e.v));
}


// Message of a tip child: P e(mask), with e the 0/1 vector of compatible states.
template <typename T>
__device__ __forceinline__ void tip_message(const T* m, int mask, T (&y)[4]) {
  T x[4];
  tip_col<T>((uint8_t)mask, x);
#pragma unroll
  for (int i = 0; i < 4; ++i) y[i] = row_dot(m, i, x);
}

// y = M^T x
template <typename T>
__device__ __forceinline__ void matvec_t_rows(const T* m, const T (&x)[4], T (&y)[4]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) y[j] = (T)0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    T r[4];
    load_row(m + 4 * i, r);
#pragma unroll
    for (int j = 0; j < 4; ++j) y[j] += r[j] * x[i];
  }
}

// Power-of-two rescale shared by the K category lanes of a pattern; returns 1 if applied.
template <int K, typename T, typename C>
__device__ __forceinline__ int rescale(const C& c, T (&y)[4]) {
  T mx = fmax(fmax(y[0], y[1]), fmax(y[2], y[3]));
  mx = group_max<K>(mx, c.lane, c.kk);
  if (mx < Scale<T>::threshold() && mx > (T)0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= Scale<T>::factor();
    return 1;
  }
  return 0;
}

// Fetch one child message of the current round.
template <typename T, int K, int NT>
__device__ __forceinline__ void fetch_child(const WaveCtx<T, K, NT>& c, int r, int kind, int ring_idx, int op,
                                            int tip_mask, T (&x)[4]) {
  if (kind == KIND_RING) s_ld(c.ring((r - 1) & 1, ring_idx), c.ring_plane(), x);
  else if (kind == KIND_GLOBAL) s_ld(c.stage(op), NT * 16, x);
  else tip_message(c.mat(r & 1, 1 + op), tip_mask, x);
}

// Write the per-slot partial sums (a, b) of this warp and, by the slot's first thread,
// the branch ids they belong to.  Finalised one round later by finalize_parts().
template <typename T, int K, int NT>
__device__ __forceinline__ void put_parts(const WaveCtx<T, K, NT>& c, int r, int id_a, int id_b, double a, double b) {
  a = seg_sum(c, a);
  b = seg_sum(c, b);
  if (!c.in_slot) return;
  unsigned char* ps = c.part(r & 1, c.s);
  if (c.head) {
    reinterpret_cast<double*>(ps + 16)[c.part_j] = a;
    reinterpret_cast<double*>(ps + 16 + 8 * c.p.pj)[c.part_j] = b;
  }
  if (c.c == 0) {
    reinterpret_cast<int*>(ps)[0] = id_a;
    reinterpret_cast<int*>(ps)[1] = id_b;
  }
}

// ids >= 0: branch gradients (a -> id_a, b -> id_b); id_a == -2: a is a log-likelihood term.
template <typename T, int K, int NT>
__device__ __forceinline__ void finalize_parts(WaveCtx<T, K, NT>& c, int r_prev) {
  if (!(c.in_slot && c.c == 0)) return;
  const unsigned char* ps = c.part(r_prev & 1, c.s);
  const int id_a = reinterpret_cast<const int*>(ps)[0], id_b = reinterpret_cast<const int*>(ps)[1];
  if (id_a == -1) return;
  const int first_warp = (c.s * c.p.ts) >> 5, last_warp = (c.s * c.p.ts + c.p.ts - 1) >> 5;
  const int nj = last_warp - first_warp + 1;
  const double* pa = reinterpret_cast<const double*>(ps + 16);
  const double* pb = reinterpret_cast<const double*>(ps + 16 + 8 * c.p.pj);
  double a = 0.0, b = 0.0;
  for (int j = 0; j < nj; ++j) { a += pa[j]; b += pb[j]; }
  if (id_a == -2) { c.ll_acc += a; return; }
  if (c.first_tile) { c.gp[id_a] = a; c.gp[id_b] = b; }
  else { c.gp[id_a] += a; c.gp[id_b] += b; }
}

// All warp shuffles (category-group and segment reductions) sit on the unconditional path:
// lanes of one warp can belong to different slots (different tasks, idle slots).
template <typename T, int K, int NT>
__device__ __forceinline__ void post_task(WaveCtx<T, K, NT>& c, const WaveEntry& e, int r, bool last_round,
                                          const T (&ma)[4], const T (&mb)[4]) {
  const WaveParams<T>& p = c.p;
  const bool live = e.v >= 0 && c.act;
  const bool is_root = e.v == 2 * p.n_tips - 2;
  T y[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) y[i] = ma[i] * mb[i];
  const int bumped = rescale<K>(c, y);
  if (live && !is_root) {
    if (bumped && c.k == 0) atomicAdd(c.scount() + c.c, 1);
    T mv[4];
    const T* m = c.mat(r & 1, 0);
#pragma unroll
    for (int i = 0; i < 4; ++i) mv[i] = row_dot(m, i, y);
    if (e.flags & F_O0_RING) s_st(c.ring(r & 1, e.ring_o0), c.ring_plane(), mv);
    if (e.flags & F_O0_GLOBAL) st_col(p.msg + c.goff(e.v), mv);
  }
  if (!last_round) return;  // uniform over the CTA
  // root (alone in the last round, so every rescale count is final): site likelihoods
  T site = (T)p.freqs[0] * y[0] + (T)p.freqs[1] * y[1] + (T)p.freqs[2] * y[2] + (T)p.freqs[3] * y[3];
  site = group_sum<K>(site * c.cw, c.lane, c.kk);
  double term = 0.0;
  if (live && is_root && c.k == 0)
    term = c.w_p * (log((double)site) - (double)(c.scount()[c.c] + bumped) * Scale<T>::log_factor());
  put_parts(c, r, (e.v >= 0 && is_root) ? -2 : -1, -1, term, 0.0);
}

// Operands of the current round -> registers (zeros for idle / inactive lanes).
template <int PASS, typename T, int K, int NT>
__device__ __forceinline__ void load_operands(const WaveCtx<T, K, NT>& c, const WaveEntry& e, int r, T (&ma)[4],
                                              T (&mb)[4], T (&u)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) ma[i] = mb[i] = u[i] = (T)0;
  if (!(e.v >= 0 && c.act)) return;
  fetch_child(c, r, e.kl, e.ring_l, 0, c.tip_l, ma);
  fetch_child(c, r, e.kr, e.ring_r, 1, c.tip_r, mb);
  if (PASS == 1) {
    if (e.kv == KIND_RING) s_ld(c.ring((r - 1) & 1, e.ring_v), c.ring_plane(), u);
    else if (e.kv == KIND_GLOBAL) s_ld(c.stage(2), NT * 16, u);
    else {
#pragma unroll
      for (int i = 0; i < 4; ++i) u[i] = (T)c.p.freqs[i];
    }
  }
}

template <typename T, int K, int NT>
__device__ __forceinline__ void pre_task(WaveCtx<T, K, NT>& c, const WaveEntry& e, int r, T (&ma)[4], T (&mb)[4],
                                         const T (&u)[4]) {
  const WaveParams<T>& p = c.p;
  const int N = p.n_tips;
  const bool live = e.v >= 0 && c.act;
  T ql[4], qr[4];
  T den = (T)0, numl = (T)0, numr = (T)0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    ql[i] = u[i] * mb[i];
    qr[i] = u[i] * ma[i];
    den += ql[i] * ma[i];
  }
  // d/dt (P x) = r Q (P x): Q comes from the constant bank
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const T dl = p.q[4 * i] * ma[0] + p.q[4 * i + 1] * ma[1] + p.q[4 * i + 2] * ma[2] + p.q[4 * i + 3] * ma[3];
    const T dr = p.q[4 * i] * mb[0] + p.q[4 * i + 1] * mb[1] + p.q[4 * i + 2] * mb[2] + p.q[4 * i + 3] * mb[3];
    numl += ql[i] * dl;
    numr += qr[i] * dr;
  }
  den = group_sum<K>(den * c.cw, c.lane, c.kk);
  numl = group_sum<K>(numl * c.cwr, c.lane, c.kk);
  numr = group_sum<K>(numr * c.cwr, c.lane, c.kk);
  double gl = 0.0, gr = 0.0;
  if (live && c.k == 0) {
    const double inv = c.w_p / (double)den;
    gl = (double)numl * inv;
    gr = (double)numr * inv;
  }
  put_parts(c, r, e.v >= 0 ? e.l : -1, e.r, gl, gr);
  // pre-order partials of the internal children: u_c = P_c^T q_c  (ma / mb are dead)
  const bool want_l = live && e.l >= N, want_r = live && e.r >= N;
  if (want_l) matvec_t_rows(c.mat(r & 1, 1), ql, ma);
  const int bl = rescale<K>(c, ma);
  (void)bl;
  if (want_l) {
    if (e.flags & F_O0_RING) s_st(c.ring(r & 1, e.ring_o0), c.ring_plane(), ma);
    if (e.flags & F_O0_GLOBAL) st_col(p.pre_g + c.goff(e.l), ma);
  }
  if (want_r) matvec_t_rows(c.mat(r & 1, 2), qr, mb);
  const int br = rescale<K>(c, mb);
  (void)br;
  if (want_r) {
    if (e.flags & F_O1_RING) s_st(c.ring(r & 1, e.ring_o1), c.ring_plane(), mb);
    if (e.flags & F_O1_GLOBAL) st_col(p.pre_g + c.goff(e.r), mb);
  }
}

// One pass over its schedule.  Round r: (1) take this round's operands out of ring /
// stage / tip registers, (2) re-arm the stage and the cooperative buffers for round r+1
// (entries for r+2), (3) compute and store, (4) fold the previous round's partial sums,
// (5) cp.async.wait_all + one CTA barrier.
template <int PASS, typename T, int K, int NT>
__device__ __forceinline__ void wave_pass(WaveCtx<T, K, NT>& c) {
  const WaveParams<T>& p = c.p;
  const WaveEntry* sched = PASS ? p.pre : p.post;
  const int R = PASS ? p.rounds_pre : p.rounds_post;
  // prologue: entries of rounds 0 and 1, then everything round 0 needs
  prefetch_entries(c, sched, 0, R);
  prefetch_entries(c, sched, 1, R);
  cp_async_wait_all();
  __syncthreads();
  prefetch_mats<PASS>(c, 0);
  prefetch_operands<PASS>(c, 0);
  cp_async_wait_all();
  __syncthreads();
  WaveEntry idle;
  idle.v = -1; idle.l = idle.r = -1; idle.kl = idle.kr = idle.kv = KIND_TIP; idle.flags = 0;
  idle.ring_l = idle.ring_r = idle.ring_v = idle.ring_o0 = idle.ring_o1 = 0;
  for (int r = 0; r < R; ++r) {
    WaveEntry e = idle;
    if (c.in_slot) e = c.ent(r)[c.s];
    // the task loads its staged operands before the prefetch below re-arms the same
    // per-thread stage slot (program order within one thread: LDS, then LDGSTS)
    T ma[4], mb[4], u[4];
    load_operands<PASS>(c, e, r, ma, mb, u);
    if (r + 1 < R) {
      prefetch_operands<PASS>(c, r + 1);
      prefetch_mats<PASS>(c, r + 1);
    }
    prefetch_entries(c, sched, r + 2, R);
    if (PASS == 0) post_task(c, e, r, r == R - 1, ma, mb); else pre_task(c, e, r, ma, mb, u);
    if (PASS == 1 && r > 0) finalize_parts(c, r - 1);
    cp_async_wait_all();
    __syncthreads();
  }
  finalize_parts(c, R - 1);
  __syncthreads();
}

template <typename T, int K, int NT, bool GRAD>
__global__ void __launch_bounds__(NT, 1) wave_kernel(const WaveParams<T> p) {
  extern __shared__ __align__(16) unsigned char wave_smem[];
  __shared__ double red[NT / 32];
  WaveCtx<T, K, NT> c(p);
  c.smem = wave_smem;
  c.t = threadIdx.x; c.lane = c.t & 31; c.warp = c.t >> 5;
  c.kk = (K > 0) ? K : p.n_cat;
  c.s = c.t / p.ts;
  const int tt = c.t - c.s * p.ts;
  c.c = (tt >> 5) * p.cpw + (tt & 31);
  c.in_slot = c.s < p.wn && (tt & 31) < p.cpw && c.c < p.ct;
  if (p.cpw == 32) c.in_slot = c.s < p.wn;  // packed mapping: ts == ct
  if (!c.in_slot) c.c = 0;
  c.k = c.c % c.kk;
  c.cw = (T)p.cat_w[c.k];
  c.cwr = (T)(p.cat_w[c.k] * p.rates[c.k]);
  // segment structure of the warp (static)
  c.same = 0;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int o = c.lane + (1 << i);
    const int so = __shfl_down_sync(0xffffffffu, c.s, 1 << i);
    if (o < 32 && so == c.s) c.same |= 1u << i;
  }
  const int sprev = __shfl_up_sync(0xffffffffu, c.s, 1);
  c.head = c.lane == 0 || sprev != c.s;
  c.part_j = c.warp - ((c.s * p.ts) >> 5);
  c.ll_acc = 0.0;
  c.gp = p.gpart + (size_t)blockIdx.x * p.n_branches;
  c.first_tile = true;

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, c.first_tile = false) {
    c.col_g = tile * p.ct + c.c;
    c.act = c.in_slot && c.col_g < p.n_cols;
    c.pat = c.act ? c.col_g / c.kk : 0;
    c.w_p = c.act ? __ldg(p.weights + c.pat) : 0.0;
    for (int i = c.t; i < p.ct; i += NT) c.scount()[i] = 0;
    __syncthreads();
    wave_pass<0>(c);
    if (GRAD) wave_pass<1>(c);
  }

  double v = warp_sum(c.ll_acc);
  if (c.lane == 0) red[c.warp] = v;
  __syncthreads();
  if (c.t == 0) {
    double s = 0.0;
    for (int w = 0; w < NT / 32; ++w) s += red[w];
    p.lpart[blockIdx.x] = s;
  }
}

}  // namespace b2p
