// Wave-scheduled traversal kernel (sm_100a) — SURVEY.md §8(a) rows a3-a6.
//
// One persistent CTA owns a tile of CT adjacent columns (column = pattern x category) and
// walks the WHOLE tree for it, so every dependency is CTA-local: no grid sync, no flags,
// one launch for both passes.  Per round the CTA executes WN independent node tasks side
// by side: thread t works on (slot t / TS, column t % TS), i.e. slots are packed at thread
// granularity, not warp granularity, so an awkward tile width (68 columns at 1000 taxa x
// 10k patterns on 148 SMs) still fills the lanes.
//
// Operands never wait on HBM/L2 latency inside a round:
//   * a block produced in round p and consumed in round p+1 is read from a shared-memory
//     RING (two parities of ring_cap blocks);
//   * a block consumed later (>= p+2) is read back from global through the STAGE: each
//     thread copies its own column with cp.async one round ahead (own column => no
//     cross-thread hazard; single-buffered: registers are loaded first, then the slot is
//     re-armed);
//   * tip states (1 byte per pattern) travel one round ahead in a register;
//   * the 4x4 transition matrices and the task descriptors of a round are staged
//     cooperatively one round ahead.
// Ring and stage blocks share one layout, so a task descriptor names every operand by a
// shared-memory block offset and the kernel does not branch on where it came from.  The
// host side of the C ABI builds the static schedule and routes operands (b2p_plan.h).
//
// Message formulation: what a node stores is not its partial x_v but the message to its
// parent  m_v = P_v x_v.  The post-order task of v is  x_v = m_l o m_r ; m_v = P_v x_v
// and the pre-order task needs no matrix products for its inputs:
//   q_l = u_v o m_r,  q_r = u_v o m_l,  L = sum q_l.m_l,
//   dL/dt_l = r_k q_l.(Q m_l),  u_l = P_l^T q_l  (same for r).
#pragma once
#include "b2p_device.cuh"

namespace b2p {

constexpr int kMaxCatW = 32;
constexpr unsigned short kNoBlock = 0xFFFFu;  // tip child / root prior / no ring copy
constexpr unsigned kNoFetch = 0xFFFFFFFFu;

enum : unsigned short { TF_ROOT = 1, TF_G0 = 2, TF_G1 = 4, TF_INT_L = 8, TF_INT_R = 16 };

struct WaveTask {  // 32 B: what the compute phase of a round needs
  int l, r;        // child node ids (= branch ids); l < 0: idle slot
  unsigned short src_l, src_r, src_v, dst0, dst1, flags;  // shared-memory block offsets / 16
  unsigned gout0, gout1;  // global column offsets ((node - N) * n_cols) of the outputs
  unsigned pad;
};
static_assert(sizeof(WaveTask) == 32, "WaveTask must be 32 bytes");

struct WaveAhead {  // 32 B: what must be known one round ahead
  unsigned g_l, g_r, g_v;  // block to stage (global column offset) or tip row offset (id * P); kNoFetch = none
  unsigned kinds;          // bit 0: l is a tip, bit 1: r is a tip
  int m[3];                // node ids whose matrices fill the slot's matrices 0..2 (-1 = none)
  int pad;
};
static_assert(sizeof(WaveAhead) == 32, "WaveAhead must be 32 bytes");

template <typename T>
struct WaveParams {
  int n_tips, n_patterns, n_cat, n_cols;
  int ct, ts, cpw, wn;  // tile columns, threads per slot, columns per warp, tasks per round
  int n_tiles, rounds_post, rounds_pre, n_branches;
  int pj, plane, stage_op_stride, pad0;  // partial sums per slot; ct*16; wn*block bytes
  int off_task, off_ahead, off_mats, off_stage, off_part, off_scount;  // smem byte offsets
  const WaveTask* post_task;    // [rounds_post][wn]
  const WaveAhead* post_ahead;  // [rounds_post][wn]
  const WaveTask* pre_task;     // [rounds_pre][wn]
  const WaveAhead* pre_ahead;   // [rounds_pre][wn]
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

// row i of a staged matrix times x
template <typename T>
__device__ __forceinline__ T row_dot(const T* m, int i, const T (&x)[4]) {
  T r[4];
  load_row(m + 4 * i, r);
  return r[0] * x[0] + r[1] * x[1] + r[2] * x[2] + r[3] * x[3];
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


// Per-thread state.  The thread -> (slot, column) map is static, so everything that depends
// only on it is computed once (shared-memory offsets, segment structure of the warp).
template <typename T, int K, int NT>
struct WaveCtx {
  const WaveParams<T>& p;
  unsigned char* smem;
  int t, lane, s;
  int c16;              // 16 * column within the tile
  int k;                // this thread's category
  bool in_slot;         // s < wn and the lane maps to a column of the tile
  bool act;             // in_slot and the column exists
  unsigned colg;        // global column
  int pat;              // global pattern
  double w_p;           // pattern weight
  unsigned same;        // bit i: lane + (1 << i) is in the same slot (segmented reductions)
  bool head;            // first lane of this slot's segment in the warp
  int part_j;           // index of this warp among the warps that touch slot s
  int mat_off;          // smem offset of this thread's (slot, category) matrix 0, parity 0
  int stage_off;        // smem offset of this thread's column in stage block (op 0, slot s)
  int tips;             // tip masks of the coming task: l | r << 8 (register prefetch)
  double ll_acc;
  bool first_tile;
  __device__ __forceinline__ WaveCtx(const WaveParams<T>& p_) : p(p_) {}
  __device__ __forceinline__ int kk() const { return K > 0 ? K : p.n_cat; }
  __device__ __forceinline__ int mats_parity_bytes() const { return p.wn * 3 * kk() * MatBytes<T>::v; }
  __device__ __forceinline__ const T* mat(int parity, int which) const {
    return reinterpret_cast<const T*>(smem + mat_off + parity * mats_parity_bytes() + which * kk() * MatBytes<T>::v);
  }
  __device__ __forceinline__ unsigned char* blk(unsigned off16) const { return smem + off16 * 16 + c16; }
  __device__ __forceinline__ unsigned char* stage(int op) const { return smem + stage_off + op * p.stage_op_stride; }
  __device__ __forceinline__ int part_stride() const { return 16 + 16 * p.pj; }
  __device__ __forceinline__ unsigned char* part(int parity, int slot) const {
    return smem + p.off_part + (parity * p.wn + slot) * part_stride();
  }
  __device__ __forceinline__ int* scount() const { return reinterpret_cast<int*>(smem + p.off_scount); }
};

// Power-of-two rescale shared by the K category lanes of a pattern; returns 1 if applied.
// fp64: partials are non-negative, so the high words order like the values; a denormal
// maximum (high word 0) is left alone — it cannot arise from one step below 2^-200.
template <int K, typename C>
__device__ __forceinline__ int rescale(const C& c, double (&y)[4]) {
  int h = max(max(__double2hiint(y[0]), __double2hiint(y[1])), max(__double2hiint(y[2]), __double2hiint(y[3])));
  h = group_max_int<K>(h, c.lane, c.kk());
  if (h < 0x33700000 && h > 0) {  // < 2^-200
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= Scale<double>::factor();
    return 1;
  }
  return 0;
}
template <int K, typename C>
__device__ __forceinline__ int rescale(const C& c, float (&y)[4]) {
  float mx = fmaxf(fmaxf(y[0], y[1]), fmaxf(y[2], y[3]));
  mx = group_max<K>(mx, c.lane, c.kk());
  if (mx < Scale<float>::threshold() && mx > 0.f) {
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= Scale<float>::factor();
    return 1;
  }
  return 0;
}

// Sum two values over the lanes of this thread's slot inside the warp; valid in `head` lanes.
template <typename C>
__device__ __forceinline__ void seg_sum2(const C& c, double& a, double& b) {
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const double oa = __shfl_down_sync(0xffffffffu, a, 1 << i);
    const double ob = __shfl_down_sync(0xffffffffu, b, 1 << i);
    if ((c.same >> i) & 1) { a += oa; b += ob; }
  }
}

// Partial sums (a, b) of this warp for its slot, and by the slot's first thread the ids
// they belong to; folded one round later by finalize_parts().
template <typename T, int K, int NT>
__device__ __forceinline__ void put_parts(const WaveCtx<T, K, NT>& c, int par, int id_a, int id_b, double a, double b) {
  seg_sum2(c, a, b);
  if (!c.in_slot) return;
  unsigned char* ps = c.part(par, c.s);
  if (c.head) {
    reinterpret_cast<double*>(ps + 16)[c.part_j] = a;
    reinterpret_cast<double*>(ps + 16 + 8 * c.p.pj)[c.part_j] = b;
  }
  if (c.c16 == 0) {
    reinterpret_cast<int*>(ps)[0] = id_a;
    reinterpret_cast<int*>(ps)[1] = id_b;
  }
}

// ids >= 0: branch gradients (a -> id_a, b -> id_b); id_a == -2: a is a log-likelihood term.
template <typename T, int K, int NT>
__device__ __forceinline__ void finalize_parts(WaveCtx<T, K, NT>& c, int par) {
  if (!(c.in_slot && c.c16 == 0)) return;
  const unsigned char* ps = c.part(par, c.s);
  const int id_a = reinterpret_cast<const int*>(ps)[0], id_b = reinterpret_cast<const int*>(ps)[1];
  if (id_a == -1) return;
  const int first_warp = (c.s * c.p.ts) >> 5, last_warp = (c.s * c.p.ts + c.p.ts - 1) >> 5;
  const int nj = last_warp - first_warp + 1;
  const double* pa = reinterpret_cast<const double*>(ps + 16);
  const double* pb = reinterpret_cast<const double*>(ps + 16 + 8 * c.p.pj);
  double a = 0.0, b = 0.0;
  for (int j = 0; j < nj; ++j) { a += pa[j]; b += pb[j]; }
  if (id_a == -2) { c.ll_acc += a; return; }
  double* gp = c.p.gpart + (size_t)blockIdx.x * c.p.n_branches;
  if (c.first_tile) { gp[id_a] = a; gp[id_b] = b; }
  else { gp[id_a] += a; gp[id_b] += b; }
}


// ---- one round ahead -----------------------------------------------------------------
// (a) this thread's own operands of the coming round: global blocks -> stage (cp.async),
//     tip masks -> register; (b) cooperative: the coming round's matrices, the coming
//     round's task descriptors and the look-ahead descriptors of the round after it.
template <int PASS, typename T, int K, int NT>
__device__ __forceinline__ void prefetch_round(WaveCtx<T, K, NT>& c, const WaveAhead* ah_smem, int par_next) {
  const WaveParams<T>& p = c.p;
  if (c.act) {
    const uint4 a = *reinterpret_cast<const uint4*>(ah_smem + c.s);
    int tl = 15, tr = 15;
    if (a.x != kNoFetch) {
      if (a.w & 1u) tl = __ldg(p.tips + a.x + c.pat);
      else s_cp(c.stage(0), p.plane, p.msg + ((size_t)a.x + c.colg) * 4);
    }
    if (a.y != kNoFetch) {
      if (a.w & 2u) tr = __ldg(p.tips + a.y + c.pat);
      else s_cp(c.stage(1), p.plane, p.msg + ((size_t)a.y + c.colg) * 4);
    }
    if (PASS == 1 && a.z != kNoFetch) s_cp(c.stage(2), p.plane, p.pre_g + ((size_t)a.z + c.colg) * 4);
    c.tips = tl | (tr << 8);
  }
  // matrices of the coming round
  constexpr int G = (int)sizeof(T);        // 16-byte granules per 4x4 matrix: 8 fp64, 4 fp32
  constexpr int EPG = 16 / (int)sizeof(T);
  const int kk = c.kk();
  const int per_slot = 3 * kk * G;
  const int total = p.wn * per_slot;
  unsigned char* dst = c.smem + p.off_mats + par_next * c.mats_parity_bytes();
  for (int g = c.t; g < total; g += NT) {
    const int sl = g / per_slot, rem = g - sl * per_slot;
    const int which = rem / (kk * G), rem2 = rem - which * (kk * G);
    const int k_ = rem2 / G, gi = rem2 - k_ * G;
    const int node = ah_smem[sl].m[which];
    if (node >= 0)
      cp_async16(dst + ((sl * 3 + which) * kk + k_) * MatBytes<T>::v + gi * 16,
                 p.pmat + ((size_t)node * kk + k_) * 16 + gi * EPG);
  }
}

// descriptors of one round: global -> shared (2 x 16 B per slot)
template <typename D, typename T, int K, int NT>
__device__ __forceinline__ void copy_descs(const WaveCtx<T, K, NT>& c, D* dst, const D* src_round) {
  const int n16 = c.p.wn * 2;
  for (int g = c.t; g < n16; g += NT)
    cp_async16(reinterpret_cast<int4*>(dst) + g, reinterpret_cast<const int4*>(src_round) + g);
}

struct TaskRegs {
  int l, r;
  unsigned src_l, src_r, src_v, dst0, dst1, flags, gout0, gout1;
};
__device__ __forceinline__ TaskRegs load_task(const WaveTask* tk) {
  const int4 a = *reinterpret_cast<const int4*>(tk);
  const int4 b = *reinterpret_cast<const int4*>(reinterpret_cast<const unsigned char*>(tk) + 16);
  TaskRegs t;
  t.l = a.x; t.r = a.y;
  t.src_l = (unsigned)a.z & 0xffffu; t.src_r = (unsigned)a.z >> 16;
  t.src_v = (unsigned)a.w & 0xffffu; t.dst0 = (unsigned)a.w >> 16;
  t.dst1 = (unsigned)b.x & 0xffffu; t.flags = (unsigned)b.x >> 16;
  t.gout0 = (unsigned)b.y; t.gout1 = (unsigned)b.z;
  return t;
}

// Operands of the current round -> registers (zeros for idle slots / inactive lanes).
template <int PASS, typename T, int K, int NT>
__device__ __forceinline__ void load_operands(const WaveCtx<T, K, NT>& c, const TaskRegs& e, bool live, int par,
                                              T (&ma)[4], T (&mb)[4], T (&u)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) ma[i] = mb[i] = u[i] = (T)0;
  if (!live) return;
  if (e.src_l != kNoBlock) s_ld(c.blk(e.src_l), c.p.plane, ma);
  else tip_message(c.mat(par, 1), c.tips & 0xff, ma);
  if (e.src_r != kNoBlock) s_ld(c.blk(e.src_r), c.p.plane, mb);
  else tip_message(c.mat(par, 2), c.tips >> 8, mb);
  if (PASS == 1) {
    if (e.src_v != kNoBlock) s_ld(c.blk(e.src_v), c.p.plane, u);
    else {
#pragma unroll
      for (int i = 0; i < 4; ++i) u[i] = (T)c.p.freqs[i];
    }
  }
}

// All warp shuffles (category-group and segment reductions) sit on the unconditional path:
// lanes of one warp can belong to different slots (different tasks, idle slots).
template <typename T, int K, int NT>
__device__ __forceinline__ void post_task(WaveCtx<T, K, NT>& c, const TaskRegs& e, bool live, int par,
                                          bool last_round, const T (&ma)[4], const T (&mb)[4]) {
  const WaveParams<T>& p = c.p;
  const bool is_root = (e.flags & TF_ROOT) != 0;
  T y[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) y[i] = ma[i] * mb[i];
  const int bumped = rescale<K>(c, y);
  if (live && !is_root) {
    if (bumped && c.k == 0) atomicAdd(c.scount() + (c.c16 >> 4), 1);
    T mv[4];
    const T* m = c.mat(par, 0);
#pragma unroll
    for (int i = 0; i < 4; ++i) mv[i] = row_dot(m, i, y);
    if (e.dst0 != kNoBlock) s_st(c.blk(e.dst0), p.plane, mv);
    st_col(p.msg + ((size_t)e.gout0 + c.colg) * 4, mv);
  }
  if (!last_round) return;  // uniform over the CTA
  // the root runs alone in the last round, so every rescale count is final
  T site = (T)p.freqs[0] * y[0] + (T)p.freqs[1] * y[1] + (T)p.freqs[2] * y[2] + (T)p.freqs[3] * y[3];
  site = group_sum<K>(site * (T)p.cat_w[c.k], c.lane, c.kk());
  double term = 0.0;
  if (live && is_root && c.k == 0)
    term = c.w_p * (log((double)site) - (double)(c.scount()[c.c16 >> 4] + bumped) * Scale<T>::log_factor());
  put_parts(c, par, (e.l >= 0 && is_root) ? -2 : -1, -1, term, 0.0);
}

template <typename T, int K, int NT>
__device__ __forceinline__ void pre_task(WaveCtx<T, K, NT>& c, const TaskRegs& e, bool live, int par, T (&ma)[4],
                                         T (&mb)[4], const T (&u)[4]) {
  const WaveParams<T>& p = c.p;
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
  if (K != 1) {
    const T cw = (T)p.cat_w[c.k], cwr = (T)(p.cat_w[c.k] * p.rates[c.k]);
    den = group_sum<K>(den * cw, c.lane, c.kk());
    numl = group_sum<K>(numl * cwr, c.lane, c.kk());
    numr = group_sum<K>(numr * cwr, c.lane, c.kk());
  }
  double gl = 0.0, gr = 0.0;
  if (live && c.k == 0) {
    // K == 1: the single category's weight cancels, its rate multiplies the numerators
    const double inv = (K == 1 ? c.w_p * p.rates[0] : c.w_p) / (double)den;
    gl = (double)numl * inv;
    gr = (double)numr * inv;
  }
  put_parts(c, par, e.l, e.r, gl, gr);
  // pre-order partials of the internal children: u_c = P_c^T q_c  (ma / mb are dead)
  const bool want_l = live && (e.flags & TF_INT_L), want_r = live && (e.flags & TF_INT_R);
  if (want_l) matvec_t_rows(c.mat(par, 1), ql, ma);
  (void)rescale<K>(c, ma);
  if (want_l) {
    if (e.dst0 != kNoBlock) s_st(c.blk(e.dst0), p.plane, ma);
    if (e.flags & TF_G0) st_col(p.pre_g + ((size_t)e.gout0 + c.colg) * 4, ma);
  }
  if (want_r) matvec_t_rows(c.mat(par, 2), qr, mb);
  (void)rescale<K>(c, mb);
  if (want_r) {
    if (e.dst1 != kNoBlock) s_st(c.blk(e.dst1), p.plane, mb);
    if (e.flags & TF_G1) st_col(p.pre_g + ((size_t)e.gout1 + c.colg) * 4, mb);
  }
}


// One pass over its schedule.  Round r: (1) task descriptor + operands -> registers,
// (2) re-arm stage / matrices / descriptors for round r+1, (3) compute and store,
// (4) fold the previous round's partial sums, (5) cp.async.wait_all + one CTA barrier.
// Descriptor rings: task[r] lives in parity r&1 (loaded during round r-1), ahead[r+1] in
// parity (r+1)&1 (loaded during round r-1, consumed by the prefetch of round r).
template <int PASS, typename T, int K, int NT>
__device__ __forceinline__ void wave_pass(WaveCtx<T, K, NT>& c) {
  const WaveParams<T>& p = c.p;
  const WaveTask* g_task = PASS ? p.pre_task : p.post_task;
  const WaveAhead* g_ahead = PASS ? p.pre_ahead : p.post_ahead;
  const int R = PASS ? p.rounds_pre : p.rounds_post;
  WaveTask* tring = reinterpret_cast<WaveTask*>(c.smem + p.off_task);     // [2][wn]
  WaveAhead* aring = reinterpret_cast<WaveAhead*>(c.smem + p.off_ahead);  // [2][wn]
  // prologue: descriptors of rounds 0 (and ahead[1]), then everything round 0 needs
  copy_descs(c, tring, g_task);
  copy_descs(c, aring, g_ahead);
  if (R > 1) copy_descs(c, aring + p.wn, g_ahead + p.wn);
  cp_async_wait_all();
  __syncthreads();
  prefetch_round<PASS>(c, aring, 0);
  cp_async_wait_all();
  __syncthreads();
  for (int r = 0; r < R; ++r) {
    const int par = r & 1;
    TaskRegs e;
    e.l = -1; e.r = -1; e.src_l = e.src_r = e.src_v = e.dst0 = e.dst1 = kNoBlock; e.flags = 0; e.gout0 = e.gout1 = 0;
    if (c.in_slot) e = load_task(tring + par * p.wn + c.s);
    const bool live = e.l >= 0 && c.act;
    T ma[4], mb[4], u[4];
    load_operands<PASS>(c, e, live, par, ma, mb, u);
    if (r + 1 < R) {
      prefetch_round<PASS>(c, aring + (par ^ 1) * p.wn, par ^ 1);
      copy_descs(c, tring + (par ^ 1) * p.wn, g_task + (size_t)(r + 1) * p.wn);
      if (r + 2 < R) copy_descs(c, aring + par * p.wn, g_ahead + (size_t)(r + 2) * p.wn);
    }
    if (PASS == 0) post_task(c, e, live, par, r == R - 1, ma, mb);
    else pre_task(c, e, live, par, ma, mb, u);
    if (PASS == 1 && r > 0) finalize_parts(c, par ^ 1);
    cp_async_wait_all();
    __syncthreads();
  }
  finalize_parts(c, (R - 1) & 1);
  __syncthreads();
}

template <typename T, int K, int NT, int MINB, bool GRAD>
__global__ void __launch_bounds__(NT, MINB) wave_kernel(const WaveParams<T> p) {
  extern __shared__ __align__(16) unsigned char wave_smem[];
  __shared__ double red[NT / 32];
  WaveCtx<T, K, NT> c(p);
  c.smem = wave_smem;
  c.t = threadIdx.x; c.lane = c.t & 31;
  const int warp = c.t >> 5;
  const int kk = c.kk();
  c.s = c.t / p.ts;
  const int tt = c.t - c.s * p.ts;
  int col = (tt >> 5) * p.cpw + (tt & 31);
  c.in_slot = c.s < p.wn && (tt & 31) < p.cpw && col < p.ct;
  if (!c.in_slot) col = 0;
  c.c16 = col * 16;
  c.k = col % kk;
  // segment structure of the warp (static)
  c.same = 0;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int so = __shfl_down_sync(0xffffffffu, c.s, 1 << i);
    if (c.lane + (1 << i) < 32 && so == c.s) c.same |= 1u << i;
  }
  const int sprev = __shfl_up_sync(0xffffffffu, c.s, 1);
  c.head = c.lane == 0 || sprev != c.s;
  c.part_j = warp - ((c.s * p.ts) >> 5);
  const int slot = c.in_slot ? c.s : 0;
  c.mat_off = p.off_mats + (slot * 3 * kk + c.k) * MatBytes<T>::v;
  c.stage_off = p.off_stage + slot * (p.ct * 4 * (int)sizeof(T)) + c.c16;
  c.tips = 15 | (15 << 8);
  c.ll_acc = 0.0;
  c.first_tile = true;

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, c.first_tile = false) {
    const int cg = tile * p.ct + col;
    c.act = c.in_slot && cg < p.n_cols;
    c.colg = c.act ? (unsigned)cg : 0u;
    c.pat = (int)c.colg / kk;
    c.w_p = c.act ? __ldg(p.weights + c.pat) : 0.0;
    for (int i = c.t; i < p.ct; i += NT) c.scount()[i] = 0;
    __syncthreads();
    wave_pass<0>(c);
    if (GRAD) wave_pass<1>(c);
  }

  double v = warp_sum(c.ll_acc);
  if (c.lane == 0) red[warp] = v;
  __syncthreads();
  if (c.t == 0) {
    double s = 0.0;
    for (int w = 0; w < NT / 32; ++w) s += red[w];
    p.lpart[blockIdx.x] = s;
  }
}

}  // namespace b2p
