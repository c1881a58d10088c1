// Traversal kernel (sm_100a) — SURVEY.md §8(a) rows a3-a6: post-order partials, root
// log-likelihood, pre-order partials, branch gradients, in ONE launch.
//
// One persistent CTA owns a tile of ct adjacent columns (column = pattern x category) and
// walks the WHOLE tree for it, so every dependency is CTA-local: no grid sync, no flags.
// Per round the CTA runs up to wn independent node tasks side by side; thread t works on
// slot t / ts and on column t % ts of the tile (with two columns per thread: t % ts and
// ts + t % ts) — slots are packed at thread granularity so an awkward tile width (68
// columns at 1000 taxa x 10k patterns on 148 SMs) still fills the lanes.  The planner lays
// the tasks of a round out by kind (no / one / two tip children), so the two slots a warp
// straddles mostly take the same branches.
//
// Data movement is entirely asynchronous and issued by the copy warps (the last warps of
// the CTA, which hold no compute threads) from lists the host planner wrote (b2p_plan.h):
// every operand block, tip-code row, matrix bundle and descriptor record of round q + depth
// is a cp.async.bulk (TMA, UBLKCP in SASS) global -> shared copy that completes on the
// mbarrier of that round; every result block is written to shared memory by the compute
// threads and leaves with a cp.async.bulk shared -> global copy.  Compute threads execute no
// global-memory instruction at all.  Shared memory is a pool of blocks allocated statically
// by the planner (the schedule is static, so lifetimes are known): a value that is re-read
// soon simply stays where it was produced.
//
// Message formulation: a node stores the message to its parent, m_v = P_v x_v.
//   post-order task of v:  x_v = m_l o m_r ;  m_v = P_v x_v
//   pre-order task of v:   q_l = u_v o m_r,  q_r = u_v o m_l,  L = q_l . m_l,
//                          dL/dt_l = r_k q_l.(Q m_l),  u_l = P_l^T q_l   (same for r)
// A tip child's message is a column of P: K1 writes, per tip branch, a 5-row table (row s =
// column s of P, row 4 = all ones for a gap) and the thread reads row `code` — one 32-byte
// read, the same instructions as an internal child's block read with another address.
// Everything a thread needs for a round comes as byte offsets in a 32-byte descriptor
// (two LDS.128) and the tip codes sit at fixed per-slot positions, so the dependent chain at
// the top of a round is: {descriptor, tip codes} -> operands -> arithmetic.
// Gradient sums over a slot's columns go through shared memory and are folded one round
// later by the copy warps (fixed order: deterministic).
#pragma once
#include "b2p_device.cuh"
#include "b2p_walk_types.h"

namespace b2p {

constexpr int kMaxCatW = 32;

template <typename T>
struct WalkParams {
  int n_tiles, rounds_run, rounds_post, n_branches, n_cat;
  int n_cols, n_patterns, n_copy_warps, wn;
  int ct, ts, cpw, stages, depth;
  int blk, plane, ct_pad, row_stride;
  int frame_bytes, off_frame, rel_cr, rel_ir, rel_tips, ir_jobs;  // ir_jobs: first job, relative to the issue record
  int off_bar, off_gbuf, off_scount, off_freq;
  unsigned boot_base, tail_base;  // tail_base: fold ids of the last round that runs
  unsigned boot_n[4], boot_tx[4];
  unsigned long long tile_stride_part;  // (N-1) * blk
  unsigned long long tile_stride_tips;  // rounds * 2 wn * ct_pad
  const unsigned char* records;  // round records (descriptors | issue record | matrix rows written by K1), tail, boot jobs
  unsigned char* msg;         // [tile][N-1] blocks: m_v
  unsigned char* pre;         // [tile][N-1] blocks: u_c that had to leave shared memory
  const unsigned char* tips;  // [tile][round][wn][2][ct_pad] state codes per column
  const double* weights;      // [P]
  double* lpart;              // [grid]
  double* gpart;              // [grid][2N-2]
  double* opart;              // outer-product mode: [grid][2N-2][K][16] per-CTA sums of (w / L) c_k q m^T
  int off_outer, outer_row;   // outer-product mode: shared-memory rows of the compute warps, [2 parities][warps][2][K * 32] doubles
  int off_ostage;             // outer-product mode: [warps] 4 KB staging rows of the tensor-core sums
  long long* trace;           // debug build (-DB2P_TRACE_CLOCKS): [2][rounds][8] clock stamps of CTA 0
  int dbg;                    // debug build (-DB2P_DBG): 1 skip the fold, 2 skip the stores, 4 skip the tasks; outer-product
                              // mode: 8 skip its fold, 16 skip its sums, 32 its tensor-core products only, 64 its staging
                              // stores only (timing experiments: results are wrong by design)
  T q[16];
  double freqs[4];
  double cat_w[kMaxCatW];
  double rates[kMaxCatW];
};

// ---- mbarrier / bulk-copy PTX ----------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "W_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra D_%=;\n"
      "bra W_%=;\n"
      "D_%=:\n"
      "}\n" ::"r"(bar), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_load(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void bulk_store(void* dst, unsigned src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- shared memory through 32-bit shared addresses ---------------------------------------
__device__ __forceinline__ uint4 lds_u4(unsigned a) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ unsigned lds_u8(unsigned a) {
  unsigned v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ int2 lds_i2(unsigned a) {
  int2 v;
  asm volatile("ld.shared.v2.s32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ double2 lds_d2(unsigned a) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_d2(unsigned a, double x, double y) {
  asm volatile("st.shared.v2.f64 [%0], {%1,%2};" ::"r"(a), "d"(x), "d"(y) : "memory");
}
// Four values whose halves sit at a and a + half (fp64; a matrix / table row: half = 16, a
// column of a block: half = the plane stride) or one 16-byte unit (fp32; half unused).
__device__ __forceinline__ void lds_vec(unsigned a, unsigned half, double (&x)[4]) {
  const double2 u = lds_d2(a), v = lds_d2(a + half);
  x[0] = u.x; x[1] = u.y; x[2] = v.x; x[3] = v.y;
}
__device__ __forceinline__ void lds_vec(unsigned a, unsigned, float (&x)[4]) {
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x[0]), "=f"(x[1]), "=f"(x[2]), "=f"(x[3]) : "r"(a));
}
__device__ __forceinline__ void sts_vec(unsigned a, unsigned half, const double (&x)[4]) {
  sts_d2(a, x[0], x[1]);
  sts_d2(a + half, x[2], x[3]);
}
__device__ __forceinline__ void sts_vec(unsigned a, unsigned, const float (&x)[4]) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(x[0]), "f"(x[1]), "f"(x[2]), "f"(x[3]) : "memory");
}

// A value the compiler must keep in a register: it cannot re-derive it from its definition
// (kernel parameters, %tid, the shared window base) inside the round loop, where every such
// re-derivation is a constant-bank load or an S2R on the critical path of the round.
__device__ __forceinline__ unsigned pinned(unsigned x) {
  asm volatile("mov.u32 %0, %0;" : "+r"(x));
  return x;
}

// Power-of-two rescale shared by the K category lanes of a pattern; returns the number of
// factors applied (0 almost always).  The count comes from the exponent of the group's maximum,
// so ONE call brings any positive value back above the threshold — a product of two messages
// whose large entries sit on different states can be far more than one factor below it (fp32:
// factor 2^40 against a range of 2^-126).  Partials are non-negative, so
// the raw bits order like the values.  Zero and denormal maxima (exponent field 0) are left
// alone: the pattern's likelihood is zero or about to be.
// `on`: this lane holds a value (lanes without one take part in the group maximum with 0 —
// for K != 1 every lane of the warp must call this: full-warp shuffles inside).
// The multiplication is rare: it sits in a loop so that it stays a branch instead of four
// predicated multiplies issued every time.
template <int K>
__device__ __forceinline__ int rescale(bool on, int lane, int kk, double (&y)[4]) {
  int h = max(max(__double2hiint(y[0]), __double2hiint(y[1])), max(__double2hiint(y[2]), __double2hiint(y[3])));
  h = group_max_int<K>(on ? h : 0, lane, kk);
  int need = 0;
#pragma unroll 1
  for (unsigned z = pinned((unsigned)(on && h < 0x33700000 && h >= 0x00100000)); z != 0; --z) {  // < 2^-200, normal
    need = (823 - (h >> 20) + 199) / 200;  // biased exponent against 823 = 1023 - 200; at most 5
    const double f = __hiloint2double((1023 + 200 * need) << 20, 0);  // factor^need, one multiplication
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= f;
  }
  return need;
}
template <int K>
__device__ __forceinline__ int rescale(bool on, int lane, int kk, float (&y)[4]) {
  int h = max(max(__float_as_int(y[0]), __float_as_int(y[1])), max(__float_as_int(y[2]), __float_as_int(y[3])));
  h = group_max_int<K>(on ? h : 0, lane, kk);
  int need = 0;
#pragma unroll 1
  for (unsigned z = pinned((unsigned)(on && h < 0x2b800000 && h >= 0x00800000)); z != 0; --z) {  // < 2^-40, normal
    need = (87 - (h >> 23) + 39) / 40;  // biased exponent against 87 = 127 - 40; at most 3
    const float f = __int_as_float((127 + 40 * need) << 23);  // factor^need, one multiplication
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= f;
  }
  return need;
}

// 1 / x for a positive, normal x (a site likelihood, rescaled to stay far from the
// denormals): hardware seed + two Newton steps, relative error ~1e-16, no slow path.
__device__ __forceinline__ double fast_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  return r;
}

// Per-thread constants (all shared-memory addresses are 32-bit).
template <typename T, int CPT>
struct WalkThread {
  unsigned sc;     // shared base + 16 * column: a block's byte offset is added to this
  unsigned dB;     // 16 * ts: the thread's next column inside a block
  unsigned plane;  // fp64: bytes between the two halves of a column
  unsigned km;     // shared base + k * row_bytes: a row's byte offset is added to this
  unsigned rs;     // bytes between rows
  int lane, kk, k, ts;
  T cw, cwr;       // K != 1: category weight, category weight x rate
  double w[CPT];   // pattern weights of the thread's columns (0 for padding columns)
};

// The 32-byte descriptor as the planner wrote it (b2p_walk_types.h: WalkDesc), unpacked.
struct TaskDesc {
  unsigned a_l, a_r, a_v, dst0, dst1, m_l, m_r, flags;
  unsigned id_l, id_r;  // pre-order: the children's branch ids (used by the outer-product mode only)
};
__device__ __forceinline__ TaskDesc unpack_desc(const uint4 d0, const uint4 d1) {
  TaskDesc d;
  d.a_l = d0.x; d.a_r = d0.y; d.a_v = d0.z; d.dst0 = d0.w;
  d.dst1 = d1.x;
  d.m_l = (d1.y & 0xffffu) * 16u; d.m_r = (d1.y >> 16) * 16u;
  d.flags = d1.z & 0xffffu;
  d.id_l = d1.w & 0xffffffu; d.id_r = (d1.w >> 24) | ((d1.z >> 16) << 8);
  return d;
}

// A child's message for one column -> registers.  tip: row `code` of its table (code from the
// round's tip rows); otherwise the thread's column of the child's block (oc: byte offset of the
// column inside a block, relative to the thread's first one).  Same instructions either way.
template <typename T, int CPT>
__device__ __forceinline__ void fetch_child(const WalkThread<T, CPT>& th, unsigned a, bool tip, unsigned code, unsigned oc,
                                            T (&m)[4]) {
  const unsigned at = tip ? th.km + a + min(code, kTipAny) * th.rs : th.sc + a + oc;
  lds_vec(at, tip ? 16u : th.plane, m);
  if (tip && code >= kTipAmbiguous) {  // rare: an ambiguity code other than "any": sum the compatible rows
#pragma unroll
    for (int i = 0; i < 4; ++i) m[i] = (T)0;
#pragma unroll 1
    for (int j = 0; j < 4; ++j)
      if ((code >> j) & 1u) {
        T r[4];
        lds_vec(th.km + a + j * th.rs, 16u, r);
#pragma unroll
        for (int i = 0; i < 4; ++i) m[i] += r[i];
      }
  }
}

// Control flow of the task functions.  K == 1 has no exchange between lanes: only lanes with
// a live task call them (`live` is true).  K != 1 exchanges over the K category lanes of a
// pattern with FULL-WARP shuffles, so every lane of the warp makes the same calls: lanes without
// a task (`live` false) run on whatever their zero descriptor points at — valid shared memory —
// enter the reductions with neutral values and store nothing.
// A thread's CPT columns belong to the same task and the same category, so they share every
// matrix row: the operands are fetched column by column (short register lifetimes), the
// products with P run over all columns with each row read once (at K = 4 a row read costs the
// shared-memory pipe as much as a whole column read: four distinct addresses per warp).

// y[c] = M x[c] (TRANS false: rows of M at `at`) or y[c] = M^T x[c] for the thread's columns.
template <typename T, int CPT, bool TRANS>
__device__ __forceinline__ void rows_times(const WalkThread<T, CPT>& th, unsigned at, const T (&x)[CPT][4], T (&y)[CPT][4]) {
  if (TRANS) {
#pragma unroll
    for (int c = 0; c < CPT; ++c)
#pragma unroll
      for (int j = 0; j < 4; ++j) y[c][j] = (T)0;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    T r[4];
    lds_vec(at + i * th.rs, 16u, r);
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      if (TRANS) {
#pragma unroll
        for (int j = 0; j < 4; ++j) y[c][j] += r[j] * x[c][i];
      } else {
        y[c][i] = r[0] * x[c][0] + r[1] * x[c][1] + r[2] * x[c][2] + r[3] * x[c][3];
      }
    }
  }
}

// ------------------------------------------------------------------ post-order task
//   x_v = m_l o m_r (rescaled when tiny) ;  m_v = P_v x_v ; root: site likelihood -> log-likelihood term
// root_round: this warp runs the root's task (warp-uniform: the root runs alone in its round).
template <typename T, int K, int CPT>
__device__ __forceinline__ void post_task(const WalkParams<T>& p, const WalkThread<T, CPT>& th, const TaskDesc& d, bool live,
                                          bool root_round, int* scount, int col, const unsigned (&code_l)[CPT],
                                          const unsigned (&code_r)[CPT], double& term) {
  T x[CPT][4];
  int bumped[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    T mb[4];
    fetch_child<T, CPT>(th, d.a_l, (d.flags & WF_TIP_L) != 0, code_l[c], c * th.dB, x[c]);
    fetch_child<T, CPT>(th, d.a_r, (d.flags & WF_TIP_R) != 0, code_r[c], c * th.dB, mb);
#pragma unroll
    for (int i = 0; i < 4; ++i) x[c][i] *= mb[i];
    bumped[c] = rescale<K>(live, th.lane, th.kk, x[c]);
  }
  if (!root_round) {
    T mv[CPT][4];
    rows_times<T, CPT, false>(th, th.km + d.a_v, x, mv);
    if (live) {
#pragma unroll
      for (int c = 0; c < CPT; ++c) {
        sts_vec(th.sc + d.dst0 + c * th.dB, th.plane, mv[c]);
        if (bumped[c] && th.k == 0) atomicAdd(scount + col + c * th.ts, bumped[c]);
      }
    }
  } else {
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      T site = (T)p.freqs[0] * x[c][0] + (T)p.freqs[1] * x[c][1] + (T)p.freqs[2] * x[c][2] + (T)p.freqs[3] * x[c][3];
      if (K != 1) site = group_sum<K>(site * th.cw, th.lane, th.kk);
      if (live && th.k == 0 && th.w[c] != 0.0)
        term += th.w[c] * (log((double)site) - (double)(scount[col + c * th.ts] + bumped[c]) * Scale<T>::log_factor());
    }
  }
}

// ------------------------------------------------------------------- pre-order task
//   q_l = u_v o m_r, q_r = u_v o m_l ; L = q_l . m_l ; dL/dt_l = r_k q_l.(Q m_l) ; u_l = P_l^T q_l (same for r)
// OUTER (b2p_branch_outer_products): the pre-order pass also sums, per task and category, the
// 4x4 outer products
//   (w / L) c_k q_l m_l^T  and  (w / L) c_k q_r m_r^T
// over the task's columns — what every model-parameter gradient is a contraction of
// (b200phylo/paramgrad.py).  One column per thread; slots are packed at thread granularity as
// usual, so a warp holds columns of one slot or of two neighbours ("A" below its boundary lane,
// "B" from it on; ts >= 32 and ts % K == 0, so a lane's category is lane % K).  Per warp the sum
// over columns runs on the fp64 tensor cores (outer_sums) and lands in the warp's rows of the
// round's parity; per slot, the warps that have it as their slot A add the rows of all its warps
// (fixed order) into the CTA's part of `opart` at the TOP of the next round (outer_fold) — the
// round's barrier has made the rows complete, the rows of the other parity are being written,
// and the additions overlap the wait for the next operands instead of sitting at a round's tail.
struct OuterSlot {
  unsigned row0;    // shared-memory row of the slot's first warp for this slot (parity 0)
  unsigned rows1;   // the A row of the slot's second warp (further warps: 2 rows on)
  int n_warps;      // warps holding columns of the slot
  int n_fold, r;    // warps that have the slot as their A slot, this warp's index among them
};
struct OuterCtx {
  unsigned st[4];   // where this lane stores the four 16-byte chunks of its column's A values (B: + 2048)
  unsigned ld[4];   // where this lane loads its A element of product d from: ld[d % 4] + (d / 4) * 1024 (B: + 2048)
  unsigned row;     // this warp's A row ([K][32] doubles) of parity 0; its B row follows
  unsigned par;     // bytes from a row of parity 0 to the same row of parity 1
  int bl;           // boundary lane: lanes below belong to slot A, the others to slot B (32: no slot B)
  OuterSlot a;      // the slot of the warp's lane 0
  double* part;     // the CTA's part of opart
  // what the warp's slot A did in the round just finished (warp-uniform): set by pre_task, consumed by outer_fold
  bool pending;
  unsigned pend_l, pend_r, pend_par;
};

// Sum over the warp's columns of (a q)(m)^T: with the columns as the contraction index,
// A = [a q_l | a q_r]^T (8 x columns) and B = [m_l | m_r] (columns x 8) give C = A B whose
// diagonal 4x4 blocks are the two outer-product sums (the off-diagonal blocks are ignored).
// Every lane puts its column's 8 + 8 values into the warp's staging rows; each mma.sync (m8n8k4)
// then takes four columns of one category.  Products whose four columns all sit on one side of
// the boundary lane go to that slot's accumulator; a mixed one runs twice with the other slot's
// columns zeroed.
// Staging layout: a column (lane) has a 64-byte row of A values and, 2048 bytes on, one of B
// values.  Row of lane l: l ^ ((l >> 2) & 1) — neighbours of a category alternate bank halves —
// and the row's 16-byte chunks are XOR-swizzled, so that the stores (8 lanes x 16 bytes per
// phase) and the loads (four rows x 32 bytes per half-warp, at K = 4 rows 8 apart) are both
// conflict-free.  Product d takes, at K = 1, columns 4 d .. 4 d + 3; at K = 4, category d % 4,
// columns d % 4 + 4 (t + 4 (d / 4)), t = 0..3; lane (g, t) loads element g of column t of the
// four.  All of this depends on the lane only: outer_addresses runs once per kernel.
__device__ __forceinline__ unsigned outer_swizzle(unsigned row) { return ((((row >> 1) ^ (row >> 3)) & 1u) << 1) | ((row >> 2) & 1u); }
template <int K>
__device__ __forceinline__ unsigned outer_column(int d, unsigned t) {
  return K == 1 ? 4u * d + t : (unsigned)(d & 3) + 4u * (t + 4u * (unsigned)(d >> 2));
}
template <int K>
__device__ __forceinline__ void outer_addresses(OuterCtx& oc, unsigned stage, int lane) {
  const unsigned row = (unsigned)lane ^ (((unsigned)lane >> 2) & 1u), sw = outer_swizzle(row);
  const unsigned g = (unsigned)lane >> 2, t = (unsigned)lane & 3u;
#pragma unroll
  for (unsigned c = 0; c < 4; ++c) {
    oc.st[c] = stage + row * 64 + ((c ^ sw) << 4);
    const unsigned col = outer_column<K>((int)c, t), r2 = col ^ ((col >> 2) & 1u);
    oc.ld[c] = stage + r2 * 64 + (((g >> 1) ^ outer_swizzle(r2)) << 4) + ((g & 1u) << 3);  // products c + 4: 16 rows = 1024 bytes on
  }
}

template <int K>
__device__ __forceinline__ void outer_sums(const OuterCtx& oc, unsigned row, int lane, const double (&aq)[8], const double (&m)[8],
                                           int dbg = 0) {
#ifdef B2P_DBG
  if (!(dbg & 64))
#endif
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    sts_d2(oc.st[c], aq[2 * c], aq[2 * c + 1]);
    sts_d2(oc.st[c] + 2048, m[2 * c], m[2 * c + 1]);
  }
  __syncwarp();
  // all sixteen loads first, then the products: two independent accumulation chains per category
  double av[8], bv[8];
#pragma unroll
  for (int d = 0; d < 8; ++d) {
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(av[d]) : "r"(oc.ld[d & 3] + (d >> 2) * 1024));
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(bv[d]) : "r"(oc.ld[d & 3] + (d >> 2) * 1024 + 2048));
  }
  constexpr int NA = K == 1 ? 2 : K;  // accumulators per slot: K = 1 splits its products into two chains
  double ca[NA][2], cb[NA][2];
#pragma unroll
  for (int k = 0; k < NA; ++k) ca[k][0] = ca[k][1] = cb[k][0] = cb[k][1] = 0.0;
#ifdef B2P_DBG
  if (dbg & 32) {
#pragma unroll
    for (int d = 0; d < 8; ++d) ca[d % NA][0] += av[d] * bv[d];
  } else
#endif
#pragma unroll
  for (int d = 0; d < 8; ++d) {
    const int ai = K == 1 ? (d & 1) : (d & 3);
    const int first = (int)outer_column<K>(d, 0u), last = (int)outer_column<K>(d, 3u);  // warp-uniform branches below
    if (last < oc.bl) {
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(ca[ai][0]), "+d"(ca[ai][1]) : "d"(av[d]), "d"(bv[d]));
    } else if (first >= oc.bl) {
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(cb[ai][0]), "+d"(cb[ai][1]) : "d"(av[d]), "d"(bv[d]));
    } else {  // the boundary lane is among the four columns: once per slot, the other slot's columns zeroed
      const bool in_a = (int)outer_column<K>(d, (unsigned)lane & 3u) < oc.bl;
      const double xa = in_a ? av[d] : 0.0, xb = in_a ? 0.0 : av[d];
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(ca[ai][0]), "+d"(ca[ai][1]) : "d"(xa), "d"(bv[d]));
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(cb[ai][0]), "+d"(cb[ai][1]) : "d"(xb), "d"(bv[d]));
    }
  }
  // lane (g, t) holds C[g][2t], C[g][2t+1]: rows 0-3 x columns 0-3 = the left branch's matrix
  // (entries 0-15 of the row), rows 4-7 x columns 4-7 = the right branch's (16-31)
  const unsigned g = (unsigned)lane >> 2, t = (unsigned)lane & 3u;
  if ((g < 4) == (t < 2)) {
    const unsigned e = (g & 3u) * 4u + (t & 1u) * 2u + (g < 4 ? 0u : 16u);
#pragma unroll
    for (int k = 0; k < K; ++k) {
      sts_d2(row + (k * 32 + e) * 8, K == 1 ? ca[0][0] + ca[1][0] : ca[k][0], K == 1 ? ca[0][1] + ca[1][1] : ca[k][1]);
      if (oc.bl < 32)
        sts_d2(row + (K * 32 + k * 32 + e) * 8, K == 1 ? cb[0][0] + cb[1][0] : cb[k][0], K == 1 ? cb[0][1] + cb[1][1] : cb[k][1]);
    }
  }
}

// Slot A's sums of the round just finished -> the CTA's part of opart (top of the next round, or
// after a tile's last round).  Warp-uniform; the rows were completed by the round's barrier.
template <int K>
__device__ __forceinline__ void outer_fold(OuterCtx& oc, int lane, bool first_tile, int outer_row, int dbg = 0) {
  if (!oc.pending) return;
  oc.pending = false;
#ifdef B2P_DBG
  if (dbg & 8) return;
#endif
  const OuterSlot& sl = oc.a;
  const unsigned row0 = sl.row0 + oc.pend_par, rows1 = sl.rows1 + oc.pend_par;
  // A lane adds one PAIR of adjacent entries per step (16-byte loads and stores): pair j of the
  // slot's K * 16 sits at byte 16 j of every row; the slot's A warps take 32 pairs each per step.
  for (int j = sl.r * 32 + lane; j < K * 16; j += sl.n_fold * 32) {
    const unsigned off = j * 16;
    double2 v = lds_d2(row0 + off);
    const double2 x = lds_d2((sl.n_warps > 1 ? rows1 : row0) + off);
    if (sl.n_warps > 1) { v.x += x.x; v.y += x.y; }
    for (int r = 2; r < sl.n_warps; ++r) {
      const double2 y = lds_d2(rows1 + (r - 1) * 2 * outer_row + off);
      v.x += y.x; v.y += y.y;
    }
    const int k = j >> 4, e = (j & 15) * 2;  // entries e, e + 1 of category k: 0-15 the left branch, 16-31 the right
    double* gp = oc.part + ((size_t)(e < 16 ? oc.pend_l : oc.pend_r) * K + k) * 16 + (e & 15);
    // a branch is written once per tile: plain store for the CTA's first tile, then adds in tile order
    if (first_tile) *reinterpret_cast<double2*>(gp) = v;
    else { atomicAdd(gp, v.x); atomicAdd(gp + 1, v.y); }
  }
}

template <typename T, int K, int CPT, bool OUTER = false>
__device__ __forceinline__ void pre_task(const WalkParams<T>& p, const WalkThread<T, CPT>& th, const TaskDesc& d, bool live,
                                         const unsigned (&code_l)[CPT], const unsigned (&code_r)[CPT], double& gl, double& gr,
                                         OuterCtx* ocp = nullptr, unsigned par = 0u) {
  T ql[CPT][4], qr[CPT][4];
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    T ma[4], mb[4], u[4];
    fetch_child<T, CPT>(th, d.a_l, (d.flags & WF_TIP_L) != 0, code_l[c], c * th.dB, ma);
    fetch_child<T, CPT>(th, d.a_r, (d.flags & WF_TIP_R) != 0, code_r[c], c * th.dB, mb);
    lds_vec(th.sc + d.a_v + c * th.dB, th.plane, u);
    T den = (T)0, numl = (T)0, numr = (T)0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      ql[c][i] = u[i] * mb[i];
      qr[c][i] = u[i] * ma[i];
      den += ql[c][i] * ma[i];
    }
    // d/dt (P x) = r Q (P x): Q comes from the constant bank.  (Outer-product mode: the branch
    // gradient is the contraction sum_k r_k <Q, G_{b,k}> of what it sums anyway — done after the
    // walk, grad_from_outer_kernel — so the numerators are not formed here.)
    if constexpr (!OUTER) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const T dl = p.q[4 * i] * ma[0] + p.q[4 * i + 1] * ma[1] + p.q[4 * i + 2] * ma[2] + p.q[4 * i + 3] * ma[3];
        const T dr = p.q[4 * i] * mb[0] + p.q[4 * i + 1] * mb[1] + p.q[4 * i + 2] * mb[2] + p.q[4 * i + 3] * mb[3];
        numl += ql[c][i] * dl;
        numr += qr[c][i] * dr;
      }
    }
    if (K != 1) {
      // the site likelihood is the sum over the pattern's K category lanes; the numerators
      // need no such exchange: every lane adds its own category's share and the fold sums
      // over all lanes of the slot anyway
      den = group_sum<K>(den * th.cw, th.lane, th.kk);
      numl *= th.cwr;
      numr *= th.cwr;
    }
    // K == 1: the single category's weight cancels, its rate multiplies the numerators.  A
    // pattern of likelihood zero contributes nothing instead of inf * 0.
    const double dd = (double)den;
    const double inv = (live && dd > 0.0) ? (K == 1 ? th.w[c] * p.rates[0] : th.w[c]) * fast_rcp(dd) : 0.0;
    gl += live ? (double)numl * inv : 0.0;
    gr += live ? (double)numr * inv : 0.0;
#ifdef B2P_DBG
    if (!(p.dbg & 16))
#endif
    if constexpr (OUTER) {
      static_assert(!OUTER || (CPT == 1 && (K == 1 || K == 4)), "outer-product mode: one column per thread, K = 1 or 4");
      const double a = (live && dd > 0.0) ? (K == 1 ? th.w[c] : th.w[c] * (double)th.cw) * fast_rcp(dd) : 0.0;
      double aq[8], m[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        aq[i] = a * (double)ql[c][i]; aq[4 + i] = a * (double)qr[c][i];
        m[i] = (double)ma[i]; m[4 + i] = (double)mb[i];
      }
      // lanes without a column: zeros (their operands are whatever their zero descriptor points
      // at — 0 x NaN would spoil the sum).  Rare, so a branch rather than sixteen selects.
#pragma unroll 1
      for (unsigned z = pinned((unsigned)!live); z != 0; --z)
#pragma unroll
        for (int i = 0; i < 8; ++i) m[i] = 0.0;
      outer_sums<K>(*ocp, ocp->row + par * ocp->par, th.lane, aq, m, p.dbg);
    }
  }
  // pre-order partials of the internal children: u_c = P_c^T q_c (rescaled when tiny).  K == 1: a
  // tip child skips the product; K != 1: every lane computes it (the rescale inside is a warp-wide
  // exchange) and only lanes with an internal child keep it.
  const bool out_l = live && !(d.flags & WF_TIP_L), out_r = live && !(d.flags & WF_TIP_R);
  if (K != 1 || out_l) {
    T uc[CPT][4];
    rows_times<T, CPT, true>(th, th.km + d.m_l, ql, uc);
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      (void)rescale<K>(out_l, th.lane, th.kk, uc[c]);
      if (out_l) sts_vec(th.sc + d.dst0 + c * th.dB, th.plane, uc[c]);
    }
  }
  if (K != 1 || out_r) {
    T uc[CPT][4];
    rows_times<T, CPT, true>(th, th.km + d.m_r, qr, uc);
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      (void)rescale<K>(out_r, th.lane, th.kk, uc[c]);
      if (out_r) sts_vec(th.sc + d.dst1 + c * th.dB, th.plane, uc[c]);
    }
  }
  if constexpr (OUTER) {
    // slot A's task of this round (lane 0 always holds a column of slot A): folded at the top of the next round
    const unsigned fa = __shfl_sync(0xffffffffu, d.flags, 0);
    ocp->pending = (fa & WF_LIVE) != 0;
    ocp->pend_l = __shfl_sync(0xffffffffu, d.id_l, 0);
    ocp->pend_r = __shfl_sync(0xffffffffu, d.id_r, 0);
    ocp->pend_par = par * ocp->par;
  }
}

// Fold the per-thread sums of the previous round (copy warps): a slot's ts pairs -> (a, b),
// added to the branch gradients / the log-likelihood.  The copy warps share the shared-memory
// and fp64 pipes with thirteen busy compute warps and get an instruction through every ~15
// cycles, so the fold is written for instruction count: a warp folds FOUR slots per pass (slots
// s0, s0 + step, ...: eight lanes each), every lane adds ts / 8 pairs — branch-free, every load
// independent — and three shuffle stages finish.  Fixed summation order: deterministic.
// ids: what each slot's sums belong to (shared memory: the issue record of the next round;
// after the last round: global memory).
template <typename T, bool IDS_SHARED>
__device__ __forceinline__ void fold_slots(const WalkParams<T>& p, unsigned gbuf, int s0, int step, int lane, bool first_tile,
                                           unsigned ids_smem, const int2* ids_gmem, double& ll_acc, long long* stamp = nullptr) {
  const bool holes = p.cpw != 32;  // K does not divide 32: lanes of a slot that hold no column never wrote their pair
  const int li = lane & 7, s2 = s0 + (lane >> 3) * step;
  int2 ids = make_int2(-1, -1);
  if (s2 < p.wn) ids = IDS_SHARED ? lds_i2(ids_smem + s2 * 8) : __ldg(ids_gmem + s2);
  const bool live = ids.x != -1;
  const unsigned row = gbuf + min(s2, p.wn - 1) * p.ts * 16;
  double a = 0.0, b = 0.0;
  // every lane always loads (a clamped, valid address); the value is zeroed when the pair does
  // not exist — guarded loads cost more in branches than in loads
  for (int c0 = li; c0 < p.ts; c0 += 32) {
    double2 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int c = c0 + 8 * u;
      v[u] = lds_d2(row + min(c, p.ts - 1) * 16);
      const bool ok = live & (c < p.ts) & !(holes & (((c & 31) >= p.cpw) | ((c >> 5) * p.cpw + (c & 31) >= p.ct)));
      v[u].x = ok ? v[u].x : 0.0;
      v[u].y = ok ? v[u].y : 0.0;
    }
    a += (v[0].x + v[1].x) + (v[2].x + v[3].x);
    b += (v[0].y + v[1].y) + (v[2].y + v[3].y);
  }
#ifdef B2P_TRACE_CLOCKS
  if (stamp) { stamp[0] = clock64(); }
#endif
  // lanes 0-3 of a slot's eight end up with the first sum, lanes 4-7 with the second
  const double oa = __shfl_xor_sync(0xffffffffu, a, 4), ob = __shfl_xor_sync(0xffffffffu, b, 4);
  double v = (li & 4) ? (b + ob) : (a + oa);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  if ((li & 3) == 0) {
    const int id = li ? ids.y : ids.x;
    if (id == -2) ll_acc += v;
    else if (id >= 0) {
      // a branch is written once per tile, always by this lane: plain store for the CTA's first
      // tile, then fire-and-forget adds in tile order (deterministic, no read latency)
      double* gp = p.gpart + (size_t)blockIdx.x * p.n_branches + id;
      if (first_tile) *gp = v; else atomicAdd(gp, v);
    }
  }
}

template <typename T, int K, int NT, int CPT, bool OUTER = false>
__global__ void __launch_bounds__(NT, NT <= 256 ? 2 : 1) walk_kernel(const __grid_constant__ WalkParams<T> p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ double red[NT / 32];
  __shared__ unsigned long long gbase[4];  // global base of every job kind for the current tile
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int kk = K > 0 ? K : p.n_cat;
  const int first_copy_warp = NT / 32 - p.n_copy_warps;
  const bool copier = warp >= first_copy_warp;
  const int s = t / p.ts;
  const int tt = t - s * p.ts;
  int col = CPT == 1 ? (tt >> 5) * p.cpw + (tt & 31) : tt;
  const bool in_slot = !copier && s < p.wn && (CPT != 1 || ((tt & 31) < p.cpw && col < p.ct));
  if (!in_slot) col = 0;
  const int S = p.stages, D = p.depth, R = p.rounds_run;
  const unsigned sbase = pinned(smem_u32(smem));
  const unsigned bar0 = sbase + p.off_bar;
  const unsigned gb = pinned(smem_u32(gbase));
  int* scount = reinterpret_cast<int*>(smem + p.off_scount);

  WalkThread<T, CPT> th;
  th.lane = lane; th.kk = kk; th.k = col % kk; th.ts = p.ts;
  th.sc = pinned(sbase + col * 16); th.dB = p.ts * 16; th.plane = p.plane;
  th.km = pinned(sbase + th.k * 4 * (unsigned)sizeof(T));
  th.rs = p.row_stride;
  th.cw = (T)p.cat_w[th.k]; th.cwr = (T)(p.cat_w[th.k] * p.rates[th.k]);
  const int sl = in_slot ? s : 0;
  const unsigned o_desc = pinned(p.rel_cr + sl * 32);
  const unsigned o_tip = pinned(p.rel_tips + sl * 2 * p.ct_pad + col);
  const unsigned o_gslot = pinned(sbase + p.off_gbuf + (sl * p.ts + tt) * 16);

  if (t == 0) {
    for (int j = 0; j < S; ++j) mbar_init(bar0 + 8 * j, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_async_smem();
  }
  // the root's parent-side partial: the frequencies for every column (same for every tile)
  for (int c = t; c < p.ct; c += NT) {
    T f[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) f[i] = (T)p.freqs[i];
    sts_vec(sbase + p.off_freq + c * 16, p.plane, f);
  }
  if (!copier && s < p.wn && !in_slot) {
    // warp-aligned slots: the lanes past the tile's last column never write their pair of sums
    sts_d2(sbase + p.off_gbuf + (s * p.ts + tt) * 16, 0.0, 0.0);
    sts_d2(sbase + p.off_gbuf + ((p.wn + s) * p.ts + tt) * 16, 0.0, 0.0);
  }
  // outer-product mode: which slots the warp's lanes belong to and who shares them (see OuterCtx)
  OuterCtx oc{};
  if constexpr (OUTER) {
    const int w0 = warp * 32, sa = w0 / p.ts, end_a = (sa + 1) * p.ts;
    const bool has_b = end_a < w0 + 32 && sa + 1 < p.wn;
    auto slot = [&](int sx) {
      OuterSlot o;
      const int wf = (sx * p.ts) / 32, wl = ((sx + 1) * p.ts - 1) / 32, wa = (sx * p.ts + 31) / 32;  // wa: first warp with A = sx
      o.n_warps = wl - wf + 1; o.n_fold = wl - wa + 1; o.r = warp - wa;
      // the first warp holds the slot's first columns as its A columns only if the slot starts on its lane 0
      o.row0 = sbase + p.off_outer + (2 * wf + ((sx * p.ts) % 32 ? 1 : 0)) * p.outer_row;
      o.rows1 = sbase + p.off_outer + 2 * (wf + 1) * p.outer_row;
      return o;
    };
    outer_addresses<K>(oc, sbase + p.off_ostage + warp * 4096, lane);
    oc.row = sbase + p.off_outer + 2 * warp * p.outer_row;
    oc.bl = has_b ? end_a - w0 : 32;
    oc.a = slot(min(sa, p.wn - 1));
    oc.par = 2 * first_copy_warp * p.outer_row;
    oc.pending = false;
    oc.part = p.opart + (size_t)blockIdx.x * p.n_branches * (K * 16);
  }
  unsigned phase_bits = 0;
  double ll_acc = 0.0;
  bool first_tile = true;
  // one bulk copy of a job list entry {goff, soff, bytes, kind}
  auto job_ptr = [&](const uint4 jb) -> unsigned char* {
    unsigned long long b;
    asm volatile("ld.shared.u64 %0, [%1];" : "=l"(b) : "r"(gb + jb.w * 8));
    return reinterpret_cast<unsigned char*>(b + jb.x);
  };

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, first_tile = false) {
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const int cg = tile * p.ct + col + c * p.ts;
      th.w[c] = (in_slot && cg < p.n_cols) ? __ldg(p.weights + cg / kk) : 0.0;
    }
    for (int i = t; i < p.ct; i += NT) scount[i] = 0;
    if (t == 0) {
      gbase[JK_MSG] = (unsigned long long)(p.msg + (unsigned long long)tile * p.tile_stride_part);
      gbase[JK_PRE] = (unsigned long long)(p.pre + (unsigned long long)tile * p.tile_stride_part);
      gbase[JK_TIPS] = (unsigned long long)(p.tips + (unsigned long long)tile * p.tile_stride_tips);
      gbase[JK_REC] = (unsigned long long)p.records;
    }
    __syncthreads();
    if (warp == NT / 32 - 1) {  // boot: loads of rounds 0 .. depth-1
      const uint4* jobs = reinterpret_cast<const uint4*>(p.records + p.boot_base);
      unsigned j0 = 0;
      for (int f = 0; f < D && f < R; ++f) {
        const unsigned bar = bar0 + 8 * (f % S);
        if (lane == 0) mbar_expect_tx(bar, p.boot_tx[f]);
        __syncwarp();
        for (unsigned j = lane; j < p.boot_n[f]; j += 32) {
          const uint4 jb = __ldg(jobs + j0 + j);
          bulk_load(sbase + jb.y, job_ptr(jb), jb.z, bar);
        }
        j0 += p.boot_n[f];
      }
    }

    int st = 0;                                // stage of the current round: q % S
    unsigned frame = sbase + p.off_frame;      // its frame
    for (int q = 0; q < R; ++q) {
      const int par = q & 1;
#ifdef B2P_TRACE_CLOCKS
      const bool tr = p.trace && blockIdx.x == 0 && (t == 0 || t == NT - 32) && tile == 0;
      long long* trow = tr ? p.trace + ((t ? 1 : 0) * (long long)R + q) * 8 : nullptr;
      if (tr) trow[0] = clock64();
#endif
      // outer-product mode: the previous round's sums of the warp's slot go out while the operands arrive
      if constexpr (OUTER)
        if (!copier) outer_fold<K>(oc, lane, first_tile, p.outer_row, p.dbg);
      mbar_wait(bar0 + 8 * st, (phase_bits >> st) & 1u);
      phase_bits ^= 1u << st;
#ifdef B2P_TRACE_CLOCKS
      if (tr) trow[1] = clock64();
#endif
      // issue record of this round: {n_loads, n_stores, tx_bytes, pass}, fold ids, jobs
      const unsigned ir = frame + p.rel_ir, jobs = ir + p.ir_jobs;
      if (copier) {
        // Copy warps: the loads of round q + depth, the stores of round q-1's results (job j belongs
        // to copy warp j % n: issuing a bulk copy blocks its warp for ~100 cycles, and the SM takes
        // one every ~70 cycles at best, so the issue is kept off the compute warps), then the fold
        // of round q-1's sums, one slot per warp at a time.  The planner never lets a load target a
        // block whose store is issued in the same round, nor read back a value before the round
        // after its store completed (rules S and L, plan_check.cpp).
        const uint4 hdr = lds_u4(ir);
        const unsigned st_next = st == 0 ? S - 1 : st - 1;  // (q + D) % S with S = D + 1
        const unsigned bar = bar0 + 8 * st_next;
        const unsigned n_loads = (q + D < R) ? hdr.x : 0u;
        const unsigned np = p.n_copy_warps, pw = warp - first_copy_warp;
        if (pw == 0 && lane == 0 && n_loads) mbar_expect_tx(bar, hdr.z);
        for (unsigned j = pw + np * lane; j < n_loads; j += 32 * np) {
          const uint4 jb = lds_u4(jobs + j * 16);
          bulk_load(sbase + jb.y, job_ptr(jb), jb.z, bar);
        }
#ifdef B2P_TRACE_CLOCKS
        if (tr) trow[5] = clock64();
#endif
#ifdef B2P_DBG
        if (!(p.dbg & 2))
#endif
        for (unsigned j = pw + np * lane; j < hdr.y; j += 32 * np) {
          const uint4 jb = lds_u4(jobs + (hdr.x + j) * 16);
          bulk_store(job_ptr(jb), sbase + jb.y, jb.z);
        }
        bulk_commit();  // one (possibly empty) group per lane per round
#ifdef B2P_TRACE_CLOCKS
        if (tr) trow[2] = clock64();
#endif
#ifdef B2P_DBG
        if (!(p.dbg & 1))
#endif
        // sums exist from the root's post-order round on (outer-product mode: only the root's term)
        if (OUTER ? q == p.rounds_post : q >= p.rounds_post) {
          for (int s0 = pw; s0 < p.wn; s0 += 4 * np)
#ifdef B2P_TRACE_CLOCKS
            fold_slots<T, true>(p, sbase + p.off_gbuf + (par ^ 1) * p.wn * p.ts * 16, s0, np, lane, first_tile, ir + 16, nullptr, ll_acc,
                                tr && s0 == (int)pw ? trow + 7 : nullptr);
#else
            fold_slots<T, true>(p, sbase + p.off_gbuf + (par ^ 1) * p.wn * p.ts * 16, s0, np, lane, first_tile, ir + 16, nullptr, ll_acc);
#endif
        }
#ifdef B2P_TRACE_CLOCKS
        if (tr) trow[6] = clock64();
#endif
        // this round's stores have read shared memory (their blocks may be rewritten next round);
        // those of the round before are complete in global memory
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 1;" ::: "memory");
      } else {
        // descriptor and tip codes: independent reads (the codes sit at fixed per-slot positions);
        // lanes outside every slot read slot 0's and ignore them
        const uint4 d0 = lds_u4(frame + o_desc), d1 = lds_u4(frame + o_desc + 16);
        unsigned code_l[CPT], code_r[CPT];
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
          code_l[c] = lds_u8(frame + o_tip + c * p.ts);
          code_r[c] = lds_u8(frame + o_tip + p.ct_pad + c * p.ts);
        }
        TaskDesc d = unpack_desc(d0, d1);
        if (!in_slot) d.flags = 0;
        bool live = (d.flags & WF_LIVE) != 0;
#ifdef B2P_DBG
        if (p.dbg & 4) live = false;
#endif
        // K == 1: lanes without a task skip it.  K != 1: the task exchanges values over the K
        // category lanes with full-warp shuffles, so a warp runs it as a whole (lanes without a
        // task compute on a zero descriptor and store nothing) or, with no task at all, skips it.
        const bool run = (K == 1 && !OUTER) ? live : __any_sync(0xffffffffu, live);
        if (run) {
          const unsigned gslot = o_gslot + par * p.wn * p.ts * 16;
          double ga = 0.0, gb2 = 0.0;  // post-order (root): log-likelihood term, 0; pre-order: the two gradient sums
          if (q < p.rounds_post) {
            const bool is_root = (d.flags & WF_ROOT) != 0;
            const bool root_round = K == 1 ? is_root : __any_sync(0xffffffffu, is_root);  // the root runs alone in its round
            post_task<T, K, CPT>(p, th, d, live, root_round, scount, col, code_l, code_r, ga);
            if (live && is_root) sts_d2(gslot, ga, 0.0);
          } else {
            pre_task<T, K, CPT, OUTER>(p, th, d, live, code_l, code_r, ga, gb2, &oc, (unsigned)par);
            if (!OUTER && live) sts_d2(gslot, ga, gb2);
          }
        }
#ifdef B2P_TRACE_CLOCKS
        if (tr) trow[2] = clock64();
#endif
      }
      // shared-memory results of this round (generic proxy) before the bulk copies (async
      // proxy) that read them or overwrite their blocks
      fence_async_smem();
#ifdef B2P_TRACE_CLOCKS
      if (tr) trow[3] = clock64();
#endif
      __syncthreads();
#ifdef B2P_TRACE_CLOCKS
      if (tr) trow[4] = clock64();
#endif
      if (++st == S) { st = 0; frame = sbase + p.off_frame; } else frame += p.frame_bytes;
    }
    if (copier) {
      // the last round's sums: their ids come from global memory (no further issue record)
      const int2* tail = reinterpret_cast<const int2*>(p.records + p.tail_base);
      if (!OUTER || R == p.rounds_post)
      for (int s0 = warp - first_copy_warp; s0 < p.wn; s0 += 4 * p.n_copy_warps)
        fold_slots<T, false>(p, sbase + p.off_gbuf + ((R - 1) & 1) * p.wn * p.ts * 16, s0, p.n_copy_warps, lane, first_tile, 0u, tail, ll_acc);
      // (the last round leaves nothing to store: the root's post-order task writes no message,
      //  and a pre-order partial is stored only when a later round loads it back)
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    } else if constexpr (OUTER) {
      outer_fold<K>(oc, lane, first_tile, p.outer_row, p.dbg);  // the tile's last round
    }
    __syncthreads();
  }

  double v = warp_sum(ll_acc);
  if (lane == 0) red[warp] = v;
  __syncthreads();
  if (t == 0) {
    double sum = 0.0;
    for (int w = 0; w < NT / 32; ++w) sum += red[w];
    p.lpart[blockIdx.x] = sum;
  }
}

}  // namespace b2p
