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
  int frame_bytes, off_frame, rel_cr, rel_mats, rel_tips, ir_jobs;  // rel_ir == 0; ir_jobs: first job of an issue record
  int off_bar, off_gbuf, off_scount, off_freq;
  unsigned boot_base, tail_base;  // tail_base: fold ids of the last round that runs
  unsigned boot_n[4], boot_tx[4];
  unsigned long long tile_stride_part;  // (N-1) * blk
  unsigned long long tile_stride_tips;  // rounds * 2 wn * ct_pad
  const unsigned char* records;
  const unsigned char* mats;  // matrix bundles of all rounds, back to back (written by K1)
  unsigned char* msg;         // [tile][N-1] blocks: m_v
  unsigned char* pre;         // [tile][N-1] blocks: u_c that had to leave shared memory
  const unsigned char* tips;  // [tile][round][wn][2][ct_pad] state codes per column
  const double* weights;      // [P]
  double* lpart;              // [grid]
  double* gpart;              // [grid][2N-2]
  long long* trace;           // debug build (-DB2P_TRACE_CLOCKS): [2][rounds][8] clock stamps of CTA 0
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

// Power-of-two rescale shared by the K category lanes of a pattern; returns 1 if applied.
// fp64: partials are non-negative, so the high words order like the values; a denormal
// maximum (high word 0) is left alone — it cannot arise from one step below 2^-200.
template <int K>
__device__ __forceinline__ int rescale(unsigned gmask, int lane, int kk, double (&y)[4]) {
  int h = max(max(__double2hiint(y[0]), __double2hiint(y[1])), max(__double2hiint(y[2]), __double2hiint(y[3])));
  h = group_max_int<K>(h, gmask, lane, kk);
  if (h < 0x33700000 && h > 0) {  // < 2^-200
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= Scale<double>::factor();
    return 1;
  }
  return 0;
}
template <int K>
__device__ __forceinline__ int rescale(unsigned gmask, int lane, int kk, float (&y)[4]) {
  float mx = fmaxf(fmaxf(y[0], y[1]), fmaxf(y[2], y[3]));
  mx = group_max<K>(mx, gmask, lane, kk);
  if (mx < Scale<float>::threshold() && mx > 0.f) {
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= Scale<float>::factor();
    return 1;
  }
  return 0;
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
  unsigned dB;     // 16 * ts: the thread's second column inside a block (CPT == 2)
  unsigned plane;  // fp64: bytes between the two halves of a column
  unsigned km;     // shared base + k * row_bytes: a row's byte offset is added to this
  unsigned rs;     // bytes between rows
  unsigned gmask;  // the K lanes of this thread's pattern
  int lane, kk, k, ts;
  T cw, cwr;       // K != 1: category weight, category weight x rate
  double w[CPT];   // pattern weights of the thread's columns (0 for padding columns)
};

// A child's message -> registers.  tip: row `code` of its table (code from the round's tip rows);
// otherwise the thread's column of the child's block.  Same instructions either way.
template <typename T, int CPT>
__device__ __forceinline__ void fetch_child(const WalkThread<T, CPT>& th, unsigned a, bool tip, const unsigned (&code)[CPT],
                                            T (&m)[CPT][4]) {
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    const unsigned row = min(code[c], kTipAny);
    const unsigned at = tip ? th.km + a + row * th.rs : th.sc + a + c * th.dB;
    lds_vec(at, tip ? 16u : th.plane, m[c]);
    if (tip && code[c] >= kTipAmbiguous) {  // rare: an ambiguity code other than "any": sum the compatible rows
#pragma unroll
      for (int i = 0; i < 4; ++i) m[c][i] = (T)0;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if ((code[c] >> j) & 1u) {
          T r[4];
          lds_vec(th.km + a + j * th.rs, 16u, r);
#pragma unroll
          for (int i = 0; i < 4; ++i) m[c][i] += r[i];
        }
    }
  }
}

// The 32-byte descriptor as the planner wrote it (b2p_walk_types.h: WalkDesc), unpacked.
struct TaskDesc {
  unsigned a_l, a_r, a_v, dst0, dst1, m_l, m_r, flags;
};
__device__ __forceinline__ TaskDesc unpack_desc(const uint4 d0, const uint4 d1) {
  TaskDesc d;
  d.a_l = d0.x; d.a_r = d0.y; d.a_v = d0.z; d.dst0 = d0.w;
  d.dst1 = d1.x;
  d.m_l = (d1.y & 0xffffu) * 16u; d.m_r = (d1.y >> 16) * 16u;
  d.flags = d1.z & 0xffffu;
  return d;
}

// ------------------------------------------------------------------ post-order task
//   x_v = m_l o m_r (rescaled when tiny) ;  m_v = P_v x_v ; root: site likelihoods -> log-likelihood term
template <typename T, int K, int CPT>
__device__ __forceinline__ void post_task(const WalkParams<T>& p, const WalkThread<T, CPT>& th, const TaskDesc& d,
                                          int* scount, int col, unsigned gslot, T (&ma)[CPT][4], T (&mb)[CPT][4]) {
  int bumped[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
#pragma unroll
    for (int i = 0; i < 4; ++i) ma[c][i] *= mb[c][i];
    bumped[c] = rescale<K>(th.gmask, th.lane, th.kk, ma[c]);
  }
  if (!(d.flags & WF_ROOT)) {
    T mv[CPT][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      T r[4];
      lds_vec(th.km + d.a_v + i * th.rs, 16u, r);
#pragma unroll
      for (int c = 0; c < CPT; ++c)
        mv[c][i] = r[0] * ma[c][0] + r[1] * ma[c][1] + r[2] * ma[c][2] + r[3] * ma[c][3];
    }
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      sts_vec(th.sc + d.dst0 + c * th.dB, th.plane, mv[c]);
      if (bumped[c] && th.k == 0) atomicAdd(scount + col + c * th.ts, 1);
    }
  } else {
    double term = 0.0;
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      T site = (T)p.freqs[0] * ma[c][0] + (T)p.freqs[1] * ma[c][1] + (T)p.freqs[2] * ma[c][2] + (T)p.freqs[3] * ma[c][3];
      if (K != 1) site = group_sum<K>(site * th.cw, th.gmask, th.lane, th.kk);
      if (th.k == 0 && th.w[c] != 0.0)
        term += th.w[c] * (log((double)site) - (double)(scount[col + c * th.ts] + bumped[c]) * Scale<T>::log_factor());
    }
    sts_d2(gslot, term, 0.0);
  }
}

// One child's side of a pre-order task: its gradient numerator r_k q.(Q m) and, for an internal
// child, its pre-order partial u_c = P_c^T q (rescaled when tiny) -> shared memory.
template <typename T, int K, int CPT>
__device__ __forceinline__ void pre_side(const WalkParams<T>& p, const WalkThread<T, CPT>& th, bool tip, unsigned mat, unsigned dst,
                                         const T (&q)[CPT][4], const T (&m)[CPT][4], T (&num)[CPT]) {
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    num[c] = (T)0;
    // d/dt (P x) = r Q (P x): Q comes from the constant bank
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const T dm = p.q[4 * i] * m[c][0] + p.q[4 * i + 1] * m[c][1] + p.q[4 * i + 2] * m[c][2] + p.q[4 * i + 3] * m[c][3];
      num[c] += q[c][i] * dm;
    }
  }
  if (!tip) {
    T u[CPT][4];
#pragma unroll
    for (int c = 0; c < CPT; ++c)
#pragma unroll
      for (int j = 0; j < 4; ++j) u[c][j] = (T)0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      T r[4];
      lds_vec(th.km + mat + i * th.rs, 16u, r);
#pragma unroll
      for (int c = 0; c < CPT; ++c)
#pragma unroll
        for (int j = 0; j < 4; ++j) u[c][j] += r[j] * q[c][i];
    }
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      (void)rescale<K>(th.gmask, th.lane, th.kk, u[c]);
      sts_vec(th.sc + dst + c * th.dB, th.plane, u[c]);
    }
  }
}

// ------------------------------------------------------------------- pre-order task
//   q_l = u_v o m_r, q_r = u_v o m_l ; L = q_l . m_l ; dL/dt_l = r_k q_l.(Q m_l) ; u_l = P_l^T q_l (same for r)
template <typename T, int K, int CPT>
__device__ __forceinline__ void pre_task(const WalkParams<T>& p, const WalkThread<T, CPT>& th, const TaskDesc& d,
                                         unsigned gslot, const T (&ma)[CPT][4], const T (&mb)[CPT][4]) {
  T ql[CPT][4], qr[CPT][4], den[CPT], numl[CPT], numr[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    T u[4];
    lds_vec(th.sc + d.a_v + c * th.dB, th.plane, u);
    den[c] = (T)0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      ql[c][i] = u[i] * mb[c][i];
      qr[c][i] = u[i] * ma[c][i];
      den[c] += ql[c][i] * ma[c][i];
    }
  }
  pre_side<T, K, CPT>(p, th, (d.flags & WF_TIP_L) != 0, d.m_l, d.dst0, ql, ma, numl);
  pre_side<T, K, CPT>(p, th, (d.flags & WF_TIP_R) != 0, d.m_r, d.dst1, qr, mb, numr);
  double gl = 0.0, gr = 0.0;
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    T dn = den[c];
    if (K != 1) {
      // the site likelihood is the sum over the pattern's K category lanes; the numerators
      // need no such exchange: every lane adds its own category's share and the fold sums
      // over all lanes of the slot anyway
      dn = group_sum<K>(dn * th.cw, th.gmask, th.lane, th.kk);
      numl[c] *= th.cwr;
      numr[c] *= th.cwr;
    }
    // K == 1: the single category's weight cancels, its rate multiplies the numerators.  A
    // pattern of likelihood zero contributes nothing instead of inf * 0.
    const double dd = (double)dn;
    const double inv = dd > 0.0 ? (K == 1 ? th.w[c] * p.rates[0] : th.w[c]) * fast_rcp(dd) : 0.0;
    gl += (double)numl[c] * inv;
    gr += (double)numr[c] * inv;
  }
  sts_d2(gslot, gl, gr);
}

// Fold one round's per-thread sums (done by the copy warps one round later): slot s2's ts
// pairs -> (a, b), added to the branch gradients / the log-likelihood in a fixed order.
// A warp folds four slots at a time, eight lanes each.  ids: what each slot's sums belong to
// (shared memory: the issue record of the next round; after the last round: global memory).
template <typename T, bool IDS_SHARED>
__device__ __forceinline__ void fold_round(const WalkParams<T>& p, unsigned sbase, int par, int pw, int lane, bool first_tile,
                                           unsigned ids_smem, const int2* ids_gmem, double& ll_acc) {
  const unsigned gbuf = sbase + p.off_gbuf + par * p.wn * p.ts * 16;
  const int li = lane & 7;
  for (int s0 = pw * 4; s0 < p.wn; s0 += p.n_copy_warps * 4) {
    const int s2 = s0 + (lane >> 3);
    int2 ids = make_int2(-1, -1);
    if (s2 < p.wn) ids = IDS_SHARED ? lds_i2(ids_smem + s2 * 8) : __ldg(ids_gmem + s2);
    double a = 0.0, b = 0.0;
    if (ids.x != -1) {
      for (int c = li; c < p.ts; c += 8) {
        // lanes of a slot that hold no column (K does not divide 32) never wrote their pair
        if ((c & 31) >= p.cpw || (c >> 5) * p.cpw + (c & 31) >= p.ct) continue;
        const double2 v = lds_d2(gbuf + (s2 * p.ts + c) * 16);
        a += v.x; b += v.y;
      }
    }
    const double oa = __shfl_xor_sync(0xffffffffu, a, 4), ob = __shfl_xor_sync(0xffffffffu, b, 4);
    double v = (li & 4) ? (b + ob) : (a + oa);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    if ((li & 3) == 0) {
      const int id = li ? ids.y : ids.x;
      if (id == -2) ll_acc += v;
      else if (id >= 0) {
        double* gp = p.gpart + (size_t)blockIdx.x * p.n_branches + id;
        *gp = first_tile ? v : *gp + v;
      }
    }
  }
}

template <typename T, int K, int NT, int CPT>
__global__ void __launch_bounds__(NT, NT <= 256 ? 2 : 1) walk_kernel(const __grid_constant__ WalkParams<T> p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ double red[NT / 32];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int kk = K > 0 ? K : p.n_cat;
  const int first_copy_warp = NT / 32 - p.n_copy_warps;
  const bool copier = warp >= first_copy_warp;
  const int s = t / p.ts;
  const int tt = t - s * p.ts;
  int col = CPT == 2 ? tt : (tt >> 5) * p.cpw + (tt & 31);
  const bool in_slot = !copier && s < p.wn && (CPT == 2 || ((tt & 31) < p.cpw && col < p.ct));
  if (!in_slot) col = 0;
  const int S = p.stages, D = p.depth, R = p.rounds_run;
  const unsigned sbase = smem_u32(smem);
  const unsigned bar0 = sbase + p.off_bar;
  int* scount = reinterpret_cast<int*>(smem + p.off_scount);

  WalkThread<T, CPT> th;
  th.lane = lane; th.kk = kk; th.k = col % kk; th.ts = p.ts;
  th.gmask = group_mask<K>(lane, kk);
  th.sc = sbase + col * 16; th.dB = p.ts * 16; th.plane = p.plane;
  th.km = sbase + th.k * 4 * (unsigned)sizeof(T);
  th.rs = p.row_stride;
  th.cw = (T)p.cat_w[th.k]; th.cwr = (T)(p.cat_w[th.k] * p.rates[th.k]);
  const int sl = in_slot ? s : 0;
  const unsigned o_desc = p.rel_cr + sl * 32;
  const unsigned o_tip = p.rel_tips + sl * 2 * p.ct_pad + col;
  const unsigned o_gslot = sbase + p.off_gbuf + (sl * p.ts + tt) * 16;

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
  unsigned phase_bits = 0;
  double ll_acc = 0.0;
  bool first_tile = true;

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, first_tile = false) {
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const int cg = tile * p.ct + col + c * p.ts;
      th.w[c] = (in_slot && cg < p.n_cols) ? __ldg(p.weights + cg / kk) : 0.0;
    }
    for (int i = t; i < p.ct; i += NT) scount[i] = 0;
    __syncthreads();
    unsigned char* base_msg = p.msg + (unsigned long long)tile * p.tile_stride_part;
    unsigned char* base_pre = p.pre + (unsigned long long)tile * p.tile_stride_part;
    const unsigned char* base_tips = p.tips + (unsigned long long)tile * p.tile_stride_tips;
    auto gptr = [&](unsigned kind, unsigned goff) -> const unsigned char* {
      const unsigned char* b = kind == JK_MSG ? base_msg : kind == JK_PRE ? base_pre : kind == JK_TIPS ? base_tips
                               : kind == JK_MATS ? p.mats : p.records;
      return b + goff;
    };
    if (warp == NT / 32 - 1) {  // boot: loads of rounds 0 .. depth-1
      const uint4* jobs = reinterpret_cast<const uint4*>(p.records + p.boot_base);
      unsigned j0 = 0;
      for (int f = 0; f < D && f < R; ++f) {
        const unsigned bar = bar0 + 8 * (f % S);
        if (lane == 0) mbar_expect_tx(bar, p.boot_tx[f]);
        __syncwarp();
        for (unsigned j = lane; j < p.boot_n[f]; j += 32) {
          const uint4 jb = __ldg(jobs + j0 + j);
          bulk_load(sbase + jb.y, gptr(jb.w, jb.x), jb.z, bar);
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
      mbar_wait(bar0 + 8 * st, (phase_bits >> st) & 1u);
      phase_bits ^= 1u << st;
#ifdef B2P_TRACE_CLOCKS
      if (tr) trow[1] = clock64();
#endif
      if (copier) {
        // Copy warps: the loads of round q + depth, then the stores of round q-1's results.
        // Issuing them from warps of their own lets the copies overlap the arithmetic; one warp
        // issues a bulk copy every ~100 cycles, so several share the lists (job j belongs to
        // copy warp j % n).  Order within the round is free: the planner never lets a load
        // target a block whose store is issued in the same round, nor read back a value before
        // the round after its store completed (rules S and L, tests/tools/plan_check.cpp).
        const uint4 hdr = lds_u4(frame);
        const unsigned st_next = st == 0 ? S - 1 : st - 1;  // (q + D) % S with S = D + 1
        const unsigned bar = bar0 + 8 * st_next;
        const unsigned n_loads = (q + D < R) ? hdr.x : 0u;
        const unsigned jobs = frame + p.ir_jobs;
        const unsigned np = p.n_copy_warps, pw = warp - first_copy_warp;
        if (pw == 0 && lane == 0 && n_loads) mbar_expect_tx(bar, hdr.z);
        // loads of round q + depth first: they are what the next round waits for; the stores only
        // have to have read shared memory by the end of this round
        for (unsigned j = pw + np * lane; j < n_loads; j += 32 * np) {
          const uint4 jb = lds_u4(jobs + j * 16);
          bulk_load(sbase + jb.y, gptr(jb.w, jb.x), jb.z, bar);
        }
#ifdef B2P_TRACE_CLOCKS
        if (tr) trow[5] = clock64();
#endif
        for (unsigned j = pw + np * lane; j < hdr.y; j += 32 * np) {
          const uint4 jb = lds_u4(jobs + (hdr.x + j) * 16);
          bulk_store(const_cast<unsigned char*>(gptr(jb.w, jb.x)), sbase + jb.y, jb.z);
        }
        bulk_commit();  // one (possibly empty) group per lane per round
#ifdef B2P_TRACE_CLOCKS
        if (tr) trow[2] = clock64();
#endif
        if (q > 0) fold_round<T, true>(p, sbase, par ^ 1, pw, lane, first_tile, frame + 16, nullptr, ll_acc);
#ifdef B2P_TRACE_CLOCKS
        if (tr) trow[6] = clock64();
#endif
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 1;" ::: "memory");
      } else if (in_slot) {
        // descriptor and tip codes: independent reads (the codes sit at fixed per-slot positions)
        const uint4 d0 = lds_u4(frame + o_desc), d1 = lds_u4(frame + o_desc + 16);
        unsigned code_l[CPT], code_r[CPT];
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
          code_l[c] = lds_u8(frame + o_tip + c * p.ts);
          code_r[c] = lds_u8(frame + o_tip + p.ct_pad + c * p.ts);
        }
        const TaskDesc d = unpack_desc(d0, d1);
        if (d.flags & WF_LIVE) {
          T ma[CPT][4], mb[CPT][4];
          fetch_child<T, CPT>(th, d.a_l, (d.flags & WF_TIP_L) != 0, code_l, ma);
          fetch_child<T, CPT>(th, d.a_r, (d.flags & WF_TIP_R) != 0, code_r, mb);
#ifdef B2P_TRACE_CLOCKS
          if (tr) trow[2] = clock64();
#endif
          const unsigned gslot = o_gslot + par * p.wn * p.ts * 16;
          if (q < p.rounds_post) post_task<T, K, CPT>(p, th, d, scount, col, gslot, ma, mb);
          else pre_task<T, K, CPT>(p, th, d, gslot, ma, mb);
        }
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
      fold_round<T, false>(p, sbase, (R - 1) & 1, warp - first_copy_warp, lane, first_tile, 0u,
                           reinterpret_cast<const int2*>(p.records + p.tail_base), ll_acc);
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
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
