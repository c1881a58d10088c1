// Traversal kernel (sm_100a) — SURVEY.md §8(a) rows a3-a6: post-order partials, root
// log-likelihood, pre-order partials, branch gradients, in ONE launch.
//
// One persistent CTA owns a tile of ct adjacent columns (column = pattern x category) and
// walks the WHOLE tree for it, so every dependency is CTA-local: no grid sync, no flags.
// Per round the CTA runs up to wn independent node tasks side by side; thread t works on
// (slot t / ts, column t % ts) — slots are packed at thread granularity so an awkward tile
// width (68 columns at 1000 taxa x 10k patterns on 148 SMs) still fills the lanes.
//
// Data movement is entirely asynchronous and issued by ONE warp (the last one, which holds
// no compute threads) from lists the host
// planner wrote (b2p_plan.h): every operand block, tip row, matrix bundle and descriptor
// record of round q + depth is a cp.async.bulk (TMA, UBLKCP in SASS) global -> shared copy
// that completes on the mbarrier of that round; every result block is written to shared
// memory by the compute threads and leaves with a cp.async.bulk shared -> global copy.
// Compute threads execute no global-memory instruction at all.  Shared memory is a pool of
// blocks allocated statically by the planner (the schedule is static, so lifetimes are
// known): a value that is re-read soon simply stays where it was produced.
//
// Message formulation: a node stores the message to its parent, m_v = P_v x_v.
//   post-order task of v:  x_v = m_l o m_r ;  m_v = P_v x_v
//   pre-order task of v:   q_l = u_v o m_r,  q_r = u_v o m_l,  L = q_l . m_l,
//                          dL/dt_l = r_k q_l.(Q m_l),  u_l = P_l^T q_l   (same for r)
// A tip child's message is a column of P picked by its state mask (stored transposed, so
// the column is one 32-byte read); ambiguity masks sum the compatible columns.
// Gradient sums over a slot's columns go through shared memory and are folded one round
// later by one warp per slot (fixed order: deterministic).
#pragma once
#include "b2p_device.cuh"
#include "b2p_walk_types.h"

namespace b2p {

constexpr int kMaxCatW = 32;

template <typename T>
struct WalkParams {
  int n_tiles, rounds_run, n_branches, n_cat;
  int n_cols, n_patterns, n_copy_warps, pad_;
  int ct, ts, cpw, wn, stages, depth;
  int blk, plane, ct_pad, mat_stride, mats_bytes, cr_bytes, ir_bytes, tips_bytes;
  int off_bar, off_cr, off_ir, off_mats, off_tips, off_gbuf, off_gids, off_scount;
  unsigned ir_base, boot_base;
  unsigned boot_n[4], boot_tx[4];
  unsigned long long tile_stride_part;  // (N-1) * blk
  unsigned long long tile_stride_tips;  // 2N * ct_pad
  const unsigned char* records;
  const unsigned char* mats;  // [rounds][wn][3][K] matrices in schedule order (written by K1)
  unsigned char* msg;         // [tile][N-1] blocks: m_v
  unsigned char* pre;         // [tile][N-1] blocks: u_c that had to leave shared memory
  const unsigned char* tips;  // [tile][2N][ct_pad] state masks per column, tip rows in schedule order
  const double* weights;      // [P]
  double* lpart;              // [grid]
  double* gpart;              // [grid][2N-2]
  long long* trace;           // debug: [2][rounds][5] clock stamps of CTA 0 (warp 0 / last warp), or null
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

// ---- one column (4 states) in shared memory ---------------------------------------------
// fp64: two 16-byte halves in separate planes (consecutive threads -> consecutive 16 B =>
// conflict-free LDS.128 / STS.128); fp32: one 16 B unit.
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

// Message of a tip child from the transposed matrix mt (mt[4*j + i] = P[i][j]).
template <typename T>
__device__ __forceinline__ void tip_message(const T* mt, unsigned mask, T (&y)[4]) {
  if (__popc(mask) == 1) {
    load_row(mt + 4 * (__ffs(mask) - 1), y);
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] = (T)0;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if ((mask >> j) & 1u) {
        T r[4];
        load_row(mt + 4 * j, r);
#pragma unroll
        for (int i = 0; i < 4; ++i) y[i] += r[i];
      }
  }
}

// y = M x, rows of M read as 16-byte units
template <typename T>
__device__ __forceinline__ void matvec_rows(const T* m, const T (&x)[4], T (&y)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    T r[4];
    load_row(m + 4 * i, r);
    y[i] = r[0] * x[0] + r[1] * x[1] + r[2] * x[2] + r[3] * x[3];
  }
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
// fp64: partials are non-negative, so the high words order like the values; a denormal
// maximum (high word 0) is left alone — it cannot arise from one step below 2^-200.
template <int K>
__device__ __forceinline__ int rescale(int lane, int kk, double (&y)[4]) {
  int h = max(max(__double2hiint(y[0]), __double2hiint(y[1])), max(__double2hiint(y[2]), __double2hiint(y[3])));
  h = group_max_int<K>(h, lane, kk);
  if (h < 0x33700000 && h > 0) {  // < 2^-200
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= Scale<double>::factor();
    return 1;
  }
  return 0;
}
template <int K>
__device__ __forceinline__ int rescale(int lane, int kk, float (&y)[4]) {
  float mx = fmaxf(fmaxf(y[0], y[1]), fmaxf(y[2], y[3]));
  mx = group_max<K>(mx, lane, kk);
  if (mx < Scale<float>::threshold() && mx > 0.f) {
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] *= Scale<float>::factor();
    return 1;
  }
  return 0;
}

template <typename T, int K, int NT>
__global__ void __launch_bounds__(NT, 1) walk_kernel(const __grid_constant__ WalkParams<T> p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ double red[NT / 32];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int kk = K > 0 ? K : p.n_cat;
  const int s = t / p.ts;
  const int tt = t - s * p.ts;
  int col = (tt >> 5) * p.cpw + (tt & 31);
  const bool in_slot = s < p.wn && (tt & 31) < p.cpw && col < p.ct;
  if (!in_slot) col = 0;
  const int k = col % kk;
  const int c16 = col * 16;
  const int plane = p.plane;
  const int S = p.stages, D = p.depth, R = p.rounds_run;
  const unsigned sbase = smem_u32(smem);
  const unsigned bar0 = sbase + p.off_bar;
  int* scount = reinterpret_cast<int*>(smem + p.off_scount);
  const T cw = (T)p.cat_w[k];
  const T cwr = (T)(p.cat_w[k] * p.rates[k]);

  if (t == 0) {
    for (int j = 0; j < S; ++j) mbar_init(bar0 + 8 * j, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_async_smem();
  }
  unsigned phase_bits = 0;
  double ll_acc = 0.0;
  bool first_tile = true;

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, first_tile = false) {
    const int cg = tile * p.ct + col;
    const double w_p = (in_slot && cg < p.n_cols) ? __ldg(p.weights + cg / kk) : 0.0;
    for (int i = t; i < p.ct; i += NT) scount[i] = 0;
    __syncthreads();
    const unsigned char* base_msg = p.msg + (unsigned long long)tile * p.tile_stride_part;
    const unsigned char* base_pre = p.pre + (unsigned long long)tile * p.tile_stride_part;
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

    int st = 0;  // stage of the current round: q % S
    for (int q = 0; q < R; ++q) {
      const int par = q & 1;
      const bool tr = p.trace && blockIdx.x == 0 && (t == 0 || t == NT - 32) && tile == 0;
      long long* trow = tr ? p.trace + ((t ? 1 : 0) * (long long)R + q) * 5 : nullptr;
      if (tr) trow[0] = clock64();
      mbar_wait(bar0 + 8 * st, (phase_bits >> st) & 1u);
      if (tr) trow[1] = clock64();
      phase_bits ^= 1u << st;
      const unsigned char* ir = smem + p.off_ir + st * p.ir_bytes;
      const uint4 hdr = *reinterpret_cast<const uint4*>(ir);
      if (warp >= NT / 32 - p.n_copy_warps) {
        // Copy warps (the last warps hold no compute threads): stores of round q-1's results
        // first — the end of this round waits until they have read shared memory — then the
        // loads of round q + depth.  Issuing them from warps of their own lets the copies
        // overlap the arithmetic; one warp issues a bulk copy every ~100 cycles, so several
        // share the list (job j belongs to copy warp j % n_copy_warps).
        const int st_next = st == 0 ? S - 1 : st - 1;  // (q + D) % S with S = D + 1
        const unsigned bar = bar0 + 8 * st_next;
        const unsigned n_loads = (q + D < R) ? hdr.x : 0u;
        const uint4* jobs = reinterpret_cast<const uint4*>(ir + 16);
        const unsigned np = p.n_copy_warps, pw = warp - (NT / 32 - np);
        if (pw == 0 && lane == 0 && n_loads) mbar_expect_tx(bar, hdr.z);
        for (unsigned j = pw + np * lane; j < hdr.y; j += 32 * np) {
          const uint4 jb = jobs[hdr.x + j];
          bulk_store(const_cast<unsigned char*>(gptr(jb.w, jb.x)), sbase + jb.y, jb.z);
        }
        bulk_commit();  // one (possibly empty) group per lane per round
        for (unsigned j = pw + np * lane; j < n_loads; j += 32 * np) {
          const uint4 jb = jobs[j];
          bulk_load(sbase + jb.y, gptr(jb.w, jb.x), jb.z, bar);
        }
      }
      if (tr) trow[2] = clock64();
      // ---- this thread's task
      unsigned tip_l = 0, tip_r = 0;
      unsigned flags = 0, src_l = kNoBlk, src_r = kNoBlk, src_v = kNoBlk, dst0 = kNoBlk, dst1 = kNoBlk;
      const unsigned char* cr = smem + p.off_cr + st * p.cr_bytes;
      if (in_slot) {
        const uint4 dw = *reinterpret_cast<const uint4*>(cr + s * 16);
        tip_l = dw.w & 0xffffu; tip_r = dw.w >> 16;
        src_l = dw.x & 0xffffu; src_r = dw.x >> 16;
        src_v = dw.y & 0xffffu; dst0 = dw.y >> 16;
        dst1 = dw.z & 0xffffu; flags = dw.z >> 16;
        if (col == 0) {
          int2 ids = *reinterpret_cast<const int2*>(cr + p.wn * 16 + s * 8);
          if (!(flags & WF_LIVE)) ids = make_int2(-1, -1);
          reinterpret_cast<int2*>(smem + p.off_gids)[par * p.wn + s] = ids;
        }
      }
      const bool live = (flags & WF_LIVE) != 0;
      if (K != 1 || live) {  // K == 1: no warp-collective inside a task, idle lanes skip it
        const unsigned char* mslot = smem + p.off_mats + st * p.mats_bytes + (s * 3 * kk + k) * p.mat_stride;
        const T* m0 = reinterpret_cast<const T*>(mslot);
        const T* m1 = reinterpret_cast<const T*>(mslot + kk * p.mat_stride);
        const T* m2 = reinterpret_cast<const T*>(mslot + 2 * kk * p.mat_stride);
        const unsigned char* tiprow = smem + p.off_tips + st * p.tips_bytes + col;
        T ma[4], mb[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) ma[i] = mb[i] = (T)0;
        if (live) {
          if (flags & WF_TIP_L) tip_message(m1, tiprow[tip_l * p.ct_pad], ma);
          else s_ld(smem + src_l * 16 + c16, plane, ma);
          if (flags & WF_TIP_R) tip_message(m2, tiprow[tip_r * p.ct_pad], mb);
          else s_ld(smem + src_r * 16 + c16, plane, mb);
        }
        double2* gslot = reinterpret_cast<double2*>(smem + p.off_gbuf) + (par * p.wn + s) * p.ct + col;
        if (hdr.w == 0) {
          // ------------------------------------------------------------ post-order task
          T y[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) y[i] = ma[i] * mb[i];
          const int bumped = rescale<K>(lane, kk, y);
          const bool is_root = (flags & WF_ROOT) != 0;
          if (live && !is_root) {
            if (bumped && k == 0) atomicAdd(scount + col, 1);
            T mv[4];
            matvec_rows(m0, y, mv);
            s_st(smem + dst0 * 16 + c16, plane, mv);
          }
          if (K != 1 || is_root) {
            T site = (T)p.freqs[0] * y[0] + (T)p.freqs[1] * y[1] + (T)p.freqs[2] * y[2] + (T)p.freqs[3] * y[3];
            site = group_sum<K>(site * cw, lane, kk);
            if (live && is_root) {
              double term = 0.0;
              if (k == 0) term = w_p * (log((double)site) - (double)(scount[col] + bumped) * Scale<T>::log_factor());
              *gslot = make_double2(term, 0.0);
            }
          }
        } else {
          // ------------------------------------------------------------- pre-order task
          T u[4];
          if (live && src_v != kNoBlk) s_ld(smem + src_v * 16 + c16, plane, u);
          else {
#pragma unroll
            for (int i = 0; i < 4; ++i) u[i] = live ? (T)p.freqs[i] : (T)0;
          }
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
            den = group_sum<K>(den * cw, lane, kk);
            numl = group_sum<K>(numl * cwr, lane, kk);
            numr = group_sum<K>(numr * cwr, lane, kk);
          }
          if (live) {
            double gl = 0.0, gr = 0.0;
            if (k == 0) {
              // K == 1: the single category's weight cancels, its rate multiplies the numerators
              const double inv = (K == 1 ? w_p * p.rates[0] : w_p) / (double)den;
              gl = (double)numl * inv;
              gr = (double)numr * inv;
            }
            *gslot = make_double2(gl, gr);
          }
          // pre-order partials of the internal children: u_c = P_c^T q_c  (ma / mb are dead)
          const bool want_l = live && (flags & WF_OUT_L), want_r = live && (flags & WF_OUT_R);
          if (K != 1 || want_l) {
            if (want_l) matvec_t_rows(m1, ql, ma);
            (void)rescale<K>(lane, kk, ma);
            if (want_l) s_st(smem + dst0 * 16 + c16, plane, ma);
          }
          if (K != 1 || want_r) {
            if (want_r) matvec_t_rows(m2, qr, mb);
            (void)rescale<K>(lane, kk, mb);
            if (want_r) s_st(smem + dst1 * 16 + c16, plane, mb);
          }
        }
      }

      // ---- fold the previous round's gradient / likelihood sums (one warp per slot)
      if (q > 0) {
        const int2* gids = reinterpret_cast<const int2*>(smem + p.off_gids) + (par ^ 1) * p.wn;
        for (int s2 = warp; s2 < p.wn; s2 += NT / 32) {
          const int2 ids = gids[s2];
          if (ids.x == -1) continue;
          const double2* gb = reinterpret_cast<const double2*>(smem + p.off_gbuf) + ((par ^ 1) * p.wn + s2) * p.ct;
          double a = 0.0, b = 0.0;
          for (int c = lane; c < p.ct; c += 32) { const double2 v = gb[c]; a += v.x; b += v.y; }
          const double oa = __shfl_xor_sync(0xffffffffu, a, 16), ob = __shfl_xor_sync(0xffffffffu, b, 16);
          double v = (lane & 16) ? (b + ob) : (a + oa);
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          if ((lane & 15) == 0) {
            const int id = lane ? ids.y : ids.x;
            if (id == -2) ll_acc += v;
            else if (id >= 0) {
              double* gp = p.gpart + (size_t)blockIdx.x * p.n_branches + id;
              *gp = first_tile ? v : *gp + v;
            }
          }
        }
      }

      fence_async_smem();
      if (warp >= NT / 32 - p.n_copy_warps) {
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 1;" ::: "memory");
      }
      if (tr) trow[3] = clock64();
      __syncthreads();
      if (tr) trow[4] = clock64();
      st = st + 1 == S ? 0 : st + 1;
    }
    // fold the last round
    {
      const int par = (R - 1) & 1;
      const int2* gids = reinterpret_cast<const int2*>(smem + p.off_gids) + par * p.wn;
      for (int s2 = warp; s2 < p.wn; s2 += NT / 32) {
        const int2 ids = gids[s2];
        if (ids.x == -1) continue;
        const double2* gb = reinterpret_cast<const double2*>(smem + p.off_gbuf) + (par * p.wn + s2) * p.ct;
        double a = 0.0, b = 0.0;
        for (int c = lane; c < p.ct; c += 32) { const double2 v = gb[c]; a += v.x; b += v.y; }
        const double oa = __shfl_xor_sync(0xffffffffu, a, 16), ob = __shfl_xor_sync(0xffffffffu, b, 16);
        double v = (lane & 16) ? (b + ob) : (a + oa);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((lane & 15) == 0) {
          const int id = lane ? ids.y : ids.x;
          if (id == -2) ll_acc += v;
          else if (id >= 0) {
            double* gp = p.gpart + (size_t)blockIdx.x * p.n_branches + id;
            *gp = first_tile ? v : *gp + v;
          }
        }
      }
    }
    if (warp >= NT / 32 - p.n_copy_warps) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
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
