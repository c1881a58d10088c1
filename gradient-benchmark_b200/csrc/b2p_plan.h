// Host-side planner of the traversal kernel (b2p_walk.cuh): tile / thread geometry, the
// shared-memory layout, the static schedule of both passes, the static allocation of the
// shared-memory block pool, and the per-round copy lists the kernel's copy warp executes.
// Plain C++ (also compiled by g++ for tests/tools/plan_check.cpp).
//
// Scheduling.  Unit tasks (one per internal node per pass) on wn slots per round.
//   post-order rounds: in-tree precedence; Hu's highest-level-first rule (level = distance
//     from the root) is makespan-optimal for unit tasks on identical slots.
//   pre-order rounds: out-tree precedence; priority = height of the node's subtree.
//   The two passes are concatenated into ONE round sequence, so the copy pipeline never
//   drains between them.
// Block pool.  Every block a round reads or writes lives in a pool block chosen here.
//   A produced value (message m_v, pre-order partial u_c) stays resident until its last
//   reader if the pool has room; under pressure the resident value whose next use is
//   farthest away is dropped (Belady) and that use is served by a bulk load issued `depth`
//   rounds ahead instead.  Messages are always stored to HBM (the pre-order pass needs all
//   of them); pre-order partials only when they are dropped.
// Timing rules the allocation obeys (checked by plan_check.cpp):
//   L  a load for round f is issued at the top of round f - depth (rounds < depth: before
//      round 0); its block must be unused from then until round f has read it.
//   S  a store of a round-p result is issued at the top of round p + 1 and has finished
//      reading shared memory before round p + 1 ends; the block may be rewritten from
//      the top of round p + 2.
//   R  a block read by tasks of round r may be rewritten from the top of round r + 1.
#pragma once
#include <algorithm>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <queue>
#include <string>
#include <vector>

#include "b2p_walk_types.h"

namespace b2p {

struct WalkGeometry {
  int nt = 0, ct = 0, ts = 0, cpw = 32, wn = 0, depth = 1, stages = 2, n_tiles = 0, grid = 0;
  int elem = 8, K = 1, n_copy_warps = 1;
  int cpt = 1;         // columns per thread (2: thread tt of a slot owns columns tt and tt + ts)
  int frame_bytes = 0; // one stage's frame: issue record | descriptors + ids | matrices | tip rows
  int off_frame = 0, rel_ir = 0, rel_cr = 0, rel_mats = 0, rel_tips = 0;
  int ct_pad = 0;      // bytes of one tip row of a tile (ct rounded up to 16)
  int blk = 0;         // bytes of one block
  int row_bytes = 0;   // one category's 4 values of a matrix row: 4 * elem
  int row_stride = 0;  // bytes between matrix rows: K * row_bytes (+ padding: bank shift between rows)
  int mats_rows = 0;   // rows one round's matrix bundle may take (the scheduler keeps every round within it)
  int mats_bytes = 0;  // mats_rows * row_stride
  int cr_bytes = 0;    // one round's descriptors
  int ir_ids = 0;      // bytes of the fold ids inside an issue record
  int ir_bytes = 0;    // one round's issue record (header + fold ids + max_jobs jobs)
  int tips_bytes = 0;  // one round's tip rows: wn * 2 rows
  int max_jobs = 0;
  int off_bar = 0, off_cr = 0, off_ir = 0, off_tips = 0, off_gbuf = 0,
      off_scount = 0, off_freq = 0, off_pool = 0;
  int n_pool = 0;
  size_t smem = 0;
  // Outer-product mode (b2p_branch_outer_products): one column per thread, at least 32 threads per
  // slot (a warp then holds columns of at most two slots); every compute warp owns 4 KB of staging
  // rows and, per round parity, two rows of K * 32 sums.
  bool outer = false;
  int off_outer = 0, outer_row = 0;  // bytes of one warp's row
  int off_ostage = 0;                // 4 KB per compute warp: the columns' operands of the tensor-core sums
};

inline int align16(int x) { return (x + 15) & ~15; }

// Rows of the round's matrix bundle a slot gets on average (B2P_MAT_ROWS overrides): a
// post-order task with two tip children takes 14, the mix of a random tree about 9.
inline int mat_rows_per_slot() {
  const char* e = getenv("B2P_MAT_ROWS");
  return e ? std::max(8, std::min(atoi(e), kRowsTaskMax)) : 11;
}

// Shared-memory layout for a given (nt, ct, wn); returns the number of pool blocks (< 0: does not fit).
inline int layout_walk(WalkGeometry& g, size_t smem_max) {
  const int S = g.stages, wn = g.wn, K = g.K;
  g.blk = g.ct * 4 * g.elem;
  g.ct_pad = align16(g.ct);
  g.row_bytes = 4 * g.elem;
  // rows of different states / matrix rows land in different banks: one category fills 32 (16) bytes,
  // so K = 1 needs no padding; otherwise shift each row by one category's width
  g.row_stride = K * g.row_bytes + (K > 1 ? g.row_bytes : 0);
  g.mats_rows = std::max(kRowsTaskMax, wn * mat_rows_per_slot());
  g.mats_bytes = g.mats_rows * g.row_stride;
  g.cr_bytes = wn * 32;
  g.max_jobs = 7 * wn + 3;
  g.ir_ids = align16(wn * 8);
  g.ir_bytes = 16 + g.ir_ids + g.max_jobs * 16;
  g.tips_bytes = wn * 2 * g.ct_pad;
  int off = 0;
  g.off_bar = off; off += 64;  // up to 8 mbarriers
  // A stage's frame: the round's record — descriptors | issue record (header, fold ids, as many
  // jobs as the round has) | the matrix rows right behind the last job — and its tip rows.  The
  // record is ONE bulk copy: it is laid out the same way in global memory (K1 writes the
  // matrices into it), and the descriptors carry absolute offsets, so nothing needs the
  // matrices at a fixed place.
  g.rel_cr = 0;
  g.rel_ir = g.cr_bytes;
  g.rel_mats = g.rel_ir + 16 + g.ir_ids;  // + 16 * (jobs of the round)
  g.rel_tips = align16(g.rel_ir + g.ir_bytes + g.mats_bytes);
  g.frame_bytes = g.rel_tips + g.tips_bytes;
  g.off_frame = off; off += S * g.frame_bytes;
  g.off_ir = g.off_frame + g.rel_ir; g.off_cr = g.off_frame + g.rel_cr;
  g.off_tips = g.off_frame + g.rel_tips;
  g.off_gbuf = off; off += 2 * wn * g.ts * 16;  // one (sum_a, sum_b) pair per compute thread
  g.off_scount = off; off += align16(g.ct * 4);
  off = (off + 127) & ~127;
  if (g.outer) {
    g.outer_row = K * 32 * 8;
    g.off_outer = off; off += 4 * (g.nt / 32 - g.n_copy_warps) * g.outer_row;  // two rows per warp, two round parities
    g.off_ostage = off; off += (g.nt / 32 - g.n_copy_warps) * 4096;
  }
  g.off_freq = off; off += g.blk;  // the root's parent-side partial: the frequencies, for every column
  off = (off + 127) & ~127;
  g.off_pool = off;
  const long long left = (long long)smem_max - off;
  if (left < g.blk) return -1;
  g.n_pool = (int)(left / g.blk);
  g.smem = (size_t)g.off_pool + (size_t)g.n_pool * g.blk;
  return g.n_pool;
}

// Choose the tile width and thread geometry.  elem = sizeof(T).  cpt = columns per thread
// (2 halves the per-column cost of descriptors, matrix reads and address arithmetic; it
// needs the 128 registers a 512-thread CTA leaves each thread).
inline bool plan_geometry_cpt(int n_tips, int n_cols, int K, int elem, int n_sms, size_t smem_max, int nt_force,
                              int ct_force, int depth, int n_copy_warps, int cpt, WalkGeometry& g, bool outer = false) {
  const bool packed = (32 % K) == 0;
  if (cpt == 2 && !packed) return false;
  if (outer && (cpt != 1 || !packed)) return false;
  const int cpw = packed ? 32 : (32 / K) * K;
  const int q = K * cpt;  // tile widths are multiples of this
  // widest tile: 144 columns (fp64) — two waves of 136-column tiles beat three of 96 at 40 000
  // columns (profiles/README.md, round 2); B2P_CT_MAX overrides
  const int ct_max = getenv("B2P_CT_MAX") ? std::max(q, atoi(getenv("B2P_CT_MAX"))) : 144, ct_min = (8 + q - 1) / q * q;
  int m = std::max(1, (int)((n_cols + (long long)n_sms * ct_max - 1) / ((long long)n_sms * ct_max)));
  int ct = (int)((n_cols + (long long)n_sms * m - 1) / ((long long)n_sms * m));
  ct = (ct + q - 1) / q * q;
  ct = std::max(ct, ct_min);
  ct = std::min(ct, std::max(q, (n_cols + q - 1) / q * q));
  if (ct_force > 0) ct = std::max(q, ct_force / q * q);
  const int ts = packed && !(outer && ct < 32) ? ct / cpt : ((ct + cpw - 1) / cpw) * 32;  // threads per slot
  const int nts1[4] = {1024, 768, 512, 256}, nts2[4] = {512, 768, 640, 256};
  for (int nt : (cpt == 2 ? nts2 : nts1)) {
    if (nt == 0) continue;
    if (nt_force > 0 && nt != nt_force) continue;
    if (ts > nt - 32 * n_copy_warps) continue;  // the last warps only issue copies
    int wn = std::min(std::min((nt - 32 * n_copy_warps) / ts, std::max(1, n_tips - 1)), outer ? 14 : 64);
    for (; wn >= 1; --wn) {
      g = WalkGeometry();
      g.nt = nt; g.ct = ct; g.ts = ts; g.cpw = cpw; g.wn = wn; g.elem = elem; g.K = K; g.cpt = cpt;
      g.depth = std::max(1, std::min(depth, 3)); g.stages = g.depth + 1; g.n_copy_warps = n_copy_warps;
      g.outer = outer;
      const int n_pool = layout_walk(g, smem_max);
      // enough blocks for one full round of operands in flight per stage plus the round's results
      if (n_pool >= std::max(3 * g.stages + 4, 4 * wn)) break;
    }
    if (wn < 1) continue;
    g.n_tiles = (n_cols + ct - 1) / ct;
    g.grid = std::min(g.n_tiles, n_sms);
    return true;
  }
  return false;
}
inline bool plan_geometry(int n_tips, int n_cols, int K, int elem, int n_sms, size_t smem_max, int nt_force,
                          int ct_force, int depth, int n_copy_warps, int cpt_want, WalkGeometry& g, bool outer = false) {
  if (outer) return plan_geometry_cpt(n_tips, n_cols, K, elem, n_sms, smem_max, nt_force, ct_force, depth, n_copy_warps, 1, g, true);
  if (cpt_want == 2 &&
      plan_geometry_cpt(n_tips, n_cols, K, elem, n_sms, smem_max, nt_force, ct_force, depth, n_copy_warps, 2, g))
    return true;
  return plan_geometry_cpt(n_tips, n_cols, K, elem, n_sms, smem_max, nt_force, ct_force, depth, n_copy_warps, 1, g);
}

struct WalkPlan {
  int rounds_post = 0, rounds_pre = 0, rounds = 0, tree_depth = 0, wn_eff = 0;
  std::vector<uint8_t> records;  // [rounds] round records, the tail fold ids, the boot jobs
  std::vector<uint32_t> rec_off;   // [rounds + 1] byte offset of every round's record
  std::vector<uint32_t> rec_mats;  // [rounds] offset of the matrix rows inside the round's record
  uint32_t boot_base = 0;
  uint32_t boot_n[4] = {0, 0, 0, 0}, boot_tx[4] = {0, 0, 0, 0};  // loads of rounds 0..depth-1
  std::vector<int32_t> tip_order;    // [rounds][wn][2] tip id of every tip row (-1: that child is not a tip)
  std::vector<int32_t> mat_dst;  // [2N-2][2]: byte offset (in `records`) of the branch's rows for the post / pre rounds
                                 // (-1: none); tip branches get the 5-row table, internal ones the 4 matrix rows
  uint32_t tail_base = 0;          // fold ids of the last round (read from global memory after the loop):
                                   // of a full run, then (+ ir_ids bytes) of a post-order-only run
  long long resident_reads = 0, loads_msg = 0, loads_pre = 0, stores_msg = 0, stores_pre = 0, evictions = 0;
  int max_jobs_used = 0;
};

namespace plan_detail {
struct Use { int round, node, op; };  // op: 0 left, 1 right, 2 parent side
struct Value {
  int prod = -1;             // producing round
  std::vector<Use> uses;     // by round
  int next = 0;              // first use not yet served
  int block = -1;            // resident pool block
  bool store = false;        // needs a copy in HBM
  int dst_block = -1;        // block it is produced into
};
struct Fetch { int vid, use; };
struct LoadRec { uint32_t kind, goff, soff, bytes; };
}  // namespace plan_detail

// triples: [N-1][3] (node, left, right), node ids N..2N-2 in post-order (validated here).
inline bool plan_walk(int N, const int32_t* triples, const WalkGeometry& g, int wn_eff, int budget, WalkPlan& out,
                      std::string& why) {
  using namespace plan_detail;
  const int wn = g.wn, D = g.depth, S = g.stages, NB = g.n_pool;
  wn_eff = std::max(1, std::min(wn_eff, wn));
  out = WalkPlan();
  out.wn_eff = wn_eff;
  const int n_nodes = 2 * N - 1, root = 2 * N - 2;
  if (N > (1 << 23)) { why = "more than 2^23 tips"; return false; }
  if ((unsigned long long)(N - 1) * (unsigned long long)g.blk > 0xFFFFFFF0ull ||
      (16ull * N + 128ull) * 2ull * (unsigned long long)g.wn * (unsigned long long)g.ct_pad > 0xFFFFFFF0ull) {
    why = "tile too large for 32-bit copy offsets";
    return false;
  }
  std::vector<int> left(n_nodes, -1), right(n_nodes, -1), parent(n_nodes, -1);
  std::vector<char> seen(n_nodes, 0);
  for (int e = 0; e < N - 1; ++e) {
    const int v = triples[3 * e], l = triples[3 * e + 1], r = triples[3 * e + 2];
    if (v != N + e) { why = "triples must list internal nodes N..2N-2 in order (post-order ids)"; return false; }
    if (l < 0 || r < 0 || l >= v || r >= v || l == r) { why = "child id out of range / not in post-order"; return false; }
    if (seen[l] || seen[r]) { why = "a node has two parents"; return false; }
    seen[l] = seen[r] = 1;
    left[v] = l; right[v] = r; parent[l] = v; parent[r] = v;
  }
  for (int v = 0; v < root; ++v)
    if (!seen[v]) { why = "tree is not connected (node without parent)"; return false; }
  // Canonical child order: a task with exactly one tip child has it on the left, so a round
  // holds three kinds of task (no tip / one tip / two tips) instead of four.  Everything below
  // (descriptors, branch ids, matrix slots) follows left[] / right[], so this is only a naming.
  const bool canon = !(getenv("B2P_PLAN_SORT") && atoi(getenv("B2P_PLAN_SORT")) == 0);
  if (canon)
    for (int v = N; v < n_nodes; ++v)
      if (right[v] < N && left[v] >= N) std::swap(left[v], right[v]);
  // task kind: number of tip children (tasks of a round are laid out by kind, so that the
  // threads of a warp — which straddles two slots — mostly take the same branches)
  auto kind_of = [&](int v) { return (left[v] < N) + (right[v] < N); };
  auto by_kind = [&](std::vector<int>& chosen) {
    if (canon) std::stable_sort(chosen.begin(), chosen.end(), [&](int a, int b) { return kind_of(a) < kind_of(b); });
  };
  std::vector<int> depth(n_nodes, 0), height(n_nodes, 0);
  int max_depth = 0;
  for (int v = root - 1; v >= 0; --v) {
    depth[v] = depth[parent[v]] + 1;
    if (v >= N) max_depth = std::max(max_depth, depth[v]);
  }
  for (int v = N; v < n_nodes; ++v) {
    int h = 0;
    if (left[v] >= N) h = std::max(h, height[left[v]] + 1);
    if (right[v] >= N) h = std::max(h, height[right[v]] + 1);
    height[v] = h;
  }
  out.tree_depth = max_depth + 1;

  // ------------------------------------------------------------------ schedule
  // Memory-aware list scheduling.  cnt[j] bounds the pool blocks that cannot be reclaimed
  // at the top of round j: operands of rounds j..j+D (loaded or resident, they must sit in
  // the pool by then) plus the results of rounds j-1 and j (being produced / being stored).
  // A task joins round f only if every window it touches stays within the budget, so the
  // allocation below always finds a free or droppable block.
  using Item = std::pair<int, int>;  // (priority, -node)
  std::vector<int> rnd_post(n_nodes, -1), slot_post(n_nodes, -1), rnd_pre(n_nodes, -1), slot_pre(n_nodes, -1);
  std::vector<std::vector<int>> tasks;  // per global round: node ids by slot
  std::vector<int> cnt;
  auto cnt_at = [&](int j) -> int& { if ((int)cnt.size() <= j) cnt.resize(j + 1, 0); return cnt[j]; };
  // try to add a task to round f: operands produced in rounds px[0..n_op), n_out results
  std::vector<int> rows_used;  // matrix rows of every round's bundle (kept within g.mats_rows)
  auto try_add = [&](int f, const int* px, int n_op, int n_out, int n_rows) -> bool {
    if ((int)rows_used.size() <= f) rows_used.resize(f + 1, 0);
    if (rows_used[f] + n_rows > g.mats_rows) return false;
    int inc[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // windows f-D .. f+1  (D <= 3)
    for (int o = 0; o < n_op; ++o)
      for (int j = std::max(f - D, 0); j <= f; ++j)
        if (j >= px[o] + 2) ++inc[j - (f - D)];
    inc[D] += n_out;      // window f
    inc[D + 1] += n_out;  // window f + 1
    for (int w = 0; w <= D + 1; ++w) {
      const int j = f - D + w;
      if (j < 0 || inc[w] == 0) continue;
      if (cnt_at(j) + inc[w] > budget) return false;
    }
    for (int w = 0; w <= D + 1; ++w) {
      const int j = f - D + w;
      if (j >= 0) cnt_at(j) += inc[w];
    }
    rows_used[f] += n_rows;
    return true;
  };
  auto child_rows = [&](int c) { return c < N ? kRowsTip : kRowsPlain; };
  {
    std::vector<int> pending(n_nodes, 0);
    std::priority_queue<Item> ready;
    for (int v = N; v < n_nodes; ++v) {
      pending[v] = (left[v] >= N) + (right[v] >= N);
      if (pending[v] == 0) ready.push({depth[v], -v});
    }
    int done = 0;
    std::vector<Item> skipped;
    while (done < N - 1) {
      std::vector<int> chosen;
      const int q = (int)tasks.size();
      skipped.clear();
      // the root runs alone in the last post-order round (every rescale count is final then)
      while ((int)chosen.size() < wn_eff && !ready.empty()) {
        const Item it = ready.top();
        const int v = -it.second;
        if (v == root && !chosen.empty()) break;
        ready.pop();
        int px[2], n_op = 0;
        if (left[v] >= N) px[n_op++] = rnd_post[left[v]];
        if (right[v] >= N) px[n_op++] = rnd_post[right[v]];
        const int n_rows = (v == root ? 0 : kRowsPlain) + (left[v] < N ? kRowsTip : 0) + (right[v] < N ? kRowsTip : 0);
        if (!try_add(q, px, n_op, v == root ? 0 : 1, n_rows)) { skipped.push_back(it); continue; }
        chosen.push_back(v);
        if (v == root) break;
      }
      for (const Item& it : skipped) ready.push(it);
      if (chosen.empty() && ready.empty()) { why = "post schedule stalled"; return false; }
      if ((int)tasks.size() > 8 * N + 64) { why = "pool exhausted (schedule)"; return false; }
      by_kind(chosen);
      for (int i = 0; i < (int)chosen.size(); ++i) { rnd_post[chosen[i]] = q; slot_post[chosen[i]] = i; }
      tasks.push_back(chosen);
      // successors become ready for the NEXT round
      for (int v : chosen) {
        ++done;
        if (v != root && --pending[parent[v]] == 0) ready.push({depth[parent[v]], -parent[v]});
      }
    }
    out.rounds_post = (int)tasks.size();
  }
  {
    // Two ready sets: `fresh` = children of the previous round's tasks (their parent-side
    // partial still sits in the pool: no store, no load), `old` = everything else.  Fresh
    // tasks go first, by subtree height; the walk becomes wn depth-first fronts in parallel.
    // B2P_PRE_LOCALITY=0 (planner experiments) restores pure height priority.
    const bool locality = !(getenv("B2P_PRE_LOCALITY") && atoi(getenv("B2P_PRE_LOCALITY")) == 0);
    std::priority_queue<Item> fresh, old;
    fresh.push({height[root], -root});
    int done = 0;
    std::vector<Item> skipped;
    while (done < N - 1) {
      std::vector<int> chosen;
      const int q = (int)tasks.size();
      skipped.clear();
      for (int pass = 0; pass < 2; ++pass) {
        std::priority_queue<Item>& ready = pass == 0 ? fresh : old;
        while ((int)chosen.size() < wn_eff && !ready.empty()) {
          const Item it = ready.top();
          const int v = -it.second;
          ready.pop();
          int px[3], n_op = 0;
          if (left[v] >= N) px[n_op++] = rnd_post[left[v]];
          if (right[v] >= N) px[n_op++] = rnd_post[right[v]];
          if (v != root) px[n_op++] = rnd_pre[parent[v]];
          if (!try_add(q, px, n_op, (left[v] >= N) + (right[v] >= N), child_rows(left[v]) + child_rows(right[v]))) {
            skipped.push_back(it);
            continue;
          }
          chosen.push_back(v);
        }
      }
      while (!fresh.empty()) { old.push(fresh.top()); fresh.pop(); }  // not taken: no longer fresh
      for (const Item& it : skipped) old.push(it);
      if (chosen.empty() && old.empty()) { why = "pre schedule stalled"; return false; }
      if ((int)tasks.size() > 16 * N + 128) { why = "pool exhausted (schedule)"; return false; }
      by_kind(chosen);
      for (int i = 0; i < (int)chosen.size(); ++i) { rnd_pre[chosen[i]] = q; slot_pre[chosen[i]] = i; }
      tasks.push_back(chosen);
      for (int v : chosen) {
        ++done;
        std::priority_queue<Item>& dst = locality ? fresh : old;
        if (left[v] >= N) dst.push({height[left[v]], -left[v]});
        if (right[v] >= N) dst.push({height[right[v]], -right[v]});
      }
    }
    out.rounds = (int)tasks.size();
    out.rounds_pre = out.rounds - out.rounds_post;
  }
  const int R = out.rounds, Rpost = out.rounds_post;

  // ------------------------------------------------------------ values and uses
  // vid: message of internal non-root v -> v - N ; pre-order partial of c -> (N-1) + (c - N)
  const int n_int = N - 1;
  std::vector<Value> val(2 * (size_t)n_int);
  auto vid_m = [&](int v) { return v - N; };
  auto vid_u = [&](int c) { return n_int + (c - N); };
  for (int v = N; v < root; ++v) {
    Value& m = val[vid_m(v)];
    const int p = parent[v], side = (right[p] == v) ? 1 : 0;
    m.prod = rnd_post[v];
    m.store = true;
    m.uses.push_back({rnd_post[p], p, side});
    m.uses.push_back({rnd_pre[p], p, side});
    Value& u = val[vid_u(v)];
    u.prod = rnd_pre[p];
    u.uses.push_back({rnd_pre[v], v, 2});
  }
  // resolved operand / result blocks per task
  std::vector<int> src_post(2 * (size_t)n_nodes, -1), src_pre(3 * (size_t)n_nodes, -1), dst_pre(2 * (size_t)n_nodes, -1),
      dst_post(n_nodes, -1);
  auto set_src = [&](const Use& u, int blk_idx) {
    if (u.round < Rpost) src_post[2 * (size_t)u.node + u.op] = blk_idx;
    else src_pre[3 * (size_t)u.node + u.op] = blk_idx;
  };

  // ------------------------------------------------------------- pool simulation
  const int INF = INT_MAX;
  std::vector<int> free_at(NB, 0);
  std::vector<int> resident;  // vids with a resident block
  std::vector<std::vector<Fetch>> fetch_req(R);
  std::vector<std::vector<LoadRec>> loads(R), stores(R + 1);
  auto drop = [&](int vid) {
    for (size_t i = 0; i < resident.size(); ++i)
      if (resident[i] == vid) { resident[i] = resident.back(); resident.pop_back(); return; }
  };
  auto alloc = [&](int i) -> int {
    for (int b = 0; b < NB; ++b)
      if (free_at[b] <= i) return b;
    int best = -1, best_next = -1;
    for (int vid : resident) {
      const Value& v = val[vid];
      if (v.prod + 2 > i) continue;
      const int nx = v.uses[v.next].round;
      if (nx < i + D + 1) continue;
      if (nx > best_next) { best_next = nx; best = vid; }
    }
    if (best < 0) return -1;
    Value& v = val[best];
    const int b = v.block;
    for (int k = v.next; k < (int)v.uses.size(); ++k) fetch_req[v.uses[k].round].push_back({best, k});
    v.block = -1;
    v.store = true;
    drop(best);
    ++out.evictions;
    return b;
  };
  for (int i = 0; i < R; ++i) {
    // loads issued at the top of round i
    const int f_lo = (i == 0) ? 0 : i + D, f_hi = std::min(R - 1, i + D);
    for (int f = f_lo; f <= f_hi; ++f) {
      for (size_t k = 0; k < fetch_req[f].size(); ++k) {
        const Fetch fr = fetch_req[f][k];
        const int b = alloc(i);
        if (b < 0) { why = "pool exhausted (loads)"; return false; }
        free_at[b] = f + 1;
        const Value& v = val[fr.vid];
        set_src(v.uses[fr.use], b);
        const bool is_m = fr.vid < n_int;
        const int node = is_m ? N + fr.vid : N + (fr.vid - n_int);
        loads[f].push_back({is_m ? JK_MSG : JK_PRE, (uint32_t)(node - N) * (uint32_t)g.blk,
                            (uint32_t)(g.off_pool + b * g.blk), (uint32_t)g.blk});
        (is_m ? out.loads_msg : out.loads_pre) += 1;
      }
    }
    // results of round i
    for (int v : tasks[i]) {
      if (i < Rpost) {
        if (v == root) continue;
        const int b = alloc(i);
        if (b < 0) { why = "pool exhausted (results)"; return false; }
        Value& m = val[vid_m(v)];
        m.block = m.dst_block = b;
        free_at[b] = INF;
        dst_post[v] = b;
        resident.push_back(vid_m(v));
        for (const Use& u : m.uses) set_src(u, b);
      } else {
        for (int side = 0; side < 2; ++side) {
          const int c = side ? right[v] : left[v];
          if (c < N) continue;
          const int b = alloc(i);
          if (b < 0) { why = "pool exhausted (results)"; return false; }
          Value& u = val[vid_u(c)];
          u.block = u.dst_block = b;
          free_at[b] = INF;
          dst_pre[2 * (size_t)v + side] = b;
          resident.push_back(vid_u(c));
          for (const Use& us : u.uses) set_src(us, b);
        }
      }
    }
    // reads of round i
    for (int v : tasks[i]) {
      auto served = [&](int vid) {
        Value& x = val[vid];
        if (x.block < 0) return;  // this use comes from a load
        if (x.uses[x.next].round != i) return;
        ++out.resident_reads;
        ++x.next;
        if (x.next == (int)x.uses.size()) {
          free_at[x.block] = x.store ? std::max(x.prod + 2, i + 1) : i + 1;  // no store: nothing reads it later
          x.block = -1;
          drop(vid);
        }
      };
      if (left[v] >= N) served(vid_m(left[v]));
      if (right[v] >= N) served(vid_m(right[v]));
      if (i >= Rpost && v != root) served(vid_u(v));
    }
    // uses of round i that were turned into loads: advance the cursor of dropped values
    for (const Fetch& fr : fetch_req[i]) {
      Value& x = val[fr.vid];
      if (x.block < 0 && x.next <= fr.use) x.next = fr.use + 1;
    }
  }
  // stores: every message; dropped pre-order partials
  for (int vid = 0; vid < (int)val.size(); ++vid) {
    const Value& x = val[vid];
    if (x.prod < 0 || !x.store) continue;
    const bool is_m = vid < n_int;
    const int node = is_m ? N + vid : N + (vid - n_int);
    stores[x.prod + 1].push_back({is_m ? JK_MSG : JK_PRE, (uint32_t)(node - N) * (uint32_t)g.blk,
                                  (uint32_t)(g.off_pool + x.dst_block * g.blk), (uint32_t)g.blk});
    (is_m ? out.stores_msg : out.stores_pre) += 1;
  }

  // -------------------------------------------------------------------- records
  // per-round loads besides the operand blocks: the tip code rows (fixed positions — row
  // 2 * slot + side of the round — so a thread reads its codes without waiting for its
  // descriptor; rows of internal children are never read for their value)
  out.tip_order.assign((size_t)R * 2 * wn, -1);
  for (int f = 0; f < R; ++f) {
    bool any = false;
    for (int s = 0; s < (int)tasks[f].size(); ++s) {
      const int v = tasks[f][s];
      for (int side = 0; side < 2; ++side) {
        const int c = side ? right[v] : left[v];
        if (c >= N) continue;
        out.tip_order[((size_t)f * wn + s) * 2 + side] = c;
        any = true;
      }
    }
    if (any)
      loads[f].push_back({JK_TIPS, (uint32_t)f * 2u * (uint32_t)wn * (uint32_t)g.ct_pad,
                          (uint32_t)(g.off_tips + (f % S) * g.frame_bytes), (uint32_t)(2 * (int)tasks[f].size() * g.ct_pad)});
  }
  // Round records: descriptors | issue header | fold ids of round q - 1 | jobs (loads of round
  // q + D, among them the record of round q + D itself; stores of round q - 1) | matrix rows of
  // the round's tasks, slot by slot —
  //   post-order task: [own matrix, 4 rows (not the root)] [left tip table, 5] [right tip table, 5]
  //   pre-order task:  [left child: tip table 5 / matrix 4] [right child: likewise]
  // mat_dst tells K1 where each branch's rows go.
  std::vector<int> n_jobs(R, 0);
  for (int q = 0; q < R; ++q) {
    const int n_l = q + D < R ? (int)loads[q + D].size() + 1 : 0;
    const int n_s = (int)stores[q].size();
    if (n_l + n_s > g.max_jobs) { why = "too many copies in one round"; return false; }
    out.max_jobs_used = std::max(out.max_jobs_used, n_l + n_s);
    n_jobs[q] = n_l + n_s;
  }
  out.mat_dst.assign(2 * (size_t)(2 * N - 2), -1);
  out.rec_off.assign(R + 1, 0);
  out.rec_mats.assign(R, 0);
  std::vector<uint32_t> mrow_own(n_nodes, 0), mrow_post(2 * (size_t)n_nodes, 0), mrow_pre(2 * (size_t)n_nodes, 0);  // row within the round
  for (int f = 0; f < R; ++f) {
    out.rec_mats[f] = (uint32_t)(g.rel_mats + 16 * n_jobs[f]);
    const uint32_t m0 = out.rec_off[f] + out.rec_mats[f];
    uint32_t rows = 0;
    for (int v : tasks[f]) {
      const int ch[2] = {left[v], right[v]};
      if (f < Rpost) {
        if (v != root) { mrow_own[v] = rows; out.mat_dst[2 * (size_t)v] = (int32_t)(m0 + rows * g.row_stride); rows += kRowsPlain; }
        for (int side = 0; side < 2; ++side)
          if (ch[side] < N) {
            mrow_post[2 * (size_t)v + side] = rows;
            out.mat_dst[2 * (size_t)ch[side]] = (int32_t)(m0 + rows * g.row_stride);
            rows += kRowsTip;
          }
      } else {
        for (int side = 0; side < 2; ++side) {
          mrow_pre[2 * (size_t)v + side] = rows;
          out.mat_dst[2 * (size_t)ch[side] + 1] = (int32_t)(m0 + rows * g.row_stride);
          rows += ch[side] < N ? kRowsTip : kRowsPlain;
        }
      }
    }
    if ((int)rows > g.mats_rows) { why = "internal error: matrix rows over budget"; return false; }
    out.rec_off[f + 1] = out.rec_off[f] + out.rec_mats[f] + rows * (uint32_t)g.row_stride;
  }
  out.tail_base = out.rec_off[R];
  out.boot_base = out.tail_base + 2u * (uint32_t)g.ir_ids;
  size_t n_boot = 0;
  for (int f = 0; f < std::min(D, R); ++f) n_boot += loads[f].size() + 1;
  out.records.assign((size_t)out.boot_base + n_boot * 16, 0);
  // what the two sums of every slot belong to: the children's branches of a pre-order task,
  // the log-likelihood term of the root's post-order task
  auto fold_ids = [&](int q, WalkIds* ids) {
    for (int s = 0; s < wn; ++s) ids[s] = WalkIds{-1, -1};
    if (q < 0) return;
    for (int s = 0; s < (int)tasks[q].size(); ++s) {
      const int v = tasks[q][s];
      if (q >= Rpost) ids[s] = WalkIds{left[v], right[v]};
      else if (v == root) ids[s] = WalkIds{-2, -1};
    }
  };
  auto rec_job = [&](int f) {
    return WalkJob{out.rec_off[f], (uint32_t)(g.off_frame + (f % S) * g.frame_bytes), out.rec_off[f + 1] - out.rec_off[f], JK_REC};
  };
  auto job_of = [&](const LoadRec& l) { return WalkJob{l.goff, l.soff, l.bytes, l.kind}; };
  for (int q = 0; q < R; ++q) {
    uint8_t* p = out.records.data() + out.rec_off[q] + g.rel_ir;
    WalkIssueHeader h{0, 0, 0, q >= Rpost ? 1u : 0u};
    fold_ids(q - 1, reinterpret_cast<WalkIds*>(p + 16));
    WalkJob* jobs = reinterpret_cast<WalkJob*>(p + 16 + g.ir_ids);
    int n = 0;
    if (q + D < R) {
      jobs[n] = rec_job(q + D); h.tx_bytes += jobs[n].bytes; ++n;  // first: the other loads are small next to it
      for (const LoadRec& l : loads[q + D]) { jobs[n++] = job_of(l); h.tx_bytes += l.bytes; }
      h.n_loads = (uint32_t)n;
    }
    for (const LoadRec& l : stores[q]) jobs[n++] = job_of(l);
    h.n_stores = (uint32_t)n - h.n_loads;
    std::memcpy(p, &h, 16);
  }
  // after the last round: of a full run, and of a run that stops after the post-order rounds
  fold_ids(R - 1, reinterpret_cast<WalkIds*>(out.records.data() + out.tail_base));
  fold_ids(Rpost - 1, reinterpret_cast<WalkIds*>(out.records.data() + out.tail_base + g.ir_ids));
  {
    WalkJob* jobs = reinterpret_cast<WalkJob*>(out.records.data() + out.boot_base);
    int n = 0;
    for (int f = 0; f < std::min(D, R); ++f) {
      const int n0 = n;
      jobs[n] = rec_job(f); out.boot_tx[f] += jobs[n].bytes; ++n;
      for (const LoadRec& l : loads[f]) { jobs[n++] = job_of(l); out.boot_tx[f] += l.bytes; }
      out.boot_n[f] = (uint32_t)(n - n0);
    }
  }
  // descriptors: byte offsets into shared memory for the stage the round runs in
  auto blk_off = [&](int b) { return (uint32_t)(g.off_pool + b * g.blk); };
  for (int q = 0; q < R; ++q) {
    WalkDesc* d = reinterpret_cast<WalkDesc*>(out.records.data() + out.rec_off[q] + g.rel_cr);
    const uint32_t mats0 = (uint32_t)(g.off_frame + (q % S) * g.frame_bytes) + out.rec_mats[q];
    auto mat_at = [&](uint32_t row) { return mats0 + row * (uint32_t)g.row_stride; };
    for (int s = 0; s < wn; ++s) std::memset(&d[s], 0, sizeof(WalkDesc));
    for (int s = 0; s < (int)tasks[q].size(); ++s) {
      const int v = tasks[q][s], l = left[v], r = right[v];
      WalkDesc& e = d[s];
      const bool post = q < Rpost;
      const std::vector<int>& src = post ? src_post : src_pre;
      const size_t so = (post ? 2 : 3) * (size_t)v;
      e.flags = WF_LIVE;
      if (l < N) e.flags |= WF_TIP_L;
      if (r < N) e.flags |= WF_TIP_R;
      if (v == root) e.flags |= WF_ROOT;
      if ((l >= N && src[so] < 0) || (r >= N && src[so + 1] < 0) || (!post && v != root && src[so + 2] < 0)) {
        why = "internal error: unresolved operand";
        return false;
      }
      const std::vector<uint32_t>& mrow = post ? mrow_post : mrow_pre;
      e.a_l = l < N ? mat_at(mrow[2 * (size_t)v]) : blk_off(src[so]);
      e.a_r = r < N ? mat_at(mrow[2 * (size_t)v + 1]) : blk_off(src[so + 1]);
      if (post) {
        if (v != root) { e.a_v = mat_at(mrow_own[v]); e.dst0 = blk_off(dst_post[v]); }
      } else {
        e.a_v = v == root ? (uint32_t)g.off_freq : blk_off(src[so + 2]);
        e.ids_lo = (uint32_t)l | ((uint32_t)r << 24);
        e.ids_hi = (uint16_t)((uint32_t)r >> 8);
        if (l >= N) { e.dst0 = blk_off(dst_pre[2 * (size_t)v]); e.m_l = (uint16_t)(mat_at(mrow_pre[2 * (size_t)v]) / 16); }
        if (r >= N) { e.dst1 = blk_off(dst_pre[2 * (size_t)v + 1]); e.m_r = (uint16_t)(mat_at(mrow_pre[2 * (size_t)v + 1]) / 16); }
      }
    }
  }
  return true;
}

// Plan against the whole pool; if the allocation still fails (the scheduler's count is an
// upper bound, so it should not), retry with a smaller budget and then fewer slots.
inline bool plan_walk_auto(int N, const int32_t* triples, const WalkGeometry& g, WalkPlan& out, std::string& why) {
  auto retry = [&]() {
    return why.find("pool exhausted") != std::string::npos || why.find("too many copies") != std::string::npos;
  };
  for (int budget = g.n_pool; budget >= 3 * g.stages + 4; budget -= std::max(1, g.n_pool / 16)) {
    if (plan_walk(N, triples, g, g.wn, budget, out, why)) return true;
    if (!retry()) return false;
  }
  for (int w = g.wn / 2; w >= 1; w /= 2) {
    if (plan_walk(N, triples, g, w, g.n_pool, out, why)) return true;
    if (!retry()) return false;
  }
  return false;
}

}  // namespace b2p
