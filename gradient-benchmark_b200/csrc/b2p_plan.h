// Host-side planner of the wave kernel (b2p_wave.cuh): tile / thread geometry, shared
// memory layout, and the static schedules of both passes with operand routing.
//
// Scheduling model.  Unit tasks (one per internal node) on WN identical slots per round.
//   post-order pass: in-tree precedence, a parent may run the round after its last child
//     (that child's message then comes from the ring; older ones from global memory).
//     Hu's highest-level-first rule (level = distance from the root) is makespan-optimal.
//   pre-order pass: out-tree precedence (a node after its parent); priority = height of the
//     node's subtree.  A task that runs exactly one round after its parent reads u_v from
//     the ring; the ring holds at most ring_cap blocks per round, so when the budget is
//     spent such tasks are deferred by a round and read u_v from global memory instead.
#pragma once
#include <algorithm>
#include <queue>
#include <string>
#include <vector>

#include "b2p_wave.cuh"

namespace b2p {

struct WaveGeometry {
  int nt = 0, ct = 0, ts = 0, cpw = 32, wn = 0, ring_cap = 0, pj = 0, n_tiles = 0, grid = 0;
  int off_task = 0, off_ahead = 0, off_mats = 0, off_stage = 0, off_ring = 0, off_part = 0, off_scount = 0;
  int blk = 0;  // bytes of one block (ct columns)
  size_t smem = 0;
};

inline int align16(int x) { return (x + 15) & ~15; }

// Choose the tile width and thread geometry.  elem = sizeof(T).  Returns false if even
// the smallest block size does not fit.
inline bool plan_geometry(int n_tips, int n_cols, int K, int elem, int n_sms, size_t smem_max, int nt_force,
                          int ct_force, int ctas_per_sm, WaveGeometry& g) {
  n_sms *= ctas_per_sm;
  const bool packed = (32 % K) == 0;
  const int cpw = packed ? 32 : (32 / K) * K;
  const int ct_max = 128, ct_min = (8 + K - 1) / K * K;
  int m = std::max(1, (int)((n_cols + (long long)n_sms * ct_max - 1) / ((long long)n_sms * ct_max)));
  int ct = (int)((n_cols + (long long)n_sms * m - 1) / ((long long)n_sms * m));
  ct = (ct + K - 1) / K * K;
  ct = std::max(ct, ct_min);
  ct = std::min(ct, std::max(K, (n_cols + K - 1) / K * K));
  if (ct_force > 0) ct = std::max(K, ct_force / K * K);
  const int ts = packed ? ct : ((ct + cpw - 1) / cpw) * 32;
  const int mat_bytes = elem == 8 ? 144 : 80;
  const int nts[4] = {1024, 768, 512, 256};
  for (int nt : nts) {
    if (nt_force > 0 && nt != nt_force) continue;
    if (ts > nt) continue;
    int wn = std::min(std::min(nt / ts, std::max(1, n_tips - 1)), 64);
    if (wn < 1) continue;
    const int pj = (ts + 30) / 32 + 1;
    int off = 0;
    const int blk = ct * 4 * elem;
    g.off_task = off; off += 2 * wn * 32;
    g.off_ahead = off; off += 2 * wn * 32;
    g.off_mats = off; off += align16(2 * wn * 3 * K * mat_bytes);
    g.off_stage = off; off += 3 * wn * blk;
    g.off_part = off; off += align16(2 * wn * (16 + 16 * pj));
    g.off_scount = off; off += align16(ct * 4);
    g.off_ring = off;
    const long long left = (long long)smem_max - off;
    int ring_cap = (int)std::min<long long>(left / (2LL * blk), 2LL * wn);
    if (ring_cap < wn) continue;
    g.blk = blk;
    g.nt = nt; g.ct = ct; g.ts = ts; g.cpw = cpw; g.wn = wn; g.ring_cap = ring_cap; g.pj = pj;
    g.n_tiles = (n_cols + ct - 1) / ct;
    g.grid = std::min(g.n_tiles, n_sms);
    g.smem = (size_t)off + 2ULL * ring_cap * blk;
    return true;
  }
  return false;
}

struct WaveSchedule {
  std::vector<WaveTask> post_task, pre_task;
  std::vector<WaveAhead> post_ahead, pre_ahead;
  int rounds_post = 0, rounds_pre = 0, depth = 0;
  long long ring_reads_post = 0, ring_reads_pre = 0, pre_global_blocks = 0;
};

inline WaveTask idle_task() {
  WaveTask t{};
  t.l = t.r = -1;
  t.src_l = t.src_r = t.src_v = t.dst0 = t.dst1 = kNoBlock;
  return t;
}
inline WaveAhead idle_ahead() {
  WaveAhead a{};
  a.g_l = a.g_r = a.g_v = kNoFetch;
  a.m[0] = a.m[1] = a.m[2] = -1;
  return a;
}

// triples: [N-1][3] (node, left, right), node ids N..2N-2 in post-order (validated here).
inline bool plan_schedule(int N, int P, int K, const int32_t* triples, const WaveGeometry& g, WaveSchedule& out,
                          std::string& why) {
  const int wn = g.wn, ring_cap = g.ring_cap;
  const unsigned n_cols = (unsigned)P * (unsigned)K;
  if ((unsigned long long)(N - 1) * n_cols > 0xFFFFFFF0ull || (unsigned long long)N * P > 0xFFFFFFF0ull) {
    why = "problem too large for 32-bit column offsets";
    return false;
  }
  auto ring_off = [&](int round, int idx) { return (unsigned short)((g.off_ring + ((round & 1) * ring_cap + idx) * g.blk) / 16); };
  auto stage_off = [&](int op, int slot) { return (unsigned short)((g.off_stage + (op * wn + slot) * g.blk) / 16); };
  auto gcol = [&](int node) { return (unsigned)(node - N) * n_cols; };
  const int n_nodes = 2 * N - 1, root = 2 * N - 2;
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
  out.depth = max_depth + 1;
  using Item = std::pair<int, int>;  // (priority, -node)

  // ---------------------------------------------------------------- post-order pass
  std::vector<int> round_post(n_nodes, -1), slot_post(n_nodes, -1), pending(n_nodes, 0);
  {
    std::priority_queue<Item> ready;
    for (int v = N; v < n_nodes; ++v) {
      pending[v] = (left[v] >= N) + (right[v] >= N);
      if (pending[v] == 0) ready.push({depth[v], -v});
    }
    int done = 0, r = 0;
    std::vector<int> chosen;
    while (done < N - 1) {
      chosen.clear();
      while ((int)chosen.size() < wn && !ready.empty()) { chosen.push_back(-ready.top().second); ready.pop(); }
      if (chosen.empty()) { why = "post schedule stalled"; return false; }
      for (int i = 0; i < (int)chosen.size(); ++i) { round_post[chosen[i]] = r; slot_post[chosen[i]] = i; }
      for (int v : chosen) {
        ++done;
        if (v != root && --pending[parent[v]] == 0) ready.push({depth[parent[v]], -parent[v]});
      }
      ++r;
    }
    out.rounds_post = r;
    out.post_task.assign((size_t)r * wn, idle_task());
    out.post_ahead.assign((size_t)r * wn, idle_ahead());
    for (int v = N; v < n_nodes; ++v) {
      const int rv = round_post[v], sv = slot_post[v];
      WaveTask& t = out.post_task[(size_t)rv * wn + sv];
      WaveAhead& a = out.post_ahead[(size_t)rv * wn + sv];
      t.l = left[v]; t.r = right[v];
      auto route = [&](int c, int op, unsigned short& src, unsigned& fetch, int& mat) {
        if (c < N) { src = kNoBlock; fetch = (unsigned)c * (unsigned)P; a.kinds |= 1u << op; mat = c; return; }
        if (round_post[c] == rv - 1) { src = ring_off(rv - 1, slot_post[c]); ++out.ring_reads_post; }
        else { src = stage_off(op, sv); fetch = gcol(c); }
      };
      route(t.l, 0, t.src_l, a.g_l, a.m[1]);
      route(t.r, 1, t.src_r, a.g_r, a.m[2]);
      if (v == root) t.flags = TF_ROOT;
      else {
        a.m[0] = v;
        t.gout0 = gcol(v);  // the pre-order pass always reads m_v from global memory
        if (round_post[parent[v]] == rv + 1) t.dst0 = ring_off(rv, sv);
      }
    }
  }

  // ----------------------------------------------------------------- pre-order pass
  std::vector<int> round_pre(n_nodes, -1), slot_pre(n_nodes, -1);
  {
    std::priority_queue<Item> ready;
    ready.push({height[root], -root});
    int done = 0, r = 0;
    std::vector<int> chosen;
    std::vector<Item> deferred;
    while (done < N - 1) {
      chosen.clear();
      deferred.clear();
      int ring_used = 0;
      while ((int)chosen.size() < wn && !ready.empty()) {
        const Item it = ready.top();
        ready.pop();
        const int v = -it.second;
        const bool needs_ring = v != root && round_pre[parent[v]] == r - 1;
        if (needs_ring && ring_used >= ring_cap) { deferred.push_back(it); continue; }
        ring_used += needs_ring;
        chosen.push_back(v);
      }
      for (const Item& it : deferred) ready.push(it);
      if (chosen.empty() && deferred.empty()) { why = "pre schedule stalled"; return false; }
      for (int i = 0; i < (int)chosen.size(); ++i) { round_pre[chosen[i]] = r; slot_pre[chosen[i]] = i; }
      for (int v : chosen) {
        ++done;
        if (left[v] >= N) ready.push({height[left[v]], -left[v]});
        if (right[v] >= N) ready.push({height[right[v]], -right[v]});
      }
      ++r;
    }
    out.rounds_pre = r;
    out.pre_task.assign((size_t)r * wn, idle_task());
    out.pre_ahead.assign((size_t)r * wn, idle_ahead());
    std::vector<int> ring_next(r, 0);  // ring blocks handed out per producer round
    for (int v = N; v < n_nodes; ++v) {
      const int rv = round_pre[v], sv = slot_pre[v];
      WaveTask& t = out.pre_task[(size_t)rv * wn + sv];
      WaveAhead& a = out.pre_ahead[(size_t)rv * wn + sv];
      t.l = left[v]; t.r = right[v];
      a.m[1] = t.l; a.m[2] = t.r;
      for (int op = 0; op < 2; ++op) {
        const int c = op ? t.r : t.l;
        unsigned short& src = op ? t.src_r : t.src_l;
        unsigned& fetch = op ? a.g_r : a.g_l;
        if (c < N) { src = kNoBlock; fetch = (unsigned)c * (unsigned)P; a.kinds |= 1u << op; }
        else { src = stage_off(op, sv); fetch = gcol(c); t.flags |= op ? TF_INT_R : TF_INT_L; }
      }
      if (v == root) t.flags |= TF_ROOT;
    }
    // route u_c from each producer (parent task) to its consumer (child task)
    for (int v = N; v < n_nodes; ++v) {
      const int rv = round_pre[v];
      WaveTask& t = out.pre_task[(size_t)rv * wn + slot_pre[v]];
      for (int side = 0; side < 2; ++side) {
        const int c = side ? t.r : t.l;
        if (c < N) continue;
        const int rc = round_pre[c], sc = slot_pre[c];
        WaveTask& ct = out.pre_task[(size_t)rc * wn + sc];
        WaveAhead& ca = out.pre_ahead[(size_t)rc * wn + sc];
        if (rc == rv + 1) {
          const int idx = ring_next[rv]++;
          if (idx >= ring_cap) { why = "internal error: ring budget exceeded"; return false; }
          ct.src_v = ring_off(rv, idx);
          (side ? t.dst1 : t.dst0) = ring_off(rv, idx);
          ++out.ring_reads_pre;
        } else {
          ct.src_v = stage_off(2, sc);
          ca.g_v = gcol(c);
          t.flags |= side ? TF_G1 : TF_G0;
          (side ? t.gout1 : t.gout0) = gcol(c);
          ++out.pre_global_blocks;
        }
      }
    }
  }
  return true;
}

}  // namespace b2p
