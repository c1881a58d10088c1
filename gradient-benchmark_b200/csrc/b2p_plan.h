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
  int off_ent = 0, off_mats = 0, off_stage = 0, off_ring = 0, off_part = 0, off_scount = 0;
  size_t smem = 0;
};

inline int align16(int x) { return (x + 15) & ~15; }

// Choose the tile width and thread geometry.  elem = sizeof(T).  Returns false if even
// the smallest block size does not fit.
inline bool plan_geometry(int n_tips, int n_cols, int K, int elem, int n_sms, size_t smem_max, int nt_force,
                          int ct_force, WaveGeometry& g) {
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
  const int planes = elem == 8 ? 2 : 1;
  const int mat_bytes = elem == 8 ? 144 : 80;
  const int nts[4] = {1024, 768, 512, 256};
  for (int nt : nts) {
    if (nt_force > 0 && nt != nt_force) continue;
    if (ts > nt) continue;
    int wn = std::min(std::min(nt / ts, std::max(1, n_tips - 1)), 64);
    if (wn < 1) continue;
    const int pj = (ts + 30) / 32 + 1;
    int off = 0;
    g.off_ent = off; off += align16(3 * wn * 32);
    g.off_mats = off; off += align16(2 * wn * 3 * K * mat_bytes);
    g.off_stage = off; off += 3 * nt * 16 * planes;
    g.off_part = off; off += align16(2 * wn * (16 + 16 * pj));
    g.off_scount = off; off += align16(ct * 4);
    g.off_ring = off;
    const int blk = ct * 4 * elem;
    const long long left = (long long)smem_max - off;
    int ring_cap = (int)std::min<long long>(left / (2LL * blk), 2LL * wn);
    if (ring_cap < wn) continue;
    g.nt = nt; g.ct = ct; g.ts = ts; g.cpw = cpw; g.wn = wn; g.ring_cap = ring_cap; g.pj = pj;
    g.n_tiles = (n_cols + ct - 1) / ct;
    g.grid = std::min(g.n_tiles, n_sms);
    g.smem = (size_t)off + 2ULL * ring_cap * blk;
    return true;
  }
  return false;
}

struct WaveSchedule {
  std::vector<WaveEntry> post, pre;
  int rounds_post = 0, rounds_pre = 0, depth = 0;
  long long ring_reads_post = 0, ring_reads_pre = 0, pre_global_blocks = 0;
};

inline WaveEntry idle_entry() {
  WaveEntry e{};
  e.v = e.l = e.r = -1;
  e.kl = e.kr = e.kv = KIND_TIP;
  return e;
}

// triples: [N-1][3] (node, left, right), node ids N..2N-2 in post-order (validated here).
inline bool plan_schedule(int N, const int32_t* triples, int wn, int ring_cap, WaveSchedule& out, std::string& why) {
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
    out.post.assign((size_t)r * wn, idle_entry());
    for (int v = N; v < n_nodes; ++v) {
      WaveEntry& e = out.post[(size_t)round_post[v] * wn + slot_post[v]];
      e.v = v; e.l = left[v]; e.r = right[v];
      auto route = [&](int c, unsigned char& kind, unsigned short& ring_idx) {
        if (c < N) { kind = KIND_TIP; return; }
        if (round_post[c] == round_post[v] - 1) { kind = KIND_RING; ring_idx = (unsigned short)slot_post[c]; ++out.ring_reads_post; }
        else kind = KIND_GLOBAL;
      };
      route(e.l, e.kl, e.ring_l);
      route(e.r, e.kr, e.ring_r);
      e.kv = KIND_ROOT;
      if (v != root) {
        e.flags = F_O0_GLOBAL;  // the pre-order pass always reads m_v from global memory
        if (round_post[parent[v]] == round_post[v] + 1) { e.flags |= F_O0_RING; e.ring_o0 = (unsigned short)slot_post[v]; }
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
    out.pre.assign((size_t)r * wn, idle_entry());
    std::vector<int> ring_next(r, 0);  // ring blocks handed out per producer round
    for (int v = N; v < n_nodes; ++v) {
      WaveEntry& e = out.pre[(size_t)round_pre[v] * wn + slot_pre[v]];
      e.v = v; e.l = left[v]; e.r = right[v];
      e.kl = e.l < N ? KIND_TIP : KIND_GLOBAL;
      e.kr = e.r < N ? KIND_TIP : KIND_GLOBAL;
      if (v == root) e.kv = KIND_ROOT;
    }
    // route u_c from each producer (parent task) to its consumer (child task)
    for (int v = N; v < n_nodes; ++v) {
      WaveEntry& e = out.pre[(size_t)round_pre[v] * wn + slot_pre[v]];
      for (int side = 0; side < 2; ++side) {
        const int c = side ? e.r : e.l;
        if (c < N) continue;
        WaveEntry& ce = out.pre[(size_t)round_pre[c] * wn + slot_pre[c]];
        if (round_pre[c] == round_pre[v] + 1) {
          const int idx = ring_next[round_pre[v]]++;
          if (idx >= ring_cap) { why = "internal error: ring budget exceeded"; return false; }
          ce.kv = KIND_RING; ce.ring_v = (unsigned short)idx;
          if (side) { e.flags |= F_O1_RING; e.ring_o1 = (unsigned short)idx; }
          else { e.flags |= F_O0_RING; e.ring_o0 = (unsigned short)idx; }
          ++out.ring_reads_pre;
        } else {
          ce.kv = KIND_GLOBAL;
          e.flags |= side ? F_O1_GLOBAL : F_O0_GLOBAL;
          ++out.pre_global_blocks;
        }
      }
    }
  }
  return true;
}

}  // namespace b2p
