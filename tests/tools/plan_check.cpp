// CPU-side checker of the traversal kernel's static plan (test tooling, no GPU).
// Replays the records b2p_plan.h emits — descriptors, copy lists — against a symbolic
// shared-memory pool (each block holds the NAME of a value, not numbers) and verifies
// every timing rule the kernel relies on:
//   * an operand read finds exactly the value the tree says it should (message of the
//     child / pre-order partial of the node), and never a block whose load is in flight;
//   * a result is never written into a block that is read or written by another task of
//     the same round, that has a load in flight, or that a store is still reading;
//   * a load reads HBM only after the store of that value has completed (>= 2 rounds);
//   * every message is stored exactly once; transaction byte counts match the job lists.
// usage: plan_check N shape seed ct K elem smem_bytes depth [nt [cpt]]   -> prints "OK ..." or "FAIL ..."
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <map>
#include <random>
#include <string>
#include <vector>

#include "../../gradient-benchmark_b200/csrc/b2p_plan.h"

using namespace b2p;

static std::vector<int32_t> random_triples(int N, int shape, unsigned seed) {
  // shape 0: random joins (coalescent-like), 1: caterpillar, 2: balanced
  std::mt19937 rng(seed);
  std::vector<int32_t> tr;
  std::vector<int> roots(N);
  for (int i = 0; i < N; ++i) roots[i] = i;
  int next = N;
  while (roots.size() > 1) {
    size_t a, b;
    if (shape == 1) { a = roots.size() - 1; b = 0; }
    else if (shape == 2) { a = 0; b = 1; }
    else { a = rng() % roots.size(); do { b = rng() % roots.size(); } while (b == a); }
    const int l = roots[a], r = roots[b];
    if (a < b) std::swap(a, b);
    roots.erase(roots.begin() + a);
    roots.erase(roots.begin() + b);
    tr.push_back(next); tr.push_back(l); tr.push_back(r);
    if (shape == 2) roots.push_back(next); else roots.insert(roots.begin(), next);
    ++next;
  }
  return tr;
}

#define FAIL(...) do { printf("FAIL "); printf(__VA_ARGS__); printf("\n"); return 1; } while (0)

int main(int argc, char** argv) {
  if (argc < 9) { printf("usage\n"); return 2; }
  const int N = atoi(argv[1]), shape = atoi(argv[2]), seed = atoi(argv[3]), ct = atoi(argv[4]), K = atoi(argv[5]),
            elem = atoi(argv[6]), smem = atoi(argv[7]), depth = atoi(argv[8]), nt = argc > 9 ? atoi(argv[9]) : 0, cpt = argc > 10 ? atoi(argv[10]) : 1;
  std::vector<int32_t> tr = random_triples(N, shape, seed);
  WalkGeometry g;
  const bool outer = argc > 11 && atoi(argv[11]) != 0;  // outer-product mode: warp-aligned slots + a row per compute warp
  if (!plan_geometry(N, ct * 148, K, elem, 148, smem, nt, ct, depth, outer ? 3 : 2, cpt, g, outer)) FAIL("geometry");
  if (outer) {
    if (g.ts < 32 || g.ts % K || g.cpt != 1 || !g.outer || g.ts < g.ct || (g.ts != g.ct && g.ts != 32)) FAIL("outer mode: threads per slot");
    const int rows_end = g.off_outer + 4 * (g.nt / 32 - g.n_copy_warps) * g.outer_row;
    if (g.outer_row != K * 32 * 8 || g.off_outer < g.off_scount + g.ct * 4 || rows_end > g.off_ostage ||
        g.off_ostage % 16 || g.off_ostage + (g.nt / 32 - g.n_copy_warps) * 4096 > g.off_freq)
      FAIL("outer mode: rows overlap their neighbours");
    if (g.wn * g.ts > g.nt - 32 * g.n_copy_warps) FAIL("outer mode: slots reach into the copy warps");
    if (g.wn > 14) FAIL("outer mode: more slots than named barriers");
  }
  WalkPlan plan;
  std::string why;
  if (!plan_walk_auto(N, tr.data(), g, plan, why)) FAIL("plan: %s", why.c_str());
  const int R = plan.rounds, Rpost = plan.rounds_post, D = g.depth, S = g.stages, wn = g.wn, root = 2 * N - 2;
  std::vector<int> left(2 * N - 1, -1), right(2 * N - 1, -1), parent(2 * N - 1, -1);
  for (int e = 0; e < N - 1; ++e) { const int v = tr[3 * e]; left[v] = tr[3 * e + 1]; right[v] = tr[3 * e + 2]; parent[left[v]] = parent[right[v]] = v; }
  // the planner's canonical child order: a lone tip child is the left one
  if (!(getenv("B2P_PLAN_SORT") && atoi(getenv("B2P_PLAN_SORT")) == 0))
    for (int v = N; v <= root; ++v) if (right[v] < N && left[v] >= N) std::swap(left[v], right[v]);
  // which node is the post-order task at a given place of the matrix bundles?  (own matrix rows)
  std::map<uint32_t, int> post_node_at;  // byte offset in the bundle array -> node
  for (int v = N; v < root; ++v) {
    if (plan.mat_dst[2 * v] < 0 || plan.mat_dst[2 * v] % 16 != 0) FAIL("mat_dst own of %d", v);
    post_node_at[(uint32_t)plan.mat_dst[2 * v]] = v;
  }
  for (int b = 0; b < 2 * N - 2; ++b) {
    if (plan.mat_dst[2 * b + 1] < 0) FAIL("branch %d has no pre-order matrix rows", b);
    if (b < N && plan.mat_dst[2 * b] < 0) FAIL("tip %d has no post-order matrix rows", b);
  }
  if ((int)plan.rec_off.size() != plan.rounds + 1 || (int)plan.rec_mats.size() != plan.rounds || (int)plan.tip_order.size() != plan.rounds * 2 * g.wn) FAIL("table sizes");
  {  // rows of different branches never overlap, every round stays within the frame's matrix region
    std::vector<std::pair<uint32_t, uint32_t>> spans;
    for (int b = 0; b < 2 * N - 2; ++b)
      for (int w = 0; w < 2; ++w)
        if (plan.mat_dst[2 * b + w] >= 0) spans.push_back({(uint32_t)plan.mat_dst[2 * b + w], (uint32_t)((b < N ? kRowsTip : kRowsPlain) * g.row_stride)});
    std::sort(spans.begin(), spans.end());
    for (size_t i = 1; i < spans.size(); ++i) if (spans[i - 1].first + spans[i - 1].second > spans[i].first) FAIL("matrix rows overlap");
    if (!spans.empty() && spans.back().first + spans.back().second > plan.rec_off.back()) FAIL("matrix rows past the end");
    for (int q = 0; q < plan.rounds; ++q) {
      const int len = (int)(plan.rec_off[q + 1] - plan.rec_off[q]);
      if (len > g.rel_tips || len % 16 || plan.rec_off[q] % 16) FAIL("round %d: record larger than the frame / misaligned", q);
      if ((int)(len - plan.rec_mats[q]) > g.mats_bytes) FAIL("round %d: more matrix rows than the budget", q);
    }
  }
  const uint8_t* rec = plan.records.data();
  // symbolic pool: tag = 1 + vid (vid: message of v -> v-N ; partial of c -> (N-1)+(c-N)), 0 = nothing
  const int NB = g.n_pool;
  std::vector<int> tag(NB, 0), inflight_until(NB, 0), store_busy(NB, -1);
  std::map<int, int> stored_at;  // vid -> iteration the store was issued
  auto blk_of = [&](uint32_t soff) { return ((int)soff - g.off_pool) / g.blk; };
  auto check_soff = [&](uint32_t soff) { return (int)soff >= g.off_pool && ((int)soff - g.off_pool) % g.blk == 0 && blk_of(soff) < NB; };
  long long n_loads = 0, n_stores = 0, n_tasks = 0, n_rec_jobs = 0;
  std::vector<int> done_post(2 * N - 1, 0), done_pre(2 * N - 1, 0);
  auto do_load = [&](const WalkJob& j, int iter, int f) -> const char* {
    if (j.bytes % 16 || j.soff % 16 || j.goff % 16) return "alignment";
    if (j.kind == JK_MSG || j.kind == JK_PRE) {
      if (!check_soff(j.soff) || (int)j.bytes != g.blk || j.goff % g.blk) return "bad block job";
      const int b = blk_of(j.soff), node = N + (int)(j.goff / g.blk);
      const int vid = j.kind == JK_MSG ? node - N : (N - 1) + (node - N);
      if (!stored_at.count(vid)) return "load of a value that was never stored";
      if (iter < stored_at[vid] + 2) return "load issued before its store completed";
      if (store_busy[b] == iter) return "load into a block a store is reading";
      if (inflight_until[b] > iter) return "load into a block with a load in flight";
      tag[b] = 1 + vid; inflight_until[b] = f;
      ++n_loads;
    } else if (j.kind == JK_TIPS) {
      if ((int)j.soff != g.off_tips + (f % S) * g.frame_bytes || (int)j.bytes > g.tips_bytes ||
          j.goff != (uint32_t)f * 2u * (uint32_t)g.wn * (uint32_t)g.ct_pad || j.bytes % (2 * g.ct_pad)) return "tips job";
    } else if (j.kind == JK_REC) {
      if (j.goff != plan.rec_off[f] || (int)j.soff != g.off_frame + (f % S) * g.frame_bytes || j.bytes != plan.rec_off[f + 1] - plan.rec_off[f]) return "record job";
      ++n_rec_jobs;
    } else return "kind";
    return nullptr;
  };
  // boot
  {
    const WalkJob* jobs = reinterpret_cast<const WalkJob*>(rec + plan.boot_base);
    int j0 = 0;
    for (int f = 0; f < D && f < R; ++f) {
      uint32_t tx = 0;
      for (uint32_t j = 0; j < plan.boot_n[f]; ++j) {
        if (const char* e = do_load(jobs[j0 + j], 0, f)) FAIL("boot round %d: %s", f, e);
        tx += jobs[j0 + j].bytes;
      }
      if (tx != plan.boot_tx[f]) FAIL("boot tx");
      j0 += plan.boot_n[f];
    }
  }
  for (int q = 0; q < R; ++q) {
    const uint8_t* ir = rec + plan.rec_off[q] + g.rel_ir;
    WalkIssueHeader h;
    memcpy(&h, ir, 16);
    const WalkJob* jobs = reinterpret_cast<const WalkJob*>(ir + 16 + g.ir_ids);
    if (h.pass != (q >= Rpost ? 1u : 0u)) FAIL("pass flag round %d", q);
    if ((int)(h.n_loads + h.n_stores) > g.max_jobs) FAIL("too many jobs");
    if (q + D >= R && h.n_loads) FAIL("loads past the end");
    // stores first mark their blocks busy for this round (both are issued at the top of round q)
    for (uint32_t j = h.n_loads; j < h.n_loads + h.n_stores; ++j) {
      const WalkJob& jb = jobs[j];
      if (!(jb.kind == JK_MSG || jb.kind == JK_PRE) || !check_soff(jb.soff) || (int)jb.bytes != g.blk) FAIL("store job");
      const int b = blk_of(jb.soff), node = N + (int)(jb.goff / g.blk);
      const int vid = jb.kind == JK_MSG ? node - N : (N - 1) + (node - N);
      if (tag[b] != 1 + vid) FAIL("round %d: store of value %d finds tag %d", q, vid, tag[b] - 1);
      if (inflight_until[b] > q) FAIL("store from a block with a load in flight");
      if (stored_at.count(vid)) FAIL("value %d stored twice", vid);
      stored_at[vid] = q;
      store_busy[b] = q;
      ++n_stores;
    }
    uint32_t tx = 0;
    for (uint32_t j = 0; j < h.n_loads; ++j) {
      if (const char* e = do_load(jobs[j], q, q + D)) FAIL("round %d load %u: %s", q, j, e);
      tx += jobs[j].bytes;
    }
    if (tx != h.tx_bytes) FAIL("tx bytes round %d", q);
    // tasks of round q
    const WalkDesc* d = reinterpret_cast<const WalkDesc*>(rec + plan.rec_off[q] + g.rel_cr);
    // what the fold of the NEXT round is told about this round's sums
    const WalkIds* ids = reinterpret_cast<const WalkIds*>(q + 1 < R ? rec + plan.rec_off[q + 1] + g.rel_ir + 16 : rec + plan.tail_base);
    if (q == Rpost - 1) {  // a run that stops here reads the second tail record: must say the same
      const WalkIds* t2 = reinterpret_cast<const WalkIds*>(rec + plan.tail_base + g.ir_ids);
      for (int s = 0; s < wn; ++s) if (t2[s].a != ids[s].a || t2[s].b != ids[s].b) FAIL("post-only tail ids");
    }
    const uint32_t frame0 = (uint32_t)(g.off_frame + (q % S) * g.frame_bytes), mats0 = frame0 + plan.rec_mats[q];
    if (plan.rec_mats[q] != (uint32_t)g.rel_mats + 16u * (h.n_loads + h.n_stores)) FAIL("round %d: matrix rows do not follow the last job", q);
    auto mat_goff = [&](uint32_t soff) { return plan.rec_off[q] + (soff - frame0); };
    auto in_mats = [&](uint32_t soff, int rows) { return soff >= mats0 && soff + (uint32_t)rows * g.row_stride <= frame0 + (plan.rec_off[q + 1] - plan.rec_off[q]); };
    std::vector<int> reads, writes;
    std::vector<std::pair<int, int>> produced;  // (block, vid)
    auto off_blk = [&](uint32_t o) { return blk_of(o); };
    for (int s = 0; s < wn; ++s) {
      const WalkDesc& e = d[s];
      if (!(e.flags & WF_LIVE)) { if (ids[s].a != -1 || ids[s].b != -1) FAIL("round %d: fold ids of an idle slot", q); continue; }
      ++n_tasks;
      int v;
      if (q < Rpost) {
        if (e.flags & WF_ROOT) { v = root; if (ids[s].a != -2 || ids[s].b != -1) FAIL("root fold ids"); }
        else {
          if (!in_mats(e.a_v, kRowsPlain)) FAIL("round %d slot %d: own matrix rows outside the bundle", q, s);
          auto it = post_node_at.find(mat_goff(e.a_v));
          if (it == post_node_at.end()) FAIL("round %d slot %d: unknown post task", q, s);
          v = it->second;
          if (ids[s].a != -1 || ids[s].b != -1) FAIL("round %d: fold ids of a post-order task", q);
        }
        if (done_post[v]++) FAIL("post task %d twice", v);
      } else {
        if (ids[s].a < 0 || ids[s].b < 0 || parent[ids[s].a] != parent[ids[s].b]) FAIL("ids");
        v = parent[ids[s].a];
        if (left[v] != ids[s].a || right[v] != ids[s].b) FAIL("ids order");
        if ((int)(e.ids_lo & 0xffffffu) != ids[s].a || (int)((e.ids_lo >> 24) | ((uint32_t)e.ids_hi << 8)) != ids[s].b) FAIL("descriptor branch ids");
        if (done_pre[v]++) FAIL("pre task %d twice", v);
        if (v != root && !done_pre[parent[v]]) FAIL("pre task %d before its parent", v);
      }
      const int l = left[v], r = right[v];
      if (((e.flags & WF_TIP_L) != 0) != (l < N) || ((e.flags & WF_TIP_R) != 0) != (r < N)) FAIL("tip flags");
      if (r < N && l >= N) FAIL("a lone tip child must be the left one");
      if (s > 0 && (d[s - 1].flags & WF_LIVE)) {  // tasks of a round are laid out by kind
        auto kind = [&](const WalkDesc& x) { return ((x.flags & WF_TIP_L) != 0) + ((x.flags & WF_TIP_R) != 0); };
        if (kind(d[s - 1]) > kind(e)) FAIL("round %d: tasks not sorted by kind", q);
      }
      const int w = q < Rpost ? 0 : 1;
      const int ch[2] = {l, r};
      const uint32_t ao[2] = {e.a_l, e.a_r};
      for (int side = 0; side < 2; ++side) {
        const int want = ch[side] < N ? ch[side] : -1;
        if (plan.tip_order[((size_t)q * wn + s) * 2 + side] != want) FAIL("round %d slot %d: tip row of side %d", q, s, side);
        if (ch[side] < N) {
          if (!in_mats(ao[side], kRowsTip) || (int)mat_goff(ao[side]) != plan.mat_dst[2 * ch[side] + w]) FAIL("round %d slot %d: tip table of side %d", q, s, side);
        }
      }
      auto need = [&](uint32_t src, int vid, const char* what) -> bool {
        if (!check_soff(src)) { printf("FAIL round %d node %d: %s operand offset\n", q, v, what); return false; }
        const int b = off_blk(src);
        if (tag[b] != 1 + vid) { printf("FAIL round %d node %d: %s operand block %d holds %d, want %d\n", q, v, what, b, tag[b] - 1, vid); return false; }
        if (inflight_until[b] > q) { printf("FAIL round %d node %d: %s operand load lands in round %d\n", q, v, what, inflight_until[b]); return false; }
        reads.push_back(b);
        return true;
      };
      if (l >= N && !need(e.a_l, l - N, "left")) return 1;
      if (r >= N && !need(e.a_r, r - N, "right")) return 1;
      if (q < Rpost) {
        if (!done_post[l >= N ? l : v] || !done_post[r >= N ? r : v]) FAIL("post task %d before a child", v);
        if (v != root) {
          if (!check_soff(e.dst0)) FAIL("round %d: message block offset", q);
          writes.push_back(off_blk(e.dst0)); produced.push_back({off_blk(e.dst0), v - N});
        }
      } else {
        if (v == root) { if ((int)e.a_v != g.off_freq) FAIL("root parent-side operand"); }
        else if (!need(e.a_v, (N - 1) + (v - N), "parent-side")) return 1;
        const uint32_t dsts[2] = {e.dst0, e.dst1}, ms[2] = {(uint32_t)e.m_l * 16u, (uint32_t)e.m_r * 16u};
        for (int side = 0; side < 2; ++side) {
          if (ch[side] < N) continue;
          if (!check_soff(dsts[side])) FAIL("round %d: partial block offset", q);
          if (!in_mats(ms[side], kRowsPlain) || (int)mat_goff(ms[side]) != plan.mat_dst[2 * ch[side] + 1]) FAIL("round %d slot %d: matrix rows of side %d", q, s, side);
          writes.push_back(off_blk(dsts[side])); produced.push_back({off_blk(dsts[side]), (N - 1) + (ch[side] - N)});
        }
      }
    }
    for (size_t i = 0; i < writes.size(); ++i) {
      const int b = writes[i];
      if (b < 0 || b >= NB) FAIL("round %d: result block out of range", q);
      if (inflight_until[b] > q) FAIL("round %d: result into a block with a load in flight (lands %d)", q, inflight_until[b]);
      if (store_busy[b] == q) FAIL("round %d: result into block %d that a store is reading", q, b);
      for (int rb : reads) if (rb == b) FAIL("round %d: block %d read and written in the same round", q, b);
      for (size_t k = 0; k < i; ++k) if (writes[k] == b) FAIL("round %d: block %d written twice", q, b);
    }
    for (auto& pr : produced) tag[pr.first] = 1 + pr.second;
  }
  if (n_rec_jobs != R) FAIL("every round's record must be loaded exactly once (%lld of %d)", n_rec_jobs, R);
  for (int v = N; v <= root; ++v) if (!done_post[v] || !done_pre[v]) FAIL("task %d missing", v);
  for (int v = N; v < root; ++v) if (!stored_at.count(v - N)) FAIL("message %d never stored", v);
  printf("OK N=%d shape=%d ct=%d cpt=%d K=%d elem=%d depth=%d nt=%d wn=%d wn_eff=%d pool=%d rounds=%d+%d tasks=%lld resident_reads=%lld loads=%lld stores=%lld evictions=%lld max_jobs=%d/%d smem=%zu\n",
         N, shape, g.ct, g.cpt, K, elem, D, g.nt, wn, plan.wn_eff, NB, Rpost, plan.rounds_pre, n_tasks, plan.resident_reads, n_loads, n_stores,
         plan.evictions, plan.max_jobs_used, g.max_jobs, g.smem);
  return 0;
}
