// Plain-C++ records shared by the host planner (b2p_plan.h) and the traversal kernel
// (b2p_walk.cuh).  No CUDA types here: the planner is also compiled by g++ for the
// CPU-side plan checker (tests/tools/plan_check.cpp).
//
// Vocabulary
//   column   one (pattern, category) pair; a node's partial for one column is 4 states.
//   tile     ct adjacent columns, owned by one CTA for the whole evaluation.
//   block    one node's 4-state vectors for the ct columns of a tile (ct*4*sizeof(T) bytes);
//            the unit of every copy between HBM and shared memory.
//   round    one step of the static schedule: up to wn node tasks run side by side.
//   pool     shared-memory blocks handed out by the planner (static allocation: the
//            schedule is known, so block lifetimes are too).
//   row      one matrix row (or tip-table row) for all K categories: K * 4 values + padding
//            (row_stride bytes); a thread of category k reads its 4 values at k * 4 * sizeof(T).
#pragma once
#include <stdint.h>

namespace b2p {

// task flags
enum : uint16_t {
  WF_LIVE = 1,    // slot holds a task this round
  WF_ROOT = 2,    // post: root task (site likelihoods, no message); pre: parent-side operand = frequency block
  WF_TIP_L = 4,   // left child is a tip: its message is a row of its table, picked by the tip's state code
  WF_TIP_R = 8,
};

// Rows a task's matrices take in the round's matrix bundle.
//   internal node / child: 4 rows (P, row i = P[i][.])
//   tip child:             5 rows (row s = column s of P for s = 0..3, row 4 = all ones: gap / unknown)
constexpr int kRowsPlain = 4, kRowsTip = 5;
constexpr int kRowsTaskMax = kRowsPlain + 2 * kRowsTip;  // post-order task with two tip children

// Tip state codes in the per-tile tip rows (tile_tips_kernel): 0..3 = the state, 4 = any state,
// 16 | mask = another ambiguity mask (summed row by row in the kernel's slow path).
constexpr unsigned kTipAny = 4, kTipAmbiguous = 16;

// What the compute threads of one slot need for one round: 32 bytes, two LDS.128.
// All offsets are byte offsets from the start of dynamic shared memory, written by the planner
// for the stage the round runs in — the kernel adds its per-thread column / category offset.
struct WalkDesc {
  uint32_t a_l, a_r;     // children: pool block of the message, or (tip) its 5-row table
  uint32_t a_v;          // post: the node's own matrix rows; pre: parent-side partial (root: frequency block)
  uint32_t dst0, dst1;   // post: dst0 = message of the node; pre: partials of the internal children
  uint16_t m_l, m_r;     // pre: matrix rows of the internal children, offset / 16
  uint16_t flags, ids_hi;
  uint32_t ids_lo;       // pre: the children's branch ids, 24 bits each: left = ids_lo & 0xffffff,
                         // right = ids_lo >> 24 | ids_hi << 8 (the outer-product mode's reduction reads them)
};
static_assert(sizeof(WalkDesc) == 32, "WalkDesc must be 32 bytes");

// Branch ids a pre-order task's two gradient sums belong to (-1: none, -2: log-likelihood term).
struct WalkIds {
  int32_t a, b;
};

// One bulk copy.  goff is relative to the per-kind base (tile-relative for MSG/PRE/TIPS).
enum : uint32_t { JK_MSG = 0, JK_PRE = 1, JK_TIPS = 2, JK_REC = 3 };
struct WalkJob {
  uint32_t goff;   // bytes
  uint32_t soff;   // shared-memory byte offset
  uint32_t bytes;  // multiple of 16
  uint32_t kind;
};
static_assert(sizeof(WalkJob) == 16, "WalkJob must be 16 bytes");

// Issue record of one round (read at the top of the round):
// header, then wn fold ids (what the sums of round q - 1 belong to), then n_loads load jobs
// (operands of round q + depth), then n_stores store jobs (results of round q - 1).
// A round's descriptors, issue record and matrix rows form ONE record — one bulk copy.
struct WalkIssueHeader {
  uint32_t n_loads, n_stores;
  uint32_t tx_bytes;  // sum of the load jobs' bytes (mbarrier transaction count)
  uint32_t pass;      // 0: post-order round, 1: pre-order round (of round q itself)
};
static_assert(sizeof(WalkIssueHeader) == 16, "WalkIssueHeader must be 16 bytes");

}  // namespace b2p
