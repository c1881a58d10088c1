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
#pragma once
#include <stdint.h>

namespace b2p {

constexpr uint16_t kNoBlk = 0xFFFFu;  // operand is a tip (child) / the root prior (parent side)

// task flags
enum : uint16_t {
  WF_LIVE = 1,    // slot holds a task this round
  WF_ROOT = 2,    // post: root task (site likelihoods, no message); pre: parent-side operand = frequencies
  WF_TIP_L = 4,   // left child is a tip: its message is a column of P^T picked by the tip's state mask
  WF_TIP_R = 8,
  WF_OUT_L = 16,  // pre: left child is internal -> its pre-order partial is produced (dst0)
  WF_OUT_R = 32,  // pre: right child internal (dst1)
};

// What the compute threads of one slot need for one round: 16 bytes, one LDS.128.
// Offsets are shared-memory byte offsets / 16.
struct WalkDesc {
  uint16_t src_l, src_r;  // children's messages (kNoBlk: tip)
  uint16_t src_v;         // pre: parent-side partial of the node (kNoBlk: root)
  uint16_t dst0, dst1;    // post: dst0 = message of the node; pre: partials of the children
  uint16_t flags;
  uint16_t pad0, pad1;    // tip rows (within the round's tip-row copy) of the left / right child
};
static_assert(sizeof(WalkDesc) == 16, "WalkDesc must be 16 bytes");

// Branch ids a pre-order task's two gradient sums belong to (-1: none, -2: log-likelihood term).
struct WalkIds {
  int32_t a, b;
};

// One bulk copy.  goff is relative to the per-kind base (tile-relative for MSG/PRE/TIPS).
enum : uint32_t { JK_MSG = 0, JK_PRE = 1, JK_TIPS = 2, JK_MATS = 3, JK_REC = 4 };
struct WalkJob {
  uint32_t goff;   // bytes
  uint32_t soff;   // shared-memory byte offset
  uint32_t bytes;  // multiple of 16
  uint32_t kind;
};
static_assert(sizeof(WalkJob) == 16, "WalkJob must be 16 bytes");

// Issue record of one round (read by the copy warp at the top of the round):
// header, then n_loads load jobs (operands of round q + depth), then n_stores store jobs
// (results of round q - 1).
struct WalkIssueHeader {
  uint32_t n_loads, n_stores;
  uint32_t tx_bytes;  // sum of the load jobs' bytes (mbarrier transaction count)
  uint32_t pass;      // 0: post-order round, 1: pre-order round (of round q itself)
};
static_assert(sizeof(WalkIssueHeader) == 16, "WalkIssueHeader must be 16 bytes");

}  // namespace b2p
