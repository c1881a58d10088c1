"""CPU checks of the traversal kernel's static plan (no GPU): the host planner
(csrc/b2p_plan.h) is compiled with g++ into tests/tools/plan_check.cpp, which replays the
emitted descriptors and copy lists against a symbolic shared-memory pool and verifies the
timing rules the kernel relies on (operand identity, loads in flight, reuse distances,
transaction byte counts, every message stored once)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_build", "plan_check")
SMEM = 232448 - 1024  # sharedMemPerBlockOptin on B200 minus static + reservation


@pytest.fixture(scope="module")
def plan_check():
    os.makedirs(os.path.dirname(BIN), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", BIN, os.path.join(ROOT, "tests", "tools", "plan_check.cpp")])
    return BIN


# N, shape (0 random joins, 1 caterpillar, 2 balanced), seed, ct, K, elem, depth, nt, cpt[, outer-product mode]
CASES = [
    (1000, 0, 1, 68, 1, 8, 1, 512, 1, 1),   # outer-product mode at the headline shape
    (50, 0, 3, 12, 4, 8, 1, 512, 1, 1),     # narrow tile: 32 threads per slot, 13 slots
    (1704, 0, 2, 48, 4, 8, 1, 512, 1, 1),
    (300, 2, 5, 136, 4, 8, 2, 512, 1, 1),
    (1000, 0, 1, 68, 1, 8, 1, 0, 2),     # headline geometry
    (1000, 0, 1, 68, 1, 8, 1, 0, 1),
    (1000, 0, 7, 68, 1, 8, 2, 0, 2),
    (1704, 0, 2, 48, 4, 8, 1, 0, 2),     # HKY+G4 flu shape
    (1704, 0, 2, 48, 4, 4, 1, 0, 1),     # fp32
    (5000, 0, 5, 64, 4, 8, 1, 0, 2),     # config 4 shard
    (2000, 1, 3, 96, 1, 8, 1, 0, 2),     # caterpillar: depth 2000
    (130, 2, 3, 16, 4, 4, 1, 0, 2),      # balanced, narrow tile, many slots
    (64, 0, 9, 9, 3, 8, 1, 0, 1),        # K does not divide 32
    (2, 0, 1, 8, 1, 8, 1, 0, 2),
    (3, 0, 1, 8, 1, 8, 3, 0, 1),
    (257, 0, 4, 128, 4, 8, 3, 256, 1),   # few threads, deep copy pipeline
]


@pytest.mark.parametrize("case", CASES)
def test_plan_is_consistent(plan_check, case):
    n, shape, seed, ct, k, elem, depth, nt, cpt = case[:9]
    outer = case[9] if len(case) > 9 else 0
    out = subprocess.run([plan_check, *map(str, (n, shape, seed, ct, k, elem, SMEM, depth, nt, cpt, outer))],
                         capture_output=True, text=True).stdout.strip()
    assert out.startswith("OK"), out
