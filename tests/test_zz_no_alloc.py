"""The no-allocation entry point ``b2p_coalescent_constant_noalloc`` (include/b200phylo.h,
csrc/b2p_noalloc.cu) against the existing ``b2p_coalescent_constant``, both through the C ABI.

No counterpart in the reference: the entry point was added
for callers that cannot call cudaMalloc.  Its scratch is a file-scope ``static __device__`` array
shared by every call, so each case makes TWO calls in a row (different heights) and checks both.
Tolerances are those of tests/test_gpu_timetree.py: the coalescent sum and d/d theta to 1e-12 /
1e-11 relative (the two kernels cut the sum into different per-CTA parts), the height gradient
-(k-1)/theta bit for bit.

The file sorts last on purpose: a failure here must not
hide the tests of the path proper.
"""
import ctypes

import numpy as np
import pytest
import torch

from b200phylo.tree import simulate_tree
from oracle import timetree_oracle as tto

pytestmark = pytest.mark.gpu


def _lib():
    from b200phylo.engine import load_library
    lib = load_library()
    vp, i32 = ctypes.c_void_p, ctypes.c_int32
    lib.b2p_coalescent_constant_noalloc.argtypes = [i32, i32, vp, vp, vp, vp, vp, vp, vp]
    lib.b2p_coalescent_constant_noalloc.restype = ctypes.c_int
    lib.b2p_noalloc_last_error.restype = ctypes.c_char_p
    lib.b2p_noalloc_max_tips.restype = i32
    return lib


def _case(n, shape, seed, ties):
    tree = simulate_tree(n, shape=shape, seed=seed)
    rng = np.random.default_rng(seed + 1)
    tips = rng.uniform(0.0, 12.0, n)
    if ties:
        tips = np.round(tips)  # many tips share a sampling date
    tips[rng.integers(n)] = 0.0
    out = []
    for _ in range(2):  # two sets of heights on the same tree: the two calls in a row differ
        params = rng.uniform(0.05, 0.95, n - 1)
        if ties:
            params[rng.integers(0, max(1, n - 2), size=max(1, n // 10))] = 0.0  # nodes sitting on their bound
        params[n - 2] = tips.max() + rng.uniform(0.5, 5.0)
        out.append(tto.heights_from_ratios(n, tree.post_triples, tips, params))
    return tree, tips, out


def cu(a):
    return torch.tensor(np.asarray(a, dtype=np.float64), device="cuda")


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


@pytest.mark.parametrize("n,shape,seed,ties", [(2, "balanced", 1, False), (3, "caterpillar", 2, False), (40, "coalescent", 3, False),
                                               (333, "coalescent", 4, True), (2000, "caterpillar", 6, True),
                                               (5000, "coalescent", 8, False)])
def test_noalloc_matches_the_existing_entry_point_twice_in_a_row(n, shape, seed, ties):
    from b200phylo.timetree import TimeTreeEngine
    lib = _lib()
    tree, tips, heights = _case(n, shape, seed, ties)
    eng = TimeTreeEngine(tree.post_triples, tips)
    d_tips, theta = cu(tips), 3.7
    d_theta = cu([theta])
    stream = torch.cuda.current_stream().cuda_stream
    try:
        for h in heights:
            d_h = cu(h)
            logp0, gh0, gt0 = eng.constant_coalescent(d_h, d_theta)  # the existing entry point
            logp, gt = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda"), torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
            gh = torch.full((n - 1,), float("nan"), dtype=torch.float64, device="cuda")
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated()
            rc = lib.b2p_coalescent_constant_noalloc(0, n, d_tips.data_ptr(), d_h.data_ptr(), d_theta.data_ptr(), logp.data_ptr(),
                                                     gh.data_ptr(), gt.data_ptr(), stream)
            assert rc == 0, lib.b2p_noalloc_last_error().decode()
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated() == before
            assert rel(float(logp), float(logp0)) < 1e-12, (float(logp), float(logp0))
            np.testing.assert_array_equal(gh.cpu().numpy(), gh0.cpu().numpy())
            assert rel(float(gt), float(gt0)) < 1e-11
            # and both against the oracle, like tests/test_gpu_timetree.py::test_constant_coalescent
            wl, wh, wt = tto.constant_coalescent(n, tips, h, theta)
            assert rel(float(logp), wl) < 1e-12 and rel(float(gt), wt) < 1e-11
            np.testing.assert_array_equal(gh.cpu().numpy(), wh)
    finally:
        eng.close()


def test_noalloc_optional_outputs_and_limits():
    lib = _lib()
    n = 50
    tree, tips, heights = _case(n, "coalescent", 11, False)
    d_tips, d_h, d_theta = cu(tips), cu(heights[0]), cu([2.5])
    logp = torch.zeros(1, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    assert lib.b2p_coalescent_constant_noalloc(0, n, d_tips.data_ptr(), d_h.data_ptr(), d_theta.data_ptr(), logp.data_ptr(), None, None, stream) == 0
    torch.cuda.synchronize()
    wl, _, _ = tto.constant_coalescent(n, tips, heights[0], 2.5)
    assert rel(float(logp), wl) < 1e-12
    # sizes beyond the fixed scratch are rejected with a message, not clamped
    limit = lib.b2p_noalloc_max_tips()
    assert limit == 16384
    rc = lib.b2p_coalescent_constant_noalloc(0, limit + 1, d_tips.data_ptr(), d_h.data_ptr(), d_theta.data_ptr(), logp.data_ptr(), None, None, stream)
    assert rc == -4 and b"exceeds the fixed limit" in lib.b2p_noalloc_last_error()
    assert lib.b2p_coalescent_constant_noalloc(0, n, None, d_h.data_ptr(), d_theta.data_ptr(), logp.data_ptr(), None, None, stream) == -1
    assert lib.b2p_coalescent_constant_noalloc(0, 1, d_tips.data_ptr(), d_h.data_ptr(), d_theta.data_ptr(), logp.data_ptr(), None, None, stream) == -1
    assert lib.b2p_coalescent_constant_noalloc(0, n, d_tips.data_ptr(), d_h.data_ptr(), d_theta.data_ptr(), None, None, None, stream) == -1


def test_noalloc_at_the_limit():
    """The largest tree the entry point takes (16384 tips: 32767 events, 1024 per-CTA partial sums
    at most), against the oracle."""
    lib = _lib()
    n = lib.b2p_noalloc_max_tips()
    rng = np.random.default_rng(3)
    tips = rng.uniform(0.0, 10.0, n)
    tips[0] = 0.0
    h = np.sort(rng.uniform(10.0, 60.0, n - 1))  # any internal heights above the tips are a valid input
    d_tips, d_h, d_theta = cu(tips), cu(h), cu([4.0])
    logp, gt = torch.zeros(1, dtype=torch.float64, device="cuda"), torch.zeros(1, dtype=torch.float64, device="cuda")
    gh = torch.zeros(n - 1, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        rc = lib.b2p_coalescent_constant_noalloc(0, n, d_tips.data_ptr(), d_h.data_ptr(), d_theta.data_ptr(), logp.data_ptr(), gh.data_ptr(),
                                                 gt.data_ptr(), stream)
        assert rc == 0, lib.b2p_noalloc_last_error().decode()
    torch.cuda.synchronize()
    wl, wh, wt = tto.constant_coalescent(n, tips, h, 4.0)
    assert rel(float(logp), wl) < 1e-12 and rel(float(gt), wt) < 1e-11
    np.testing.assert_array_equal(gh.cpu().numpy(), wh)
