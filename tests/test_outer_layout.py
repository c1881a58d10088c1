"""CPU check of the staging layout of the outer-product sums (csrc/b2p_walk.cuh: outer_addresses,
outer_column, outer_swizzle — restated here line by line): every lane's stores land where the
lanes of the tensor-core products look for them, nothing overlaps, and both directions are free
of shared-memory bank conflicts (8 lanes x 16 bytes per store phase, 16 lanes x 8 bytes per load
phase, 32 banks of 4 bytes)."""
import pytest


def swizzle(row):
    return ((((row >> 1) ^ (row >> 3)) & 1) << 1) | ((row >> 2) & 1)


def column(k_cat, d, t):
    return 4 * d + t if k_cat == 1 else (d & 3) + 4 * (t + 4 * (d >> 2))


def store_addresses(lane):
    """Byte offsets (in the warp's 2 KB A half) of the four 16-byte chunks lane `lane` stores."""
    row = lane ^ ((lane >> 2) & 1)
    sw = swizzle(row)
    return [row * 64 + ((c ^ sw) << 4) for c in range(4)]


def load_address(k_cat, lane, d):
    """Byte offset lane `lane` loads its A element of product d from: ld[d % 4] + (d / 4) * 1024."""
    g, t = lane >> 2, lane & 3
    c = d & 3
    col = column(k_cat, c, t)
    r2 = col ^ ((col >> 2) & 1)
    return r2 * 64 + (((g >> 1) ^ swizzle(r2)) << 4) + ((g & 1) << 3) + (d >> 2) * 1024


def test_stores_tile_the_staging_half_without_overlap():
    seen = set()
    for lane in range(32):
        for a in store_addresses(lane):
            assert a % 16 == 0 and 0 <= a < 2048
            assert a not in seen
            seen.add(a)
    assert len(seen) == 128


@pytest.mark.parametrize("k_cat", [1, 4])
def test_loads_find_what_the_owning_lane_stored(k_cat):
    """Lane (g, t) of product d must read element g (of the 8 A values) of column column(d, t):
    that is chunk g // 2 of the owning lane's row, half g % 2."""
    for d in range(8):
        cols = set()
        for lane in range(32):
            g, t = lane >> 2, lane & 3
            col = column(k_cat, d, t)
            assert 0 <= col < 32
            cols.add(col)
            want = store_addresses(col)[g >> 1] + 8 * (g & 1)
            assert load_address(k_cat, lane, d) == want, (k_cat, d, lane)
        assert len(cols) == 4
        if k_cat == 4:  # the four columns of a product share a category: lane % 4 of their owners
            assert {c % 4 for c in cols} == {d & 3}
    # every column is used by exactly one product (per category at K = 4: two products of four columns)
    used = sorted(column(k_cat, d, t) for d in range(8) for t in range(4))
    assert used == list(range(32))


def _conflicts(addresses_and_sizes):
    """Worst number of distinct 4-byte words that fall on one bank."""
    per_bank = {}
    for a, size in addresses_and_sizes:
        for w in range(a // 4, (a + size) // 4):
            per_bank.setdefault(w % 32, set()).add(w)
    return max(len(v) for v in per_bank.values())


def test_store_phases_are_conflict_free():
    for c in range(4):
        for phase in range(4):  # a 16-byte store of a warp goes out in four phases of 8 lanes
            lanes = range(8 * phase, 8 * phase + 8)
            assert _conflicts([(store_addresses(l)[c], 16) for l in lanes]) == 1


@pytest.mark.parametrize("k_cat", [1, 4])
def test_load_phases_are_conflict_free(k_cat):
    for d in range(8):
        for phase in range(2):  # an 8-byte load of a warp goes out in two phases of 16 lanes
            lanes = range(16 * phase, 16 * phase + 16)
            assert _conflicts([(load_address(k_cat, l, d), 8) for l in lanes]) == 1, (k_cat, d, phase)
