"""GPU: the device radix sort (stable, payload-carrying) on its own, across sizes that
exercise empty/one/partial/many tiles, constant keys, sparse bit masks and both key widths."""
import pytest

from msprime_b200 import _native

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(native_built):
    c = _native.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("key_bytes", [4, 8])
@pytest.mark.parametrize("n", [0, 1, 2, 31, 4096, 4097, 100_003, 3_000_001])
def test_sort_sizes(ctx, key_bytes, n):
    ctx.selftest_sort(n, key_bytes=key_bytes, seed=n + 1)


@pytest.mark.parametrize("mask,expect", [
    (0x0, 0),                        # all keys equal
    (0x1, 1),                        # one bit
    (0xFF00, 1),
    (0x0FFFFFFF, 4),                 # 28 bits -> 4 passes of 7
    (0xFF000000000000FF, 2),         # far apart digits: the constant middle is skipped
    (0xFFFFFFFFFFFFFFFF, 8),
])
def test_sort_masks(ctx, mask, expect):
    for measure in (True, False):
        if mask == 0 and not measure:
            continue
        passes = ctx.selftest_sort(200_000, key_bytes=8, key_mask=mask, seed=7, measure_bits=measure)
        assert passes == expect


def test_sort_many_duplicates_is_stable(ctx):
    # 3 distinct keys over 1e6 elements: long equal runs across many tiles
    ctx.selftest_sort(1_000_000, key_bytes=4, key_mask=0x3, seed=11)
