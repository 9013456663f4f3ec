"""Seed generator and strict flags (msprime_b200/core.py), after the reference's tests/test_core.py:
default seeds are valid, do not repeat, and stay distinct across forked worker processes (each
process gets its own generator, so a forked child never replays its parent's stream)."""
import multiprocessing

import pytest

from msprime_b200 import core


def _seed_in_worker(_):
    return core.get_random_seed()


def test_default_seeds_are_valid_and_distinct():
    seeds = [core.get_random_seed() for _ in range(1000)]
    assert all(0 < s < 2**32 for s in seeds)
    assert len(set(seeds)) == len(seeds)
    assert all(core.check_seed(s) == s for s in seeds)


def test_forked_workers_draw_different_seeds():
    core.get_random_seed()  # the parent's generator exists before the fork
    with multiprocessing.get_context("fork").Pool(4) as pool:
        seeds = pool.map(_seed_in_worker, range(100))
    assert len(set(seeds)) == len(seeds)


def test_seeding_the_seed_generator_is_reproducible():
    core.set_seed_rng_seed(5)
    a = [core.get_random_seed() for _ in range(5)]
    core.set_seed_rng_seed(5)
    assert a == [core.get_random_seed() for _ in range(5)]


@pytest.mark.parametrize("bad", [0, -1, 2**32, 2**40])
def test_seed_range(bad):
    with pytest.raises(ValueError, match="greater than 0 and less than"):
        core.check_seed(bad)
    assert core.check_seed(2**32 - 1) == 2**32 - 1 and core.check_seed(1) == 1


def test_flags_must_be_real_booleans():
    assert core.parse_flag(None, default=True) is True and core.parse_flag(None, default=False) is False
    assert core.parse_flag(False, default=True) is False
    for bad in (0, 1, "", "yes", [], {}, 1.0):
        with pytest.raises(TypeError, match="Boolean flag"):
            core.parse_flag(bad, default=True)
