"""Small helpers shared by the host layer (re-authored from the behaviour of
``msprime/core.py:58-124``: per-process seed generator and strict boolean flags)."""
from __future__ import annotations

import os
import random

__version__ = "0.1.0"

_seed_rngs: dict[int, random.Random] = {}


def get_random_seed() -> int:
    """A fresh seed in 1 .. 2**32 - 1, from a generator private to this process
    (``msprime/core.py:72-81``)."""
    pid = os.getpid()
    rng = _seed_rngs.get(pid)
    if rng is None:
        rng = _seed_rngs[pid] = random.Random()
    return rng.randint(1, 2**32 - 1)


def set_seed_rng_seed(seed: int) -> None:
    _seed_rngs[os.getpid()] = random.Random(seed)


def check_seed(seed: int) -> int:
    """``_msprime.RandomGenerator`` accepts 0 < seed < 2**32 (``_msprimemodule.c:255-274``)."""
    seed = int(seed)
    if seed <= 0 or seed >= 2**32:
        raise ValueError("seeds must be greater than 0 and less than 2^32")
    return seed


def parse_flag(value, *, default: bool) -> bool:
    """None -> default; anything that is not a ``bool`` is a TypeError
    (``msprime/core.py:107-124``): truthy values are deliberately not accepted."""
    if value is None:
        return default
    if not isinstance(value, bool):
        raise TypeError("Boolean flag must be True, False, or None (the default value)")
    return value
