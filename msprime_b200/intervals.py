"""``RateMap`` for the mutation path.

The surface the overlay needs from ``msprime/intervals.py`` (:35-97 constructor contract, :292
``uniform``, :447 ``asdict``, the read-only interval properties and the two point queries):
keyword-only ``position`` (n + 1 break points, first one 0, strictly increasing) and ``rate``
(n non-negative values, NaN = missing data).  Slicing, hapmap I/O and ``RecombinationMap`` are
out of scope.  All derived arrays are computed once, in the constructor, and frozen.
"""
from __future__ import annotations

import collections.abc

import numpy as np


def _frozen(values) -> np.ndarray:
    a = np.array(values, dtype=np.float64)
    a.setflags(write=False)
    return a


def _indexes(mask, shift=0):
    return np.flatnonzero(mask) + shift


class RateMap(collections.abc.Mapping):
    """Piecewise-constant non-negative rate over ``[0, sequence_length)``: interval ``i`` is
    ``[position[i], position[i + 1])`` with rate ``rate[i]``."""

    def __init__(self, *, position, rate):
        pos, r = _frozen(position), _frozen(rate)
        if pos.size < 2:
            raise ValueError("Must have at least two positions")
        if r.size + 1 != pos.size:
            raise ValueError("Rate array must have one less entry than the position array")
        if pos[0] != 0:
            raise ValueError("First position must be zero")
        widths = np.diff(pos)
        if (widths <= 0).any():
            raise ValueError(f"Position values not strictly increasing at indexes {_indexes(widths <= 0, 1)}")
        if (r < 0).any():
            raise ValueError(f"Rate values negative at indexes {_indexes(r < 0)}")
        gaps = np.isnan(r)
        if gaps.all():
            raise ValueError("All intervals are missing data")
        self._cols = {
            "position": pos,
            "rate": r,
            "left": pos[:-1],
            "right": pos[1:],
            "span": _frozen(widths),
            "mid": _frozen(pos[:-1] + widths / 2),
            "mass": _frozen(r * widths),
            "missing": gaps,
        }
        self._cumulative_mass = _frozen(np.concatenate([[0.0], np.nancumsum(r * widths)]))

    @staticmethod
    def uniform(sequence_length, rate) -> "RateMap":
        """One interval covering the whole sequence."""
        return RateMap(position=[0, sequence_length], rate=[rate])

    # read-only columns
    position = property(lambda self: self._cols["position"])
    rate = property(lambda self: self._cols["rate"])
    left = property(lambda self: self._cols["left"])
    right = property(lambda self: self._cols["right"])
    span = property(lambda self: self._cols["span"])
    mid = property(lambda self: self._cols["mid"])
    mass = property(lambda self: self._cols["mass"])
    missing = property(lambda self: self._cols["missing"])
    num_intervals = property(lambda self: int(self._cols["rate"].size))
    sequence_length = property(lambda self: self._cols["position"][-1])
    total_mass = property(lambda self: self._cumulative_mass[-1])

    @property
    def mean_rate(self) -> float:
        known = ~self.missing
        return float(self.mass[known].sum() / self.span[known].sum())

    def get_cumulative_mass(self, x):
        """Mass of ``[0, x)``, linear inside an interval."""
        q = np.asarray(x, dtype=np.float64)
        if ((q < 0) | (q > self.sequence_length)).any():
            raise ValueError(f"Cannot have positions < 0 or > {self.sequence_length}")
        return np.interp(q, self.position, self._cumulative_mass)

    def get_rate(self, x):
        """Rate of the interval that contains ``x``."""
        q = np.asarray(x, dtype=np.float64)
        if ((q < 0) | (q >= self.sequence_length)).any():
            raise ValueError("Rate only defined for 0 <= x < sequence_length")
        return self.rate[np.searchsorted(self.position, q, side="right") - 1]

    def asdict(self):
        """The two arrays the C call takes (``{"position": ..., "rate": ...}``)."""
        return {name: self._cols[name] for name in ("position", "rate")}

    def copy(self) -> "RateMap":
        return RateMap(position=np.array(self.position), rate=np.array(self.rate))

    # Mapping protocol: keys are the midpoints of the intervals that have data
    def __iter__(self):
        return iter(self.mid[~self.missing])

    def __len__(self):
        return int((~self.missing).sum())

    def __getitem__(self, key):
        if isinstance(key, slice):
            raise TypeError("RateMap slicing is out of scope for msprime_b200")
        return self.get_rate(key)

    def __repr__(self):
        return f"RateMap(position={self.position!r}, rate={self.rate!r})"
