"""``RateMap``: the part of ``msprime/intervals.py:35-97,292,447`` that the mutation
path needs (constructor validation, ``uniform``, ``asdict`` and the basic read-only
properties).  Slicing, hapmap I/O and ``RecombinationMap`` are out of scope."""
from __future__ import annotations

import collections.abc

import numpy as np


class RateMap(collections.abc.Mapping):
    """Non-negative rates on consecutive half-open intervals ``[position[i], position[i+1])``.

    Same constructor contract as the reference: keyword-only ``position`` (n + 1
    values starting at 0, strictly increasing) and ``rate`` (n values, NaN = missing).
    """

    def __init__(self, *, position, rate):
        self._position = np.array(position, dtype=float)
        self._position.flags.writeable = False
        self._rate = np.array(rate, dtype=float)
        self._rate.flags.writeable = False
        size = len(self._position)
        if size < 2:
            raise ValueError("Must have at least two positions")
        if len(self._rate) != size - 1:
            raise ValueError("Rate array must have one less entry than the position array")
        if self._position[0] != 0:
            raise ValueError("First position must be zero")
        span = self.span
        if np.any(span <= 0):
            bad = np.where(span <= 0)[0] + 1
            raise ValueError(f"Position values not strictly increasing at indexes {bad}")
        if np.any(self._rate < 0):
            bad = np.where(self._rate < 0)[0]
            raise ValueError(f"Rate values negative at indexes {bad}")
        self._missing = np.isnan(self._rate)
        if int(np.sum(self._missing)) == len(self._rate):
            raise ValueError("All intervals are missing data")
        self._cumulative_mass = np.insert(np.nancumsum(self.mass), 0, 0)
        self._cumulative_mass.flags.writeable = False

    @staticmethod
    def uniform(sequence_length, rate) -> "RateMap":
        return RateMap(position=[0, sequence_length], rate=[rate])

    @property
    def position(self):
        return self._position

    @property
    def rate(self):
        return self._rate

    @property
    def left(self):
        return self._position[:-1]

    @property
    def right(self):
        return self._position[1:]

    @property
    def mid(self):
        return self.left + self.span / 2

    @property
    def span(self):
        return self.right - self.left

    @property
    def mass(self):
        return self._rate * self.span

    @property
    def missing(self):
        return self._missing

    @property
    def num_intervals(self) -> int:
        return len(self._rate)

    @property
    def sequence_length(self):
        return self._position[-1]

    @property
    def total_mass(self):
        return self._cumulative_mass[-1]

    @property
    def mean_rate(self):
        span = self.span[~self._missing]
        return float(np.sum(self.mass[~self._missing]) / np.sum(span))

    def get_cumulative_mass(self, x):
        x = np.array(x)
        if np.any(x < 0) or np.any(x > self.sequence_length):
            raise ValueError(f"Cannot have positions < 0 or > {self.sequence_length}")
        return np.interp(x, self._position, self._cumulative_mass)

    def get_rate(self, x):
        x = np.array(x)
        if np.any(x < 0) or np.any(x >= self.sequence_length):
            raise ValueError("Rate only defined for 0 <= x < sequence_length")
        return self._rate[np.searchsorted(self._position, x, side="right") - 1]

    def asdict(self):
        return {"position": self.position, "rate": self.rate}

    def copy(self) -> "RateMap":
        return RateMap(position=self.position.copy(), rate=self.rate.copy())

    # Mapping protocol over interval midpoints, like the reference
    def __iter__(self):
        yield from self.mid[~self._missing]

    def __len__(self):
        return int(self.num_intervals - np.sum(self._missing))

    def __getitem__(self, key):
        if isinstance(key, slice):
            raise TypeError("RateMap slicing is out of scope for msprime_b200")
        return self.get_rate(key)

    def __repr__(self):
        return f"RateMap(position={self.position!r}, rate={self.rate!r})"
