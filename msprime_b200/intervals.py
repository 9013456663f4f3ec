"""``RateMap`` for the mutation path.

The in-memory surface of ``msprime/intervals.py:35-430``: keyword-only ``position`` (n + 1 break
points, first one 0, strictly increasing) and ``rate`` (n non-negative values, NaN = missing
data); read-only interval columns; point queries (``get_rate``, ``get_cumulative_mass``,
``find_index``); the Mapping protocol over the midpoints of the known intervals, with interval
slices (``rate_map[a:b]``, ``slice``); ``uniform``, ``copy``, ``asdict``, ``read_hapmap``; the text /
HTML table.  The deprecated ``RecombinationMap`` is out of scope.  All derived arrays are computed
once, in the constructor, and frozen.
"""
from __future__ import annotations

import collections.abc
import numbers

import numpy as np


def _frozen(values) -> np.ndarray:
    a = np.array(values, dtype=np.float64)
    a.setflags(write=False)
    return a


def _indexes(mask, shift=0):
    return np.flatnonzero(mask) + shift


def _rebuild(position, rate):
    return RateMap(position=position, rate=rate)


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
            "non_missing": ~gaps,
        }
        self._cumulative_mass = _frozen(np.concatenate([[0.0], np.nancumsum(r * widths)]))

    @staticmethod
    def uniform(sequence_length, rate) -> "RateMap":
        """One interval covering the whole sequence."""
        return RateMap(position=[0, sequence_length], rate=[rate])

    @staticmethod
    def read_hapmap(fileobj, sequence_length=None, *, has_header=True, position_col=None, rate_col=None,
                    map_col=None) -> "RateMap":
        """A rate map from a white-space separated HapMap-style table (``msprime/intervals.py:452-609``):
        one column of physical positions in bp (default: the second) and either a column of cumulative
        map positions in cM (``map_col``, default: the fourth) or one of rates in cM/Mb (``rate_col``), each
        row's rate holding up to the next row's position.  The stretch before the first row is missing
        data unless the map there is already above 0 (then the first rate extends back to 0); a
        ``sequence_length`` beyond the last row adds a stretch of missing data, one short of it is an error.
        ``fileobj`` goes to :func:`numpy.loadtxt` (file names ending in .gz / .bz2 are decompressed)."""
        if rate_col is not None and map_col is not None:
            raise ValueError("Cannot specify both rate_col and map_col")
        by_rate = rate_col is not None
        value_col = rate_col if by_rate else (3 if map_col is None else map_col)
        position_col = 1 if position_col is None else position_col
        if position_col == value_col:
            raise ValueError("Cannot specify the same columns for position_col and rate_col or map_col")
        value, pos = np.loadtxt(fileobj, skiprows=1 if has_header else 0, usecols=[value_col, position_col],
                                dtype=[("value", float), ("pos", int)], unpack=True)
        value, pos = np.atleast_1d(value), np.atleast_1d(pos)
        if by_rate:
            if len(value) and value[-1] != 0:
                raise ValueError("The last entry in the 'rate' column must be zero")
            # cM/Mb times Mb, accumulated, then cM -> Morgans
            morgans = np.insert(np.cumsum(value[:-1] * np.diff(pos / 1e6)), 0, 0) / 100
        else:
            morgans = value / 100
        if len(morgans) == 0:
            raise ValueError("Empty hapmap file")
        first, last = pos[0], pos[-1]
        if morgans[0] > 0 and first == 0:
            raise ValueError("The map distance at the start of the chromosome must be zero")
        known_from = first
        if first > 0:
            if morgans[0] > 0:
                known_from = 0  # the map is already above zero at the first row: its rate reaches back to 0
            pos, morgans = np.insert(pos, 0, 0), np.insert(morgans, 0, 0)
        if sequence_length is not None:
            if sequence_length < last:
                raise ValueError(f"The sequence_length cannot be less than the last physical position  ({pos[-1]})")
            if sequence_length > last:
                pos, morgans = np.append(pos, sequence_length), np.append(morgans, morgans[-1])
        rate = np.diff(morgans) / np.diff(pos)
        if known_from != 0:
            rate[0] = np.nan
        if last != pos[-1]:
            rate[-1] = np.nan
        return RateMap(position=pos, rate=rate)

    # read-only columns
    position = property(lambda self: self._cols["position"])
    rate = property(lambda self: self._cols["rate"])
    left = property(lambda self: self._cols["left"])
    right = property(lambda self: self._cols["right"])
    span = property(lambda self: self._cols["span"])
    mid = property(lambda self: self._cols["mid"])
    mass = property(lambda self: self._cols["mass"])
    missing = property(lambda self: self._cols["missing"])
    non_missing = property(lambda self: self._cols["non_missing"])
    num_intervals = property(lambda self: int(self._cols["rate"].size))
    num_missing_intervals = property(lambda self: int(self._cols["missing"].sum()))
    num_non_missing_intervals = property(lambda self: int(self._cols["non_missing"].sum()))
    sequence_length = property(lambda self: self._cols["position"][-1])
    total_mass = property(lambda self: self._cumulative_mass[-1])

    @property
    def mean_rate(self):
        """Total mass over the span of the intervals that have data."""
        return self.total_mass / np.sum(self.span[self.non_missing])

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
            raise ValueError("position out of bounds")
        return self.rate[np.searchsorted(self.position, q, side="right") - 1]

    def find_index(self, x) -> int:
        """Index ``i`` with ``left[i] <= x < right[i]``; KeyError outside ``[0, sequence_length)``."""
        if not 0 <= x < self.sequence_length:
            raise KeyError(f"Position {x} out of bounds")
        return int(np.searchsorted(self.position, x, side="right")) - 1

    def missing_intervals(self):
        """``(num_missing_intervals, 2)`` array of the ``[left, right)`` without data."""
        gaps = self.missing
        return np.column_stack([self.left[gaps], self.right[gaps]])

    def asdict(self):
        """The two arrays the C call takes (``{"position": ..., "rate": ...}``)."""
        return {name: self._cols[name] for name in ("position", "rate")}

    def copy(self) -> "RateMap":
        return RateMap(position=np.array(self.position), rate=np.array(self.rate))

    def __reduce__(self):  # pickling goes back through the constructor, so the copy is frozen too
        return _rebuild, (np.array(self.position), np.array(self.rate))

    def slice(self, left=None, right=None, *, trim=False) -> "RateMap":  # noqa: A003
        """The part of the map inside ``[left, right)``.  With ``trim`` the result is shifted to
        start at 0 and is ``right - left`` long; otherwise coordinates stay and everything outside
        the window becomes missing data."""
        L = self.sequence_length
        lo = 0 if left is None else left
        hi = L if right is None else right
        if not 0 <= lo < hi <= L:
            raise KeyError(f"Invalid slice: left={lo}, right={hi}")
        first = self.find_index(lo)
        last = int(np.searchsorted(self.position, hi, side="left")) - 1  # interval that ends the window
        breaks = np.concatenate([[lo], self.position[first + 1:last + 1], [hi]]).astype(np.float64)
        rates = np.array(self.rate[first:last + 1])
        if trim:
            return RateMap(position=breaks - lo, rate=rates)
        if lo != 0:  # flank on the left: widen a missing first interval, else add one
            if np.isnan(rates[0]):
                breaks[0] = 0
            else:
                breaks, rates = np.concatenate([[0], breaks]), np.concatenate([[np.nan], rates])
        if hi != L:
            if np.isnan(rates[-1]):
                breaks[-1] = L
            else:
                breaks, rates = np.concatenate([breaks, [L]]), np.concatenate([rates, [np.nan]])
        return RateMap(position=breaks, rate=rates)

    # Mapping protocol: keys are the midpoints of the intervals that have data
    def __iter__(self):
        return iter(self.mid[~self.missing])

    def __len__(self):
        return int((~self.missing).sum())

    def __getitem__(self, key):
        if isinstance(key, slice):
            if key.step is not None:
                raise TypeError("Only interval slicing is supported")
            return self.slice(key.start, key.stop)
        if not isinstance(key, numbers.Number):
            raise KeyError(f"Key {key} not in map")
        value = self.rate[self.find_index(key)]
        if np.isnan(value):  # consistent with __iter__: positions without data are not "in" the map
            raise KeyError(f"Position {key} is within a missing interval")
        return value

    def __repr__(self):
        return f"RateMap(position={self.position!r}, rate={self.rate!r})"

    # display: every interval up to 39 of them, else the first and last ten around an ellipsis row
    def _display_table(self):
        def rows(sel):
            return [[f"{l:.10g}", f"{r:.10g}", f"{m:.10g}", f"{w:.10g}", f"{v:.2g}"] for l, r, m, w, v in zip(
                self.left[sel], self.right[sel], self.mid[sel], self.span[sel], self.rate[sel])]

        if self.num_intervals < 40:
            body = rows(slice(None))
        else:
            body = rows(slice(0, 10)) + [["⋯"] * 5] + rows(slice(-10, None))
        return ["left", "right", "mid", "span", "rate"], body

    def __str__(self):
        titles, body = self._display_table()
        inner = [max(len(t), *(len(row[j]) for row in body)) + 2 for j, t in enumerate(titles)]
        rule = "─" * (sum(inner) + len(inner) - 1)

        def line(cells):
            aligned = [f"{c:<{w}}" if j < 2 else f"{c:>{w}}" for j, (c, w) in enumerate(zip(cells, inner))]
            return "│" + "│".join(aligned) + "│\n"

        return f"\n┌{rule}┐\n{line(titles)}├{rule}┤\n" + "".join(line(r) for r in body) + f"└{rule}┘\n"

    def _repr_html_(self):
        titles, body = self._display_table()
        head = "".join(f"<th>{t}</th>" for t in titles)
        rows = "".join("<tr>" + "".join(f"<td>{c}</td>" for c in r) + "</tr>" for r in body)
        return ('<div><table border="1" class="tskit-table"><thead><tr>' + head + "</tr></thead><tbody>" + rows
                + "</tbody></table></div>")
