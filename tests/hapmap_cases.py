"""HapMap-style inputs for RateMap.read_hapmap: the shapes the reference's suite uses
(tests/test_intervals.py:662-835) and generated tables.  Shared by tests/golden/make_hapmap_golden.py
(run where the reference is) and tests/test_ratemap_golden.py."""
import numpy as np


def generated(seed, rows, by_rate, start_at_zero, map_starts_positive=False):
    rng = np.random.default_rng(seed)
    pos = np.sort(rng.choice(np.arange(1, 5_000_000), size=rows, replace=False))
    if start_at_zero:
        pos[0] = 0
    rate = rng.gamma(0.5, 2.0, size=rows).round(4)  # cM / Mb
    rate[rng.random(rows) < 0.2] = 0
    rate[-1] = 0
    cm = np.concatenate([[0], np.cumsum(rate[:-1] * np.diff(pos) / 1e6)])
    if map_starts_positive:
        cm = cm + 0.25
    lines = ["Chromosome\tPosition(bp)\tRate(cM/Mb)\tMap(cM)"]
    lines += [f"chr10\t{p}\t{r:.4f}\t{c:.9f}" for p, r, c in zip(pos, rate, cm)]
    return "\n".join(lines) + "\n", dict(rate_col=2) if by_rate else {}


SIMPLE = "HEADER\nchr1 1 x 0\nchr1 2 x 0.000001 x\nchr1 3 x 0.000006 x x"
FROM_ZERO = "HEADER\nchr1 0 x 0\nchr1 1 x 0.000001 x\nchr1 2 x 0.000006 x x x"

CASES = {
    "simple": (SIMPLE, {}),
    "columns_swapped": ("HEADER\n0 0\n0.000001 1 x\n0.000006 2 x x", dict(position_col=1, map_col=0)),
    "rate_column": ("HEADER\nchr1 1 5 x\nchr1 2 0 x x x", dict(rate_col=2)),
    "rate_column_from_zero": ("HEADER\nchr1 0 1\nchr1 1 5.5\nchr1 2 0\n", dict(rate_col=2)),
    "map_above_zero_at_first_row": ("HEADER\nchr1 1 x 0.000001\nchr1 2 x 0.000001 x\nchr1 3 x 0.000006 x x x", {}),
    "same_length": (FROM_ZERO, dict(sequence_length=2)),
    "longer": (FROM_ZERO, dict(sequence_length=10)),
    "no_header": (FROM_ZERO.split("\n", 1)[1], dict(has_header=False)),
    "both_columns": (FROM_ZERO, dict(rate_col=2, map_col=3)),
    "same_column": ("HEADER\n0 0\n0.000001 1 x\n0.000006 2 x x", dict(map_col=1)),
    "last_rate_not_zero": ("HEADER\nchr1 0 5 x\nchr1 2 1 x x x", dict(rate_col=2)),
    "map_above_zero_at_position_zero": ("HEADER\nchr1 0 x 0.0000005\nchr1 1 x 0.000001 x\nchr1 2 x 0.000006 x x x", {}),
    "too_short": (FROM_ZERO, dict(sequence_length=1.999)),
    "header_read_as_data": ("chr pos rate cM\n" + FROM_ZERO.split("\n", 1)[1], dict(has_header=False)),
}
for _k, (_seed, _rows, _by_rate, _zero, _positive) in enumerate([
        (1, 40, False, False, False), (2, 40, True, False, False), (3, 25, False, True, False),
        (4, 25, True, True, False), (5, 60, False, False, True), (6, 300, False, False, False)]):
    _text, _kw = generated(_seed, _rows, _by_rate, _zero, _positive)
    CASES[f"generated_{_k}"] = (_text, _kw)
    CASES[f"generated_{_k}_padded"] = (_text, dict(_kw, sequence_length=6_000_000))
