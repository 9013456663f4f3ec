"""Inputs shared by tests/golden/make_ratemap_golden.py (run where the reference is) and
tests/test_ratemap_golden.py: rate maps, query points, slice windows and membership probes."""
import numpy as np

nan = float("nan")

MAPS = {
    "one": dict(position=[0, 10], rate=[0.1]),
    "tiny_rate": dict(position=[0, 10], rate=[1e-6]),
    "zero": dict(position=[0, 1], rate=[0]),
    "two": dict(position=[0, 1, 2], rate=[0.1, 0.2]),
    "zero_first": dict(position=[0, 1, 2], rate=[0, 0.2]),
    "three": dict(position=[0, 10, 20, 30], rate=[0.1, 0.2, 0.3]),
    "ramp": dict(position=list(range(100)), rate=list(range(99))),
    "fractional": dict(position=[0, 0.25, 0.5, 0.75, 1], rate=[0.125, 0.25, 0.5, 0.75]),
    "gap_first": dict(position=[0, 1, 2], rate=[nan, 0]),
    "gap_last": dict(position=[0, 1, 2], rate=[0, nan]),
    "gap_middle": dict(position=[0, 10, 20, 30], rate=[0.1, nan, 0.3]),
    "gap_ends": dict(position=[0, 0.5, 1, 2, 4.5], rate=[nan, 0, 1, nan]),
    "genome": dict(position=[0, 1000.5, 2500, 2500.25, 1e5], rate=[1e-8, 0, 2.5e-7, 1e-9]),
}


def probes(kwargs):
    """Points to query: every break point, every midpoint, just inside both ends."""
    pos = np.asarray(kwargs["position"], dtype=float)
    mids = (pos[:-1] + pos[1:]) / 2
    pts = np.concatenate([pos[:-1], mids, [np.nextafter(pos[-1], 0)]])
    return np.unique(pts)


def windows(kwargs):
    """(left, right) pairs for slice(): None ends, break points, interior points, and bad ones."""
    pos = np.asarray(kwargs["position"], dtype=float)
    L = pos[-1]
    inner = sorted({float(x) for x in np.concatenate([pos, (pos[:-1] + pos[1:]) / 2, [L / 3, 2 * L / 3]])})
    out = [(None, None)]
    for a in inner[:6]:
        out += [(a, None), (None, a)]
    for i, a in enumerate(inner[:7]):
        for b in inner[i::3][:4]:
            out.append((a, b))
    out += [(-1, L), (0, L + 1), (L, L + 1), (L / 2, L / 2), (L / 2, L / 4)]
    return out


MEMBERSHIP = [0, 0.5, 1, 2.9999, -1, -0.0001, 3, 3.1, 1e9, None, "sdf", "123"]
MEMBERSHIP_SLICES = [(0, 0.5), (0, 1), (0, 2), (2, 3), (0, 3), (-0.001, 1), (0, 3.0001), (2.9999, 3.0001), (3, 4),
                     (-2, -1)]
