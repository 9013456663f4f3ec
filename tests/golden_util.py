"""Loads tests/golden/mutgen_ref.npz and rebuilds each case's inputs."""
import functools
import os

import numpy as np

import cases
import golden_cases

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "mutgen_ref.npz")


@functools.lru_cache(maxsize=1)
def store():
    return np.load(PATH)


def golden(name):
    z = store()
    out = {"sites": {}, "mutations": {}}
    prefix = name + "/"
    for key in z.files:
        if key.startswith(prefix):
            parts = key[len(prefix):].split("/")
            if len(parts) == 2:
                out[parts[0]][parts[1]] = z[key]
            else:
                out[parts[0]] = int(z[key])
    return out


def case_inputs(c):
    """(tables dict incl. kept rows, stream) of a golden case."""
    ts = golden_cases.build_ts(c["ts"])
    if c["base"] is not None:
        ts = cases.with_kept(ts, golden(c["base"]))
    stream = np.random.default_rng(c["seed"]).random(c["n_stream"])
    return ts, stream


def diff_tables(want, got):
    out = []
    for tbl in ("sites", "mutations"):
        for col in want[tbl]:
            x, y = np.asarray(want[tbl][col]), np.asarray(got[tbl][col])
            if x.shape != y.shape:
                out.append(f"{tbl}.{col}: shape {x.shape} vs {y.shape}")
                continue
            if x.size == 0:
                continue
            if x.dtype.kind == "f":
                same = np.array_equal(x.view(np.uint64), np.asarray(y, dtype=np.float64).view(np.uint64))
            else:
                same = np.array_equal(x.astype(np.int64), y.astype(np.int64))
            if not same:
                out.append(f"{tbl}.{col}: values differ")
    return out
