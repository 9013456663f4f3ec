"""Writes tests/golden/ratemap_ref.json: what the REFERENCE's RateMap (msprime/intervals.py, imported
from /root/reference by scripts/ref_intervals.py) answers for the cases of tests/ratemap_cases.py.
Run here (the reference is not on the GPU box); the JSON is committed."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", "..", "scripts"))
import ratemap_cases as rc  # noqa: E402
import ref_intervals  # noqa: E402


def listed(a):
    return np.asarray(a, dtype=float).tolist()


def attempt(fn):
    try:
        return fn()
    except (KeyError, ValueError, TypeError) as e:
        return {"error": type(e).__name__}


def main():
    RateMap = ref_intervals.load().RateMap
    out = {}
    for name, kwargs in rc.MAPS.items():
        m = RateMap(**kwargs)
        pts = rc.probes(kwargs)
        rec = {col: listed(getattr(m, col)) for col in ("position", "rate", "left", "right", "mid", "span", "mass")}
        rec["missing"] = [bool(x) for x in m.missing]
        rec["counts"] = [int(m.num_intervals), int(m.num_missing_intervals), int(m.num_non_missing_intervals)]
        rec["total_mass"] = float(m.total_mass)
        rec["mean_rate"] = float(m.mean_rate)
        rec["sequence_length"] = float(m.sequence_length)
        rec["probes"] = listed(pts)
        rec["get_rate"] = listed(m.get_rate(pts))
        rec["cumulative_mass"] = listed(m.get_cumulative_mass(pts))
        rec["find_index"] = [int(m.find_index(x)) for x in pts]
        rec["getitem"] = [attempt(lambda x=x: float(m[x])) for x in pts]
        rec["keys"] = listed(list(m.keys()))
        rec["values"] = listed(list(m.values()))
        rec["len"] = int(len(m))
        rec["missing_intervals"] = listed(m.missing_intervals())
        rec["str"] = str(m)
        rec["repr"] = repr(m)
        rec["slices"] = []
        for left, right in rc.windows(kwargs):
            for trim in (False, True):
                got = attempt(lambda: m.slice(left, right, trim=trim))
                if not isinstance(got, dict):
                    got = {"position": listed(got.position), "rate": listed(got.rate)}
                rec["slices"].append(got)
        out[name] = rec
    m = RateMap(position=[0, 1, 2, 3], rate=[0.1, 0.2, 0.3])
    out["_membership"] = [bool(x in m) for x in rc.MEMBERSHIP]
    out["_membership_slices"] = [bool(slice(a, b) in m) for a, b in rc.MEMBERSHIP_SLICES]
    with open(os.path.join(HERE, "ratemap_ref.json"), "w") as f:
        json.dump(out, f)
    print("wrote", len(out), "records,", sum(len(v["slices"]) for k, v in out.items() if k[0] != "_"), "slices")


if __name__ == "__main__":
    main()
