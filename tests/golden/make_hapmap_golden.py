"""Writes tests/golden/hapmap_ref.json: what the REFERENCE's RateMap.read_hapmap (msprime/intervals.py,
imported from /root/reference by scripts/ref_intervals.py) makes of the inputs in tests/hapmap_cases.py.
Run here (the reference is not on the GPU box); the JSON is committed."""
import io
import json
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", "..", "scripts"))
import hapmap_cases  # noqa: E402
import ref_intervals  # noqa: E402


def main():
    RateMap = ref_intervals.load().RateMap
    out = {}
    for name, (text, kwargs) in hapmap_cases.CASES.items():
        try:
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                m = RateMap.read_hapmap(io.StringIO(text), **kwargs)
            out[name] = {"position": np.asarray(m.position).tolist(), "rate": np.asarray(m.rate).tolist()}
        except ValueError as e:
            out[name] = {"error": "ValueError", "message": str(e)}
    with open(os.path.join(HERE, "hapmap_ref.json"), "w") as f:
        json.dump(out, f)
    print({k: ("error" if "error" in v else len(v["rate"])) for k, v in out.items()})


if __name__ == "__main__":
    main()
