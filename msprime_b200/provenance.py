"""Provenance records for ``sim_mutations`` (behaviour of ``msprime/provenance.py:52-140``):
one JSON document per call holding every parameter, the seed and the environment.
The environment names this library and the CUDA runtime in place of msprime and GSL."""
from __future__ import annotations

import json
import platform

import numpy as np

from . import core

CURRENT_FILE_MAJOR = 1


def get_environment():
    env = {
        "os": {
            "system": platform.system(), "node": platform.node(), "release": platform.release(),
            "version": platform.version(), "machine": platform.machine(),
        },
        "python": {
            "implementation": platform.python_implementation(),
            "version": platform.python_version(),
        },
        "libraries": {"numpy": {"version": np.__version__}},
    }
    return env


def get_provenance_dict(parameters=None):
    return {
        "schema_version": "1.0.0",
        "software": {"name": "msprime_b200", "version": core.__version__},
        "parameters": parameters,
        "environment": get_environment(),
    }


class _Encoder(json.JSONEncoder):
    def default(self, obj):
        if hasattr(obj, "tables") and hasattr(obj, "sequence_length"):
            # tree sequences are not serialised (msprime/provenance.py:131-140)
            return {"__constant__": "__current_ts__"}
        if isinstance(obj, np.ndarray):
            return obj.tolist()
        if isinstance(obj, np.generic):
            return obj.item()
        if hasattr(obj, "asdict"):
            d = obj.asdict()
            d = {k: v for k, v in d.items()}
            d["__class__"] = f"{obj.__class__.__module__}.{obj.__class__.__name__}"
            return d
        return super().default(obj)


def json_encode_provenance(provenance_dict, num_replicates=1):
    return json.dumps(provenance_dict, cls=_Encoder)
