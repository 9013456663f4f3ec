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
    try:  # the native library and the CUDA runtime it was built against, where the reference names GSL
        from . import _native

        lib = _native.lib()
        cuda = int(lib.mspb_cuda_runtime_version())
        env["libraries"]["msprime_b200_native"] = {"version": lib.mspb_version().decode()}
        env["libraries"]["cuda_runtime"] = {"version": f"{cuda // 1000}.{cuda % 1000 // 10}"}
    except Exception:  # noqa: BLE001 - a provenance record must not fail for want of a version string
        pass
    return env


def get_provenance_dict(parameters=None):
    return {
        "schema_version": "1.0.0",
        "software": {"name": "msprime_b200", "version": core.__version__},
        "parameters": parameters,
        "environment": get_environment(),
    }


CURRENT_TREE_SEQUENCE = object()  # what the placeholder for "the input tree sequence" decodes to


class _Encoder(json.JSONEncoder):
    """The reference's encoding (``msprime/provenance.py:97-176``): the input tree sequence is a
    placeholder, arrays and numpy scalars carry their dtype, objects with ``asdict`` carry their class."""

    def default(self, obj):
        if hasattr(obj, "tables") and hasattr(obj, "sequence_length"):
            # tree sequences are not serialised: the provenance row before this one describes them
            return {"__constant__": "__current_ts__"}
        if isinstance(obj, np.ndarray):
            return {"__ndarray__": obj.tolist(), "dtype": obj.dtype.str}
        if isinstance(obj, np.generic):
            return {"__npgeneric__": str(obj), "dtype": obj.dtype.str}
        if hasattr(obj, "asdict"):
            d = dict(obj.asdict())
            d["__class__"] = f"{obj.__class__.__module__}.{obj.__class__.__name__}"
            return d
        raise TypeError(f"Object of type {obj.__class__.__name__} is not JSON serialisable. Please provide an "
                        "`asdict` method that returns the object's constructor arguments.")


def json_encode_provenance(provenance_dict, num_replicates=1):
    return json.dumps(provenance_dict, cls=_Encoder)


def json_decode_provenance(record: str):
    """Inverse of :func:`json_encode_provenance`: arrays, numpy scalars and model / rate map objects
    come back as such (classes are looked up in the module named in the record)."""
    def resolve(qualified):
        # Only this package's own model and rate-map classes are ever built from a record: a row in an
        # untrusted .trees file must not be able to name an arbitrary callable.
        from . import intervals, mutations

        module, _, name = qualified.rpartition(".")
        allowed = {"msprime_b200.mutations": (mutations, mutations.MutationModel),
                   "msprime_b200.intervals": (intervals, intervals.RateMap)}
        if module not in allowed:
            raise ValueError(f"provenance record names a class outside this package: {qualified}")
        mod, base = allowed[module]
        cls = getattr(mod, name, None)
        if not (isinstance(cls, type) and issubclass(cls, base)):
            raise ValueError(f"provenance record names an unknown class: {qualified}")
        return cls

    def hook(obj):
        if "__constant__" in obj:
            return {"__current_ts__": CURRENT_TREE_SEQUENCE}[obj["__constant__"]]
        if "__ndarray__" in obj:
            return np.asarray(obj["__ndarray__"], dtype=obj["dtype"])
        if "__npgeneric__" in obj:
            return np.array([obj["__npgeneric__"]]).astype(obj["dtype"])[0]
        if "__class__" in obj:
            return resolve(obj.pop("__class__"))(**obj)
        return obj

    return json.JSONDecoder(object_hook=hook).decode(record)


def parse_provenance(provenance, current_ts):
    """``(command, parameters)`` of a provenance row written by this package, ready to run again:
    ``getattr(msprime_b200, command)(**parameters)`` (``msprime/provenance.py:202-220``).  ``provenance``
    is a row with a ``record`` attribute (tskit) or the JSON text itself."""
    record = getattr(provenance, "record", provenance)
    if isinstance(record, dict):
        record = record["record"]
    name = json.loads(record).get("software", {}).get("name")  # checked before any object is rebuilt
    if name != "msprime_b200":
        raise ValueError(f"Only msprime_b200 provenances can be parsed, found {name}")
    decoded = json_decode_provenance(record)
    parameters = decoded["parameters"]
    command = parameters.pop("command")
    return command, {k: (current_ts if v is CURRENT_TREE_SEQUENCE else v) for k, v in parameters.items()}
