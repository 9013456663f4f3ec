"""Imports the REFERENCE's pure-Python model classes (msprime/mutations.py) in this container,
where the compiled extension `_msprime` cannot be built (no GSL): the extension, tskit and the
sibling modules are replaced by inert stand-ins, so only the numpy code that builds alleles,
root distributions and transition matrices runs.  Used by the generators of
tests/golden/model_matrices.npz and msprime_b200/data/amino_acid_models.npz.  Needs
/root/reference; nothing at run time imports this."""
import importlib.util
import sys
import types

REF = "/root/reference/msprime/mutations.py"


class _Holder:
    """Stand-in for the C model types: keeps what the Python subclass hands to __init__."""

    def __init__(self, *args, **kwargs):
        names = ("alleles", "root_distribution", "transition_matrix")
        for k, v in zip(names, args):
            setattr(self, k, v)
        for k, v in kwargs.items():
            setattr(self, k, v)


def load():
    pkg = types.ModuleType("msprime")
    pkg.__path__ = []
    ext = types.ModuleType("msprime._msprime")
    for name in ("MatrixMutationModel", "SLiMMutationModel", "InfiniteAllelesMutationModel", "BaseMutationModel"):
        setattr(ext, name, type(name, (_Holder,), {}))
    ext.sim_mutations = None
    pkg._msprime = ext
    stubs = {"msprime": pkg, "msprime._msprime": ext, "_msprime": ext, "tskit": types.ModuleType("tskit")}
    for sub in ("core", "intervals", "provenance"):
        m = types.ModuleType(f"msprime.{sub}")
        setattr(pkg, sub, m)
        stubs[f"msprime.{sub}"] = m
    saved = {k: sys.modules.get(k) for k in stubs}
    sys.modules.update(stubs)
    try:
        spec = importlib.util.spec_from_file_location("msprime.mutations", REF)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod
