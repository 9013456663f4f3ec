"""TEST INFRASTRUCTURE (oracle): tree sequences from the reference's OWN ancestry simulator.

``oracle/_ref/libancestry_ref.so`` (``make -C oracle ancestry``) is the unmodified ``lib/msprime.c`` (Hudson
coalescent with recombination) compiled against the GSL shim; :func:`sim_ancestry` runs it with the arguments
of ``msprime.sim_ancestry(samples, sequence_length=, recombination_rate=, population_size=, ploidy=,
discrete_genome=, random_seed=)`` (msprime/ancestry.py:1014-1340; the seed goes to ``gsl_rng_set`` as
``RandomGenerator`` does, msprime/_msprimemodule.c) and returns the node and edge tables in the layout of
``msprime_b200.synth.native_treeseq``.  Only tests and the generating script of the golden fixtures
(``tests/make_ref_ancestry.py``) use it; the product path never does.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ANCESTRY_SO = os.path.join(HERE, "_ref", "libancestry_ref.so")
_lib = None


def have():
    return os.path.exists(ANCESTRY_SO)


def _load():
    global _lib
    if _lib is None:
        lib = C.CDLL(ANCESTRY_SO)
        lib.mspb_ref_ancestry_run.restype = C.c_int
        lib.mspb_ref_ancestry_run.argtypes = [
            C.c_uint32, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, C.c_ulong,
            C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_void_p)]
        lib.mspb_ref_ancestry_copy.restype = None
        lib.mspb_ref_ancestry_copy.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        lib.mspb_ref_ancestry_free.restype = None
        lib.mspb_ref_ancestry_free.argtypes = [C.c_void_p]
        _lib = lib
    return _lib


def cached_input_path(w):
    """Where scripts/make_reference_input.py keeps the simulator's output for a bench workload (same trees for every
    workload with the same ancestry arguments)."""
    key = "n%d_L%g_r%g_ne%g_seed%d" % (w["samples"], w["L"], w["recomb"], w["ne"], w["ts_seed"])
    return os.path.join(HERE, "_ref", "inputs", key + ".npz")


def load_cached(w):
    """The cached tree sequence of a bench workload, or None."""
    path = cached_input_path(w)
    if not os.path.exists(path):
        return None
    from msprime_b200 import synth

    z = np.load(path)
    left, right = z["edge_left"], z["edge_right"]
    return {
        "sequence_length": float(z["sequence_length"]), "time_units": "generations",
        "nodes": {"time": z["node_time"], "flags": z["node_flags"]},
        "edges": {"left": left, "right": right, "parent": z["edge_parent"], "child": z["edge_child"]},
        "sites": synth.empty_sites(), "mutations": synth.empty_mutations(),
        "num_trees": int(len(np.unique(np.concatenate([left, right]))) - 1),
        "generator_seconds": float(z["seconds"]),
    }


def sim_ancestry(samples, *, sequence_length, recombination_rate, population_size, random_seed, ploidy=2,
                 discrete_genome=True):
    lib = _load()
    nn, ne, handle = C.c_uint64(), C.c_uint64(), C.c_void_p()
    ret = lib.mspb_ref_ancestry_run(int(samples) * int(ploidy), float(sequence_length), float(recombination_rate),
                                    float(population_size), int(ploidy), int(bool(discrete_genome)),
                                    int(random_seed), C.byref(nn), C.byref(ne), C.byref(handle))
    if ret != 0:
        raise RuntimeError(f"the reference's ancestry simulator returned {ret}")
    try:
        N, E = nn.value, ne.value
        flags, time = np.empty(N, np.uint32), np.empty(N, np.float64)
        left, right = np.empty(E, np.float64), np.empty(E, np.float64)
        parent, child = np.empty(E, np.int32), np.empty(E, np.int32)
        lib.mspb_ref_ancestry_copy(handle, *(a.ctypes.data for a in (flags, time, left, right, parent, child)))
    finally:
        lib.mspb_ref_ancestry_free(handle)
    from msprime_b200 import synth

    return {
        "sequence_length": float(sequence_length), "time_units": "generations",
        "nodes": {"time": time, "flags": flags},
        "edges": {"left": left, "right": right, "parent": parent, "child": child},
        "sites": synth.empty_sites(), "mutations": synth.empty_mutations(),
        "num_trees": int(len(np.unique(np.concatenate([left, right]))) - 1),
    }
