"""TEST INFRASTRUCTURE.  A stand-in for ``_native.Context`` that runs on the CPU oracle, for tests of
the HOST layer where no GPU exists (the same device-free pattern as tests/test_sharding_gloo.py):
it takes the very ``_native.Params`` struct the device call would get, decodes it, and answers with
the oracle's C restatement under MT19937.  Nothing in the product imports this."""
import ctypes

import numpy as np

from msprime_b200 import _native
from oracle import oracle


def _array(address, n, dtype):
    if n == 0 or not address:
        return np.zeros(0, dtype=dtype)
    ctype = np.ctypeslib.as_ctypes_type(dtype)
    return np.ctypeslib.as_array(ctypes.cast(address, ctypes.POINTER(ctype)), shape=(n,)).copy()


def decode_model(p):
    if p.model_kind == _native.MODEL_MATRIX:
        n = int(p.num_alleles)
        off = _array(p.allele_offset, n + 1, np.uint64).astype(np.int64)
        data = _array(p.alleles, int(off[-1]), np.uint8).tobytes()
        return {"kind": "matrix", "alleles": [data[off[i]:off[i + 1]].decode("utf8") for i in range(n)],
                "root_distribution": _array(p.root_distribution, n, np.float64),
                "transition_matrix": _array(p.transition_matrix, n * n, np.float64).reshape(n, n)}
    if p.model_kind == _native.MODEL_SLIM:
        return {"kind": "slim", "type": int(p.slim_mutation_type_id), "next_id": int(p.start_allele),
                "slim_generation": int(p.slim_generation)}
    return {"kind": "infinite_alleles", "start_allele": int(p.start_allele)}


class OracleContext:
    """load_tables / generate / fetch with the meaning they have on ``_native.Context``."""

    def __init__(self, device=0):
        oracle.build(("port",))
        self.device, self.counts, self.calls = device, None, 0

    def load_tables(self, tables_dict):
        self.tables = tables_dict

    def generate(self, params):
        p = params.struct
        n = int(p.map_size)
        self.calls += 1
        try:
            r = oracle.sim_mutations(
                self.tables, which="port", map_position=_array(p.map_position, n + 1, np.float64),
                map_rate=_array(p.map_rate, n, np.float64), model=decode_model(p),
                discrete_genome=bool(p.discrete_genome), keep=bool(p.keep), start_time=p.start_time,
                end_time=p.end_time, seed=int(p.seed))
        except oracle.OracleError as e:
            _native.check(e.status)  # the library's exception for the same code
            raise
        self.result = {"sites": r["sites"], "mutations": r["mutations"]}
        self.counts = {"num_sites": len(r["sites"]["position"]), "num_mutations": len(r["mutations"]["site"]),
                       "next_allele": r["next_allele"], "offset_bytes": 8}
        return self.counts

    def fetch(self, out=None):
        return self.result
