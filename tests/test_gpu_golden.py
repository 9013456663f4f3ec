"""GPU: the CUDA path through the C ABI against the committed golden vectors of the
unmodified reference mutgen (no oracle involved at run time).  Bit-exact."""
import numpy as np
import pytest

import cases
import golden_cases
import golden_util
from msprime_b200 import _native

pytestmark = pytest.mark.gpu
ALL = golden_cases.all_cases()


@pytest.fixture(scope="module")
def ctx(native_built):
    c = _native.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("offsets32", [False, True], ids=["off64", "off32"])
@pytest.mark.parametrize("c", ALL, ids=[c["name"] for c in ALL])
def test_cuda_matches_reference_golden(ctx, c, offsets32):
    ts, stream = golden_util.case_inputs(c)
    want = golden_util.golden(c["name"])
    ctx.load_tables(ts)
    counts = ctx.generate(_native.Params(
        map_position=c["map_position"], map_rate=c["map_rate"], model=cases.native_model(cases.MODELS[c["model"]]),
        discrete_genome=c["discrete"], keep=c["keep"], start_time=c["start_time"], end_time=c["end_time"],
        replay_stream=stream, offsets32=offsets32))
    got = ctx.fetch()
    assert got["mutations"]["derived_state_offset"].dtype == (np.uint32 if offsets32 else np.uint64)
    assert not golden_util.diff_tables(want, got)
    if cases.MODELS[c["model"]]["kind"] == "matrix":
        assert counts["draws"] == want["draws"]
    else:
        assert counts["next_allele"] == want["next_allele"]
