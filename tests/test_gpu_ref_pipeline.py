"""GPU: the CUDA path against tables the REFERENCE ITSELF produced end to end (tests/golden/ref_pipeline.npz).

Every case is a tree sequence from the reference's unmodified ancestry simulator with the sites and mutations
the reference's unmodified lib/mutgen.c laid over it under ``gsl_rng_set(seed)`` -- config 1 of BASELINE.json
(sim_ancestry(1000, sequence_length=1e6, recombination_rate=1e-8, population_size=1e4, random_seed=1), then
sim_mutations(rate=1e-8, random_seed=2)) among them.  The device has no Mersenne Twister: it is handed the
generator's doubles as a replay stream (tests/test_ref_pipeline.py shows on the CPU that this IS the seeded
run) and must return the committed tables bit for bit, every column, with the same number of draws.  No oracle
runs here: the comparison is against the fixture alone.
"""
import numpy as np
import pytest

import ref_pipeline
import test_gpu_replay as replay
from msprime_b200 import _native

pytestmark = pytest.mark.gpu
CASES = ref_pipeline.load()


@pytest.mark.parametrize("name", sorted(CASES))
def test_device_reproduces_the_reference_pipeline(native_built, name):
    c = CASES[name]
    L = c["ts"]["sequence_length"]
    stream = ref_pipeline.mt19937_doubles(c["seed"], c["draws"] + 1024)
    with _native.Context(0) as ctx:
        got = replay.gpu_run(ctx, c["ts"], model=c["model"], stream=stream, map_position=[0.0, L],
                             map_rate=[c["rate"]], discrete=c["discrete"], keep=False)
    diffs = replay.describe_diff(c["want"], got)
    assert not diffs, diffs
    assert got["draws"] == c["draws"]
    assert len(got["mutations"]["site"]) == len(c["want"]["mutations"]["site"]) > 100


def test_one_shot_entry_point_on_config1(native_built):
    """The same through ``mspb_sim_mutations`` (the one-call boundary a maintainer would bind)."""
    c = CASES["config1"]
    L = c["ts"]["sequence_length"]
    stream = ref_pipeline.mt19937_doubles(c["seed"], c["draws"] + 1024)
    import cases

    params = _native.Params(map_position=[0.0, L], map_rate=[c["rate"]], model=cases.native_model(c["model"]),
                            discrete_genome=True, keep=False, start_time=-cases.FMAX, end_time=cases.FMAX,
                            replay_stream=stream)
    res, counts = _native.sim_mutations_oneshot(c["ts"], params, device=0)
    assert not replay.describe_diff(c["want"], res)
    assert counts["draws"] == c["draws"]
