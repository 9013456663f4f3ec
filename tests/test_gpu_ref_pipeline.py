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


@pytest.mark.parametrize("case", range(24))
def test_device_on_random_trees_of_the_reference_simulator(native_built, oracle_mod, case):
    """The replay sweep of tests/test_oracle_fuzz.py on the device: random tree sequences from the reference's own
    ancestry simulator (made on the spot by oracle/_ref/libancestry_ref.so), random models and rate maps, two layers;
    the CUDA path against the reference's algorithm, every column, or the same error code."""
    import cases
    import test_oracle_fuzz as fuzz
    from oracle import ref_ancestry

    if not ref_ancestry.have():
        pytest.skip("oracle/_ref/libancestry_ref.so not built")
    rng = np.random.default_rng(9500 + case)
    L = float(rng.choice([50, 1e3, 2e4]))
    ts = ref_ancestry.sim_ancestry(
        int(rng.integers(1, 25)), sequence_length=L, recombination_rate=float(rng.choice([0.0, 1e-4, 1e-3])) / 4,
        population_size=float(rng.choice([10, 100, 1000])), random_seed=int(rng.integers(1, 2**31)),
        ploidy=int(rng.integers(1, 3)), discrete_genome=bool(rng.random() < 0.7))
    branch = float(np.sum((ts["edges"]["right"] - ts["edges"]["left"])
                          * (ts["nodes"]["time"][ts["edges"]["parent"]] - ts["nodes"]["time"][ts["edges"]["child"]])))
    target = float(rng.choice([5, 50, 400, 2000]))
    model = fuzz.random_model(rng)
    with _native.Context(0) as ctx:
        for layer in range(2):
            position, rate = fuzz.random_map(rng, L, 2 * target / (branch or 1.0))  # (one haploid sample: no edges)
            discrete = bool(rng.random() < 0.7)
            stream = rng.random(int(40 * target) + 10_000)
            want = fuzz.run(oracle_mod, "port", ts, dict(map_position=position, map_rate=rate, model=model,
                                                         discrete_genome=discrete, keep=bool(layer == 1), stream=stream))
            try:
                got = replay.gpu_run(ctx, ts, model=model, stream=stream, map_position=position, map_rate=rate,
                                     discrete=discrete, keep=bool(layer == 1))
            except (_native.LibraryError, _native.InputError) as e:
                got = e.status
            if isinstance(want, int) or isinstance(got, int):
                assert want == got, (want, got)
                return
            assert not replay.describe_diff(want, got)
            assert got["draws"] == want["draws"]
            ts = cases.with_kept(ts, want)
