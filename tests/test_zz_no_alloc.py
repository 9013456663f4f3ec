"""GPU: the allocation-free entry point `mspb_site_rows_noalloc` (include/msprime_b200.h; no counterpart in
the reference) against the existing entry points: the `site`
column that `mspb_generate` writes IS the site row of every position-sorted record, so the new call, fed the
positions of the resident result, must reproduce that column and the number of sites exactly (integer work:
bit-exact, like every other parity test here).  Two calls in a row on different inputs, then the first input
again: the calls share one static scratch array that keeps the previous call's contents.

Collected last (the file name sorts after every other test file)."""
import numpy as np
import pytest

import cases
from msprime_b200 import _native

pytestmark = pytest.mark.gpu
FMAX = cases.FMAX


def overlay(ctx, ts, rate, seed):
    L = ts["sequence_length"]
    counts = ctx.generate(_native.Params(
        map_position=[0, L], map_rate=[rate], model=cases.native_model(cases.INF_ALLELES), discrete_genome=True,
        keep=False, start_time=-FMAX, end_time=FMAX, seed=seed))
    return counts, ctx.device_tensors()


def site_rows(torch, positions_i32):
    """The new entry point on a device tensor of sorted positions (torch owns inputs and outputs; the
    library allocates nothing)."""
    n = positions_i32.numel()
    rows = torch.full((max(n, 1),), -7, dtype=torch.int32, device=positions_i32.device)
    num_sites = torch.full((1,), -7, dtype=torch.int64, device=positions_i32.device)
    stream = torch.cuda.current_stream(positions_i32.device)
    _native.site_rows_noalloc(positions_i32.device.index, stream.cuda_stream, positions_i32.data_ptr() if n else 0, n,
                              rows.data_ptr(), num_sites.data_ptr())
    stream.synchronize()
    return rows[:n], int(num_sites.item())


def test_no_alloc_site_rows_match_generate(native_built):
    import torch

    dev = torch.device("cuda", 0)
    ts = cases.smc_ts(50, 1e5, 1)
    with _native.Context(0) as ctx:
        ctx.load_tables(ts)
        inputs = []
        # ~4e5 records over 1e5 positions (most sites hit several times, ~100 tiles), then a sparse
        # overlay that ends inside a tile
        for rate, seed in ((2e-5, 11), (3e-7, 12)):
            counts, res = overlay(ctx, ts, rate, seed)
            site = res["mutations"]["site"].clone()
            pos = res["sites"]["position"][site.long()].to(torch.int32).contiguous()
            inputs.append((pos, site, counts["num_sites"]))
            assert counts["num_mutations"] == pos.numel() > 1000
        torch.cuda.synchronize(dev)
        assert inputs[0][0].numel() > 50 * 4096 and inputs[0][2] < inputs[0][0].numel()
        # two calls in a row, then the first input again (the scratch holds the second call's state)
        for pos, site, num_sites in (inputs[0], inputs[1], inputs[0]):
            rows, n_sites = site_rows(torch, pos)
            assert n_sites == num_sites
            assert torch.equal(rows, site)
    # the size limit is an error, not a truncated result; nothing is read at that size
    with pytest.raises(_native.LibraryError) as e:
        _native.site_rows_noalloc(0, 0, inputs[0][0].data_ptr(), _native.NOALLOC_MAX_RECORDS + 1,
                                  inputs[0][1].data_ptr(), inputs[0][1].data_ptr())
    assert e.value.status == _native.ERR_NOALLOC_TOO_LARGE
    assert "MSPB_NOALLOC_MAX_RECORDS" in str(e.value)
    # no records: zero sites
    rows, n_sites = site_rows(torch, torch.empty(0, dtype=torch.int32, device=dev))
    assert n_sites == 0 and rows.numel() == 0
    # hand-made input with a known answer, first record alone in its site, last tile partial
    pos = torch.tensor(np.repeat(np.arange(3000, dtype=np.int32) * 3, 3)[1:], device=dev)
    rows, n_sites = site_rows(torch, pos)
    assert n_sites == 3000
    assert torch.equal(rows.cpu(), torch.tensor(np.repeat(np.arange(3000, dtype=np.int32), 3)[1:]))
