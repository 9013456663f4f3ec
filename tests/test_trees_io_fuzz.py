"""CPU: damaged `.trees` files.  Bytes of the reference-written fixture are overwritten at random (type
codes, lengths and offsets in the descriptors, key bytes, array bytes, truncation) and the file is
given to our loader and to the reference's real tskit:

* our loader either returns tables or raises FileFormatError, nothing else (no crash, no stray
  exception: the reader maps the file and must bound every access);
* we never accept a file tskit rejects (its column rules, tables.c:100-262 and 11588-11800, are all
  applied: required columns, column pairs, element types, row counts, offsets);
* where we reject at the schema level and tskit accepts, it is text that is not UTF-8 (tskit fails when
  the value is read, we at load time).

Files our reader rejects at the CONTAINER level ("not in kastore format": descriptors or keys damaged)
are not shown to tskit at all: the reference's kastore trusts more of the descriptor table than we do
(keys out of order are simply not found) and can crash the process on such files, which is how this
test first stopped.  Being stricter there is deliberate."""
import os
import struct

import numpy as np
import pytest

import tskit_path

tskit = tskit_path.load()
if tskit is None:
    pytest.skip("the reference's tskit package is not available here", allow_module_level=True)

from msprime_b200 import build, trees_io  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "ref_overlay.trees")
EDGE_VALUES = lambda n: [0, 1, 2**32, 2**63, 2**64 - 1, n, n + 1, n - 1]  # noqa: E731


def damaged(good, desc_end, rng):
    b = bytearray(good)
    for _ in range(int(rng.integers(1, 3))):
        region = rng.random()
        if region < 0.6:
            pos = int(rng.integers(0, desc_end))
        elif region < 0.85:
            pos = int(rng.integers(desc_end, min(len(b), desc_end + 4096)))  # keys and the first arrays
        else:
            pos = int(rng.integers(0, len(b)))
        mode = int(rng.integers(0, 3))
        if mode == 0:
            b[pos] = int(rng.integers(0, 256))
        elif mode == 1:
            b[pos] ^= 1 << int(rng.integers(0, 8))
        elif pos + 8 <= len(b):
            struct.pack_into("<Q", b, pos, EDGE_VALUES(len(b))[int(rng.integers(0, 8))])
    if rng.random() < 0.1:
        b = b[:int(rng.integers(0, len(b)))]
    return bytes(b)


@pytest.mark.parametrize("seed", range(4))
def test_damaged_files(tmp_path, seed):
    build.build_treesio()
    good = open(GOLDEN, "rb").read()
    desc_end = 64 + 64 * struct.unpack_from("<I", good, 12)[0]
    rng = np.random.default_rng(seed)
    path = str(tmp_path / "damaged.trees")
    accepted = rejected = stricter = 0
    for _ in range(600):
        with open(path, "wb") as f:
            f.write(damaged(good, desc_end, rng))
        reason = None
        try:
            d = trees_io.load_tables_dict(path)
            for tbl in ("nodes", "edges", "sites", "mutations"):
                for v in d[tbl].values():
                    if isinstance(v, np.ndarray) and v.size:
                        v[-1]  # the views are readable to their ends
        except trees_io.FileFormatError as e:
            reason = str(e)
        if reason is not None and "kastore format" in reason:
            rejected += 1
            continue  # container-level damage: see the module docstring
        try:
            tskit.TableCollection.load(path)
            theirs_ok = True
        except Exception:  # noqa: BLE001
            theirs_ok = False
        if reason is None:
            assert theirs_ok, "accepted a file tskit rejects"
            accepted += 1
        else:
            rejected += 1
            if theirs_ok:
                assert "UTF-8" in reason, reason
                stricter += 1
    assert accepted > 50 and rejected > 50 and stricter < 30
