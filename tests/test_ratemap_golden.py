"""RateMap (msprime_b200/intervals.py) against what the reference's own class answers
(tests/golden/ratemap_ref.json, made by tests/golden/make_ratemap_golden.py from
/root/reference/msprime/intervals.py), plus the behaviours the reference's suite asserts
(tests/test_intervals.py:38-535).  Floating-point results are compared exactly: the rate map's
arithmetic feeds the overlay's position draws."""
import decimal
import fractions
import json
import os
import pickle
import xml.etree.ElementTree

import numpy as np
import pytest

import msprime_b200 as msp
import ratemap_cases as rc

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "ratemap_ref.json")) as f:
    REF = json.load(f)


def same(a, b):
    return np.array_equal(np.asarray(a, dtype=float), np.asarray(b, dtype=float), equal_nan=True)


def attempt(fn):
    try:
        return fn()
    except (KeyError, ValueError, TypeError) as e:
        return {"error": type(e).__name__}


@pytest.mark.parametrize("name", sorted(rc.MAPS))
def test_columns_and_queries(name):
    m, ref = msp.RateMap(**rc.MAPS[name]), REF[name]
    for col in ("position", "rate", "left", "right", "mid", "span", "mass"):
        assert same(getattr(m, col), ref[col]), col
    assert [bool(x) for x in m.missing] == ref["missing"]
    assert [bool(x) for x in m.non_missing] == [not x for x in ref["missing"]]
    assert [m.num_intervals, m.num_missing_intervals, m.num_non_missing_intervals] == ref["counts"]
    assert float(m.total_mass) == ref["total_mass"]
    assert float(m.mean_rate) == ref["mean_rate"]
    assert float(m.sequence_length) == ref["sequence_length"]
    pts = rc.probes(rc.MAPS[name])
    assert same(pts, ref["probes"])
    assert same(m.get_rate(pts), ref["get_rate"])
    assert same(m.get_cumulative_mass(pts), ref["cumulative_mass"])
    assert [m.find_index(x) for x in pts] == ref["find_index"]
    got = [attempt(lambda x=x: float(m[x])) for x in pts]
    assert got == ref["getitem"]
    assert same(list(m.keys()), ref["keys"]) and same(list(m.values()), ref["values"])
    assert len(m) == ref["len"]
    assert same(m.missing_intervals(), np.reshape(ref["missing_intervals"], (-1, 2)))
    assert m.missing_intervals().shape == (m.num_missing_intervals, 2)
    assert str(m) == ref["str"]
    assert repr(m) == ref["repr"]


@pytest.mark.parametrize("name", sorted(rc.MAPS))
def test_slices(name):
    m, ref = msp.RateMap(**rc.MAPS[name]), iter(REF[name]["slices"])
    for left, right in rc.windows(rc.MAPS[name]):
        for trim in (False, True):
            want = next(ref)
            got = attempt(lambda: m.slice(left, right, trim=trim))
            if "error" in want:
                assert got == want, (left, right, trim)
            else:
                assert same(got.position, want["position"]) and same(got.rate, want["rate"]), (left, right, trim)


def test_membership():
    m = msp.RateMap(position=[0, 1, 2, 3], rate=[0.1, 0.2, 0.3])
    assert [x in m for x in rc.MEMBERSHIP] == REF["_membership"]
    assert [slice(a, b) in m for a, b in rc.MEMBERSHIP_SLICES] == REF["_membership_slices"]
    for other in ({}, [], Exception):
        assert other not in m
    gap = msp.RateMap(position=[0, 1, 2], rate=[np.nan, 0.2])
    assert 0 not in gap and 0.999 not in gap and 1 in gap
    with pytest.raises(KeyError, match="within a missing interval"):
        gap[0.999]


@pytest.mark.parametrize("position, rate", [([], []), ([0], []), ([0], [0]), ([1, 2], [0]), ([0, -1], [0]),
                                            ([0, 1], [-1])])
def test_constructor_rejects(position, rate):
    with pytest.raises(ValueError):
        msp.RateMap(position=position, rate=rate)


def test_constructor_messages():
    with pytest.raises(ValueError, match=r"at indexes \[2 4\]"):
        msp.RateMap(position=[0, 1, 1, 2, 2, 3], rate=[0, 0, 0, 0, 0])
    with pytest.raises(ValueError, match="one less entry"):
        msp.RateMap(position=np.array([0, 1, 2]), rate=np.array([0, 1, 2]))
    with pytest.raises(ValueError, match="First position"):
        msp.RateMap(position=np.array([1, 2, 3]), rate=np.array([1, 1]))
    with pytest.raises(ValueError, match="negative.*1"):
        msp.RateMap(position=np.array([0, 1, 2]), rate=np.array([1, -1]))
    with pytest.raises(ValueError, match="negative.*1"):
        msp.RateMap(position=np.array([0, 1, 2]), rate=np.array([np.nan, -1]))
    for rate in ([np.nan], [np.nan, np.nan]):
        with pytest.raises(ValueError, match="missing data"):
            msp.RateMap(position=np.linspace(0, 10, len(rate) + 1), rate=rate)


def test_everything_is_read_only():
    positions = np.array([0, 0.25, 0.5, 0.75, 1])
    rates = np.array([0.125, 0.25, 0.5, 0.75])
    m = msp.RateMap(position=positions, rate=rates)
    for name in ("rate", "position", "left", "right", "mid", "span", "mass"):
        with pytest.raises(AttributeError):
            setattr(m, name, 1234)
        with pytest.raises(ValueError):
            getattr(m, name)[0] = 1
    positions[1] = 0.3  # the map took copies
    assert m.position[1] == 0.25
    with pytest.raises(TypeError, match="item assignment"):
        m[0] = 1
    with pytest.raises(TypeError, match="item deletion"):
        del m[0]


def test_find_index_and_point_query_bounds():
    m = msp.RateMap(position=[0, 10], rate=[0.1])
    for bad in (-1, -0.0001, 10, 10.0001, 1e9):
        with pytest.raises(KeyError, match="out of bounds"):
            m.find_index(bad)
    assert m.find_index(0) == m.find_index(0.0) == m.find_index(np.zeros(1)[0]) == 0
    m = msp.RateMap(position=np.array([0, 1, 2]), rate=np.array([1, 4]))
    with pytest.raises(ValueError, match="out of bounds"):
        m.get_rate([1, -0.1])
    with pytest.raises(ValueError, match="out of bounds"):
        m.get_rate([2])
    for bad in ([1, -0.1], [1, 2.1]):
        with pytest.raises(ValueError, match="positions"):
            m.get_cumulative_mass(bad)
    assert m.get_cumulative_mass([2]) == m.total_mass


def test_mapping_protocol():
    m = msp.RateMap(position=[0, 1, 2, 3], rate=[0.1, 0.2, 0.3])
    assert list(m.items()) == [(0.5, 0.1), (1.5, 0.2), (2.5, 0.3)]
    assert list(m.keys()) == [0.5, 1.5, 2.5] and list(m.values()) == [0.1, 0.2, 0.3]
    two = msp.RateMap(position=[0, 1, 2], rate=[0.1, 0.2])
    assert two == two and two == msp.RateMap(position=[0, 1, 2], rate=[0.1, 0.2])
    assert two != msp.RateMap(position=[0, 1, 3], rate=[0.1, 0.2])
    assert msp.RateMap(position=[0, 1], rate=[0.1]) != msp.RateMap(position=[0, 1], rate=[0.2])
    assert msp.RateMap(position=[0, 1], rate=[0.1]) != msp.RateMap(position=[0, 10], rate=[0.1])
    for key, want in ((0, 0.1), (0.5, 0.1), (1, 0.2), (1.999, 0.2), (np.float32(1), 0.2), (np.int32(1), 0.2),
                      (fractions.Fraction(1, 3), 0.1), (decimal.Decimal(1), 0.2)):
        assert two[key] == want
    assert two[:] == two.copy() and two[:] is not two
    assert two[1:] == two.slice(left=1) and two[:1.5] == two.slice(right=1.5)
    assert two[0.5:1.5] == two.slice(left=0.5, right=1.5)
    with pytest.raises(TypeError, match="interval slicing"):
        two[0:3:1]


@pytest.mark.parametrize("rate", [[0.1, 0.2, 0.3], [0.1, np.nan, 0.3]])
def test_pickle_round_trip(rate):
    m = msp.RateMap(position=[0, 1, 2, 3], rate=rate)
    again = pickle.loads(pickle.dumps(m))
    assert again == m and same(again.rate, m.rate) and same(again.position, m.position)
    with pytest.raises(ValueError):
        again.rate[0] = 1


def test_tables_for_display():
    m = msp.RateMap(position=range(101), rate=[0.1] * 100)
    titles, rows = m._display_table()
    assert len(titles) == 5 and len(rows) == 21 and int(rows[0][0]) == 0 and int(rows[-1][0]) == 99
    m = msp.RateMap(position=range(11), rate=[0.1] * 10)
    titles, rows = m._display_table()
    assert len(rows) == 10 and int(rows[-1][0]) == 9
    root = xml.etree.ElementTree.fromstring(msp.RateMap(position=[0, 10], rate=[0.1])._repr_html_())
    assert root.tag == "div" and len(list(root.find("table").find("tbody"))) == 1


# ---- RateMap.read_hapmap against the reference's own reader (tests/golden/make_hapmap_golden.py)
import gzip  # noqa: E402
import io  # noqa: E402
import warnings  # noqa: E402

import hapmap_cases  # noqa: E402

with open(os.path.join(HERE, "golden", "hapmap_ref.json")) as f:
    HAPMAP_REF = json.load(f)


@pytest.mark.parametrize("name", sorted(hapmap_cases.CASES))
def test_read_hapmap(name):
    text, kwargs = hapmap_cases.CASES[name]
    want = HAPMAP_REF[name]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # numpy warns about the empty / ragged inputs of the error cases
        if "error" in want:
            with pytest.raises(ValueError) as e:
                msp.RateMap.read_hapmap(io.StringIO(text), **kwargs)
            if want["message"].split()[0] in ("Cannot", "The", "Empty"):  # our own messages, not numpy's
                assert str(e.value) == want["message"]
            return
        m = msp.RateMap.read_hapmap(io.StringIO(text), **kwargs)
    assert same(m.position, want["position"]) and same(m.rate, want["rate"])


def test_read_hapmap_from_files(tmp_path):
    text, _ = hapmap_cases.CASES["simple"]
    plain = tmp_path / "hapfile.txt"
    plain.write_text(text)
    m = msp.RateMap.read_hapmap(plain)
    assert same(m.position, [0, 1, 2, 3]) and same(m.rate, HAPMAP_REF["simple"]["rate"])
    packed = str(tmp_path / "hapmap.txt.gz")
    with gzip.GzipFile(packed, "wb") as gz:
        gz.write(b"HEADER\nchr1 0 1\nchr1 1 5.5\nchr1 2 0\n")
    m = msp.RateMap.read_hapmap(packed, rate_col=2)
    assert same(m.position, [0, 1, 2]) and same(m.rate, [1e-8, 5.5e-8])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with pytest.raises(ValueError, match="Empty"):
            msp.RateMap.read_hapmap(io.StringIO("HEADER"))
    # a single row (the reference's reader does not get this far: len() of a 0-d array)
    one = msp.RateMap.read_hapmap(io.StringIO("HEADER\nchr1 5 x 0.000001\n"), sequence_length=9)
    assert same(one.position, [0, 5, 9]) and same(one.rate, [1e-8 / 5, np.nan])
    with pytest.raises(ValueError, match="missing data"):
        msp.RateMap.read_hapmap(io.StringIO("HEADER\nchr1 5 x 0\n"), sequence_length=9)
