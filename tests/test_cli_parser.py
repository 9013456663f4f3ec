"""The command line's argument surface, without a device: what the reference's suite asserts of
`msp mutations` parsing (tests/test_cli.py:1261-1299) plus the failures that must surface before
or instead of a device call."""
import subprocess
import sys

import pytest

from msprime_b200 import cli


def parse(*argv):
    return cli.get_parser().parse_args(list(argv))


def test_defaults():
    args = parse("mutations", "0.1")
    assert args.mutation_rate == 0.1
    assert args.input is None and args.output is None  # both mean the standard streams
    assert args.random_seed is None and args.start_time is None and args.end_time is None
    assert args.model == "jc69" and args.device == 0
    assert args.runner is cli.run_mutations


def test_standard_streams_are_spelled_with_a_dash():
    assert parse("mutations", "10", "-").input == "-"
    assert parse("mutations", "10", "-o", "-").output == "-"
    assert parse("mutations", "10", "in.trees", "--output", "out.trees").output == "out.trees"


@pytest.mark.parametrize("flag", ["--model", "-m"])
@pytest.mark.parametrize("name", ["binary", "blosum62", "infinite_alleles", "jc69", "pam"])
def test_model_is_selectable(flag, name):
    assert parse("mutations", "1", flag, name).model == name


def test_numeric_options():
    args = parse("mutations", "1e-8", "in.trees", "-s", "42", "--start-time", "0", "--end-time", "2.5")
    assert (args.mutation_rate, args.random_seed, args.start_time, args.end_time) == (1e-8, 42, 0.0, 2.5)
    assert parse("mutations", "1", "--random-seed", "7").random_seed == 7


@pytest.mark.parametrize("argv", [
    ["mutations", "--model", "not-a-model", "out.trees", "out2.trees"],
    ["mutations"],  # the rate is required
    ["mutations", "fast"],
    ["mutations", "1", "-s", "1.5"],
    ["ancestry", "10"],  # only the mutation sub-command exists here
    [],
])
def test_bad_command_lines_exit(argv, capsys):
    with pytest.raises(SystemExit) as e:
        cli.get_parser().parse_args(argv)
    assert e.value.code == 2
    assert "usage" in capsys.readouterr().err


def test_missing_input_file_is_reported_before_any_device_work(tmp_path):
    r = subprocess.run([sys.executable, "-m", "msprime_b200", "mutations", "1", str(tmp_path / "nope.trees"),
                        "-o", str(tmp_path / "out.trees")], capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and not (tmp_path / "out.trees").exists()
    assert "nope.trees" in r.stderr or "No such file" in r.stderr
