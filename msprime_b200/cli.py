"""Command line for the overlay: ``python -m msprime_b200 mutations RATE INPUT -o OUTPUT``.

The ``mutations`` sub-command of the reference's ``msp`` program (``msprime/cli.py:1043-1055,
1132-1173``) with the same arguments and defaults: the input is read from a ``.trees`` file,
mutations are added on the GPU with ``keep=True`` and a discrete genome, the result is written
as a ``.trees`` file.  Only this sub-command exists here; ancestry simulation is out of scope.
"""
from __future__ import annotations

import argparse
import os
import sys
import tempfile

from . import mutations, trees_io


def _input_path(name, stack):
    """``-`` / omitted means standard input (the reference accepts a stream there): the bytes
    are spooled to a temporary file because the reader maps its input."""
    if name not in (None, "-"):
        return name
    tmp = tempfile.NamedTemporaryFile(suffix=".trees", delete=False)
    stack.append(tmp.name)
    with tmp:
        while True:
            chunk = sys.stdin.buffer.read(1 << 24)
            if not chunk:
                break
            tmp.write(chunk)
    return tmp.name


def run_mutations(args):
    cleanup = []
    try:
        input_ts = trees_io.load(_input_path(args.input, cleanup))
        output_ts = mutations.sim_mutations(
            tree_sequence=input_ts,
            rate=args.mutation_rate,
            random_seed=args.random_seed,
            keep=True,
            start_time=args.start_time,
            end_time=args.end_time,
            discrete_genome=True,
            model=args.model,
            device=args.device,
        )
        if args.output in (None, "-"):
            with tempfile.NamedTemporaryFile(suffix=".trees", delete=False) as tmp:
                cleanup.append(tmp.name)
            trees_io.dump(output_ts, tmp.name)
            with open(tmp.name, "rb") as f:
                while True:
                    chunk = f.read(1 << 24)
                    if not chunk:
                        break
                    sys.stdout.buffer.write(chunk)
            sys.stdout.buffer.flush()
        else:
            trees_io.dump(output_ts, args.output)
    finally:
        for name in cleanup:
            try:
                os.unlink(name)
            except OSError:
                pass


def get_parser():
    top = argparse.ArgumentParser(
        prog="msprime_b200", description="Command line interface for the B200 mutation overlay.")
    sub = top.add_subparsers(dest="subcommand")
    sub.required = True
    p = sub.add_parser("mutations", help="Simulate mutations on a tree sequence.")
    p.add_argument("mutation_rate", type=float, help="The mutation rate per base")
    p.add_argument("input", nargs="?", default=None,
                   help="The input tree sequence. If omitted or '-' read from standard input")
    p.add_argument("-o", "--output", default=None,
                   help="Where to write the output tree sequence file. If omitted or '-', "
                        "write to standard output")
    p.add_argument("--random-seed", "-s", type=int, default=None,
                   help="The random seed; If not specified, one is chosen randomly.")
    p.add_argument("--start-time", type=float, default=None,
                   help="The minimum time ago at which a mutation can occur")
    p.add_argument("--end-time", type=float, default=None,
                   help="The maximum time ago at which a mutation can occur")
    p.add_argument("--model", "-m", type=str, default="jc69", choices=sorted(mutations.MODEL_MAP.keys()),
                   help="The mutation model to use (default=JC69)")
    p.add_argument("--device", type=int, default=0, help="CUDA device ordinal (default 0)")
    p.set_defaults(runner=run_mutations)
    return top


def main(arg_list=None):
    args = get_parser().parse_args(arg_list)
    args.runner(args)
