"""Structural checks of an overlay result against its input tree sequence (numpy only): the
checks live in ``msprime_b200/validate.py`` (bench.py runs them on its last timed overlay too)."""
from msprime_b200.validate import (check_count, check_infinite_alleles_numbering, check_result,  # noqa: F401
                                   edge_of, expected_new_mutations, ragged)
