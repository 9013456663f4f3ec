"""``sim_mutations`` and the mutation models: the Python surface of the path.

Re-authored drop-in for ``msprime/mutations.py``: same function signatures, argument
meaning, defaults and error behaviour (``sim_mutations`` :1551-1702, ``mutate``
:1423-1505, ``mutation_model_factory`` :1522-1548, model classes :78-201, 516-850),
with the C call ``_msprime.sim_mutations`` (``_msprimemodule.c:3047``) replaced by the
CUDA library behind ``include/msprime_b200.h``.
"""
from __future__ import annotations

import inspect
import math
import sys
import threading

import numpy as np

from . import _native, core, intervals, provenance, tables as _tables
from ._native import InputError, LibraryError  # noqa: F401  (re-exported)

_ACGT_ALLELES = ["A", "C", "G", "T"]
BINARY = 0
NUCLEOTIDES = 1


class MutationModel:
    """Abstract superclass of the mutation models (``msprime/mutations.py:60-75``)."""

    def asdict(self):
        return {
            key: getattr(self, key)
            for key in inspect.signature(self.__init__).parameters.keys()
            if hasattr(self, key)
        }


def _almost_equal(a, b, eps):
    # doubles_almost_equal, lib/util.c:372-376 (Knuth's relative comparison)
    if abs(a) < eps and abs(b) < eps:
        return True
    mx = a if abs(a) > abs(b) else b
    _, exponent = math.frexp(mx)
    delta = math.ldexp(eps, exponent)
    return not (a - b > delta) and not (a - b < -delta)


def _valid_probabilities(p):
    total, low = 0.0, 0.0
    for x in p:
        total += float(x)
        low = min(low, float(x))
    return _almost_equal(total, 1.0, 1e-12) and low >= 0.0


class MatrixMutationModel(MutationModel):
    """Finite-state model: ``alleles``, ``root_distribution`` and a row-stochastic
    ``transition_matrix``.  Validation follows ``matrix_mutation_model_alloc``
    (``lib/mutgen.c:501-524``) and the argument parsing of
    ``_msprimemodule.c:477-560`` (shape mismatches are ValueErrors)."""

    def __init__(self, alleles, root_distribution, transition_matrix):
        if isinstance(alleles, (str, bytes)):
            raise TypeError("alleles must be a list of strings")
        alleles = list(alleles)
        for a in alleles:
            if not isinstance(a, str):
                raise TypeError("alleles must be unicode strings")
        root = np.array(root_distribution, dtype=np.float64)
        matrix = np.array(transition_matrix, dtype=np.float64)
        n = len(alleles)
        if root.ndim != 1 or root.shape[0] != n:
            raise ValueError("root_distribution must have num_alleles elements")
        if matrix.ndim != 2 or matrix.shape != (n, n):
            raise ValueError("transition_matrix must be a square matrix with num_alleles rows")
        if n < 2:
            raise LibraryError(_native_message(-52), -52)
        if not _valid_probabilities(root):
            raise LibraryError(_native_message(-53), -53)
        for row in matrix:
            if not _valid_probabilities(row):
                raise LibraryError(_native_message(-54), -54)
        self._alleles = alleles
        self._root_distribution = root
        self._transition_matrix = matrix

    alleles = property(lambda self: list(self._alleles))
    root_distribution = property(lambda self: self._root_distribution.copy())
    transition_matrix = property(lambda self: self._transition_matrix.copy())

    def _native(self):
        return {
            "kind": _native.MODEL_MATRIX,
            "alleles": self._alleles,
            "root_distribution": self._root_distribution,
            "transition_matrix": self._transition_matrix,
        }

    def __str__(self):
        s = f"Mutation model with alleles {self.alleles}\n"
        s += "  root distribution: {}\n".format(" ".join(map(str, self.root_distribution)))
        s += "  transition matrix:\n"
        for row in self.transition_matrix:
            s += "     {}\n".format(" ".join(map(str, row)))
        return s


_MESSAGES = {
    -52: "Must have at least two alleles.",
    -53: "Root probabilities must be nonnegative and sum to one.",
    -54: "Each row of the transition matrix must be nonnegative and sum to one.",
    -56: "SLiM mutation IDs and mutation type IDs must be nonnegative.",
}


def _native_message(code):
    return _MESSAGES[code]


class InfiniteAlleles(MutationModel):
    """Every new allele (ancestral or derived) is the next unused integer.  The counter
    lives on the model object and persists across calls (``lib/mutgen.c:431-481``,
    ``_msprimemodule.c:876-890``)."""

    def __init__(self, start_allele=0):
        start_allele = int(start_allele)
        if start_allele < 0 or start_allele >= 2**64:
            raise OverflowError("start_allele must fit an unsigned 64 bit integer")
        self._start_allele = start_allele
        self._next_allele = start_allele

    start_allele = property(lambda self: self._start_allele)
    next_allele = property(lambda self: self._next_allele)

    def _native(self):
        return {"kind": _native.MODEL_INFINITE_ALLELES, "start_allele": self._next_allele}

    def __str__(self):
        return (
            f"Infinite alleles mutation model, beginning with allele"
            f" {self.start_allele}\n    next allele: {self.next_allele}\n"
        )


class SLiMMutationModel(MutationModel):
    """An infinite-alleles model producing "SLiM-style" mutations (``msprime/mutations.py:109-151``,
    ``lib/mutgen.c:257-413``): a new site's ancestral state is the empty string, a derived
    state is the parent's state with the next mutation id appended after a comma, and the
    metadata stacks one 17-byte ``MutationMetadataRec`` per mutation (mutation type ``type``,
    selection coefficient 0, subpopulation -1, origin generation ``slim_generation -
    floor(time)``, nucleotide -1).  The id counter lives on the model object and persists
    across calls (``_msprimemodule.c:748-760``).  ``block_size`` is accepted for
    compatibility; the device path has no block allocator."""

    def __init__(self, type, next_id=0, slim_generation=1, block_size=0):  # noqa: A002
        for name, v in (("type", type), ("next_id", next_id), ("slim_generation", slim_generation),
                        ("block_size", block_size)):
            if isinstance(v, bool) or not isinstance(v, (int, np.integer)):
                raise TypeError(f"{name} must be an integer")
        if not -2**63 <= int(next_id) < 2**63 or not -2**31 <= int(type) < 2**31 \
                or not -2**31 <= int(slim_generation) < 2**31:
            raise OverflowError("SLiM model parameter out of range")
        if int(next_id) < 0 or int(type) < 0:
            # slim_mutator_check_validity, lib/mutgen.c:326-343 (MSP_ERR_BAD_SLIM_PARAMETERS)
            raise _native.LibraryError(_native_message(-56))
        self._type = int(type)
        self._next_id = int(next_id)
        self._slim_generation = int(slim_generation)

    type = property(lambda self: self._type)  # noqa: A003
    next_id = property(lambda self: self._next_id)
    slim_generation = property(lambda self: self._slim_generation)

    def _native(self):
        return {"kind": _native.MODEL_SLIM, "type": self._type, "next_id": self._next_id,
                "slim_generation": self._slim_generation}

    def asdict(self):
        return {"type": self._type, "next_id": self._next_id, "slim_generation": self._slim_generation}

    def __str__(self):
        return (
            f"Mutation model for SLiM mutations of type m{self.type}\n"
            f"  next ID: {self.next_id}\n"
        )


class BinaryMutationModel(MatrixMutationModel):
    """Alleles "0"/"1", ancestral state always "0" (``msprime/mutations.py:179-201``)."""

    def __init__(self, state_independent=False):
        if state_independent:
            matrix = [[0.5, 0.5], [0.5, 0.5]]
        else:
            matrix = [[0.0, 1.0], [1.0, 0.0]]
        self.state_independent = state_independent
        super().__init__(["0", "1"], [1, 0], matrix)


class JC69(MatrixMutationModel):
    """Jukes-Cantor: all nucleotide changes equally likely (``msprime/mutations.py:516-547``)."""

    def __init__(self, state_independent=False):
        matrix = np.zeros((4, 4))
        if state_independent:
            matrix[:] = 1 / 4
        else:
            matrix[:] = 1 / 3
            np.fill_diagonal(matrix, 0)
        self.state_independent = state_independent
        super().__init__(_ACGT_ALLELES, [0.25, 0.25, 0.25, 0.25], matrix)


class HKY(MatrixMutationModel):
    """Hasegawa-Kishino-Yano: transitions ``kappa`` times as likely as transversions,
    scaled by the equilibrium frequencies (``msprime/mutations.py:550-649``).  The
    arithmetic order below matches the reference so the matrix entries are identical."""

    def __init__(self, kappa, equilibrium_frequencies=None, root_distribution=None,
                 state_independent=False):
        self.kappa = kappa
        self.state_independent = state_independent
        if equilibrium_frequencies is None:
            freqs = np.array([0.25, 0.25, 0.25, 0.25], dtype="float64")
        else:
            freqs = np.array(equilibrium_frequencies, dtype="float64")
        self.equilibrium_frequencies = freqs
        if root_distribution is None:
            root_distribution = freqs.copy()
        q = np.full((4, 4), 1.0)
        q[((0, 1, 2, 3), (2, 3, 0, 1))] = kappa  # A<->G, C<->T
        q *= freqs
        if state_independent:
            np.fill_diagonal(q, kappa * freqs)
        else:
            np.fill_diagonal(q, 0.0)
        q /= max(q.sum(axis=1, dtype="float64"))
        sums = q.sum(axis=1, dtype="float64")
        np.fill_diagonal(q, 1.0 - (sums - np.diag(q)))
        super().__init__(_ACGT_ALLELES, root_distribution, q)


class F84(MatrixMutationModel):
    """Felsenstein 1984 (``msprime/mutations.py:652-760``)."""

    def __init__(self, kappa, equilibrium_frequencies=None, root_distribution=None,
                 state_independent=False):
        self.kappa = kappa
        self.state_independent = state_independent
        if equilibrium_frequencies is None:
            equilibrium_frequencies = [0.25, 0.25, 0.25, 0.25]
        self.equilibrium_frequencies = equilibrium_frequencies
        if root_distribution is None:
            root_distribution = equilibrium_frequencies.copy()
        f = equilibrium_frequencies
        q = np.full((4, 4), 1.0, dtype="float64")
        gamma = np.float64(1 + (kappa - 1) / np.array([f[0] + f[2], f[1] + f[3]]))
        q[((0, 2), (2, 0))] = gamma[0]
        q[((1, 3), (3, 1))] = gamma[1]
        q *= f
        if state_independent:
            np.fill_diagonal(q, f * np.concatenate([gamma, gamma]))
        else:
            np.fill_diagonal(q, 0.0)
        q = q / max(q.sum(axis=1))
        sums = q.sum(axis=1, dtype="float64")
        np.fill_diagonal(q, 1.0 - (sums - np.diag(q)))
        super().__init__(_ACGT_ALLELES, root_distribution, q)


class GTR(MatrixMutationModel):
    """General time-reversible model; ``relative_rates`` are A<->C, A<->G, A<->T, C<->G,
    C<->T, G<->T (``msprime/mutations.py:763-850``)."""

    def __init__(self, relative_rates, equilibrium_frequencies=None, root_distribution=None,
                 state_independent=False):
        assert len(relative_rates) == 6
        self.relative_rates = relative_rates
        self.state_independent = state_independent
        if equilibrium_frequencies is None:
            equilibrium_frequencies = [0.25, 0.25, 0.25, 0.25]
        self.equilibrium_frequencies = equilibrium_frequencies
        if root_distribution is None:
            root_distribution = equilibrium_frequencies
        q = np.zeros((4, 4))
        q[np.triu_indices_from(q, k=1)] = relative_rates
        q += q.T
        if state_independent:
            np.fill_diagonal(q, 1.0)
        q *= equilibrium_frequencies
        q = q / max(q.sum(axis=1))
        sums = q.sum(axis=1, dtype="float64")
        np.fill_diagonal(q, np.diag(q) + 1.0 - sums)
        super().__init__(_ACGT_ALLELES, root_distribution, q)


def general_microsat_rate_matrix(s, u, v, p, m, lo, hi):
    """Jump-chain transition matrix of the microsatellite model of Sainudiin et al. (2004) on
    repeat numbers ``lo .. hi`` (``msprime/mutations.py:204-304``).

    From allele ``i`` the chain moves up with weight ``alpha(i)`` and down with ``1 - alpha(i)``,
    ``alpha(i) = clip(u - v (i - lo), 0, 1)``; the overall rate of ``i`` is ``beta(i) = 1 + s (i - lo)``;
    a move is a single step with probability ``p`` and otherwise a geometric(``m``) number of steps
    truncated to stay inside ``[lo, hi]``.  The rates are divided by the largest row sum and the
    diagonal takes up the rest of each row.  Built here as whole-array numpy expressions in the
    reference's order of operations, so the entries are the same doubles."""
    s, u, v, p, m = (float(x) for x in (s, u, v, p, m))
    lo, hi = int(lo), int(hi)
    n = hi - lo + 1
    q = 1.0 - m
    # (1 - m) ** k twice: the reference raises to a numpy integer in the numerator (numpy's pow)
    # and to a Python integer in the denominator (libm's pow); the two differ in the last bit
    powers_np = np.array([q ** np.int64(k) for k in range(n + 1)], dtype=np.float64)
    powers = np.array([q ** k for k in range(n + 1)], dtype=np.float64)
    idx = np.arange(n)
    i, j = np.meshgrid(idx, idx, indexing="ij")
    dist = np.abs(i - j)
    up, down = j > i, j < i
    # truncated geometric step sizes: m (1-m)^(d-1) / (1 - (1-m)^(room)), room = hi-i going up, i-lo going down
    num = m * powers_np[np.maximum(dist - 1, 0)]
    room = np.where(up, (n - 1) - i, i)
    with np.errstate(divide="ignore", invalid="ignore"):
        gamma = num / (1.0 - powers[room])
    gamma[~(up | down)] = 0.0  # the diagonal (and its 0 / 0 when m = 1) is not a move
    alpha = np.array([max(0.0, min(1.0, u - (v * k))) for k in range(n)], dtype=np.float64)
    beta = np.array([1.0 + (k * s) for k in range(n)], dtype=np.float64)
    direction = np.where(up, (beta * alpha)[:, None], (beta * (1.0 - alpha))[:, None])
    single = direction * (p + ((1.0 - p) * gamma))
    multi = direction * (1.0 - p) * gamma
    matrix = np.where(dist == 1, single, multi)
    matrix[~(up | down)] = 0.0
    matrix = np.ascontiguousarray(matrix, dtype=np.float64)
    matrix /= np.max(matrix.sum(axis=1, dtype="float64"))
    row_sums = matrix.sum(axis=1, dtype="float64")
    np.fill_diagonal(matrix, np.fmax(0.0, 1.0 - row_sums))
    return matrix


class MicrosatMutationModel(MatrixMutationModel):
    """Microsatellite repeat-number model with the five parameters of Sainudiin et al. (2004):
    alleles ``"lo" .. "hi"``, transition matrix :func:`general_microsat_rate_matrix`, root
    distribution = the chain's stationary distribution unless given
    (``msprime/mutations.py:307-398``; same defaults and argument errors)."""

    def __init__(self, *, s=None, u=None, v=None, p=None, m=None, lo=None, hi=None, root_distribution=None):
        import numbers

        s = 0.0 if s is None else s
        u = 0.5 if u is None else u
        v = 0.0 if v is None else v
        p = 1.0 if p is None else p
        m = 1.0 if m is None else m
        lo = 1 if lo is None else lo
        hi = 50 if hi is None else hi
        params = [s, u, v, p, m, lo, hi]
        if any(not math.isfinite(x) for x in params):
            raise ValueError("All parameters must be finite")
        if any(not isinstance(x, numbers.Number) for x in params):
            raise ValueError("All parameters must be numeric")
        lo, hi = int(lo), int(hi)
        if lo > hi:
            raise ValueError("lo > hi")
        if lo < 0:
            raise ValueError("lo must be >= 0")
        if s <= -1 / (hi - lo + 1):
            raise ValueError("s must be > -1/(hi-lo+1)")
        if u <= 0 or u >= 1:
            raise ValueError("u must be between 0 and 1")
        alleles = [str(k) for k in range(lo, hi + 1)]
        matrix = general_microsat_rate_matrix(s, u, v, p, m, lo, hi)
        if root_distribution is None:
            # left eigenvector of eigenvalue 1
            values, vectors = np.linalg.eig(matrix.T)
            vectors = np.real_if_close(vectors, tol=1)
            stationary = np.array(vectors[:, np.where(np.abs(values - 1.0) < 1e-8)[0][0]])
            root_distribution = stationary / np.sum(stationary)
        self.s, self.u, self.v, self.p, self.m, self.lo, self.hi = s, u, v, p, m, lo, hi
        super().__init__(alleles, root_distribution, matrix)


class SMM(MicrosatMutationModel):
    """Stepwise mutation model (Ohta & Kimura 1978): repeat number changes by +/- 1 with equal
    probability, reflecting at ``lo`` and ``hi`` (``msprime/mutations.py:401-425``)."""

    def __init__(self, lo=None, hi=None, *, root_distribution=None):
        super().__init__(s=0, u=0.5, v=0, p=0, m=1, lo=lo, hi=hi, root_distribution=root_distribution)


class TPM(MicrosatMutationModel):
    """Two-phase model (Di Rienzo et al. 1994): single steps with probability ``p``, otherwise
    geometric(``m``) multi-steps (``msprime/mutations.py:428-463``)."""

    def __init__(self, *, p, m, lo=None, hi=None, root_distribution=None):
        if not (0 < p < 1.0):
            raise ValueError("p must be between 0 and 1")
        if not (0 < m < 1.0):
            raise ValueError("m must be between 0 and 1")
        super().__init__(s=0, u=0.5, v=0, p=p, m=m, lo=lo, hi=hi, root_distribution=root_distribution)


class EL2(MicrosatMutationModel):
    """Equal-rate, linear-bias, two-phase model (Garza et al. 1995 as parameterised by Sainudiin
    et al. 2004): multi-steps only, bias ``u - v (i - lo)`` (``msprime/mutations.py:466-513``)."""

    def __init__(self, *, m, u, v, lo=None, hi=None, root_distribution=None):
        if not (0 < m < 1.0):
            raise ValueError("m must be between 0 and 1")
        if not (0 < u < 1.0):
            raise ValueError("u must be between 0 and 1")
        super().__init__(s=0, u=u, v=v, p=0, m=m, lo=lo, hi=hi, root_distribution=root_distribution)


_AA_DATA = None


def _amino_acid_model(name):
    """(alleles, root distribution, transition matrix) of an empirical amino-acid model from
    ``data/amino_acid_models.npz`` (published tables; ``scripts/make_aa_models.py`` says where
    the numbers come from)."""
    global _AA_DATA
    if _AA_DATA is None:
        import os

        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "amino_acid_models.npz")
        with np.load(path) as z:
            _AA_DATA = {k: z[k] for k in z.files}
    return ([str(a) for a in _AA_DATA[f"{name}_alleles"]], _AA_DATA[f"{name}_root_distribution"].copy(),
            _AA_DATA[f"{name}_transition_matrix"].copy())


class BLOSUM62(MatrixMutationModel):
    """BLOSUM62 amino-acid substitution model (Henikoff & Henikoff 1992), 20 alleles in
    Seq-Gen's order (``msprime/mutations.py:853-1114``)."""

    def __init__(self):
        super().__init__(*_amino_acid_model("blosum62"))


class PAM(MatrixMutationModel):
    """Dayhoff PAM amino-acid substitution model (Dayhoff et al. 1978), 20 alleles in Seq-Gen's
    order (``msprime/mutations.py:1117-1387``)."""

    def __init__(self):
        super().__init__(*_amino_acid_model("pam"))


class InfiniteSites(MatrixMutationModel):
    """Legacy model object accepted by ``mutate`` (``msprime/mutations.py:1390-1400``)."""

    def __init__(self, alphabet=BINARY):
        self.alphabet = alphabet
        models = {BINARY: BinaryMutationModel, NUCLEOTIDES: JC69}
        if alphabet not in models:
            raise ValueError("Bad alphabet")
        model = models[alphabet]()
        super().__init__(model.alleles, model.root_distribution, model.transition_matrix)


MODEL_MAP = {
    "infinite_alleles": InfiniteAlleles,
    "binary": BinaryMutationModel,
    "jc69": JC69,
    "blosum62": BLOSUM62,
    "pam": PAM,
}


def mutation_model_factory(model):
    """None -> JC69, a name -> that model, an instance -> itself; otherwise TypeError."""
    if model is None:
        return JC69()
    if isinstance(model, str):
        lower = model.lower()
        if lower not in MODEL_MAP:
            raise ValueError(
                "Model '{}' unknown. Choose from {}".format(model, sorted(MODEL_MAP.keys())))
        return MODEL_MAP[lower]()
    if isinstance(model, MutationModel):
        return model
    raise TypeError("Mutation model must be a string or an instance of MutationModel")


_contexts: dict[int, _native.Context] = {}
# One context per device is shared by every caller, and load_tables / generate / fetch are three
# ctypes calls that release the GIL: the sequence runs under the device's lock, so that two threads
# calling sim_mutations on one device cannot generate over, or fetch, each other's tables (the
# reference keeps all state per call, msprime/_msprimemodule.c:3047-3120).
_contexts_guard = threading.Lock()
_device_locks: dict[int, threading.Lock] = {}


def _device_lock(device: int) -> threading.Lock:
    with _contexts_guard:
        lock = _device_locks.get(device)
        if lock is None:
            lock = _device_locks[device] = threading.Lock()
        return lock


def _context(device: int) -> _native.Context:
    with _contexts_guard:
        ctx = _contexts.get(device)
        if ctx is None:
            ctx = _contexts[device] = _native.Context(device)
        return ctx


def _overlay(ctx, tables_dict, *, rate_map, model, discrete_genome, keep, start_time, end_time,
             seed, replay_stream=None, slice_left=0.0, slice_right=0.0, load=True):
    """Runs the device overlay and returns ``(result_columns, counts)``; updates the
    model's allele counter like the C call does."""
    if load:
        ctx.load_tables(tables_dict)
    params = _native.Params(
        map_position=rate_map["position"], map_rate=rate_map["rate"], model=model._native(),
        discrete_genome=discrete_genome, keep=keep, start_time=start_time, end_time=end_time,
        seed=seed, replay_stream=replay_stream, slice_left=slice_left, slice_right=slice_right)
    counts = ctx.generate(params)
    if isinstance(model, InfiniteAlleles):
        model._next_allele = counts["next_allele"]
    elif isinstance(model, SLiMMutationModel):
        model._next_id = counts["next_allele"]
    return ctx.fetch(), counts


def _merge_result(tables_dict, result):
    """New table dict: everything as in the input, sites and mutations replaced
    (``_msprime.sim_mutations`` touches nothing else, tests/test_mutations.py:593-598)."""
    out = dict(tables_dict)
    s, m = result["sites"], result["mutations"]
    out["sites"] = {
        "position": s["position"],
        "ancestral_state": s["ancestral_state"].view(np.int8),
        "ancestral_state_offset": s["ancestral_state_offset"],
        "metadata": s["metadata"].view(np.int8),
        "metadata_offset": s["metadata_offset"],
    }
    for key in ("metadata_schema",):
        if tables_dict.get("sites") is not None and key in tables_dict["sites"]:
            out["sites"][key] = tables_dict["sites"][key]
    out["mutations"] = {
        "site": m["site"],
        "node": m["node"],
        "parent": m["parent"],
        "time": m["time"],
        "derived_state": m["derived_state"].view(np.int8),
        "derived_state_offset": m["derived_state_offset"],
        "metadata": m["metadata"].view(np.int8),
        "metadata_offset": m["metadata_offset"],
    }
    for key in ("metadata_schema",):
        if tables_dict.get("mutations") is not None and key in tables_dict["mutations"]:
            out["mutations"][key] = tables_dict["mutations"][key]
    # edge indexes stay valid (edges are untouched); tskit rebuilds them otherwise
    return out


def sim_mutations(
    tree_sequence,
    rate=None,
    *,
    random_seed=None,
    model=None,
    start_time=None,
    end_time=None,
    discrete_genome=None,
    keep=None,
    record_provenance=True,
    device=0,
):
    """Simulates mutations on ``tree_sequence`` on a B200 and returns the new tree sequence.

    Same contract as ``msprime.sim_mutations`` (``msprime/mutations.py:1551-1702``):
    ``rate`` is a float or a :class:`RateMap` (default 0); ``model`` is a name, a
    :class:`MutationModel` or None (JC69); ``discrete_genome`` and ``keep`` default to
    True and must be real booleans; ``start_time``/``end_time`` bound the mutation
    times; ``random_seed`` lies in 1 .. 2**32 - 1.  The only extra argument is
    ``device`` (CUDA ordinal).  The input is never modified.

    The random stream is Philox4x32-10 keyed by the seed, so results are reproducible
    for a seed but differ from stock msprime's MT19937 stream; the distribution is
    the same (tests/test_statistics.py).
    """
    tables_dict, kind = _tables.from_tree_sequence(tree_sequence)
    if random_seed is None:
        seed = core.get_random_seed()
    else:
        seed = int(random_seed)

    provenance_dict = None
    if record_provenance:
        provenance_dict = provenance.get_provenance_dict(dict(
            command="sim_mutations", tree_sequence=tree_sequence, rate=rate, model=model,
            start_time=start_time, end_time=end_time, discrete_genome=discrete_genome,
            keep=keep, random_seed=seed))

    if rate is None:
        rate = 0
    try:
        rate = float(rate)
        rate_map = intervals.RateMap.uniform(tree_sequence.sequence_length, rate)
    except TypeError:
        rate_map = rate
    if not isinstance(rate_map, intervals.RateMap):
        raise TypeError("rate must be a float or a RateMap")

    if tables_dict.get("time_units") == _tables.TIME_UNITS_UNCALIBRATED:
        raise ValueError("Simulating mutations doesn't make sense when time is uncalibrated")

    start_time = -sys.float_info.max if start_time is None else float(start_time)
    end_time = sys.float_info.max if end_time is None else float(end_time)
    if start_time > end_time:
        raise ValueError("start_time must be <= end_time")
    discrete_genome = core.parse_flag(discrete_genome, default=True)
    keep = core.parse_flag(keep, default=True)
    model = mutation_model_factory(model)
    seed = core.check_seed(seed)

    with _device_lock(device):
        result, _ = _overlay(
            _context(device), tables_dict, rate_map=rate_map.asdict(), model=model,
            discrete_genome=discrete_genome, keep=keep, start_time=start_time, end_time=end_time,
            seed=seed)
    out = _merge_result(tables_dict, result)
    record = None
    if provenance_dict is not None:
        record = provenance.json_encode_provenance(provenance_dict)
    return _tables.to_tree_sequence(out, kind, record)


def mutate(tree_sequence, rate=None, random_seed=None, model=None, keep=False,
           start_time=None, end_time=None):
    """Legacy interface (``msprime/mutations.py:1423-1505``): infinite-sites mutations on
    a continuous genome, binary alleles by default, ``keep=False`` by default."""
    if model is not None and not isinstance(model, InfiniteSites):
        raise ValueError(
            "This is a legacy interface which does not support general "
            "mutation models. Please use the sim_ancestry function instead.")
    if model is None:
        model = BinaryMutationModel()
    return sim_mutations(
        tree_sequence, rate=rate, random_seed=random_seed, model=model, keep=keep,
        start_time=start_time, end_time=end_time, discrete_genome=False)
