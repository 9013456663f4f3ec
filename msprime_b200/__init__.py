"""msprime_b200: B200-native mutation overlay with the ``msprime.sim_mutations`` surface."""
from .core import __version__  # noqa: F401
from .intervals import RateMap  # noqa: F401
from .mutations import (  # noqa: F401
    BINARY,
    BLOSUM62,
    EL2,
    NUCLEOTIDES,
    PAM,
    SMM,
    TPM,
    F84,
    GTR,
    HKY,
    JC69,
    BinaryMutationModel,
    InfiniteAlleles,
    InfiniteSites,
    InputError,
    LibraryError,
    MatrixMutationModel,
    MicrosatMutationModel,
    MutationModel,
    SLiMMutationModel,
    mutate,
    mutation_model_factory,
    sim_mutations,
)
