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

# the names of the reference's ``msprime.__all__`` that belong to this path, plus the two error types
__all__ = [
    "BINARY", "BLOSUM62", "BinaryMutationModel", "EL2", "F84", "GTR", "HKY", "InfiniteAlleles", "InfiniteSites",
    "InputError", "JC69", "LibraryError", "MatrixMutationModel", "MicrosatMutationModel", "MutationModel",
    "NUCLEOTIDES", "PAM", "RateMap", "SLiMMutationModel", "SMM", "TPM", "mutate", "sim_mutations",
]
