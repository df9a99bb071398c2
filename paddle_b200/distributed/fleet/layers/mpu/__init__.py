from . import mp_layers, mp_ops, random  # noqa: F401
from .mp_layers import (ColumnParallelLinear, ParallelCrossEntropy, RowParallelLinear,  # noqa: F401
                        VocabParallelEmbedding)
from .random import RNGStatesTracker, get_rng_state_tracker, model_parallel_random_seed  # noqa: F401
