"""evox_b200 - B200-native drop-in for EvoX's multi-objective environmental-selection hot path.

    from evox_b200 import operators, utils, problems
    rank = operators.non_dominated_sort(fitness)            # evox.operators.non_dominated_sort
    dist = operators.crowding_distance(fitness, mask)       # evox.operators.crowding_distance
    F, state = problems.numerical.DTLZ2(d=12, m=3).evaluate(state, X)

All compute runs in evox_b200/libevoxmo.so (hand-written CUDA for sm_100a, C ABI in
include/evoxmo.h).  There is no CPU fallback: without the library or a CUDA device the calls raise.
"""
from . import operators, problems, utils  # noqa: F401
from ._lib import EvoxMoError, lib  # noqa: F401

__version__ = "0.1.0"
