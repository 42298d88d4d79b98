"""Mirror of `evox.operators` for the hot path (src/evox/operators/__init__.py:1-5)."""
from . import selection  # noqa: F401
from .selection.non_dominate import (  # noqa: F401
    crowding_distance,
    crowding_distance_sort,
    non_dominated_sort,
)
