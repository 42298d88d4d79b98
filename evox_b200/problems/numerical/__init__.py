"""Mirror of the `evox.problems.numerical` entries on the hot path
(src/evox/problems/numerical/__init__.py:8-35)."""
from .dtlz import DTLZ1, DTLZ2, DTLZ3, DTLZ4, DTLZ5, DTLZ6, DTLZ7  # noqa: F401
from .lsmop import LSMOP1, LSMOP2, LSMOP3, LSMOP4, LSMOP5, LSMOP6, LSMOP7, LSMOP8, LSMOP9  # noqa: F401
from .maf import MaF1  # noqa: F401
