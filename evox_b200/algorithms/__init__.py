from .nsga2 import NSGA2, run_generations  # noqa: F401
from . import spea2  # noqa: F401
from . import nsga3  # noqa: F401
