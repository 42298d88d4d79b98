from .non_dominate import NonDominate, non_dominate  # noqa: F401  (src/evox/operators/selection/__init__.py:7)
