from . import numerical  # noqa: F401
