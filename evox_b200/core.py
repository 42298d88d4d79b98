"""The calling convention of the reference's problems (src/evox/core/problem.py:9-31,
src/evox/core/state.py:28-65): `evaluate(state, X) -> (F, state)` with the state passed through
untouched.  Only what that convention needs is mirrored here; the reference's state tree, jit
helpers and workflows stay where they are."""


class State(dict):
    """Minimal stand-in for evox.State: an immutable-by-convention mapping with attribute access."""

    def __init__(self, **kw):
        super().__init__(**kw)

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def replace(self, **kw):
        out = State(**self)
        out.update(kw)
        return out


class Problem:
    def init(self, key=None):
        return self.setup(key)

    def setup(self, key):
        return State(key=key)

    def evaluate(self, state, pop):
        raise NotImplementedError
