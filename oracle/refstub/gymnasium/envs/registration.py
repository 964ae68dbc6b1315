"""TEST INFRASTRUCTURE -- ``register`` is a no-op here."""


def register(id=None, entry_point=None, **kw):
    pass
