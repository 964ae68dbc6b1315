"""TEST INFRASTRUCTURE -- ``munch`` stand-in: attribute-access dicts."""


class Munch(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def munchify(x):
    if isinstance(x, dict):
        return Munch((k, munchify(v)) for k, v in x.items())
    if isinstance(x, list):
        return [munchify(v) for v in x]
    return x
