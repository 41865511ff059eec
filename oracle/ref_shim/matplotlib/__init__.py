"""Stub of matplotlib (not installed here); only what the reference touches at import time."""


class _RC(dict):
    def update(self, *a, **k):
        pass


rcParams = _RC()


def cycler(*a, **k):
    return None


def use(*a, **k):
    pass
