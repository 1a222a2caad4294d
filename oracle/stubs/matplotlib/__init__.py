"""Inert stand-in for matplotlib so the reference modules import without a plotting stack.

TEST INFRASTRUCTURE ONLY (see oracle/refharness.py). The reference imports
matplotlib at module import time (modulation.py:1, distortion.py:1, utilities.py:9-10,
plot_settings.py:3-5) but never needs it for the link arithmetic.
"""


class _Inert:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Inert()

    def __getattr__(self, name):
        return _Inert()

    def __iter__(self):
        return iter((_Inert(), _Inert()))

    def __getitem__(self, k):
        return _Inert()

    def __setitem__(self, k, v):
        pass


rcParams = {}
rcParamsDefault = {}


def rc(*a, **k):
    pass


def use(*a, **k):
    pass


def cycler(*a, **k):
    return _Inert()
