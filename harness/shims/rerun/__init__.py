"""Stand-in for `rerun` (viewer SDK, absent; no graphics stack on this box).  dpvo/dpvo.py:7 imports it at module top and only
touches it when DPVO(..., rerun=True); every attribute is a no-op callable."""


class _Nop:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return self

    def __getattr__(self, name):
        return _Nop()


def __getattr__(name):
    return _Nop()
