"""TEST INFRASTRUCTURE - no-op pyplot (SSCS_maker.py:372-375 calls bar/xlim/savefig)."""


def bar(*a, **k):
    return None


def xlim(*a, **k):
    return None


def savefig(*a, **k):
    return None
