"""TEST INFRASTRUCTURE - no-op pyplot (SSCS_maker.py:372-375 calls bar/xlim/savefig)."""


def bar(*a, **k):
    return None


def xlim(*a, **k):
    return None


def savefig(*a, **k):
    return None


# extract_barcodes.py:372-396 (barcode list mode) draws a bar chart of the barcode counts
class _Anything:
    def __getattr__(self, name):
        return lambda *a, **k: _Anything()

    def __getitem__(self, i):
        return _Anything()

    def __iter__(self):
        return iter(())


def subplots(*a, **k):
    return _Anything(), _Anything()


def gcf(*a, **k):
    return _Anything()


def tick_params(*a, **k):
    return None


def ylabel(*a, **k):
    return None
