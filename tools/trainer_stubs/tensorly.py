"""tensorly stand-in: examples/lib_bilagrid.py selects a backend at import; the bilateral grid is not used here."""


def set_backend(*a, **k):
    pass
