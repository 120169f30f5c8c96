def get_cmap(*a, **k):
    raise RuntimeError("matplotlib is not available offline")
