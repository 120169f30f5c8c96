"""viser stand-in (the trainer runs with disable_viewer=True)."""


class ViserServer:
    def __init__(self, *a, **k):
        raise RuntimeError("viewer is not available offline: run with disable_viewer=True")
