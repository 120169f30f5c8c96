"""nerfview stand-in (the trainer runs with disable_viewer=True)."""


class CameraState:
    pass


class Viewer:
    def __init__(self, *a, **k):
        raise RuntimeError("viewer is not available offline: run with disable_viewer=True")


def apply_float_colormap(*a, **k):
    raise RuntimeError("viewer is not available offline")
