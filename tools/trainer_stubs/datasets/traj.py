"""examples/datasets/traj.py stand-in: trajectory rendering (a video) is skipped in this run."""


def _unavailable(*a, **k):
    raise RuntimeError("trajectory rendering is skipped offline (needs a video writer)")


generate_interpolated_path = generate_ellipse_path_z = generate_spiral_path = _unavailable
