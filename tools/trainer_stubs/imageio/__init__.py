"""imageio stand-in: images are written as .npy next to the requested name; videos are dropped."""
import numpy as np


def imwrite(path, array, *a, **k):
    np.save(str(path) + ".npy", np.asarray(array))


def imread(path, *a, **k):
    return np.load(str(path) + ".npy")


class _Writer:
    def append_data(self, *a, **k):
        pass

    def close(self):
        pass


def get_writer(*a, **k):
    return _Writer()
