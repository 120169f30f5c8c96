from . import get_writer, imread, imwrite  # noqa: F401
