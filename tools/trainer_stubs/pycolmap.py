"""pycolmap stand-in: nothing here reads a COLMAP reconstruction (datasets/colmap.py is synthetic)."""
