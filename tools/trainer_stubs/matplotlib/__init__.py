"""matplotlib stand-in: examples/utils.py imports pyplot / colormaps for depth colouring, unused in this run."""
colormaps = {}
