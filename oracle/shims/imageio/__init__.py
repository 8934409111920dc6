"""Empty stand-in: the reference imports imageio in datasets.py but the hot path never calls it."""
