"""Empty stand-in for imageio.v2 (see __init__.py)."""
