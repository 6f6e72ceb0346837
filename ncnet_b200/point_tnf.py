"""Drop-in for the hot-path part of the reference's ``lib/point_tnf.py``."""
from .functional import corr_to_matches  # noqa: F401   (lib/point_tnf.py:12-80)


def normalize_axis(x, L):
    """pixel coordinate (1-based) -> [-1, 1]   (lib/point_tnf.py:6-7)"""
    return (x - 1 - (L - 1) / 2) * 2 / (L - 1)


def unnormalize_axis(x, L):
    """[-1, 1] -> pixel coordinate (1-based)   (lib/point_tnf.py:9-10)"""
    return x * (L - 1) / 2 + 1 + (L - 1) / 2
