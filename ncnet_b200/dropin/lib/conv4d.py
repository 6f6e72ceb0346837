"""lib.conv4d -> ncnet_b200.conv4d (same names as the reference's lib/conv4d.py)."""
from ncnet_b200.conv4d import Conv4d, conv4d  # noqa: F401
