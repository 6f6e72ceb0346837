"""lib.point_tnf -> ncnet_b200.point_tnf (same names as the reference's lib/point_tnf.py)."""
from ncnet_b200.point_tnf import (PointsToPixelCoords, PointsToUnitCoords, bilinearInterpPointTnf,  # noqa: F401
                                  corr_to_matches, normalize_axis, unnormalize_axis)
