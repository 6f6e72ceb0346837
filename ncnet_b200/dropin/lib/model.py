"""lib.model -> ncnet_b200.model (same names as the reference's lib/model.py)."""
from ncnet_b200.model import (FeatureCorrelation, FeatureExtraction, ImMatchNet, MutualMatching,  # noqa: F401
                              NeighConsensus, featureL2Norm, maxpool4d)
from ncnet_b200.conv4d import Conv4d  # noqa: F401
