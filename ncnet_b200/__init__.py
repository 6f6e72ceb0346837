"""ncnet_b200 -- B200-native matching head of Neighbourhood Consensus Networks.

Public surface mirrors the reference's ``lib.model`` / ``lib.conv4d`` / ``lib.point_tnf`` names for
the hot path; everything computes through libncnet_b200.so (include/ncnet_b200.h).
"""
from .conv4d import Conv4d, conv4d  # noqa: F401
from .model import (FeatureCorrelation, FeatureExtraction, ImMatchNet, MutualMatching, NeighConsensus,  # noqa: F401
                    featureL2Norm, maxpool4d)
from .point_tnf import corr_to_matches, normalize_axis, unnormalize_axis  # noqa: F401

__version__ = "0.1.0"
