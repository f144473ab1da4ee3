"""B200-native (sm_100a CUDA) implementation of line2Dup::Detector::match behind the reference's API.

    from shape_based_matching_b200 import Detector
"""
from ._capi import DEVICE_REC, FEATURE, MATCH_REC, TEMPLATE_DESC, SbmError  # noqa: F401
from .detector import Detector, Feature, Match, Template  # noqa: F401
