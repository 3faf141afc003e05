"""visual_slam_b200 -- B200-native match -> RANSAC pose -> bundle-adjustment hot path
behind the call signatures of juuso-oskari/visual_slam src/v2."""
from ._cabi import VslamError  # noqa: F401
from .matching import FeatureMatcher, match_batch  # noqa: F401
