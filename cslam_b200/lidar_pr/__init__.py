"""B200-native mirror of ``cslam.lidar_pr`` (only the matcher)."""
from .scancontext_matching import ScanContextMatching  # noqa: F401
