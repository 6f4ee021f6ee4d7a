"""cslam_b200 — B200-native global-descriptor nearest-neighbour search for cslam.

Only the hot path of nhahn/cslam named in BASELINE.json is implemented here:
``cslam.nns_matching.NearestNeighborsMatching`` (add_item / search / search_best) behind the
reference's own Python API, backed by hand-written sm_100a CUDA through a C-ABI shared library.
"""
from .nns_matching import NearestNeighborsMatching  # noqa: F401

__all__ = ["NearestNeighborsMatching"]
