"""B200-native exhaustive descriptor matching: the BruteForce hot path of Tarboucha/3DReconstruction
(`EnhancedBFMatcher`, FeatureMatchingExtraction/feature_matchers.py:184-252) behind the reference's
own plugin API.  The directory name starts with a digit, so import it with
`importlib.import_module("3dreconstruction_b200")` or through the `b200match` alias at the repo root.
"""
from . import _lib, contracts, synth                      # noqa: F401
from .engine import (NORM_HAMMING, NORM_L2, DeviceDescriptors, Engine, Knn2Arrays, MatchArrays,  # noqa: F401
                     default_engine, mode_of)
from .matcher import B200BFMatcher, install, uninstall    # noqa: F401

__all__ = ["B200BFMatcher", "Engine", "MatchArrays", "Knn2Arrays", "DeviceDescriptors", "install", "uninstall",
           "default_engine", "mode_of", "NORM_L2", "NORM_HAMMING", "contracts", "synth"]
