"""B200-native drop-in for audio_feature_extraction_toolkit's AudioFeatureExtractor hot path.

Export surface mirrors /root/reference/audio_feature_extraction_toolkit/__init__.py:1-6.
"""
from .core.feature_extractor import AudioFeatureExtractor, FeatureExtractionError
from .evaluation.evaluator import FeatureEvaluator

__version__ = "0.1.0"

__all__ = ["AudioFeatureExtractor", "FeatureEvaluator", "FeatureExtractionError"]
