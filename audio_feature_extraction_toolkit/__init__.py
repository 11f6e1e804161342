"""Import-path shim: `from audio_feature_extraction_toolkit import AudioFeatureExtractor, FeatureEvaluator`
(the reference's public import, audio_feature_extraction_toolkit/__init__.py:1-6) resolves to the
B200-native implementation in `audio_feature_extraction_b200`."""
from audio_feature_extraction_b200 import AudioFeatureExtractor, FeatureEvaluator, FeatureExtractionError, __version__

__all__ = ["AudioFeatureExtractor", "FeatureEvaluator"]
