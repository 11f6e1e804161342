"""CPU oracle for the AudioFeatureExtractor hot path.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the reference computes every number with librosa 0.11.0, which is not
vendored in /root/reference and is not installable in this image (no network), and the
reference's own tests hold no golden vector for the extractor
(`tests/test_feature_extractor.py` is an empty file).  This package is a numpy/scipy
restatement of the librosa 0.11.0 algorithms called at
`audio_feature_extraction_toolkit/core/feature_extractor.py:69,72,87-94,127-138,164-168`,
cross-checked block by block against independent implementations that *are* installed
(scipy.signal, scipy.fft, torchaudio, transformers.audio_utils) in `tests/test_oracle.py`.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu-baseline / `--impl reference`
legs may import this package.  The product (`audio_feature_extraction_b200`) never does.
"""
from .librosa_restated import *  # noqa: F401,F403
from .extractor import OracleFeatureExtractor  # noqa: F401
