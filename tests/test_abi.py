"""The C-ABI library loads, exports every symbol include/afx_b200.h declares, and fails loudly
(no CPU fallback) when no sm_100a device is present.  No compute calls: runs without a GPU."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def declared_symbols():
    text = (ROOT / "include" / "afx_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(afx_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for s in ("afx_create", "afx_destroy", "afx_extract", "afx_extract_host", "afx_extract_host_packed",
              "afx_workspace_bytes", "afx_summary_width", "afx_frame_width", "afx_last_error"):
        assert s in syms


def test_library_exports_every_declared_symbol(built_lib):
    for s in declared_symbols():
        assert hasattr(built_lib, s), f"libafx_b200.so does not export {s}"
    from audio_feature_extraction_b200 import _native as N
    assert sorted(N.EXPORTS) == sorted(declared_symbols())


def test_abi_version_and_defaults(built_lib):
    from audio_feature_extraction_b200 import _native as N
    assert built_lib.afx_abi_version() == N.ABI_VERSION
    cfg = N.AfxConfig()
    built_lib.afx_config_default(C.byref(cfg))
    # AudioFeatureExtractor.__init__ defaults, core/feature_extractor.py:10-17
    assert (cfg.sr, cfg.frame_length, cfg.hop_length, cfg.n_mfcc) == (22050, 1024, 256, 13)
    assert cfg.f0_min == pytest.approx(65.40639132514966, rel=0, abs=1e-12)
    assert cfg.f0_max == pytest.approx(2093.004522404789, rel=0, abs=1e-9)
    assert cfg.pre_emphasis == pytest.approx(0.97)
    assert (cfg.trim_top_db, cfg.n_mels, cfg.mfcc_top_db) == (30.0, 128, 80.0)


def test_invalid_config_is_rejected_before_touching_cuda(built_lib):
    from audio_feature_extraction_b200 import _native as N
    cfg = N.AfxConfig()
    built_lib.afx_config_default(C.byref(cfg))
    cfg.frame_length = 1000
    h = C.c_void_p()
    rc = built_lib.afx_create(C.byref(cfg), C.byref(h))
    assert rc in (N.ERR_UNSUPPORTED, N.ERR_INVALID_ARG)
    assert b"frame_length" in built_lib.afx_last_error()


def test_no_cpu_fallback_without_a_device(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
    import numpy as np
    ex = AudioFeatureExtractor()
    with pytest.raises(N.AfxError) as ei:
        ex._h()
    assert ei.value.code == N.ERR_CUDA


def test_constructor_rejects_unsupported_geometry_up_front(built_lib):
    """the reference's constructor accepts any frame_length / hop_length; where no kernel exists the drop-in
    says so in __init__ (host-only check), not at the first extraction call."""
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    for kw in (dict(frame_length=400, hop_length=160), dict(frame_length=1024, hop_length=600), dict(frame_length=1024, hop_length=0),
               dict(n_mfcc=65), dict(n_mfcc=0)):
        with pytest.raises(ValueError):
            AudioFeatureExtractor(**kw)
    for kw in (dict(sr=16000, frame_length=512, hop_length=160), dict(frame_length=2048, hop_length=441), dict(frame_length=1024, hop_length=33),
               dict(n_mfcc=64)):
        AudioFeatureExtractor(**kw)   # no device touched yet


def test_product_never_imports_the_oracle():
    pkg = ROOT / "audio_feature_extraction_b200"
    for p in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")) + list((ROOT / "audio_feature_extraction_toolkit").rglob("*.py")):
        txt = p.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), p


def test_stock_configurations_select_the_unrolled_kernel_variant(built_lib):
    """MelCnt<NFFT, true> (csrc/afx_warp.cuh) must carry the run lengths of the mel tables the host builds for the stock
    pairings; any other sample rate or hop takes the table-driven variant, unsupported pairings none."""
    runs = (C.c_int32 * 5)()
    for sr, n_fft, hop, expect in ((22050, 1024, 256, (2, 3, 6, 14, 2)), (16000, 512, 128, (1, 2, 3, 7, 1)),
                                   (44100, 2048, 512, (2, 4, 12, 33, 2))):
        assert built_lib.afx_clip_kernel_variant(sr, n_fft, hop, runs) == 1, (sr, n_fft, hop, list(runs))
        assert tuple(runs) == expect
    assert built_lib.afx_clip_kernel_variant(22050, 1024, 512, None) == 0     # stock tables, other hop
    assert built_lib.afx_clip_kernel_variant(16000, 1024, 256, None) == 0     # other sample rate: table-driven walk
    assert built_lib.afx_clip_kernel_variant(8000, 512, 128, None) == 0
    assert built_lib.afx_clip_kernel_variant(22050, 1024, 100, None) == 2      # any other hop <= n_fft / 2: run-time-hop variant
    assert built_lib.afx_clip_kernel_variant(16000, 512, 160, None) == 2
    assert built_lib.afx_clip_kernel_variant(22050, 1024, 600, None) == -1
    assert built_lib.afx_clip_kernel_variant(22050, 1000, 250, None) == -1
