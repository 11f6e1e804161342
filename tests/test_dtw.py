"""DTW stage.  Parity here is PINNED: tests/golden/dtw_golden.npz was produced by the
reference's own FastDTWAligner.compute_distance_matrix / _compute_accumulated_cost (tests/golden/gen_dtw_golden.py)."""
from pathlib import Path

import numpy as np
import pytest

GOLDEN = Path(__file__).resolve().parent / "golden" / "dtw_golden.npz"
DIST_ATOL = 1e-12      # float64 sums over <= 39 feature dimensions in a different order than numpy / BLAS


def _cases():
    g = np.load(GOLDEN)
    return [(g[f"x{c}"], g[f"y{c}"], g[f"D{c}"], g[f"A{c}"]) for c in range(int(g["n_cases"]))]


def test_oracle_matches_reference_golden():
    from oracle import dtw
    for x, y, D, A in _cases():
        np.testing.assert_allclose(dtw.frame_distance_matrix(x, y), D, rtol=0, atol=DIST_ATOL)
        np.testing.assert_array_equal(dtw.accumulated_cost(D), A)          # same operation order: bit-exact


@pytest.mark.gpu
def test_cuda_matches_reference_golden(torch_cuda):
    from audio_feature_extraction_b200.dtw import FastDTWAligner
    al = FastDTWAligner()
    cases = _cases()
    Ds, As, dist = [], [], []
    for dim in sorted({c[0].shape[1] for c in cases}):                       # one batch per feature dimension
        group = [c for c in cases if c[0].shape[1] == dim]
        d_, a_, t_ = al.cost_matrices_batch([(x, y) for x, y, _, _ in group])
        Ds += d_; As += a_; dist += list(t_)
    cases = sorted(cases, key=lambda c: c[0].shape[1])                       # stable: the order of the batches above
    for (x, y, D, A), gD, gA, d in zip(cases, Ds, As, dist):
        np.testing.assert_allclose(gD, D, rtol=0, atol=DIST_ATOL)
        np.testing.assert_allclose(gA, A, rtol=1e-12, atol=1e-10)            # on the device's own D (differs in the last bits)
        assert d == gA[-1, -1]
    # accumulated cost on the reference's D: bit-exact, through the reference's method names too
    got = al.accumulated_cost_batch([D for _, _, D, _ in cases])
    for (_, _, D, A), g in zip(cases, got):
        np.testing.assert_array_equal(g, A)
    np.testing.assert_array_equal(al._compute_accumulated_cost(cases[0][2]), cases[0][3])
    np.testing.assert_allclose(al.compute_distance_matrix(cases[0][0], cases[0][1]), cases[0][2], rtol=0, atol=DIST_ATOL)


@pytest.mark.gpu
def test_cuda_matches_oracle_on_feature_sized_pairs(torch_cuda):
    """Ragged batch at the size of real clips (39-dim frames, up to 1300 frames: several row strips of the wavefront
    kernel), a constant frame (NaN correlation, as np.corrcoef gives) and an empty pair."""
    from audio_feature_extraction_b200.dtw import FastDTWAligner
    from oracle import dtw
    rng = np.random.default_rng(11)
    shapes = [(862, 870), (431, 300), (1300, 1100), (1025, 64), (2, 1500), (257, 258), (0, 5)]
    pairs = []
    for n, m in shapes:
        x = rng.standard_normal((n, 39)); y = rng.standard_normal((m, 39)) * 0.5 + 0.1
        pairs.append((x, y))
    pairs[1][0][7] = 3.0                                                   # constant frame
    al = FastDTWAligner()
    Ds, As, dist = al.cost_matrices_batch(pairs)
    for (x, y), gD, gA in zip(pairs, Ds, As):
        if x.shape[0] == 0:
            assert gD.shape == (0, 5) and gA.shape == (0, 5)
            continue
        with np.errstate(invalid="ignore", divide="ignore"):
            rD = dtw.frame_distance_matrix(x, y)
        assert np.array_equal(np.isnan(gD), np.isnan(rD))
        np.testing.assert_allclose(gD, rD, rtol=0, atol=DIST_ATOL, equal_nan=True)
        np.testing.assert_array_equal(gA, dtw.accumulated_cost(gD))         # bit-exact on the same D, NaN rows included
