"""CPU restatement of the two DTW stages downstream of the feature extractor (TEST INFRASTRUCTURE ONLY: imported by
tests/, never by the product).

Follows /root/reference/05_dtw_alignment_experiment/dtw_alignment.py:
  frame_distance_matrix  :532-552  composite frame distance (euclidean / cosine similarity / Pearson correlation with
                                   similarity-dependent weights), WITHOUT the positional penalties of
                                   _apply_local_constraints (:559-668: an in-place serial sweep whose np.mean over the
                                   whole matrix changes after every cell, and whose :570 broadcast only works for n == m)
  accumulated_cost       :930-969  weighted DP: borders with the 1 + 0.1 k penalty, interior
                                   D + min(1.2 up, 0.8 diagonal, 1.2 left)
Parity of this restatement is PINNED: tests/golden/dtw_golden.npz holds outputs of the reference's own functions
(tests/golden/gen_dtw_golden.py), checked in tests/test_dtw.py.
"""
from __future__ import annotations

import numpy as np


def frame_distance_matrix(x, y):
    """dtw_alignment.py:539-552 for every (i, j); x [n, dim], y [m, dim], float64."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    eucl = np.empty((x.shape[0], y.shape[0]))
    for i0 in range(0, x.shape[0], 128):                                                  # row blocks bound the temporary
        diff = x[i0:i0 + 128, None, :] - y[None, :, :]
        eucl[i0:i0 + 128] = np.sqrt(np.sum(diff ** 2, axis=2))                            # :542
    cos = (x @ y.T) / (np.linalg.norm(x, axis=1)[:, None] * np.linalg.norm(y, axis=1)[None, :] + 1e-8)   # :543
    xc = x - x.mean(axis=1, keepdims=True)
    yc = y - y.mean(axis=1, keepdims=True)
    with np.errstate(invalid="ignore", divide="ignore"):
        # np.corrcoef (:544): covariance divided by the two standard deviations one after the other, clipped to [-1, 1]
        corr = (xc @ yc.T) / np.sqrt(np.sum(xc * xc, axis=1))[:, None] / np.sqrt(np.sum(yc * yc, axis=1))[None, :]
    corr = np.clip(corr, -1.0, 1.0)
    w2 = 0.25 * (1 + cos)                                                                 # :548
    w3 = 0.15 * (1 + np.abs(corr))                                                        # :549
    return 0.6 * eucl - w2 * cos - w3 * corr                                              # :552


def accumulated_cost(D):
    """dtw_alignment.py:930-969, same operation order (bit-exact against the reference)."""
    D = np.asarray(D, dtype=np.float64)
    n, m = D.shape
    A = np.zeros_like(D)
    A[0, 0] = D[0, 0]
    for i in range(1, n):
        A[i, 0] = A[i - 1, 0] + D[i, 0] * (1 + 0.1 * i)
    for j in range(1, m):
        A[0, j] = A[0, j - 1] + D[0, j] * (1 + 0.1 * j)
    for i in range(1, n):
        up = A[i - 1]
        row = A[i]
        d = D[i]
        left = row[0]
        for j in range(1, m):
            left = d[j] + min(up[j] * 1.2, up[j - 1] * 0.8, left * 1.2)
            row[j] = left
    return A
