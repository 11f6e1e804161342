"""Batched DTW cost matrices on the device: the compute-heavy stage downstream of the extractor
(/root/reference/05_dtw_alignment_experiment/dtw_alignment.py, class FastDTWAligner).

`compute_distance_matrix` (:532-552, the composite frame distance, without the serial positional penalties of
`_apply_local_constraints` :559-668) and `_compute_accumulated_cost` (:930-969) keep the reference's names and argument
meaning for ONE pair; `cost_matrices_batch` runs many (teacher, student) pairs in two launches.  float64, as numpy
computes them.  No CPU fallback: the arithmetic is in csrc/afx_dtw.cuh behind afx_dtw_frame_distance /
afx_dtw_accumulated_cost.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

from . import _native as N
from .core.feature_extractor import AudioFeatureExtractor


class FastDTWAligner:
    def __init__(self, feature_dir: str = ".", device: int = 0):
        self.feature_dir = feature_dir          # reference :83-84
        self.device = device
        self._ex = AudioFeatureExtractor(device=device)   # owns the native handle (device, header staging)

    # ------------------------------------------------------------------ batch entry point
    def cost_matrices_batch(self, pairs: Sequence[Tuple[np.ndarray, np.ndarray]], accumulate: bool = True):
        """pairs: [(x [n, dim], y [m, dim]), ...] with one common dim.  Returns (D list, A list or None, distances):
        D[p] = frame distances [n, m], A[p] = accumulated cost [n, m], distances[p] = A[p][-1, -1] (align_features :847)."""
        import torch

        if len(pairs) == 0:
            return [], ([] if accumulate else None), np.zeros(0)
        L = N.lib()
        h = self._ex._h()
        xs = [np.ascontiguousarray(p[0], dtype=np.float64) for p in pairs]
        ys = [np.ascontiguousarray(p[1], dtype=np.float64) for p in pairs]
        dim = int(xs[0].shape[1]) if xs[0].ndim == 2 else -1
        for x, y in zip(xs, ys):
            if x.ndim != 2 or y.ndim != 2 or x.shape[1] != dim or y.shape[1] != dim:
                raise ValueError("every sequence must be [frames, dim] with one common dim")
        n = np.asarray([x.shape[0] for x in xs], dtype=np.int32)
        m = np.asarray([y.shape[0] for y in ys], dtype=np.int32)
        x_off = np.concatenate([[0], np.cumsum(n.astype(np.int64) * dim)]).astype(np.int64)
        y_off = np.concatenate([[0], np.cumsum(m.astype(np.int64) * dim)]).astype(np.int64)
        D_off = np.concatenate([[0], np.cumsum(n.astype(np.int64) * m.astype(np.int64))]).astype(np.int64)
        dev = torch.device("cuda", self.device)
        hx = torch.from_numpy(np.concatenate([x.reshape(-1) for x in xs] + [np.zeros(1)]))
        hy = torch.from_numpy(np.concatenate([y.reshape(-1) for y in ys] + [np.zeros(1)]))
        dx, dy = hx.to(dev), hy.to(dev)
        dD = torch.empty(int(D_off[-1]) + 1, dtype=torch.float64, device=dev)
        s = torch.cuda.current_stream(dev).cuda_stream
        N.check(L.afx_dtw_frame_distance(h, dx.data_ptr(), x_off.ctypes.data, n.ctypes.data, dy.data_ptr(), y_off.ctypes.data,
                                         m.ctypes.data, dim, len(pairs), dD.data_ptr(), D_off.ctypes.data, s))
        dA = None
        if accumulate:
            dA = torch.empty_like(dD)
            N.check(L.afx_dtw_accumulated_cost(h, dD.data_ptr(), D_off.ctypes.data, n.ctypes.data, m.ctypes.data, len(pairs),
                                               dA.data_ptr(), s))
        hD = dD.cpu().numpy()
        Ds = [hD[D_off[p]:D_off[p + 1]].reshape(int(n[p]), int(m[p])) for p in range(len(pairs))]
        if not accumulate:
            return Ds, None, None
        hA = dA.cpu().numpy()
        As = [hA[D_off[p]:D_off[p + 1]].reshape(int(n[p]), int(m[p])) for p in range(len(pairs))]
        dist = np.asarray([a[-1, -1] if a.size else np.nan for a in As])
        return Ds, As, dist

    def accumulated_cost_batch(self, mats: Sequence[np.ndarray]) -> List[np.ndarray]:
        """_compute_accumulated_cost for a list of distance matrices (any origin), one launch."""
        import torch

        if len(mats) == 0:
            return []
        L = N.lib()
        h = self._ex._h()
        mats = [np.ascontiguousarray(d, dtype=np.float64) for d in mats]
        n = np.asarray([d.shape[0] for d in mats], dtype=np.int32)
        m = np.asarray([d.shape[1] for d in mats], dtype=np.int32)
        D_off = np.concatenate([[0], np.cumsum(n.astype(np.int64) * m.astype(np.int64))]).astype(np.int64)
        dev = torch.device("cuda", self.device)
        dD = torch.from_numpy(np.concatenate([d.reshape(-1) for d in mats] + [np.zeros(1)])).to(dev)
        dA = torch.empty_like(dD)
        N.check(L.afx_dtw_accumulated_cost(h, dD.data_ptr(), D_off.ctypes.data, n.ctypes.data, m.ctypes.data, len(mats),
                                           dA.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
        hA = dA.cpu().numpy()
        return [hA[D_off[p]:D_off[p + 1]].reshape(int(n[p]), int(m[p])) for p in range(len(mats))]

    # ------------------------------------------------------------------ the reference's per-pair methods
    def compute_distance_matrix(self, x, y) -> np.ndarray:
        """reference :532-552 (frame distances of one pair)."""
        return self.cost_matrices_batch([(x, y)], accumulate=False)[0][0]

    def _compute_accumulated_cost(self, distance_matrix) -> np.ndarray:
        """reference :930-969."""
        return self.accumulated_cost_batch([distance_matrix])[0]
