"""Device timing of the DTW kernels on P pairs of 39-dim sequences resident in HBM,
next to the CPU oracle on a bounded sample.  One JSON line per kernel.  Usage: python tools/bench_dtw.py [P] [frames]"""
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def main():
    import torch
    from audio_feature_extraction_b200 import _native as N
    from audio_feature_extraction_b200.dtw import FastDTWAligner
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 862
    dim = 39
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    hbm = float(peaks.get("hbm_gbs", 6541.5))
    al = FastDTWAligner()
    L, h = N.lib(), al._ex._h()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(0)
    x = torch.randn(P * T * dim + 1, dtype=torch.float64, device=dev, generator=g)
    y = torch.randn(P * T * dim + 1, dtype=torch.float64, device=dev, generator=g) * 0.5 + 0.1
    n = np.full(P, T, np.int32); m = np.full(P, T, np.int32)
    xo = (np.arange(P, dtype=np.int64) * T * dim); Do = (np.arange(P, dtype=np.int64) * T * T)
    D = torch.empty(P * T * T + 1, dtype=torch.float64, device=dev)
    A = torch.empty_like(D)
    s = torch.cuda.current_stream(dev).cuda_stream

    def dist():
        N.check(L.afx_dtw_frame_distance(h, x.data_ptr(), xo.ctypes.data, n.ctypes.data, y.data_ptr(), xo.ctypes.data,
                                         m.ctypes.data, dim, P, D.data_ptr(), Do.ctypes.data, s))

    def acc():
        N.check(L.afx_dtw_accumulated_cost(h, D.data_ptr(), Do.ctypes.data, n.ctypes.data, m.ctypes.data, P, A.data_ptr(), s))

    out = []
    for name, fn, bytes_per_pair in (("dtw_distance_kernel", dist, 2 * T * dim * 8 + T * T * 8),
                                     ("dtw_accumulate_kernel", acc, 2 * T * T * 8)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        out.append({"kernel": name, "pairs": P, "frames": T, "dim": dim, "ms": ms, "pairs_per_s": P / ms * 1e3,
                    "gcells_per_s": P * T * T / ms / 1e6,
                    "roofline": {"bound": "hbm", "achieved": P * bytes_per_pair / ms / 1e6, "peak": hbm, "unit": "GB/s",
                                 "frac": P * bytes_per_pair / ms / 1e6 / hbm}})
    # CPU oracle on a bounded sample (numpy distance; the DP is the reference's own pure-Python double loop)
    from oracle import dtw
    xs, ys = x[:T * dim].cpu().numpy().reshape(T, dim), y[:T * dim].cpu().numpy().reshape(T, dim)
    t0 = time.perf_counter(); Dc = dtw.frame_distance_matrix(xs, ys); t1 = time.perf_counter()
    Ac = dtw.accumulated_cost(Dc); t2 = time.perf_counter()
    out[0]["cpu_port_pairs_per_s"] = 1.0 / (t1 - t0)
    out[1]["cpu_port_pairs_per_s"] = 1.0 / (t2 - t1)
    got = A[:T * T].cpu().numpy().reshape(T, T)
    out[1]["matches_oracle_on_own_D"] = bool(np.array_equal(got, dtw.accumulated_cost(D[:T * T].cpu().numpy().reshape(T, T))))
    for o in out:
        print(json.dumps(o))


if __name__ == "__main__":
    main()
