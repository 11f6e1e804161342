"""Golden vectors for the DTW kernels, produced by the REFERENCE's own code (run in the build container only).

Imports /root/reference/05_dtw_alignment_experiment/dtw_alignment.py (its two third-party imports that are absent here,
fastdtw and pywt, are stubbed: neither is touched by the functions used) and runs, on seeded inputs,
  FastDTWAligner.compute_distance_matrix   (:532-557) with _apply_local_constraints (:559) replaced by the identity, so the
                                           output is the composite frame distance of :539-552
  FastDTWAligner._compute_accumulated_cost (:930-969)
and stores inputs and outputs in tests/golden/dtw_golden.npz.  Usage: python tests/golden/gen_dtw_golden.py
"""
import importlib.util
import os
import sys
import tempfile
import types
from pathlib import Path

import numpy as np

REF = "/root/reference/05_dtw_alignment_experiment/dtw_alignment.py"
OUT = Path(__file__).resolve().parent / "dtw_golden.npz"


def load_reference():
    for name in ("fastdtw", "pywt"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.fastdtw = lambda *a, **k: (_ for _ in ()).throw(RuntimeError("stub"))
            sys.modules[name] = m
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp())          # the module opens a log file in the working directory at import
    try:
        spec = importlib.util.spec_from_file_location("ref_dtw_alignment", REF)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        os.chdir(cwd)
    return mod


def main():
    mod = load_reference()
    al = mod.FastDTWAligner(".")
    al._apply_local_constraints = lambda D: D
    rng = np.random.default_rng(2024)
    out = {}
    cases = [(23, 31, 39), (40, 40, 39), (1, 7, 13), (9, 1, 13), (64, 50, 6)]
    for c, (n, m, dim) in enumerate(cases):
        x = rng.standard_normal((n, dim))
        y = 0.7 * rng.standard_normal((m, dim)) + 0.2
        if c == 1:
            y = x + 0.05 * rng.standard_normal((n, dim))      # near-identical sequences (teacher / student)
            y[3] = x[3]                                       # an exactly equal frame: euclidean distance 0
        D = al.compute_distance_matrix(x, y)
        A = al._compute_accumulated_cost(D)
        out[f"x{c}"], out[f"y{c}"], out[f"D{c}"], out[f"A{c}"] = x, y, D, A
    out["n_cases"] = np.asarray(len(cases))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
