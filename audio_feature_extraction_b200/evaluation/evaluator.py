"""FeatureEvaluator: consumer of the feature dicts (host side, not on the GPU hot path).

Written from the behaviour documented in the reference's docs/api.md and
audio_feature_extraction_toolkit/evaluation/evaluator.py (:16 statistics, :57 quality rates,
:101 JSON + CSV report, :149 distributions) so that `from audio_feature_extraction_b200 import
FeatureEvaluator` keeps a caller such as examples/basic_usage.py working unchanged.  It is outside
the accelerated path; numpy on the host is the right tool for a few hundred scalars.
"""
from __future__ import annotations

import csv
import json
import logging
from pathlib import Path
from typing import Any, Dict, Iterable, List

import numpy as np

_MFCC_STD_THRESHOLD = 0.5     # reference :84
_ENERGY_STD_THRESHOLD = 0.1   # reference :92


def _flatten(features_list: Iterable[Dict[str, Any]], key: str) -> List[float]:
    vals: List[float] = []
    for feats in features_list:
        if key in feats:
            v = feats[key]
            vals.extend(v) if isinstance(v, list) else vals.append(v)
    return vals


def _describe(values) -> Dict[str, Any]:
    arr = np.asarray(values)
    q25, q50, q75 = (float(np.percentile(arr, p)) for p in (25, 50, 75))
    return {"mean": float(arr.mean()), "std": float(arr.std()), "percentiles": {"25": q25, "50": q50, "75": q75}}


class FeatureEvaluator:
    def __init__(self):
        logging.basicConfig(level=logging.INFO)
        self.logger = logging.getLogger(__name__)

    def calculate_feature_statistics(self, features_list: List[Dict[str, Any]]) -> Dict[str, Any]:
        """{key}_min/_max/_mean/_std over all files for every key but 'file_path'; list values are flattened."""
        if not features_list:
            return {}
        keys = set().union(*(f.keys() for f in features_list)) - {"file_path"}
        stats: Dict[str, Any] = {}
        for key in keys:
            vals = _flatten(features_list, key)
            if not vals:
                continue
            arr = np.asarray(vals)
            for suffix, fn in (("min", np.min), ("max", np.max), ("mean", np.mean), ("std", np.std)):
                stats[f"{key}_{suffix}"] = float(fn(arr))
        return stats

    def evaluate_feature_quality(self, features_list: List[Dict[str, Any]]) -> Dict[str, float]:
        if not features_list:
            return {}
        n = len(features_list)
        f0_rate = sum(f.get("f0_quality", 0) for f in features_list) / n * 100
        mfcc_ok = sum(1 for f in features_list if np.mean(f.get("mfcc_std", [1.0])) < _MFCC_STD_THRESHOLD)
        energy_ok = sum(1 for f in features_list if f.get("energy_std", 1.0) < _ENERGY_STD_THRESHOLD)
        return {
            "total_files": n,
            "feature_integrity_rate": 100.0,
            "f0_quality_rate": f0_rate,
            "mfcc_stability_rate": mfcc_ok / n * 100,
            "energy_stability_rate": energy_ok / n * 100,
        }

    def generate_evaluation_report(self, features_list: List[Dict[str, Any]],
                                   output_dir: str = "feature_evaluation") -> Dict[str, Any]:
        try:
            out = Path(output_dir)
            out.mkdir(parents=True, exist_ok=True)
            quality = self.evaluate_feature_quality(features_list)
            report = {
                "statistics": self.calculate_feature_statistics(features_list),
                "quality_metrics": quality,
                "features_list": features_list,
            }
            with open(out / "evaluation_detailed.json", "w", encoding="utf-8") as f:
                json.dump(report, f, indent=2, ensure_ascii=False)
            with open(out / "evaluation_summary.csv", "w", encoding="utf-8", newline="") as f:
                w = csv.writer(f)
                w.writerow(["Metric", "Value"])
                w.writerows(quality.items())
            self.logger.info("評估報告生成完成")
            return report
        except Exception as e:
            self.logger.error(f"生成評估報告失敗: {e}")
            raise

    def analyze_feature_distribution(self, features_list: List[Dict[str, Any]]) -> Dict[str, Any]:
        if not features_list:
            return {}
        dist: Dict[str, Any] = {}
        f0 = [f.get("f0_mean", 0) for f in features_list if f.get("f0_mean", 0) > 0]
        if f0:
            dist["f0_distribution"] = _describe(f0)
        mfcc = _flatten(features_list, "mfcc_mean")
        if mfcc:
            dist["mfcc_distribution"] = _describe(mfcc)
        energy = [f.get("energy_mean", 0) for f in features_list]
        if energy:
            dist["energy_distribution"] = _describe(energy)
        return dist
