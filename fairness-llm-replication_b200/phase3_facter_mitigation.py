"""Drop-in for the conformal-calibration functions of the reference's
``src/phase3_facter_mitigation.py`` (same names, parameters and return types), plus the
embedding form of the nonconformity score that BASELINE.json's north_star prescribes.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np
import pandas as pd
import torch

from . import ops, utils

_COLUMNS = ["profile_id", "group", "movie", "confidence", "actual_rating", "nonconformity"]


def generate_calibration_data(profiles: List[Dict], recommendations: Dict[str, Dict],
                              ground_truth_ratings: Optional[Dict] = None) -> pd.DataFrame:
    """One calibration record per (profile, rank) -- reference:
    src/phase3_facter_mitigation.py:109-150.

    confidence = 1 - 0.05*rank; actual = the supplied rating, else clip(confidence + N(0, 0.1),
    0, 1) drawn from the global ``np.random`` stream in record order (so a seeded caller gets the
    reference's numbers); nonconformity = |confidence - actual|.  This simulated scalar score is
    data generation, not scoring; the scored form is ``embedding_nonconformity`` below.
    """
    pids, groups, movies, ranks = [], [], [], []
    for profile in profiles:
        pid = profile["id"]
        if pid not in recommendations or "recommendations" not in recommendations[pid]:
            continue
        recs = recommendations[pid]["recommendations"]
        if not recs:
            continue
        group = profile["gender"]
        pids.extend([pid] * len(recs))
        groups.extend([group] * len(recs))
        movies.extend(recs)
        ranks.extend(range(len(recs)))
    if not pids:
        return pd.DataFrame([])
    confidence = 1.0 - (np.asarray(ranks, dtype=np.float64) * 0.05)
    if ground_truth_ratings:
        rated = np.fromiter((m in ground_truth_ratings for m in movies), dtype=bool, count=len(movies))
    else:
        rated = np.zeros(len(movies), dtype=bool)
    actual = np.empty(len(movies), dtype=np.float64)
    n_sim = int((~rated).sum())
    if n_sim:
        noise = np.random.normal(0, 0.1, size=n_sim)
        actual[~rated] = np.clip(confidence[~rated] + noise, 0, 1)
    if rated.any():
        actual[rated] = [ground_truth_ratings[m] for m, r in zip(movies, rated) if r]
    return pd.DataFrame({"profile_id": pids, "group": groups, "movie": movies, "confidence": confidence,
                         "actual_rating": actual, "nonconformity": np.abs(confidence - actual)},
                        columns=_COLUMNS)


def embedding_nonconformity(anchor: torch.Tensor, counterfactual: torch.Tensor) -> torch.Tensor:
    """s = max(0, 1 - cos(anchor[u], counterfactual[u, v])) on the GPU (fp64 accumulation of fp32
    products, one fp32 rounding).  anchor float32 [U, D]; counterfactual [U, V, D] or [U, D]."""
    return ops.cosine_nonconformity(anchor, counterfactual)


def calculate_conformal_thresholds(calibration_data: pd.DataFrame, alpha: float = 0.1) -> Dict[str, float]:
    """Group-specific conformal thresholds -- reference: src/phase3_facter_mitigation.py:152-182.

    Per group the score at sorted index ceil((n+1)(1-alpha)) - 1, clamped to [0, n-1]; the
    selected element is found by an exact radix select over the float64 bit patterns, so it is
    the very element ``np.sort(scores)[index]`` returns.
    """
    thresholds: Dict[str, float] = {}
    if len(calibration_data) == 0:
        calibration_data["group"]           # same KeyError as the reference on a column-less frame
        return thresholds
    codes, uniques = pd.factorize(calibration_data["group"], use_na_sentinel=True)
    scores = np.array(calibration_data["nonconformity"].values, dtype=np.float64)      # own, writable copy
    dev = utils._device()
    t_scores = torch.from_numpy(scores).to(dev)
    names = list(uniques)
    for base in range(0, len(names), 255):
        chunk = names[base:base + 255]
        local = codes - base
        local = np.where((local >= 0) & (local < len(chunk)), local, 255).astype(np.uint8)
        thr, _ = ops.radix_select(t_scores, alpha=alpha, groups=torch.from_numpy(local).to(dev),
                                  n_groups=len(chunk))
        for name, value in zip(chunk, thr.cpu().numpy()):
            thresholds[name] = value
    if (codes < 0).any():
        # a NaN group label equals nothing, so the reference sees an empty group there (:169-171)
        thresholds[calibration_data["group"][codes < 0].iloc[0]] = 0.5
    return thresholds


def apply_conformal_filtering(recommendations: Dict[str, Dict], profiles: List[Dict],
                              thresholds: Dict[str, float]) -> Dict[str, List[str]]:
    """Keep the ranks whose simulated confidence 1 - 0.05*rank reaches the group threshold, falling
    back to the first three -- reference: src/phase3_facter_mitigation.py:184-222.  The kept set
    is a prefix, so the GPU returns one length per profile."""
    profile_dict = {p["id"]: p for p in profiles}
    pids, recs_list, thr_list = [], [], []
    for pid, rec_data in recommendations.items():
        if pid not in profile_dict:
            continue
        group = profile_dict[pid]["gender"]
        pids.append(pid)
        thr_list.append(float(thresholds.get(group, 0.5)))
        recs_list.append(rec_data.get("recommendations", []))
    if not pids:
        return {}
    dev = utils._device()
    lens = torch.tensor([len(r) for r in recs_list], dtype=torch.int32, device=dev)
    uniq, inverse = np.unique(np.asarray(thr_list, dtype=np.float64), return_inverse=True)
    if len(uniq) <= 255:
        keep = ops.conformal_keep_len(lens, torch.from_numpy(uniq).to(dev),
                                      torch.from_numpy(inverse.astype(np.uint8)).to(dev))
    else:                                              # one threshold per profile
        keep = torch.empty_like(lens)
        thr_all = torch.tensor(thr_list, dtype=torch.float64, device=dev)
        for i in range(0, len(pids), 255):
            n = min(255, len(pids) - i)
            keep[i:i + n] = ops.conformal_keep_len(lens[i:i + n].contiguous(), thr_all[i:i + n].contiguous(),
                                                   torch.arange(n, dtype=torch.uint8, device=dev))
    keep = keep.cpu().tolist()
    return {pid: recs[:k] for pid, recs, k in zip(pids, recs_list, keep)}


def measure_quality_preservation(original_recs: Dict, mitigated_recs: Dict) -> Dict:
    """Mean Jaccard of the first 10 original vs mitigated recommendations -- reference:
    src/phase3_facter_mitigation.py:333-376."""
    a, b = [], []
    for pid in original_recs.keys():
        if pid not in mitigated_recs:
            continue
        orig, mitg = original_recs[pid], mitigated_recs[pid]
        if isinstance(orig, dict):
            orig = orig.get("recommendations", [])
        if isinstance(mitg, dict):
            mitg = mitg.get("recommendations", [])
        a.append(orig[:10])
        b.append(mitg[:10])
    inter, union = utils.pair_set_counts(a, b)
    overlaps = utils.jaccard_from_counts(inter, union, 1.0)
    avg_overlap = np.mean(overlaps) if overlaps else 0
    return {"average_overlap": avg_overlap, "quality_preservation_pct": avg_overlap * 100,
            "num_comparisons": len(overlaps)}


def measure_bias_reduction(original_recs: Dict, mitigated_recs: Dict, profiles: List[Dict]) -> Dict:
    """Demographic parity by gender before and after mitigation and the relative bias reduction --
    reference: src/phase3_facter_mitigation.py:280-331 (the phase-3 caller of N1).  Lists are grouped by the
    profile's gender (ids without a profile and values that are neither a list nor a dict with
    'recommendations' are skipped, as there); the two parity scores come from the GPU histogram / distance
    kernels behind ``utils.calculate_demographic_parity``."""
    profile_dict = {p["id"]: p for p in profiles}

    def organize_recs(recs_dict):
        by_gender: Dict[str, List[List[str]]] = {}
        for pid, recs in recs_dict.items():
            if pid not in profile_dict:
                continue
            gender = profile_dict[pid]["gender"]
            if isinstance(recs, dict) and "recommendations" in recs:
                rec_list = recs["recommendations"]
            elif isinstance(recs, list):
                rec_list = recs
            else:
                continue
            by_gender.setdefault(gender, []).append(rec_list)
        return by_gender

    original_dp, _ = utils.calculate_demographic_parity(organize_recs(original_recs))
    mitigated_dp, _ = utils.calculate_demographic_parity(organize_recs(mitigated_recs))
    original_bias = 1 - original_dp
    mitigated_bias = 1 - mitigated_dp
    bias_reduction_rate = (original_bias - mitigated_bias) / original_bias * 100 if original_bias > 0 else 0
    return {"original_fairness": original_dp, "mitigated_fairness": mitigated_dp, "original_bias": original_bias,
            "mitigated_bias": mitigated_bias, "bias_reduction_rate": bias_reduction_rate}
