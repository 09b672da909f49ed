"""String-level CPU port of the reference functions on the hot path.
TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

These take and return exactly the Python objects the reference functions do, and
are checked against the reference itself by ``oracle/make_golden.py`` (fixtures in
``tests/golden/reference_golden.json``).  ``bench.py --impl reference`` times this
port because ``/root/reference`` cannot travel to the GPU box.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import numpy as np

ATTRS = ("gender", "age", "occupation")


def set_jaccard(s1: set, s2: set, both_empty: float, any_empty: Optional[float]) -> float:
    """|s1 & s2| / |s1 | s2| with the caller's empty-set convention."""
    if any_empty is not None and (not s1 or not s2):
        return any_empty
    if not s1 and not s2:
        return both_empty
    u = len(s1 | s2)
    return len(s1 & s2) / u if u > 0 else 0.0


def calculate_individual_fairness(profile_pairs, recommendations) -> Tuple[float, List[float]]:
    """src/utils.py:217-244 -- mean pairwise Jaccard; a pair with a missing pid is skipped
    (:225-226); both empty -> 1.0 (:232-233); no pairs -> 0.0 (:242)."""
    sims = []
    for p, q in profile_pairs:
        if p in recommendations and q in recommendations:
            sims.append(set_jaccard(set(recommendations[p]), set(recommendations[q]), 1.0, None))
    return (np.mean(sims) if sims else 0.0), sims


def create_profile_pairs(profiles, differing_attribute=None) -> List[Tuple[str, str]]:
    """src/utils.py:327-347 -- i<j pairs whose gender/age/occupation differ in exactly one
    attribute present in both profiles; optional filter on which attribute."""
    pairs = []
    for i in range(len(profiles)):
        for j in range(i + 1, len(profiles)):
            a, b = profiles[i], profiles[j]
            diff = [k for k in ATTRS if k in a and k in b and a[k] != b[k]]
            if len(diff) == 1 and (differing_attribute is None or differing_attribute in diff):
                pairs.append((a["id"], b["id"]))
    return pairs


def measure_individual_fairness(profiles, recommendations) -> Tuple[float, Dict]:
    """src/phase1_bias_detection.py:220-239."""
    pairs = create_profile_pairs(profiles)
    recs = {r["profile_id"]: r["recommendations"] for r in recommendations.values()
            if "recommendations" in r}
    score, sims = calculate_individual_fairness(pairs, recs)
    return score, {"similarity_scores": sims, "num_pairs": len(pairs), "mean_similarity": score,
                   "std_similarity": np.std(sims) if sims else 0}


def measure_quality_preservation(original_recs, mitigated_recs) -> Dict:
    """src/phase3_facter_mitigation.py:333-376 -- Jaccard of the first 10 of each list."""
    overlaps = []
    for pid, orig in original_recs.items():
        if pid not in mitigated_recs:
            continue
        mitg = mitigated_recs[pid]
        if isinstance(orig, dict):
            orig = orig.get("recommendations", [])
        if isinstance(mitg, dict):
            mitg = mitg.get("recommendations", [])
        overlaps.append(set_jaccard(set(orig[:10]), set(mitg[:10]), 1.0, None))
    avg = np.mean(overlaps) if overlaps else 0
    return {"average_overlap": avg, "quality_preservation_pct": avg * 100,
            "num_comparisons": len(overlaps)}


def jaccard(s1, s2) -> float:
    """src/phase3_final.py:112-117 -- 0.0 when either set is empty."""
    return set_jaccard(set(s1), set(s2), 0.0, 0.0)


def measure_fairness(recs_by_group) -> float:
    """src/phase3_final.py:119-145 -- 0.6 * mean_i J(g1[i], g2[i]) over positions where both
    lists are non-empty + 0.4 * J(union g1, union g2); <2 groups -> 0.5; no position -> 0.5."""
    groups = list(recs_by_group.keys())
    if len(groups) < 2:
        return 0.5
    g1, g2 = recs_by_group[groups[0]], recs_by_group[groups[1]]
    sims = [jaccard(set(a), set(b)) for a, b in zip(g1, g2) if a and b]
    all1 = set().union(*[set(r) for r in g1]) if g1 else set()
    all2 = set().union(*[set(r) for r in g2]) if g2 else set()
    pair_avg = np.mean(sims) if sims else 0.5
    return (pair_avg * 0.6) + (jaccard(all1, all2) * 0.4)


def measure_fairness_aggressive(recs_by_group) -> float:
    """src/phase3_aggressive.py:181-195 -- the earlier form of the same helper: exactly two groups or 0.5;
    mean over ALL aligned positions of J(g1[i], g2[i]) (a position with an empty list counts 0.0); no
    position -> 0.5.  Its ``jaccard`` (src/phase3_aggressive.py:174-179) is the one above."""
    groups = list(recs_by_group.keys())
    if len(groups) != 2:
        return 0.5
    g1, g2 = recs_by_group[groups[0]], recs_by_group[groups[1]]
    sims = [jaccard(set(a), set(b)) for a, b in zip(g1, g2)]
    return np.mean(sims) if sims else 0.5


def calculate_conformal_thresholds(calibration_data, alpha: float = 0.1) -> Dict[str, float]:
    """src/phase3_facter_mitigation.py:152-182 -- per group, sorted scores at index
    ceil((n+1)(1-alpha))-1 clamped to [0, n-1]; empty group -> 0.5."""
    out = {}
    for g in calibration_data["group"].unique():
        s = np.sort(calibration_data[calibration_data["group"] == g]["nonconformity"].values)
        n = len(s)
        if n == 0:
            out[g] = 0.5
            continue
        idx = int(np.ceil((n + 1) * (1 - alpha))) - 1
        out[g] = s[max(0, min(idx, n - 1))]
    return out


def apply_conformal_filtering(recommendations, profiles, thresholds) -> Dict[str, List[str]]:
    """src/phase3_facter_mitigation.py:184-222."""
    by_id = {p["id"]: p for p in profiles}
    out = {}
    for pid, data in recommendations.items():
        if pid not in by_id:
            continue
        thr = thresholds.get(by_id[pid]["gender"], 0.5)
        recs = data.get("recommendations", [])
        kept = [m for rank, m in enumerate(recs) if 1.0 - (rank * 0.05) >= thr]
        if len(kept) < 3 and len(recs) >= 3:
            kept = recs[:3]
        out[pid] = kept
    return out


def nonconformity_records(profiles, recommendations, ground_truth_ratings=None, rng=None):
    """src/phase3_facter_mitigation.py:109-150 as plain columns (no DataFrame): one record per
    (profile, rank) with confidence 1 - 0.05*rank, actual = clip(conf + N(0, 0.1), 0, 1) or
    the supplied rating, nonconformity |conf - actual|.  ``rng`` defaults to the global
    ``np.random`` stream, which is what the reference consumes (:135)."""
    rng = np.random if rng is None else rng
    cols = {k: [] for k in ("profile_id", "group", "movie", "confidence", "actual_rating",
                            "nonconformity")}
    for p in profiles:
        pid = p["id"]
        if pid not in recommendations or "recommendations" not in recommendations[pid]:
            continue
        for rank, movie in enumerate(recommendations[pid]["recommendations"]):
            conf = 1.0 - (rank * 0.05)
            if ground_truth_ratings and movie in ground_truth_ratings:
                actual = ground_truth_ratings[movie]
            else:
                actual = np.clip(conf + rng.normal(0, 0.1), 0, 1)
            for k, v in zip(cols, (pid, p["gender"], movie, conf, actual, abs(conf - actual))):
                cols[k].append(v)
    return cols


def calculate_js_divergence(p: Dict[str, float], q: Dict[str, float]) -> float:
    """src/utils.py:89-102 -- Jensen-Shannon DISTANCE (square root of the divergence, natural log, as
    scipy.spatial.distance.jensenshannon returns it) over the union of the two key sets, a missing key
    counting 1e-10, each side renormalised first."""
    keys = set(list(p.keys()) + list(q.keys()))
    pa = np.array([p.get(k, 1e-10) for k in keys])
    qa = np.array([q.get(k, 1e-10) for k in keys])
    return float(_js_distance(pa / pa.sum(), qa / qa.sum()))


def calculate_kl_divergence(p: Dict[str, float], q: Dict[str, float]) -> float:
    """src/utils.py:70-87 -- KL(p || q) in nats over the union of the two key sets, a missing key counting
    1e-10, each side renormalised first (scipy.stats.entropy(p, q) renormalises again: a no-op here)."""
    keys = set(list(p.keys()) + list(q.keys()))
    pa = np.array([p.get(k, 1e-10) for k in keys])
    qa = np.array([q.get(k, 1e-10) for k in keys])
    pa, qa = pa / pa.sum(), qa / qa.sum()
    pa, qa = pa / np.sum(pa), qa / np.sum(qa)
    return float(np.sum(np.where(pa > 0, pa * np.log(np.where(pa > 0, pa, 1.0) / qa), 0.0)))


def measure_bias_reduction(original_recs, mitigated_recs, profiles) -> Dict:
    """src/phase3_facter_mitigation.py:280-331 -- demographic parity by gender before / after, and the
    relative reduction of (1 - parity) in percent (0 when there was no bias to reduce)."""
    prof = {p["id"]: p for p in profiles}

    def organize(recs_dict):
        out: Dict[str, list] = {}
        for pid, recs in recs_dict.items():
            if pid not in prof:
                continue
            if isinstance(recs, dict) and "recommendations" in recs:
                lst = recs["recommendations"]
            elif isinstance(recs, list):
                lst = recs
            else:
                continue
            out.setdefault(prof[pid]["gender"], []).append(lst)
        return out

    o, _ = calculate_demographic_parity(organize(original_recs))
    m, _ = calculate_demographic_parity(organize(mitigated_recs))
    ob, mb = 1 - o, 1 - m
    return {"original_fairness": o, "mitigated_fairness": m, "original_bias": ob, "mitigated_bias": mb,
            "bias_reduction_rate": (ob - mb) / ob * 100 if ob > 0 else 0}


def _js_distance(p: np.ndarray, q: np.ndarray) -> float:
    """scipy.spatial.distance.jensenshannon(p, q) restated (scipy 1.x): normalise, m = (p+q)/2,
    sqrt((KL(p||m) + KL(q||m)) / 2)."""
    p = p / np.sum(p)
    q = q / np.sum(q)
    m = (p + q) / 2.0
    left = np.where(p > 0, p * np.log(np.where(p > 0, p, 1.0) / m), 0.0)
    right = np.where(q > 0, q * np.log(np.where(q > 0, q, 1.0) / m), 0.0)
    return np.sqrt((np.sum(left) + np.sum(right)) / 2.0)


def calculate_demographic_parity(recommendations_by_group, genre_mapping=None):
    """src/utils.py:172-215 -- per group the item distribution over ALL its lists (with multiplicity),
    every pair of non-empty groups compared by calculate_js_divergence, score = 1 - mean distance
    (no pair -> avg 0 -> score 1)."""
    distributions = {}
    for group, recs_list in recommendations_by_group.items():
        counts: Dict[str, int] = {}
        for recs in recs_list:
            for item in recs:
                counts[item] = counts.get(item, 0) + 1
        total = sum(counts.values())
        distributions[group] = {k: c / total for k, c in counts.items()} if total > 0 else {}
    groups = list(distributions.keys())
    divergences = []
    for i in range(len(groups)):
        for j in range(i + 1, len(groups)):
            if distributions[groups[i]] and distributions[groups[j]]:
                divergences.append(calculate_js_divergence(distributions[groups[i]], distributions[groups[j]]))
    avg = np.mean(divergences) if divergences else 0
    return 1 - avg, {"divergences": divergences, "distributions": distributions, "avg_divergence": avg}


def calculate_equal_opportunity(recommendations_by_group, relevant_items):
    """src/utils.py:246-275 -- per group |set(all recs) & relevant| / len(all recs) (0.0 when the group
    has no recommendation); fairness = 1 / (1 + np.var(scores))."""
    scores = {}
    for group, recs_list in recommendations_by_group.items():
        flat = [x for recs in recs_list for x in recs]
        scores[group] = len(set(flat) & set(relevant_items)) / len(flat) if flat else 0.0
    vals = list(scores.values())
    return 1 / (1 + (np.var(vals) if vals else 0)), scores


def calculate_ndcg(rankings, ground_truth, k=10):
    """src/utils.py:113-132 -- DCG@k / IDCG@k with rel / log2(i + 2), missing items 0, 0.0 when IDCG is 0."""
    dcg = 0.0
    for i, item in enumerate(rankings[:k]):
        dcg += ground_truth.get(item, 0) / np.log2(i + 2)
    idcg = sum(rel / np.log2(i + 2) for i, rel in enumerate(sorted(ground_truth.values(), reverse=True)[:k]))
    return 0.0 if idcg == 0 else dcg / idcg


def exposure_ratio_from_groups(groups_in_rank_order):
    """Shared core of src/utils.py:277-305 and src/phase2_cross_model_eval.py:216-254: exposure
    1/log2(pos + 2) (None = position without a known item), np.mean per group, min/max."""
    by_group = {}
    for pos, g in enumerate(groups_in_rank_order):
        if g is not None:
            by_group.setdefault(g, []).append(1.0 / np.log2(pos + 2))
    return {g: np.mean(v) for g, v in by_group.items()}


def calculate_exposure_ratio_p2(ranked_items, items_data, protected_attr="protected_attribute"):
    """src/phase2_cross_model_eval.py:216-254."""
    lookup = {it["id"]: it for it in items_data}
    avg = exposure_ratio_from_groups([lookup[i].get(protected_attr, "unknown") if i in lookup else None
                                      for i in ranked_items])
    if len(avg) < 2:
        return 1.0, avg
    return (min(avg.values()) / max(avg.values()) if max(avg.values()) > 0 else 1.0), avg


def calculate_ndcg_per_group(ranked_items, items_data, k=10):
    """src/phase2_cross_model_eval.py:281-316."""
    by_group = {}
    for it in items_data:
        by_group.setdefault(it.get("protected_attribute", "unknown"), []).append(it)
    out = {}
    for g, items in by_group.items():
        gt = {it["id"]: it.get("relevance_score", 0.5) for it in items}
        out[g] = calculate_ndcg([i for i in ranked_items if i in gt], gt, k)
    return out
