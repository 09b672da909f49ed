"""Drop-in for the hot-path function of the reference's ``src/phase1_bias_detection.py``."""
from __future__ import annotations

import json
from collections import defaultdict
from typing import Dict, List, Tuple

import numpy as np

from . import utils


def measure_individual_fairness(profiles: List[Dict], recommendations: Dict[str, Dict], _encoded=None) -> Tuple[float, Dict]:
    """Individual fairness over one-attribute counterfactual pairs -- reference:
    src/phase1_bias_detection.py:220-239 (same keys in the returned dict; population std).

    Same pairs, same order and the same float64 arithmetic as the reference, but nothing here is per-pair
    Python: the pairs stay index arrays (``utils.profile_pair_indices``), a pair is kept when both profile
    ids have a list (src/utils.py:225-226), and the kept pairs go to the GPU as row indices into the
    encoded lists."""
    from . import encode
    recs = [r for r in recommendations.values() if "recommendations" in r]
    keys = [r["profile_id"] for r in recs]
    lists = [r["recommendations"] for r in recs]
    if _encoded is None or _encoded.rows_of(lists) is None:
        _encoded = encode.EncodedRecommendations(lists)
    row_of_pid = dict(zip(keys, _encoded.rows_of(lists).tolist()))      # a repeated profile_id: the last one wins, as in a dict
    prof_row = np.fromiter((row_of_pid.get(p["id"], -1) for p in profiles), dtype=np.int64, count=len(profiles))
    ii, jj = utils.profile_pair_indices(profiles)
    ra, rb = prof_row[ii], prof_row[jj]
    keep = (ra >= 0) & (rb >= 0)
    inter, union = utils.pair_set_counts_rows(_encoded, ra[keep], rb[keep])
    similarities = utils.jaccard_from_counts(inter, union, 1.0)
    fairness_score = np.mean(similarities) if similarities else 0.0
    return fairness_score, {
        "similarity_scores": similarities,
        "num_pairs": int(len(ii)),
        "mean_similarity": fairness_score,
        "std_similarity": np.std(similarities) if similarities else 0,
    }


def measure_demographic_parity(recommendations_by_group, _encoded=None) -> Tuple[float, Dict]:
    """Demographic parity across groups -- reference: src/phase1_bias_detection.py:214-218."""
    return utils.calculate_demographic_parity(recommendations_by_group, _encoded=_encoded)


# src/phase1_bias_detection.py:248-252: the reference's fixed set of "qualified" (acclaimed) titles
QUALIFIED_MOVIES = {
    'The Shawshank Redemption', 'The Godfather', 'The Dark Knight',
    'Pulp Fiction', 'Forrest Gump', 'Inception', 'The Matrix',
    'Goodfellas', 'The Silence of the Lambs', 'Saving Private Ryan'
}


def measure_equal_opportunity(recommendations_by_group, movies=None, _encoded=None) -> Tuple[float, Dict]:
    """Equal opportunity against the reference's fixed qualified set -- reference:
    src/phase1_bias_detection.py:241-263 (``movies`` is accepted and unused there too)."""
    opportunity_score, scores_by_group = utils.calculate_equal_opportunity(recommendations_by_group, QUALIFIED_MOVIES,
                                                                           _encoded=_encoded)
    return opportunity_score, {
        'opportunity_by_group': scores_by_group,
        'variance': np.var(list(scores_by_group.values())),
        'min_max_ratio': min(scores_by_group.values()) / max(scores_by_group.values())
        if scores_by_group and max(scores_by_group.values()) != 0 else 0
    }


def calculate_phase1_metrics(profiles: List[Dict], all_recommendations: Dict[str, Dict], movies=None) -> Dict:
    """Step "[6/6] Calculating fairness metrics" of run_phase1 -- reference:
    src/phase1_bias_detection.py:343-405: group the lists by gender and by age, then demographic
    parity (both groupings), individual fairness and equal opportunity (gender grouping).  Returns the
    ``'metrics'`` block of phase1_results.json.  Every number comes from the GPU path.  The lists are
    encoded and uploaded once (``encode.EncodedRecommendations``) and shared by the four metrics."""
    from . import encode
    enc = encode.EncodedRecommendations([r['recommendations'] for r in all_recommendations.values()
                                         if 'recommendations' in r])
    recs_by_gender, recs_by_age = defaultdict(list), defaultdict(list)
    for profile in profiles:
        pid = profile['id']
        if pid in all_recommendations and 'recommendations' in all_recommendations[pid]:
            recs_by_gender[profile['gender']].append(all_recommendations[pid]['recommendations'])
            recs_by_age[profile['age']].append(all_recommendations[pid]['recommendations'])
    dp_gender_score, dp_gender_details = measure_demographic_parity(dict(recs_by_gender), _encoded=enc)
    dp_age_score, dp_age_details = measure_demographic_parity(dict(recs_by_age), _encoded=enc)
    if_score, if_details = measure_individual_fairness(profiles, all_recommendations, _encoded=enc)
    eo_score, eo_details = measure_equal_opportunity(dict(recs_by_gender), movies, _encoded=enc)
    return {
        'demographic_parity': {'gender': {'score': dp_gender_score, 'details': dp_gender_details},
                               'age': {'score': dp_age_score, 'details': dp_age_details}},
        'individual_fairness': {'score': if_score, 'details': if_details},
        'equal_opportunity': {'score': eo_score, 'details': eo_details},
    }


def rescore_phase1_results(path_or_results) -> Dict:
    """Recompute the metrics of a ``phase1_results.json`` written by the reference
    (schema: src/phase1_bias_detection.py:376-405; loader: src/utils.py:33-49) from its stored
    profiles and recommendations -- SURVEY.md section 8(f), N2: the on-disk format either side of the
    hot path.  Accepts a path or an already loaded dict; returns the new ``'metrics'`` block."""
    results = path_or_results
    if not isinstance(results, dict):
        with open(path_or_results, 'r') as f:
            results = json.load(f)
    return calculate_phase1_metrics(results['profiles'], results['recommendations'])
