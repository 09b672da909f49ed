"""Drop-in for the hot-path function of the reference's ``src/phase1_bias_detection.py``."""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np

from . import utils


def measure_individual_fairness(profiles: List[Dict], recommendations: Dict[str, Dict]) -> Tuple[float, Dict]:
    """Individual fairness over one-attribute counterfactual pairs -- reference:
    src/phase1_bias_detection.py:220-239 (same keys in the returned dict; population std)."""
    pairs = utils.create_profile_pairs(profiles)
    recs_dict = {r["profile_id"]: r["recommendations"]
                 for r in recommendations.values() if "recommendations" in r}
    fairness_score, similarities = utils.calculate_individual_fairness(pairs, recs_dict)
    return fairness_score, {
        "similarity_scores": similarities,
        "num_pairs": len(pairs),
        "mean_similarity": fairness_score,
        "std_similarity": np.std(similarities) if similarities else 0,
    }


def measure_demographic_parity(recommendations_by_group) -> Tuple[float, Dict]:
    """Demographic parity across groups -- reference: src/phase1_bias_detection.py:214-218."""
    return utils.calculate_demographic_parity(recommendations_by_group)


# src/phase1_bias_detection.py:248-252: the reference's fixed set of "qualified" (acclaimed) titles
QUALIFIED_MOVIES = {
    'The Shawshank Redemption', 'The Godfather', 'The Dark Knight',
    'Pulp Fiction', 'Forrest Gump', 'Inception', 'The Matrix',
    'Goodfellas', 'The Silence of the Lambs', 'Saving Private Ryan'
}


def measure_equal_opportunity(recommendations_by_group, movies=None) -> Tuple[float, Dict]:
    """Equal opportunity against the reference's fixed qualified set -- reference:
    src/phase1_bias_detection.py:241-263 (``movies`` is accepted and unused there too)."""
    opportunity_score, scores_by_group = utils.calculate_equal_opportunity(recommendations_by_group, QUALIFIED_MOVIES)
    return opportunity_score, {
        'opportunity_by_group': scores_by_group,
        'variance': np.var(list(scores_by_group.values())),
        'min_max_ratio': min(scores_by_group.values()) / max(scores_by_group.values())
        if scores_by_group and max(scores_by_group.values()) != 0 else 0
    }
