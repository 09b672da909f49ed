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
