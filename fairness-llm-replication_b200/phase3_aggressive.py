"""Drop-in for ``jaccard`` / ``measure_fairness`` of the reference's ``src/phase3_aggressive.py`` (lines
174-195), the earlier form of the helpers in ``src/phase3_final.py``: same ``jaccard``; ``measure_fairness``
wants exactly two groups and averages over every aligned position, an empty list counting 0.0."""
from __future__ import annotations

import numpy as np

from .phase3_final import _jaccard_any_empty_zero, jaccard  # noqa: F401  (jaccard: src/phase3_aggressive.py:174-179)


def measure_fairness(recs_by_group) -> float:
    """Mean position-wise Jaccard of the two groups' lists -- reference: src/phase3_aggressive.py:181-195.
    Any number of groups other than two -> 0.5; no aligned position -> 0.5.  One GPU call for all positions."""
    groups = list(recs_by_group.keys())
    if len(groups) != 2:
        return 0.5
    g1, g2 = recs_by_group[groups[0]], recs_by_group[groups[1]]
    n = min(len(g1), len(g2))
    if n == 0:
        return 0.5
    sims = _jaccard_any_empty_zero([list(x) for x in g1[:n]], [list(y) for y in g2[:n]])
    return np.mean(sims)
