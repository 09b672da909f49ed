"""Drop-in for ``jaccard`` / ``measure_fairness`` of the reference's ``src/phase3_final.py``
(the same two helpers appear in ``src/phase3_aggressive.py:174-195`` with a simpler blend)."""
from __future__ import annotations

import numpy as np

from . import utils


def _jaccard_any_empty_zero(pairs_a, pairs_b):
    """src/phase3_final.py:112-117 over aligned pairs: 0.0 when either set is empty."""
    inter, union = utils.pair_set_counts(pairs_a, pairs_b)
    # either empty -> inter == 0, so inter/union is already 0.0 whenever union > 0
    return utils.jaccard_from_counts(inter, union, 0.0)


def jaccard(s1, s2) -> float:
    """|s1 & s2| / |s1 | s2|, 0.0 if either is empty -- reference: src/phase3_final.py:112-117."""
    return _jaccard_any_empty_zero([list(s1)], [list(s2)])[0]


def measure_fairness(recs_by_group) -> float:
    """0.6 * mean position-wise Jaccard of the first two groups' lists + 0.4 * Jaccard of the
    two groups' item unions -- reference: src/phase3_final.py:119-145.  Fewer than two groups
    -> 0.5; no position with both lists non-empty -> pairwise term 0.5."""
    groups = list(recs_by_group.keys())
    if len(groups) < 2:
        return 0.5
    g1, g2 = recs_by_group[groups[0]], recs_by_group[groups[1]]
    a = [x for x, y in zip(g1, g2) if x and y]
    b = [y for x, y in zip(g1, g2) if x and y]
    all_g1 = list(dict.fromkeys(t for recs in g1 for t in recs))
    all_g2 = list(dict.fromkeys(t for recs in g2 for t in recs))
    sims = _jaccard_any_empty_zero([list(x) for x in a] + [all_g1], [list(y) for y in b] + [all_g2])
    pairwise_sims, overall_sim = sims[:-1], sims[-1]
    pairwise_avg = np.mean(pairwise_sims) if pairwise_sims else 0.5
    return (pairwise_avg * 0.6) + (overall_sim * 0.4)
