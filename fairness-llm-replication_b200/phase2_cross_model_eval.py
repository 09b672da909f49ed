"""Drop-in for the ranking-fairness metrics of the reference's ``src/phase2_cross_model_eval.py``
(SURVEY.md section 8(f), N3).  The reference ranks <= 20 synthetic documents per call, so there is
nothing to speed up at its own shapes; the point of these is that a phase-2 run can stay on the same
library, with results identical to the reference's."""
from __future__ import annotations

from collections import defaultdict
from typing import Dict, List, Tuple

import torch

from . import ops, utils


def calculate_exposure_ratio(ranked_items: List[str], items_data: List[Dict],
                             protected_attr: str = 'protected_attribute') -> Tuple[float, Dict]:
    """Exposure ratio across protected groups -- reference: src/phase2_cross_model_eval.py:216-254.
    Items missing from ``items_data`` are skipped but keep their rank position; fewer than two groups
    -> ratio 1.0."""
    item_dict = {item['id']: item for item in items_data}
    names: List = []
    code: Dict = {}
    ids = []
    for item_id in ranked_items:
        if item_id not in item_dict:
            ids.append(255)                                   # position kept, exposure not counted
            continue
        group = item_dict[item_id].get(protected_attr, 'unknown')
        if group not in code:
            code[group] = len(names)
            names.append(group)
        ids.append(code[group])
    avg_exposure: Dict = {}
    if names:
        if len(ids) > 128 or len(names) > 254:
            raise ValueError("exposure ratio: at most 128 ranked items and 254 groups on the GPU path")
        dev = utils._device()
        mean, _ = ops.exposure_by_group(torch.tensor([ids], dtype=torch.uint8, device=dev),
                                        torch.tensor([len(ids)], dtype=torch.int32, device=dev), len(names))
        vals = mean.cpu().numpy()[0]
        avg_exposure = {name: vals[i] for i, name in enumerate(names)}
    if not avg_exposure or len(avg_exposure) < 2:
        return 1.0, avg_exposure
    max_exp = max(avg_exposure.values())
    min_exp = min(avg_exposure.values())
    ratio = min_exp / max_exp if max_exp > 0 else 1.0
    return ratio, avg_exposure


def calculate_pairwise_preference_ratio(comparisons: List[Dict]) -> Dict[str, float]:
    """Win rate per protected group over pairwise comparisons -- reference:
    src/phase2_cross_model_eval.py:256-279.  Pure counting over a handful of dicts: host code."""
    group_wins = defaultdict(int)
    group_total = defaultdict(int)
    for comp in comparisons:
        if comp['winner'] == 'A':
            group_wins[comp['item_a_attr']] += 1
        elif comp['winner'] == 'B':
            group_wins[comp['item_b_attr']] += 1
        group_total[comp['item_a_attr']] += 1
        group_total[comp['item_b_attr']] += 1
    return {group: group_wins[group] / group_total[group] if group_total[group] > 0 else 0
            for group in group_total.keys()}


def calculate_ndcg_per_group(ranked_items: List[str], items_data: List[Dict], k: int = 10) -> Dict[str, float]:
    """NDCG@k of the ranking restricted to each protected group -- reference:
    src/phase2_cross_model_eval.py:281-316.  All groups go to the GPU in one batched call."""
    items_by_group = defaultdict(list)
    for item in items_data:
        items_by_group[item.get('protected_attribute', 'unknown')].append(item)
    groups = list(items_by_group.keys())
    if not groups:
        return {}
    rels, ideals = [], []
    for group in groups:
        ground_truth = {item['id']: item.get('relevance_score', 0.5) for item in items_by_group[group]}
        group_ranking = [item_id for item_id in ranked_items if item_id in ground_truth]
        rels.append([float(ground_truth.get(i, 0)) for i in group_ranking[:k]])
        ideals.append([float(v) for v in sorted(ground_truth.values(), reverse=True)[:k]])
    n = max([len(r) for r in rels] + [1])
    m = max([len(r) for r in ideals] + [1])
    dev = utils._device()
    pad = lambda rows, w: torch.tensor([r + [0.0] * (w - len(r)) for r in rows], dtype=torch.float64, device=dev)
    out = ops.ndcg(pad(rels, n), torch.tensor([len(r) for r in rels], dtype=torch.int32, device=dev),
                   pad(ideals, m), torch.tensor([len(r) for r in ideals], dtype=torch.int32, device=dev), 1 << 30)   # rows are already cut to k
    vals = out.cpu().tolist()
    return {group: (v if v != 0.0 else 0.0) for group, v in zip(groups, vals)}
