"""Drop-in for the hot-path functions of the reference's ``src/utils.py``.

Same names, parameters, return shapes and Python types; the per-pair set arithmetic runs
in the sm_100a list-metric kernel (``ops.list_metrics``).  Intersection and union sizes come
back as exact integers and the ratio is taken in float64 on the host with the reference's own
expression, so results are bit-identical to the reference's.
"""
from __future__ import annotations

import json
import os
import re
from typing import Any, Dict, Hashable, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import encode, ops

_ATTRS = ("gender", "age", "occupation")


def _device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("fairness_llm_replication_b200 needs a CUDA device (sm_100a); "
                           "there is no CPU implementation of the scoring path")
    return torch.device("cuda", torch.cuda.current_device())


def pair_set_counts(lists_a: Sequence[Sequence[Hashable]], lists_b: Sequence[Sequence[Hashable]],
                    k_cap: Optional[int] = None, encoded: Optional[encode.EncodedRecommendations] = None
                    ) -> Tuple[np.ndarray, np.ndarray]:
    """|set(a[:k]) & set(b[:k])| and |set(a[:k]) | set(b[:k])| for aligned list pairs, on the GPU.
    Returns two int64 numpy arrays.  ``encoded``: lists already on the device
    (``encode.EncodedRecommendations``); only their row indices travel.  One upload, one kernel
    launch, one read-back per call."""
    P = len(lists_a)
    if P == 0:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    dev = _device()
    ia = ib = None
    if encoded is not None:
        ia, ib = encoded.rows_of(lists_a), encoded.rows_of(lists_b)
    counts = torch.empty((2, P), dtype=torch.int32, device=dev)
    if ia is None or ib is None:
        # lists not seen before: encode them now; rows are gathered on the host and everything the kernel
        # needs goes up in ONE copy
        enc = encode.EncodedRecommendations(list(lists_a) + list(lists_b))
        seq = np.array(enc.input_rows, dtype=np.int64)
        K = enc.table.shape[1]
        packed = np.concatenate([enc.table[seq].reshape(-1), enc.lens[seq]])
        d = torch.from_numpy(packed).to(dev)
        rows, rlen, vocab = d[:2 * P * K].view(2 * P, K), d[2 * P * K:], enc.vocab_size
    else:
        return pair_set_counts_rows(encoded, ia, ib, k_cap)
    ops.list_metrics_counts(rows[:P], rlen[:P], rows[P:].view(P, 1, K), rlen[P:].view(P, 1), vocab=vocab,
                            k_cap=K if k_cap is None else min(int(k_cap), K), out_inter=counts[0], out_union=counts[1])
    host = counts.cpu().numpy().astype(np.int64)
    return host[0], host[1]


def pair_set_counts_rows(encoded: encode.EncodedRecommendations, rows_a: np.ndarray, rows_b: np.ndarray,
                         k_cap: Optional[int] = None) -> Tuple[np.ndarray, np.ndarray]:
    """``pair_set_counts`` for lists that are already on the device: pair p compares rows ``rows_a[p]`` and
    ``rows_b[p]`` of ``encoded``.  Only the two index arrays go up and the two count arrays come back."""
    P = len(rows_a)
    if P == 0:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    dev = _device()
    table, lens = encoded.on_device(dev)
    K = table.shape[1]
    idx = torch.from_numpy(np.concatenate([rows_a, rows_b]).astype(np.int64, copy=False)).to(dev)
    rows, rlen = table.index_select(0, idx), lens.index_select(0, idx)
    counts = torch.empty((2, P), dtype=torch.int32, device=dev)
    ops.list_metrics_counts(rows[:P], rlen[:P], rows[P:].view(P, 1, K), rlen[P:].view(P, 1), vocab=encoded.vocab_size,
                            k_cap=K if k_cap is None else min(int(k_cap), K), out_inter=counts[0], out_union=counts[1])
    host = counts.cpu().numpy().astype(np.int64)
    return host[0], host[1]


def jaccard_from_counts(inter: np.ndarray, union: np.ndarray, both_empty: float) -> List[float]:
    """``intersection / union if union > 0`` (src/utils.py:237) in float64; union == 0 means
    both sets were empty."""
    safe = np.where(union > 0, union, 1)
    sims = np.where(union > 0, inter / safe, both_empty)
    return sims.tolist()


def calculate_individual_fairness(profile_pairs: List[Tuple[str, str]],
                                  recommendations: Dict[str, List[str]],
                                  _encoded: Optional[encode.EncodedRecommendations] = None) -> Tuple[float, List[float]]:
    """Mean pairwise Jaccard similarity -- reference: src/utils.py:217-244.

    Pairs with a pid missing from ``recommendations`` are skipped (:225-226); two empty lists
    score 1.0 (:232-233); no scored pair gives 0.0 (:242).
    """
    kept = [(p, q) for p, q in profile_pairs if p in recommendations and q in recommendations]
    inter, union = pair_set_counts([recommendations[p] for p, _ in kept],
                                   [recommendations[q] for _, q in kept], encoded=_encoded)
    similarity_scores = jaccard_from_counts(inter, union, 1.0)
    fairness_score = np.mean(similarity_scores) if similarity_scores else 0.0
    return fairness_score, similarity_scores


def create_profile_pairs(profiles: List[Dict], differing_attribute: Optional[str] = None) -> List[Tuple[str, str]]:
    """Pairs (i < j) of profiles that differ in exactly one of gender / age / occupation among the
    attributes present in both -- reference: src/utils.py:327-347, same pair order.

    Host logic (it decides which pairs are scored, it scores nothing).  The reference's
    double Python loop is replaced by blocked integer comparisons, so 10^5 profiles are usable.
    """
    ii, jj = profile_pair_indices(profiles, differing_attribute)
    ids = [p["id"] for p in profiles]
    return [(ids[i], ids[j]) for i, j in zip(ii.tolist(), jj.tolist())]


def profile_pair_indices(profiles: List[Dict], differing_attribute: Optional[str] = None) -> Tuple[np.ndarray, np.ndarray]:
    """``create_profile_pairs`` as two int64 index arrays into ``profiles`` (same pairs, same order): what
    a caller that goes on to score the pairs needs, without a Python tuple per pair."""
    n = len(profiles)
    empty = np.zeros(0, np.int64)
    if n < 2:
        return empty, empty
    codes = np.zeros((len(_ATTRS), n), np.int64)
    for a, attr in enumerate(_ATTRS):
        table: Dict[Hashable, int] = {}
        for i, p in enumerate(profiles):
            if attr in p:
                codes[a, i] = table.setdefault(p[attr], len(table) + 1)
            # 0 = attribute absent
    want = None if differing_attribute is None else (
        _ATTRS.index(differing_attribute) if differing_attribute in _ATTRS else -1)
    out_i, out_j = [], []
    block = max(1, min(n, (1 << 24) // n))
    for s in range(0, n, block):
        e = min(n, s + block)
        ci = codes[:, s:e, None]
        cj = codes[:, None, :]
        differs = (ci != cj) & (ci != 0) & (cj != 0)               # [attr, i, j]
        ok = differs.sum(axis=0) == 1
        if want is not None:
            ok &= differs[want] if want >= 0 else False
        ok &= np.arange(n)[None, :] > np.arange(s, e)[:, None]     # j > i
        ii, jj = np.nonzero(ok)
        out_i.append(ii + s)
        out_j.append(jj)
    return np.concatenate(out_i).astype(np.int64), np.concatenate(out_j).astype(np.int64)


def _aligned_distributions(p: Dict[Hashable, float], q: Dict[Hashable, float]) -> Tuple[np.ndarray, np.ndarray]:
    """The reference's preamble (src/utils.py:75-84, :93-100): union of the keys, 1e-10 for a missing key,
    each side renormalised."""
    keys = set(list(p.keys()) + list(q.keys()))
    pa = np.array([p.get(k, 1e-10) for k in keys])
    qa = np.array([q.get(k, 1e-10) for k in keys])
    return pa / pa.sum(), qa / qa.sum()


def calculate_kl_divergence(p: Dict[Hashable, float], q: Dict[Hashable, float]) -> float:
    """KL(p || q) in nats between two distributions given as dicts -- reference: src/utils.py:70-87
    (``scipy.stats.entropy(p, q)``).  Host arithmetic: two small dicts are not GPU work; the batched form of
    the divergences is ``ops.js_distance_matrix`` behind ``calculate_demographic_parity``."""
    pa, qa = _aligned_distributions(p, q)
    pa, qa = pa / np.sum(pa), qa / np.sum(qa)           # scipy.stats.entropy normalises once more
    return float(np.sum(np.where(pa > 0, pa * np.log(np.where(pa > 0, pa, 1.0) / qa), 0.0)))


def calculate_js_divergence(p: Dict[Hashable, float], q: Dict[Hashable, float]) -> float:
    """Jensen-Shannon DISTANCE (the square root of the divergence, natural log -- what
    ``scipy.spatial.distance.jensenshannon`` returns) between two distributions given as dicts --
    reference: src/utils.py:89-102.  Host arithmetic, as ``calculate_kl_divergence``."""
    pa, qa = _aligned_distributions(p, q)
    pa, qa = pa / np.sum(pa), qa / np.sum(qa)
    m = (pa + qa) / 2.0
    left = np.where(pa > 0, pa * np.log(np.where(pa > 0, pa, 1.0) / m), 0.0)
    right = np.where(qa > 0, qa * np.log(np.where(qa > 0, qa, 1.0) / m), 0.0)
    return float(np.sqrt((np.sum(left) + np.sum(right)) / 2.0))


def calculate_demographic_parity(recommendations_by_group: Dict[str, List[List[str]]],
                                 genre_mapping: Optional[Dict[str, List[str]]] = None,
                                 _encoded: Optional[encode.EncodedRecommendations] = None) -> Tuple[float, Dict]:
    """Demographic parity -- reference: src/utils.py:172-215 (``genre_mapping`` is accepted and unused
    there too).

    Per group the distribution of recommended items over all its lists (every occurrence counts), the
    Jensen-Shannon distance between every two non-empty groups, score = 1 - mean distance.  Counting
    and the distances run on the GPU (``flr_group_item_histogram``, ``flr_js_distance_matrix``); the
    probabilities in ``'distributions'`` are count / total in float64, exactly the reference's values.
    The distances agree with the reference to ~1e-15 relative: the reference itself sums in the
    arbitrary order of a Python ``set`` of titles.
    """
    names, items, index_of, hist = _group_histogram(recommendations_by_group, _encoded)
    G = len(names)
    counts = np.zeros((G, max(len(items), 1)), np.int64)
    divergences: List[float] = []
    if hist is not None:
        if G >= 2:
            dist, valid = ops.js_distance_matrix(hist)
            divergences = [float(d) for d, ok in zip(dist.cpu().tolist(), valid.cpu().tolist()) if ok]
        counts = hist.cpu().numpy()
    distributions = {}
    for g, name in enumerate(names):
        total = int(counts[g].sum())
        nz = np.flatnonzero(counts[g])
        distributions[name] = ({items[x]: int(c) / total for x, c in zip(nz.tolist(), counts[g][nz].tolist())}
                               if total > 0 else {})
    avg_divergence = np.mean(divergences) if divergences else 0
    parity_score = 1 - avg_divergence
    return parity_score, {"divergences": divergences, "distributions": distributions,
                          "avg_divergence": avg_divergence}


def _group_histogram(recommendations_by_group, encoded: Optional[encode.EncodedRecommendations] = None):
    """(group names, items in id order, item -> id lookup, int64 [G, vocab] occurrence counts on the GPU or
    None if there are no rows).  With ``encoded`` the lists are already on the device and only
    (row, group) indices are sent; otherwise the lists of this call are encoded first."""
    names = list(recommendations_by_group.keys())
    if len(names) > 256:
        raise ValueError("more than 256 groups are not supported on the GPU path")
    lists, row_group = [], []
    for g, name in enumerate(names):
        for recs in recommendations_by_group[name]:
            lists.append(recs)
            row_group.append(g)
    rows = encoded.rows_of(lists) if encoded is not None else None
    if rows is None:
        encoded = encode.EncodedRecommendations(lists)
        rows = encoded.rows_of(lists)
    items = encoded.items
    index_of = _ItemIndex(items)
    if not lists:
        return names, items, index_of, None
    dev = _device()
    table, lens = encoded.on_device(dev)
    sel = torch.from_numpy(np.concatenate([rows, np.asarray(row_group, np.int64)])).to(dev)
    R = len(lists)
    hist = ops.group_item_histogram(table.index_select(0, sel[:R]), lens.index_select(0, sel[:R]),
                                    vocab=encoded.vocab_size, n_groups=len(names), groups=sel[R:].to(torch.uint8))
    return names, items, index_of, hist


class _ItemIndex:
    """item -> id, built on first use (equal opportunity looks up a handful of titles)."""

    def __init__(self, items):
        self._items, self._index = items, None

    def get(self, item):
        if self._index is None:
            self._index = {x: i for i, x in enumerate(self._items)}
        return self._index.get(item)


def calculate_equal_opportunity(recommendations_by_group: Dict[str, List[List[str]]],
                                relevant_items: set,
                                _encoded: Optional[encode.EncodedRecommendations] = None) -> Tuple[float, Dict[str, float]]:
    """Equal opportunity -- reference: src/utils.py:246-275.

    Per group: distinct relevant items recommended / recommendations made (0.0 for a group without
    recommendations); fairness = 1 / (1 + population variance of the group scores).  The two integer
    counts per group come from the GPU; the ratios, ``np.var`` and the final expression are the
    reference's own, so the result is bit-identical.
    """
    names, _, index_of, hist = _group_histogram(recommendations_by_group, _encoded)
    opportunity_scores: Dict[str, float] = {}
    if hist is None:
        opportunity_scores = {name: 0.0 for name in names}
    else:
        flags = np.zeros(hist.shape[1], np.uint8)
        for item in relevant_items:
            i = index_of.get(item)
            if i is not None:
                flags[i] = 1
        rel, tot = ops.equal_opportunity_counts(hist, torch.from_numpy(flags).to(hist.device))
        both = torch.stack([rel, tot]).cpu().tolist()            # one read-back
        for name, r, n in zip(names, both[0], both[1]):
            opportunity_scores[name] = r / n if n > 0 else 0.0
    scores = list(opportunity_scores.values())
    variance = np.var(scores) if scores else 0
    fairness_score = 1 / (1 + variance)
    return fairness_score, opportunity_scores


def _exposure_means(groups_in_rank_order: Sequence[Hashable]):
    """(group names in order of first appearance, {name: np.float64 mean exposure}) on the GPU."""
    names: List[Hashable] = []
    code: Dict[Hashable, int] = {}
    ids = []
    for g in groups_in_rank_order:
        if g not in code:
            code[g] = len(names)
            names.append(g)
        ids.append(code[g])
    if not ids:
        return names, {}
    if len(ids) > 128 or len(names) > 255:
        raise ValueError("exposure ratio: at most 128 ranked items and 255 groups on the GPU path")
    dev = _device()
    mean, _ = ops.exposure_by_group(torch.tensor([ids], dtype=torch.uint8, device=dev),
                                    torch.tensor([len(ids)], dtype=torch.int32, device=dev), len(names))
    vals = mean.cpu().numpy()[0]
    return names, {name: vals[i] for i, name in enumerate(names)}


def calculate_exposure_ratio(rankings: List[Tuple[str, str]], protected_attribute: str = 'gender') -> Tuple[float, Dict[str, float]]:
    """Exposure ratio of a ranking of (item, group) tuples -- reference: src/utils.py:277-305: exposure
    1/log2(position + 2), mean per group, min/max of the means (1.0 when there is nothing to compare)."""
    _, avg_exposure = _exposure_means([group for _, group in rankings])
    if not avg_exposure:
        return 1.0, {}
    max_exposure = max(avg_exposure.values())
    min_exposure = min(avg_exposure.values())
    exposure_ratio = min_exposure / max_exposure if max_exposure > 0 else 1.0
    return exposure_ratio, avg_exposure


def calculate_ndcg(rankings: List[str], ground_truth: Dict[str, float], k: int = 10) -> float:
    """NDCG@k -- reference: src/utils.py:113-132 (missing items count 0; IDCG from the k largest
    ground-truth values; 0.0 when IDCG is 0).  Bit-identical: same additions in the same order."""
    rel = [float(ground_truth.get(item, 0)) for item in rankings[:k]]
    ideal = [float(v) for v in sorted(ground_truth.values(), reverse=True)[:k]]
    if not ideal:
        return 0.0                       # IDCG is 0
    dev = _device()
    n = max(len(rel), 1)
    out = ops.ndcg(torch.tensor([rel + [0.0] * (n - len(rel))], dtype=torch.float64, device=dev),
                   torch.tensor([len(rel)], dtype=torch.int32, device=dev),
                   torch.tensor([ideal], dtype=torch.float64, device=dev),
                   torch.tensor([len(ideal)], dtype=torch.int32, device=dev), 1 << 30)   # already cut to k
    v = float(out.cpu().numpy()[0])
    return v if v != 0.0 else 0.0


# ---- the data formats either side of the path: host text/JSON only ----
_LEADING_NUMBER = re.compile(r'^\d+[\.\)]\s*')
_LEADING_BULLET = re.compile(r'^[\-\*]\s*')
_YEAR_SUFFIX = re.compile(r'\s*\(\d{4}\)')


def load_results(filename: str, results_dir: str = '../results') -> Any:
    """Read ``results_dir/filename`` as JSON -- reference: src/utils.py:33-49 (the reference logs and
    re-raises; the exception a caller sees is the same FileNotFoundError / JSONDecodeError)."""
    with open(os.path.join(results_dir, filename), 'r') as f:
        return json.load(f)


def parse_recommendation_list(text: str) -> List[str]:
    """One title per non-blank line of a model reply -- reference: src/utils.py:350-375.  In this
    order: trim, drop one leading "12." / "12)" counter, drop one leading "-" or "*" bullet, trim
    straight quotes from both ends; a line left empty is dropped.  These strings are what the
    string->id encoder (``encode.Vocabulary``) sees, so equality of titles is decided here."""
    out = []
    for raw in text.strip().split('\n'):
        line = raw.strip()
        if not line:
            continue
        line = _LEADING_BULLET.sub('', _LEADING_NUMBER.sub('', line)).strip('"\'')
        if line:
            out.append(line)
    return out


def extract_movie_title(text: str) -> str:
    """Title without its "(1999)" year tags -- reference: src/utils.py:377-382."""
    return _YEAR_SUFFIX.sub('', text).strip()
