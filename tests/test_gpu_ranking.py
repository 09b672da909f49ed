"""N3: phase-2 ranking metrics through the GPU drop-ins, bit-identical to the reference's golden outputs."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_dropin_ranking_metrics(golden):
    from fairness_llm_replication_b200 import phase2_cross_model_eval as p2, utils
    rm = golden["ranking_metrics"]
    assert utils.calculate_ndcg(["i2", "i1", "i3"], {"i1": 5, "i2": 3, "i3": 1}) == rm["ndcg_known"]
    assert utils.calculate_ndcg(["a"], {}) == 0.0
    assert p2.calculate_exposure_ratio([], []) == (1.0, {})
    assert [[a, float(b)] for a, b in p2.calculate_pairwise_preference_ratio(rm["comparisons"]).items()] == rm["win_rates"]
    for c in rm["cases"]:
        ratio, avg = p2.calculate_exposure_ratio(c["ranked"], c["items"])
        assert float(ratio) == c["ratio"]
        assert [[a, float(b)] for a, b in avg.items()] == c["avg"]
        nd = p2.calculate_ndcg_per_group(c["ranked"], c["items"], c["k"])
        assert [[a, float(b)] for a, b in nd.items()] == c["ndcg_per_group"]
        r2, avg2 = utils.calculate_exposure_ratio([tuple(t) for t in c["tuples"]])
        assert float(r2) == c["ratio_tuples"] and [[a, float(b)] for a, b in avg2.items()] == c["avg_tuples"]
        gt = {it["id"]: it.get("relevance_score", 0.5) for it in c["items"]}
        assert float(utils.calculate_ndcg(c["ranked"], gt, c["k"])) == c["ndcg_all"]


def test_batched_ranking_kernels_match_numpy():
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(4)
    B, n, G = 500, 37, 5
    group = rng.integers(0, G + 1, size=(B, n)).astype(np.uint8)
    group[group == G] = 255
    lens = rng.integers(0, n + 1, size=B).astype(np.int32)
    mean, cnt = ops.exposure_by_group(torch.from_numpy(group).cuda(), torch.from_numpy(lens).cuda(), G)
    w = 1.0 / np.log2(np.arange(n) + 2)
    for b in range(0, B, 7):
        for g in range(G):
            sel = w[:lens[b]][group[b, :lens[b]] == g]
            assert cnt[b, g].item() == sel.size
            assert mean[b, g].item() == (float(np.mean(sel)) if sel.size else 0.0)     # numpy's own summation order
    rel = rng.random((B, n)); ideal = -np.sort(-rng.random((B, n)), axis=1)
    out = ops.ndcg(torch.from_numpy(rel).cuda(), torch.from_numpy(lens).cuda(), torch.from_numpy(ideal).cuda(),
                   torch.full((B,), n, dtype=torch.int32).cuda(), 10).cpu().numpy()
    for b in range(0, B, 11):
        dcg = 0.0
        for i in range(min(lens[b], 10)):
            dcg += rel[b, i] / np.log2(i + 2)
        idcg = sum(ideal[b, i] / np.log2(i + 2) for i in range(10))
        assert out[b] == dcg / idcg
