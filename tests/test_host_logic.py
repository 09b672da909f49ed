"""Host-side logic of the drop-in layer that needs no GPU: pair enumeration, encoding,
calibration-record generation, sharding, the pipeline's summary arithmetic."""
import numpy as np
import pytest
import torch

from fairness_llm_replication_b200 import encode, pipeline, utils
from fairness_llm_replication_b200 import phase3_facter_mitigation as p3
from oracle import np_oracle as O


def test_create_profile_pairs_matches_reference(golden):
    for c in golden["profile_pairs"]:
        got = utils.create_profile_pairs(c["profiles"], c["attr"])
        assert [list(p) for p in got] == c["pairs"]
        assert all(isinstance(p, tuple) for p in got)


def test_create_profile_pairs_blocked_path():
    profs = [{"id": i, "gender": i % 3, "age": (i // 3) % 5, "occupation": "x"} for i in range(300)]
    got = utils.create_profile_pairs(profs)
    want = [(a["id"], b["id"]) for i, a in enumerate(profs) for b in profs[i + 1:]
            if sum(a[k] != b[k] for k in ("gender", "age", "occupation")) == 1]
    assert got == want
    assert utils.create_profile_pairs([]) == [] and utils.create_profile_pairs(profs[:1]) == []
    assert utils.create_profile_pairs(profs, "race") == []


def test_encode_pair_lists():
    a, la, b, lb, vocab = encode.encode_pair_lists([["x", "y", "x"], []], [["y", " x"], ["x"]])
    assert a.shape == (2, 3) and b.shape == (2, 3) and vocab == 3
    assert la.tolist() == [3, 0] and lb.tolist() == [2, 1]
    assert a[0].tolist() == [0, 1, 0] and b[0, :2].tolist() == [1, 2] and b[1, 0] == 0
    assert a[1].tolist() == [-1, -1, -1]
    a, la, b, lb, vocab = encode.encode_pair_lists([[]], [[]])
    assert a.shape == (1, 1) and vocab == 1


def test_generate_calibration_data_matches_reference(golden):
    c = golden["calibration_data"]
    np.random.seed(c["seed"])
    df = p3.generate_calibration_data(c["profiles"], c["recs"], c["ratings"])
    assert list(df.columns) == list(c["columns"].keys()) or set(df.columns) == set(c["columns"].keys())
    for k, want in c["columns"].items():
        got = [v if isinstance(v, str) else float(v) for v in df[k]]
        assert got == want
    empty = p3.generate_calibration_data([], {})
    assert len(empty) == 0 and len(empty.columns) == 0


def test_generate_calibration_data_keyerror_like_reference():
    with pytest.raises(KeyError):
        p3.generate_calibration_data([{"id": "u"}], {"u": {"recommendations": ["m"]}})


def test_shard_range_partitions():
    for n in (0, 1, 7, 8, 6040, 10_000_000):
        for w in (1, 2, 3, 8):
            spans = [pipeline.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_summarize_matches_oracle_spreads():
    cfg = pipeline.PipelineConfig(list_len=20, rounds=3)
    V, R = cfg.n_variants, cfg.rounds
    rng = np.random.default_rng(0)
    U = 17
    jac = rng.random((U, V)); serp_num = rng.integers(0, 210, (U, V)); prag_num = rng.integers(0, 190, (U, V))
    sums = np.stack([jac.sum(0), serp_num.sum(0), prag_num.sum(0), np.full(V, U)], 1).astype(np.float64)
    counts = rng.integers(0, 50, R); vcounts = rng.integers(0, 5, (V, R))
    red = torch.from_numpy(np.concatenate([sums.ravel(), counts, vcounts.ravel()]).astype(np.float64))
    st = torch.zeros((1, 8), dtype=torch.int64); st[0, 2] = U * V; st[0, 3] = 5
    res = pipeline.StepResult(None, None, None, None, None, None, torch.tensor([.1, .09, .08, .07]), st, red,
                              torch.zeros(1, dtype=torch.int32))
    s = pipeline.ScoringPipeline(cfg).summarize(res)
    want = O.snsr_snsv(jac.mean(0))
    for attr in want:
        assert s["spreads"][attr]["jaccard"]["snsr"] == pytest.approx(want[attr]["snsr"], rel=1e-12)
        assert s["spreads"][attr]["jaccard"]["snsv"] == pytest.approx(want[attr]["snsv"], rel=1e-12)
    assert np.allclose(s["variant_means"][:, 1], serp_num.mean(0) / (20 * 21 / 2))
    assert np.allclose(s["variant_means"][:, 2], prag_num.mean(0) / (20 * 21))
    assert s["violation_counts"].tolist() == counts.tolist() and s["n_scores"] == U * V


def test_ops_refuse_cpu_tensors():
    from fairness_llm_replication_b200 import ops
    with pytest.raises(TypeError):
        ops.cosine_nonconformity(torch.zeros(2, 8), torch.zeros(2, 3, 8))
    with pytest.raises(TypeError):
        ops.list_metrics(torch.zeros(2, 4, dtype=torch.int32), torch.zeros(2, dtype=torch.int32),
                         torch.zeros(2, 1, 4, dtype=torch.int32), torch.zeros(2, 1, dtype=torch.int32), vocab=8)


def test_text_parsing_matches_reference(golden, tmp_path):
    """N2: reply text -> titles (src/utils.py:350-382) and the JSON loader (src/utils.py:33-49)."""
    import json
    from fairness_llm_replication_b200 import utils
    for c in golden["text_parsing"]:
        lines = utils.parse_recommendation_list(c["text"])
        assert lines == c["lines"]
        assert [utils.extract_movie_title(x) for x in lines] == c["titles"]
    (tmp_path / "phase1").mkdir()
    payload = {"recommendations": {"p1": {"recommendations": ["A", "B"]}}, "n": 3}
    (tmp_path / "phase1" / "r.json").write_text(json.dumps(payload))
    assert utils.load_results("phase1/r.json", results_dir=str(tmp_path)) == payload
    with pytest.raises(FileNotFoundError):
        utils.load_results("missing.json", results_dir=str(tmp_path))
