"""End-to-end pipeline on the GPU against the oracle, host-buffer streaming path included."""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O

pytestmark = pytest.mark.gpu


def _oracle_pipeline(cfg, hb):
    neutral, nlen, lists, lens, en, e = (t.numpy() for t in hb.tensors())
    lm = O.list_metrics(neutral, nlen, lists, lens, k_cap=cfg.k_cap, vocab=cfg.vocab)
    scores = O.cosine_nonconformity_mirror(en, e)
    q0 = O.conformal_threshold(scores, cfg.alpha)
    v = O.violation_rounds(scores, q0, cfg.gamma, cfg.rounds, cfg.q_min)
    return lm, scores, v


@pytest.mark.parametrize("U,K", [(512, 50), (333, 20)])
def test_pipeline_matches_oracle(U, K):
    from fairness_llm_replication_b200 import pipeline
    cfg = pipeline.PipelineConfig(list_len=K)
    hb = pipeline.synth_batch_host(cfg, 1000, U)
    db = pipeline.synth_batch(cfg, 1000, U, "cuda:0")
    for h, d in zip(hb.tensors(), db.tensors()):
        assert torch.equal(h, d.cpu())                                     # host and device generators agree
    pipe = pipeline.ScoringPipeline(cfg)
    res = pipe.run(db)
    summ = pipe.summarize(res)
    lm, scores, v = _oracle_pipeline(cfg, hb)
    assert np.array_equal(res.inter.cpu().numpy(), lm["inter"])            # intersections bit-exact
    np.testing.assert_allclose(res.jaccard.cpu().numpy(), lm["jaccard"], rtol=1e-6)
    np.testing.assert_allclose(res.serp.cpu().numpy(), lm["serp"], rtol=1e-6)
    np.testing.assert_allclose(res.prag.cpu().numpy(), lm["prag"], rtol=1e-6)
    assert np.array_equal(res.scores.cpu().numpy(), scores)
    assert np.array_equal(res.thresholds.cpu().numpy(), v["thresholds"])   # selected order statistic bit-exact
    assert np.array_equal(res.flags.cpu().numpy(), v["flags"])             # violation flags bit-exact
    assert summ["violation_counts"].tolist() == v["counts"].tolist()
    assert np.array_equal(summ["variant_violation_counts"], v["variant_counts"])
    assert summ["n_scores"] == U * cfg.n_variants and summ["n_users"] == U and summ["oov_ids"] == 0
    np.testing.assert_allclose(summ["variant_means"][:, 0], lm["jaccard"].mean(0), rtol=1e-12)
    np.testing.assert_allclose(summ["variant_means"][:, 1], lm["serp"].mean(0), rtol=1e-12)
    np.testing.assert_allclose(summ["variant_means"][:, 2], lm["prag"].mean(0), rtol=1e-12)
    for i, m in enumerate(("jaccard", "serp", "prag")):
        want = O.snsr_snsv(lm[m].mean(0))
        for attr in want:
            assert summ["spreads"][attr][m]["snsr"] == pytest.approx(want[attr]["snsr"], rel=1e-9, abs=1e-15)
            assert summ["spreads"][attr][m]["snsv"] == pytest.approx(want[attr]["snsv"], rel=1e-9, abs=1e-15)


def test_run_host_streams_to_the_same_result():
    from fairness_llm_replication_b200 import pipeline
    cfg = pipeline.PipelineConfig(list_len=50)
    U = 2500
    hb = pipeline.synth_batch_host(cfg, 0, U, pin=True)
    pipe = pipeline.ScoringPipeline(cfg)
    ref = pipe.run(pipeline.Batch(*[t.cuda() for t in hb.tensors()]))
    for chunk in (700, 1024, 4096):
        res, summ = pipe.run_host(hb, "cuda:0", chunk_users=chunk)
        assert torch.equal(res.inter, ref.inter) and torch.equal(res.prag, ref.prag)
        assert torch.equal(res.scores, ref.scores) and torch.equal(res.flags, ref.flags)
        assert torch.equal(res.thresholds, ref.thresholds)
        assert torch.equal(res.reduced[31 * 4:], ref.reduced[31 * 4:])
        assert torch.allclose(res.reduced[:31 * 4], ref.reduced[:31 * 4], rtol=1e-12)
    assert summ["n_users"] == U


def test_pipeline_repeatable():
    from fairness_llm_replication_b200 import pipeline
    cfg = pipeline.PipelineConfig(list_len=50)
    db = pipeline.synth_batch(cfg, 0, 20000, "cuda:0")
    pipe = pipeline.ScoringPipeline(cfg)
    a = pipe.run(db)
    b = pipe.run(db)
    assert torch.equal(a.reduced, b.reduced) and torch.equal(a.flags, b.flags) and torch.equal(a.thresholds, b.thresholds)


@pytest.mark.parametrize("K,k_cap,vocab,V", [(50, 10, 3706, 31), (20, 7, 200, 31), (30, 30, 3706, 12), (50, 50, 130, 31)])
def test_pipeline_other_shapes(K, k_cap, vocab, V):
    """k_cap below K, small vocabularies (many shared items, vocab just above 2K + 2), V != 31."""
    from fairness_llm_replication_b200 import pipeline
    cfg = pipeline.PipelineConfig(list_len=K, k_cap=k_cap, vocab=vocab, n_variants=V, emb_dim=128, rounds=3,
                                  attribute_slices={"all": (0, V)})
    U = 257
    hb = pipeline.synth_batch_host(cfg, 7, U)
    pipe = pipeline.ScoringPipeline(cfg)
    res = pipe.run(pipeline.Batch(*[t.cuda() for t in hb.tensors()]))
    summ = pipe.summarize(res)
    lm, scores, v = _oracle_pipeline(cfg, hb)
    assert np.array_equal(res.inter.cpu().numpy(), lm["inter"])
    np.testing.assert_allclose(res.serp.cpu().numpy(), lm["serp"], rtol=1e-6)
    np.testing.assert_allclose(res.prag.cpu().numpy(), lm["prag"], rtol=1e-6)
    assert np.array_equal(res.scores.cpu().numpy(), scores)
    assert np.array_equal(res.thresholds.cpu().numpy(), v["thresholds"])
    assert np.array_equal(res.flags.cpu().numpy(), v["flags"])
    np.testing.assert_allclose(summ["variant_means"][:, 2], lm["prag"].mean(0), rtol=1e-12)
    assert set(summ["spreads"]) == {"all"}
