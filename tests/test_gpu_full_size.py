"""Parity at BASELINE.json's full sizes (configs[2], [3] and the per-GPU shard of configs[4]).

The oracle cannot restate 38.75 M pairs in seconds, so at these sizes the checks are
  * windows of users regenerated on the host (the synthetic generator is counter-based: any user range
    can be produced alone) and scored by the oracle -- bit-exact against the same rows of the big run;
  * the global steps redone by the oracle on the GPU's own score array (sort, order statistic, the
    round recurrence): thresholds, counts and every flag bit-exact;
  * properties that need no oracle: a sub-range scored alone equals the slice of the big run, the
    per-variant sums of two halves add up to the whole, similarity matrices are symmetric with a
    unit diagonal.
"""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O

pytestmark = pytest.mark.gpu


def _need_gb(gb):
    free, _ = torch.cuda.mem_get_info()
    if free < gb * 1e9:
        pytest.skip(f"needs {gb} GB of free device memory, {free / 1e9:.0f} GB available")


def test_config5_shard_full_size():
    """1.25 M users x 31 variants, K = 50, D = 384 (69.6 GB of inputs): the shard one GPU of eight owns."""
    from fairness_llm_replication_b200 import ops, pipeline
    _need_gb(95)
    cfg = pipeline.PipelineConfig()
    U, user0 = 1_250_000, 3_750_000                        # rank 3 of 8 in the 10 M-user sweep
    V, R = cfg.n_variants, cfg.rounds
    db = pipeline.synth_batch(cfg, user0, U, "cuda:0")
    pipe = pipeline.ScoringPipeline(cfg)
    res = pipe.run(db)
    summ = pipe.summarize(res)
    torch.cuda.synchronize()
    assert summ["n_users"] == U and summ["n_scores"] == U * V and summ["oov_ids"] == 0

    # -- windows of users against the oracle (first / last users, a tile boundary, the middle) --
    for off, n in ((0, 48), (4093, 37), (617_001, 64), (U - 50, 50)):
        hb = pipeline.synth_batch_host(cfg, user0 + off, n)
        for h, d in zip(hb.tensors(), db.tensors()):
            assert torch.equal(h, d[off:off + n].cpu())
        neutral, nlen, lists, lens, en, e = (t.numpy() for t in hb.tensors())
        lm = O.list_metrics(neutral, nlen, lists, lens, vocab=cfg.vocab)
        sl = slice(off, off + n)
        assert np.array_equal(res.inter[sl].cpu().numpy(), lm["inter"])
        np.testing.assert_allclose(res.jaccard[sl].cpu().numpy(), lm["jaccard"], rtol=1e-6)
        np.testing.assert_allclose(res.serp[sl].cpu().numpy(), lm["serp"], rtol=1e-6)
        np.testing.assert_allclose(res.prag[sl].cpu().numpy(), lm["prag"], rtol=1e-6)
        assert np.array_equal(res.scores[sl].cpu().numpy(), O.cosine_nonconformity_mirror(en, e))

    # -- the global steps, redone by the oracle on the 38.75 M scores the GPU produced --
    scores = res.scores.cpu().numpy()
    q0 = O.conformal_threshold(scores, cfg.alpha)
    v = O.violation_rounds(scores, q0, cfg.gamma, R, cfg.q_min)
    assert np.array_equal(res.thresholds.cpu().numpy(), v["thresholds"])
    assert summ["violation_counts"].tolist() == v["counts"].tolist()
    assert np.array_equal(summ["variant_violation_counts"], v["variant_counts"])
    assert np.array_equal(res.flags.cpu().numpy(), v["flags"])
    del scores, v

    # -- a sub-range scored alone equals the slice; halves add up to the whole --
    lo, hi = 600_003, 611_114
    part = ops.list_metrics(db.neutral[lo:hi], db.neutral_len[lo:hi], db.lists[lo:hi], db.lens[lo:hi],
                            vocab=cfg.vocab, want_union=False)
    assert torch.equal(part.inter, res.inter[lo:hi]) and torch.equal(part.prag, res.prag[lo:hi])
    assert torch.equal(part.serp, res.serp[lo:hi]) and torch.equal(part.jaccard, res.jaccard[lo:hi])
    assert torch.equal(ops.cosine_nonconformity(db.emb_neutral[lo:hi], db.emb[lo:hi]), res.scores[lo:hi])
    half = U // 2 + 1
    sums = [ops.list_metrics(db.neutral[a:b], db.neutral_len[a:b], db.lists[a:b], db.lens[a:b], vocab=cfg.vocab,
                             want_union=False).variant_sums.cpu().numpy() for a, b in ((0, half), (half, U))]
    whole = res.reduced[:V * 4].view(V, 4).cpu().numpy()
    np.testing.assert_allclose(sums[0] + sums[1], whole, rtol=1e-12)
    assert np.array_equal(sums[0][:, 1:] + sums[1][:, 1:], whole[:, 1:])      # integer numerators and counts: exact
    # the same sums from the per-pair outputs
    np.testing.assert_allclose(res.jaccard.double().sum(0).cpu().numpy(), whole[:, 0], rtol=1e-7)
    assert whole[:, 3].tolist() == [float(U)] * V


def test_config3_full_size():
    """1,000,000 calibration pairs of 384-d embeddings: score, exact quantile, violation rounds."""
    from fairness_llm_replication_b200 import ops
    from fairness_llm_replication_b200 import phase3_facter_mitigation as p3
    _need_gb(8)
    N, D, seed = 1_000_000, 384, 20251203
    a, b = ops.synth_embeddings(seed, 0, N, 1, D, "cuda:0")           # a [N, D], b [N, 1, D]
    s = p3.embedding_nonconformity(a, b[:, 0])
    assert s.shape == (N,)
    for off, n in ((0, 200), (499_999, 130), (N - 77, 77)):
        ha, hb = ops.synth_embeddings_host(seed, off, n, 1, D)
        assert np.array_equal(s[off:off + n].cpu().numpy(), O.cosine_nonconformity_mirror(ha, hb)[:, 0])
    scores = s.cpu().numpy()
    for alpha in (0.1, 0.01, 0.5):
        thr, state = ops.radix_select(s, alpha=alpha)
        assert thr.cpu().numpy()[0] == O.conformal_threshold(scores, alpha)      # the very element np.sort picks
    thr0, state = ops.radix_select(s, alpha=0.1)
    thr = ops.threshold_schedule(thr0, state, gamma=0.95, q_min=0.0, rounds=4)
    flags, counts, _ = ops.violation_detect(s, thr, rounds=4)
    v = O.violation_rounds(scores.reshape(N, 1), O.conformal_threshold(scores, 0.1), 0.95, 4, 0.0)
    assert np.array_equal(thr.cpu().numpy(), v["thresholds"])
    assert counts.cpu().numpy().tolist() == v["counts"].tolist()
    assert np.array_equal(flags.cpu().numpy().reshape(N, 1), v["flags"])
    # a sub-range alone equals the slice
    assert torch.equal(p3.embedding_nonconformity(a[123_457:234_568], b[123_457:234_568, 0]), s[123_457:234_568])


def test_config4_full_size():
    """6,040 users x 124 lists / embeddings (4 models x 31 variants), K = 20, D = 384."""
    from fairness_llm_replication_b200 import ops
    _need_gb(8)
    U, M, K, D, vocab, seed = 6040, 124, 20, 384, 3706, 20251203
    # 124 rows per user = 4 consecutive synthetic users' 31 variant rows
    _, e = ops.synth_embeddings(seed, 0, U * 4, 31, D, "cuda:0")
    x = e.reshape(U, M, D)
    xh = ops.normalize_rows_bf16(x)
    S = ops.pair_similarity(xh)
    torch.cuda.synchronize()
    assert tuple(S.shape) == (U, M, M)
    assert float((S - S.transpose(1, 2)).abs().max()) <= 1e-6
    assert float((torch.diagonal(S, dim1=1, dim2=2) - 1.0).abs().max()) < 1e-2
    for u0, n in ((0, 6), (3017, 5), (U - 4, 4)):
        got_x = xh[u0:u0 + n].float().cpu().numpy()
        want_x = O.normalize_rows_bf16(x[u0:u0 + n].cpu().numpy())
        assert np.abs(got_x - want_x).max() <= 2 ** -8
        assert np.abs(S[u0:u0 + n].cpu().numpy() - O.pair_similarity(got_x)).max() <= 2e-3
    # all-pairs list overlap over the same 124 lists
    _, _, lists, lens = ops.synth_lists(seed, 0, U * 4, 31, K, vocab, "cuda:0")
    lists, lens = lists.reshape(U, M, K).contiguous(), lens.reshape(U, M).contiguous()
    inter, sizes = ops.pair_jaccard(lists, lens, vocab=vocab)
    torch.cuda.synchronize()
    assert torch.equal(inter, inter.transpose(1, 2))
    assert torch.equal(torch.diagonal(inter, dim1=1, dim2=2), sizes)
    for u0, n in ((0, 3), (4001, 3), (U - 2, 2)):
        want = O.pair_jaccard(lists[u0:u0 + n].cpu().numpy(), lens[u0:u0 + n].cpu().numpy())
        assert np.array_equal(inter[u0:u0 + n].cpu().numpy(), want["inter"])
        assert np.array_equal(sizes[u0:u0 + n].cpu().numpy(), want["sizes"])
