"""world_size-2 run of the sharded path's host logic on CPU (gloo): contiguous user shards,
the summed radix histograms give every rank the GLOBAL order statistic, and the fused
float64 buffer of variant sums + violation counts reduces to the single-process values.
The kernels themselves are stood in for by the oracle (no GPU here)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fairness_llm_replication_b200 import pipeline
from oracle import np_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _inputs(U=96, V=31, K=8, D=16, seed=0):
    rng = np.random.default_rng(seed)
    neutral = np.stack([rng.permutation(60)[:K] for _ in range(U)]).astype(np.int32)
    lists = np.where(rng.random((U, V, K)) < 0.4, rng.integers(0, 60, (U, V, K)), neutral[:, None, :]).astype(np.int32)
    nlen = np.full(U, K, np.int32); lens = rng.integers(0, K + 1, (U, V)).astype(np.int32)
    en = rng.normal(size=(U, D)).astype(np.float32)
    e = (en[:, None] + 0.3 * rng.normal(size=(U, V, D))).astype(np.float32)
    return neutral, nlen, lists, lens, en, e


def _local(cfg, arrs, lo, hi):
    neutral, nlen, lists, lens, en, e = (a[lo:hi] for a in arrs)
    lm = O.list_metrics(neutral, nlen, lists, lens)
    scores = O.cosine_nonconformity_mirror(en, e)
    sums = np.stack([lm["jaccard"].sum(0), lm["serp_num"].sum(0), lm["prag_num"].sum(0),
                     np.full(cfg.n_variants, hi - lo)], 1).astype(np.float64)
    return sums, scores


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = pipeline.PipelineConfig(list_len=8, emb_dim=16, rounds=4)
        V, R = cfg.n_variants, cfg.rounds
        arrs = _inputs()
        lo, hi = pipeline.shard_range(arrs[0].shape[0], rank, world)
        pipe = pipeline.ScoringPipeline(cfg, distributed=True)
        sums, scores = _local(cfg, arrs, lo, hi)

        def all_reduce(h):
            t = torch.from_numpy(h.copy())
            pipe._all_reduce(t)
            return t.numpy()

        thr0, n, k, less, eq = O.radix_select_sharded([scores], cfg.alpha, all_reduce)
        sched = O.threshold_schedule(thr0, n - less - eq, cfg.gamma, R, cfg.q_min)
        flags = np.stack([(scores > sched[r]) for r in range(R)], -1)
        counts = flags.sum((0, 1)); vcounts = flags.sum(0)
        red = torch.from_numpy(np.concatenate([sums.ravel(), counts, vcounts.ravel()]).astype(np.float64))
        pipe._all_reduce(red)
        st = torch.zeros((1, 8), dtype=torch.int64); st[0, 2] = n; st[0, 3] = k
        res = pipeline.StepResult(None, None, None, None, None, None, torch.from_numpy(sched), st, red,
                                  torch.zeros(1, dtype=torch.int32))
        s = pipe.summarize(res)
        torch.save({"thr": sched, "means": s["variant_means"], "counts": s["violation_counts"],
                    "vcounts": s["variant_violation_counts"], "n": s["n_scores"], "k": s["select_rank"]},
                   f"{out}.{rank}")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_rank_shards_reproduce_the_global_result(tmp_path):
    world, port, out = 2, _free_port(), str(tmp_path / "r")
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    got = [torch.load(f"{out}.{r}", weights_only=False) for r in range(world)]
    cfg = pipeline.PipelineConfig(list_len=8, emb_dim=16, rounds=4)
    arrs = _inputs()
    U = arrs[0].shape[0]
    sums, scores = _local(cfg, arrs, 0, U)
    q0 = O.conformal_threshold(scores, cfg.alpha)
    v = O.violation_rounds(scores, q0, cfg.gamma, cfg.rounds, cfg.q_min)
    for g in got:
        assert np.array_equal(g["thr"], v["thresholds"])                    # bit-exact global quantile + schedule
        assert g["counts"].tolist() == v["counts"].tolist()
        assert np.array_equal(g["vcounts"], v["variant_counts"])
        assert g["n"] == scores.size and g["k"] == O.conformal_index(scores.size, cfg.alpha)
        assert np.allclose(g["means"][:, 0], sums[:, 0] / U, rtol=1e-13)
        assert np.array_equal(g["means"][:, 1], sums[:, 1] / U / (8 * 9 / 2.0))   # integer sums: exact
    assert np.array_equal(got[0]["means"], got[1]["means"])
