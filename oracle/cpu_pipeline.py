"""The whole scoring pipeline on host cores, from the oracle's numpy restatement.
TEST INFRASTRUCTURE / CPU BASELINE ONLY (see ``oracle/__init__.py``): used by
``bench.py``'s ``cpu_baseline`` leg and its ``--impl reference`` arm, and by smoke().

The reference itself is single-threaded CPython over string sets
(src/utils.py:224-239) and has no code at all for SERP*/PRAG*/cosine/violations, and
/root/reference does not exist on the GPU box; so the timed CPU arm is this port, vectorised
and spread over all host cores (``kind: "port"``).
"""
from __future__ import annotations

import multiprocessing as mp
import os
import time

import numpy as np

from . import np_oracle as O

_G = {}


def _score_slice(span):
    lo, hi = span
    a = _G["arrays"]
    cfg = _G["cfg"]
    neutral, nlen, lists, lens, en, e = (x[lo:hi] for x in a)
    lm = O.list_metrics(neutral, nlen, lists, lens, k_cap=cfg["k_cap"], chunk_users=256)
    scores = O.cosine_nonconformity(en, e).astype(np.float32)
    sums = np.stack([lm["jaccard"].sum(0), lm["serp_num"].sum(0), lm["prag_num"].sum(0)], 1)
    return sums, scores


def run(arrays, k_cap=None, alpha=0.1, gamma=0.95, rounds=4, q_min=0.0, workers=None):
    """Score one batch (numpy arrays: neutral, neutral_len, lists, lens, emb_neutral, emb).
    Returns (summary dict, seconds of scoring time)."""
    U = arrays[0].shape[0]
    workers = max(1, min(workers or (os.cpu_count() or 1), U))
    _G["arrays"] = arrays
    _G["cfg"] = {"k_cap": k_cap}
    bounds = np.linspace(0, U, workers * 4 + 1).astype(int)
    spans = [(int(a), int(b)) for a, b in zip(bounds[:-1], bounds[1:]) if b > a]
    if workers == 1:
        t0 = time.perf_counter()
        parts = [_score_slice(s) for s in spans]
    else:
        ctx = mp.get_context("fork")                 # children inherit the arrays copy-free
        with ctx.Pool(workers) as pool:
            pool.map(_noop, range(workers))          # fork + import cost stays outside the timing
            t0 = time.perf_counter()
            parts = pool.map(_score_slice, spans)
    sums = sum(p[0] for p in parts)
    scores = np.concatenate([p[1] for p in parts], 0)
    q0 = O.conformal_threshold(scores, alpha)
    v = O.violation_rounds(scores, q0, gamma, rounds, q_min)
    dt = time.perf_counter() - t0
    return {"sums": sums, "thresholds": v["thresholds"], "counts": v["counts"], "scores": scores,
            "flags": v["flags"]}, dt


def _noop(_):
    return 0
