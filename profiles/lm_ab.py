"""A/B timing of the list-metric kernels on one GPU (development aid, not a bench line).

    python profiles/lm_ab.py [--users 300000] [--k 50] [--variants reg,warp] [--lib other/libflr_b200.so]

FLR_LM_KERNEL=reg selects the previous register kernel, anything else the current default.
Outputs of every variant are compared bit for bit with the first one; times are CUDA events over
`--iters` launches after a warm-up (the inputs are far larger than L2).
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fairness_llm_replication_b200 import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--users", type=int, default=300000)
    ap.add_argument("--k", type=int, default=50)
    ap.add_argument("--v", type=int, default=31)
    ap.add_argument("--vocab", type=int, default=3706)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--variants", default="reg,tas")
    ap.add_argument("--peak", type=float, default=6546.9)
    ap.add_argument("--lib", default=None, help="another build of libflr_b200.so to time (same-board A/B of two builds)")
    a = ap.parse_args()
    if a.lib:
        from fairness_llm_replication_b200 import _lib
        _lib.LIB_PATH = os.path.abspath(a.lib)
    dev = torch.device("cuda", 0)
    neutral, nlen, lists, lens = ops.synth_lists(20251203, 0, a.users, a.v, a.k, a.vocab, dev)
    pairs = a.users * a.v
    bytes_per_pair = a.k * 4 * (a.v + 1) / a.v + 4 * (a.v + 1) / a.v + 16
    first = None
    for name in a.variants.split(","):
        os.environ["FLR_LM_KERNEL"] = name
        res = ops.list_metrics(neutral, nlen, lists, lens, vocab=a.vocab, want_union=True)
        torch.cuda.synchronize()
        cur = [res.inter, res.union, res.jaccard, res.serp, res.prag, res.variant_sums, res.status]
        if first is None:
            first = [t.clone() for t in cur]
            same = True
        else:
            same = all(torch.equal(x, y) for x, y in zip(first, cur))
        for _ in range(3):
            ops.list_metrics(neutral, nlen, lists, lens, vocab=a.vocab)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(a.iters):
            ops.list_metrics(neutral, nlen, lists, lens, vocab=a.vocab)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.iters
        gbs = pairs * bytes_per_pair / (ms * 1e-3) / 1e9
        print(json.dumps({"variant": name, "users": a.users, "K": a.k, "ms": round(ms, 4),
                          "GBps": round(gbs, 1), "frac_of_measured_hbm": round(gbs / a.peak, 4),
                          "identical_to_first": same, "lib": os.path.basename(os.path.dirname(os.path.abspath(a.lib))) if a.lib else "in-tree"}))


if __name__ == "__main__":
    main()
