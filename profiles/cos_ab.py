"""Timing of the cosine nonconformity kernel on one GPU (development aid, not a bench line).

    python profiles/cos_ab.py [--pairs 1000000] [--v 1] [--d 384] [--iters 10] [--lib other/libflr_b200.so]

V = 1 is the calibration form (configs[2]: N pairs of rows), V = 31 the counterfactual form of the bench step.
Times are CUDA events around single launches, L2 flushed before each one.
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
    ap.add_argument("--pairs", type=int, default=1_000_000)
    ap.add_argument("--v", type=int, default=1)
    ap.add_argument("--d", type=int, default=384)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--peak", type=float, default=6546.9)
    ap.add_argument("--lib", default=None)
    a = ap.parse_args()
    if a.lib:
        from fairness_llm_replication_b200 import _lib
        _lib.LIB_PATH = os.path.abspath(a.lib)
    dev = torch.device("cuda", 0)
    users = a.pairs // a.v
    ea, eb = ops.synth_embeddings(20251203, 0, users, a.v, a.d, dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = ops.cosine_nonconformity(ea, eb)
    ref = out.clone()
    times = []
    for _ in range(a.iters + 2):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.cosine_nonconformity(ea, eb, out=out)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = sorted(times[2:])[len(times[2:]) // 2]
    nbytes = users * (a.v + 1) * a.d * 4 + users * a.v * 4
    print(json.dumps({"pairs": users * a.v, "V": a.v, "D": a.d, "ms": round(ms, 4), "GBps": round(nbytes / ms / 1e6, 1),
                      "frac_of_measured_hbm": round(nbytes / ms / 1e6 / a.peak, 4), "identical": bool(torch.equal(out, ref)),
                      "checksum": float(out.double().sum()), "lib": "other" if a.lib else "in-tree"}))


if __name__ == "__main__":
    main()
