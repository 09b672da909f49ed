"""Repeated launches of the list-metric kernels at HBM-resident sizes (warp kernel at K = 50 / 20 / 64, chunked,
few variants) with unrelated traffic on a second stream: every run must reproduce the first bit for bit
(compute-sanitizer is closed on this pool; this is the race hunt at scale -- tests/ holds a smaller one).

    python profiles/stress_list_metrics.py
"""
import sys, torch
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fairness_llm_replication_b200 import ops
dev = torch.device("cuda", 0)
bad = 0
for (U, V, K) in [(300000, 31, 50), (400000, 31, 20), (100000, 64, 50), (300000, 8, 20), (200000, 31, 64)]:
    n, nl, l, ln = ops.synth_lists(20251203, 0, U, V, K, 3706, dev)
    ref = ops.list_metrics(n, nl, l, ln, vocab=3706, want_union=True)
    torch.cuda.synchronize()
    keep = [t.clone() for t in (ref.inter, ref.union, ref.jaccard, ref.serp, ref.prag, ref.variant_sums, ref.status)]
    side = torch.cuda.Stream()
    junk = torch.empty(64 << 20, dtype=torch.uint8, device=dev)
    for it in range(60):
        with torch.cuda.stream(side):
            junk.fill_(it & 255)                       # unrelated traffic on another stream
        r = ops.list_metrics(n, nl, l, ln, vocab=3706, want_union=True)
        cur = (r.inter, r.union, r.jaccard, r.serp, r.prag, r.variant_sums, r.status)
        if not all(torch.equal(a, b) for a, b in zip(keep, cur)):
            bad += 1
    torch.cuda.synchronize()
    print((U, V, K), "mismatching runs:", bad)
print("STRESS", "FAILED" if bad else "ok")
