"""Times the config-4 kernels (6040 users x 124 rows x 384 dims) with CUDA events, L2 flushed between runs."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fairness_llm_replication_b200 import ops, pipeline

dev = torch.device("cuda", 0)
U, M, K, D = 6040, 124, 20, 384
_, _, l4, n4 = ops.synth_lists(pipeline.SEED, 0, U, M, K, 3706, dev)
_, e4 = ops.synth_embeddings(pipeline.SEED, 0, U, M, D, dev)
xh = ops.normalize_rows_bf16(e4)
flush = torch.zeros(128 * 1024 * 1024, dtype=torch.float32, device=dev)

def timed(fn, reps=20):
    for _ in range(3): fn()
    ts = []
    for _ in range(reps):
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))

peak = 6546.9
ms = timed(lambda: ops.pair_similarity(xh)); by = U * (M * D * 2 + M * M * 4)
print(json.dumps({"kernel": "pair_similarity_kernel", "ms": ms, "gbs": by / ms / 1e6, "frac": by / ms / 1e6 / peak}))
ms = timed(lambda: ops.pair_jaccard(l4, n4, vocab=3706)); by = U * (M * K * 4 + M * 4 + M * M * 4 + M * 4)
print(json.dumps({"kernel": "pair_jaccard_kernel", "ms": ms, "gbs": by / ms / 1e6, "frac": by / ms / 1e6 / peak}))
ms = timed(lambda: ops.normalize_rows_bf16(e4)); by = U * M * D * 6
print(json.dumps({"kernel": "normalize_rows_bf16_kernel", "ms": ms, "gbs": by / ms / 1e6, "frac": by / ms / 1e6 / peak}))
