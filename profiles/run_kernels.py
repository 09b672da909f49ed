"""Launch each kernel of the library once at a benchmark-like size, for `ncu --set full` captures
(profiles/README.md has the command).  Not a bench: nothing here is timed."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fairness_llm_replication_b200 import ops, pipeline  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    cfg = pipeline.PipelineConfig(list_len=50)
    U = int(os.environ.get("FLR_PROFILE_USERS", "200000"))
    # the step: list metrics, cosine (+ pass-0 histogram), select passes, threshold schedule, violations
    batch = pipeline.synth_batch(cfg, 0, U, dev)
    pipe = pipeline.ScoringPipeline(cfg)
    pipe.run(batch)
    torch.cuda.synchronize()
    # N1: dense group histogram + JS distances on the same lists
    hist = ops.group_item_histogram(batch.lists, batch.lens, vocab=cfg.vocab, n_groups=31)
    ops.js_distance_matrix(hist)
    # a9 with 31 groups (the reference's per-group thresholds), f32
    N = 1_000_000
    g = torch.Generator(device=dev); g.manual_seed(1)
    sc = torch.rand(N, device=dev, generator=g)
    grp = torch.randint(0, 31, (N,), device=dev, generator=g, dtype=torch.int32).to(torch.uint8)
    ops.radix_select(sc, alpha=0.1, groups=grp, n_groups=31)
    del batch
    # config 4: 6040 users x 124 lists / embeddings
    Uc, M, K, D = 6040, 124, 20, 384
    _, _, l4, n4 = ops.synth_lists(pipeline.SEED, 0, Uc, M, K, 3706, dev)
    _, e4 = ops.synth_embeddings(pipeline.SEED, 0, Uc, M, D, dev)
    xh = ops.normalize_rows_bf16(e4)
    ops.pair_similarity(xh)
    ops.pair_jaccard(l4, n4, vocab=3706)
    # K = 20 list metrics at an HBM-resident size
    n2, nl2, l2, ln2 = ops.synth_lists(pipeline.SEED, 0, 500000, 31, 20, 3706, dev)
    ops.list_metrics(n2, nl2, l2, ln2, vocab=3706)
    del n2, nl2, l2, ln2
    # config 3: the calibration form of the cosine kernel, 1M pairs of 384-d rows
    ea, eb = ops.synth_embeddings(pipeline.SEED, 0, 1_000_000, 1, 384, dev)
    ops.cosine_nonconformity(ea, eb)
    del ea, eb
    # more than 32 variants per user: the warp kernel once per chunk of 32 (two launches here)
    n3, nl3, l3, ln3 = ops.synth_lists(pipeline.SEED, 0, 100000, 64, 50, 3706, dev)
    ops.list_metrics(n3, nl3, l3, ln3, vocab=3706)
    torch.cuda.synchronize()
    print("done")


if __name__ == "__main__":
    main()
