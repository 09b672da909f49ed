import sys, torch
sys.path.insert(0, ".")
from fairness_llm_replication_b200 import ops
from torch.profiler import profile, ProfilerActivity
for N in (6_200_000, 38_750_000):
    s = torch.rand(N, device="cuda") * 0.3
    thr = torch.tensor([0.27, 0.25, 0.24, 0.23, 0.2], device="cuda")
    for _ in range(3): f, c, vc = ops.violation_detect(s, thr, rounds=4, n_variants=31)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(10): f, c, vc = ops.violation_detect(s, thr, rounds=4, n_variants=31)
        torch.cuda.synchronize()
    for e in prof.key_averages():
        if "violation" in e.key: print(N, e.key[:40], "%.1f us avg" % (e.device_time_total / e.count))
    assert int(c[0]) == int((s > 0.27).sum()) and int(vc.sum()) == int(c.sum())
# cosine, calibration form
for N in (1_000_000,):
    a, b = ops.synth_embeddings(20251203, 0, N, 1, 384, "cuda:0")
    b2 = b[:, 0].contiguous()
    for _ in range(3): sc = ops.cosine_nonconformity(a, b2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): sc = ops.cosine_nonconformity(a, b2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("cosine V=1 N=%d: %.3f ms %.0f GB/s" % (N, ms, N * 3076 / ms / 1e6))
