"""The bench step (list metrics, cosine, select, schedule, violations) over the 1.25 M-user shard, timed per
kernel with CUDA events, for one build of the library -- to compare two builds back to back on one board,
power cap included (development aid, not a bench line).

    python profiles/step_ab.py [--users 1250000] [--steps 40] [--warmup 5] [--lib other/libflr_b200.so]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fairness_llm_replication_b200 import pipeline  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--users", type=int, default=1_250_000)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--lib", default=None)
    a = ap.parse_args()
    if a.lib:
        from fairness_llm_replication_b200 import _lib
        _lib.LIB_PATH = os.path.abspath(a.lib)
    dev = torch.device("cuda", 0)
    cfg = pipeline.PipelineConfig(list_len=50)
    batch = pipeline.synth_batch(cfg, 0, a.users, dev)
    pipe = pipeline.ScoringPipeline(cfg)
    for _ in range(a.warmup):
        pipe.run(batch)
    torch.cuda.synchronize()
    pipe.kernel_events = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        pipe.run(batch)
    e1.record()
    torch.cuda.synchronize()
    per = {}
    for name, s, e in pipe.kernel_events:
        per.setdefault(name, []).append(s.elapsed_time(e))
    out = {"lib": "other" if a.lib else "in-tree", "users": a.users, "steps": a.steps,
           "ms_per_step": round(e0.elapsed_time(e1) / a.steps, 4)}
    for name, v in per.items():
        out[name] = round(sum(v) / len(v), 4)
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(0)
        out["sm_mhz_after"] = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
    except Exception:
        pass
    print(json.dumps(out))


if __name__ == "__main__":
    main()
