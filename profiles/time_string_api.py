"""Wall clock of the drop-in string API beside the string-level CPU port (development aid)."""
import os, sys, time, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fairness_llm_replication_b200 import utils as U, phase1_bias_detection as P1
from oracle import ref_port as R

rng = np.random.default_rng(0)
profiles = [{"id": "p%d" % i, "gender": g, "age": a, "occupation": o}
            for i, (g, a, o) in enumerate((g, a, o) for g in "mfn" for a in (18, 25, 35, 45, 56) for o in ("x", "y", "z"))]
recs = {p["id"]: ["Movie %d" % m for m in rng.choice(60, size=10, replace=False)] for p in profiles}
pairs = U.create_profile_pairs(profiles)
for name, fn in (("ours", lambda: U.calculate_individual_fairness(pairs, recs)), ("port", lambda: R.calculate_individual_fairness(pairs, recs))):
    fn(); ts = []
    for _ in range(21):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    print(json.dumps({"case": "45 profiles, %d pairs" % len(pairs), "impl": name, "ms": 1e3 * float(np.median(ts))}))
allrecs = {p["id"]: {"profile_id": p["id"], "recommendations": recs[p["id"]]} for p in profiles}
P1.calculate_phase1_metrics(profiles, allrecs); ts = []
for _ in range(11):
    t0 = time.perf_counter(); P1.calculate_phase1_metrics(profiles, allrecs); ts.append(time.perf_counter() - t0)
print(json.dumps({"case": "phase-1 metric step, 45 profiles", "impl": "ours", "ms": 1e3 * float(np.median(ts))}))
