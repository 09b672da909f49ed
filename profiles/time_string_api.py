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

# the whole phase-1 metric step on a results-file-sized input (as bench.py's other_configs does)
titles = ["Title %d (19%02d)" % (i, i % 100) for i in range(300)]
w = 1.0 / np.arange(1, 301)
big_profiles = [{"id": "q%d" % i, "gender": g, "age": a, "occupation": "professional"}
                for i, (g, a, _) in enumerate((g, a, r) for g in ("male", "female", "non-binary")
                                              for a in ("18-24", "25-34", "35-44", "45-54", "55+") for r in range(67))]
big_recs = {p["id"]: {"profile_id": p["id"], "recommendations":
                      [titles[t] for t in rng.choice(300, size=10, replace=False, p=w / w.sum())]} for p in big_profiles}
P1.calculate_phase1_metrics(big_profiles[:30], big_recs)
ts = []
for _ in range(5):
    t0 = time.perf_counter(); m = P1.calculate_phase1_metrics(big_profiles, big_recs); ts.append(time.perf_counter() - t0)
t0 = time.perf_counter(); ref = R.measure_individual_fairness(big_profiles, big_recs); t_port_if = time.perf_counter() - t0
print(json.dumps({"case": "phase-1 metric step, 1005 profiles, %d pairs" % m["individual_fairness"]["details"]["num_pairs"],
                  "ours_s": float(np.median(ts)), "port_individual_fairness_alone_s": t_port_if,
                  "identical": bool(float(m["individual_fairness"]["score"]) == float(ref[0])
                                    and m["individual_fairness"]["details"]["similarity_scores"] == ref[1]["similarity_scores"])}))
