#!/usr/bin/env python
"""Benchmark of the scoring hot path (BASELINE.json: user-variant pairs/s, list metrics +
conformal scoring).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step is one pass of the full pipeline (list metrics -> cosine nonconformity -> exact global
quantile -> threshold schedule -> violation flags) over one shard of synthetic users of the
10M-user sweep (BASELINE.json configs[4]; configs[1] is 15 MB and launch-bound, SURVEY.md
section 7): 1.25M users x 31 variants per GPU, top-50 lists, 384-d fp32 embeddings, i.e.
exactly the 10M-user job at 8 GPUs and its per-GPU shard below that (weak scaling).

Prints ONE JSON line (rank 0).  `value` is device-resident throughput, `e2e` is the same
pipeline called with pinned HOST buffers (H2D and the D2H of the summary inside the timing).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "user_variant_pairs_per_sec"
UNIT = "pairs/s"
FALLBACK_HBM_GBS = 6650.0          # /opt/skills/guides/B200_PROFILING.md fallback


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--users-per-gpu", type=int, default=1_250_000)
    ap.add_argument("--e2e-users", type=int, default=65_536)
    ap.add_argument("--cpu-sample-users", type=int, default=16384, help="users per CPU-port batch")
    ap.add_argument("--ref-sample-users", type=int, default=4096, help="users per step of the reference arm")
    ap.add_argument("--parity-users", type=int, default=8192, help="users per rank of the multi-GPU parity check")
    ap.add_argument("--cpu-sample-seconds", type=float, default=10.0,
                    help="cpu_baseline keeps scoring batches until this much scoring time has accumulated")
    ap.add_argument("--list-len", type=int, default=50)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true")
    return ap.parse_args()


def workload_config(args, n_gpus):
    return {
        "workload": "cfg5-shard: 10M-user sweep sharded by users (1.25M users/GPU at the default), "
                    "31 variants, top-%d lists over 3706 items, 384-d fp32 embeddings, full pipeline" % args.list_len,
        "users_per_gpu": args.users_per_gpu, "n_variants": 31, "list_len": args.list_len, "emb_dim": 384,
        "vocab": 3706, "alpha": 0.1, "gamma": 0.95, "rounds": 4,
        "parallelism": "user-sharded dp%d" % n_gpus,
        "e2e_users_per_gpu_per_step": min(args.e2e_users, args.users_per_gpu),
        "e2e_note": "the e2e leg streams a smaller batch per step than `value` scores (it is PCIe-bound, so its "
                    "rate does not depend on the batch size; a 1.25M-user batch would pin 70 GB of host memory per rank)",
        "l2": "inputs (%.1f GB per GPU) are larger than the 126 MB L2; no flush needed"
              % (args.users_per_gpu * (32 * args.list_len * 4 + 32 * 4 + 32 * 384 * 4) / 1e9),
    }


# ---------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop = [], set(), None, threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            self.ok = False
        self.thread = threading.Thread(target=self._loop, daemon=True)

    def _loop(self):
        while not self._stop.is_set():
            try:
                self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
            except Exception:
                pass
            try:
                mask = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.ok:
            self.thread.start()

    def stop(self):
        self._stop.set()
        if self.ok:
            self.thread.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel, pairs):
    """DRAM bytes per launch of `kernel` (dram__bytes_read.sum + dram__bytes_write.sum) from the committed
    `ncu --set full` capture, scaled from the captured launch's pair count to this launch's."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        per_pair = d["dram_bytes_per_pair"]
        key = kernel if kernel in per_pair else kernel.split("<")[0]
        return per_pair[key] * pairs
    except Exception:
        return None


def bind_to_gpu_numa_node(local):
    """Keep this rank on the CPUs next to its GPU's PCIe root, so that the pinned host buffers of the e2e
    leg are first touched -- and therefore allocated -- on that NUMA node.  Returns a short description."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)
        name = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        with open("/sys/bus/pci/devices/%s/local_cpulist" % name) as f:
            text = f.read().strip()
        cpus = set()
        for part in text.split(","):
            if part:
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0)
        target = cpus & allowed
        if target and target != allowed:
            os.sched_setaffinity(0, target)
            return "gpu %s: bound to %d of %d allowed cpus (%s)" % (name, len(target), len(allowed), text)
        return "gpu %s: no narrower cpu set (local %s, allowed %d cpus)" % (name, text, len(allowed))
    except Exception as exc:                                  # no sysfs entry, no permission: run unbound
        return "unbound (%s)" % exc


# ---------------------------------------------------------------------------------- CPU arms
# Nothing below imports the product package: inputs come from oracle/synth_ref (plain C, byte-identical
# to the product's generator -- tests/test_oracle_golden.py), so a reference-arm process never maps
# libflr_b200.so.
def cpu_sample(args, n_users, user0=0):
    """Time the oracle pipeline (oracle/cpu_pipeline.py) on a bounded sample, all host cores."""
    from oracle import cpu_pipeline, synth_ref
    V, K, D = 31, args.list_len, 384
    lists = synth_ref.lists(20251203, user0, n_users, V, K, 3706)
    embs = synth_ref.embeddings(20251203, user0, n_users, V, D)
    cores = os.cpu_count() or 1
    _, dt = cpu_pipeline.run((*lists, *embs), workers=cores)
    return n_users * V / dt, dt, cores


def reference_modules():
    """The reference's own utils / phase3_facter_mitigation, staged by oracle/build_ref.py; None if absent."""
    from oracle import ref_loader
    if not ref_loader.available(ref_loader.STAGED_SRC):
        return None
    return (ref_loader.load("utils", ref_loader.STAGED_SRC),
            ref_loader.load("phase3_facter_mitigation", ref_loader.STAGED_SRC))


def reference_sample(args, mods, n_users, user0=0):
    """Time the UNMODIFIED reference on a bounded sample of the workload, one CPython thread (what it is):
    utils.calculate_individual_fairness (src/utils.py:217-244) over the sample's (user, variant) pairs as
    title-string lists, then calculate_conformal_thresholds (src/phase3_facter_mitigation.py:152-182) over
    the sample's nonconformity scores.  The reference has no SERP*/PRAG*/cosine/violation code (SURVEY.md
    Fact 2): the scores it thresholds are computed by the numpy oracle outside the timing, and the parts it
    lacks are simply not in its time -- which flatters the reference, never this repository."""
    import pandas as pd
    from oracle import np_oracle, synth_ref
    ref_utils, ref_p3 = mods
    V, K, D = 31, args.list_len, 384
    neutral, nlen, lists, lens = synth_ref.lists(20251203, user0, n_users, V, K, 3706)
    en, e = synth_ref.embeddings(20251203, user0, n_users, V, D)
    recs, pairs = {}, []
    for u in range(n_users):
        recs["u%d" % u] = ["m%d" % i for i in neutral[u, :nlen[u]]]
        for v in range(V):
            pid = "u%d_%d" % (u, v)
            recs[pid] = ["m%d" % i for i in lists[u, v, :lens[u, v]]]
            pairs.append(("u%d" % u, pid))
    scores = np_oracle.cosine_nonconformity(en, e).astype(np.float32).ravel()
    calib = pd.DataFrame({"profile_id": [p[1] for p in pairs], "group": "all", "nonconformity": scores})
    t0 = time.perf_counter()
    ref_utils.calculate_individual_fairness(pairs, recs)
    t1 = time.perf_counter()
    ref_p3.calculate_conformal_thresholds(calib, alpha=0.1)
    t2 = time.perf_counter()
    return len(pairs) / (t2 - t0), t2 - t0, {"jaccard_s": t1 - t0, "threshold_s": t2 - t1}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import logging
    logging.disable(logging.INFO)                      # the reference logs every threshold dict
    cfg = workload_config(args, args.gpus)
    steps, warm = max(1, args.steps), max(0, args.warmup)
    mods = reference_modules()
    # bounded: about a second of CPU work per step
    if mods is not None:
        n = args.ref_sample_users
        for _ in range(min(warm, 1)):
            reference_sample(args, mods, min(n, 64))
        times, pairs, split = [], 0, {"jaccard_s": 0.0, "threshold_s": 0.0}
        for s in range(steps):
            _, dt, parts = reference_sample(args, mods, n, user0=s * n)
            times.append(dt)
            pairs += n * 31
            for k in split:
                split[k] += parts[k]
        value, kind, cores = pairs / sum(times), "reference", 1
        sample = ("%d users x 31 variants per step through the reference's own calculate_individual_fairness "
                  "(title-string sets) and calculate_conformal_thresholds; SERP*/PRAG*/cosine/violations have "
                  "no reference code and are not in its time (jaccard %.1f s, thresholds %.1f s in all)"
                  % (n, split["jaccard_s"], split["threshold_s"]))
        note = "unmodified reference modules staged in oracle/_ref by oracle/build_ref.py, one CPython thread"
        _, dt_port, cores_port = cpu_sample(args, args.cpu_sample_users)
        port = {"value": args.cpu_sample_users * 31 / dt_port, "unit": UNIT, "cores": cores_port, "kind": "port",
                "sample": "%d users x 31 variants, full pipeline incl. SERP*/PRAG*/cosine/violations, numpy port"
                          % args.cpu_sample_users}
    else:
        n = args.cpu_sample_users
        for _ in range(min(warm, 1)):
            cpu_sample(args, min(n, 512))
        times, pairs, cores = [], 0, 1
        for s in range(steps):
            _, dt, cores = cpu_sample(args, n, user0=s * n)
            times.append(dt)
            pairs += n * 31
        value, kind, port = pairs / sum(times), "port", None
        sample = "%d users x 31 variants per step (full pipeline: list metrics + cosine + quantile + violations)" % n
        note = "oracle/_ref is not staged (run oracle/build_ref.py where /root/reference exists): numpy oracle port"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": 1e3 * sum(times) / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
                         "port_all_cores": port},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "note": note,
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------- other configs
def other_configs(dev, peak):
    """BASELINE.json configs[1], [2], [3] timed briefly on one GPU (CUDA events, after warm-up).
    They are parity-test cases, not the headline; the numbers are extra evidence only."""
    import torch
    from fairness_llm_replication_b200 import ops, pipeline

    def timed(fn, reps=10, flush=None):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            if flush is not None:
                flush.add_(1.0)                      # 512 MB write: evicts the 126 MB L2
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    out = {}
    # configs[0]: the reference's own scale -- 45 profiles, K = 10 title strings, 405 pairs -- through the
    # string API (host encoding + one kernel launch) beside the string-level CPU port of the same function
    from fairness_llm_replication_b200 import utils as fast_utils
    from oracle import ref_port
    rng = np.random.default_rng(0)
    profiles = [{"id": "p%d" % i, "gender": g, "age": a, "occupation": o}
                for i, (g, a, o) in enumerate((g, a, o) for g in "mfn" for a in (18, 25, 35, 45, 56) for o in ("x", "y", "z"))]
    recs = {p_["id"]: ["Movie %d" % m for m in rng.choice(60, size=10, replace=False)] for p_ in profiles}
    pairs = fast_utils.create_profile_pairs(profiles)
    ours, port = [], []
    for _ in range(7):
        t0 = time.perf_counter(); r_ours = fast_utils.calculate_individual_fairness(pairs, recs); ours.append(time.perf_counter() - t0)
        t0 = time.perf_counter(); r_port = ref_port.calculate_individual_fairness(pairs, recs); port.append(time.perf_counter() - t0)
    out["cfg1_string_api_45_profiles"] = {
        "pairs": len(pairs), "ours_ms": 1e3 * float(np.median(ours)), "cpu_port_ms": 1e3 * float(np.median(port)),
        "identical": bool(r_ours[0] == r_port[0] and list(r_ours[1]) == list(r_port[1])),
        "note": "wall clock incl. string->id encoding, H2D, launch, D2H; at this size the host work is all there is"}
    # the whole metric step of phase 1 (src/phase1_bias_detection.py:343-405) through the drop-in functions on a
    # results-file-sized input: 1005 profiles (3 genders x 5 ages x 67), top-10 title lists; beside the same
    # step composed from the string-level CPU port
    from fairness_llm_replication_b200 import phase1_bias_detection as fast_p1
    titles = ["Title %d (19%02d)" % (i, i % 100) for i in range(300)]
    w = 1.0 / np.arange(1, 301)
    big_profiles = [{"id": "q%d" % i, "gender": g, "age": a, "occupation": "professional"}
                    for i, (g, a, _) in enumerate((g, a, r) for g in ("male", "female", "non-binary")
                                                  for a in ("18-24", "25-34", "35-44", "45-54", "55+") for r in range(67))]
    big_recs = {p_["id"]: {"profile_id": p_["id"], "recommendations":
                           [titles[t] for t in rng.choice(300, size=10, replace=False, p=w / w.sum())]}
                for p_ in big_profiles}
    fast_p1.calculate_phase1_metrics(big_profiles[:30], big_recs)              # warm-up (module load, workspace)
    t0 = time.perf_counter(); m_ours = fast_p1.calculate_phase1_metrics(big_profiles, big_recs); t_ours = time.perf_counter() - t0
    t0 = time.perf_counter()
    by_gender, by_age = {}, {}
    for p_ in big_profiles:
        by_gender.setdefault(p_["gender"], []).append(big_recs[p_["id"]]["recommendations"])
        by_age.setdefault(p_["age"], []).append(big_recs[p_["id"]]["recommendations"])
    dp_g = ref_port.calculate_demographic_parity(by_gender)[0]
    dp_a = ref_port.calculate_demographic_parity(by_age)[0]
    if_port = ref_port.measure_individual_fairness(big_profiles, big_recs)
    eo_port = ref_port.calculate_equal_opportunity(by_gender, fast_p1.QUALIFIED_MOVIES)[0]
    t_port = time.perf_counter() - t0
    out["string_api_phase1_step_1005_profiles"] = {
        "profiles": len(big_profiles), "pairs": int(m_ours["individual_fairness"]["details"]["num_pairs"]),
        "ours_s": t_ours, "cpu_port_s": t_port, "speedup": t_port / t_ours,
        "identical_individual_fairness": bool(float(m_ours["individual_fairness"]["score"]) == float(if_port[0])),
        "demographic_parity_rel_diff": abs(float(m_ours["demographic_parity"]["gender"]["score"]) - float(dp_g)) / float(dp_g),
        "note": "wall clock of calculate_phase1_metrics (pairs, 2 x demographic parity, individual fairness, equal "
                "opportunity; lists encoded and uploaded once) vs the same step on the string-level CPU port"}
    flush = torch.zeros(128 * 1024 * 1024, dtype=torch.float32, device=dev)
    # configs[1]: 6040 users x 31 variants, K = 20, all phase-1 metrics (15 MB: launch/latency-bound)
    cfg = pipeline.PipelineConfig(list_len=20)
    neutral, nlen, lists, lens = ops.synth_lists(pipeline.SEED, 0, 6040, 31, 20, 3706, dev)
    f = lambda: ops.list_metrics(neutral, nlen, lists, lens, vocab=3706)
    ms_cold, ms_warm = timed(f, flush=flush), timed(f)
    out["cfg2_phase1_metrics_6040x31_k20"] = {
        "pairs": 6040 * 31, "ms_l2_flushed": ms_cold, "ms_l2_warm": ms_warm, "pairs_per_s": 6040 * 31 / (ms_cold * 1e-3),
        "note": "15 MB working set: bounded by launch + tile latency, not HBM (SURVEY.md section 7)"}
    del neutral, nlen, lists, lens
    # the same K = 20 shape at a size that is HBM-resident, not L2-resident (1M users, 3.2 GB of lists)
    Ub = 1_000_000
    neutral, nlen, lists, lens = ops.synth_lists(pipeline.SEED, 0, Ub, 31, 20, 3706, dev)
    ms_big = timed(lambda: ops.list_metrics(neutral, nlen, lists, lens, vocab=3706), reps=5)
    k20_bytes = Ub * 31 * (20 * 4 * 32 / 31 + 4 * 32 / 31 + 16)
    out["list_metrics_k20_1M_users"] = {
        "pairs": Ub * 31, "ms": ms_big, "pairs_per_s": Ub * 31 / (ms_big * 1e-3), "gbs": k20_bytes / (ms_big * 1e-3) / 1e9,
        "frac_of_measured_hbm": k20_bytes / (ms_big * 1e-3) / 1e9 / peak}
    del neutral, nlen, lists, lens
    # configs[2]: 1M calibration pairs of 384-d fp32 embeddings: cosine + exact quantile + violations
    N = 1_000_000
    ea, eb = ops.synth_embeddings(pipeline.SEED, 0, N, 1, 384, dev)
    def conformal():
        h = torch.zeros(2048, dtype=torch.int64, device=dev)
        sc = ops.cosine_nonconformity(ea, eb, hist0=h)
        thr0, st = ops.radix_select(sc, alpha=0.1, hist0=h)
        thr = ops.threshold_schedule(thr0, st, rounds=4)
        return ops.violation_detect(sc, thr, rounds=4, n_variants=1, per_variant=False)
    ms = timed(conformal, flush=flush)
    ms_cos = timed(lambda: ops.cosine_nonconformity(ea, eb), flush=flush)
    out["cfg3_conformal_1M_pairs"] = {
        "pairs": N, "ms": ms, "pairs_per_s": N / (ms * 1e-3), "cosine_ms": ms_cos,
        "cosine_gbs": N * 3076 / (ms_cos * 1e-3) / 1e9, "cosine_frac_of_measured_hbm": N * 3076 / (ms_cos * 1e-3) / 1e9 / peak}
    del ea, eb
    # configs[3]: 6040 users x (4 models x 31 variants) = 124 lists / embeddings per user
    U, M, K, D = 6040, 124, 20, 384
    _, _, l4, n4 = ops.synth_lists(pipeline.SEED, 0, U, M, K, 3706, dev)
    _, e4 = ops.synth_embeddings(pipeline.SEED, 0, U, M, D, dev)
    xh = ops.normalize_rows_bf16(e4)
    ms_gemm = timed(lambda: ops.pair_similarity(xh), flush=flush)
    ms_norm = timed(lambda: ops.normalize_rows_bf16(e4), flush=flush)
    ms_pj = timed(lambda: ops.pair_jaccard(l4, n4, vocab=3706), flush=flush)
    gemm_bytes = U * (M * D * 2 + M * M * 4)
    out["cfg4_cross_model_6040x124"] = {
        "users": U, "gemm_ms": ms_gemm, "gemm_gbs": gemm_bytes / (ms_gemm * 1e-3) / 1e9,
        "gemm_frac_of_measured_hbm": gemm_bytes / (ms_gemm * 1e-3) / 1e9 / peak,
        "gemm_tflops": U * 2.0 * M * M * D / (ms_gemm * 1e-3) / 1e12,
        "normalize_ms": ms_norm, "pair_jaccard_ms": ms_pj,
        "pair_jaccard_list_pairs_per_s": U * M * (M - 1) / 2 / (ms_pj * 1e-3),
        "normalize_gbs": U * M * D * 6 / (ms_norm * 1e-3) / 1e9,
        "normalize_frac_of_measured_hbm": U * M * D * 6 / (ms_norm * 1e-3) / 1e9 / peak,
        "pair_jaccard_gbs": U * (M * K * 4 + M * 4 + M * M * 4 + M * 4) / (ms_pj * 1e-3) / 1e9,
        "pair_jaccard_frac_of_measured_hbm": U * (M * K * 4 + M * 4 + M * M * 4 + M * 4) / (ms_pj * 1e-3) / 1e9 / peak,
        "note": "GEMM is HBM-bound (75 flop/B): tensor-pipe utilisation is low by construction"}
    del l4, n4, e4, xh
    # a9 with the reference's per-group semantics (src/phase3_facter_mitigation.py:159-178): 1M scores in 31 groups
    N = 1_000_000
    g = torch.Generator(device=dev); g.manual_seed(1)
    sc32 = torch.rand(N, device=dev, generator=g)
    grp = torch.randint(0, 31, (N,), device=dev, generator=g, dtype=torch.int32).to(torch.uint8)
    sc64 = sc32.double()
    ms32 = timed(lambda: ops.radix_select(sc32, alpha=0.1, groups=grp, n_groups=31), flush=flush)
    ms64 = timed(lambda: ops.radix_select(sc64, alpha=0.1, groups=grp, n_groups=31), flush=flush)
    out["a9_select_per_group_1M_x_31_groups"] = {
        "f32_ms": ms32, "f32_gbs": 3 * N * 5 / (ms32 * 1e-3) / 1e9, "f32_frac_of_measured_hbm": 3 * N * 5 / (ms32 * 1e-3) / 1e9 / peak,
        "f64_ms": ms64, "f64_gbs": 6 * N * 9 / (ms64 * 1e-3) / 1e9, "f64_frac_of_measured_hbm": 6 * N * 9 / (ms64 * 1e-3) / 1e9 / peak,
        "note": "whole select (3 passes f32 / 6 passes f64: histogram + scan each, torch allocations included); "
                "5 MB of scores + group ids: latency-bound, the fraction of HBM is for the record"}
    del sc32, sc64, grp
    # N1: demographic-parity item histogram over the bench shard's lists (dense [U, V, K] layout)
    Uh = 1_250_000
    _, _, lh, nh = ops.synth_lists(pipeline.SEED, 0, Uh, 31, 50, 3706, dev)
    ms_h = timed(lambda: ops.group_item_histogram(lh, nh, vocab=3706, n_groups=31), reps=5)
    hb_bytes = Uh * 31 * (50 * 4 + 4)
    out["n1_group_item_histogram_1250k_users"] = {
        "rows": Uh * 31, "ms": ms_h, "gbs": hb_bytes / (ms_h * 1e-3) / 1e9,
        "frac_of_measured_hbm": hb_bytes / (ms_h * 1e-3) / 1e9 / peak, "kernel": "dense_group_histogram_kernel"}
    hist = ops.group_item_histogram(lh, nh, vocab=3706, n_groups=31)
    out["n1_js_distance_matrix_31_groups_ms"] = timed(lambda: ops.js_distance_matrix(hist))
    del lh, nh
    return out

# ---------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from fairness_llm_replication_b200 import ops, pipeline

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the scoring path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(local)
    distributed = world > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not distributed:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    cfg = pipeline.PipelineConfig(list_len=args.list_len)
    V, K, D = cfg.n_variants, cfg.list_len, cfg.emb_dim
    U = args.users_per_gpu
    batch = pipeline.synth_batch(cfg, rank * U, U, dev)
    torch.cuda.synchronize()
    pipe = pipeline.ScoringPipeline(cfg, distributed=distributed)
    steps, warm = max(1, args.steps), max(3, args.warmup)

    for _ in range(warm):
        res = pipe.run(batch)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    pipe.kernel_events = []
    launches0 = pipe.n_launches
    barrier()
    ev[0].record()
    for _ in range(steps):
        res = pipe.run(batch)
    ev[1].record()
    barrier()
    clocks = sampler.stop()
    ms_total = max_over_ranks(ev[0].elapsed_time(ev[1]))
    launches = pipe.n_launches - launches0
    kernel_ms = {}
    for name, a, b in pipe.kernel_events:
        kernel_ms.setdefault(name, []).append(a.elapsed_time(b))
    pipe.kernel_events = None
    summary = pipe.summarize(res)
    pairs_per_step = U * V * n_gpus
    value = pairs_per_step * steps / (ms_total * 1e-3)

    # ---- end to end: pinned host buffers in, summary out ----
    e2e = None
    if not args.no_e2e:
        E = min(args.e2e_users, U)
        hb = pipeline.synth_batch_host(cfg, rank * U, E, pin=True)
        for _ in range(2):
            pipe.run_host(hb, dev)
        barrier()
        t0 = time.perf_counter()
        e2e_steps = max(3, min(steps, 10))
        for _ in range(e2e_steps):
            _, summ = pipe.run_host(hb, dev)
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        d2h = res.reduced.numel() * 8 + res.thresholds.numel() * 4 + res.select_state.numel() * 8 + 4
        e2e = {"value": E * V * n_gpus * e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": hb.nbytes(),
               "d2h_bytes_per_step": d2h, "users_per_gpu_per_step": E, "steps": e2e_steps,
               "h2d_gbs_per_gpu": hb.nbytes() * e2e_steps / dt / 1e9,      # PCIe is the bound here (~55 GB/s on Gen5 x16)
               "note": "ScoringPipeline.run_host: chunked H2D from pinned memory overlapped with the kernels",
               "cpu_binding_rank0": numa}
        # ---- what the host->device path of this box can do with no kernel at all: the same pinned batch,
        #      the same chunks, copies only; all ranks at once (they share the host's memory and PCIe roots)
        bufs = pipe._staging(dev, 16384)
        copy = pipe._copy_stream(dev)
        def copies_only():
            with torch.cuda.stream(copy):
                for c in range((E + 16383) // 16384):
                    lo, hi = c * 16384, min(E, (c + 1) * 16384)
                    for dst, src in zip(bufs[c & 1], hb.tensors()):
                        dst[:hi - lo].copy_(src[lo:hi], non_blocking=True)
            copy.synchronize()
        copies_only()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            copies_only()
        barrier()
        dt_copy = max_over_ranks(time.perf_counter() - t0)
        e2e["h2d_ceiling_gbs_per_gpu"] = hb.nbytes() * e2e_steps / dt_copy / 1e9
        e2e["h2d_frac_of_ceiling"] = e2e["h2d_gbs_per_gpu"] / e2e["h2d_ceiling_gbs_per_gpu"]
        e2e["ceiling_note"] = ("copies only, no kernels, all %d ranks at once: what is left between this and the "
                               "e2e rate is software, what is below the link rate is the host" % n_gpus)
        del hb

    # ---- multi-GPU correctness witness (outside every timed region): sharded run vs the union on rank 0
    parity = None
    if distributed:
        parity = pipeline.multi_gpu_parity(cfg, rank, world, dev, users_per_rank=args.parity_users)
        barrier()

    extra = None
    if rank == 0 and n_gpus == 1 and not args.no_other_configs:
        del batch
        torch.cuda.empty_cache()
        extra = other_configs(dev, hbm_peak()[0])
    if rank == 0:
        peak, peak_src = hbm_peak()
        pairs_gpu = U * V
        lm_name, cos_name = pipe.kernel_names()
        records = []
        for name, bytes_per_pair in ((cos_name, D * 4 * (V + 1) / V + 4),
                                     (lm_name, K * 4 * (V + 1) / V + 4 * (V + 1) / V + 16)):
            if kernel_ms.get(name):
                t = float(np.mean(kernel_ms[name])) * 1e-3
                ach = pairs_gpu * bytes_per_pair / t / 1e9
                records.append({"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                "traffic": ncu_traffic(name, pairs_gpu), "kernel": name, "avg_launch_ms": t * 1e3,
                                "algorithmic_bytes_per_launch": pairs_gpu * bytes_per_pair,
                                "share_of_step_kernel_time": None,
                                "peak_source": peak_src, "frac_of_nominal_8000": ach / 8000.0})
        t_all = sum(r["avg_launch_ms"] for r in records) or 1.0
        for r in records:
            r["share_of_step_kernel_time"] = r["avg_launch_ms"] / t_all
        # `roofline` is the dominant kernel's record (most of the step's time and bytes); the records of every
        # step kernel, weakest included, ride inside it so that they stay with the line
        roofline = None
        if records:
            roofline = dict(max(records, key=lambda r: r["avg_launch_ms"]))
            roofline["step_kernels"] = records
            roofline["worst_frac"] = min(r["frac"] for r in records)
            roofline["worst_kernel"] = min(records, key=lambda r: r["frac"])["kernel"]
        cpu = None
        if n_gpus == 1 and not args.no_cpu_baseline:
            # bounded sample: batches of --cpu-sample-users users until about 10 s of scoring time
            n, t_sum, reps, cores = args.cpu_sample_users, 0.0, 0, 1
            while t_sum < args.cpu_sample_seconds and reps < 64:
                _, dt, cores = cpu_sample(args, n, user0=reps * n)
                t_sum += dt
                reps += 1
            cpu = {"value": reps * n * V / t_sum, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": "%d batches of %d users x %d variants, full pipeline, %.1f s of scoring on %d cores"
                             % (reps, n, V, t_sum, cores)}
            mods = reference_modules()
            if mods is not None:       # the reference's own functions on one thread, beside the port
                v_ref, dt_ref, parts = reference_sample(args, mods, args.ref_sample_users)
                cpu["reference_single_thread"] = {
                    "value": v_ref, "unit": UNIT, "cores": 1, "kind": "reference",
                    "sample": "%d users x %d variants: calculate_individual_fairness %.2f s + "
                              "calculate_conformal_thresholds %.2f s (no SERP*/PRAG*/cosine in the reference)"
                              % (args.ref_sample_users, V, parts["jaccard_s"], parts["threshold_s"])}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": steps, "warmup": warm,
            "ms_per_step": ms_total / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args, n_gpus), "e2e": e2e,
            "gpu_launches": launches, "clocks": clocks,
            "roofline": roofline,
            "cpu_baseline": cpu,
            "step_kernel_ms": {k: float(np.mean(v)) for k, v in kernel_ms.items()},
            "other_configs": extra,
            "check": dict({"threshold": float(summary["thresholds"][0]),
                           "violations_round0": int(summary["violation_counts"][0]),
                           "mean_jaccard_v0": float(summary["variant_means"][0, 0])}, **(parity or {})),
        }
        print(json.dumps(line))
    if distributed:
        dist.destroy_process_group()


def main():
    args = parse_args()
    # Native libraries (NCCL's version banner, for one) write to file descriptor 1.  Keep the real stdout
    # for the JSON line alone: everything else that lands on fd 1 goes to stderr.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    try:
        if args.impl == "reference":
            run_reference(args)
        else:
            run_ours(args)
    finally:
        real_stdout.flush()


if __name__ == "__main__":
    main()
