"""The CPU oracle against the golden vectors produced by the unmodified reference
(oracle/make_golden.py -> tests/golden/reference_golden.json).  No GPU."""
import json
import os

import numpy as np
import pandas as pd
import pytest

from oracle import np_oracle as O
from oracle import ref_port as R


def _encode(recs):
    vocab = {}
    return {k: [vocab.setdefault(t, len(vocab)) for t in v] for k, v in recs.items()}, max(len(vocab), 1)


def test_individual_fairness_port_matches_reference(golden):
    for c in golden["individual_fairness"]:
        score, sims = R.calculate_individual_fairness([tuple(p) for p in c["pairs"]], c["recs"])
        assert sims == c["sims"]
        assert float(score) == c["score"]


def test_survey_known_answers(golden):
    c = golden["individual_fairness"][0]           # SURVEY.md section 4, first row
    assert c["score"] == 0.5416666666666666
    assert c["sims"] == [0.5, 1.0, 0.0, 0.6666666666666666]
    t = golden["conformal_thresholds"][0]["thresholds"]
    assert t == {"m": 0.1, "f": 0.9}
    q = golden["quality_preservation"][0]
    assert q["average_overlap"] == 0.23076923076923078


def test_numpy_jaccard_matches_reference(golden):
    """Vectorised integer oracle on id-encoded lists == the reference's string sets."""
    for c in golden["individual_fairness"]:
        recs, vocab = _encode(c["recs"])
        kept = [(p, q) for p, q in c["pairs"] if p in recs and q in recs]
        if not kept:
            assert c["sims"] == []
            continue
        K = max([len(v) for v in recs.values()] + [1])
        A = np.full((len(kept), K), -1, np.int32)
        B = np.full((len(kept), 1, K), -1, np.int32)
        la = np.zeros(len(kept), np.int32)
        lb = np.zeros((len(kept), 1), np.int32)
        for i, (p, q) in enumerate(kept):
            A[i, :len(recs[p])] = recs[p]; la[i] = len(recs[p])
            B[i, 0, :len(recs[q])] = recs[q]; lb[i, 0] = len(recs[q])
        r = O.list_metrics(A, la, B, lb, vocab=vocab)
        assert r["jaccard"][:, 0].tolist() == c["sims"]


def test_profile_pairs_port(golden):
    for c in golden["profile_pairs"]:
        assert [list(p) for p in R.create_profile_pairs(c["profiles"], c["attr"])] == c["pairs"]
    assert len(golden["profile_pairs"][0]["pairs"]) == 405        # 45 profiles -> 405 pairs


def test_measure_individual_fairness_port(golden):
    for c in golden["measure_individual_fairness"]:
        score, det = R.measure_individual_fairness(c["profiles"], c["recs"])
        assert float(score) == c["score"] and det["similarity_scores"] == c["sims"]
        assert det["num_pairs"] == c["num_pairs"] and float(det["std_similarity"]) == c["std"]


def test_quality_preservation_port(golden):
    for c in golden["quality_preservation"]:
        r = R.measure_quality_preservation(c["orig"], c["mitg"])
        assert float(r["average_overlap"]) == c["average_overlap"]
        assert float(r["quality_preservation_pct"]) == c["quality_preservation_pct"]
        assert r["num_comparisons"] == c["num_comparisons"]


def test_final_jaccard_and_fairness_port(golden):
    for c in golden["final_jaccard"]:
        assert R.jaccard(set(c["s1"]), set(c["s2"])) == c["jaccard"]
    for c in golden["measure_fairness"]:
        assert float(R.measure_fairness(c["recs_by_group"])) == c["fairness"]


def test_aggressive_jaccard_and_fairness_port(golden):
    for c in golden["aggressive"]["jaccard"]:
        assert R.jaccard(set(c["s1"]), set(c["s2"])) == c["jaccard"]
    for c in golden["aggressive"]["measure_fairness"]:
        assert float(R.measure_fairness_aggressive(c["recs_by_group"])) == c["fairness"]


def test_divergence_and_bias_reduction_port(golden):
    """KL / JS helpers (src/utils.py:70-102) and measure_bias_reduction (src/phase3_facter_mitigation.py:280):
    the reference sums in the order of a Python set, so 1e-12 relative rather than equality."""
    for c in golden["divergences"]:
        assert R.calculate_kl_divergence(c["p"], c["q"]) == pytest.approx(c["kl"], rel=1e-12, abs=1e-15)
        assert R.calculate_js_divergence(c["p"], c["q"]) == pytest.approx(c["js"], rel=1e-12, abs=1e-8)
    for c in golden["bias_reduction"]:
        got = R.measure_bias_reduction(c["original"], c["mitigated"], c["profiles"])
        assert set(got) == set(c["result"])
        for k, v in c["result"].items():
            assert float(got[k]) == pytest.approx(v, rel=1e-9, abs=1e-12)


def test_conformal_index_rule(golden):
    for c in golden["conformal_index"]:
        assert O.conformal_index(c["n"], c["alpha"]) == c["index"]


def test_conformal_thresholds(golden):
    for c in golden["conformal_thresholds"]:
        df = pd.DataFrame({"group": c["group"], "nonconformity": c["nonconformity"]})
        assert {k: float(v) for k, v in R.calculate_conformal_thresholds(df, c["alpha"]).items()} == c["thresholds"]
        scores = np.asarray(c["nonconformity"])
        groups = np.asarray(c["group"])
        for g, want in c["thresholds"].items():
            assert float(O.conformal_threshold(scores[groups == g], c["alpha"])) == want


def test_conformal_filtering(golden):
    for c in golden["conformal_filtering"]:
        assert R.apply_conformal_filtering(c["recs"], c["profiles"], c["thresholds"]) == c["filtered"]
        by_id = {p["id"]: p for p in c["profiles"]}
        for pid, kept in c["filtered"].items():
            n = len(c["recs"][pid].get("recommendations", []))
            thr = c["thresholds"].get(by_id[pid]["gender"], 0.5)
            assert int(O.conformal_keep_len(np.array([n]), thr)[0]) == len(kept)


def test_calibration_records_port(golden):
    c = golden["calibration_data"]
    np.random.seed(c["seed"])
    cols = R.nonconformity_records(c["profiles"], c["recs"], c["ratings"])
    for k, want in c["columns"].items():
        assert [v if isinstance(v, str) else float(v) for v in cols[k]] == want


def test_vectorised_list_oracle_vs_scalar():
    rng = np.random.default_rng(7)
    U, V, K = 60, 4, 14
    A = rng.integers(0, 25, size=(U, K)).astype(np.int32)
    B = rng.integers(0, 25, size=(U, V, K)).astype(np.int32)
    la = rng.integers(0, K + 1, size=U).astype(np.int32)
    lb = rng.integers(0, K + 1, size=(U, V)).astype(np.int32)
    for kc in (None, 10, 1):
        r = O.list_metrics(A, la, B, lb, k_cap=kc)
        kn = K if kc is None else kc
        for u in range(U):
            for v in range(V):
                p = O.list_metrics_pair_python(A[u], la[u], B[u, v], lb[u, v], kn)
                for key in ("inter", "union", "n_b", "serp_num", "prag_num"):
                    assert r[key][u, v] == p[key]


def test_identical_lists_closed_forms():
    K = 9
    A = np.arange(K, dtype=np.int32)[None]
    r = O.list_metrics(A, [K], A[None], [[K]])
    assert r["jaccard"][0, 0] == 1.0
    assert r["serp"][0, 0] == 1.0                                   # sum (K-j) = K(K+1)/2
    assert r["prag"][0, 0] == (K * (K - 1) / 2) / (K * (K + 1))     # all ordered pairs agree
    rev = A[:, ::-1].copy()
    r = O.list_metrics(A, [K], rev[None], [[K]])
    assert r["jaccard"][0, 0] == 1.0 and r["prag_num"][0, 0] == 0


def test_empty_policies():
    A = np.zeros((3, 4), np.int32)
    B = np.zeros((3, 1, 4), np.int32)
    la, lb = [0, 2, 0], [[0], [0], [2]]
    r1 = O.list_metrics(A, la, B, lb, empty_policy=O.EMPTY_BOTH_IS_ONE)
    r0 = O.list_metrics(A, la, B, lb, empty_policy=O.EMPTY_ANY_IS_ZERO)
    assert r1["jaccard"][:, 0].tolist() == [1.0, 0.0, 0.0]
    assert r0["jaccard"][:, 0].tolist() == [0.0, 0.0, 0.0]


def test_oov_rejected():
    A = np.array([[5, 1]], np.int32)
    with pytest.raises(ValueError):
        O.list_metrics(A, [2], A[None], [[2]], vocab=4)
    O.list_metrics(A, [1], A[None], [[0]], vocab=6)


def test_snsr_snsv():
    sim = np.linspace(0.2, 0.8, 31)
    s = O.snsr_snsv(sim)
    assert s["gender"]["snsr"] == pytest.approx(sim[2] - sim[0])
    assert s["age"]["snsv"] == pytest.approx(np.std(sim[3:10]))
    assert set(s) == {"gender", "age", "occupation"}


def test_cosine_mirror_close_to_float64():
    rng = np.random.default_rng(3)
    a = rng.normal(size=(16, 384)).astype(np.float32)
    b = (a[:, None, :] + rng.normal(size=(16, 6, 384)).astype(np.float32) * np.linspace(0.01, 0.6, 6)[None, :, None]
         ).astype(np.float32)
    b[0, 0] = 0.0                        # zero-norm row -> 1
    b[1, 1] = a[1]                       # identical -> 0 (clamped)
    s = O.cosine_nonconformity(a, b)
    m = O.cosine_nonconformity_mirror(a, b)
    assert s[0, 0] == 1.0 and m[0, 0] == 1.0
    assert s[1, 1] <= 1e-15 and m[1, 1] == 0.0
    cos_err = np.abs((1.0 - s) - (1.0 - m.astype(np.float64)))
    assert cos_err.max() <= 1e-6 * np.abs(1.0 - s).max()
    assert np.abs(s - m).max() <= 6e-8


def test_violation_rounds_equals_schedule():
    rng = np.random.default_rng(5)
    s = np.abs(rng.normal(0.05, 0.03, size=(50, 31))).astype(np.float32)
    for q0 in (float(np.sort(s.ravel())[1400]), float(s.max()), 0.0):
        v = O.violation_rounds(s, q0, gamma=0.95, rounds=4, q_min=0.01)
        n_greater = int((s > np.float32(q0)).sum())
        sched = O.threshold_schedule(q0, n_greater, gamma=0.95, rounds=4, q_min=0.01)
        assert v["thresholds"].tolist() == sched.tolist()
        for r in range(4):
            assert v["counts"][r] == int((s > sched[r]).sum())
            assert np.array_equal((v["flags"] >> r) & 1, (s > sched[r]).astype(np.uint8))
        assert np.array_equal(v["variant_counts"].sum(0), v["counts"])


def test_radix_restatement_matches_sort():
    rng = np.random.default_rng(11)
    for n in (1, 2, 3, 10, 1000, 30000):
        s = np.abs(rng.normal(0.05, 0.03, size=n)).astype(np.float32)
        s[rng.integers(0, n, size=n // 4)] = np.float32(0.0625)      # heavy ties
        for alpha in (0.1, 0.5, 0.01):
            thr, nn, k, less, eq = O.radix_select_sharded(np.array_split(s, 3), alpha)
            assert thr == O.conformal_threshold(s, alpha) and nn == n and k == O.conformal_index(n, alpha)
            assert (less, eq) == O.select_counts(s, thr)[:2]
    assert O.radix_select_sharded([np.zeros(0, np.float32)])[0] == np.float32(0.5)


def test_bf16_round_and_pair_similarity():
    rng = np.random.default_rng(2)
    x = rng.normal(size=(2, 5, 64)).astype(np.float32)
    xh = O.normalize_rows_bf16(x)
    assert np.all((xh.view(np.uint32) & 0xFFFF) == 0)
    S = O.pair_similarity(xh)
    assert np.allclose(np.diagonal(S, axis1=1, axis2=2), 1.0, atol=1e-2)
    assert np.allclose(S, np.swapaxes(S, 1, 2))


def test_demographic_parity_port_and_counts_oracle(golden):
    for c in golden["js_divergence"]:
        assert R.calculate_js_divergence(c["p"], c["q"]) == pytest.approx(c["js"], rel=1e-13)
    for c in golden["demographic_parity"]:
        score, det = R.calculate_demographic_parity(dict(c["groups"]))
        assert det["distributions"] == c["distributions"]
        assert det["divergences"] == pytest.approx(c["divergences"], rel=1e-12, abs=1e-15)
        assert float(score) == pytest.approx(c["score"], rel=1e-12)
        # numpy restatement on id-encoded rows
        vocab, rows, grp = {}, [], []
        for g, (name, lists) in enumerate(c["groups"]):
            for l in lists:
                rows.append([vocab.setdefault(x, len(vocab)) for x in l]); grp.append(g)
        G = len(c["groups"])
        if not rows:
            continue
        K = max([len(r) for r in rows] + [1])
        arr = np.full((len(rows), K), -1, np.int32)
        for i, r in enumerate(rows):
            arr[i, :len(r)] = r
        counts = O.group_item_counts(arr, [len(r) for r in rows], grp, G, max(len(vocab), 1))
        s2, div = O.demographic_parity_from_counts(counts)
        assert div == pytest.approx(c["divergences"], rel=1e-12, abs=1e-15)
        assert s2 == pytest.approx(c["score"], rel=1e-12)


def test_equal_opportunity_port(golden):
    for c in golden["equal_opportunity"]:
        score, by_group = R.calculate_equal_opportunity(dict(c["groups"]), set(c["relevant"]))
        assert [[k, float(v)] for k, v in by_group.items()] == c["by_group"]
        assert float(score) == c["score"]


def test_ranking_metrics_port(golden):
    rm = golden["ranking_metrics"]
    assert rm["ndcg_known"] == 0.9001539923801699
    assert R.calculate_ndcg(["i2", "i1", "i3"], {"i1": 5, "i2": 3, "i3": 1}) == rm["ndcg_known"]
    assert R.calculate_ndcg(["a"], {}) == rm["ndcg_empty"] == 0.0
    for c in rm["cases"]:
        ratio, avg = R.calculate_exposure_ratio_p2(c["ranked"], c["items"])
        assert float(ratio) == c["ratio"] and [[a, float(b)] for a, b in avg.items()] == c["avg"]
        nd = R.calculate_ndcg_per_group(c["ranked"], c["items"], c["k"])
        assert [[a, float(b)] for a, b in nd.items()] == c["ndcg_per_group"]
        gt = {it["id"]: it.get("relevance_score", 0.5) for it in c["items"]}
        assert float(R.calculate_ndcg(c["ranked"], gt, c["k"])) == c["ndcg_all"]


def test_cpu_input_generator_matches_the_products_host_generator(lib):
    """oracle/synth_ref.c (plain C, used by the CPU arms of bench.py) restates the product's generator:
    identical bytes for lists, lengths and embeddings, at the benchmark shape and at an odd one."""
    from fairness_llm_replication_b200 import ops
    from oracle import synth_ref
    for seed, user0, U, V, K, vocab in ((20251203, 12345, 257, 31, 50, 3706), (7, 0, 64, 5, 13, 90)):
        mine = synth_ref.lists(seed, user0, U, V, K, vocab)
        theirs = ops.synth_lists_host(seed, user0, U, V, K, vocab)
        for a, b in zip(mine, theirs):
            assert a.dtype == b.dtype and np.array_equal(a, b)
    for a, b in zip(synth_ref.embeddings(20251203, 99, 33, 31, 384), ops.synth_embeddings_host(20251203, 99, 33, 31, 384)):
        assert a.dtype == b.dtype and np.array_equal(a, b)


def test_reference_arm_runs_without_the_cuda_library():
    """`bench.py --impl reference` must not map libflr_b200.so: run it in a fresh process and
    read its own /proc/self/maps from inside."""
    import subprocess, sys, textwrap
    code = textwrap.dedent("""
        import sys, json, io, contextlib
        sys.argv = ["bench.py", "--impl", "reference", "--steps", "1", "--warmup", "0",
                    "--ref-sample-users", "64", "--cpu-sample-users", "64"]
        import bench
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            bench.run_reference(bench.parse_args())
        line = json.loads(buf.getvalue())
        maps = open("/proc/self/maps").read()
        print(json.dumps({"kind": line["cpu_baseline"]["kind"], "impl": line["impl"], "value": line["value"],
                          "maps_flr": "libflr_b200" in maps, "torch": "torch" in sys.modules}))
    """)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["impl"] == "reference" and out["value"] > 0
    assert out["kind"] in ("reference", "port")
    assert not out["maps_flr"] and not out["torch"]
