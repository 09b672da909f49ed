"""Generate tests/golden/reference_golden.json by running the UNMODIFIED reference
functions from /root/reference/src on seeded inputs.  Run in the build container:

    python -m oracle.make_golden                      (everything)
    python -m oracle.make_golden --append-missing     (only the sections the committed file lacks)

TEST INFRASTRUCTURE ONLY.  The JSON holds inputs and the reference's outputs; floats
round-trip exactly through ``repr``.  (The reference's demographic-parity distances are summed in the
iteration order of a Python ``set`` of titles, so their last digit can change from run to run; the tests
compare them at 1e-12.)  The reference has no tests of its own
(SURVEY.md section 4), so these vectors are what pins the oracle and the CUDA path.
"""
from __future__ import annotations

import json
import os
import random

import numpy as np
import pandas as pd

from . import ref_loader

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "tests", "golden", "reference_golden.json")

TITLES = ["Toy Story", "Heat", "Casino", "Se7en", "Babe", "Clueless", "Fargo", "Twister",
          "Scream", "Alien", "Aliens", "Psycho", "Jaws", "Rocky", "Gandhi", "Platoon",
          "Brazil", "Big", "Ghost", "Speed", "Rudy", "Nell", "Congo", "Kids", "Crumb",
          "Friday", "Nixon", "Sabrina", "Jumanji", "GoldenEye", " Heat", "heat"]


def _rand_list(rng, kmax, p_empty=0.1, p_dup=0.2):
    if rng.random() < p_empty:
        return []
    k = rng.randint(1, kmax)
    lst = [rng.choice(TITLES) for _ in range(k)] if rng.random() < p_dup else rng.sample(TITLES, k)
    return lst


def _profiles(genders, ages, occs, repeats):
    out = []
    for g in genders:
        for a in ages:
            for o in occs:
                for r in range(repeats):
                    out.append({"id": f"p_{g}_{a}_{o}_{r}", "gender": g, "age": a, "occupation": o})
    return out


def aggressive_cases():
    """a5, second copy: jaccard / measure_fairness of src/phase3_aggressive.py:174-195 (own seed, so that
    adding this section left every earlier vector as it was)."""
    agg = ref_loader.load_script_functions("phase3_aggressive.py")
    rng = random.Random(174195)
    inputs = [{"g1": [["a"]]}, {"g1": [[]], "g2": [[]]}, {"g1": [], "g2": []}, {"a": [["x"]], "b": [["x"]], "c": [["x"]]},
              {"m": [["a", "b"], []], "f": [["b", "c"], ["q"]]}]
    for _ in range(10):
        n1, n2 = rng.randint(0, 8), rng.randint(0, 8)
        inputs.append({"male": [_rand_list(rng, 10, p_empty=0.2) for _ in range(n1)],
                       "female": [_rand_list(rng, 10, p_empty=0.2) for _ in range(n2)]})
    return {"jaccard": [{"s1": s1, "s2": s2, "jaccard": float(agg.jaccard(set(s1), set(s2)))}
                        for s1, s2 in [([], []), (["a"], []), ([], ["b"]), (["a", "b"], ["b", "c"]), (["a"], ["a"])]],
            "measure_fairness": [{"recs_by_group": d, "fairness": float(agg.measure_fairness(d))} for d in inputs]}


def divergence_and_bias_reduction_cases():
    """utils.calculate_kl_divergence / calculate_js_divergence (src/utils.py:70-102) and
    phase3_facter_mitigation.measure_bias_reduction (:280-331), own seed."""
    utils = ref_loader.load("utils")
    p3 = ref_loader.load("phase3_facter_mitigation")
    rng = random.Random(280331)

    def dist(n):
        keys = rng.sample(TITLES, n)
        w = [rng.random() + 0.01 for _ in keys]
        return {k: x / sum(w) for k, x in zip(keys, w)}
    pairs = [({"a": 1.0}, {"a": 1.0}), ({"a": 1.0}, {"b": 1.0}), ({"a": 0.5, "b": 0.5}, {"a": 0.9, "b": 0.1}),
             ({"a": 2, "b": 6}, {"b": 1, "c": 3})]                      # unnormalised inputs are renormalised
    pairs += [(dist(rng.randint(1, 12)), dist(rng.randint(1, 12))) for _ in range(10)]
    div = [{"p": p, "q": q, "kl": float(utils.calculate_kl_divergence(p, q)),
            "js": float(utils.calculate_js_divergence(p, q))} for p, q in pairs]
    cases = []
    for n in (0, 1, 6, 30):
        profs = [{"id": f"u{i}", "gender": rng.choice(["male", "female", "non-binary"]), "age": "25-34",
                  "occupation": "other"} for i in range(n)]
        def recs(p_dict):
            out = {}
            for p in profs:
                r = _rand_list(rng, 10, p_empty=0.1, p_dup=0.2)
                x = rng.random()
                out[p["id"]] = {"recommendations": r} if x < p_dict else r if x < 0.95 else "n/a"
            out["ghost"] = ["Heat"]                                      # no such profile: skipped
            return out
        o, m = recs(0.5), recs(0.3)
        r = p3.measure_bias_reduction(o, m, profs)
        cases.append({"profiles": profs, "original": o, "mitigated": m, "result": {k: float(v) for k, v in r.items()}})
    return {"divergences": div, "bias_reduction": cases}


def append_missing():
    """Add the sections the committed file does not have yet, leaving the others untouched."""
    g = json.load(open(OUT))
    if "aggressive" not in g:
        g["aggressive"] = aggressive_cases()
    if "bias_reduction" not in g:
        g.update(divergence_and_bias_reduction_cases())
    with open(OUT, "w") as f:
        json.dump(g, f, indent=0, sort_keys=True)
    print("updated", OUT)


def main():
    rng = random.Random(20251203)
    nrng = np.random.default_rng(20251203)
    utils = ref_loader.load("utils")
    p1 = ref_loader.load("phase1_bias_detection")
    p3 = ref_loader.load("phase3_facter_mitigation")
    fin = ref_loader.load_script_functions("phase3_final.py")
    g = {"generator": "oracle/make_golden.py", "reference": "/root/reference/src (unmodified)",
         "numpy": np.__version__, "pandas": pd.__version__}

    # ---- a1: calculate_individual_fairness (src/utils.py:217) ----
    cases = []
    x, y, z, w = "x", "y", "z", "w"
    fixed = ([("a", "b"), ("c", "d"), ("a", "c"), ("a", "e"), ("a", "zz")],
             {"a": [x, y, z], "b": [y, z, w], "c": [], "d": [], "e": [x, x, y]})
    inputs = [fixed, ([], {}), ([("a", "b")], {"a": [], "b": []})]
    for _ in range(12):
        pids = [f"u{i}" for i in range(rng.randint(2, 14))]
        recs = {p: _rand_list(rng, 20) for p in pids if rng.random() > 0.1}
        pairs = [(rng.choice(pids), rng.choice(pids)) for _ in range(rng.randint(1, 40))]
        inputs.append((pairs, recs))
    for pairs, recs in inputs:
        score, sims = utils.calculate_individual_fairness(pairs, recs)
        cases.append({"pairs": [list(p) for p in pairs], "recs": recs,
                      "score": float(score), "sims": [float(s) for s in sims]})
    g["individual_fairness"] = cases

    # ---- a3: create_profile_pairs (src/utils.py:327) ----
    cases = []
    prof45 = _profiles(["male", "female", "non-binary"], ["18-24", "25-34", "35-44", "45-54", "55+"],
                       ["professional"], 3)
    sets = [prof45, _profiles(["m", "f"], ["a", "b"], ["x", "y", "z"], 1)]
    odd = _profiles(["m", "f"], ["a", "b"], ["x"], 2)
    del odd[3]["age"]
    del odd[5]["occupation"]
    sets.append(odd)
    for profs in sets:
        for attr in (None, "gender", "age", "occupation"):
            pairs = utils.create_profile_pairs(profs, attr)
            cases.append({"profiles": profs, "attr": attr, "pairs": [list(p) for p in pairs]})
    g["profile_pairs"] = cases

    # ---- a2: measure_individual_fairness (src/phase1_bias_detection.py:220) ----
    cases = []
    for _ in range(3):
        recs = {p["id"]: {"profile_id": p["id"], "recommendations": _rand_list(rng, 10)}
                for p in prof45 if rng.random() > 0.05}
        if recs:
            first = next(iter(recs))
            recs[first] = {"profile_id": first, "error": "boom"}      # row without 'recommendations'
        score, det = p1.measure_individual_fairness(prof45, recs)
        cases.append({"profiles": prof45, "recs": recs, "score": float(score),
                      "sims": [float(s) for s in det["similarity_scores"]],
                      "num_pairs": int(det["num_pairs"]), "std": float(det["std_similarity"])})
    g["measure_individual_fairness"] = cases

    # ---- a4: measure_quality_preservation (src/phase3_facter_mitigation.py:333) ----
    cases = []
    letters = list("abcdefghijklmnop")
    inputs = [({"u": letters[:12]}, {"u": ["a", "b", "c", "x", "y", "z"]}), ({}, {}),
              ({"u": []}, {"u": []}), ({"u": ["a"]}, {"v": ["a"]})]
    for _ in range(6):
        pids = [f"u{i}" for i in range(rng.randint(1, 12))]
        o = {p: (_rand_list(rng, 16) if rng.random() < 0.5 else {"recommendations": _rand_list(rng, 16)})
             for p in pids}
        m = {p: (_rand_list(rng, 16) if rng.random() < 0.5 else {"recommendations": _rand_list(rng, 16)})
             for p in pids if rng.random() > 0.15}
        inputs.append((o, m))
    for o, m in inputs:
        r = p3.measure_quality_preservation(o, m)
        cases.append({"orig": o, "mitg": m, "average_overlap": float(r["average_overlap"]),
                      "quality_preservation_pct": float(r["quality_preservation_pct"]),
                      "num_comparisons": int(r["num_comparisons"])})
    g["quality_preservation"] = cases

    # ---- a5: phase3_final.jaccard / measure_fairness (src/phase3_final.py:112,119) ----
    cases = []
    for s1, s2 in [([], []), (["a"], []), (["a", "b"], ["b", "c"]), (["a"], ["a"])]:
        cases.append({"s1": s1, "s2": s2, "jaccard": float(fin.jaccard(set(s1), set(s2)))})
    g["final_jaccard"] = cases
    cases = []
    inputs = [{"g1": [["a"]]}, {"g1": [[]], "g2": [[]]}, {"g1": [], "g2": []}]
    for _ in range(8):
        n1, n2 = rng.randint(1, 8), rng.randint(1, 8)
        d = {"male": [_rand_list(rng, 10) for _ in range(n1)],
             "female": [_rand_list(rng, 10) for _ in range(n2)]}
        if rng.random() < 0.4:
            d["non-binary"] = [_rand_list(rng, 10)]
        inputs.append(d)
    for d in inputs:
        cases.append({"recs_by_group": d, "fairness": float(fin.measure_fairness(d))})
    g["measure_fairness"] = cases

    # ---- a9: calculate_conformal_thresholds (src/phase3_facter_mitigation.py:152) ----
    cases = []
    fixed = pd.DataFrame({"group": ["m"] * 10 + ["f"] * 3,
                          "nonconformity": [.05, .01, .09, .03, .07, .02, .10, .04, .08, .06, .5, .2, .9]})
    frames = [(fixed, 0.1)]
    for _ in range(8):
        n = rng.randint(1, 400)
        groups = nrng.choice(["male", "female", "non-binary"], size=n)
        sc = np.abs(nrng.normal(0, 0.1, size=n))
        if rng.random() < 0.5:
            sc = np.round(sc, 2)                      # heavy ties
        frames.append((pd.DataFrame({"group": groups, "nonconformity": sc}),
                       rng.choice([0.1, 0.05, 0.2, 0.5, 0.01])))
    for df, alpha in frames:
        thr = p3.calculate_conformal_thresholds(df, alpha)
        cases.append({"group": list(df["group"]), "nonconformity": [float(v) for v in df["nonconformity"]],
                      "alpha": alpha, "thresholds": {k: float(v) for k, v in thr.items()}})
    g["conformal_thresholds"] = cases
    g["conformal_index"] = [{"n": n, "alpha": a, "index": max(0, min(int(np.ceil((n + 1) * (1 - a))) - 1, n - 1))}
                            for n in (1, 2, 3, 9, 10, 11, 19, 20, 99, 100, 1000, 187240, 1000000, 310000000)
                            for a in (0.1, 0.05, 0.2)]

    # ---- a10: apply_conformal_filtering (src/phase3_facter_mitigation.py:184) ----
    cases = []
    profs = [{"id": f"u{i}", "gender": ["male", "female", "other"][i % 3]} for i in range(9)]
    for thr in ({"male": 0.8, "female": 0.95}, {"male": 0.0, "female": 1.0, "other": 0.55},
                {"male": 0.5000000001, "female": 0.85, "other": 0.15}):
        recs = {f"u{i}": {"recommendations": [f"m{j}" for j in range(rng.randint(0, 22))]} for i in range(9)}
        recs["ghost"] = {"recommendations": ["a", "b", "c"]}
        recs["u2"] = {}
        out = p3.apply_conformal_filtering(recs, profs, thr)
        cases.append({"recs": recs, "profiles": profs, "thresholds": thr, "filtered": out})
    g["conformal_filtering"] = cases

    # ---- a8: generate_calibration_data (src/phase3_facter_mitigation.py:109), seeded ----
    profs = [{"id": f"u{i}", "gender": ["male", "female"][i % 2]} for i in range(6)]
    recs = {f"u{i}": {"recommendations": [f"m{j}" for j in range(i + 1)]} for i in range(5)}
    recs["u3"] = {"error": "x"}
    np.random.seed(1234)
    df = p3.generate_calibration_data(profs, recs, {"m0": 0.75})
    g["calibration_data"] = {"profiles": profs, "recs": recs, "ratings": {"m0": 0.75}, "seed": 1234,
                             "columns": {c: [v if isinstance(v, str) else float(v) for v in df[c]]
                                         for c in df.columns}}

    # ---- N1: calculate_demographic_parity / calculate_js_divergence (src/utils.py:172, :89) ----
    # (appended last so that every seeded case above keeps its values)
    cases = []
    inputs = [{"g1": [["x", "y"], ["x", "z"]], "g2": [["x", "y"], ["y", "w"]]}, {}, {"solo": [["a"]]},
              {"a": [[]], "b": [["m"]]}, {"a": [["m", "m", "n"]], "b": [["m"]], "c": [["q", "q"]]}]
    for _ in range(8):
        names = rng.sample(["male", "female", "non-binary", "18-24", "25-34", "35-44", "55+"], rng.randint(2, 6))
        inputs.append({n: [_rand_list(rng, 12, p_empty=0.15, p_dup=0.3) for _ in range(rng.randint(0, 9))]
                       for n in names})
    for d in inputs:
        score, det = utils.calculate_demographic_parity(d)
        # group order matters (pairs are enumerated i < j) and json.dump sorts keys: keep it as a list
        cases.append({"groups": [[k, v] for k, v in d.items()], "score": float(score),
                      "divergences": [float(x) for x in det["divergences"]],
                      "avg_divergence": float(det["avg_divergence"]),
                      "distributions": {k: {i: float(v) for i, v in dist.items()} for k, dist in det["distributions"].items()}})
    g["demographic_parity"] = cases
    g["js_divergence"] = [{"p": {"a": 1}, "q": {"b": 1}, "js": float(utils.calculate_js_divergence({"a": 1}, {"b": 1}))},
                          {"p": {"a": 0.5, "b": 0.3, "c": 0.2}, "q": {"a": 0.4, "b": 0.4, "c": 0.2},
                           "js": float(utils.calculate_js_divergence({"a": 0.5, "b": 0.3, "c": 0.2}, {"a": 0.4, "b": 0.4, "c": 0.2}))}]

    # ---- N4: calculate_equal_opportunity / measure_equal_opportunity (src/utils.py:246, phase1:241) ----
    cases = []
    qualified = ["Heat", "Fargo", "Alien", "Jaws", "The Matrix"]
    for c in g["demographic_parity"]:
        d = {k: v for k, v in c["groups"]}
        score, by_group = utils.calculate_equal_opportunity(d, set(qualified))
        entry = {"groups": c["groups"], "relevant": qualified, "score": float(score),
                 "by_group": [[k, float(v)] for k, v in by_group.items()]}
        if d:
            s2, det = p1.measure_equal_opportunity(d, None)
            entry["measure"] = {"score": float(s2), "variance": float(det["variance"]),
                                "min_max_ratio": float(det["min_max_ratio"])}
        cases.append(entry)
    d = {"a": [["The Matrix", "Heat", "The Matrix"], ["Inception"]], "b": [["Heat"]], "c": [[]]}
    score, by_group = utils.calculate_equal_opportunity(d, {"The Matrix", "Inception", "Unseen"})
    s2, det = p1.measure_equal_opportunity(d, None)
    cases.append({"groups": [[k, v] for k, v in d.items()], "relevant": ["The Matrix", "Inception", "Unseen"],
                  "score": float(score), "by_group": [[k, float(v)] for k, v in by_group.items()],
                  "measure": {"score": float(s2), "variance": float(det["variance"]),
                              "min_max_ratio": float(det["min_max_ratio"])}})
    g["equal_opportunity"] = cases

    # ---- N2: the metric step of run_phase1 on a results dict (src/phase1_bias_detection.py:343-374) ----
    from collections import defaultdict
    titles = TITLES + ["The Matrix", "Inception", "Goodfellas", "Pulp Fiction"]
    recs = {p["id"]: {"profile_id": p["id"], "recommendations": rng.sample(titles, rng.randint(0, 10))}
            for p in prof45 if rng.random() > 0.1}
    by_gender, by_age = defaultdict(list), defaultdict(list)
    for p in prof45:
        if p["id"] in recs:
            by_gender[p["gender"]].append(recs[p["id"]]["recommendations"])
            by_age[p["age"]].append(recs[p["id"]]["recommendations"])
    dpg, _ = p1.measure_demographic_parity(dict(by_gender))
    dpa, _ = p1.measure_demographic_parity(dict(by_age))
    ifs, ifd = p1.measure_individual_fairness(prof45, recs)
    eos, eod = p1.measure_equal_opportunity(dict(by_gender), None)
    g["phase1_metric_step"] = {"profiles": prof45, "recommendations": recs, "dp_gender": float(dpg), "dp_age": float(dpa),
                               "if_score": float(ifs), "if_std": float(ifd["std_similarity"]), "eo_score": float(eos),
                               "eo_by_group": [[k, float(v)] for k, v in eod["opportunity_by_group"].items()]}

    # ---- N3: ranking metrics of phase 2 (src/phase2_cross_model_eval.py:216-316, src/utils.py:113,277) ----
    p2 = ref_loader.load("phase2_cross_model_eval")
    cases = []
    for _ in range(10):
        n = rng.randint(1, 24)
        items = [{"id": f"i{j}", "protected_attribute": rng.choice(["male", "female", "other"]),
                  "relevance_score": round(rng.random(), 3)} for j in range(n)]
        if rng.random() < 0.3:
            del items[0]["relevance_score"]
        if rng.random() < 0.3:
            del items[-1]["protected_attribute"]
        ranked = [it["id"] for it in items]
        rng.shuffle(ranked)
        if rng.random() < 0.5:
            ranked.insert(rng.randint(0, len(ranked)), "ghost")             # an id the model made up
        ratio, avg = p2.calculate_exposure_ratio(ranked, items)
        k = rng.choice([3, 5, 10, 50])
        nd = p2.calculate_ndcg_per_group(ranked, items, k)
        tuples = [(it["id"], it.get("protected_attribute", "unknown")) for it in items]
        r2, avg2 = utils.calculate_exposure_ratio(tuples)
        gt = {it["id"]: it.get("relevance_score", 0.5) for it in items}
        cases.append({"ranked": ranked, "items": items, "k": k, "ratio": float(ratio),
                      "avg": [[a, float(b)] for a, b in avg.items()], "ndcg_per_group": [[a, float(b)] for a, b in nd.items()],
                      "tuples": [list(t) for t in tuples], "ratio_tuples": float(r2),
                      "avg_tuples": [[a, float(b)] for a, b in avg2.items()],
                      "ndcg_all": float(utils.calculate_ndcg(ranked, gt, k))})
    comps = [{"winner": rng.choice(["A", "B", "tie"]), "item_a_attr": rng.choice(["m", "f"]), "item_b_attr": rng.choice(["m", "f"])}
             for _ in range(25)]
    g["ranking_metrics"] = {"cases": cases, "comparisons": comps,
                            "win_rates": [[a, float(b)] for a, b in p2.calculate_pairwise_preference_ratio(comps).items()],
                            "ndcg_known": float(utils.calculate_ndcg(["i2", "i1", "i3"], {"i1": 5, "i2": 3, "i3": 1})),
                            "ndcg_empty": float(utils.calculate_ndcg(["a"], {})),
                            "exposure_empty": float(p2.calculate_exposure_ratio([], [])[0])}

    # ---- N2: text -> list of titles, the step before the string->id encoder (src/utils.py:350-382) ----
    texts = [
        "1. The Matrix (1999)\n2. Inception\n3) \"Toy Story (1995)\"\n",
        "- Heat\n* 'Alien'\n\n   \n-- Up\n12.Se7en\n",
        "",
        "   \n\n",
        "Here are my picks:\n1.  Fargo (1996)\n2.\tBrazil\n3. 2001: A Space Odyssey (1968)\n10) 10 Things I Hate About You",
        "\"\"\n'\n1.\n- \n* *Stars*\n1. 2. Nested",
        "No numbering at all\r\nSecond line\r\n",
        "1.5 Stars\n(1999)\nMovie (19999)\nMovie (1999) (2001)  \n  Am\u00e9lie (2001)",
    ]
    g["text_parsing"] = [{"text": t, "lines": utils.parse_recommendation_list(t),
                          "titles": [utils.extract_movie_title(x) for x in utils.parse_recommendation_list(t)]}
                         for t in texts]

    g["aggressive"] = aggressive_cases()
    g.update(divergence_and_bias_reduction_cases())

    with open(OUT, "w") as f:
        json.dump(g, f, indent=0, sort_keys=True)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    import sys
    append_missing() if "--append-missing" in sys.argv else main()
