"""GPU parity of the conformal path (through the C ABI): cosine nonconformity, exact radix
select, threshold schedule, violation flags, conformal filtering."""
import numpy as np
import pandas as pd
import pytest
import torch

from oracle import np_oracle as O

pytestmark = pytest.mark.gpu
COS_RTOL = 1e-6      # north_star: fp32 cosine scores within 1e-6 relative (stated on cos)


def _emb(rng, U, V, D, sig_hi=0.6):
    a = rng.normal(size=(U, D)).astype(np.float32)
    sig = np.linspace(0.01, sig_hi, V)[None, :, None]
    b = (a[:, None, :] + sig * rng.normal(size=(U, V, D))).astype(np.float32)
    return a, b


@pytest.mark.parametrize("U,V,D", [(301, 31, 384), (64, 1, 384), (50, 33, 128), (40, 64, 260), (17, 2, 1024),
                                   (9, 5, 4), (1000, 31, 384), (23, 124, 384), (150, 31, 256), (5, 300, 128),
                                   (20000, 1, 128), (5000, 1, 256), (333, 1, 200), (41, 1, 1024), (70000, 1, 384)])
def test_cosine_matches_oracle(U, V, D):
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(U + V + D)
    a, b = _emb(rng, U, V, D)
    if U > 8:
        b[0, 0] = 0.0
        a[1] = 0.0
        b[2, V - 1] = a[2]
        b[3, 0] = -a[3]
    hist = torch.zeros(2048, dtype=torch.int64, device="cuda:0")
    s = ops.cosine_nonconformity(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), hist0=hist).cpu().numpy()
    want = O.cosine_nonconformity(a, b)
    cos_got, cos_want = 1.0 - s.astype(np.float64), 1.0 - want
    assert np.all(np.abs(cos_got - cos_want) <= COS_RTOL * np.maximum(np.abs(cos_want), 1e-30) + 1e-7 * (cos_want == 0))
    np.testing.assert_allclose(s, want, rtol=2e-7, atol=1e-9)          # fp64 accumulation: ~1 ulp of fp32 on s itself
    assert np.array_equal(s, O.cosine_nonconformity_mirror(a, b))       # bit-exact vs the arithmetic mirror
    if U > 8:
        assert s[0, 0] == 1.0 and np.all(s[1] == 1.0) and s[2, V - 1] == 0.0 and abs(s[3, 0] - 2.0) < 1e-6
    want_hist = np.bincount((O.radix_key_f32(s.ravel()) >> np.uint32(21)).astype(np.int64), minlength=2048)
    assert np.array_equal(hist.cpu().numpy(), want_hist)


def test_cosine_calibration_form_2d():
    from fairness_llm_replication_b200 import phase3_facter_mitigation as p3
    rng = np.random.default_rng(0)
    a, b = _emb(rng, 5000, 1, 384)
    s = p3.embedding_nonconformity(torch.from_numpy(a).cuda(), torch.from_numpy(b[:, 0]).cuda())
    assert s.shape == (5000,)
    assert np.array_equal(s.cpu().numpy(), O.cosine_nonconformity_mirror(a, b)[:, 0])


def _select(scores, alpha=0.1, groups=None, G=1):
    from fairness_llm_replication_b200 import ops
    t = torch.from_numpy(scores).cuda()
    g = None if groups is None else torch.from_numpy(groups.astype(np.uint8)).cuda()
    thr, state = ops.radix_select(t, alpha=alpha, groups=g, n_groups=G)
    return thr.cpu().numpy(), state.cpu().numpy()


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n", [1, 2, 3, 10, 11, 1000, 187240, 1_000_003])
def test_select_matches_sorted_index(dtype, n):
    rng = np.random.default_rng(n)
    s = np.abs(rng.normal(0.05, 0.03, size=n)).astype(dtype)
    for alpha in (0.1, 0.05, 0.5):
        thr, st = _select(s, alpha)
        want = O.conformal_threshold(s, alpha)
        assert thr[0] == want and thr.dtype == dtype
        less, eq, _ = O.select_counts(s, want)
        assert (st[0, 2], st[0, 3], st[0, 4], st[0, 5]) == (n, O.conformal_index(n, alpha), less, eq)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_select_ties_signs_specials(dtype):
    rng = np.random.default_rng(3)
    s = rng.normal(size=50_000).astype(dtype)
    s[rng.integers(0, s.size, 20_000)] = dtype(0.25)                # one value holds 40% of the mass
    s[:7] = [np.inf, -np.inf, 0.0, -0.0, np.nan, 1e-40, -1e-40]
    for alpha in (0.9, 0.5, 0.3, 0.1, 0.0001, 0.99999):
        thr, st = _select(s, alpha)
        want = np.sort(s)[O.conformal_index(s.size, alpha)]
        assert thr[0] == want or (np.isnan(thr[0]) and np.isnan(want))
    const = np.full(4097, 0.125, dtype)
    thr, st = _select(const)
    assert thr[0] == 0.125 and st[0, 4] == 0 and st[0, 5] == 4097


@pytest.mark.parametrize("n", [100_000, 1_000_000])
def test_select_per_group_and_empty_group(n):
    """Per-group thresholds, the reference's own semantics (src/phase3_facter_mitigation.py:159-178), at the
    size BASELINE.md timed on the CPU (1M scores in 31 groups) and below."""
    rng = np.random.default_rng(8)
    G = 31
    s = np.abs(rng.normal(0.05, 0.03, size=n))
    g = rng.integers(0, G - 1, size=n)                               # group G-1 stays empty
    for dtype in (np.float32, np.float64):
        thr, st = _select(s.astype(dtype), 0.1, g, G)
        want = O.conformal_thresholds_by_group(s.astype(dtype), g, G, 0.1)
        assert [float(x) for x in thr] == [float(x) for x in want]
        assert thr[G - 1] == 0.5 and st[G - 1, 6] == 1
    thr, st = _select(s.astype(np.float32), 0.2, g % 3, 3)          # G <= 8: shared-memory histograms
    assert [float(x) for x in thr] == [float(x) for x in O.conformal_thresholds_by_group(s.astype(np.float32), g % 3, 3, 0.2)]


def test_select_chunked_histograms_equal_single_shot():
    """Simulated world_size: summing per-shard histograms gives the single-GPU result (bit-exact)."""
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(5)
    s = np.abs(rng.normal(0.05, 0.03, size=300_000)).astype(np.float32)
    whole, st_whole = _select(s)
    shards = [torch.from_numpy(x).cuda() for x in np.array_split(s, 4)]
    lib, P = ops._lib, ops._ptr
    state = torch.zeros((1, 8), dtype=torch.int64, device="cuda:0")
    thr = torch.empty(1, dtype=torch.float32, device="cuda:0")
    for p in range(3):
        parts = []
        for sh in shards:
            h = torch.zeros((1, 2048), dtype=torch.int64, device="cuda:0")
            lib.call("flr_select_hist", P(sh), 0, None, sh.numel(), 1, p, P(state), P(h), ops._stream())
            parts.append(h)
        total = torch.stack(parts).sum(0)
        lib.call("flr_select_scan", 0, 1, p, 0.9, None, P(total), P(state), P(thr), ops._stream())
    assert thr.cpu().numpy()[0] == whole[0] and np.array_equal(state.cpu().numpy(), st_whole)


def test_schedule_and_violation_match_round_recurrence():
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(6)
    U, V, R = 4001, 31, 4
    s = np.abs(rng.normal(0.05, 0.03, size=(U, V))).astype(np.float32)
    t = torch.from_numpy(s).cuda()
    for alpha, qmin in ((0.1, 0.0), (0.1, 0.09), (1e-9, 0.0)):      # last: threshold = max -> no violation
        thr0, state = ops.radix_select(t, alpha=alpha)
        thr = ops.threshold_schedule(thr0, state, gamma=0.95, q_min=qmin, rounds=R)
        flags, counts, vcounts = ops.violation_detect(t, thr, rounds=R, n_variants=V)
        want = O.violation_rounds(s, O.conformal_threshold(s, alpha), 0.95, R, qmin)
        assert np.array_equal(thr.cpu().numpy(), want["thresholds"])
        assert np.array_equal(flags.cpu().numpy(), want["flags"])          # flags bit-exact
        assert counts.cpu().tolist() == want["counts"].tolist()
        assert np.array_equal(vcounts.cpu().numpy(), want["variant_counts"])
    with pytest.raises(ops._lib.FlrError):
        ops.threshold_schedule(thr0, state, gamma=1.5)
    # a negative order statistic with scores above it: the closed form does not hold, NaN says so
    neg = torch.from_numpy((s - 1.0).astype(np.float32)).cuda()
    thr0, state = ops.radix_select(neg, alpha=0.1)
    assert bool(torch.isnan(ops.threshold_schedule(thr0, state, rounds=R)).all())


def test_violation_small_and_odd_sizes():
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(7)
    for (U, V, R) in [(1, 1, 1), (3, 31, 8), (1000, 7, 3), (257, 124, 4)]:
        s = rng.random((U, V)).astype(np.float32)
        thr = np.sort(rng.random(R).astype(np.float32))[::-1].copy()
        flags, counts, vc = ops.violation_detect(torch.from_numpy(s).cuda(), torch.from_numpy(thr).cuda(),
                                                 rounds=R, n_variants=V)
        want = np.zeros((U, V), np.uint8)
        for r in range(R):
            want |= ((s > thr[r]).astype(np.uint8) << r).astype(np.uint8)
        assert np.array_equal(flags.cpu().numpy(), want)
        assert counts.cpu().tolist() == [int((s > thr[r]).sum()) for r in range(R)]
        assert np.array_equal(vc.cpu().numpy(), np.stack([(s > thr[r]).sum(0) for r in range(R)], 1))


def test_dropin_conformal_thresholds(golden):
    from fairness_llm_replication_b200 import phase3_facter_mitigation as p3
    for c in golden["conformal_thresholds"]:
        df = pd.DataFrame({"group": c["group"], "nonconformity": c["nonconformity"]})
        got = p3.calculate_conformal_thresholds(df, c["alpha"])
        assert {k: float(v) for k, v in got.items()} == c["thresholds"]
        assert list(got.keys()) == list(pd.unique(df["group"]))
    with pytest.raises(KeyError):
        p3.calculate_conformal_thresholds(pd.DataFrame([]))


def test_dropin_conformal_filtering(golden):
    from fairness_llm_replication_b200 import phase3_facter_mitigation as p3
    for c in golden["conformal_filtering"]:
        got = p3.apply_conformal_filtering(c["recs"], c["profiles"], c["thresholds"])
        assert got == c["filtered"] and list(got.keys()) == list(c["filtered"].keys())


def test_keep_len_exhaustive_thresholds():
    from fairness_llm_replication_b200 import ops
    lens = np.arange(0, 40, dtype=np.int32)
    for thr in [0.0, 0.05, 0.5, 0.55, 0.6000000000000001, 0.8, 0.85, 0.95, 1.0, 1.01, -0.3, 1 - 3 * 0.05, 1 - 7 * 0.05]:
        got = ops.conformal_keep_len(torch.from_numpy(lens).cuda(), torch.tensor([thr], dtype=torch.float64).cuda())
        assert got.cpu().tolist() == O.conformal_keep_len(lens, thr).tolist()
