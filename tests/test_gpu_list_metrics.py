"""GPU parity of the list-metric kernels (through the C ABI) against the CPU oracle and the
reference's golden vectors.  Integer outputs bit-exact; float ratios within 1e-6 relative."""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O

pytestmark = pytest.mark.gpu
RTOL = 1e-6          # north_star: rank metrics within 1e-6 relative


def _run(neutral, nlen, lists, lens, **kw):
    from fairness_llm_replication_b200 import ops
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    res = ops.list_metrics(t(neutral), t(nlen), t(lists), t(lens), want_union=True, **kw)
    torch.cuda.synchronize()
    return res


def _check(neutral, nlen, lists, lens, vocab, k_cap=None, empty_policy=1, rank=True):
    res = _run(neutral, nlen, lists, lens, vocab=vocab, k_cap=k_cap, empty_policy=empty_policy, rank_metrics=rank)
    want = O.list_metrics(neutral, nlen, lists, lens, k_cap=k_cap, empty_policy=empty_policy)
    assert np.array_equal(res.inter.cpu().numpy(), want["inter"])
    assert np.array_equal(res.union.cpu().numpy(), want["union"])
    np.testing.assert_allclose(res.jaccard.cpu().numpy(), want["jaccard"], rtol=RTOL, atol=0)
    sums = res.variant_sums.cpu().numpy()
    np.testing.assert_allclose(sums[:, 0], want["jaccard"].sum(0), rtol=1e-12)
    assert np.array_equal(sums[:, 3], np.full(lists.shape[1], lists.shape[0]))
    if rank:
        np.testing.assert_allclose(res.serp.cpu().numpy(), want["serp"], rtol=RTOL, atol=0)
        np.testing.assert_allclose(res.prag.cpu().numpy(), want["prag"], rtol=RTOL, atol=0)
        assert np.array_equal(sums[:, 1], want["serp_num"].sum(0))          # integer sums: exact
        assert np.array_equal(sums[:, 2], want["prag_num"].sum(0))
        means = res.variant_means().cpu().numpy()
        np.testing.assert_allclose(means[:, 1], want["serp"].mean(0), rtol=1e-12)
    assert int(res.status.item()) == 0
    return res, want


def _random_case(rng, U, V, K, vocab, p_dup=0.3):
    neutral = rng.integers(0, vocab, size=(U, K)).astype(np.int32)
    lists = np.where(rng.random((U, V, K)) < 0.5, neutral[:, None, :], rng.integers(0, vocab, size=(U, V, K)))
    lists = lists.astype(np.int32)
    shuffle = rng.random((U, V)) < 0.5
    perm = rng.permuted(np.broadcast_to(np.arange(K), (U, V, K)), axis=-1)
    lists = np.where(shuffle[..., None], np.take_along_axis(lists, perm, -1), lists)
    nlen = np.where(rng.random(U) < 0.2, rng.integers(0, K + 1, U), K).astype(np.int32)
    lens = np.where(rng.random((U, V)) < 0.2, rng.integers(0, K + 1, (U, V)), K).astype(np.int32)
    if p_dup == 0:
        pass
    return neutral, nlen, lists, lens


@pytest.mark.parametrize("forced", ["", "r"])
@pytest.mark.parametrize("V,K", [(31, 20), (31, 50), (31, 64), (32, 10), (7, 13), (16, 3), (2, 1), (64, 8)])
def test_table_kernel_random(V, K, forced, monkeypatch):
    """Every shape twice: with the planner's own choice, and with FLR_LM_KERNEL=r, which keeps the tile kernels
    (register / table) in use for shapes the warp kernel would otherwise take."""
    if forced:
        monkeypatch.setenv("FLR_LM_KERNEL", forced)
    else:
        monkeypatch.delenv("FLR_LM_KERNEL", raising=False)
    rng = np.random.default_rng(V * 100 + K)
    vocab = 3706 if K > 8 else 40            # small vocab => many duplicates and hits
    case = _random_case(rng, 203, V, K, vocab)
    _check(*case, vocab=vocab)
    _check(*case, vocab=vocab, empty_policy=0)
    _check(*case, vocab=vocab, rank=False)
    if K > 4:
        _check(*case, vocab=vocab, k_cap=K // 2)


@pytest.mark.parametrize("V,K,vocab", [(17, 33, 3706), (24, 62, 3706), (31, 64, 3706), (19, 63, 300), (29, 50, 200), (31, 32, 64), (20, 11, 5000), (31, 20, 8000), (32, 45, 16384),
                                       (32, 64, 2000), (31, 56, 2001)])
def test_warp_kernel_shapes(V, K, vocab):
    """The warp kernel's corners: idle lanes (V < 32), rows longer than one 32-wide half (K > 32), rows below
    the unroll capacity (K between capacities), tiny vocabularies (nearly every row has a repeat and goes
    through the exact warp-cooperative pass), the largest vocabulary the table kernels take, vocabularies
    that end on a 16-byte boundary of the rank table (the spare id of position 63 is the last byte looked up)."""
    rng = np.random.default_rng(V * 1000 + K)
    case = _random_case(rng, 131, V, K, min(vocab, 4 * K))
    _check(*case, vocab=vocab)
    _check(*case, vocab=vocab, empty_policy=0, k_cap=max(1, K - 7))
    _check(*case, vocab=vocab, rank=False)


@pytest.mark.parametrize("V,K,vocab", [(33, 20, 3706), (40, 20, 3706), (64, 50, 3706), (65, 7, 500), (97, 33, 3706),
                                       (124, 20, 3706), (128, 64, 2000), (100, 48, 8000), (45, 19, 90)])
def test_warp_kernel_chunked_shapes(V, K, vocab):
    """More than 32 variants per user: the warp kernel runs once per chunk of 32 variants (a last chunk of
    one variant, odd K and so unaligned chunk rows, padded rows, the widest user the table kernels take).
    Per-variant sums land in each chunk's own slice; foreign ids of the neutral list count once per user."""
    from fairness_llm_replication_b200 import ops
    assert ops.list_metrics_kernel_name(V, K, vocab).startswith("list_metrics_warp_kernel") and \
        ops.list_metrics_kernel_name(V, K, vocab).endswith(",1>")
    rng = np.random.default_rng(V * 1000 + K)
    case = _random_case(rng, 157, V, K, min(vocab, 4 * K))
    _check(*case, vocab=vocab)
    _check(*case, vocab=vocab, empty_policy=0, k_cap=max(1, K - 5))
    _check(*case, vocab=vocab, rank=False)


def test_random_shapes_against_the_oracle():
    """Forty seeded random shapes across every kernel family (V 1..40, K 1..64, vocabularies from a few ids
    to beyond the table kernels' limit, random lengths and k_cap, both empty-set conventions)."""
    rng = np.random.default_rng(2024)
    for _ in range(40):
        V = int(rng.integers(1, 41))
        K = int(rng.integers(1, 65))
        vocab = int(rng.choice([max(2 * K, 8), 300, 3706, 9000, 20000]))
        U = int(rng.integers(1, 90))
        case = _random_case(rng, U, V, K, min(vocab, max(8, 3 * K)))
        k_cap = None if rng.random() < 0.5 else int(rng.integers(1, K + 1))
        _check(*case, vocab=vocab, k_cap=k_cap, empty_policy=int(rng.integers(0, 2)), rank=bool(rng.random() < 0.7))


def _fuzz_case(rng):
    """One random shape with everything that has ever mattered turned on at random: vocabularies on and next
    to 16-byte boundaries of the rank table, repeats, short and empty rows, foreign ids on either side."""
    V = int(rng.choice([1, 2, 3, 7, 12, 16, 17, 20, 31, 32, 33, 40, 63, 64, 65, 96, 124, 128, 130]))
    K = int(rng.choice([1, 2, 5, 9, 10, 11, 16, 19, 20, 21, 31, 32, 33, 40, 48, 50, 51, 56, 62, 63, 64]))
    vocab = int(rng.choice([K + 3, 64, 90, 255, 256, 257, 2000, 2001, 3706, 4095, 4096, 8191, 16383, 16384, 16385, 50000]))
    U = int(rng.integers(1, 70))
    ids = min(vocab, int(rng.choice([max(K // 2, 1), 2 * K, 6 * K, vocab])))
    neutral, nlen, lists, lens = _random_case(rng, U, V, K, ids)
    if rng.random() < 0.3:          # foreign ids: negative and beyond the vocabulary
        bad_l = rng.random(lists.shape) < 0.02
        bad_n = rng.random(neutral.shape) < 0.03
        lists = np.where(bad_l, rng.choice([-1, -7, vocab, vocab + 1, vocab + 70, 2 ** 31 - 1], size=lists.shape), lists).astype(np.int32)
        neutral = np.where(bad_n, rng.choice([-3, vocab, vocab + 64, 2 ** 30], size=neutral.shape), neutral).astype(np.int32)
    k_cap = None if rng.random() < 0.5 else int(rng.integers(1, K + 1))
    return (neutral, nlen, lists, lens), dict(vocab=vocab, k_cap=k_cap, empty_policy=int(rng.integers(0, 2)),
                                              rank=bool(rng.random() < 0.7))


def _check_with_foreign_ids(neutral, nlen, lists, lens, vocab, k_cap, empty_policy, rank):
    """Like _check, for inputs that may hold ids outside [0, vocab): those are skipped wherever they stand and
    counted in `status`; intersections and unions are checked against Python sets."""
    U, V, K = lists.shape
    kc = K if k_cap is None else k_cap
    ok_n, ok_l = (neutral >= 0) & (neutral < vocab), (lists >= 0) & (lists < vocab)
    if ok_n.all() and ok_l.all():
        _check(neutral, nlen, lists, lens, vocab=vocab, k_cap=k_cap, empty_policy=empty_policy, rank=rank)
        return
    res = _run(neutral, nlen, lists, lens, vocab=vocab, k_cap=k_cap, empty_policy=empty_policy, rank_metrics=rank)
    inter, union = res.inter.cpu().numpy(), res.union.cpu().numpy()
    pos = np.arange(K)
    in_n = pos[None, :] < np.minimum(nlen, kc)[:, None]
    in_l = pos[None, None, :] < np.minimum(lens, kc)[:, :, None]
    assert int(res.status.item()) == int((~ok_n & in_n).sum() + (~ok_l & in_l).sum())
    for u in range(U):
        a = {int(x) for x, o, i in zip(neutral[u], ok_n[u], in_n[u]) if o and i}
        for v in range(V):
            b = {int(x) for x, o, i in zip(lists[u, v], ok_l[u, v], in_l[u, v]) if o and i}
            assert inter[u, v] == len(a & b) and union[u, v] == len(a | b), (u, v)


def test_fuzz_shapes_against_the_oracle():
    """Random shapes over every kernel family and plan boundary (FLR_FUZZ_SHAPES of them, default 200; the
    default takes a few seconds, 3000 have been run).  Foreign-id cases check the counts against Python sets."""
    import os
    rng = np.random.default_rng(int(os.environ.get("FLR_FUZZ_SEED", "77")))
    for i in range(int(os.environ.get("FLR_FUZZ_SHAPES", "200"))):
        case, kw = _fuzz_case(rng)
        forced = rng.random() < 0.3                      # keep the tile kernels in the mix
        os.environ["FLR_LM_KERNEL"] = "r" if forced else ""
        if kw["rank"] and (kw["k_cap"] or case[2].shape[2]) > 64:
            kw["rank"] = False
        try:
            _check_with_foreign_ids(*case, **kw)
        except AssertionError as e:
            raise AssertionError("shape %d: U, V, K = %s, %s, tile kernels forced: %s: %s" % (i, case[2].shape, kw, forced, e)) from e
        finally:
            os.environ.pop("FLR_LM_KERNEL", None)


@pytest.mark.parametrize("V,K,vocab", [(1, 10, 500), (1, 20, 100000), (3, 64, 50), (124, 20, 3706), (5, 12, 70000)])
def test_compare_kernel_shapes(V, K, vocab):
    """V = 1 pair lists (string API), huge vocabularies and V > 128 go through the compare kernel."""
    rng = np.random.default_rng(K)
    case = _random_case(rng, 150, V, K, min(vocab, 60))
    _check(*case, vocab=vocab)
    _check(*case, vocab=vocab, empty_policy=0, k_cap=max(1, K - 3))


def test_jaccard_only_long_lists():
    rng = np.random.default_rng(9)
    case = _random_case(rng, 40, 2, 200, 300)
    _check(*case, vocab=300, rank=False)
    from fairness_llm_replication_b200 import _lib
    with pytest.raises(_lib.FlrError):
        _run(*case, vocab=300, rank_metrics=True)            # SERP/PRAG need k_cap <= 64


def test_heavy_duplicates_and_empties():
    rng = np.random.default_rng(1)
    U, V, K = 97, 31, 20
    neutral = rng.integers(0, 6, size=(U, K)).astype(np.int32)        # 6 distinct ids in 20 slots
    lists = rng.integers(0, 8, size=(U, V, K)).astype(np.int32)
    nlen = rng.integers(0, 3, U).astype(np.int32) * 10
    lens = rng.integers(0, 3, (U, V)).astype(np.int32) * 10
    for pol in (0, 1):
        _check(neutral, nlen, lists, lens, vocab=3706, empty_policy=pol)


def test_padding_is_never_read_as_ids():
    rng = np.random.default_rng(4)
    neutral, nlen, lists, lens = _random_case(rng, 64, 31, 20, 3706)
    pos = np.arange(20)
    g_neutral = np.where(pos[None] < nlen[:, None], neutral, -1).astype(np.int32)
    g_lists = np.where(pos[None, None] < lens[..., None], lists, 2 ** 31 - 1).astype(np.int32)
    a, _ = _check(neutral, nlen, lists, lens, vocab=3706)
    b, _ = _check(g_neutral, nlen, g_lists, lens, vocab=3706)
    assert torch.equal(a.inter, b.inter) and torch.equal(a.prag, b.prag)


def test_out_of_vocab_ids_are_counted_and_skipped():
    neutral = np.array([[1, 2, 3, 4]], np.int32)
    lists = np.array([[[1, 99, 2, -5]], [[0, 0, 0, 0]]], np.int32).reshape(1, 2, 4)
    res = _run(neutral, np.array([4], np.int32), lists, np.array([[4, 1]], np.int32), vocab=10)
    assert res.inter.cpu().tolist() == [[2, 0]] and int(res.status.item()) == 2
    assert res.union.cpu().tolist() == [[4, 5]]


@pytest.mark.parametrize("V,K,vocab", [(31, 20, 3706), (8, 12, 3706), (31, 20, 20000), (3, 64, 50), (70, 20, 3706)])
def test_foreign_ids_in_the_neutral_list_count_once_per_user(V, K, vocab):
    """Ids outside [0, vocab) are skipped wherever they stand and counted once per list row: a neutral row
    belongs to the user, not to each of its V pairs (nor to each chunk of 32 variants the warp kernel is
    launched for).  Every kernel family (warp, register, table, compare) must report the same number for the
    same input."""
    rng = np.random.default_rng(V + K)
    U = 97
    neutral, nlen, lists, lens = _random_case(rng, U, V, K, min(vocab, 60))
    bad_n = rng.random((U, K)) < 0.07
    bad_l = rng.random((U, V, K)) < 0.03
    neutral = np.where(bad_n, vocab + rng.integers(0, 5, (U, K)), neutral).astype(np.int32)
    lists = np.where(bad_l, -1 - rng.integers(0, 5, (U, V, K)), lists).astype(np.int32)
    res = _run(neutral, nlen, lists, lens, vocab=vocab)
    in_n = np.arange(K)[None, :] < nlen[:, None]
    in_l = np.arange(K)[None, None, :] < lens[:, :, None]
    expect = int((bad_n & in_n).sum() + (bad_l & in_l).sum())
    assert int(res.status.item()) == expect
    inter, union = res.inter.cpu().numpy(), res.union.cpu().numpy()
    ok = lambda x: 0 <= x < vocab
    for u in range(U):
        a = {int(x) for x in neutral[u, :nlen[u]] if ok(x)}
        for v in range(V):
            b = {int(x) for x in lists[u, v, :lens[u, v]] if ok(x)}
            assert inter[u, v] == len(a & b) and union[u, v] == len(a | b)


def test_tail_tiles_and_tiny_inputs():
    rng = np.random.default_rng(12)
    for U in (1, 2, 3, 4, 5, 7, 9, 301):
        _check(*_random_case(rng, U, 31, 20, 3706), vocab=3706)


def test_unaligned_views_take_the_plain_load_path():
    rng = np.random.default_rng(13)
    from fairness_llm_replication_b200 import ops
    dev = torch.device("cuda", 0)
    def off(a):                      # same data at a 4-byte offset: not 16-byte aligned
        flat = torch.empty(a.size + 1, dtype=torch.int32, device=dev)
        flat[1:] = torch.from_numpy(a.ravel()).to(dev)
        return flat[1:].view(a.shape)
    for V, K in ((31, 10), (40, 10), (70, 33)):          # one launch; chunks of 32 variants, even and odd K
        neutral, nlen, lists, lens = _random_case(rng, 41, V, K, 3706)
        res = ops.list_metrics(off(neutral), off(nlen), off(lists), off(lens), vocab=3706, want_union=True)
        want = O.list_metrics(neutral, nlen, lists, lens)
        assert np.array_equal(res.inter.cpu().numpy(), want["inter"])
        np.testing.assert_allclose(res.prag.cpu().numpy(), want["prag"], rtol=RTOL)


def test_synthetic_config2_full_size_matches_oracle():
    """BASELINE.json configs[1]: 6040 users x 31 variants, K=20 -- small enough for the oracle."""
    from fairness_llm_replication_b200 import ops
    U, V, K, vocab = 6040, 31, 20, 3706
    host = ops.synth_lists_host(20251203, 0, U, V, K, vocab)
    devt = ops.synth_lists(20251203, 0, U, V, K, vocab, "cuda:0")
    for h, d in zip(host, devt):
        assert np.array_equal(h, d.cpu().numpy())                  # device generator == host generator
    res, want = _check(*host, vocab=vocab)
    spreads = O.snsr_snsv(want["jaccard"].mean(0))
    got = O.snsr_snsv(res.variant_means().cpu().numpy()[:, 0])
    for a in spreads:
        assert got[a]["snsr"] == pytest.approx(spreads[a]["snsr"], rel=1e-9)
        assert got[a]["snsv"] == pytest.approx(spreads[a]["snsv"], rel=1e-9)


def test_large_k50_properties():
    """Config-5 shape (K=50) at a size the oracle cannot sweep: size-independent properties."""
    from fairness_llm_replication_b200 import ops
    U, V, K, vocab = 200_000, 31, 50, 3706
    neutral, nlen, lists, lens = ops.synth_lists(20251203, 0, U, V, K, vocab, "cuda:0")
    res = ops.list_metrics(neutral, nlen, lists, lens, vocab=vocab, want_union=True)
    # (1) a slice must equal the oracle
    sl = slice(123_400, 123_700)
    want = O.list_metrics(neutral[sl].cpu().numpy(), nlen[sl].cpu().numpy(), lists[sl].cpu().numpy(),
                          lens[sl].cpu().numpy())
    assert np.array_equal(res.inter[sl].cpu().numpy(), want["inter"])
    np.testing.assert_allclose(res.prag[sl].cpu().numpy(), want["prag"], rtol=RTOL)
    # (2) every list against itself: Jaccard 1, SERP/PRAG closed forms of the distinct prefix
    same = ops.list_metrics(neutral, nlen, neutral.view(U, 1, K).expand(U, V, K).contiguous(),
                            nlen.view(U, 1).expand(U, V).contiguous(), vocab=vocab, want_union=True)
    assert torch.equal(same.inter, same.union)
    full = (nlen == K) & (same.inter[:, 0] == K)
    assert torch.all(same.serp[full] == 1.0)
    assert torch.allclose(same.prag[full], torch.tensor((K * (K - 1) / 2) / (K * (K + 1))))
    # (3) sums are the column sums of the per-pair outputs
    assert torch.equal(res.variant_sums[:, 3], torch.full((V,), float(U), dtype=torch.float64, device="cuda:0"))
    assert torch.allclose(res.variant_sums[:, 0], res.jaccard.double().sum(0), rtol=1e-6)
    # (4) 0 <= inter <= min sizes, union consistent
    assert int(res.inter.min()) >= 0 and torch.all(res.inter <= res.union)
    # (5) idempotent
    again = ops.list_metrics(neutral, nlen, lists, lens, vocab=vocab, want_union=True)
    assert torch.equal(again.inter, res.inter) and torch.equal(again.prag, res.prag)
    assert torch.equal(again.variant_sums, res.variant_sums)       # fixed-order reduction: bit-reproducible


def test_repeated_runs_are_bit_identical_under_stress():
    """compute-sanitizer is not available on the GPU pool, so shared-memory races are hunted the blunt
    way: inputs that exercise every rare path (short rows, repeats, foreign ids, empty neutral lists)
    at a size that keeps all CTAs busy over many tiles, run several times; any race shows up as a
    difference between runs or against the oracle."""
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(77)
    U, V, K, vocab = 20011, 31, 50, 3706
    neutral = rng.integers(0, 300, size=(U, K)).astype(np.int32)                 # many repeats
    lists = np.where(rng.random((U, V, K)) < 0.6, neutral[:, None, :], rng.integers(-3, 3800, size=(U, V, K)))
    lists = lists.astype(np.int32)                                               # some ids outside [0, vocab)
    nlen = np.where(rng.random(U) < 0.3, rng.integers(0, K + 1, U), K).astype(np.int32)
    lens = np.where(rng.random((U, V)) < 0.3, rng.integers(0, K + 1, (U, V)), K).astype(np.int32)
    t = lambda a: torch.from_numpy(a).cuda()
    args = (t(neutral), t(nlen), t(lists), t(lens))
    runs = [ops.list_metrics(*args, vocab=vocab, want_union=True) for _ in range(5)]
    torch.cuda.synchronize()
    for r in runs[1:]:
        assert torch.equal(r.inter, runs[0].inter) and torch.equal(r.union, runs[0].union)
        assert torch.equal(r.serp, runs[0].serp) and torch.equal(r.prag, runs[0].prag)
        assert torch.equal(r.variant_sums, runs[0].variant_sums) and torch.equal(r.status, runs[0].status)
    sl = slice(0, 1500)
    live = np.arange(K)[None, None, :] < lens[sl][:, :, None]
    foreign = ((lists[sl] < 0) | (lists[sl] >= vocab)) & live
    clean = np.where(foreign, -1, lists[sl])
    # the oracle refuses foreign ids, the kernel skips and counts them: compare on rows without any
    ok_rows = ~foreign.any(-1)
    want = O.list_metrics(neutral[sl], nlen[sl], np.where(live, clean, 0).astype(np.int32) * ok_rows[..., None],
                          lens[sl] * ok_rows)
    got_inter = runs[0].inter[sl].cpu().numpy()
    got_prag = runs[0].prag[sl].cpu().numpy()
    assert np.array_equal(got_inter[ok_rows], want["inter"][ok_rows])
    np.testing.assert_allclose(got_prag[ok_rows], want["prag"][ok_rows], rtol=RTOL)
    n_foreign = int((((lists < 0) | (lists >= vocab)) & (np.arange(K)[None, None, :] < lens[:, :, None])).sum())
    assert int(runs[0].status.item()) == n_foreign


def test_dense_histogram_matches_per_group_kernel_and_bincount():
    """The all-groups-resident kernel (dense [U, V, K] layout) against the per-group kernel (explicit group
    ids) and numpy, on rows with short lengths, foreign ids and repeats; K = 50 (one row per warp pass),
    K = 20 (three rows side by side) and an odd K (scalar loads)."""
    from fairness_llm_replication_b200 import ops
    for U, V, K, vocab, k_cap in ((5003, 31, 50, 3706, None), (4001, 31, 20, 3706, 7), (1500, 7, 9, 500, None)):
        rng = np.random.default_rng(U)
        lists = rng.integers(-2, vocab + 40, size=(U, V, K)).astype(np.int32)
        lists[rng.random((U, V, K)) < 0.3] = 3                                   # repeats of one popular id
        lens = np.where(rng.random((U, V)) < 0.3, rng.integers(0, K + 1, (U, V)), K).astype(np.int32)
        tl, tn = torch.from_numpy(lists).cuda(), torch.from_numpy(lens).cuda()
        dense = ops.group_item_histogram(tl, tn, vocab=vocab, n_groups=V, k_cap=k_cap)
        groups = (torch.arange(U * V, device="cuda") % V).to(torch.uint8)
        per_group = ops.group_item_histogram(tl.view(U * V, K), tn.view(-1), vocab=vocab, n_groups=V, groups=groups,
                                             k_cap=k_cap)
        assert torch.equal(dense, per_group)
        kc = K if k_cap is None else k_cap
        live = (np.arange(K)[None, None, :] < np.minimum(lens, kc)[:, :, None]) & (lists >= 0) & (lists < vocab)
        want = np.stack([np.bincount(lists[:, g][live[:, g]], minlength=vocab) for g in range(V)])
        assert np.array_equal(dense.cpu().numpy(), want)


def test_dense_histogram_counter_halves_do_not_overflow():
    """Every position of every row holds the same id: a 16-bit counter half would wrap if the kernel did not
    flush at its chunk boundaries (every 1310 users at K = 50; each CTA here walks 2800)."""
    from fairness_llm_replication_b200 import ops
    sms = ops.device_info()["sm_count"]
    U, V, K, vocab = sms * 2800, 31, 50, 3706          # 70,000 per half per CTA without the flushes
    lists = torch.full((U, V, K), 7, dtype=torch.int32, device="cuda")
    lists[:, :, 1::2] = 1853 + 7                                                 # the other half of the same word
    lens = torch.full((U, V), K, dtype=torch.int32, device="cuda")
    h = ops.group_item_histogram(lists, lens, vocab=vocab, n_groups=V)
    want = torch.zeros((V, vocab), dtype=torch.int64, device="cuda")
    want[:, 7] = U * (K // 2)
    want[:, 1860] = U * (K // 2)
    assert torch.equal(h, want)


# ---- string API against the reference's golden outputs ------------------------------------------
def test_dropin_individual_fairness(golden):
    from fairness_llm_replication_b200 import utils
    for c in golden["individual_fairness"]:
        score, sims = utils.calculate_individual_fairness([tuple(p) for p in c["pairs"]], c["recs"])
        assert sims == c["sims"] and float(score) == c["score"]
        assert all(type(s) is float for s in sims)


def test_dropin_measure_individual_fairness(golden):
    from fairness_llm_replication_b200 import phase1_bias_detection as p1
    for c in golden["measure_individual_fairness"]:
        score, det = p1.measure_individual_fairness(c["profiles"], c["recs"])
        assert float(score) == c["score"] and det["similarity_scores"] == c["sims"]
        assert det["num_pairs"] == c["num_pairs"] and float(det["std_similarity"]) == c["std"]
        assert float(det["mean_similarity"]) == c["score"]


def test_dropin_quality_preservation(golden):
    from fairness_llm_replication_b200 import phase3_facter_mitigation as p3
    for c in golden["quality_preservation"]:
        r = p3.measure_quality_preservation(c["orig"], c["mitg"])
        assert float(r["average_overlap"]) == c["average_overlap"]
        assert float(r["quality_preservation_pct"]) == c["quality_preservation_pct"]
        assert r["num_comparisons"] == c["num_comparisons"]


def test_dropin_phase3_final(golden):
    from fairness_llm_replication_b200 import phase3_final as fin
    for c in golden["final_jaccard"]:
        assert fin.jaccard(set(c["s1"]), set(c["s2"])) == c["jaccard"]
    for c in golden["measure_fairness"]:
        assert float(fin.measure_fairness(c["recs_by_group"])) == c["fairness"]


def test_dropin_phase3_aggressive(golden):
    """The earlier copy of the two helpers (src/phase3_aggressive.py:174-195): own golden vectors."""
    from fairness_llm_replication_b200 import phase3_aggressive as agg
    for c in golden["aggressive"]["jaccard"]:
        assert agg.jaccard(set(c["s1"]), set(c["s2"])) == c["jaccard"]
    for c in golden["aggressive"]["measure_fairness"]:
        assert float(agg.measure_fairness(c["recs_by_group"])) == c["fairness"]


def test_pair_jaccard_bitsets():
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(21)
    # K = 300: intersections above 255 (16-bit staging); M odd: the last 2 x 2 block is clipped
    for (U, M, K, vocab) in [(5, 124, 20, 3706), (3, 7, 5, 40), (2, 33, 50, 3706), (3, 9, 300, 1000), (301, 124, 20, 3706)]:
        lists = rng.integers(0, min(vocab, 200 if K < 100 else 400), size=(U, M, K)).astype(np.int32)
        lens = rng.integers(0, K + 1, size=(U, M)).astype(np.int32)
        inter, sizes = ops.pair_jaccard(torch.from_numpy(lists).cuda(), torch.from_numpy(lens).cuda(), vocab=vocab)
        want = O.pair_jaccard(lists, lens)
        assert np.array_equal(inter.cpu().numpy(), want["inter"])
        assert np.array_equal(sizes.cpu().numpy(), want["sizes"])


# ---- N1: demographic parity ---------------------------------------------------------------------
def test_group_item_histogram_and_js_distances():
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(31)
    U, V, K, vocab = 3001, 31, 20, 3706
    _, _, lists, lens = ops.synth_lists_host(20251203, 0, U, V, K, vocab)
    lists = lists.copy()
    lists[5, 3, :4] = 7                                            # multiplicity must be counted
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    hist = ops.group_item_histogram(t(lists), t(lens), vocab=vocab, n_groups=V)          # implicit groups r % V
    groups = np.tile(np.arange(V, dtype=np.uint8), U)
    want = O.group_item_counts(lists.reshape(-1, K), lens.reshape(-1), groups, V, vocab)
    assert np.array_equal(hist.cpu().numpy(), want)
    perm = rng.permutation(U * V)                                   # explicit group ids, shuffled rows
    hist2 = ops.group_item_histogram(t(lists.reshape(-1, K)[perm]), t(lens.reshape(-1)[perm]), vocab=vocab,
                                     n_groups=V, groups=t(groups[perm]))
    assert torch.equal(hist, hist2)
    hist3 = ops.group_item_histogram(t(lists), t(lens), vocab=vocab, n_groups=V, k_cap=10)
    assert np.array_equal(hist3.cpu().numpy(),
                          O.group_item_counts(lists.reshape(-1, K), lens.reshape(-1), groups, V, vocab, k_cap=10))
    hist[4] = 0                                                     # an empty group: its pairs are skipped
    dist, valid = ops.js_distance_matrix(hist)
    score, div = O.demographic_parity_from_counts(hist.cpu().numpy())
    got = [d for d, ok in zip(dist.cpu().tolist(), valid.cpu().tolist()) if ok]
    assert len(got) == len(div) == 30 * 29 // 2
    np.testing.assert_allclose(got, div, rtol=1e-12)


def test_dropin_demographic_parity(golden):
    from fairness_llm_replication_b200 import utils, phase1_bias_detection as p1
    for c in golden["demographic_parity"]:
        score, det = utils.calculate_demographic_parity(dict(c["groups"]))
        assert det["distributions"] == c["distributions"]              # count / total: exact
        assert det["divergences"] == pytest.approx(c["divergences"], rel=1e-12, abs=1e-15)
        assert float(score) == pytest.approx(c["score"], rel=1e-12)
        assert float(det["avg_divergence"]) == pytest.approx(c["avg_divergence"], rel=1e-12, abs=1e-15)
        s2, _ = p1.measure_demographic_parity(dict(c["groups"]))
        assert float(s2) == float(score)


def test_dropin_divergences_and_bias_reduction(golden):
    """utils.calculate_kl_divergence / calculate_js_divergence (host helpers, src/utils.py:70-102) and the phase-3
    caller of demographic parity, measure_bias_reduction (src/phase3_facter_mitigation.py:280-331)."""
    from fairness_llm_replication_b200 import phase3_facter_mitigation as p3, utils
    for c in golden["divergences"]:
        assert utils.calculate_kl_divergence(c["p"], c["q"]) == pytest.approx(c["kl"], rel=1e-12, abs=1e-15)
        assert utils.calculate_js_divergence(c["p"], c["q"]) == pytest.approx(c["js"], rel=1e-12, abs=1e-8)
    for c in golden["bias_reduction"]:
        got = p3.measure_bias_reduction(c["original"], c["mitigated"], c["profiles"])
        assert set(got) == set(c["result"])
        for k, v in c["result"].items():
            assert float(got[k]) == pytest.approx(v, rel=1e-9, abs=1e-12)


def test_dropin_equal_opportunity(golden):
    """N4: bit-identical to the reference (integer counts from the GPU, the reference's own float ops)."""
    from fairness_llm_replication_b200 import utils, phase1_bias_detection as p1
    for c in golden["equal_opportunity"]:
        d = dict(c["groups"])
        score, by_group = utils.calculate_equal_opportunity(d, set(c["relevant"]))
        assert [[k, float(v)] for k, v in by_group.items()] == c["by_group"]
        assert float(score) == c["score"]
        if "measure" in c and c["relevant"][0] == "Heat":
            continue                      # measure_* uses the reference's own fixed qualified set
    d = dict(golden["equal_opportunity"][-1]["groups"])
    s2, det = p1.measure_equal_opportunity(d, None)
    m = golden["equal_opportunity"][-1]["measure"]
    assert float(s2) == m["score"] and float(det["variance"]) == m["variance"]
    assert float(det["min_max_ratio"]) == m["min_max_ratio"]


def test_equal_opportunity_counts_kernel():
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(5)
    G, vocab = 31, 3706
    hist = rng.integers(0, 4, size=(G, vocab)) * (rng.random((G, vocab)) < 0.2)
    hist[7] = 0
    relevant = (rng.random(vocab) < 0.05).astype(np.uint8)
    rel, tot = ops.equal_opportunity_counts(torch.from_numpy(hist.astype(np.int64)).cuda(), torch.from_numpy(relevant).cuda())
    assert rel.cpu().tolist() == [int(((hist[g] > 0) & (relevant > 0)).sum()) for g in range(G)]
    assert tot.cpu().tolist() == hist.sum(1).tolist()


def test_phase1_metric_step_on_a_results_file(golden, tmp_path):
    """N2: a phase1_results.json in the reference's schema, re-scored end to end on the GPU path."""
    import json
    from fairness_llm_replication_b200 import phase1_bias_detection as p1
    c = golden["phase1_metric_step"]
    path = tmp_path / "phase1_results.json"
    path.write_text(json.dumps({"metadata": {"model": "x"}, "profiles": c["profiles"],
                                "recommendations": c["recommendations"], "metrics": {}}))
    m = p1.rescore_phase1_results(str(path))
    assert float(m["demographic_parity"]["gender"]["score"]) == pytest.approx(c["dp_gender"], rel=1e-12)
    assert float(m["demographic_parity"]["age"]["score"]) == pytest.approx(c["dp_age"], rel=1e-12)
    assert float(m["individual_fairness"]["score"]) == c["if_score"]
    assert float(m["individual_fairness"]["details"]["std_similarity"]) == c["if_std"]
    assert float(m["equal_opportunity"]["score"]) == c["eo_score"]
    assert [[k, float(v)] for k, v in m["equal_opportunity"]["details"]["opportunity_by_group"].items()] == c["eo_by_group"]


def test_dropin_edge_inputs_match_the_port():
    """Inputs the golden file does not hold: very long lists (Jaccard-only compare kernel), non-ASCII and
    near-identical titles, non-string items, the same list object on both sides."""
    from fairness_llm_replication_b200 import utils, phase3_final as fin
    from fairness_llm_replication_b200 import phase3_facter_mitigation as p3
    from oracle import ref_port as R
    rng = np.random.default_rng(8)
    long_a = [f"title {i}" for i in rng.integers(0, 900, 300)]
    long_b = [f"title {i}" for i in rng.integers(0, 900, 412)]
    shared = ["Amélie", "Amelie", "amelie ", "千と千尋の神隠し", "", " ", "Se7en", "Se7en"]
    recs = {"a": long_a, "b": long_b, "c": shared, "d": list(reversed(shared)), "e": [1, 2.0, (3, 4), None, 1],
            "f": [2, None, "1"], "g": shared}
    pairs = [("a", "b"), ("b", "a"), ("c", "d"), ("c", "g"), ("e", "f"), ("a", "c"), ("g", "g")]
    got = utils.calculate_individual_fairness(pairs, recs)
    want = R.calculate_individual_fairness(pairs, recs)
    assert got[1] == want[1] and float(got[0]) == float(want[0])
    assert fin.jaccard(set(long_a), set(long_b)) == R.jaccard(set(long_a), set(long_b))
    q = p3.measure_quality_preservation({"u": long_a, "v": {"recommendations": shared}}, {"u": long_b, "v": shared})
    w = R.measure_quality_preservation({"u": long_a, "v": {"recommendations": shared}}, {"u": long_b, "v": shared})
    assert float(q["average_overlap"]) == float(w["average_overlap"]) and q["num_comparisons"] == w["num_comparisons"]
    d = {"g1": [long_a, shared], "g2": [long_b, shared[:3]], "g3": [["x"]]}
    s_got, det = utils.calculate_demographic_parity(d)
    s_want, det_w = R.calculate_demographic_parity(d)
    assert det["distributions"] == det_w["distributions"]
    assert det["divergences"] == pytest.approx(det_w["divergences"], rel=1e-12)
