"""flr_list_metrics_static: list metrics for hosts that cannot provide scratch memory.

Some users of this library do not get to allocate on the GPU at all: their process runs inside
a memory plan made elsewhere (one pool owns all of HBM, or the user shard was sized to fill the
device), or their step loop must not hit the device-wide synchronisation that cudaMalloc
causes.  They own input and output buffers and nothing more, so this entry point keeps its
partial sums in a fixed array inside the library instead of asking for a scratch pointer.

The test goes through the C ABI exactly as such a host would (raw pointers, no scratch
argument) and compares every output with the regular entry point, flr_list_metrics, run on
the same inputs: integers must be identical, floats within the 1e-6 relative tolerance the
other list-metric tests use.  It does so on two calls in a row with different shapes, because
all calls share the one array and the second call starts from whatever the first left in it.

This file is named so that pytest collects it last.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def _inputs(seed, U, V, K, vocab):
    rng = np.random.default_rng(seed)
    neutral = rng.integers(0, vocab, size=(U, K)).astype(np.int32)
    lists = np.where(rng.random((U, V, K)) < 0.5, neutral[:, None, :], rng.integers(0, vocab, size=(U, V, K)))
    nlen = np.where(rng.random(U) < 0.1, rng.integers(0, K + 1, U), K).astype(np.int32)
    lens = np.where(rng.random((U, V)) < 0.1, rng.integers(0, K + 1, (U, V)), K).astype(np.int32)
    dev = torch.device("cuda", 0)
    return [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (neutral, nlen, lists.astype(np.int32), lens)]


def _outputs(U, V):
    dev = torch.device("cuda", 0)
    mk = lambda dt: torch.full((U, V), -7, dtype=dt, device=dev)
    return {"inter": mk(torch.int32), "union": mk(torch.int32), "jac": mk(torch.float32),
            "serp": mk(torch.float32), "prag": mk(torch.float32),
            "sums": torch.full((V, 4), -7.0, dtype=torch.float64, device=dev),
            "status": torch.zeros(1, dtype=torch.int32, device=dev)}


def _call(static, tensors, U, V, K, vocab, k_cap, policy, rank):
    """One call through the C ABI; `static` picks the entry point that takes no scratch."""
    from fairness_llm_replication_b200 import _lib
    lib = _lib.load()
    p = lambda t: C.c_void_p(t.data_ptr())
    o = _outputs(U, V)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    head = [p(t) for t in tensors] + [U, V, K, k_cap, vocab, policy, rank,
                                      p(o["inter"]), p(o["union"]), p(o["jac"]), p(o["serp"]), p(o["prag"]),
                                      p(o["sums"]), p(o["status"])]
    if static:
        _lib.call("flr_list_metrics_static", *head, stream)
    else:
        nbytes = lib.flr_list_metrics_scratch_bytes(U, V, K, vocab)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")
        _lib.call("flr_list_metrics", *head, p(scratch), nbytes, stream)
    torch.cuda.synchronize()
    return o


def _same(a, b, rank):
    assert torch.equal(a["inter"], b["inter"]) and torch.equal(a["union"], b["union"])
    assert torch.equal(a["status"], b["status"])
    torch.testing.assert_close(a["jac"], b["jac"], rtol=RTOL, atol=0)
    if rank:
        torch.testing.assert_close(a["serp"], b["serp"], rtol=RTOL, atol=0)
        torch.testing.assert_close(a["prag"], b["prag"], rtol=RTOL, atol=0)
    torch.testing.assert_close(a["sums"], b["sums"], rtol=1e-12, atol=0)


def test_static_scratch_entry_matches_regular_entry_twice_in_a_row():
    """Two consecutive calls of the no-scratch entry point, different shapes, each checked against
    flr_list_metrics.  The second call reuses the array the first one wrote into."""
    shapes = [dict(seed=1, U=5003, V=31, K=50, vocab=3706, k_cap=50, policy=1, rank=1),    # table kernel
              dict(seed=2, U=777, V=31, K=20, vocab=3706, k_cap=10, policy=0, rank=1),     # fewer CTAs, stale tail
              dict(seed=3, U=20000, V=1, K=12, vocab=900, k_cap=12, policy=1, rank=0),     # compare kernel
              dict(seed=4, U=901, V=70, K=50, vocab=3706, k_cap=50, policy=1, rank=1)]     # three chunks of variants
    for s in shapes:
        tensors = _inputs(s["seed"], s["U"], s["V"], s["K"], s["vocab"])
        args = (tensors, s["U"], s["V"], s["K"], s["vocab"], s["k_cap"], s["policy"], s["rank"])
        got = _call(True, *args)
        want = _call(False, *args)
        _same(got, want, s["rank"])
        again = _call(True, *args)                     # and once more on the same inputs
        _same(again, want, s["rank"])


def test_static_scratch_entry_matches_the_oracle():
    """The no-scratch entry point against the CPU oracle directly (not only against its sibling):
    exact intersections, unions and integer sums, ratios within 1e-6."""
    from oracle import np_oracle as O
    U, V, K, vocab = 1201, 31, 50, 3706
    tensors = _inputs(11, U, V, K, vocab)
    got = _call(True, tensors, U, V, K, vocab, K, 1, 1)
    neutral, nlen, lists, lens = (t.cpu().numpy() for t in tensors)
    want = O.list_metrics(neutral, nlen, lists, lens, k_cap=K, empty_policy=1)
    assert np.array_equal(got["inter"].cpu().numpy(), want["inter"])
    assert np.array_equal(got["union"].cpu().numpy(), want["union"])
    np.testing.assert_allclose(got["serp"].cpu().numpy(), want["serp"], rtol=RTOL, atol=0)
    np.testing.assert_allclose(got["prag"].cpu().numpy(), want["prag"], rtol=RTOL, atol=0)
    sums = got["sums"].cpu().numpy()
    assert np.array_equal(sums[:, 1], want["serp_num"].sum(0)) and np.array_equal(sums[:, 2], want["prag_num"].sum(0))
    np.testing.assert_allclose(sums[:, 0], want["jaccard"].sum(0), rtol=1e-12)


def test_static_scratch_entry_allocates_nothing():
    """What the entry point is for: once the first call on the device has looked up the array's address,
    further calls leave the device's free memory and torch's allocator untouched."""
    from fairness_llm_replication_b200 import _lib
    U, V, K, vocab = 4000, 31, 20, 3706
    tensors = _inputs(12, U, V, K, vocab)
    o = _outputs(U, V)
    p = lambda t: C.c_void_p(t.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    args = [p(t) for t in tensors] + [U, V, K, K, vocab, 1, 1, p(o["inter"]), p(o["union"]), p(o["jac"]), p(o["serp"]),
                                      p(o["prag"]), p(o["sums"]), p(o["status"]), stream]
    _lib.call("flr_list_metrics_static", *args)          # first call: address look-up, module load
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    stats0 = torch.cuda.memory_stats()
    for _ in range(5):
        _lib.call("flr_list_metrics_static", *args)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    stats1 = torch.cuda.memory_stats()
    assert free1 == free0
    for key in ("allocation.all.allocated", "segment.all.allocated", "reserved_bytes.all.current"):
        assert stats1.get(key) == stats0.get(key)


def test_static_scratch_entry_serves_large_compare_shapes_and_refuses_what_cannot_fit():
    """Compare-kernel shapes (V = 1 pair lists of the string API) run with as many blocks as the array has
    room for, whatever the number of pairs; a shape whose first V blocks alone overflow the array is turned
    down with a message that names both sizes, and nothing is launched."""
    from fairness_llm_replication_b200 import _lib
    U, V, K, vocab = 200_000, 1, 8, 500
    tensors = _inputs(4, U, V, K, vocab)
    _same(_call(True, tensors, U, V, K, vocab, K, 1, 0), _call(False, tensors, U, V, K, vocab, K, 1, 0), 0)
    U, V, K, vocab = 3, 600, 4, 100000          # 600 blocks x 128 threads x 32 B > 2 MiB
    tensors = _inputs(6, U, V, K, vocab)
    with pytest.raises(_lib.FlrError) as err:
        _call(True, tensors, U, V, K, vocab, K, 1, 0)
    assert err.value.code == _lib.E_UNSUPPORTED
    assert str(_lib.LIST_METRICS_STATIC_SCRATCH_BYTES) in str(err.value) and "scratch" in str(err.value)
    _same(_call(True, _inputs(5, 50, 31, 20, 3706), 50, 31, 20, 3706, 20, 1, 1),
          _call(False, _inputs(5, 50, 31, 20, 3706), 50, 31, 20, 3706, 20, 1, 1), 1)   # still usable afterwards
