"""GPU parity of the tcgen05 similarity GEMM (config 4) against the float64 oracle on the same
bf16-rounded inputs.  Tolerance 2e-3 absolute (north_star: bf16 GEMM similarity)."""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O

pytestmark = pytest.mark.gpu
ATOL = 2e-3


def _bf16_to_f32(t):
    return t.float().cpu().numpy()


@pytest.mark.parametrize("U,M,D", [(5, 124, 384), (301, 124, 384), (3, 128, 64), (7, 50, 512), (4, 33, 128),
                                   (2, 1, 64), (149, 124, 384)])
def test_pair_similarity_matches_oracle(U, M, D):
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(U * 1000 + M + D)
    base = rng.normal(size=(U, 1, D)).astype(np.float32)
    x = (base + 0.5 * rng.normal(size=(U, M, D))).astype(np.float32)
    xh = ops.normalize_rows_bf16(torch.from_numpy(x).cuda())
    assert xh.dtype == torch.bfloat16 and tuple(xh.shape) == (U, M, D)
    got_x = _bf16_to_f32(xh)
    want_x = O.normalize_rows_bf16(x)
    assert np.abs(got_x - want_x).max() <= 2 ** -8              # at most one bf16 ulp apart
    assert (got_x != want_x).mean() < 1e-3
    S = ops.pair_similarity(xh)
    torch.cuda.synchronize()
    want = O.pair_similarity(got_x)
    got = S.cpu().numpy()
    assert got.shape == (U, M, M)
    assert np.abs(got - want).max() <= ATOL
    assert np.abs(got - want).max() <= 2e-5                      # fp32 accumulation of exact bf16 products
    assert np.allclose(got, np.swapaxes(got, 1, 2), atol=1e-6)
    assert np.abs(np.diagonal(got, axis1=1, axis2=2) - 1.0).max() < 1e-2


def test_pair_similarity_ignores_neighbouring_users():
    """Rows 124..127 of a tile belong to the next user (or are out of bounds): they must not leak."""
    from fairness_llm_replication_b200 import ops
    rng = np.random.default_rng(3)
    x = rng.normal(size=(6, 124, 384)).astype(np.float32)
    xh = ops.normalize_rows_bf16(torch.from_numpy(x).cuda())
    full = ops.pair_similarity(xh)
    poisoned = xh.clone()
    poisoned[3] = float("nan")
    part = ops.pair_similarity(poisoned)
    torch.cuda.synchronize()
    for u in (0, 1, 2, 4, 5):
        assert torch.equal(full[u], part[u])
    assert torch.isnan(part[3]).all()


def test_pair_similarity_rejects_unsupported_shapes():
    from fairness_llm_replication_b200 import ops, _lib
    x = torch.zeros((2, 130, 64), dtype=torch.bfloat16, device="cuda")
    with pytest.raises(_lib.FlrError):
        ops.pair_similarity(x)
    with pytest.raises(_lib.FlrError):
        ops.pair_similarity(torch.zeros((2, 8, 96), dtype=torch.bfloat16, device="cuda"))
