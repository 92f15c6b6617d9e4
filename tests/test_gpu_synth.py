"""Device twin of the synthetic genotype generator and the staging kernels."""
import numpy as np
import pytest
import torch

from gnn_b200 import ops
from oracle import synth

pytestmark = pytest.mark.gpu


def test_synth_matches_oracle_bit_for_bit():
    for n, p, row0, p_total in ((37, 101, 0, 101), (64, 256, 1000, 256), (5, 7, 3, 50)):
        X = ops.synth_genotypes(n, p, "cuda", row0=row0, p_total=p_total)
        want = synth.genotypes(n, p, row0=row0, p_total=p_total)
        assert np.array_equal(X.cpu().numpy(), want)


def test_gather_sqnorm_convert():
    X = ops.synth_genotypes(100, 77, "cuda")
    rows = torch.tensor([5, 99, 0, 5, 42], device="cuda")
    G = ops.gather_rows(X, rows)
    assert G.stride(0) % 16 == 0 and torch.equal(G, X[rows])
    assert torch.equal(ops.row_sqnorm(X), (X.long() ** 2).sum(1))
    assert torch.equal(ops.i8_to_f32(X), X.float())
    A = torch.randn(1000, 100, device="cuda")
    torch.testing.assert_close(ops.colsum(A), A.sum(0), rtol=1e-4, atol=1e-4)
