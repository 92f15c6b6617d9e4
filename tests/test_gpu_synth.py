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


@pytest.mark.parametrize("n,p", [(1, 1), (3, 5), (17, 64), (100, 1279), (257, 4099)])
def test_pack_unpack_2bit_round_trip(n, p):
    """The PCIe format of the markers: host packer, device packer and device unpacker agree bit for bit."""
    import numpy as np
    rng = np.random.default_rng(n * 31 + p)
    X = rng.integers(0, 4, size=(n, p)).astype(np.int8)
    host = ops.pack_2bit_host(X)
    assert host.shape == (n, ops.packed_stride(p)) and host.dtype == np.uint8
    for j in (0, p // 2, p - 1):
        assert np.array_equal((host[:, j // 4] >> (2 * (j % 4))) & 3, X[:, j])
    Xd = torch.from_numpy(X).cuda()
    dev_packed = ops.pack_2bit(Xd)
    assert np.array_equal(dev_packed.cpu().numpy(), host)
    back = ops.unpack_2bit(torch.from_numpy(host).cuda(), p)
    assert back.stride(0) % 16 == 0 and np.array_equal(back.cpu().numpy(), X)
    # unaligned destination takes the scalar kernel
    out = torch.full((n, p + 1), 7, dtype=torch.int8, device="cuda")
    ops.unpack_2bit(dev_packed, p, out=out[:, 1:])
    assert np.array_equal(out[:, 1:].cpu().numpy(), X) and bool((out[:, 0] == 7).all())
    with pytest.raises(ValueError):
        ops.pack_2bit(torch.full((2, 8), 5, dtype=torch.int8, device="cuda"))
    with pytest.raises(ValueError):
        ops.pack_2bit_host(np.full((2, 8), -1, dtype=np.int8))


@pytest.mark.parametrize("n,p", [(3, 1), (5, 79), (4, 80), (7, 81), (33, 5003), (64, 50000), (300, 1279)])
def test_base3_host_format_recodes_to_the_2bit_form(n, p):
    """Five 0/1/2 codes per byte across PCIe, re-coded on the device into the 2-bit form the projection kernels read: equal,
    byte for byte (padding included), to packing the int8 codes directly."""
    X = torch.from_numpy(synth.genotypes(n, p, seed=n * 1000 + p)).cuda()
    P3 = ops.pack_base3(X)
    assert P3.shape == (n, ops.base3_stride(p)) and int(P3.max()) <= 242
    # the host side of the format, restated in numpy
    Xh = X.cpu().numpy().astype(np.int64)
    q = (p + 4) // 5
    padded = np.zeros((n, q * 5), dtype=np.int64)
    padded[:, :p] = Xh
    want = (padded.reshape(n, q, 5) * np.array([1, 3, 9, 27, 81])).sum(axis=2)
    assert np.array_equal(P3[:, :q].cpu().numpy(), want)
    got = ops.unpack_base3(P3, p)
    assert torch.equal(got, ops.pack_2bit(X))
    assert torch.equal(ops.unpack_2bit(got, p), X)
    with pytest.raises(ValueError):
        ops.pack_base3(torch.full((2, 7), 3, dtype=torch.int8, device="cuda"))
