"""K1 on the B200 through the C ABI: canonical kNN lists must be bit-exact against the oracle."""
import numpy as np
import pytest
import torch

from gnn_b200 import ops
from oracle import restate, synth

pytestmark = pytest.mark.gpu
DEV = "cuda"


def dev_i8(X, pad=True):
    X = np.ascontiguousarray(X, dtype=np.int8)
    n, p = X.shape
    ld = (p + 15) // 16 * 16 if pad else p
    buf = torch.zeros((n, ld), dtype=torch.int8, device=DEV)
    buf[:, :p] = torch.from_numpy(X).to(DEV)
    return buf[:, :p]


def run_knn(X, k, extra=None, impl=None, pad=True):
    idx, d2 = ops.knn(dev_i8(X, pad), k, extra=None if extra is None else torch.from_numpy(extra), impl=impl)
    torch.cuda.synchronize()
    return idx.cpu().numpy().astype(np.int64), d2.cpu().numpy()


def check(X, k, extra=None, impl=None, pad=True):
    idx, d2 = run_knn(X, k, extra, impl, pad)
    ridx, rd2 = restate.canonical_knn(X, k, extra=extra)
    bad = np.nonzero((idx != ridx).any(axis=1) | (d2 != rd2).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} rows differ, first {bad[:5]}: got {idx[bad[0]]} {d2[bad[0]]} want {ridx[bad[0]]} {rd2[bad[0]]}"


@pytest.mark.parametrize("impl", [1, 2, 0])
def test_wheat_golden(wheat, golden, impl):
    X, y = wheat
    g = golden("wheat_knn.npz")
    idx, d2 = run_knn(X, 32, impl=impl)
    assert np.array_equal(idx, g["idx_all"]) and np.array_equal(d2, g["d2_all"].astype(np.float64))
    idx, d2 = run_knn(X[:479], 32, impl=impl)
    assert np.array_equal(idx, g["idx_train"]) and np.array_equal(d2, g["d2_train"].astype(np.float64))
    for k in (1, 5, 10):                                   # shorter lists are prefixes of the k = 32 list
        a, b = run_knn(X, k, impl=impl)
        assert np.array_equal(a, g["idx_all"][:, :k]) and np.array_equal(b, g["d2_all"][:, :k].astype(np.float64))


@pytest.mark.parametrize("impl", [1, 2, 0])
def test_wheat_with_phenotype_column(wheat, golden, impl):
    """The partition graph keeps the float phenotype column (util.py:52-53): float64 epilogue, bit-exact."""
    X, y = wheat
    g = golden("wheat_knn.npz")
    idx, d2 = run_knn(X, 10, extra=y[:, None], impl=impl)
    assert np.array_equal(idx, g["idx_with_pheno"])
    assert np.array_equal(d2, g["d2_with_pheno"])


@pytest.mark.parametrize("n,p,k", [(130, 64, 1), (257, 100, 5), (1000, 777, 32), (64, 13, 63), (65, 1279, 64), (2, 5, 1),
                                   (3000, 256, 10)])
@pytest.mark.parametrize("impl", [1, 2, 0])
def test_synthetic_shapes(n, p, k, impl):
    check(synth.genotypes(n, p, seed=11 + n), k, impl=impl)


@pytest.mark.parametrize("n,p,k", [(128, 128, 3), (129, 257, 32), (1000, 4096, 33), (2500, 130, 64), (5000, 1279, 16),
                                   (640, 16384 + 96, 8), (1300, 20000, 32), (300, 16384, 64)])
def test_tensor_core_path_equals_cuda_core_path(n, p, k):
    """impl=2 (tcgen05 int8 tiles) and impl=1 (dp4a tiles) must produce the same bits.  p >= 16384 runs the clusters of two
    CTA pairs with multicast key tiles (odd pair counts leave the last cluster with an empty second pair)."""
    X = dev_i8(synth.genotypes(n, p, seed=n + p))
    a_i, a_d = ops.knn(X, k, impl=1)
    b_i, b_d = ops.knn(X, k, impl=2)
    torch.cuda.synchronize()
    assert torch.equal(a_i, b_i) and torch.equal(a_d, b_d)


def test_unpadded_rows_and_general_int8_codes():
    rng = np.random.default_rng(5)
    X = rng.integers(0, 128, size=(300, 203)).astype(np.int8)
    check(X, 7, impl=1, pad=False)                          # ld = 203: byte-wise loads
    check(X, 7, impl=0, pad=True)


def test_duplicates_and_heavy_ties():
    X = synth.genotypes(200, 40, seed=3)
    X[10] = X[4]; X[11] = X[4]; X[150:160] = X[0]
    check(X, 12, impl=1)
    check(X, 12, impl=2)
    check(np.zeros((90, 33), dtype=np.int8), 20, impl=0)
    check(np.zeros((300, 64), dtype=np.int8), 20, impl=2)    # every distance equal: order is by index alone


def test_two_float_columns():
    X = synth.genotypes(400, 90, seed=8)
    extra = np.random.default_rng(1).standard_normal((400, 2))
    check(X, 9, extra=extra, impl=1)
    check(X, 9, extra=extra, impl=2)


def test_streaming_key_blocks_equal_one_pass():
    """merge=1 folds key blocks into the running lists: the result must not depend on the blocking."""
    X = synth.genotypes(700, 300, seed=21)
    Xd = dev_i8(X)
    sq = ops.row_sqnorm(Xd)
    k = 16
    one_i, one_d = ops.knn(Xd, k)
    best_d = torch.empty((700, k), dtype=torch.float64, device=DEV)
    best_i = torch.empty((700, k), dtype=torch.int32, device=DEV)
    bounds = [0, 130, 131, 500, 700]
    order = [2, 0, 3, 1]                                     # any order of key blocks
    for t, b in enumerate(order):
        lo, hi = bounds[b], bounds[b + 1]
        ops.knn_block(Xd, sq, 0, Xd[lo:hi], sq[lo:hi], lo, k, best_d, best_i, merge=int(t > 0))
    torch.cuda.synchronize()
    assert torch.equal(best_i, one_i) and torch.equal(best_d, one_d)
    # query blocks with an offset (what a rank owning rows [q0, q1) does)
    q0, q1 = 200, 333
    bd = torch.empty((q1 - q0, k), dtype=torch.float64, device=DEV)
    bi = torch.empty((q1 - q0, k), dtype=torch.int32, device=DEV)
    ops.knn_block(Xd[q0:q1], sq[q0:q1], q0, Xd, sq, 0, k, bd, bi, merge=0)
    assert torch.equal(bi, one_i[q0:q1]) and torch.equal(bd, one_d[q0:q1])


def test_unfilled_slots_and_errors():
    X = dev_i8(synth.genotypes(5, 20))
    sq = ops.row_sqnorm(X)
    bd = torch.empty((5, 4), dtype=torch.float64, device=DEV)
    bi = torch.empty((5, 4), dtype=torch.int32, device=DEV)
    ops.knn_block(X, sq, 0, X[:2], sq[:2], 0, 4, bd, bi, merge=0)     # only 2 keys: rows 0/1 see one, others two
    torch.cuda.synchronize()
    assert torch.isinf(bd[:, 2:]).all() and (bi[:, 2:] == -1).all()
    assert torch.isinf(bd[0, 1]) and bi[0, 0] == 1
    with pytest.raises(ValueError):
        ops.knn(X, 5)
    from gnn_b200._lib import GnnbError
    with pytest.raises(GnnbError, match="1..64"):
        ops.knn_block(X, sq, 0, X, sq, 0, 65, bd, bi, merge=0)


def test_large_property_checks():
    """20k x 2k: sortedness, self exclusion, and a random sample of rows against brute force."""
    n, p, k = 20000, 2048, 32
    X = ops.synth_genotypes(n, p, DEV)
    idx, d2 = ops.knn(X, k)
    torch.cuda.synchronize()
    assert (d2[:, 1:] >= d2[:, :-1]).all()
    ties = d2[:, 1:] == d2[:, :-1]
    assert (idx[:, 1:][ties] > idx[:, :-1][ties]).all()
    assert not (idx == torch.arange(n, device=DEV, dtype=torch.int32)[:, None]).any()
    Xh = X.cpu().numpy().astype(np.float64)
    rows = np.random.default_rng(0).choice(n, 48, replace=False)
    d = ((Xh[rows] ** 2).sum(1)[:, None] + (Xh ** 2).sum(1)[None, :] - 2.0 * Xh[rows] @ Xh.T)
    d[np.arange(48), rows] = np.inf
    order = np.argsort(d, axis=1, kind="stable")[:, :k]
    assert np.array_equal(order, idx.cpu().numpy()[rows])
    assert np.array_equal(np.take_along_axis(d, order, 1), d2.cpu().numpy()[rows])


def test_row_sharded_entry_point_single_rank():
    """parallel.knn_sharded with one rank equals ops.knn; the 2-rank NCCL run is scripts/check_knn_sharded.py
    (bit-identical on 2 B200s, recorded in profiles/README.md)."""
    from gnn_b200 import parallel
    X = dev_i8(synth.genotypes(900, 300, seed=4))
    a_i, a_d = parallel.knn_sharded(X, 0, 900, 12, 0, 1)
    b_i, b_d = ops.knn(X, 12)
    assert torch.equal(a_i, b_i) and torch.equal(a_d, b_d)


@pytest.mark.parametrize("n,p,k", [(300, 600, 5), (1000, 4096, 10), (3534, 5003, 10), (2000, 20001, 32), (129, 512, 3)])
def test_fp4_tensor_path_equals_int8_tensor_path(n, p, k):
    """Genotype codes 0 / 1 / 2 are exact e2m1 values and their dot products exact in f32, so the search on the FP4 tensor
    path (tcgen05 kind::mxf4 with constant block scales, key tiles of 224) returns the int8 path's lists bit for bit --
    with a float64 side column too, folding key blocks into running lists too."""
    X = ops.synth_genotypes(n, p, "cuda", row0=3)
    a_i, a_d = ops.knn(X, k, impl=2)
    b_i, b_d = ops.knn(X, k, impl=3)
    assert torch.equal(a_i, b_i) and torch.equal(a_d, b_d)
    extra = torch.randn(n, 1, device="cuda", dtype=torch.float64)
    a_i, a_d = ops.knn(X, k, extra=extra, impl=2)
    b_i, b_d = ops.knn(X, k, extra=extra, impl=3)
    assert torch.equal(a_i, b_i) and torch.equal(a_d, b_d)
    # streaming: two key blocks folded one after the other
    X4, sq = ops.pack_fp4(X), ops.row_sqnorm(X)
    assert X4 is not None and X4.shape == (n, ops.fp4_stride(p))
    d2 = torch.empty((n, k), dtype=torch.float64, device="cuda")
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda")
    cut = (n // 2 + 15) // 16 * 16
    ops.knn_block_fp4(X4, sq, 0, X4[cut:], sq[cut:], cut, p, k, d2, idx, 0)
    ops.knn_block_fp4(X4, sq, 0, X4[:cut], sq[:cut], 0, p, k, d2, idx, 1)
    want_i, want_d = ops.knn(X, k, impl=2)
    assert torch.equal(idx, want_i) and torch.equal(d2, want_d)


def test_fp4_tensor_path_is_refused_for_other_codes():
    X = torch.randint(0, 4, (300, 600), dtype=torch.int8, device="cuda")            # a code 3 somewhere
    X[5, 7] = 3
    assert ops.pack_fp4(X) is None
    with pytest.raises(Exception, match="codes in 0..2"):
        ops.knn(X, 4, impl=3)
    a = ops.knn(X, 4)                                                                # "pick" falls back to the int8 tiles
    b = ops.knn(X, 4, impl=1)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
