"""Stateless synthetic genotypes / phenotypes (numpy) -- TEST INFRASTRUCTURE.

Follows SURVEY.md section 8(d): any block of the genotype matrix can be
regenerated identically on host and device because element (i, j) depends only
on (seed, i, j).  All arithmetic is integer so numpy and CUDA agree bit for bit:

    h_ij  = splitmix64(seed ^ (i * p + j))
    t_j   = T0 + ((splitmix64(seed2 ^ j) >> 32) * T1 >> 32)   # allele freq in [0.05, 0.5) as a 32-bit threshold
    x_ij  = [lo32(h_ij) < t_j] + [hi32(h_ij) < t_j]          # Binomial(2, f_j) in {0, 1, 2}

The device twin of this generator is gnn_b200/csrc/synth.cu; tests compare them.
The shapes stand in for the datasets the reference names but does not ship
(.MISSING_LARGE_BLOBS:1-3).
"""
import numpy as np

SEED_DATA = 20240601
SEED_FREQ = 20240602
SEED_PHENO = 20240603
T0 = 214748365      # floor(0.05 * 2**32)
T1 = 1932735283     # floor(0.45 * 2**32)
_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(z):
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def thresholds(p, cols=None, seed2=SEED_FREQ):
    j = np.arange(p, dtype=np.uint64) if cols is None else np.asarray(cols, dtype=np.uint64)
    h = splitmix64(np.uint64(seed2) ^ j)
    return (np.uint64(T0) + (((h >> np.uint64(32)) * np.uint64(T1)) >> np.uint64(32))).astype(np.uint64)


def genotypes(n, p, row0=0, cols=None, seed=SEED_DATA, seed2=SEED_FREQ, p_total=None):
    """int8 [n, len(cols) or p] block of the virtual [N, p_total] genotype matrix."""
    p_total = p if p_total is None else p_total
    j = np.arange(p, dtype=np.uint64) if cols is None else np.asarray(cols, dtype=np.uint64)
    t = thresholds(p_total, cols=j, seed2=seed2)
    out = np.empty((n, j.size), dtype=np.int8)
    step = max(1, (1 << 22) // max(1, j.size))
    with np.errstate(over="ignore"):
        for r in range(0, n, step):
            i = np.arange(row0 + r, row0 + min(n, r + step), dtype=np.uint64)[:, None]
            h = splitmix64(np.uint64(seed) ^ (i * np.uint64(p_total) + j[None, :]))
            lo = h & np.uint64(0xFFFFFFFF)
            hi = h >> np.uint64(32)
            out[r:r + i.shape[0]] = ((lo < t[None, :]).astype(np.int8) + (hi < t[None, :]).astype(np.int8))
    return out


def phenotypes(n, p, traits=1, n_causal=100, h2=0.5, seed=SEED_DATA, seed2=SEED_FREQ, seed3=SEED_PHENO):
    """float64 [n, traits], centred: y = X beta + eps on a shared causal set, noise scaled for h^2."""
    rng = np.random.default_rng(seed3)
    causal = np.sort(rng.choice(p, size=min(n_causal, p), replace=False))
    xc = genotypes(n, p, cols=causal, seed=seed, seed2=seed2, p_total=p).astype(np.float64)
    beta = rng.standard_normal((causal.size, traits))
    g = xc @ beta
    var_g = g.var(axis=0)
    eps = rng.standard_normal((n, traits)) * np.sqrt(var_g * (1.0 - h2) / h2)
    y = g + eps
    return y - y.mean(axis=0)
