"""Seeded synthetic clouds for the parity tests and bench.py.

Everything is generated on the CPU with torch.Generator so the CPU oracle and the GPU
path see identical bits.  Seeds: cloud 1 = 1000*c, cloud 2 = 1000*c+1, grads = 1000*c+2
where c is the BASELINE.json config number (or any test-chosen integer).
"""
from __future__ import annotations

import torch

KINDS = ("uniform", "clustered", "lattice", "duplicates", "identical", "crosscopy", "large", "tiny")


def _gen(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def uniform(b: int, n: int, seed: int) -> torch.Tensor:
    """xyz uniform in [0,1)^3."""
    return torch.rand(b, n, 3, generator=_gen(seed), dtype=torch.float32)


def clustered(b: int, n: int, seed: int, centres: int = 64, sigma: float = 0.01) -> torch.Tensor:
    """64 centres uniform in [0,1)^3, isotropic normal sigma=0.01, equal cluster sizes, fixed permutation."""
    g = _gen(seed)
    c = torch.rand(b, centres, 3, generator=g, dtype=torch.float32)
    which = torch.arange(n) % centres
    pts = c[:, which, :] + sigma * torch.randn(b, n, 3, generator=g, dtype=torch.float32)
    perm = torch.randperm(n, generator=g)
    return pts[:, perm, :].contiguous()


def lattice(b: int, n: int, seed: int, cells: int = 64) -> torch.Tensor:
    """Points snapped to a cells^3 lattice: thousands of exact equal-distance ties and duplicates."""
    return (torch.floor(uniform(b, n, seed) * cells) / cells).contiguous()


def duplicates(b: int, n: int, seed: int) -> torch.Tensor:
    """Second half of the cloud is a copy of the first half."""
    x = uniform(b, n, seed)
    h = n // 2
    if h > 0:
        x[:, h:2 * h, :] = x[:, :h, :]
    return x


def make_pair(kind: str, b: int, n: int, m: int, config: int = 0):
    """(xyz1[b,n,3], xyz2[b,m,3]) float32 CPU tensors for one of KINDS."""
    s1, s2 = 1000 * config, 1000 * config + 1
    if kind == "uniform":
        return uniform(b, n, s1), uniform(b, m, s2)
    if kind == "clustered":
        return clustered(b, n, s1), clustered(b, m, s2)
    if kind == "lattice":
        return lattice(b, n, s1), lattice(b, m, s2)
    if kind == "duplicates":
        return duplicates(b, n, s1), duplicates(b, m, s2)
    if kind == "identical":  # xyz2 = xyz1 (needs n == m): all dist 0, idx = lowest duplicate index
        x = duplicates(b, n, s1)
        return x, x.clone()
    if kind == "crosscopy":  # 25 % of xyz2 rows are copies of random xyz1 rows
        x1, x2 = uniform(b, n, s1), uniform(b, m, s2)
        g = _gen(s2 + 7)
        k = max(1, m // 4)
        rows = torch.randperm(m, generator=g)[:k]
        src = torch.randint(0, n, (k,), generator=g)
        x2[:, rows, :] = x1[:, src, :]
        return x1, x2
    if kind == "large":  # coordinates ~1e3..1e4: stresses rounding of the squares and sums
        return uniform(b, n, s1) * 9000.0 + 1000.0, uniform(b, m, s2) * 9000.0 + 1000.0
    if kind == "tiny":  # squares underflow into the denormal range
        return uniform(b, n, s1) * 1e-20, uniform(b, m, s2) * 1e-20
    raise ValueError(f"unknown cloud kind {kind!r}; choose from {KINDS}")


def grad_pair(b: int, n: int, m: int, config: int = 0, mean: bool = False):
    """graddist1[b,n], graddist2[b,m]: random for parity, 1/(B*N) for timing (mean chamfer)."""
    if mean:
        return (torch.full((b, n), 1.0 / max(1, b * n), dtype=torch.float32),
                torch.full((b, m), 1.0 / max(1, b * m), dtype=torch.float32))
    g = _gen(1000 * config + 2)
    return (torch.rand(b, n, generator=g, dtype=torch.float32), torch.rand(b, m, generator=g, dtype=torch.float32))
