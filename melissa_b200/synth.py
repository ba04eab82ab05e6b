"""Synthetic single-cell methylomes of a named shape.

Follows the reference's generator (inst/data-raw/create_synth_data.R:81-156):
cluster labels ~ Multinomial(pi) with uniform pi; per region, with probability
`cluster_var` every cluster gets its own profile w ~ N(0, 1.2^2), otherwise one
shared profile plus N(0, 0.05^2) per cluster (:100-110); CpG loci are integers
in -1000..1000 min-max scaled to [-1, 1] (:145-149); y ~ Bernoulli(clip(Phi(h.w)
+ N(0, 0.05^2), 1e-10, 1 - 1e-10)) (:150-154).  Coverage replaces the
generator's Binomial(25, .8): every region has `sites_per_region` candidate
loci (fixed across cells, as CpG sites are on a genome) and cell n covers each
with its own rate c_n ~ U(cov_lo, cov_hi), so I_nm ~ Binomial(sites, c_n)
(mean 25 at the defaults) and a pair with I_nm = 0 is uncovered.

Everything global (labels, profiles, loci) depends on `seed` only; everything
per cell on (seed, global cell index), so any cell sharding generates the same
data set.
"""
from __future__ import annotations

import math

import numpy as np
from scipy.special import ndtr

from .data import MelissaData, create_rbf_object


def _design(x, basis):
    mus = np.asarray(basis["mus"], np.float64)
    H = np.ones(x.shape + (mus.shape[0] + 1,))
    if mus.shape[0]:
        d = x[..., None] - mus
        H[..., 1:] = np.exp(-basis["gamma"] * d * d)
    return H


def generate(N, M, K, n_basis=4, seed=1, sites_per_region=200, cov_lo=0.05, cov_hi=0.20,
             cluster_var=0.5, cell_lo=0, cell_hi=None, N_total=None, pinned=False):
    """Generate cells [cell_lo, cell_hi) of an N_total x M data set.

    Returns (MelissaData, truth) where truth has W [M, D, K], labels [N_total]
    (0-based), C_true one-hot of the generated cells, and w0 -- the
    deterministic non-degenerate initialisation 0.5 * W + N(0, 0.05^2).
    `pinned=True` allocates x / y in page-locked host memory (needs torch).
    """
    N_total = N if N_total is None else N_total
    cell_hi = N_total if cell_hi is None else cell_hi
    basis = create_rbf_object(n_basis)
    D = n_basis + 1
    g = np.random.Generator(np.random.Philox(key=seed))
    labels = g.integers(0, K, size=N_total)
    W = np.empty((M, D, K))
    variable = g.random(M) < cluster_var
    shared = g.normal(0.0, 1.2, size=(M, D))
    own = g.normal(0.0, 1.2, size=(M, D, K))
    jitter = g.normal(0.0, 0.05, size=(M, D, K))
    W[variable] = own[variable]
    W[~variable] = shared[~variable][:, :, None] + jitter[~variable]
    w0 = 0.5 * W + g.normal(0.0, 0.05, size=W.shape)
    # candidate loci per region: sorted sample without replacement from -1000..1000
    loci = np.sort(np.argsort(g.random((M, 2001)), axis=1)[:, :sites_per_region], axis=1) - 1000
    x_sites = (loci.astype(np.float64) + 1000.0) / 2000.0 * 2.0 - 1.0
    H_sites = _design(x_sites, basis)                                  # [M, S, D]
    p_sites = ndtr(np.einsum("msd,mdk->msk", H_sites, W))              # [M, S, K]

    n_loc = cell_hi - cell_lo
    counts = np.zeros((n_loc, M), dtype=np.int32)
    xs, ys = [], []
    for j, n in enumerate(range(cell_lo, cell_hi)):
        gc = np.random.Generator(np.random.Philox(key=seed + 1, counter=[n, 0, 0, 0]))
        c_n = gc.uniform(cov_lo, cov_hi)
        mask = gc.random((M, sites_per_region), dtype=np.float32) < c_n
        counts[j] = mask.sum(axis=1)
        p = p_sites[:, :, labels[n]][mask] + gc.normal(0.0, 0.05, size=int(counts[j].sum()))
        np.clip(p, 1e-10, 1 - 1e-10, out=p)
        ys.append((gc.random(p.shape[0]) < p).astype(np.uint8))
        xs.append(x_sites[mask])
    covered = counts > 0
    if not covered.any(axis=1).all():
        raise ValueError("a generated cell covers no region; raise the coverage")
    pair_cell, pair_region = np.nonzero(covered)
    pair_ptr = np.concatenate([[0], np.cumsum(counts[covered], dtype=np.int64)])
    nnz = int(pair_ptr[-1])
    if pinned:
        import torch
        xt = torch.empty(nnz, dtype=torch.float64).pin_memory()
        yt = torch.empty(nnz, dtype=torch.uint8).pin_memory()
        x, y = xt.numpy(), yt.numpy()
        np.concatenate(xs, out=x)
        np.concatenate(ys, out=y)
        keep = (xt, yt)
    else:
        x, y, keep = np.concatenate(xs), np.concatenate(ys), None
    data = MelissaData(N=n_loc, M=M, pair_ptr=pair_ptr.astype(np.int64), pair_cell=pair_cell.astype(np.int32),
                       pair_region=pair_region.astype(np.int32), x=x, y=y)
    data.extras["pinned_owner"] = keep
    C_true = np.zeros((n_loc, K))
    C_true[np.arange(n_loc), labels[cell_lo:cell_hi]] = 1.0
    truth = dict(W=W, w0=w0, labels=labels, C_true=C_true, basis=basis, x_sites=x_sites)
    return data, truth


def generate_device(N, M, K, n_basis=4, seed=1, sites_per_region=200, cov_lo=0.05, cov_hi=0.20,
                    cluster_var=0.5, cell_lo=0, cell_hi=None, N_total=None, device=0, chunk_cells=64):
    """Same recipe as `generate`, drawn with torch on the GPU (needed at the
    20,000 x 5,000 and larger shapes, where NumPy would take minutes).  Cells are
    generated in fixed global chunks of `chunk_cells`, each with its own seeded
    generator, so shards whose boundaries are chunk multiples see identical data.
    Returns (dict of CUDA tensors ready for Engine / melissa_pack_device, truth).
    """
    import torch
    N_total = N if N_total is None else N_total
    cell_hi = N_total if cell_hi is None else cell_hi
    if cell_lo % chunk_cells:
        raise ValueError("cell_lo must be a multiple of chunk_cells")
    dev = torch.device(f"cuda:{device}")
    basis = create_rbf_object(n_basis)
    D = n_basis + 1
    g = np.random.Generator(np.random.Philox(key=seed))
    labels = g.integers(0, K, size=N_total)
    W = np.empty((M, D, K))
    variable = g.random(M) < cluster_var
    shared = g.normal(0.0, 1.2, size=(M, D))
    own = g.normal(0.0, 1.2, size=(M, D, K))
    jitter = g.normal(0.0, 0.05, size=(M, D, K))
    W[variable] = own[variable]
    W[~variable] = shared[~variable][:, :, None] + jitter[~variable]
    w0 = 0.5 * W + g.normal(0.0, 0.05, size=W.shape)
    loci = np.sort(np.argsort(g.random((M, 2001)), axis=1)[:, :sites_per_region], axis=1) - 1000
    x_sites = (loci.astype(np.float64) + 1000.0) / 2000.0 * 2.0 - 1.0
    p_sites = ndtr(np.einsum("msd,mdk->msk", _design(x_sites, basis), W))      # [M, S, K]
    x_sites_t = torch.from_numpy(x_sites).to(dev)
    p_sites_t = torch.from_numpy(np.ascontiguousarray(p_sites.transpose(2, 0, 1))).to(dev)   # [K, M, S]
    labels_t = torch.from_numpy(labels).to(dev)
    xs, ys, cnts = [], [], []
    for c0 in range(cell_lo, cell_hi, chunk_cells):
        c1 = min(c0 + chunk_cells, cell_hi)
        gen = torch.Generator(device=dev)
        gen.manual_seed(int(seed) * 1000003 + c0)
        nc = c1 - c0
        c_n = cov_lo + (cov_hi - cov_lo) * torch.rand(chunk_cells, device=dev, generator=gen)[:nc]
        mask = torch.rand((chunk_cells, M, sites_per_region), device=dev, generator=gen)[:nc] < c_n[:, None, None]
        cnts.append(mask.sum(dim=2, dtype=torch.int32))
        idx = mask.nonzero(as_tuple=True)                                       # sorted (cell, region, site)
        p = p_sites_t[labels_t[c0:c1][idx[0]], idx[1], idx[2]]
        p = p + 0.05 * torch.randn(p.shape, device=dev, dtype=torch.float64, generator=gen)
        p.clamp_(1e-10, 1 - 1e-10)
        ys.append((torch.rand(p.shape, device=dev, dtype=torch.float64, generator=gen) < p).to(torch.uint8))
        xs.append(x_sites_t[idx[1], idx[2]])
        del mask, idx, p
    counts = torch.cat(cnts)                                                    # [n_loc, M]
    covered = counts > 0
    if not bool(covered.any(dim=1).all()):
        raise ValueError("a generated cell covers no region; raise the coverage")
    pc, pr = covered.nonzero(as_tuple=True)
    lens = counts[covered].to(torch.int64)
    pair_ptr = torch.zeros(lens.shape[0] + 1, dtype=torch.int64, device=dev)
    torch.cumsum(lens, 0, out=pair_ptr[1:])
    data = dict(on_device=True, N=cell_hi - cell_lo, M=M, pair_ptr=pair_ptr, pair_cell=pc.to(torch.int32).contiguous(),
                pair_region=pr.to(torch.int32).contiguous(), x=torch.cat(xs), y=torch.cat(ys))
    n_loc = cell_hi - cell_lo
    C_true = np.zeros((n_loc, K))
    C_true[np.arange(n_loc), labels[cell_lo:cell_hi]] = 1.0
    truth = dict(W=W, w0=w0, labels=labels, C_true=C_true, basis=basis)
    return data, truth
