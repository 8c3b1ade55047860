"""Deterministic synthetic graphs / features of the BASELINE.json shapes.

Bench and test INPUT generation only -- not part of the propagation path.  Everything is a
counter-based hash (splitmix64 of ``(seed, stream, index)``) evaluated with torch int64 tensor
ops, so the same call gives the same graph on CPU and on a GPU, on any rank, with no file I/O
(the Reddit / ogbn files the reference expects are not shipped: /root/reference/README.md:68).

Graph model ``G(N, E, alpha, cap, lam, seed)``: Chung-Lu with Pareto(alpha) node weights capped at
``cap``; every node draws ``~w_i * E / sum(w)`` out-neighbours from the weight distribution (a
fraction ``lam`` of them forced into the node's own 1/``blocks`` block, for partition locality);
self-loops removed, duplicates merged; optionally symmetrised to a binary undirected graph.
Generation is per source row, so a rank can build just its own row slice (``row_range``).
"""
import math

import torch

_M64 = (1 << 64)


def _s64(c):
    """Python int (unsigned 64-bit constant) -> the signed value torch.int64 holds."""
    c &= _M64 - 1
    return c - _M64 if c >= (1 << 63) else c


_GOLDEN = _s64(0x9E3779B97F4A7C15)
_C1 = _s64(0xBF58476D1CE4E5B9)
_C2 = _s64(0x94D049BB133111EB)


def _lsr(z, s):
    return (z >> s) & ((1 << (64 - s)) - 1)


def mix64(x):
    """splitmix64 output function on an int64 tensor (wrap-around arithmetic)."""
    z = x + _GOLDEN
    z = (z ^ _lsr(z, 30)) * _C1
    z = (z ^ _lsr(z, 27)) * _C2
    return z ^ _lsr(z, 31)


def _key(seed, stream):
    # scalar pre-mix of (seed, stream) done in Python ints
    z = (seed * 0x9E3779B97F4A7C15 + stream * 0xD1B54A32D192ED03 + 0x632BE59BD9B4E019) & (_M64 - 1)
    z ^= z >> 30
    z = (z * 0xBF58476D1CE4E5B9) & (_M64 - 1)
    z ^= z >> 27
    z = (z * 0x94D049BB133111EB) & (_M64 - 1)
    z ^= z >> 31
    return _s64(z)


def hash_uniform(seed, stream, idx):
    """U[0,1) fp64 for every int64 index in ``idx`` (53 random bits)."""
    z = mix64(idx ^ _key(seed, stream))
    return _lsr(z, 11).to(torch.float64) * (1.0 / (1 << 53))


def hash_normal(seed, stream, idx):
    """N(0,1) fp64 (Box-Muller on two hashed uniforms)."""
    u1 = hash_uniform(seed, 2 * stream, idx)
    u2 = hash_uniform(seed, 2 * stream + 1, idx)
    return torch.sqrt(-2.0 * torch.log1p(-u1)) * torch.cos((2.0 * math.pi) * u2)


def node_weights(n, alpha, cap, seed, device, draws_per_unit_weight=None):
    """Pareto(alpha) weights.  ``cap`` bounds the EXPECTED DEGREE of a node: weights are clipped
    at the value ``wc`` for which ``wc * draws / sum(min(w, wc)) == cap`` (4 fixed-point steps),
    where ``draws`` is the number of stored entries a unit of weight attracts."""
    idx = torch.arange(n, dtype=torch.int64, device=device)
    u = hash_uniform(seed, 0, idx)
    w = torch.pow(1.0 - u, -1.0 / alpha)
    if draws_per_unit_weight is None:
        return torch.clamp(w, max=float(cap))
    wc = float(cap)
    for _ in range(4):
        total = torch.clamp(w, max=wc).sum().item()
        wc = float(cap) * total / float(draws_per_unit_weight)
    return torch.clamp(w, max=wc)


def powerlaw_graph(n, n_edges, alpha, cap, seed=0, lam=0.0, blocks=8, symmetric=True,
                   row_range=None, device="cpu", value=1.0):
    """Returns canonical CSR ``(row_ptr int64 [rows+1], col int32 [nnz], val fp64 [nnz])``.

    ``n_edges`` is the number of sampled (src,dst) pairs over the whole graph (before duplicate
    and self-loop removal; each pair yields two stored entries when ``symmetric``).
    ``row_range=(r0, r1)`` builds only rows [r0, r1) -- supported for ``symmetric=False`` (a
    directed graph is generated row by row, so shards are independent).
    """
    dev = torch.device(device)
    w = node_weights(n, alpha, cap, seed, dev,
                     draws_per_unit_weight=(2 if symmetric else 1) * float(n_edges))
    cum = torch.cumsum(w, 0)
    total = cum[-1]
    ids = torch.arange(n, dtype=torch.int64, device=dev)
    # stochastic rounding of the expected out-degree
    expdeg = w * (float(n_edges) / total)
    deg = torch.floor(expdeg + hash_uniform(seed, 1, ids)).to(torch.int64)
    off = torch.cumsum(deg, 0) - deg  # exclusive
    if row_range is not None:
        if symmetric:
            raise ValueError("row_range requires symmetric=False")
        r0, r1 = row_range
    else:
        r0, r1 = 0, n
    deg_s = deg[r0:r1]
    m = int(deg_s.sum().item())
    src = torch.repeat_interleave(ids[r0:r1], deg_s)
    eid = torch.arange(m, dtype=torch.int64, device=dev) + (int(off[r0].item()) if m else 0)
    t = hash_uniform(seed, 2, eid) * total
    if lam > 0.0:
        bsz = (n + blocks - 1) // blocks
        b0 = (src // bsz) * bsz
        b1 = torch.clamp(b0 + bsz, max=n)
        lo = torch.where(b0 > 0, cum[torch.clamp(b0 - 1, min=0)], torch.zeros_like(total))
        hi = cum[b1 - 1]
        local = hash_uniform(seed, 3, eid) < lam
        t = torch.where(local, lo + hash_uniform(seed, 2, eid) * (hi - lo), t)
    dst = torch.clamp(torch.searchsorted(cum, t, right=True), max=n - 1)
    keep = src != dst
    src, dst = src[keep], dst[keep]
    if symmetric:
        src, dst = torch.cat([src, dst]), torch.cat([dst, src])
    key = torch.unique(src * n + dst)  # sorted, duplicates merged
    del src, dst
    rows = key // n
    col = (key - rows * n).to(torch.int32)
    counts = torch.bincount(rows - r0, minlength=r1 - r0)
    row_ptr = torch.zeros(r1 - r0 + 1, dtype=torch.int64, device=dev)
    row_ptr[1:] = torch.cumsum(counts, 0)
    val = torch.full((col.numel(),), float(value), dtype=torch.float64, device=dev)
    return row_ptr, col, val


def gaussian_features(n, f, seed=1, scaled=False, row_range=None, device="cpu", chunk_rows=1 << 16):
    """fp32 [rows, f] features.  ``scaled=False``: N(0,1).  ``scaled=True`` (Reddit-shaped,
    non-standardised so the fused standardisation is exercised): ``x = N(0,1)*s_f + m_f`` with
    ``s_f ~ U(0.1, 5)``, ``m_f ~ U(-3, 3)`` per column."""
    dev = torch.device(device)
    r0, r1 = (0, n) if row_range is None else row_range
    out = torch.empty((r1 - r0, f), dtype=torch.float32, device=dev)
    cols = torch.arange(f, dtype=torch.int64, device=dev)
    if scaled:
        s = 0.1 + 4.9 * hash_uniform(seed, 100, cols)
        m = -3.0 + 6.0 * hash_uniform(seed, 101, cols)
    for a in range(r0, r1, chunk_rows):
        b = min(a + chunk_rows, r1)
        idx = torch.arange(a, b, dtype=torch.int64, device=dev)[:, None] * f + cols[None, :]
        z = hash_normal(seed, 10, idx)
        if scaled:
            z = z * s[None, :] + m[None, :]
        out[a - r0:b - r0] = z.to(torch.float32)
    return out


def tfidf_features(n, f, density=0.1, seed=1, device="cpu"):
    """Pubmed-like features: Bernoulli(density) * U(0,1), rows normalised to sum 1 (fp32)."""
    dev = torch.device(device)
    idx = torch.arange(n * f, dtype=torch.int64, device=dev).view(n, f)
    mask = hash_uniform(seed, 20, idx) < density
    x = torch.where(mask, hash_uniform(seed, 21, idx), torch.zeros((), dtype=torch.float64, device=dev))
    x = x.to(torch.float32)
    rs = x.sum(1, keepdim=True)
    rs[rs == 0] = 1.0
    return x / rs


def planted_labels(n, n_classes, blocks_seed=2, device="cpu"):
    """Class = contiguous block id (planted communities for the Reddit-shaped F1 parity run)."""
    ids = torch.arange(n, dtype=torch.int64, device=device)
    return (ids * n_classes) // n


# Named shapes from BASELINE.json `configs`.
SHAPES = {
    # name: dict(n, pairs, alpha, cap, f, k, symmetric, value)
    "reddit": dict(n=232_965, pairs=12_070_000, alpha=1.3, cap=22_000, f=602, k=2,
                   symmetric=True, value=2.0, scaled=True),
    "products": dict(n=2_449_029, pairs=61_970_000, alpha=1.5, cap=20_000, f=100, k=3,
                     symmetric=True, value=1.0, scaled=False),
    "papers100m": dict(n=111_059_956, pairs=1_640_000_000, alpha=1.5, cap=50_000, f=128, k=3,
                       symmetric=False, value=1.0, scaled=False),
}


def make_shape(name, device="cpu", scale=1.0, lam=0.0, row_range=None, seed=0, with_features=True):
    """Graph (+ features) of a named BASELINE shape, optionally scaled down by ``scale``
    (nodes and pairs scaled together, feature width kept)."""
    s = SHAPES[name]
    n = max(8, int(round(s["n"] * scale)))
    pairs = max(8, int(round(s["pairs"] * scale)))
    cap = max(4, min(s["cap"], n // 4))
    row_ptr, col, val = powerlaw_graph(n, pairs, s["alpha"], cap, seed=seed, lam=lam,
                                       symmetric=s["symmetric"], row_range=row_range,
                                       device=device, value=s["value"])
    x = None
    if with_features:
        x = gaussian_features(n, s["f"], seed=seed + 1, scaled=s["scaled"], row_range=row_range,
                              device=device)
    return dict(n=n, f=s["f"], k=s["k"], row_ptr=row_ptr, col=col, val=val, x=x)


def write_reddit_npz(out_dir, n=24_000, f=602, n_classes=41, pairs_per_node=25, lam=0.85, seed=0,
                     signal=1.5, splits=(0.66, 0.10, 0.24)):
    """Synthetic Reddit in the file layout `utils.loadRedditFromNPZ` reads (/root/reference/utils.py:104-108),
    so that the reference's reddit.py and this repo's drop-in run on identical files:

        reddit_adj.npz   scipy CSR (save_npz), symmetric 0/1 adjacency, no self-loops
        reddit.npz       feats fp32 [N, F] (NOT standardised), y_train / y_val / y_test (int labels),
                         train_index / val_index / test_index (int64 node ids)

    41 planted communities: class = contiguous block id (`planted_labels`), a fraction `lam` of every node's
    neighbours drawn from its own block; features = per-column scale/offset as in `gaussian_features(scaled=True)`
    plus a class centroid of strength `signal` (so that K-hop propagated features are linearly separable but
    not trivially so).  Splits: random permutation with seed+2 (~66 % / 10 % / 24 %)."""
    import os

    import numpy as np
    import scipy.sparse as sp
    row_ptr, col, _ = powerlaw_graph(n, n * pairs_per_node, alpha=1.3, cap=max(8, n // 12), seed=seed, lam=lam,
                                     blocks=n_classes, symmetric=True, device="cpu")
    adj = sp.csr_matrix((np.ones(col.numel(), dtype=np.float64), col.numpy(), row_ptr.numpy()), shape=(n, n))
    labels = planted_labels(n, n_classes).numpy()
    x = gaussian_features(n, f, seed=seed + 1, scaled=True, device="cpu").numpy()
    cols = torch.arange(f, dtype=torch.int64)
    scale = (0.1 + 4.9 * hash_uniform(seed + 1, 100, cols)).numpy()           # the per-column s_f of gaussian_features
    cent = hash_normal(seed + 3, 7, torch.arange(n_classes * f, dtype=torch.int64)).view(n_classes, f).numpy()
    x = (x + signal * cent[labels] * scale[None, :]).astype(np.float32)
    perm = np.random.RandomState(seed + 2).permutation(n)
    n_tr, n_va = int(splits[0] * n), int(splits[1] * n)
    tr, va, te = np.sort(perm[:n_tr]), np.sort(perm[n_tr:n_tr + n_va]), np.sort(perm[n_tr + n_va:])
    os.makedirs(out_dir, exist_ok=True)
    sp.save_npz(os.path.join(out_dir, "reddit_adj.npz"), adj)
    np.savez(os.path.join(out_dir, "reddit.npz"), feats=x, y_train=labels[tr], y_val=labels[va], y_test=labels[te],
             train_index=tr.astype(np.int64), val_index=va.astype(np.int64), test_index=te.astype(np.int64))
    return {"n": n, "f": f, "nnz": int(adj.nnz), "classes": n_classes, "train": int(tr.size), "val": int(va.size),
            "test": int(te.size)}
