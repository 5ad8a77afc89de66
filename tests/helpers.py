"""Shared helpers for the GPU parity tests (oracle = checker only)."""
import numpy as np
import torch

from oracle import knn_smallest


def reps(seed, b, d, kind="tanh"):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(b, d, generator=g)
    if kind == "tanh":
        return torch.tanh(x)
    if kind == "unit":
        return torch.nn.functional.normalize(x, dim=1, p=2)
    return x


def assert_knn_matches(src, tgt, k, gpu_dist, gpu_idx, rtol=1e-5, tie_rtol=4e-6, oracle=None):
    """Index SETS must equal the oracle's except where the oracle's own distances of the
    disputed neighbours tie with its k-th distance to within `tie_rtol` (a different but equally
    valid fp32 summation order may swap them); distances within `rtol`."""
    od, oi = oracle if oracle is not None else knn_smallest(src, tgt, k)
    gd = gpu_dist.cpu()
    gi = gpu_idx.cpu().long()
    assert gd.shape == od.shape and gi.shape == oi.shape
    assert torch.all(gi >= 0) and torch.all(gi < tgt.size(0))
    np.testing.assert_allclose(gd.numpy(), od.numpy(), rtol=rtol, atol=1e-6)
    assert torch.all(gd[:, 1:] >= gd[:, :-1]), "distances must be ascending"
    so, sg = oi.sort(dim=1).values, gi.sort(dim=1).values
    bad_rows = (so != sg).any(dim=1).nonzero().flatten().tolist()
    n_tie_rows = 0
    for r in bad_rows:
        a, b = set(oi[r].tolist()), set(gi[r].tolist())
        kth = float(od[r, -1])
        for j in a ^ b:
            dj = float(torch.norm(src[r].reshape(-1) - tgt[j].reshape(-1), p=2))
            assert abs(dj - kth) <= tie_rtol * max(kth, 1e-30), (
                f"row {r}: neighbour {j} at {dj} is not a near-tie of the k-th distance {kth}")
        n_tie_rows += 1
    # within a row, indices must be distinct
    assert all(len(set(row)) == k for row in gi.tolist())
    return n_tie_rows
