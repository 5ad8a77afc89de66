"""Torch-CPU restatement of the reference's particle-based-entropy reward.

Test infrastructure only (see oracle/__init__.py).  Every function cites the reference lines it
follows.  The arithmetic is expressed with the same ATen ops as the reference so that on the same
torch build the results are bitwise equal to the reference's (checked by tests/test_oracle_pins.py
against tests/golden/ produced from the unmodified reference).
"""
from dataclasses import dataclass, field

import torch


@dataclass
class RmsState:
    """Running moments carried by the reference's ``RMS`` object.

    Follows /root/reference/utils/utils.py:259-262 (M = zeros(shape), S = ones(shape),
    n = epsilon as a python float).
    """
    M: torch.Tensor = field(default_factory=lambda: torch.zeros(1))
    S: torch.Tensor = field(default_factory=lambda: torch.ones(1))
    n: float = 1e-4


def rms_step(state: RmsState, x: torch.Tensor):
    """One running-moment update; mutates and returns ``state``.

    Follows /root/reference/utils/utils.py:264-276: batch mean and *unbiased* variance along
    dim 0, Chan-style combination with the carried (M, S, n), then n += bs.
    """
    bs = x.size(0)
    tot = state.n + bs
    delta = torch.mean(x, dim=0) - state.M
    m_new = state.M + delta * bs / tot
    s_new = (state.S * state.n + torch.var(x, dim=0) * bs
             + torch.square(delta) * state.n * bs / tot) / tot
    state.M, state.S, state.n = m_new, s_new, tot
    return state


def knn_smallest(source: torch.Tensor, target: torch.Tensor, k: int, q_tile: int = 128,
                 t_tile: int = 4096):
    """k smallest L2 distances (ascending) and their target indices for every source row.

    Follows /root/reference/utils/utils.py:293-300 — the direct-difference form
    ``norm(source[:, None, :] - target[None, :, :], dim=-1, p=2)`` followed by
    ``topk(k, largest=False, sorted=True)`` — evaluated in (q_tile x t_tile) tiles with a running
    top-k so that memory stays bounded.  SURVEY.md §8c: the tiled distance values are bitwise
    equal to the untiled expression.  The reference discards the indices (utils.py:297); they
    are returned here because parity is defined on index sets.
    Raises RuntimeError when k > number of target rows, like torch.topk in the reference.
    """
    b1, b2 = source.size(0), target.size(0)
    if k > b2:
        raise RuntimeError("selected index k out of range")
    src = source.reshape(b1, -1)
    tgt = target.reshape(b2, -1)
    out_d = torch.empty(b1, k, dtype=src.dtype)
    out_i = torch.empty(b1, k, dtype=torch.int64)
    for q0 in range(0, b1, q_tile):
        s = src[q0:q0 + q_tile]
        best_d = None
        best_i = None
        for t0 in range(0, b2, t_tile):
            t = tgt[t0:t0 + t_tile]
            d = torch.norm(s[:, None, :] - t[None, :, :], dim=-1, p=2)
            idx = torch.arange(t0, t0 + t.size(0)).expand_as(d)
            if best_d is not None:
                d = torch.cat([best_d, d], dim=1)
                idx = torch.cat([best_i, idx], dim=1)
            kk = min(k, d.size(1))
            best_d, pos = d.topk(kk, dim=1, largest=False, sorted=True)
            best_i = torch.gather(idx, 1, pos)
        out_d[q0:q0 + q_tile] = best_d
        out_i[q0:q0 + q_tile] = best_i
    return out_d, out_i


def pbe_reward(source, target, state: RmsState, knn_k, knn_avg, knn_rms, knn_clip,
               knn=None):
    """APT particle-based reward for every source row, shape (b1, 1).

    Follows /root/reference/utils/utils.py:289-319.  ``knn`` may carry precomputed
    (dists, idx) from :func:`knn_smallest`.  Branches:
      * k-th only (utils.py:301-308): column k-1, RMS over b1 values;
      * average   (utils.py:309-317): all b1*k values through RMS, mean over k;
    then in both: divide by the *updated* running mean when knn_rms, ``max(r - clip, 0)`` when
    clip >= 0, and ``log(r + 1)`` (utils.py:318).
    """
    b1 = source.size(0)
    dists = knn[0] if knn is not None else knn_smallest(source, target, knn_k)[0]
    # Views, not copies: in the k-th branch the reference hands RMS a stride-k view
    # (utils.py:302-304) and ATen's mean/var summation order depends on the strides.
    if knn_avg:
        r = dists.reshape(-1, 1)
    else:
        r = dists[:, -1].reshape(-1, 1)
    if knn_rms:
        r = r / rms_step(state, r).M
    if knn_clip >= 0.0:
        r = torch.maximum(r - knn_clip, torch.zeros_like(r))
    if knn_avg:
        r = r.reshape(b1, knn_k).mean(dim=1, keepdim=True)
    return torch.log(r + 1.0)


def proto_enqueue(queue: torch.Tensor, ptr: int, rows: torch.Tensor):
    """FIFO write of candidate rows, returns the new pointer.

    Follows /root/reference/agents/unsupervised_learning/proto.py:110-112.
    """
    n = rows.size(0)
    queue[ptr:ptr + n] = rows
    return (ptr + n) % queue.size(0)


def proto_candidates(z: torch.Tensor, protos_weight: torch.Tensor, noise: torch.Tensor):
    """One batch row per prototype, with the sampling noise made explicit.

    Follows /root/reference/agents/unsupervised_learning/proto.py:105-107:
    ``scores = protos(z).T`` (bias-free nn.Linear), ``prob = softmax(scores, dim=1)``,
    ``Categorical(prob).sample()``.  Categorical renormalises its probs (``probs /
    probs.sum(-1, keepdim=True)``) and torch.multinomial with one sample is the exponential
    race ``argmax(probs / q)`` with ``q = empty_like(probs).exponential_(1)`` drawn from the
    generator; ``noise`` is that q (tests/test_oracle_pins.py checks the restatement against the
    reference lines run under the seed that produced the recorded q).
    """
    scores = torch.nn.functional.linear(z, protos_weight).T
    prob = torch.nn.functional.softmax(scores, dim=1)
    prob = prob / prob.sum(-1, keepdim=True)
    return torch.argmax(prob / noise, dim=1), prob


def proto_reward(z: torch.Tensor, queue: torch.Tensor, topk: int):
    """ProtoRL intrinsic reward: distance to the topk-th nearest queue entry, shape (B, 1).

    Follows /root/reference/agents/unsupervised_learning/proto.py:115-118.
    """
    d = torch.norm(z[:, None, :] - queue[None, :, :], dim=2, p=2)
    near, _ = torch.topk(d, topk, dim=1, largest=False)
    return near[:, -1:]


def rnd_reward(prediction: torch.Tensor, target: torch.Tensor, state: RmsState, rnd_scale=1.0):
    """RND prediction error and normalised reward.
    /root/reference/agents/unsupervised_learning/rnd.py:52-54 (error) and :94-99 (reward)."""
    prediction_error = torch.square(target.detach() - prediction).mean(dim=-1, keepdim=True)  # :52
    var = rms_step(state, prediction_error).S                                                 # :96
    reward = rnd_scale * prediction_error / (torch.sqrt(var) + 1e-8)                          # :97
    return reward, prediction_error
