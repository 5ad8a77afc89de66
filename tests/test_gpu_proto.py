"""ProtoRL candidate sampling fused with the queue write (exb_proto_sample_f32):
parity with the reference's proto.py:105-112 on its own recorded sampling noise, the device Philox
stream's distribution, and the whole compute_intr_reward data path at BASELINE config 2 sizes."""
import numpy as np
import pytest
import torch

import exorl_b200 as xb
from helpers import reps
from oracle import proto_candidates, proto_reward

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _assert_candidates(got, z, w, noise, want=None):
    """Equal to the oracle's argmax except where the oracle's own best two ratios are a near-tie
    (a different fp32 summation order of the 128-term scores may swap them)."""
    cand, prob = proto_candidates(z, w, noise)
    if want is not None:
        assert np.array_equal(cand.numpy(), want)
    ratio = prob / noise
    got = got.cpu()
    bad = (got != cand).nonzero().flatten().tolist()
    for p in bad:
        a, b = float(ratio[p, cand[p]]), float(ratio[p, got[p]])
        assert abs(a - b) <= 1e-5 * abs(a), (p, int(cand[p]), int(got[p]), a, b)
    assert len(bad) <= max(1, len(cand) // 50)


@pytest.mark.parametrize("tag", ["small", "peaked", "odd"])
def test_reference_fixture_candidates_and_queue(golden, tag):
    g = golden("proto_sampling_reference.npz")
    z, w = torch.from_numpy(g[f"{tag}_z"]), torch.from_numpy(g[f"{tag}_w"])
    noise = torch.from_numpy(g[f"{tag}_noise"])
    queue = torch.from_numpy(g[f"{tag}_queue_before"]).to(DEV)
    ptr, new_ptr = (int(v) for v in g[f"{tag}_ptr"])
    got_ptr, cand = xb.proto_sample_enqueue(z.to(DEV), w.to(DEV), queue, ptr, noise=noise.to(DEV))
    assert got_ptr == new_ptr and cand.dtype == torch.int64 and cand.shape == (w.size(0),)
    _assert_candidates(cand, z, w, noise, want=g[f"{tag}_candidates"])
    if np.array_equal(cand.cpu().numpy(), g[f"{tag}_candidates"]):
        assert np.array_equal(queue.cpu().numpy(), g[f"{tag}_queue_after"])
    # rows outside the slice are untouched; rows inside are the chosen z rows, bit for bit
    q = queue.cpu()
    before = torch.from_numpy(g[f"{tag}_queue_before"])
    n = w.size(0)
    assert torch.equal(q[:ptr], before[:ptr]) and torch.equal(q[ptr + n:], before[ptr + n:])
    assert torch.equal(q[ptr:ptr + n], z[cand.cpu()])


def test_config2_sizes_with_module_and_recorded_noise():
    """1024 z x 512 prototypes x 128 (proto.py defaults) with the prototype nn.Linear passed as
    the reference holds it, noise drawn by torch on the host (the reference's stream)."""
    z = reps(1, 1024, 128, "unit")
    protos = torch.nn.Linear(128, 512, bias=False)
    with torch.no_grad():
        protos.weight.copy_(reps(2, 512, 128, "unit") * 8.0)
    torch.manual_seed(3)
    noise = torch.empty(512, 1024).exponential_(1)
    queue = torch.zeros(2048, 128, device=DEV)
    ptr, cand = xb.proto_sample_enqueue(z.to(DEV), protos.to(DEV), queue, 1536,
                                        noise=noise.to(DEV))
    assert ptr == 0                                       # (1536 + 512) % 2048, proto.py:112
    _assert_candidates(cand, z, protos.weight.detach().cpu(), noise)
    assert torch.equal(queue[1536:].cpu(), z[cand.cpu()]) and not queue[:1536].any()
    # then the reward of proto.py:115-118 against the updated queue
    r = xb.proto_knn_reward(z.to(DEV), queue, 3)
    np.testing.assert_allclose(r.cpu().numpy(), proto_reward(z, queue.cpu(), 3).numpy(), rtol=1e-5,
                               atol=1e-6)


def test_philox_stream_is_deterministic_and_follows_the_softmax():
    """Without `noise` the Exp(1) variates come from the device Philox stream: same (seed, call)
    -> same candidates; over many calls the empirical frequencies approach the softmax."""
    z = reps(5, 64, 32, "unit")
    w = reps(6, 16, 32, "unit") * 4.0
    zd, wd = z.to(DEV), w.to(DEV)
    queue = torch.zeros(64, 32, device=DEV)
    _, a = xb.proto_sample_enqueue(zd, wd, queue, 0, seed=7, call=3)
    _, b = xb.proto_sample_enqueue(zd, wd, queue, 0, seed=7, call=3)
    _, c = xb.proto_sample_enqueue(zd, wd, queue, 0, seed=7, call=4)
    assert torch.equal(a, b) and not torch.equal(a, c)
    counts = torch.zeros(16, 64)
    calls = 4000
    for i in range(calls):
        _, cand = xb.proto_sample_enqueue(zd, wd, queue, 0, seed=11, call=i)
        counts[torch.arange(16), cand.cpu()] += 1
    prob = torch.softmax((z @ w.T).T, dim=1)
    err = (counts / calls - prob).abs().max()
    assert err < 4.5 * (0.25 / calls) ** 0.5, float(err)      # 4.5 sigma of a binomial at p <= 1/2
    # default call counter: successive calls draw fresh noise
    _, d1 = xb.proto_sample_enqueue(zd, wd, queue, 0, seed=1)
    _, d2 = xb.proto_sample_enqueue(zd, wd, queue, 0, seed=1)
    assert not torch.equal(d1, d2)


def test_argument_errors_mirror_the_reference():
    z, w = reps(1, 32, 16, "unit").to(DEV), reps(2, 8, 16, "unit").to(DEV)
    queue = torch.zeros(20, 16, device=DEV)
    with pytest.raises(RuntimeError, match="expanded size"):
        xb.proto_sample_enqueue(z, w, queue, 16)               # slice clipped at the queue end
    with pytest.raises(RuntimeError, match="cannot be multiplied"):
        xb.proto_sample_enqueue(z, w[:, :8].contiguous(), queue, 0)
    with pytest.raises(RuntimeError, match="no CPU path"):
        xb.proto_sample_enqueue(z.cpu(), w, queue, 0)


def test_compiled_fullgraph_proto_update():
    """compute_intr_reward's data path (sampling + enqueue + kNN reward) in one compiled region."""
    z = reps(9, 256, 64, "unit").to(DEV)
    w = (reps(10, 32, 64, "unit") * 5.0).to(DEV)
    noise = torch.empty(32, 256).exponential_(1).to(DEV)
    q1 = torch.zeros(128, 64, device=DEV)
    q2 = torch.zeros(128, 64, device=DEV)

    def step(zz, queue):
        ptr, cand = xb.proto_sample_enqueue(zz, w, queue, 32, noise=noise)
        return xb.proto_knn_reward(zz, queue, 3), cand

    want_r, want_c = step(z, q1)
    got_r, got_c = torch.compile(step, fullgraph=True, backend="aot_eager")(z, q2)
    assert torch.equal(got_c, want_c) and torch.equal(q1, q2) and torch.equal(got_r, want_r)
