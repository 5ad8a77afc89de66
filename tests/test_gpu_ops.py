"""torch.ops.exorl_b200.* (exorl_b200/ops.py): schema/fake-kernel checks with torch.library.opcheck
and the reward call sites compiled with torch.compile(fullgraph=True) — icm_apt.py:96-100 /
aps.py:167-172 (PBE(rep)) and proto.py:115-118 — against the oracle."""
import types

import numpy as np
import pytest
import torch

import exorl_b200 as xb
from helpers import reps
from oracle import RmsState, pbe_reward, proto_reward, rnd_reward, rms_step

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _state():
    r = xb.RMS(DEV)
    return r.M, r.S, r._n


def test_opcheck_schemas_and_fake_kernels():
    x = reps(1, 256, 64).to(DEV)
    y = reps(2, 300, 64).to(DEV)
    checks = ("test_schema", "test_faketensor", "test_autograd_registration")
    torch.library.opcheck(torch.ops.exorl_b200.knn_topk.default, (x, y, 5, 0, "exact", False),
                          test_utils=checks)
    torch.library.opcheck(torch.ops.exorl_b200.knn_topk.default, (x, x, 5, 7, "screened", True),
                          test_utils=checks)
    torch.library.opcheck(torch.ops.exorl_b200.apt_reward.default,
                          (x, x, *_state(), 12, True, True, 0.0, "auto", True), test_utils=checks)
    torch.library.opcheck(torch.ops.exorl_b200.proto_reward.default, (x, y, 3, "auto"),
                          test_utils=checks)
    torch.library.opcheck(torch.ops.exorl_b200.rms_update.default,
                          (x[:, 0].contiguous(), *_state()), test_utils=checks)
    torch.library.opcheck(torch.ops.exorl_b200.rnd_reward.default, (x, x * 0.5, *_state(), 1.0),
                          test_utils=checks)


def test_replay_sample_op_opcheck_and_equals_buffer_call():
    rng = np.random.default_rng(5)
    buf = xb.DeviceReplayBuffer(DEV, 10 ** 4, 3, 0.99)
    for e, L in enumerate((20, 7, 33)):
        n = L + 1
        buf.add_episode({"observation": rng.standard_normal((n, 24)).astype(np.float32),
                         "action": rng.uniform(-1, 1, (n, 6)).astype(np.float32),
                         "reward": rng.uniform(0, 1, (n, 1)).astype(np.float32),
                         "discount": np.ones((n, 1), np.float32)}, key=f"episode_{e}_{L}.npz")
    _, starts, lens = buf._tables()
    obs, act = buf.cols["observation"], buf.cols["action"]
    args = ([obs, act, obs], [-1, 0, 2], buf.cols["reward"], buf.cols["discount"], starts, lens,
            128, 99, 4, 3, 1, 0.99)
    torch.library.opcheck(torch.ops.exorl_b200.replay_sample.default, args,
                          test_utils=("test_schema", "test_faketensor"))
    out = torch.ops.exorl_b200.replay_sample(*args)
    want, rows = buf.sample_philox(128, 99, 4, return_rows=True)
    assert torch.equal(out[0], want[0]) and torch.equal(out[1], want[1])
    assert torch.equal(out[2], want[4]) and torch.equal(out[3], want[2])
    assert torch.equal(out[4], want[3]) and torch.equal(out[5], rows)


def test_compiled_fullgraph_reward_call_sites_match_oracle():
    """compute_intr_reward of the APT/APS agents (`reward = self.pbe(rep)`) and of ProtoRL
    (z-to-queue kNN) inside ONE torch.compile(fullgraph=True) region: no graph break at the
    reward, RMS state advanced on the device by every call of the compiled function."""
    pbe = xb.PBE(xb.RMS(DEV), 0.0, 12, True, True, DEV)
    rms_rnd = xb.RMS(DEV)

    def rewards(rep, z, queue, pred, tgt):
        r_apt = pbe(rep)                                   # icm_apt.py:98 / aps.py:171
        r_proto = xb.proto_knn_reward(z, queue, 3)         # proto.py:115-118
        r_rnd, _ = xb.rnd_intr_reward(pred, tgt, rms_rnd, 0.5)      # rnd.py:94-99
        return r_apt * 2.0, r_proto + 1.0, r_rnd

    compiled = torch.compile(rewards, fullgraph=True, backend="aot_eager")
    st, st_rnd = RmsState(), RmsState()
    for call in range(3):
        rep = reps(40 + call, 1024, 512)
        z, q = reps(50 + call, 256, 128, "unit"), reps(60 + call, 512, 128, "unit")
        pred, tgt = reps(70 + call, 1024, 64, "raw"), reps(80 + call, 1024, 64, "raw")
        got = compiled(*(t.to(DEV) for t in (rep, z, q, pred, tgt)))
        want_apt = pbe_reward(rep, rep, st, 12, True, True, 0.0) * 2.0
        want_proto = proto_reward(z, q, 3) + 1.0
        want_rnd, _ = rnd_reward(pred, tgt, st_rnd, 0.5)
        np.testing.assert_allclose(got[0].cpu().numpy(), want_apt.numpy(), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(got[1].cpu().numpy(), want_proto.numpy(), rtol=1e-5)
        np.testing.assert_allclose(got[2].cpu().numpy(), want_rnd.numpy(), rtol=2e-5)
        assert pbe.rms.n == st.n and rms_rnd.n == st_rnd.n
        np.testing.assert_allclose(float(pbe.rms.M), float(st.M), rtol=2e-5)


def test_compiled_compute_apt_reward_source_ne_target():
    args = types.SimpleNamespace(knn_k=5, knn_avg=False, knn_rms=True, knn_clip=0.5)
    rms = xb.RMS(DEV)

    def f(s, t):
        return xb.compute_apt_reward(s, t, args, rms=rms)

    cf = torch.compile(f, fullgraph=True, backend="aot_eager")
    s, t = reps(7, 300, 50), reps(8, 700, 50)
    st = RmsState()
    for _ in range(2):
        got = cf(s.to(DEV), t.to(DEV))
        want = pbe_reward(s, t, st, 5, False, True, 0.5)
        np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=1e-5, atol=1e-7)
    assert rms.n == st.n


def test_rms_state_is_updated_in_place_and_n_reads_back():
    """M, S and n live on the device and are updated in place (same tensor objects across calls);
    `.n` reads the fp64 word back and is assignable; pickling carries all three."""
    import pickle
    r = xb.RMS(DEV)
    m_obj, s_obj = r.M, r.S
    st = RmsState()
    for i, n in enumerate((1000, 3, 50000)):
        x = reps(90 + i, n, 1, "raw") * 2 + 5
        m, s = r(x.to(DEV))
        rms_step(st, x)
        assert m is m_obj and s is s_obj and r.M is m_obj
        assert r.n == st.n
        np.testing.assert_allclose([float(m), float(s)], [float(st.M), float(st.S)], rtol=2e-5)
    q = pickle.loads(pickle.dumps(r))
    assert q.n == r.n and torch.equal(q.M, r.M) and torch.equal(q.S, r.S)
    q.n = 7.5
    assert q.n == 7.5 and r.n == st.n
