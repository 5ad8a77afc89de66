"""exb_pbe_reward_noalloc (the allocation-free RMS + PBE epilogue, static device scratch) against
exb_pbe_reward, both through the C ABI, on two calls in a row.

The file is named so that pytest collects it last: the test
passes alone in a fresh process and after the other GPU tests (the entry point clears the
scratch words that must start at zero itself)."""
import numpy as np
import pytest
import torch

from exorl_b200 import _lib
from oracle import RmsState, pbe_reward

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
RTOL = 1e-5          # the reward tolerance of tests/test_gpu_knn.py (north_star: rel 1e-5)
MAX_VALUES = 1 << 24  # EXB_PBE_NOALLOC_MAX_VALUES, include/exorl_b200.h


def _dist(nq, k, seed):
    g = torch.Generator().manual_seed(seed)
    d = torch.rand(nq, k, generator=g) * 3.0 + 0.01
    return torch.sort(d, dim=1).values.contiguous()


def _state():
    return (torch.zeros(1, device=DEV), torch.ones(1, device=DEV),
            torch.full((1,), 1e-4, dtype=torch.float64, device=DEV))


def _existing(lib, dist, k, avg, use_rms, clip, st, ws):
    out = torch.empty(dist.size(0), dtype=torch.float32, device=DEV)
    _lib.check(lib.exb_pbe_reward(_lib.ptr(dist), dist.size(0), k, avg, use_rms, clip,
                                  _lib.ptr(st[0]), _lib.ptr(st[1]), _lib.ptr(st[2]),
                                  _lib.ptr(out), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(DEV)))
    return out


def _noalloc(lib, dist, k, avg, use_rms, clip, st):
    out = torch.empty(dist.size(0), dtype=torch.float32, device=DEV)
    _lib.check(lib.exb_pbe_reward_noalloc(_lib.ptr(dist), dist.size(0), k, avg, use_rms, clip,
                                          _lib.ptr(st[0]), _lib.ptr(st[1]), _lib.ptr(st[2]),
                                          _lib.ptr(out), _lib.stream_ptr(DEV)))
    return out


def _same(a, b, rtol):
    np.testing.assert_allclose(a.double().cpu().numpy(), b.double().cpu().numpy(), rtol=rtol,
                               atol=1e-7)


@pytest.mark.parametrize("nq,k,avg,clip", [
    (8192, 12, 1, 0.0),      # both entry points run the multi-block moments kernel
    (300, 12, 1, 0.0005),    # the existing one takes its single-block kernel here
    (50000, 3, 0, -1.0),     # column k-1 only, no clip
])
def test_noalloc_matches_existing_twice(nq, k, avg, clip):
    lib = _lib.load()
    with torch.cuda.device(DEV):
        ws = torch.zeros(lib.exb_pbe_workspace_bytes(nq, k), dtype=torch.uint8, device=DEV)
        st_a, st_b = _state(), _state()
        for call in range(2):
            dist = _dist(nq, k, 100 * call + nq % 97).to(DEV)
            want = _existing(lib, dist, k, avg, 1, clip, st_a, ws)
            got = _noalloc(lib, dist, k, avg, 1, clip, st_b)
            torch.cuda.synchronize()
            assert torch.isfinite(got).all()
            _same(got, want, RTOL)
            for a, b in zip(st_a, st_b):     # RMS state (M, S, n) carried into the next call
                _same(b, a, 1e-6)
        assert float(st_b[2]) == pytest.approx(1e-4 + 2 * (nq * k if avg else nq))


@pytest.mark.parametrize("use_rms", [0, 1])
def test_noalloc_matches_oracle(use_rms):
    lib = _lib.load()
    state = RmsState()
    with torch.cuda.device(DEV):
        st = _state() if use_rms else (None, None, None)
        for call in range(2):
            dist = _dist(2048, 12, 7 + call)
            got = _noalloc(lib, dist.to(DEV), 12, 1, use_rms, 0.0, st)
            torch.cuda.synchronize()
            want = pbe_reward(dist, dist, state, 12, True, bool(use_rms), 0.0, knn=(dist, None))
            _same(got, want.flatten(), RTOL)


def test_noalloc_rejects_sizes_beyond_the_limit():
    lib = _lib.load()
    with torch.cuda.device(DEV):
        dist = _dist(64, 12, 1).to(DEV)
        st = _state()
        out = torch.empty(64, dtype=torch.float32, device=DEV)
        nq_big = MAX_VALUES // 12 + 1     # validated before anything is launched or read
        with pytest.raises(ValueError, match="exceeds the static scratch"):
            _lib.check(lib.exb_pbe_reward_noalloc(_lib.ptr(dist), nq_big, 12, 1, 1, 0.0,
                                                  _lib.ptr(st[0]), _lib.ptr(st[1]),
                                                  _lib.ptr(st[2]), _lib.ptr(out),
                                                  _lib.stream_ptr(DEV)))
        # the limit itself is accepted by the argument check (no launch with these dummies:
        # use_rms = 0 and a real, small call afterwards still works)
        got = _noalloc(lib, dist, 12, 1, 1, 0.0, st)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
