"""Native .npz episode reader (exb_npz_*, host-only) against numpy's own reader on files in the
reference's format (np.savez_compressed of one array per spec name,
utils/replay_buffer.py:18-23) — bit-exact, every dtype the reference stores."""
import io
import pathlib

import numpy as np
import pytest
import torch

import exorl_b200 as xb
from exorl_b200 import ingest
from oracle import write_episode_npz, read_episode_npz


def _episode(rng, L, pixels=False):
    n = L + 1
    obs = (rng.integers(0, 256, (n, 9, 84, 84), dtype=np.uint8) if pixels
           else rng.standard_normal((n, 24)).astype(np.float32))
    return {"observation": obs,
            "action": rng.uniform(-1, 1, (n, 6)).astype(np.float32),
            "reward": rng.uniform(0, 1, (n, 1)).astype(np.float32),
            "discount": np.ones((n, 1), np.float32),
            "skill": rng.standard_normal((n, 16)).astype(np.float32),
            "constraint": rng.integers(0, 2, (n, 1)).astype(bool),
            "done": np.zeros((n, 1), bool),
            "physics": rng.standard_normal((n, 18)).astype(np.float64),
            "step_id": np.arange(n, dtype=np.int64)}


def _same(got, want):
    assert set(got) == set(want)
    for k, w in want.items():
        g = got[k].numpy()
        assert g.dtype == w.dtype and g.shape == w.shape, k
        assert np.array_equal(g, w), k


@pytest.mark.parametrize("pixels", [False, True])
def test_reference_format_bit_exact(tmp_path, pixels):
    rng = np.random.default_rng(0)
    ep = _episode(rng, 20 if pixels else 1000, pixels)
    fn = tmp_path / "episode_0_20.npz"
    write_episode_npz(ep, fn)
    _same(ingest.read_episode(fn, pin=False), read_episode_npz(fn))
    _same(ingest.read_episode(fn, pin=False), ep)


def test_stored_members_and_key_filter(tmp_path):
    rng = np.random.default_rng(1)
    ep = _episode(rng, 7)
    fn = tmp_path / "plain.npz"
    np.savez(fn, **ep)                                  # ZIP_STORED members
    _same(ingest.read_episode(fn, pin=False), ep)
    sub = ingest.read_episode(fn, pin=False, keys=("reward", "done"))
    assert set(sub) == {"reward", "done"} and sub["done"].dtype == torch.bool


def test_zero_length_and_scalar_shapes(tmp_path):
    fn = tmp_path / "odd.npz"
    want = {"empty": np.zeros((0, 5), np.float32), "one": np.float32(3.5).reshape(()),
            "vec": np.arange(3, dtype=np.int32)}
    np.savez_compressed(fn, **want)
    got = ingest.read_episode(fn, pin=False)
    assert got["empty"].shape == (0, 5) and got["one"].shape == () and float(got["one"]) == 3.5
    assert np.array_equal(got["vec"].numpy(), want["vec"])


def test_parallel_reader_keeps_order(tmp_path):
    rng = np.random.default_rng(2)
    eps, paths = [], []
    for e in range(23):
        ep = _episode(rng, 5 + e)
        fn = tmp_path / f"episode_{e:03d}_{5 + e}.npz"
        write_episode_npz(ep, fn)
        eps.append(ep)
        paths.append(fn)
    out = list(ingest.read_episodes(paths, workers=4, pin=False))
    assert [p for p, _ in out] == paths
    for (_, got), want in zip(out, eps):
        _same(got, want)


def test_corrupt_and_unsupported_files_raise(tmp_path):
    rng = np.random.default_rng(3)
    fn = tmp_path / "episode_0_9.npz"
    write_episode_npz(_episode(rng, 9), fn)
    raw = fn.read_bytes()
    (tmp_path / "trunc.npz").write_bytes(raw[: len(raw) // 2])      # half-written file
    with pytest.raises((ValueError, RuntimeError)):
        ingest.read_episode(tmp_path / "trunc.npz", pin=False)
    (tmp_path / "junk.npz").write_bytes(b"not a zip archive at all" * 10)
    with pytest.raises((ValueError, RuntimeError)):
        ingest.read_episode(tmp_path / "junk.npz", pin=False)
    with pytest.raises((ValueError, RuntimeError)):
        ingest.read_episode(tmp_path / "missing.npz", pin=False)
    # big-endian / object arrays are outside the reference's format: refused, not misread
    np.savez_compressed(tmp_path / "be.npz", x=np.arange(4, dtype=">f4"))
    with pytest.raises((ValueError, RuntimeError)):
        ingest.read_episode(tmp_path / "be.npz", pin=False)
    # flipped payload byte: inflate either fails or (stored data) differs — never crashes
    bad = bytearray(raw)
    bad[len(raw) // 3] ^= 0xFF
    (tmp_path / "flip.npz").write_bytes(bytes(bad))
    try:
        ingest.read_episode(tmp_path / "flip.npz", pin=False)
    except (ValueError, RuntimeError):
        pass


def test_golden_replay_fixture_roundtrip(tmp_path, golden):
    """Episodes of the committed reference fixture, written in the reference's format and read
    back by the native reader, equal the fixture arrays."""
    g = golden("replay_reference.npz")
    names = sorted({k.split("_", 3)[3] for k in g.files if k.startswith("meta_n3_ep0_")})
    ep = {n: g[f"meta_n3_ep0_{n}"] for n in names}
    fn = tmp_path / "episode_0_x.npz"
    write_episode_npz(ep, fn)
    _same(ingest.read_episode(fn, pin=False), ep)
