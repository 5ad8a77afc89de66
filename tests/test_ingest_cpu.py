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


def _write_npz(path, arrays, level, strategy=None):
    """An .npz whose members are deflated at a chosen level (np.savez_compressed is level 6)."""
    import zipfile
    import zlib
    with zipfile.ZipFile(path, "w") as z:
        for name, a in arrays.items():
            buf = io.BytesIO()
            np.save(buf, a)
            if strategy is None:
                z.writestr(name + ".npy", buf.getvalue(),
                           compress_type=zipfile.ZIP_DEFLATED if level else zipfile.ZIP_STORED,
                           compresslevel=level or None)
            else:   # fixed-Huffman / RLE / Huffman-only streams: written by hand into the member
                co = zlib.compressobj(level, zlib.DEFLATED, -15, 8, strategy)
                comp = co.compress(buf.getvalue()) + co.flush()
                zi = zipfile.ZipInfo(name + ".npy")
                zi.compress_type = zipfile.ZIP_DEFLATED
                # zipfile would deflate again: write a stored member, then patch nothing — instead
                # emit through the private hook that takes pre-compressed bytes
                zi.file_size = len(buf.getvalue())
                zi.compress_size = len(comp)
                zi.CRC = zlib.crc32(buf.getvalue())
                zi.header_offset = z.fp.tell()
                z.fp.write(zi.FileHeader())
                z.fp.write(comp)
                z.filelist.append(zi)
                z.NameToInfo[zi.filename] = zi
                z.start_dir = z.fp.tell()
                z._didModify = True


@pytest.mark.parametrize("level", [0, 1, 6, 9])
def test_deflate_levels_and_data_shapes_bit_exact(tmp_path, level):
    """Stored, fast, default and best deflate; incompressible, constant, periodic and sparse data
    (long matches at distance 1, short distances, literal runs) — equal to np.load."""
    rng = np.random.default_rng(level)
    arrays = {"noise": rng.standard_normal((2001, 24)).astype(np.float32),
              "zeros": np.zeros((5000, 7), np.float32),
              "ramp": (np.arange(300000) % 251).astype(np.uint8),
              "frames": rng.integers(0, 4, (30, 9, 84, 84), dtype=np.uint8),
              "sparse": (rng.random((4000, 50)) < 0.01).astype(np.float64),
              "tiny": np.array([1.5], np.float32),
              "empty": np.zeros((0, 3), np.float32)}
    fn = tmp_path / f"episode_0_{level}.npz"
    _write_npz(fn, arrays, level)
    with np.load(fn) as z:
        want = {k: z[k] for k in z.files}
    _same(ingest.read_episode(fn, pin=False), want)


@pytest.mark.parametrize("strategy", ["fixed", "rle", "huffman_only"])
def test_deflate_block_types_bit_exact(tmp_path, strategy):
    import zlib
    rng = np.random.default_rng(7)
    arrays = {"a": rng.integers(0, 16, (70000,), dtype=np.uint8),
              "b": np.repeat(rng.standard_normal(300).astype(np.float32), 40)}
    fn = tmp_path / "episode_0_1.npz"
    _write_npz(fn, arrays, 6, {"fixed": zlib.Z_FIXED, "rle": zlib.Z_RLE,
                               "huffman_only": zlib.Z_HUFFMAN_ONLY}[strategy])
    with np.load(fn) as z:        # the hand-written members are valid for the reference reader
        want = {k: z[k] for k in z.files}
    _same(ingest.read_episode(fn, pin=False), want)


@pytest.mark.parametrize("level", [0, 6])
def test_payload_corruption_is_detected_by_crc(tmp_path, level):
    """np.load raises BadZipFile('Bad CRC-32') on a damaged member; so does the native reader —
    for stored members too, where no decoder would notice."""
    import zipfile
    rng = np.random.default_rng(11)
    fn = tmp_path / "episode_0_2.npz"
    _write_npz(fn, {"observation": rng.standard_normal((500, 24)).astype(np.float32)}, level)
    raw = bytearray(fn.read_bytes())
    with zipfile.ZipFile(fn) as z:
        info = z.infolist()[0]
    at = info.header_offset + 30 + len(info.filename) + 200 + info.compress_size // 2
    raw[at] ^= 0x10
    bad = tmp_path / "episode_0_3.npz"
    bad.write_bytes(bytes(raw))
    with pytest.raises(Exception):
        with np.load(bad) as z:
            z["observation"]
    with pytest.raises((ValueError, RuntimeError)):
        ingest.read_episode(bad, pin=False)


def test_zlib_path_gives_the_same_bytes(tmp_path):
    """exb_npz_set_inflate_mode: 2 = own decoder only (no zlib behind it), 1 = default, 0 = zlib
    only — the same arrays from all three."""
    import hashlib
    from exorl_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(5)
    fn = tmp_path / "episode_0_1000.npz"
    write_episode_npz(_episode(rng, 1000), fn)
    outs = []
    prev = lib.exb_npz_set_inflate_mode(1)
    try:
        for mode in (2, 1, 0):
            lib.exb_npz_set_inflate_mode(mode)
            ep = ingest.read_episode(fn, pin=False)
            outs.append(hashlib.sha256(b"".join(ep[k].numpy().tobytes()
                                                for k in sorted(ep))).hexdigest())
        assert lib.exb_npz_set_inflate_mode(7) == 0      # out of range: unchanged
    finally:
        lib.exb_npz_set_inflate_mode(prev)
    assert outs[0] == outs[1] == outs[2] and len(outs[0]) == 64
