"""CPU checks of the drop-in boundary: the C-ABI library builds/loads and exports every symbol
include/exorl_b200.h declares; host-side argument errors; no compute without a GPU."""
import pathlib
import re

import pytest
import torch

import exorl_b200
from exorl_b200 import _lib

HEADER = pathlib.Path(__file__).resolve().parent.parent / "include" / "exorl_b200.h"


def _declared_symbols():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(exb_[a-z0-9_]+)\s*\(", text)))


def test_library_loads_and_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert lib.exb_abi_version() == 2


def test_binding_table_matches_header():
    assert sorted(_lib.SIGNATURES) == _declared_symbols()


def test_workspace_queries_need_no_gpu():
    lib = _lib.load()
    assert lib.exb_pbe_workspace_bytes(1024, 12) > 0
    assert lib.exb_knn_exact_workspace_bytes(0, 10, 4, 3) == 0      # empty input
    assert lib.exb_knn_exact_workspace_bytes(10, 10, 4, 99) == 0    # k beyond EXB_MAX_K


def test_cpu_tensors_are_rejected_loudly():
    x = torch.randn(8, 4)
    with pytest.raises(RuntimeError, match="no CPU path"):
        exorl_b200.knn_topk(x, x, 2)
    rms = exorl_b200.RMS("cpu")
    pbe = exorl_b200.PBE(rms, 0.0, 2, True, True, "cpu")
    with pytest.raises(RuntimeError, match="no CPU path"):
        pbe(x)
    with pytest.raises(RuntimeError, match="CUDA device"):
        exorl_b200.DeviceReplayBuffer("cpu", 100, 3, 0.99)


def test_product_does_not_import_oracle():
    root = pathlib.Path(exorl_b200.__file__).parent
    for f in root.rglob("*.py"):
        assert "oracle" not in f.read_text().replace("# oracle", ""), f
    for f in (root / "csrc").glob("*"):
        assert "oracle" not in f.read_text(), f


def test_shard_bounds_partition():
    from exorl_b200.relabel import shard_bounds
    for n, w in ((10, 3), (1_000_000, 8), (7, 7), (5, 2)):
        b = [shard_bounds(n, w, r) for r in range(w)]
        assert b[0][0] == 0 and b[-1][1] == n
        assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
        sizes = [hi - lo for lo, hi in b]
        assert max(sizes) - min(sizes) <= 1


def test_rms_general_shape_cpu_formula():
    """RMS with a non-(1,) shape is not on the kNN path and uses the reference formula with torch
    ops on whatever device the state lives on."""
    r = exorl_b200.RMS("cpu", shape=(3,))
    x = torch.arange(12.).reshape(4, 3)
    m, s = r(x)
    assert m.shape == (3,) and r.n == pytest.approx(4.0001)
    assert torch.allclose(m, x.mean(0) * 4 / 4.0001, rtol=1e-6)


def test_symmetric_multi_gpu_spans_cover_every_tile_pair_exactly_once():
    """relabel.sym_span: over all ranks, every unordered pair of bank tiles {I, J} (I != J) is
    scanned exactly once (by a pair-CTA of tile I scanning tile J or the other way round) and
    every diagonal tile exactly once; the per-rank work is balanced."""
    from exorl_b200.relabel import sym_tile_blocks, sym_span
    for n, world in ((1_000_000, 8), (1_000_000, 2), (1_000_000, 4), (300_000, 3), (70_000, 5),
                     (9_000, 7), (4_096, 2), (262_144, 6), (5_000, 1)):
        tiles, tb, rb = sym_tile_blocks(n, world)
        assert rb[0] == 0 and rb[-1] == n and tb[-1] == tiles
        seen = {}
        work = []
        for r in range(world):
            lo, hi, span_end, mid, extra = sym_span(tb, world, r)
            assert (lo, hi) == (tb[r], tb[r + 1])
            w = 0
            for u in range(lo, hi):
                length = span_end - u + (extra if u >= mid else 0)
                assert 1 <= length <= tiles
                for j in range(length):
                    t = (u + j) % tiles
                    key = (min(u, t), max(u, t))
                    seen[key] = seen.get(key, 0) + 1
                w += length
            work.append(w)
        assert len(seen) == tiles * (tiles + 1) // 2, (n, world)
        assert set(seen.values()) == {1}, (n, world)
        assert max(work) <= 1.05 * (sum(work) / world) + 2 * tiles / world + 8, (n, world, work)


def test_symmetric_rate_balanced_blocks_cover_once_and_equalise_time():
    """relabel.sym_balanced_blocks: blocks re-cut from per-rank times (slow GPUs get fewer tiles)
    still give every unordered tile pair to exactly one rank, stay within the skew bound the hit
    buffers are sized for, and equalise work x time-per-unit-of-work across the ranks."""
    from exorl_b200.relabel import sym_balanced_blocks, sym_span, SYM_BALANCE_MAX_SKEW
    cases = [(600, 8, [7.13, 6.82, 7.77, 7.80, 7.52, 7.68, 7.26, 7.69]),
             (391, 2, [29.6, 27.0]), (400, 4, [1.0, 1.3, 0.9, 1.0]), (350, 3, [1.0, 1.05, 1.2]),
             (256, 8, [1.0] * 8), (300, 5, [1.0, 5.0, 1.0, 1.0, 1.0])]
    for tiles, world, ms in cases:
        tb = sym_balanced_blocks(tiles, world, ms)
        assert tb == sym_balanced_blocks(tiles, world, list(ms))        # deterministic
        assert tb[0] == 0 and tb[-1] == tiles and all(b > a for a, b in zip(tb, tb[1:]))
        eq = tiles / world
        for a, b in zip(tb, tb[1:]):
            assert eq * (1 - SYM_BALANCE_MAX_SKEW) - 1.5 <= b - a <= eq * (1 + SYM_BALANCE_MAX_SKEW) + 1.5
        seen, work = {}, []
        for r in range(world):
            lo, hi, span_end, mid, extra = sym_span(tb, world, r)
            w = 0
            for u in range(lo, hi):
                length = span_end - u + (extra if u >= mid else 0)
                assert 1 <= length <= tiles
                for j in range(length):
                    t = (u + j) % tiles
                    key = (min(u, t), max(u, t))
                    seen[key] = seen.get(key, 0) + 1
                w += length
            work.append(w)
        assert len(seen) == tiles * (tiles + 1) // 2 and set(seen.values()) == {1}, (tiles, world)
        t_bal = [w * m for w, m in zip(work, ms)]
        if max(ms) / min(ms) < 1 + SYM_BALANCE_MAX_SKEW:       # inside the clamp: times equalised
            assert max(t_bal) <= 1.03 * min(t_bal) + 3 * tiles * max(ms) / world, (tiles, world)
        eq_tb = [(r * tiles) // world for r in range(world + 1)]
        t_eq = []
        for r in range(world):
            lo, hi, span_end, mid, extra = sym_span(eq_tb, world, r)
            t_eq.append(sum(span_end - u + (extra if u >= mid else 0) for u in range(lo, hi)) * ms[r])
        assert max(t_bal) <= max(t_eq) * 1.005, (tiles, world, max(t_bal), max(t_eq))


def test_sample_parts_partition_the_strided_sample():
    """relabel.sample_parts: the per-rank pieces of rep[::stride] (each rank uploads the sample
    rows of its own shard, the pieces are all-gathered) are disjoint, ordered and complete."""
    from exorl_b200.relabel import sample_parts, shard_bounds
    for n, world, stride in ((1_000_000, 8, 16), (150_000, 2, 16), (160_000, 8, 16), (1000, 4, 7),
                             (999, 3, 16), (64, 8, 16), (10, 1, 3)):
        first, counts = sample_parts(n, world, stride)
        rows = []
        for r in range(world):
            lo, hi = shard_bounds(n, world, r)
            part = [first[r] + j * stride for j in range(counts[r])]
            assert all(lo <= x < hi for x in part), (n, world, stride, r)
            rows += part
        assert rows == list(range(0, n, stride)), (n, world, stride)


def test_feed_chunk_cuts_cover_the_rows():
    from exorl_b200.relabel import feed_chunk_cuts
    for rows in (1, 5000, 119999, 120000, 125000, 130000, 250000, 500000, 1_000_000, 3_000_001):
        cuts = feed_chunk_cuts(rows)
        assert cuts[0] == 0 and cuts[-1] == rows and all(b > a for a, b in zip(cuts, cuts[1:]))
        assert len(cuts) <= 2 + rows // 32768
        if rows >= 120000:
            assert cuts[1] == 32768 and max(b - a for a, b in zip(cuts, cuts[1:])) <= 2 * 262144
            # growth <= 1.4 (the last chunk takes the remainder: up to 1.5 of its nominal size)
            assert all(c - b <= 2.1 * (b - a) + 1 for a, b, c in zip(cuts, cuts[1:], cuts[2:]))
