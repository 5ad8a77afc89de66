"""Scale-size oracle fixtures for BASELINE config 5 (SURVEY.md §8c "scalable oracle for large N").

Run from the repo root (≈40 min on 8 host cores):  python oracle/gen_scale_golden.py

Evaluates the reference's own expression — `torch.norm(source[:, None, :] - target[None, :, :],
dim=-1, p=2)` then `topk(k, largest=False, sorted=True)`, /root/reference/utils/utils.py:293-300 —
in tiles (oracle.knn_smallest, pinned bitwise to the untiled reference by
tests/test_oracle_pins.py) on

  (i)  ALL 32 768 queries of the N = 32 768, D = 512 set   -> tests/golden/knn_scale_32k.npz
  (ii) a fixed 4 096-query slice of the seed-5 10^6 x 512 bank, against the whole bank
                                                           -> tests/golden/knn_scale_1m.npz

The inputs are seeded CPU tensors (`tanh(randn(n, 512, generator=manual_seed(5)))`, SURVEY §8d
C5), which the GPU tests regenerate; `bank_sum` (fp64 sum of the bank) lets a test tell a
different random stream from a kNN defect.  Test infrastructure only.
"""
import pathlib
import sys
import time

import numpy as np
import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle.pbe_oracle import knn_smallest  # noqa: E402

OUT = ROOT / "tests" / "golden"
K = 12
DIM = 512


def bank(n, seed=5):
    g = torch.Generator().manual_seed(seed)
    return torch.tanh(torch.randn(n, DIM, generator=g))


def slice_rows(n, count):
    """The fixed query slice: `count` rows spread evenly over the bank."""
    return torch.arange(count, dtype=torch.int64) * (n // count)


def run(tag, n, rows):
    x = bank(n)
    t0 = time.time()
    parts_d, parts_i = [], []
    step = 512
    for lo in range(0, rows.numel(), step):
        d, i = knn_smallest(x[rows[lo:lo + step]], x, K)
        parts_d.append(d)
        parts_i.append(i)
        if (lo // step) % 8 == 0:
            print(f"{tag}: {lo + step}/{rows.numel()} queries, {time.time() - t0:.0f} s", flush=True)
    dist, idx = torch.cat(parts_d), torch.cat(parts_i)
    assert torch.equal(idx[:, 0], rows) and not dist[:, 0].any()     # self-match first
    np.savez(OUT / f"knn_scale_{tag}.npz", n=np.array(n), dim=np.array(DIM), k=np.array(K),
             seed=np.array(5), rows=rows.numpy(), dist=dist.numpy(),
             idx=idx.numpy().astype(np.int32), bank_sum=np.array(float(x.double().sum())),
             seconds=np.array(time.time() - t0), threads=np.array(torch.get_num_threads()))
    print(f"{tag}: done in {time.time() - t0:.0f} s")


if __name__ == "__main__":
    torch.set_num_threads(int(sys.argv[sys.argv.index("--threads") + 1])
                          if "--threads" in sys.argv else torch.get_num_threads())
    if "--only-1m" not in sys.argv:
        run("32k", 32768, torch.arange(32768, dtype=torch.int64))
    if "--only-32k" not in sys.argv:
        run("1m", 1_000_000, slice_rows(1_000_000, 4096))
