"""Episode ingestion throughput on the host cores: native parallel .npz reader vs the reference's
np.load loop (utils/replay_buffer.py:26-29), same files.  One JSON line."""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from exorl_b200 import ingest  # noqa: E402
from oracle import write_episode_npz, read_episode_npz  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--episodes", type=int, default=200)
ap.add_argument("--pixels", action="store_true")
ap.add_argument("--workers", type=int, default=os.cpu_count())
a = ap.parse_args()
rng = np.random.default_rng(0)
L = 1000 if not a.pixels else 250
with tempfile.TemporaryDirectory() as d:
    paths = []
    for e in range(a.episodes if not a.pixels else max(a.episodes // 20, 4)):
        n = L + 1
        obs = (rng.integers(0, 64, (n, 9, 84, 84), dtype=np.uint8) if a.pixels
               else rng.standard_normal((n, 24)).astype(np.float32))
        ep = {"observation": obs, "action": rng.uniform(-1, 1, (n, 6)).astype(np.float32),
              "reward": rng.uniform(0, 1, (n, 1)).astype(np.float32),
              "discount": np.ones((n, 1), np.float32)}
        p = os.path.join(d, f"episode_{e:06d}_{L}.npz")
        write_episode_npz(ep, p)
        paths.append(p)
    raw = sum(sum(v.nbytes for v in read_episode_npz(p).values()) for p in paths[:1]) * len(paths)
    t0 = time.perf_counter()
    for p in paths:
        read_episode_npz(p)
    t_ref = time.perf_counter() - t0
    t0 = time.perf_counter()
    for _ in ingest.read_episodes(paths, workers=1, pin=False):
        pass
    t_one = time.perf_counter() - t0
    t0 = time.perf_counter()
    for _ in ingest.read_episodes(paths, workers=a.workers, pin=False):
        pass
    t_par = time.perf_counter() - t0
    print(json.dumps({"case": "episode ingestion " + ("pixels" if a.pixels else "walker states"),
                      "episodes": len(paths), "raw_mb": raw / 1e6,
                      "np_load_s": t_ref, "native_1thread_s": t_one,
                      f"native_{a.workers}threads_s": t_par,
                      "np_load_mb_s": raw / 1e6 / t_ref, "native_mb_s": raw / 1e6 / t_par,
                      "cores": os.cpu_count()}))
