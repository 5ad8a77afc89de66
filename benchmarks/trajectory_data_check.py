"""How the screen behaves on trajectory-like data (consecutive states are near-duplicates), not
the i.i.d. synthetic reps of the headline config: undecided rows, candidates per row, time.
obs = per-episode random walk (1000 steps), rep = ICM-trunk-shaped tanh(LayerNorm(obs W^T))."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exorl_b200 as xb  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--episodes", type=int, default=200)
    ap.add_argument("--step", type=float, default=0.05)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    L = 1000
    for step in (a.step, a.step / 10, 0.0):
        start = torch.randn(a.episodes, 1, 24, generator=g, device=dev)
        walk = torch.cumsum(step * torch.randn(a.episodes, L, 24, generator=g, device=dev), dim=1)
        obs = (start + walk).reshape(-1, 24)
        if step == 0.0:   # i.i.d. control
            obs = torch.randn(a.episodes * L, 24, generator=g, device=dev)
        w = torch.randn(512, 24, generator=g, device=dev) / 24 ** 0.5
        z = torch.zeros(512, device=dev)
        p = xb.icm_get_rep((w, z, torch.ones(512, device=dev), z), obs)
        for _ in range(2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            d, i = xb.knn_topk(p, p, 12)
            e1.record()
            torch.cuda.synchronize()
        und, invalid, pairs, _ = xb.knn.screened_status(dev)
        why = None
        if os.environ.get("EXB_SCREEN_TRACE"):
            import ctypes
            from exorl_b200 import _lib
            ws, _ = _lib.Workspace.get(dev, "knn_screened", 1024)
            out = (ctypes.c_int64 * 16)()
            _lib.check(_lib.load().exb_debug_trace_sync(_lib.ptr(ws), out, _lib.stream_ptr(dev)))
            t = list(out)
            why = {"undecided": t[0], "list_overflow": t[1], "fewer_than_k": t[2],
                   "fp16_range": t[3], "survivors_gt_64": t[4],
                   "last": {"row": t[12], "why": t[13], "total": t[14], "n_surv": t[15]}}
        d_e, i_e = xb.knn_topk(p.rep, p.rep, 12, mode="exact")
        same = bool(torch.allclose(d, d_e, rtol=1e-6, atol=1e-6))
        print(json.dumps({"rows": obs.size(0), "walk_step": step, "ms": e0.elapsed_time(e1),
                          "undecided_rows": und, "overflowed_device_fallback": invalid,
                          "rerank_pairs_per_row": pairs / obs.size(0),
                          "median_knn_dist": float(d[:, 1:].median()),
                          "equals_exact_kernel": same, "why": why}), flush=True)


def per_update():
    """Per-update chain (B = 1024) on the same kinds of data: one whole episode (1024 consecutive
    states) as the batch."""
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(1)
    w = torch.randn(512, 24, generator=g, device=dev) / 24 ** 0.5
    z = torch.zeros(512, device=dev)
    import itertools
    for block_scan, step in itertools.product((False, True), (0.05, 0.005, 0.0)):
        xb.knn.set_per_update_scan(block_scan)
        obs = torch.randn(1, 24, generator=g, device=dev) + torch.cumsum(
            step * torch.randn(1024, 24, generator=g, device=dev), dim=0)
        if step == 0.0:
            obs = torch.randn(1024, 24, generator=g, device=dev)
        rep = xb.icm_get_rep((w, z, torch.ones(512, device=dev), z), obs).rep
        pbe = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev)
        ts = []
        for _ in range(20):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = pbe(rep)
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        und = xb.knn.screened_status(dev, key="apt_reward")[0]
        pe = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev)
        pe.mode = "exact"
        te = []
        for _ in range(20):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            re_ = pe(rep)
            e1.record()
            e1.synchronize()
            te.append(e0.elapsed_time(e1) * 1e3)
        same = bool(torch.allclose(r, re_, rtol=1e-5, atol=1e-7))
        print(json.dumps({"per_update_batch": 1024, "walk_step": step,
                          "block_cooperative_scan": block_scan, "us_median": sorted(ts)[10],
                          "exact_path_us_median": sorted(te)[10],
                          "rows_recomputed_inline": und, "equals_exact_path": same}), flush=True)


if __name__ == "__main__":
    import sys
    per_update()
    if "--per-update-only" in sys.argv:
        sys.exit(0)
    main()
