"""Screen-kernel time vs bank-shard size at a fixed query count (strong-scaling model of C5):
kernel_ms(nb) = a * nb + b; b is the per-shard fixed cost (threshold warm-up) that limits the
G-GPU efficiency.  One JSON line per shard size.  EXB_SCREEN_DEBUG=1 skips the filter (MMA only).
"""
import argparse
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exorl_b200 as xb  # noqa: E402
from exorl_b200 import _lib  # noqa: E402
from exorl_b200 import knn as xknn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--k", type=int, default=12)
    ap.add_argument("--shards", default="1,2,4,8")
    ap.add_argument("--reps", type=int, default=4)
    ap.add_argument("--query-blocks", action="store_true",
                    help="shard the QUERIES instead (block of n/G queries against the whole bank)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    g = torch.Generator(device=dev).manual_seed(5)
    rep = torch.tanh(torch.randn(a.n, a.dim, generator=g, device=dev))
    for G in [int(s) for s in a.shards.split(",")]:
        nb = a.n // G
        lo = (G // 2) * nb if G > 1 else 0           # a middle shard
        bank = rep if G == 1 else rep[lo:lo + nb]
        src = rep
        if a.query_blocks:
            src, bank, lo = bank, rep, 0
        xb.knn_topk(src, bank, a.k, idx_base=lo, mode="screened")     # warm-up
        torch.cuda.synchronize()
        from bench import ClockSampler
        sampler = ClockSampler(0)
        _lib.check(lib.exb_profile_begin(1))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            xb.knn_topk(src, bank, a.k, idx_base=lo, mode="screened")
        e1.record()
        torch.cuda.synchronize()
        kms = (ctypes.c_float * 64)()
        kn = ctypes.c_int(0)
        _lib.check(lib.exb_profile_collect_sync(kms, 64, ctypes.byref(kn)))
        _lib.check(lib.exb_profile_begin(0))
        clocks = sampler.stop()
        und, invalid, pairs, _ = xknn.screened_status(dev)
        k_ms = sum(kms[i] for i in range(kn.value)) / max(kn.value, 1)
        extra = {}
        if os.environ.get("EXB_SCREEN_TRACE"):
            ws, _ = _lib.Workspace.get(dev, "knn_screened", 1024)
            out = (ctypes.c_int64 * 16)()
            _lib.check(lib.exb_debug_trace_sync(_lib.ptr(ws), out, _lib.stream_ptr(dev)))
            t = list(out)
            n_cta = max(t[7], 1)
            wpc = 4 if os.environ.get("EXB_SCREEN_FILTER", "3") in ("0", "2") else 8
            n_warp = 2 * wpc * n_cta    # epilogue warps per CTA x 2 CTAs per issuing leader
            extra = {"per_epilogue_warp_kcycles": {"wait_acc": t[0] / n_warp / 1e3,
                                                   "filter": t[1] / n_warp / 1e3,
                                                   "slow_events": t[2] / n_warp,
                                                   "slow": t[3] / n_warp / 1e3},
                     "per_issuer_kcycles": {"wait_free_acc": t[4] / n_cta / 1e3,
                                            "wait_tma": t[5] / n_cta / 1e3,
                                            "total": t[6] / n_cta / 1e3}, "issuers": t[7],
                     "per_insert": {"inserts_per_row": t[12] / (32 * n_warp),
                                    "store_cycles": t[13] / max(t[12], 1),
                                    "repacks_per_row": t[14] / (32 * n_warp),
                                    "chain_refresh_cycles": t[15] / max(t[12], 1)}}
        print(json.dumps({"shards": G, "nb": bank.size(0), "nq": src.size(0),
                          "sharded": "queries" if a.query_blocks else "bank",
                          "wave_splits": os.environ.get("EXB_SCREEN_WAVESPLITS", "auto"),
                          "screen_ms": k_ms,
                          "local_knn_ms": e0.elapsed_time(e1) / a.reps,
                          "tflops": 2.0 * src.size(0) * bank.size(0) * a.dim / (k_ms * 1e-3) / 1e12,
                          "undecided": und, "invalid": invalid, "rerank_pairs": pairs, "sm_mhz": clocks.get("sm_mhz"),
                          "power_w": clocks.get("power_w_max"),
                          "debug": os.environ.get("EXB_SCREEN_DEBUG", "0"), **extra}), flush=True)


if __name__ == "__main__":
    main()
