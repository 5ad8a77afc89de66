"""Symmetric screen (csrc/knn_sym.cuh) against the full screen and the exact kernel: equality of
the results and CUDA-event times, per problem size."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exorl_b200 as xb  # noqa: E402
from exorl_b200 import knn as xknn  # noqa: E402


def timed(fn, iters):
    fn()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2], out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", type=int, nargs="+", default=[65536, 131072])
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--k", type=int, default=12)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--exact-max", type=int, default=140000)
    ap.add_argument("--kind", default="tanh")
    ap.add_argument("--sample-max", type=int, default=0, help="override knn.SYM_SAMPLE_MAX")
    ap.add_argument("--no-full", action="store_true", help="skip the full-screen comparison")
    ap.add_argument("--knob", action="append", default=[], help="NAME=VALUE for exb_debug_knob_set")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    import ctypes
    from exorl_b200 import _lib
    lib = _lib.load()
    lib.exb_debug_knob_set(b"EXB_SYM_STATS", 1)
    for kv in a.knob:
        name, val = kv.split("=")
        lib.exb_debug_knob_set(name.encode(), int(val))
    for n in a.sizes:
        g = torch.Generator(device=dev).manual_seed(5)
        x = torch.randn(n, a.dim, generator=g, device=dev)
        if a.kind == "tanh":
            x = torch.tanh(x)
        elif a.kind == "walk":
            w = torch.randn(24, a.dim, generator=g, device=dev) / 24 ** 0.5
            obs = torch.randn(n // 1000 + 1, 1, 24, generator=g, device=dev) + torch.cumsum(
                0.05 * torch.randn(n // 1000 + 1, 1000, 24, generator=g, device=dev), dim=1)
            x = torch.tanh(torch.nn.functional.layer_norm(obs.reshape(-1, 24)[:n] @ w, (a.dim,)))
        xknn.SYM_MIN_ROWS = 1024
        if a.sample_max:
            xknn.SYM_SAMPLE_MAX = a.sample_max
        xknn.SYM_ENABLED = True
        xknn.sym_last.clear()
        t_sym, (d_s, i_s) = timed(lambda: xb.knn_topk(x, x, a.k, mode="screened"), a.iters)
        info = dict(xknn.sym_last)
        # one more call with the library's event brackets around the two screen launches
        kms, kn = (ctypes.c_float * 64)(), ctypes.c_int(0)
        _lib.check(lib.exb_profile_begin(1))
        xb.knn_topk(x, x, a.k, mode="screened")
        _lib.check(lib.exb_profile_collect_sync(kms, 64, ctypes.byref(kn)))
        _lib.check(lib.exb_profile_begin(0))
        info["screen_launch_ms"] = [round(kms[i], 3) for i in range(kn.value)]
        ws, _ = _lib.Workspace.get(dev, "knn_sym", 1024)
        tr = (ctypes.c_int64 * 16)()
        _lib.check(lib.exb_debug_trace_sync(_lib.ptr(ws), tr, _lib.stream_ptr(dev)))
        t = list(tr)
        info["counts"] = {"direct_mean_per_half": t[0] / (2 * n), "direct_max": t[1],
                          "transposed_appended": t[2], "append_capacity": t[3],
                          "received_mean": t[4] / n, "received_max": t[5],
                          "direct_overflowed": t[6]}
        xknn.SYM_ENABLED = False
        out = {"n": n, "dim": a.dim, "k": a.k, "kind": a.kind, "sym_ms": t_sym, "sym_info": info}
        if not a.no_full:
            t_full, (d_f, i_f) = timed(lambda: xb.knn_topk(x, x, a.k, mode="screened"), a.iters)
            out.update({"full_ms": t_full, "speedup": t_full / t_sym,
                        "dist_equal_full": bool(torch.equal(d_s, d_f)),
                        "idx_rows_differing_from_full": int((i_s != i_f).any(dim=1).sum()),
                        "max_abs_dist_diff_vs_full": float((d_s - d_f).abs().max())})
        if n <= a.exact_max:
            d_e, i_e = xb.knn_topk(x, x, a.k, mode="exact")
            out["max_abs_dist_diff_vs_exact"] = float((d_s - d_e).abs().max())
            out["idx_rows_differing_from_exact"] = int((i_s != i_e).any(dim=1).sum())
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
