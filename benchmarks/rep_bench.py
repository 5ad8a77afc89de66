"""Rep producer measurement at the C5 scale: 10^6 walker observations (24)
-> ICM trunk rep (512) -> prepared kNN operand.

  fused      exb_rep_trunk_f32: one kernel writes the fp32 rep + fp16 operand + norms
  unfused    torch CUDA ops (addmm, layer_norm, tanh) + exb_knn_prep_f16 — what the path costs
             when the rep comes from the framework and the kNN prepares it itself
Algorithmic bytes per row: read 4*in_dim, write 4*D + 2*Dpad + 8  (3176 B at 24 -> 512).
Inputs + outputs (3.2 GB) exceed L2, so no flush.  One JSON line per case; CUDA events.
"""
import argparse
import json
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import exorl_b200 as xb  # noqa: E402
from exorl_b200 import _lib  # noqa: E402

DEV = torch.device("cuda:0")
HBM_GBS = 6537.6
try:
    HBM_GBS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def time_gpu(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts), min(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--in-dim", type=int, default=24)
    ap.add_argument("--dim", type=int, default=512)
    a = ap.parse_args()
    lib = _lib.load()
    g = torch.Generator(device=DEV).manual_seed(0)
    obs = torch.randn(a.n, a.in_dim, device=DEV, generator=g)
    lin = torch.nn.Linear(a.in_dim, a.dim).to(DEV)
    ln = torch.nn.LayerNorm(a.dim).to(DEV)
    params = (lin.weight.detach(), lin.bias.detach(), ln.weight.detach(), ln.bias.detach())
    out = xb.PreparedRep(a.n, a.dim, DEV)
    st = _lib.stream_ptr(DEV)
    dpad = (a.dim + 63) // 64 * 64
    bytes_row = 4 * a.in_dim + 4 * a.dim + 2 * dpad + 8

    def fused():
        _lib.check(lib.exb_rep_trunk_f32(_lib.ptr(obs), a.n, a.in_dim, *map(_lib.ptr, params),
                                         1e-5, a.dim, _lib.ptr(out.rep), out.operand, st))

    def fused_op_only():
        _lib.check(lib.exb_rep_trunk_f32(_lib.ptr(obs), a.n, a.in_dim, *map(_lib.ptr, params),
                                         1e-5, a.dim, None, out.operand, st))

    def unfused():
        with torch.no_grad():
            rep = torch.tanh(ln(lin(obs)))
        _lib.check(lib.exb_knn_prep_f16(_lib.ptr(rep), a.n, a.dim, _lib.ptr(out.f16),
                                        _lib.ptr(out.half_norm), _lib.ptr(out.round_err),
                                        _lib.ptr(out.scalars), st))

    def prep_only():
        _lib.check(lib.exb_knn_prep_f16(_lib.ptr(out.rep), a.n, a.dim, _lib.ptr(out.f16),
                                        _lib.ptr(out.half_norm), _lib.ptr(out.round_err),
                                        _lib.ptr(out.scalars), st))

    for name, fn, nbytes in (("fused trunk -> rep + operand", fused, bytes_row * a.n),
                             ("fused trunk -> operand only", fused_op_only,
                              (bytes_row - 4 * a.dim) * a.n),
                             ("torch trunk + exb_knn_prep_f16", unfused, bytes_row * a.n),
                             ("exb_knn_prep_f16 alone", prep_only,
                              (4 * a.dim + 2 * dpad + 8) * a.n)):
        med, mn = time_gpu(fn)
        print(json.dumps({"case": name, "n": a.n, "in_dim": a.in_dim, "dim": a.dim,
                          "us_median": med, "us_min": mn, "algorithmic_bytes": nbytes,
                          "achieved_gbs": nbytes / med / 1e3, "hbm_peak_gbs": HBM_GBS,
                          "frac_of_hbm_peak": nbytes / med / 1e3 / HBM_GBS,
                          "trunk_r": os.environ.get("EXB_TRUNK_R", "4")}), flush=True)


if __name__ == "__main__":
    main()
