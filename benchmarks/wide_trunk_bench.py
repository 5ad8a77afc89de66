"""Wide ICM trunk (pixel APT: 39 200 encoder features -> 512, icm_apt.py:17-18): the 3xTF32
tcgen05 GEMM + LayerNorm/Tanh/operand pass (exb_rep_trunk_wide_f32) against torch's fp32 and tf32
library GEMMs followed by the fused LN pass.  Prints one JSON line per batch size with times
(CUDA events, L2 flushed between iterations), the tensor-pipe fraction of the GEMM kernel against
the measured bf16 peak / 2 (tf32 runs at half the bf16 rate) and errors against an fp64 reference.
"""
import argparse
import json
import pathlib
import sys

import torch

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import exorl_b200 as xb  # noqa: E402
from exorl_b200 import rep as xrep  # noqa: E402


def timed(fn, iters, flush):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(iters)]
    for _ in range(3):
        fn()
    for a, b in ev:
        flush.fill_(1.0)
        flush.sum()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    return ts[len(ts) // 2] * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--in-dim", type=int, default=39200)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--batches", type=int, nargs="+", default=[1024, 256, 4096])
    ap.add_argument("--iters", type=int, default=30)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    peaks = json.loads((pathlib.Path(__file__).resolve().parent.parent
                        / "MEASURED_PEAKS.json").read_text())
    bf16 = float(peaks.get("bf16_tflops", 1650.4))
    g = torch.Generator().manual_seed(0)
    w = (torch.randn(a.dim, a.in_dim, generator=g) / a.in_dim ** 0.5).to(dev)
    b = (0.1 * torch.randn(a.dim, generator=g)).to(dev)
    lnw = (1 + 0.1 * torch.randn(a.dim, generator=g)).to(dev)
    lnb = (0.1 * torch.randn(a.dim, generator=g)).to(dev)
    flush = torch.empty(64 * 1024 * 1024, device=dev)
    for n in a.batches:
        obs = torch.relu(torch.randn(n, a.in_dim, generator=g)).to(dev)   # encoder output: ReLU
        ref_pre = (obs.double() @ w.double().t() + b.double())
        ref = torch.tanh(torch.nn.functional.layer_norm(ref_pre, (a.dim,), lnw.double(),
                                                        lnb.double(), 1e-5))
        ours = lambda: xb.icm_get_rep((w, b, lnw, lnb), obs)
        lib32 = lambda: xrep.ln_tanh_rep(torch.addmm(b, obs, w.t()), lnw, lnb)

        def libtf32():
            torch.backends.cuda.matmul.allow_tf32 = True
            try:
                return xrep.ln_tanh_rep(torch.addmm(b, obs, w.t()), lnw, lnb)
            finally:
                torch.backends.cuda.matmul.allow_tf32 = False
        out = {"n": n, "in_dim": a.in_dim, "dim": a.dim}
        for name, fn in (("tcgen05_3xtf32", ours), ("torch_fp32", lib32), ("torch_tf32", libtf32)):
            rep = fn().rep.double()
            out[name] = {"us": timed(fn, a.iters, flush),
                         "max_abs_err": float((rep - ref).abs().max()),
                         "max_rel_err": float(((rep - ref).abs() / (ref.abs() + 2e-1)).max())}
        flops = 2.0 * n * a.in_dim * a.dim
        out["algorithmic_tflops"] = flops / (out["tcgen05_3xtf32"]["us"] * 1e-6) / 1e12
        out["executed_tf32_tflops"] = 3 * out["algorithmic_tflops"]
        out["tf32_peak_tflops"] = bf16 / 2
        out["tensor_frac_executed"] = out["executed_tf32_tflops"] / (bf16 / 2)
        out["speedup_vs_torch_fp32"] = out["torch_fp32"]["us"] / out["tcgen05_3xtf32"]["us"]
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
