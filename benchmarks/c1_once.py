"""Three C1 calls per mode (for an ncu launch list of the per-update path)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exorl_b200 as xb
dev = "cuda:0"
g = torch.Generator().manual_seed(0)
x = torch.tanh(torch.randn(1024, 512, generator=g)).to(dev)
for mode in ("screened", "exact"):
    p = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev); p.mode = mode
    for _ in range(3):
        r = p(x)
    torch.cuda.synchronize()
print("ok", float(r.sum()))
