"""Three C1 / C2 calls per mode (for an ncu launch list of the per-update path) + screen status."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exorl_b200 as xb
from exorl_b200 import knn as xknn
dev = "cuda:0"
g = torch.Generator().manual_seed(0)
x = torch.tanh(torch.randn(1024, 512, generator=g)).to(dev)
for mode in ("screened", "exact"):
    p = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev); p.mode = mode
    for _ in range(3):
        r = p(x)
    torch.cuda.synchronize()
    if mode == "screened":
        print("C1 screened status (undecided, invalid, pairs, _):", xknn.screened_status(torch.device(dev), key="apt_reward"))
z = torch.nn.functional.normalize(torch.randn(1024, 128, generator=g), dim=1).to(dev)
q = torch.nn.functional.normalize(torch.randn(2048, 128, generator=g), dim=1).to(dev)
for _ in range(3):
    r2 = xb.proto_knn_reward(z, q, 3, mode="screened")
torch.cuda.synchronize()
print("C2 screened status:", xknn.screened_status(torch.device(dev)))
print("ok", float(r.sum()), float(r2.sum()))
if os.environ.get("EXB_SCREEN_TRACE"):
    import ctypes
    from exorl_b200 import _lib
    lib = _lib.load()
    pf = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev); pf.mode = "screened"
    for name, key, fn in (("C1 knn", "knn_screened", lambda: xb.knn_topk(x, x, 12, mode="screened")),
                          ("C1 fused PBE", "apt_reward", lambda: pf(x)),
                          ("C2 knn", "knn_screened", lambda: xb.knn_topk(z, q, 3, mode="screened"))):
        for _ in range(300):
            fn()
        torch.cuda.synchronize()
        ws, _ = _lib.Workspace.get(torch.device(dev), key, 1024)
        out = (ctypes.c_int64 * 16)()
        _lib.check(lib.exb_debug_trace_sync(_lib.ptr(ws), out, _lib.stream_ptr(torch.device(dev))))
        t = list(out)
        print(name, "trace ns rel. screen start:", [v - t[0] if v else None for v in t[:8]],
              "rerank:", [v - t[8] if v else None for v in t[8:12]], "gap:", t[8] - t[6])
