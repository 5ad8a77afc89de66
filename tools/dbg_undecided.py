import os, sys, ctypes, torch
sys.path.insert(0, os.getcwd())
import exorl_b200 as xb
from exorl_b200 import _lib
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(5)
rep = torch.tanh(torch.randn(1_000_000, 512, generator=g, device=dev))
lib = _lib.load()
d, i = xb.knn_topk(rep, rep, 12, mode="screened")
torch.cuda.synchronize()
ws, _ = _lib.Workspace.get(dev, "knn_screened", 1024)
out = (ctypes.c_int64 * 16)()
_lib.check(lib.exb_debug_trace_sync(_lib.ptr(ws), out, _lib.stream_ptr(dev)))
t = list(out)
print("undecided row", t[12], "why", t[13], "total", t[14], "n_surv", t[15])
from exorl_b200 import knn as xknn
print(xknn.screened_status(dev))
row = int(t[12])
print("self idx in result:", (i[row] == row).any().item(), d[row][:4].tolist())
