"""Phase timeline of the fused per-update chain kernel (csrc/knn_chain.cuh): globaltimer stamps of
block 0 and of the last block — start, prep done, barrier 1 passed, screen done, barrier 2 passed,
rerank done, barrier 3 passed, end — for C1 (PBE 1024 x 512), hot and with an L2 flush in front."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exorl_b200 as xb  # noqa: E402
from exorl_b200 import _lib  # noqa: E402

lib = _lib.load()
lib.exb_debug_knob_set(b"EXB_CHAIN_FUSED", 1)
lib.exb_debug_knob_set(b"EXB_CHAIN_TRACE", 1)
for kv in sys.argv[1:]:
    name, val = kv.split("=")
    lib.exb_debug_knob_set(name.encode(), int(val))
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
x = torch.tanh(torch.randn(1024, 512, generator=g)).to(dev)
pbe = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev)
pbe.mode = "screened"
flush = torch.empty(64 * 1024 * 1024, device=dev)
names = ["prep", "bar1", "screen", "bar2", "rerank", "bar3", "finish"]
for cold in (False, True):
    rows = []
    for it in range(40):
        if cold:
            flush.fill_(1.0)
            flush.sum()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        pbe(x)
        e1.record()
        torch.cuda.synchronize()
        ws, _ = _lib.Workspace.get(dev, "apt_reward", 1024)
        out = (ctypes.c_int64 * 16)()
        _lib.check(lib.exb_debug_trace_sync(_lib.ptr(ws), out, _lib.stream_ptr(dev)))
        t = list(out)
        rows.append((e0.elapsed_time(e1) * 1e3, t))
    rows = rows[10:]
    rows.sort(key=lambda r: r[0])
    us, t = rows[len(rows) // 2]
    first = [(t[i + 1] - t[i]) / 1e3 for i in range(7)]
    last = [(t[8 + i + 1] - t[8 + i]) / 1e3 for i in range(7)]
    print("L2 flushed" if cold else "hot loop", "event us %.1f" % us,
          "| block 0:", {n: round(v, 2) for n, v in zip(names, first)},
          "total %.2f" % ((t[7] - t[0]) / 1e3),
          "| last block:", {n: round(v, 2) for n, v in zip(names, last)},
          "start skew %.2f" % ((t[8] - t[0]) / 1e3))
