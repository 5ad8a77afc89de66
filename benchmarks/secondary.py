"""Secondary measurements for BASELINE.json configs 1-4 (the headline config 5 is bench.py).

  C1  PBE 1024 x 512, k=12, avg+rms, clip 0         per-call latency (target < 50 us)
  C2  Proto 1024 z (128) vs 2048 queue, topk 3      per-call latency
  C3  walker-state replay gather B=1024, nstep 3    per-batch latency
  C4  pixel replay gather B=1024 from a 1M-transition uint8 (9,84,84) buffer    GB/s vs HBM peak
Each with the CPU port (oracle) timed beside it on the host cores.  CUDA events, >= 5 warm-ups,
L2 flushed (256 MB write) between timed iterations for the small cases.  One JSON line per case.
"""
import argparse
import json
import os
import statistics
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import exorl_b200 as xb  # noqa: E402
from oracle import RmsState, pbe_reward, proto_reward, nstep_transition  # noqa: E402

DEV = torch.device("cuda:0")
HBM_GBS = 6537.6
try:
    HBM_GBS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def time_gpu(fn, iters=50, warm=10, flush=True):
    # L2 flush = write 256 MB, then read another 256 MB so that the lines left in L2 are clean
    # (otherwise the timed kernel pays the write-back of the flush's own dirty lines)
    scratch = torch.empty(256 << 20, dtype=torch.uint8, device=DEV) if flush else None
    scratch2 = torch.zeros(64 << 20, dtype=torch.int32, device=DEV) if flush else None
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        if flush:
            scratch.zero_()
            scratch2.max()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts), min(ts)


def time_graph(fn, iters=50):
    """Same work replayed from a CUDA graph (launch-bound inner loops)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    return time_gpu(g.replay, iters=iters, warm=5)


def time_cpu(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t0) * 1e6)
    return statistics.median(ts)


def emit(**kw):
    print(json.dumps(kw), flush=True)


def c1():
    g = torch.Generator().manual_seed(0)
    x = torch.tanh(torch.randn(1024, 512, generator=g))
    xd = x.to(DEV)
    cpu_us = time_cpu(lambda: pbe_reward(x, x, RmsState(), 12, True, True, 0.0))
    for mode in ("screened", "exact"):
        p = xb.PBE(xb.RMS(DEV), 0.0, 12, True, True, DEV)
        p.mode = mode
        med, best = time_gpu(lambda: p(xd))
        try:
            gmed, gbest = time_graph(lambda: p(xd))
        except Exception as e:  # graph capture is optional evidence
            gmed = gbest = None
            print("graph capture failed:", e, file=sys.stderr)
        emit(case="C1 PBE 1024x512 k=12 avg+rms", mode=mode, gpu_us_median=med, gpu_us_min=best,
             graph_us_median=gmed, graph_us_min=gbest, cpu_us_median=cpu_us,
             cpu_cores=torch.get_num_threads(), flops=2 * 1024 * 1024 * 512, target_us=50)


def c1b():
    """compute_intr_reward of icm_apt.py:96-100 from observations: trunk (Linear -> LayerNorm ->
    Tanh, 24 -> 512) then PBE.  'torch trunk' = the reference's nn.Sequential on the GPU feeding
    xb.PBE; 'fused producer' = xb.icm_get_rep (one kernel: rep + prepared operand) feeding xb.PBE."""
    g = torch.Generator().manual_seed(0)
    obs = torch.randn(1024, 24, generator=g).to(DEV)
    trunk = torch.nn.Sequential(torch.nn.Linear(24, 512), torch.nn.LayerNorm(512),
                                torch.nn.Tanh()).to(DEV)
    pa = xb.PBE(xb.RMS(DEV), 0.0, 12, True, True, DEV)
    pa.mode = "screened"
    pb = xb.PBE(xb.RMS(DEV), 0.0, 12, True, True, DEV)

    def torch_trunk():
        with torch.no_grad():
            return pa(trunk(obs))

    def fused():
        return pb(xb.icm_get_rep(trunk, obs))

    def producer_only():
        return xb.icm_get_rep(trunk, obs)

    def torch_trunk_only():
        with torch.no_grad():
            return trunk(obs)

    for name, fn in (("torch trunk + PBE", torch_trunk), ("fused producer + PBE", fused),
                     ("fused producer alone", producer_only),
                     ("torch trunk alone", torch_trunk_only)):
        med, best = time_gpu(fn)
        try:
            gmed, gbest = time_graph(fn)
        except Exception as e:
            gmed = gbest = None
            print("graph capture failed:", e, file=sys.stderr)
        emit(case="C1b obs (1024x24) -> ICM trunk -> PBE k=12 avg+rms", path=name,
             gpu_us_median=med, gpu_us_min=best, graph_us_median=gmed, graph_us_min=gbest)


def c2():
    g = torch.Generator().manual_seed(1)
    z = torch.nn.functional.normalize(torch.randn(1024, 128, generator=g), dim=1)
    g = torch.Generator().manual_seed(2)
    q = torch.nn.functional.normalize(torch.randn(2048, 128, generator=g), dim=1)
    zd, qd = z.to(DEV), q.to(DEV)
    cpu_us = time_cpu(lambda: proto_reward(z, q, 3))
    for mode in ("screened", "exact"):
        med, best = time_gpu(lambda: xb.proto_knn_reward(zd, qd, 3, mode=mode))
        emit(case="C2 Proto 1024x2048x128 topk=3", mode=mode, gpu_us_median=med, gpu_us_min=best,
             cpu_us_median=cpu_us, cpu_cores=torch.get_num_threads(), flops=2 * 1024 * 2048 * 128)


def c3():
    rng = np.random.default_rng(3)
    n_eps, L = 1000, 1000
    buf = xb.DeviceReplayBuffer(DEV, 10 ** 6, 3, 0.99, capacity_rows=n_eps * (L + 1))
    eps = []
    for e in range(n_eps):
        ep = {"observation": rng.standard_normal((L + 1, 24)).astype(np.float32),
              "action": rng.uniform(-1, 1, (L + 1, 6)).astype(np.float32),
              "reward": rng.uniform(0, 1, (L + 1, 1)).astype(np.float32),
              "discount": np.ones((L + 1, 1), np.float32)}
        buf.add_episode(ep, key=f"episode_{e:06d}_{L}.npz")
        if e < 8:
            eps.append(ep)
    call = [0]

    def draw_and_gather():
        call[0] += 1
        return buf.sample_philox(1024, 7, call[0])
    med, best = time_gpu(draw_and_gather)

    def draw_then_gather():
        call[0] += 1
        return buf.gather(buf.draw_rows_philox(1024, 7, call[0]))
    med2, best2 = time_gpu(draw_then_gather)
    rows = buf.draw_rows_philox(1024, 7, 0)
    gmed, gbest = time_gpu(lambda: buf.gather(rows))
    # CPU port: the reference's per-sample arithmetic (python loop) for one 1024-batch
    def cpu_batch():
        for s in range(1024):
            nstep_transition(eps[s % 8], 1 + (s * 37) % (L - 3), 3, 0.99)
    cpu_us = time_cpu(cpu_batch, iters=3, warm=1)
    emit(case="C3 walker-state gather B=1024 nstep=3 (1M-transition buffer)",
         sample_us_median=med, sample_us_min=best, two_launch_draw_gather_us_median=med2,
         two_launch_draw_gather_us_min=best2, gather_only_us_median=gmed,
         gather_only_us_min=gbest, algorithmic_bytes=464 * 1024, cpu_us_median=cpu_us, cpu_cores=1,
         note="latency-bound (475 KB per batch); CPU = reference per-sample python arithmetic")


    # BASELINE config 3 end to end: one APT update's data path — sample a 1024-batch from the
    # device buffer, ICM trunk on obs (24 -> 512, fused producer), PBE reward
    trunk = torch.nn.Sequential(torch.nn.Linear(24, 512), torch.nn.LayerNorm(512),
                                torch.nn.Tanh()).to(DEV)
    pbe = xb.PBE(xb.RMS(DEV), 0.0, 12, True, True, DEV)

    def update_path():
        call[0] += 1
        batch = buf.sample_philox(1024, 7, call[0])
        return pbe(xb.icm_get_rep(trunk, batch[0]))
    umed, ubest = time_gpu(update_path)
    try:
        ugmed, ugbest = time_graph(update_path)
    except Exception as e:
        ugmed = ugbest = None
        print("graph capture failed:", e, file=sys.stderr)
    emit(case="C3 end to end: sample B=1024 nstep=3 -> ICM trunk 24->512 -> PBE k=12",
         gpu_us_median=umed, gpu_us_min=ubest, graph_us_median=ugmed, graph_us_min=ugbest,
         launches="replay sample, trunk producer, latency screen, rerank+moments, finish")


def c4(n_trans):
    rows_total = n_trans + n_trans // 1000 + 1
    obs_bytes = 9 * 84 * 84
    L = 1000
    n_eps = n_trans // L
    buf = xb.DeviceReplayBuffer(DEV, n_trans, 3, 0.99, capacity_rows=n_eps * (L + 1))
    # synthetic pixel buffer written straight into the device columns (headless)
    buf.cols["observation"] = torch.empty((n_eps * (L + 1), 9, 84, 84), dtype=torch.uint8, device=DEV)
    g = torch.Generator(device=DEV).manual_seed(4)
    chunk = 20000
    for r0 in range(0, n_eps * (L + 1), chunk):
        r1 = min(r0 + chunk, n_eps * (L + 1))
        buf.cols["observation"][r0:r1] = torch.randint(0, 256, (r1 - r0, 9, 84, 84), generator=g,
                                                       dtype=torch.uint8, device=DEV)
    buf.cols["action"] = torch.rand((n_eps * (L + 1), 6), generator=g, device=DEV) * 2 - 1
    buf.cols["reward"] = torch.rand((n_eps * (L + 1), 1), generator=g, device=DEV)
    buf.cols["discount"] = torch.ones((n_eps * (L + 1), 1), device=DEV)
    buf.episodes = [{"key": f"episode_{e:06d}_{L}.npz", "start": e * (L + 1), "len": L}
                    for e in range(n_eps)]
    buf.size = n_eps * L
    buf._tables_dirty = True
    for batch in (1024, 256):
        call = [0]

        def draw_and_gather():
            call[0] += 1
            return buf.sample_philox(batch, 11, call[0])
        med, best = time_gpu(draw_and_gather, iters=30, warm=5, flush=True)
        alg = (2 * obs_bytes + 24 + 3 * 8 + 2 * obs_bytes + 24 + 8) * batch
        emit(case=f"C4 pixel gather B={batch} nstep=3 from {n_trans} uint8(9,84,84) transitions",
             us_median=med, us_min=best, algorithmic_bytes=alg,
             achieved_gbs=alg / (med * 1e-6) / 1e9, hbm_peak_gbs=HBM_GBS,
             frac_of_hbm_peak=alg / (med * 1e-6) / 1e9 / HBM_GBS,
             buffer_gb=n_eps * (L + 1) * obs_bytes / 1e9,
             note="one launch (fused draw + gather); fresh random rows every iteration; a 256 MB L2 flush runs "
                  "before every timed iteration (it also lets the host run ahead, so the figure "
                  "is device time, not python launch overhead)")
    # correctness spot check against the oracle on the last batch
    rows = buf.draw_rows_philox(64, 11, 999)
    out = buf.gather(rows)
    obs_cpu = buf.cols["observation"]
    r = rows.cpu().numpy()
    for s in range(0, 64, 7):
        assert torch.equal(out[0][s], obs_cpu[r[s] - 1]) and torch.equal(out[4][s], obs_cpu[r[s] + 2])


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="c1,c2,c3,c4")
    ap.add_argument("--c4-transitions", type=int, default=1_000_000)
    a = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    for c in a.cases.split(","):
        {"c1": c1, "c1b": c1b, "c2": c2, "c3": c3, "c4": lambda: c4(a.c4_transitions)}[c]()
