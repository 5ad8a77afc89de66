"""Multi-GPU parity of the sharded relabel (run under torchrun, one rank per GPU, NCCL):
the bank-sharded result (both exchanges) must equal the single-shard result computed on rank 0's
GPU from the same reps: identical index sets outside fp32 near-ties, rewards rel 1e-5, same RMS.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29511 benchmarks/relabel_nccl_check.py [--rows 20000] [--dim 512]
"""
import argparse
import os
import sys
import types

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exorl_b200 as xb  # noqa: E402
from exorl_b200.knn import knn_topk, _reward_from_knn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", dest="n", type=int, default=20000)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--k", type=int, default=12)
    ap.add_argument("--sym-min-rows", type=int, default=1024,
                    help="knn.SYM_MIN_ROWS for the 'sym' legs (the symmetric screen over the ranks)")
    ap.add_argument("--allow-a2a", action="store_true",
                    help="do not insist that the 'sym' legs used the peer-memory exchange")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    g = torch.Generator().manual_seed(5)
    rep = torch.tanh(torch.randn(a.n, a.dim, generator=g)).to(dev)
    args = types.SimpleNamespace(knn_k=a.k, knn_avg=True, knn_rms=True, knn_clip=0.0)

    # single-shard truth on this rank's GPU (no collectives involved; full screen)
    from exorl_b200 import knn as xknn
    xknn.SYM_ENABLED = False
    d0, i0 = knn_topk(rep, rep, a.k, mode="screened")
    rms0 = xb.RMS(dev)
    r0 = _reward_from_knn(d0, a.k, True, True, 0.0, rms0.M, rms0.S, rms0._n)

    ok = True
    # "prepared": the same set handed over as a PreparedRep (exb_knn_prep_f16 here; the rep
    # producers emit the same operand), query-block partition
    rep_host = rep.cpu().pin_memory()
    from exorl_b200 import relabel as xrel
    # "balanced" runs three times so that the rate-proportional query blocks (QueryBalance) have
    # moved away from the equal split by the last run; "from_host" is the end-to-end pipeline
    # "sym": the symmetric screen over the ranks with the hit exchange inside the kernel (NVLink
    # peer memory), twice (buffers reused); "sym_a2a": the same with the all-to-all exchange
    # "sym_from_host": the end-to-end form — rows uploaded in chunks behind the sample pre-pass
    for exchange in ("sym", "sym", "sym_a2a", "sym_from_host", "sym_from_host", "queries", "balanced", "balanced", "balanced", "from_host",
                     "prepared", "lists", "auto", "all_to_all", "all_gather"):
        rms = xb.RMS(dev)
        xknn.SYM_ENABLED = exchange in ("sym", "sym_a2a", "sym_from_host")
        xknn.SYM_MIN_ROWS = a.sym_min_rows
        if exchange in ("sym", "sym_a2a"):
            xknn.sym_last.clear()
            xrel.SYM_PEER_MEMORY = exchange == "sym"
            out = xb.relabel_apt_reward(rep, args, rms=rms, exchange="queries", mode="screened",
                                        return_knn=True)
            xrel.SYM_PEER_MEMORY = True
            assert xknn.sym_last.get("world") == world, xknn.sym_last
            assert "fell_back_to_full_screen" not in xknn.sym_last, xknn.sym_last
            want = "peer_memory" if exchange == "sym" and not a.allow_a2a else None
            assert want is None or xknn.sym_last.get("exchange") == want, xknn.sym_last
            if rank == 0:
                print(f"[nccl check] symmetric exchange: {xknn.sym_last.get('exchange')}",
                      flush=True)
        elif exchange == "sym_from_host":
            xknn.sym_last.clear()
            scratch = torch.zeros_like(rep)            # the rows must really come from the host
            out = xb.relabel_apt_reward_from_host(rep_host, args, rms=rms, rep_out=scratch,
                                                  return_knn=True)
            assert xknn.sym_last.get("world") == world, xknn.sym_last
            assert "fell_back_to_full_screen" not in xknn.sym_last, xknn.sym_last
            assert torch.equal(scratch, rep)
            del scratch
        elif exchange == "from_host":
            out = xb.relabel_apt_reward_from_host(rep_host, args, rms=rms, return_knn=True)
        elif exchange == "prepared":
            out = xb.relabel_apt_reward(xb.prepare_rep(rep), args, rms=rms, exchange="queries",
                                        return_knn=True)
        elif exchange == "balanced":
            out = xb.relabel_apt_reward(rep, args, rms=rms, exchange="queries", mode="screened",
                                        return_knn=True, balance=True)
        else:
            out = xb.relabel_apt_reward(rep, args, rms=rms, exchange=exchange, mode="screened",
                                        return_knn=True)
        r, d, i = out
        torch.cuda.synchronize()
        np.testing.assert_allclose(r.cpu().numpy(), r0.cpu().numpy(), rtol=1e-5, atol=1e-7)
        assert abs(float(rms.M) - float(rms0.M)) <= 1e-6 * abs(float(rms0.M))
        assert rms.n == rms0.n
        if exchange == "all_gather":            # every rank holds all rows
            lo, hi = 0, a.n
        elif exchange in ("sym", "sym_a2a", "sym_from_host", "queries", "balanced", "auto", "prepared", "from_host"):
            lo, hi = xrel.last_query_block["lo"], xrel.last_query_block["hi"]   # balanced block
        else:                                    # rank r finalised query block r
            lo, hi = xrel.shard_bounds(a.n, world, rank)
        np.testing.assert_allclose(d.cpu().numpy(), d0[lo:hi].cpu().numpy(), rtol=1e-6)
        same = (i.sort(dim=1).values == i0[lo:hi].sort(dim=1).values).all(dim=1)
        bad = int((~same).sum())
        assert bad <= max(1, (hi - lo) // 1000), f"{bad} rows differ"   # fp32 near-ties only
        # identical on every rank
        chk = r.double().sum().reshape(1)
        lst = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(lst, chk)
        assert all(float(x) == float(lst[0]) for x in lst), "ranks disagree"
        if rank == 0:
            print(f"[nccl check] world={world} n={a.n} exchange={exchange}: OK "
                  f"(rows with near-tie swaps: {bad}; rank-0 block [{lo}, {hi}))", flush=True)
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
