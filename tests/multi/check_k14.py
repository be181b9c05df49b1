"""K14 probe on real peers: a weight matrix partitioned over the ranks' shards, consumed by the GEMM whose TMA
producer fetches each B tile from its owner over NVLink, unicast per CTA (TMA multicast loads read only this
GPU's memory).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tests/multi/check_k14.py [--time]

Each direction prints its result before the next one starts."""
import argparse
import math
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from b200bench import ops as _ops  # noqa: E402
from b200bench.parallel.collectives import DistCtx  # noqa: E402
from b200bench.parallel.peer import PeerComm  # noqa: E402

BF16 = torch.bfloat16


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--M", type=int, default=2048)
    a = ap.parse_args()
    rank, lr, ws = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device(f"cuda:{lr}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    ops = _ops.load()
    comm = PeerComm(DistCtx(rank, ws, "nccl"), dev)
    M, N, K = a.M, 4096, 1024
    torch.manual_seed(1)
    w = (torch.randn(N, K, device=dev) * 0.05).to(BF16)            # same on every rank (same seed)
    shard = (N * K + ws - 1) // ws
    shard = (shard + 64 * K - 1) // (64 * K) * (64 * K)
    pb = comm.alloc("w", shard, BF16)
    flat = torch.zeros(ws * shard, dtype=BF16, device=dev)
    flat[:N * K] = w.view(-1)
    pb.tensor.copy_(flat[rank * shard:(rank + 1) * shard])
    torch.cuda.synchronize(dev)
    dist.barrier()
    if rank == 0:
        print(f"[k14] {ws} ranks, shard {shard * 2 / 2**20:.1f} MiB, allocation "
              f"{'symmetric memory' if pb.handle is not None else 'CUDA IPC'}", flush=True)
    peer = dict(table=pb.table, ws=ws, flat_off=0, shard=shard)
    x = torch.randn(M, K, device=dev).to(BF16)
    dy = torch.randn(M, N, device=dev).to(BF16)
    y = torch.empty(M, N, dtype=BF16, device=dev)
    dx = torch.empty(M, K, dtype=BF16, device=dev)
    ref_y = x.float() @ w.float().t()
    ref_dx = dy.float() @ w.float()
    fails = 0
    for units in (1,):
        for name, fn, out, ref, tol in (
                ("fwd", lambda: ops.gemm(x, None, y, M, N, K, peer=peer), y, ref_y, 0.05 * math.sqrt(K / 64)),
                ("dgrad", lambda: ops.gemm(dy, None, dx, M, K, N, b_mn=True, peer=peer), dx, ref_dx,
                 0.05 * math.sqrt(N / 64))):
            out.zero_()
            fn()
            torch.cuda.synchronize(dev)
            err = float((out.float() - ref).abs().max())
            ok = err < tol + 0.02 * float(ref.abs().max())
            msg = f"[k14] rank {rank} {name}: {'ok' if ok else 'FAIL'} max_err={err:.4f}"
            if a.time:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(20):
                    fn()
                e1.record()
                torch.cuda.synchronize(dev)
                us = e0.elapsed_time(e1) / 20 * 1e3
                remote = N * K * 2 * (ws - 1) / ws
                msg += f"  {us:.1f} us/call, {remote / us / 1e3:.0f} GB/s of remote weight bytes if fetched once"
            print(msg, flush=True)
            fails += 0 if ok else 1
            dist.barrier()
    if a.time:      # local-weight yardstick
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.gemm(x, w, y, M, N, K)
        e0.record()
        for _ in range(20):
            ops.gemm(x, w, y, M, N, K)
        e1.record()
        torch.cuda.synchronize(dev)
        if rank == 0:
            print(f"[k14] same GEMM on a local weight: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us/call", flush=True)
    comm.close()
    dist.destroy_process_group()
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
