"""Multi-GPU checks for the peer-memory runtime and the fused collectives.  Run under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/multi/check_comm.py [--bench]

Every check compares a fused kernel against the same arithmetic done with NCCL collectives and the
single-GPU kernels; timings are CUDA-event, max over ranks.
"""
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


def timed(fn, iters, ws, dev):
    torch.cuda.synchronize(dev); dist.barrier(); torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bench", action="store_true")
    a = ap.parse_args()
    rank, lr, ws = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device(f"cuda:{lr}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    ctx = DistCtx(rank, ws, "nccl")
    ops = _ops.load()
    comm = PeerComm(ctx, dev)
    fails = 0

    def report(name, ok, extra=""):
        nonlocal fails
        flag = torch.tensor([0 if ok else 1], device=dev)
        dist.all_reduce(flag)
        if rank == 0:
            print(f"[check] {name}: {'ok' if flag.item() == 0 else 'FAIL'} {extra}", flush=True)
        fails += int(flag.item() != 0)

    # ---- K13 all-reduce ------------------------------------------------------------------------
    n = 8 * 1024 * 1024 + 128 * 3
    gb = comm.alloc("g", n, torch.bfloat16)
    if rank == 0:
        print(f"[check] allocation: {'symmetric memory' if gb.handle is not None else 'cudaMalloc + CUDA IPC'}, "
              f"multicast address {'yes (NVLS kernels)' if gb.mc_ptr else 'no (peer load/store kernels)'}", flush=True)
    torch.manual_seed(100 + rank)
    src = (torch.randn(n, device=dev) * 0.5).to(torch.bfloat16)
    gb.tensor.copy_(src)
    ref = src.float().clone()
    dist.all_reduce(ref)
    ref /= ws
    off, cnt = 128, n - 256
    nvls_default = comm.use_nvls
    for mode in (["p2p", "nvls"] if gb.mc_ptr else ["p2p"]):       # both data paths, whatever the default is
        comm.use_nvls = mode == "nvls"
        gb.tensor.copy_(src)
        comm.allreduce(gb, off, cnt, 1.0 / ws, side_stream=False)
        torch.cuda.synchronize(dev)
        got = gb.tensor.float()
        err = (got[off:off + cnt] - ref[off:off + cnt]).abs().max().item()
        untouched = torch.equal(gb.tensor[:off], src[:off]) and torch.equal(gb.tensor[off + cnt:], src[off + cnt:])
        report(f"K13 allreduce_scale ({mode})", err < 0.02 and untouched, f"max_err={err:.4f}")
    comm.use_nvls = nvls_default
    # second launch (epoch advance) on the side stream with fork/join
    gb.tensor.copy_(src)
    comm.fork(); comm.allreduce(gb, 0, n, 1.0 / ws); comm.join()
    torch.cuda.synchronize(dev)
    err = (gb.tensor.float() - ref).abs().max().item()
    report("K13 second launch (side stream)", err < 0.02, f"max_err={err:.4f}")

    # ---- K15 reduce-scatter + AdamW + push ---------------------------------------------------------
    for clip in (0.0, 0.5):
        wb = comm.alloc(f"w{clip}", n, torch.bfloat16)
        shard = n // ws // 8 * 8
        lo = rank * shard
        torch.manual_seed(7)
        w0 = torch.randn(n, device=dev)
        wb.tensor.copy_(w0.to(torch.bfloat16))
        gb.tensor.copy_(src)
        master = w0[lo:lo + shard].clone(); m = torch.zeros(shard, device=dev); v = torch.zeros(shard, device=dev)
        gsh = torch.zeros(shard, dtype=torch.bfloat16, device=dev)
        hp = torch.zeros(16, device=dev)
        ops.adamw_set_hp(hp, 1e-2, 0.9, 0.999, 1e-8, 0.01, 1 / (1 - 0.9), 1 / math.sqrt(1 - 0.999), 1.0 / ws, clip)
        gn = torch.zeros(2, device=dev)      # K15 writes the norm and its square
        comm.rs_adamw(gb, wb, lo, shard, master, m, v, gsh, hp, push_all=True, gnorm_out=gn)
        torch.cuda.synchronize(dev)
        # reference: NCCL all-reduce, then the single-GPU AdamW kernel on the shard
        gsum = src.float().clone(); dist.all_reduce(gsum)
        gmean_bf = (gsum[lo:lo + shard]).to(torch.bfloat16)      # kernel keeps the reduced grad in bf16 when clipping
        rm = w0[lo:lo + shard].clone(); rmm = torch.zeros(shard, device=dev); rv = torch.zeros(shard, device=dev)
        rw = torch.empty(shard, dtype=torch.bfloat16, device=dev)
        ss = torch.zeros(1, device=dev)
        if clip > 0:
            used = shard * ws
            ss.fill_(float(((gsum[:used] / ws) ** 2).sum()))
        ops.adamw_flat(rm, rmm, rv, gmean_bf if clip > 0 else gsum[lo:lo + shard].to(torch.bfloat16), rw, hp,
                       ss if clip > 0 else None)
        torch.cuda.synchronize(dev)
        e_master = (master - rm).abs().max().item()
        full = torch.empty(shard * ws, dtype=torch.bfloat16, device=dev)
        dist.all_gather_into_tensor(full, rw)
        e_push = (wb.tensor[:shard * ws].float() - full.float()).abs().max().item()
        extra = f"master_err={e_master:.2e} push_err={e_push:.2e}"
        if clip > 0:
            extra += f" norm={gn[0].item():.3f} ref={math.sqrt(ss.item()):.3f}"
        report(f"K15 rs_adamw clip={clip}", e_master < 2e-3 and e_push < 2e-2, extra)

    # ---- K16 all-gather as one collective: both data paths, two launches each (epoch advance) ---------------------
    shard = n // ws // 8 * 8
    torch.manual_seed(11)
    wfull = torch.randn(shard * ws, device=dev).to(torch.bfloat16)          # same on every rank
    ws_pb = comm.alloc("k16_shard", shard, torch.bfloat16)
    ws_pb.tensor.copy_(wfull[rank * shard:(rank + 1) * shard])
    st_pb = comm.alloc("k16_stage", shard * ws + 256, torch.bfloat16)
    tab = comm.adjusted_table(ws_pb, shard)
    torch.cuda.synchronize(dev); dist.barrier()
    for mc in ((1, 0) if st_pb.mc_ptr else (0,)):
        os.environ["B200BENCH_GATHER_MC"] = str(mc)
        okk = True
        for g_off, g_n, d_off in ((0, shard * ws, 128), (shard - 8 * 40, 8 * 1000, 8)):
            st_pb.tensor.fill_(-1.0)
            torch.cuda.synchronize(dev); dist.barrier()
            comm.gather_bcast(tab, st_pb, d_off, g_off, g_n, shard)
            torch.cuda.synchronize(dev); dist.barrier()
            t = st_pb.tensor
            okk = okk and torch.equal(t[d_off:d_off + g_n], wfull[g_off:g_off + g_n]) \
                and bool((t[:d_off] == -1).all()) and bool((t[d_off + g_n:] == -1).all())
        report(f"K16 gather_bcast ({'multicast' if mc else 'peer stores'})", okk)
    os.environ.pop("B200BENCH_GATHER_MC", None)

    # ---- engine: fused vs NCCL schedule, same data, dropout 0 ----------------------------------------
    from b200bench.config import StrategyConfig, get_model_config, load_deepspeed_config
    from b200bench.engine import NativeEngine

    # fsdp / zero3 are checked with both weight all-gather schedules (inside the GEMM, and per block on the
    # comm stream); each fused run is compared with the NCCL schedule of the same strategy
    # the last case accumulates over 2 micro-steps: the second one reuses the blocks gathered by the first
    # "stream+pull" / "stream+k16": the block gathers as plain peer pulls or as the K16 collective (both forced, whatever
    # the default for this world size is)
    for strat, pg, accum in (("ddp", None, 1), ("zero2", None, 1), ("fsdp", "stream+pull", 1), ("fsdp", "stream+k16", 1),
                             ("fsdp", "gemm", 1), ("zero3", "stream+pull", 1), ("zero3", "stream+k16", 1),
                             ("zero3", "gemm", 1), ("zero3", "stream+k16", 2)):
        cfgm = get_model_config("T", 128)
        label = pg
        if pg and "+" in pg:
            os.environ["B200BENCH_GATHER_COLL"] = "1" if pg.endswith("k16") else "0"
            pg = "stream"
        else:
            os.environ.pop("B200BENCH_GATHER_COLL", None)
        sds = []
        for mode in ("fused", "nccl"):
            if strat in ("zero2", "zero3"):
                sc = load_deepspeed_config(os.path.join(ROOT, f"configs/deepspeed/{strat}.json"), 2, accum, ws)
            else:
                sc = StrategyConfig()
                sc.fsdp_reshard_after_forward = True      # two staging slots: blocks are gathered again in backward
            eng = NativeEngine(cfgm, strat, sc, ctx, dev, dropout=0.0, batch=2, seq_len=128, comm=mode,
                               use_graph=(mode == "fused"), param_gather=pg, grad_accum=accum)
            g = torch.Generator(device=dev); g.manual_seed(5 + rank)
            for _ in range(5 * accum + (accum - 1) * 3):
                tok = torch.randint(0, 512, (2, 128), device=dev, dtype=torch.int32, generator=g)
                loss = eng.train_step(tok, tok)
            torch.cuda.synchronize(dev)
            sds.append((eng.state_dict(), float(loss.item())))
            eng.close()
        worst = max((sds[0][0][k] - sds[1][0][k]).abs().max().item() for k in sds[0][0])
        # replicas must agree across ranks as well
        w = sds[0][0]["transformer.h.0.mlp.0.weight"].clone()
        w2 = w.clone(); dist.broadcast(w2, 0)
        report(f"engine {strat}{'/' + label if label else ''}{' accum ' + str(accum) if accum > 1 else ''}: fused == nccl", worst < 5e-3 and torch.equal(w, w2),
               f"max_param_diff={worst:.2e} loss={sds[0][1]:.4f}/{sds[1][1]:.4f}")

    os.environ.pop("B200BENCH_GATHER_COLL", None)

    # ---- bandwidth / latency probes -----------------------------------------------------------------
    if a.bench:
        from b200bench.utils.clocks import ClockSampler
        sampler = ClockSampler(lr).start()            # SM clocks / throttle reasons while the probes below run
        nbytes = 256 * 1024 * 1024
        big = comm.alloc("big", nbytes // 2, torch.bfloat16)
        dst = torch.empty(nbytes // 2, dtype=torch.bfloat16, device=dev)
        peer = (rank + 1) % ws
        for ctas in (32, 64, 128):
            ms = timed(lambda: comm.peer_copy(big.peer_ptrs[peer], dst, nbytes, ctas), 10, ws, dev)
            if rank == 0:
                print(f"[bench] P2P read {nbytes/2**20:.0f} MiB, {ctas} CTAs: {ms*1e3:.0f} us  {nbytes/ms/1e6:.0f} GB/s", flush=True)
        for mb in (1, 8, 25, 64, 256):
            cnt = mb * 1024 * 1024 // 2
            t = big.tensor[:cnt]
            ms_n = timed(lambda: dist.all_reduce(t), 10, ws, dev)
            for ctas in (8, 16, 32, 64, 128):
                comm.comm_ctas = comm.nvls_ctas = ctas
                res = {}
                for mode in ("nvls", "p2p"):
                    if mode == "nvls" and not big.mc_ptr:
                        continue
                    comm.use_nvls = mode == "nvls"
                    res[mode] = timed(lambda: comm.allreduce(big, 0, cnt, 1.0 / ws, side_stream=False), 10, ws, dev)
                comm.use_nvls = nvls_default
                if rank == 0:
                    txt = "  ".join(f"{k} {v*1e3:8.1f} us (busbw {2 * (ws - 1) / ws * cnt * 2 / v / 1e6:5.0f} GB/s)"
                                    for k, v in res.items())
                    print(f"[bench] allreduce {mb:4d} MiB ctas={ctas:3d}: {txt}   nccl {ms_n*1e3:8.1f} us", flush=True)
        # K15 on a tier-A-sized buffer (236.4 M elements): reduce-scatter + AdamW (+ push to every peer)
        nk = 236_400_000 // (8 * ws) * (8 * ws)
        gk = comm.alloc("gk", nk, torch.bfloat16)
        wk = comm.alloc("wk", nk, torch.bfloat16)
        shard = nk // ws
        master = torch.zeros(shard, device=dev); m1 = torch.zeros(shard, device=dev); m2 = torch.zeros(shard, device=dev)
        gsh = torch.zeros(shard, dtype=torch.bfloat16, device=dev)
        hp = torch.zeros(16, device=dev)
        ops.adamw_set_hp(hp, 1e-4, 0.9, 0.999, 1e-8, 0.01, 10.0, 31.6, 1.0 / ws, 0.0)
        for push in (True, False):
            ms = timed(lambda: comm.rs_adamw(gk, wk, rank * shard, shard, master, m1, m2, gsh, hp, push_all=push),
                       5, ws, dev)
            nvl = (ws - 1) * shard * 2 * (2 if push else 1)        # bytes pulled (+ pushed) over NVLink per GPU
            hbm = shard * 28 + ws * shard * 2                       # optimizer traffic + local reads of the ws slices
            if rank == 0:
                print(f"[bench] K15 rs_adamw push_all={int(push)} {nk/1e6:.1f} M elements: {ms*1e3:7.1f} us  "
                      f"NVLink {nvl/ms/1e6:5.0f} GB/s per direction-equivalent, roofline max(NVLink {nvl/770e9*1e6:.0f} us, "
                      f"HBM {hbm/6454.6e9*1e6:.0f} us)", flush=True)
        # block gather (the fsdp / zero3 weight all-gather): 25 MB = one tier-A block, 403 MB = one 7B-shaped block
        table = comm.adjusted_table(wk, shard)
        for mb in (25, 403):
            cnt = min(mb * 1000 * 1000 // 2 // 8 * 8, nk)
            dstg = torch.empty(cnt, dtype=torch.bfloat16, device=dev)
            for ctas in (32, 64, 128):
                ms = timed(lambda: comm.gather_range(table, dstg, 0, cnt, shard, ctas=ctas), 10, ws, dev)
                if rank == 0:
                    print(f"[bench] gather_range {cnt*2/1e6:6.1f} MB ctas={ctas:3d}: {ms*1e3:7.1f} us  "
                          f"{cnt*2/ms/1e6:5.0f} GB/s", flush=True)
            # K16: the same range as one collective (1/ws fetched per rank, written to every rank through the switch
            # or by peer stores) into a symmetric destination
            for mc in ((1, 0) if gk.mc_ptr else (0,)):
                os.environ["B200BENCH_GATHER_MC"] = str(mc)
                for ctas in (8, 16, 32, 64):
                    ms = timed(lambda: comm.gather_bcast(table, gk, 0, 0, cnt, shard, ctas=ctas), 10, ws, dev)
                    if rank == 0:
                        print(f"[bench] K16 gather_bcast {cnt*2/1e6:6.1f} MB {'multicast' if mc else 'peer st  '} "
                              f"ctas={ctas:3d}: {ms*1e3:7.1f} us  {cnt*2/ms/1e6:5.0f} GB/s", flush=True)
            os.environ.pop("B200BENCH_GATHER_MC", None)
        ck = sampler.stop()
        allck = [None] * ws
        dist.all_gather_object(allck, {k: ck.get(k) for k in ("sm_mhz", "sm_max_mhz", "reasons")})
        if rank == 0:
            print(f"[clocks] bandwidth probes, per rank: {allck}", flush=True)
    comm.close()
    dist.barrier()
    if rank == 0:
        print(f"[check] done, {fails} failure(s)", flush=True)
    dist.destroy_process_group()
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
