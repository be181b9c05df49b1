"""Peer-memory communication runtime: symmetric buffers mapped into every rank, flag words, and
the Python front-end of the fused collective kernels in ``csrc/comm.cu``.

The reference's whole communication layer is ``init_process_group('nccl', tcp://...)`` plus what
DDP / FSDP / DeepSpeed do internally (``benchmarking/train_harness.py:186-204,217-271``; SURVEY.md
§2.7 N1-N10, §5.8).  Here ``torch.distributed`` is kept for rendezvous only (N1): it carries the
64-byte CUDA IPC handles of each rank's buffers.  After that every rank holds device pointers to
every peer's gradient buffer, weight buffer and flag words, and the collectives are single kernels
that read and write those pointers over NVLink (K13 all-reduce+scale, K15 reduce-scatter+AdamW+push).

Allocation: first choice is torch's symmetric memory (``cuMemCreate`` + POSIX file-descriptor
exchange + an NVSwitch multicast object bound to every rank's copy, ``torch.distributed._symmetric_memory``), which
gives the peer pointers AND a multicast address per buffer - the address the NVLS kernels (``multimem.ld_reduce`` /
``multimem.st``) work on.  If that is not available (no multicast support, a container without the needed
syscalls, ``B200BENCH_SYMM_MEM=0``) buffers come from ``cudaMalloc`` inside the native library and are exchanged as
CUDA IPC handles; there is then no multicast address and the collectives use plain peer loads / stores.
"""
from __future__ import annotations

import ctypes
import os
import sys
from ctypes import c_float, c_int, c_size_t, c_void_p
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .. import ops as _ops
from .collectives import DistCtx


class _RawCuda:
    """Minimal ``__cuda_array_interface__`` holder so torch can view raw device memory."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


class PeerBuffer:
    def __init__(self, name: str, local_ptr: int, nbytes: int, peer_ptrs: List[int], tensor: torch.Tensor,
                 mc_ptr: int = 0, handle=None):
        self.name, self.local_ptr, self.nbytes, self.peer_ptrs, self.tensor = name, local_ptr, nbytes, peer_ptrs, tensor
        self.table = (c_void_p * len(peer_ptrs))(*peer_ptrs)      # host array of device pointers
        self.mc_ptr = int(mc_ptr or 0)        # multicast (NVLS) address of the buffer, 0 = none
        self.handle = handle                  # symmetric-memory handle (keeps the mapping alive), None = CUDA IPC


class PeerComm:
    def __init__(self, ctx: DistCtx, device: torch.device, comm_ctas: int = 64):
        if not ctx.active:
            raise RuntimeError("PeerComm needs an initialised process group (world_size > 1)")
        self.ctx, self.device = ctx, device
        self.rank, self.ws = ctx.rank, ctx.world_size
        if self.ws not in (2, 4, 8):
            raise NotImplementedError("fused collectives are instantiated for world sizes 2, 4 and 8")
        self.ops = _ops.load()
        L = self.ops.lib
        vp, i, f, sz = c_void_p, c_int, c_float, c_size_t
        L.b200_peer_alloc.argtypes = [sz, ctypes.POINTER(vp), vp]
        L.b200_peer_open.argtypes = [vp, ctypes.POINTER(vp)]
        L.b200_peer_close.argtypes = [vp]
        L.b200_peer_free.argtypes = [vp]
        L.b200_comm_flag_bytes.restype = sz
        L.b200_comm_allreduce.argtypes = [vp, vp, i, i, sz, sz, f, i, vp]
        L.b200_comm_rs_adamw.argtypes = [vp, vp, vp, i, i, sz, sz, vp, vp, vp, vp, vp, i, vp, vp, i, vp]
        L.b200_comm_gather_range.argtypes = [vp, i, vp, ctypes.c_ulonglong, ctypes.c_ulonglong, ctypes.c_ulonglong, i, vp]
        L.b200_comm_gather_range_from.argtypes = [vp, i, i, vp, ctypes.c_ulonglong, ctypes.c_ulonglong,
                                                  ctypes.c_ulonglong, i, vp]
        ull = ctypes.c_ulonglong
        L.b200_comm_gather_bcast.argtypes = [vp, vp, vp, vp, i, i, ull, ull, ull, ull, i, vp]
        L.b200_comm_allreduce_nvls.argtypes = [vp, vp, i, i, sz, sz, f, i, vp]
        L.b200_comm_rs_adamw_nvls.argtypes = [vp, vp, vp, vp, i, i, sz, sz, vp, vp, vp, vp, vp, i, vp, vp, i, vp]
        L.b200_peer_copy.argtypes = [vp, vp, sz, i, vp]
        self.buffers: Dict[str, PeerBuffer] = {}
        self.comm_ctas = comm_ctas
        self.use_symm = os.environ.get("B200BENCH_SYMM_MEM", "1") != "0"
        # NVLS (in-switch reduction / broadcast): from 4 ranks on by default.  At 2 ranks there is nothing to save
        # (each rank pulls half of the bucket from its one peer either way) and the peer load/store kernel measured
        # faster (profiles/round2/README.md).  B200BENCH_NVLS=1 / 0 forces it on / off.
        nv = os.environ.get("B200BENCH_NVLS", "")
        self.use_nvls = (self.ws >= 4) if nv == "" else (nv != "0")
        self.nvls_ctas = int(os.environ.get("B200BENCH_NVLS_CTAS", "0")) or 16
        # very large buffers (the 13 GB shards / gradients of the 7B shape) stay on cudaMalloc + IPC
        self.symm_max_bytes = int(os.environ.get("B200BENCH_SYMM_MAX_BYTES", str(4 << 30)))
        self.flags = self.alloc("flags", int(L.b200_comm_flag_bytes()), torch.uint8)
        self.stream = torch.cuda.Stream(device=device)
        self._ev_ready = torch.cuda.Event()
        self._ev_done = torch.cuda.Event()
        self._forked = False
        self.bytes_allocated = 0

    # ---- symmetric allocation -------------------------------------------------------------------
    def alloc(self, name: str, numel: int, dtype: torch.dtype) -> PeerBuffer:
        """Every rank must call this with the same arguments, in the same order."""
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        nbytes = (nbytes + 255) // 256 * 256
        if self.use_symm and nbytes <= self.symm_max_bytes:
            pb = self._alloc_symm(name, numel, nbytes, dtype)
            if pb is not None:
                return pb
        ptr = c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        rc = self.ops.lib.b200_peer_alloc(nbytes, ctypes.byref(ptr), ctypes.cast(handle, c_void_p))
        if rc:
            raise _ops.NativeError(f"peer_alloc({name}, {nbytes}) failed with CUDA error {rc}")
        handles: List[Optional[bytes]] = [None] * self.ws
        dist.all_gather_object(handles, bytes(handle))
        peer_ptrs: List[int] = []
        for r, h in enumerate(handles):
            if r == self.rank:
                peer_ptrs.append(ptr.value)
                continue
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            pp = c_void_p()
            rc = self.ops.lib.b200_peer_open(ctypes.cast(buf, c_void_p), ctypes.byref(pp))
            if rc:
                raise _ops.NativeError(f"peer_open({name}, rank {r}) failed with CUDA error {rc}")
            peer_ptrs.append(pp.value)
        t = torch.as_tensor(_RawCuda(ptr.value, nbytes), device=self.device).view(dtype)[:numel]
        pb = PeerBuffer(name, ptr.value, nbytes, peer_ptrs, t)
        self.buffers[name] = pb
        self.bytes_allocated = getattr(self, "bytes_allocated", 0) + nbytes
        dist.barrier()
        return pb

    def _alloc_symm(self, name: str, numel: int, nbytes: int, dtype: torch.dtype) -> Optional[PeerBuffer]:
        """Symmetric-memory allocation; every rank must agree on success, so the outcome is all-reduced."""
        ok, t, hdl = 1, None, None
        try:
            import torch.distributed._symmetric_memory as symm_mem

            t = symm_mem.empty(nbytes, dtype=torch.uint8, device=self.device)
            try:
                hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
            except TypeError:
                hdl = symm_mem.rendezvous(t, group=dist.group.WORLD)
            t.zero_()
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            if len(ptrs) != self.ws or ptrs[self.rank] != t.data_ptr():
                ok = 0
        except Exception as e:      # noqa: BLE001 - any failure means "use the IPC path"
            ok = 0
            if self.rank == 0 and not getattr(self, "_symm_warned", False):
                print(f"[b200bench] symmetric memory unavailable ({type(e).__name__}: {str(e)[:160]}); "
                      f"falling back to cudaMalloc + CUDA IPC (no NVLS)", file=sys.stderr, flush=True)
                self._symm_warned = True
        flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            self.use_symm = False
            return None
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        pb = PeerBuffer(name, t.data_ptr(), nbytes, ptrs, t.view(dtype)[:numel], mc_ptr=mc, handle=(hdl, t))
        self.buffers[name] = pb
        self.bytes_allocated = getattr(self, "bytes_allocated", 0) + nbytes
        torch.cuda.synchronize(self.device)
        dist.barrier()
        return pb

    @property
    def has_multicast(self) -> bool:
        return all(pb.mc_ptr != 0 for pb in self.buffers.values()) and len(self.buffers) > 0

    # ---- stream plumbing: comm kernels run on a side stream, forked from / joined to compute ------
    def fork(self) -> None:
        """Make the comm stream wait for everything issued so far on the current stream."""
        self._ev_ready.record(torch.cuda.current_stream())
        self.stream.wait_event(self._ev_ready)
        self._forked = True

    def join(self) -> None:
        if self._forked:
            self._ev_done.record(self.stream)
            torch.cuda.current_stream().wait_event(self._ev_done)
            self._forked = False

    # ---- collectives -------------------------------------------------------------------------------
    def allreduce(self, buf: PeerBuffer, off: int, n: int, scale: float, side_stream: bool = True) -> None:
        s = self.stream if side_stream else torch.cuda.current_stream()
        if self.use_nvls and buf.mc_ptr:
            rc = self.ops.lib.b200_comm_allreduce_nvls(c_void_p(buf.mc_ptr), self.flags.table, self.rank, self.ws,
                                                       off, n, scale, self.nvls_ctas, c_void_p(s.cuda_stream))
        else:
            rc = self.ops.lib.b200_comm_allreduce(buf.table, self.flags.table, self.rank, self.ws, off, n, scale,
                                                  self.comm_ctas, c_void_p(s.cuda_stream))
        if rc:
            raise _ops.NativeError(f"comm_allreduce failed with code {rc}")
        self.ops.launches += 2

    def rs_adamw(self, grads: PeerBuffer, weights: PeerBuffer, lo: int, n: int, master, m, v, gshard, hp,
                 push_all: bool, gnorm_out=None, extra_norm_sq=None, ctas: int = 128, weights_table=None) -> None:
        s = torch.cuda.current_stream()
        p = lambda t: c_void_p(t.data_ptr()) if t is not None else c_void_p(0)  # noqa: E731
        wt = weights_table if weights_table is not None else weights.table
        if self.use_nvls and push_all and weights.mc_ptr and weights_table is None:
            rc = self.ops.lib.b200_comm_rs_adamw_nvls(grads.table, wt, c_void_p(weights.mc_ptr), self.flags.table,
                                                      self.rank, self.ws, lo, n, p(master), p(m), p(v), p(gshard),
                                                      p(hp), int(push_all), p(gnorm_out), p(extra_norm_sq), ctas,
                                                      c_void_p(s.cuda_stream))
        else:
            rc = self.ops.lib.b200_comm_rs_adamw(grads.table, wt, self.flags.table, self.rank, self.ws,
                                                 lo, n, p(master), p(m), p(v), p(gshard), p(hp), int(push_all),
                                                 p(gnorm_out), p(extra_norm_sq), ctas, c_void_p(s.cuda_stream))
        if rc:
            raise _ops.NativeError(f"comm_rs_adamw failed with code {rc}")
        self.ops.launches += 2

    def adjusted_table(self, buf: PeerBuffer, shard_elems: int, elem_size: int = 2):
        """Pointer table for a partition buffer, biased so that FLAT offsets index it: entry r =
        (rank r's partition base) - r * shard_elems."""
        return (c_void_p * self.ws)(*[pp - r * shard_elems * elem_size for r, pp in enumerate(buf.peer_ptrs)])

    def gather_range(self, table, dst: torch.Tensor, flat_off: int, n: int, shard_elems: int, ctas: int = 128) -> None:
        rc = self.ops.lib.b200_comm_gather_range_from(table, self.ws, self.rank, c_void_p(dst.data_ptr()), flat_off, n,
                                                      shard_elems, ctas,
                                                 c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc:
            raise _ops.NativeError(f"comm_gather_range failed with code {rc}")
        self.ops.launches += 1

    def gather_bcast(self, table, dst: PeerBuffer, dst_off: int, flat_off: int, n: int, shard_elems: int,
                     ctas: int = 0) -> None:
        """K16: all-gather of the flat range into element offset dst_off of EVERY rank's copy of ``dst`` - a
        collective (two rank barriers): every rank must call it with the same arguments in the same order.  Through
        the NVSwitch multicast address of ``dst`` when it has one (B200BENCH_GATHER_MC=0: peer stores)."""
        mc = dst.mc_ptr if os.environ.get("B200BENCH_GATHER_MC", "1") != "0" else 0
        rc = self.ops.lib.b200_comm_gather_bcast(table, dst.table, c_void_p(mc) if mc else None, self.flags.table,
                                                 self.rank, self.ws, flat_off, dst_off, n, shard_elems,
                                                 ctas or (16 if mc else 32),
                                                 c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc:
            raise _ops.NativeError(f"comm_gather_bcast failed with code {rc}")
        self.ops.launches += 2

    def peer_copy(self, src_ptr: int, dst: torch.Tensor, nbytes: int, ctas: int = 64) -> None:
        rc = self.ops.lib.b200_peer_copy(c_void_p(src_ptr), c_void_p(dst.data_ptr()), nbytes, ctas,
                                         c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc:
            raise _ops.NativeError(f"peer_copy failed with code {rc}")
        self.ops.launches += 1

    def extra_memory_bytes(self) -> int:
        return int(self.bytes_allocated)

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        if dist.is_initialized():
            dist.barrier()
        for pb in self.buffers.values():
            if pb.handle is not None:
                continue
            for r, pp in enumerate(pb.peer_ptrs):
                if r != self.rank:
                    self.ops.lib.b200_peer_close(c_void_p(pp))
        if dist.is_initialized():
            dist.barrier()
        for pb in self.buffers.values():
            pb.tensor = None
            if pb.handle is None:
                self.ops.lib.b200_peer_free(c_void_p(pb.local_ptr))
            pb.handle = None                  # symmetric memory: dropping the tensor and the handle unmaps it
        self.buffers.clear()
