"""Loader and Python front-end of the native library (``b200bench/lib/libb200bench.so``).

The library is plain C ABI (``include/b200bench.h``) and is called through ``ctypes`` with raw
device pointers and the current CUDA stream, so there is no dependence on PyTorch's C++ ABI and
every launch is capturable in a CUDA graph.  ``load()`` raises if the library is missing: on a GPU
box a silent fallback to PyTorch ops would defeat the purpose of the native path.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from ctypes import c_float, c_int, c_size_t, c_uint, c_void_p
from typing import Optional

import torch

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
# B200BENCH_LIB selects another build of the same ABI (e.g. `make debug-comm` -> libb200bench_debug.so)
LIB_PATH = os.environ.get("B200BENCH_LIB") or os.path.join(_ROOT, "b200bench", "lib", "libb200bench.so")

EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_BIAS_RESID, EPI_DGELU, EPI_ACCUM, EPI_ROWDOT, EPI_BIAS_GELU_GRAD, EPI_MUL = range(9)
ATTN_BWD_HAVE_D, ATTN_BWD_ACC_IS_ZERO = 1, 2


class GemmEx(ctypes.Structure):
    """Mirror of ``B200GemmEx`` (include/b200bench.h)."""
    _fields_ = [("dbias", c_void_p), ("dbias_accum", c_int), ("rowdot", c_void_p), ("rd_T", c_int),
                ("rd_hd", c_int), ("b_shards", c_void_p), ("b_ws", c_int), ("b_flat_off", ctypes.c_ulonglong),
                ("b_shard_elems", ctypes.c_ulonglong), ("mc_units", c_int)]

_ops: Optional["NativeOps"] = None


def build(verbose: bool = False) -> str:
    """Compile every .cu for sm_100a into the in-tree shared library (needs nvcc, no GPU)."""
    r = subprocess.run(["make", "-j8", "-C", _ROOT, "all"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("native build failed (make all)")
    return LIB_PATH


def _p(t) -> c_void_p:
    if t is None:
        return c_void_p(0)
    if isinstance(t, int):
        return c_void_p(t)
    return c_void_p(t.data_ptr())


def _stream() -> c_void_p:
    return c_void_p(torch.cuda.current_stream().cuda_stream)


class NativeError(RuntimeError):
    pass


class NativeOps:
    def __init__(self, path: str = LIB_PATH):
        if not os.path.exists(path):
            raise NativeError(f"{path} not found: run `make` (or __graft_entry__.build()) first")
        self.path = path
        self.lib = ctypes.CDLL(path)
        L = self.lib
        vp, i, f, u, sz = c_void_p, c_int, c_float, c_uint, c_size_t
        L.b200_gemm_bf16.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, i, vp, vp, vp, f, u, i, i, vp]
        L.b200_gemm_bf16_ex.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, i, vp, vp, vp, f, u, i, i, vp,
                                        ctypes.POINTER(GemmEx)]
        L.b200_tmap_cache_stats.argtypes = [ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(ctypes.c_ulonglong)]
        L.b200_attention_fwd.argtypes = [vp, vp, vp, i, i, i, i, i, f, u, vp]
        L.b200_attention_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, f, u, vp]
        L.b200_attention_bwd_ex.argtypes = [vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, f, u, vp, i]
        L.b200_layernorm_bwd_fused_scratch_bytes.argtypes = [i, i]
        L.b200_layernorm_bwd_fused_scratch_bytes.restype = sz
        L.b200_layernorm_bwd_fused.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, f, u, vp, sz, i, i, vp]
        L.b200_layernorm_param_fold.argtypes = [vp, i, i, vp, vp, i, vp]
        L.b200_layernorm_fwd.argtypes = [vp, vp, vp, vp, vp, vp, i, i, f, vp]
        L.b200_layernorm_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, i, i, vp]
        L.b200_colreduce_scratch_bytes.argtypes = [i, i, i]
        L.b200_colreduce_scratch_bytes.restype = sz
        L.b200_layernorm_bwd_params.argtypes = [vp, vp, vp, vp, vp, vp, i, i, i, vp, sz, vp]
        L.b200_colsum.argtypes = [vp, vp, i, i, i, i, vp, sz, vp]
        L.b200_embedding_fwd.argtypes = [vp, vp, vp, vp, i, i, i, f, u, vp]
        L.b200_embedding_bwd.argtypes = [vp, vp, vp, vp, i, i, i, f, u, i, vp]
        L.b200_cross_entropy.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, f, i, i, f, vp]
        L.b200_adamw_set_hp.argtypes = [vp, f, f, f, f, f, f, f, f, f, vp]
        L.b200_adamw.argtypes = [vp, vp, vp, vp, vp, sz, vp, vp, vp]
        L.b200_sumsq.argtypes = [vp, sz, f, vp, vp]
        L.b200_dropout_apply.argtypes = [vp, vp, i, i, f, u, vp]
        L.b200_step_seed_register.argtypes = [vp]
        L.b200_step_seed_advance.argtypes = [vp]
        self.launches = 0          # kernel launches issued through this object (bench.py reports it)

    def _ck(self, rc: int, what: str) -> None:
        if rc != 0:
            raise NativeError(f"{what} failed with code {rc}")

    # ---- GEMM -------------------------------------------------------------------------------
    def gemm(self, A, B, D, M, N, K, a_mn=False, b_mn=False, epi=EPI_NONE, bias=None, aux=None, D2=None,
             drop_p=0.0, seed=0, lda=None, ldb=None, ldd=None, bn=0, max_ctas=0, dbias=None, dbias_accum=False,
             rowdot=None, rd_T=0, rd_hd=0, peer=None, mc_units=0) -> None:
        """``peer`` (K14): dict(table=<c_void_p array of the owners' shard bases>, ws, flat_off, shard) - B is then
        fetched tile by tile from the owning ranks and the ``B`` argument is only used for its shape."""
        if peer is not None and mc_units > 1:
            raise NativeError("cluster multicast is for a local B only: TMA multicast loads read only this GPU's memory")
        lda = lda or (M if a_mn else K)
        ldb = ldb or (N if b_mn else K)
        ldd = ldd or N
        if dbias is None and rowdot is None and peer is None and not mc_units:
            self._ck(self.lib.b200_gemm_bf16(_p(A), _p(B), _p(D), M, N, K, lda, ldb, ldd, int(a_mn), int(b_mn),
                                             epi, _p(bias), _p(aux), _p(D2), drop_p, seed, bn, max_ctas,
                                             _stream()), "gemm")
            self.launches += 1
            return
        ex = GemmEx()
        ex.dbias = dbias.data_ptr() if dbias is not None else None
        ex.dbias_accum = int(bool(dbias_accum))
        ex.rowdot = rowdot.data_ptr() if rowdot is not None else None
        ex.rd_T, ex.rd_hd, ex.mc_units = int(rd_T), int(rd_hd), int(mc_units)
        if peer is not None:
            ex.b_shards = ctypes.cast(peer["table"], c_void_p)
            ex.b_ws, ex.b_flat_off, ex.b_shard_elems = int(peer["ws"]), int(peer["flat_off"]), int(peer["shard"])
        self._ck(self.lib.b200_gemm_bf16_ex(_p(A), _p(None if peer is not None else B), _p(D), M, N, K, lda, ldb,
                                            ldd, int(a_mn), int(b_mn), epi, _p(bias), _p(aux), _p(D2), drop_p,
                                            seed, bn, max_ctas, _stream(), ctypes.byref(ex)), "gemm_ex")
        self.launches += 1

    def linear_fwd(self, x, w, out, bias=None, epi=None, aux=None, out2=None, drop_p=0.0, seed=0, peer=None) -> None:
        """out[M,N] = epi(x[M,K] @ w[N,K]^T).  ``w`` may be a (N, K) shape tuple when ``peer`` names the shards."""
        M, K = x.shape
        N = w[0] if isinstance(w, tuple) else w.shape[0]
        if epi is None:
            epi = EPI_BIAS if bias is not None else EPI_NONE
        self.gemm(x, None if isinstance(w, tuple) else w, out, M, N, K, epi=epi, bias=bias, aux=aux, D2=out2,
                  drop_p=drop_p, seed=seed, peer=peer)

    def linear_dgrad(self, dy, w, dx, epi=EPI_NONE, aux=None, peer=None, rowdot=None, rd_T=0, rd_hd=0) -> None:
        """dx[M,K] = epi(dy[M,N] @ w[N,K])"""
        M, N = dy.shape
        K = w[1] if isinstance(w, tuple) else w.shape[1]
        self.gemm(dy, None if isinstance(w, tuple) else w, dx, M, K, N, b_mn=True, epi=epi, aux=aux, peer=peer,
                  rowdot=rowdot, rd_T=rd_T, rd_hd=rd_hd)

    def linear_wgrad(self, dy, x, dw, accumulate=False, dbias=None) -> None:
        """dw[N,K] (+)= dy[M,N]^T @ x[M,K];  dbias[N] (+)= column sums of dy (same main loop)."""
        M, N = dy.shape
        K = x.shape[1]
        self.gemm(dy, x, dw, N, K, M, a_mn=True, b_mn=True, epi=EPI_ACCUM if accumulate else EPI_NONE,
                  dbias=dbias, dbias_accum=accumulate)

    def tmap_cache_stats(self):
        h, m = ctypes.c_ulonglong(0), ctypes.c_ulonglong(0)
        self.lib.b200_tmap_cache_stats(ctypes.byref(h), ctypes.byref(m))
        return int(h.value), int(m.value)

    # ---- attention ----------------------------------------------------------------------------
    def attention_head_dims(self):
        """Head widths the attention kernels are instantiated for."""
        return (64,)

    def attention_fwd(self, qkv, out, lse, B, T, H, hd, causal=False, drop_p=0.0, seed=0) -> None:
        self._ck(self.lib.b200_attention_fwd(_p(qkv), _p(out), _p(lse), B, T, H, hd, int(causal), drop_p,
                                             seed, _stream()), "attention_fwd")
        self.launches += 1

    def set_attention_fwd_variant(self, lazy: int) -> int:
        """1 = lazy rescale with O in tensor memory (default), 0 = exact per-block rescale; returns the old value."""
        return int(self.lib.b200_set_attention_fwd_variant(int(lazy)))

    def attention_bwd(self, qkv, out, dout, lse, dqkv, D, dq_acc, B, T, H, hd, causal=False, drop_p=0.0,
                      seed=0, have_d=False, acc_is_zero=False) -> None:
        """``have_d``: D was already produced (EPI_ROWDOT of the out-proj dgrad); ``acc_is_zero``: dq_acc holds
        zeros (every call leaves it zeroed).  With both, two launches: main kernel + dQ conversion."""
        flags = (ATTN_BWD_HAVE_D if have_d else 0) | (ATTN_BWD_ACC_IS_ZERO if acc_is_zero else 0)
        self._ck(self.lib.b200_attention_bwd_ex(_p(qkv), _p(out), _p(dout), _p(lse), _p(dqkv), _p(D), _p(dq_acc),
                                                B, T, H, hd, int(causal), drop_p, seed, _stream(), flags),
                 "attention_bwd")
        self.launches += 2 + (0 if have_d else 1) + (0 if acc_is_zero else 1)

    # ---- LayerNorm ----------------------------------------------------------------------------
    def layernorm_fwd(self, x, gamma, beta, y, mean, rstd, eps=1e-5) -> None:
        M, d = x.shape
        self._ck(self.lib.b200_layernorm_fwd(_p(x), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), M, d, eps,
                                             _stream()), "layernorm_fwd")
        self.launches += 1

    def layernorm_bwd(self, dy, x, gamma, mean, rstd, dx, dres=None) -> None:
        M, d = x.shape
        self._ck(self.lib.b200_layernorm_bwd(_p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dres), _p(dx),
                                             M, d, _stream()), "layernorm_bwd")
        self.launches += 1

    def layernorm_bwd_fused_scratch_bytes(self, M, d) -> int:
        return int(self.lib.b200_layernorm_bwd_fused_scratch_bytes(M, d))

    def layernorm_bwd_fused(self, dy, x, gamma, mean, rstd, dx, dres=None, dx_drop=None, drop_p=0.0, seed=0,
                            partials=None, dgamma=None, dbeta=None, accumulate=False) -> None:
        """dx (+ dres), optional dropout-masked copy dx_drop, optional per-CTA partial column sums for the
        parameter gradients - one pass, one launch.  With dgamma / dbeta the fold is launched right behind it on
        the current stream; callers that want it on another stream pass only ``partials`` and call
        ``layernorm_param_fold`` themselves."""
        M, d = x.shape
        nbytes = partials.numel() * partials.element_size() if partials is not None else 0
        self._ck(self.lib.b200_layernorm_bwd_fused(_p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dres), _p(dx),
                                                   _p(dx_drop), drop_p, seed, _p(partials), nbytes, M, d, _stream()),
                 "layernorm_bwd_fused")
        self.launches += 1
        if dgamma is not None:
            self.layernorm_param_fold(partials, M, d, dgamma, dbeta, accumulate)

    def layernorm_param_fold(self, partials, M, d, dgamma, dbeta, accumulate=False) -> None:
        self._ck(self.lib.b200_layernorm_param_fold(_p(partials), M, d, _p(dgamma), _p(dbeta), int(accumulate),
                                                    _stream()), "layernorm_param_fold")
        self.launches += 1

    def colreduce_scratch_bytes(self, M, N, outputs) -> int:
        return int(self.lib.b200_colreduce_scratch_bytes(M, N, outputs))

    def layernorm_bwd_params(self, dy, x, mean, rstd, dgamma, dbeta, scratch, accumulate=False) -> None:
        M, d = x.shape
        self._ck(self.lib.b200_layernorm_bwd_params(_p(dy), _p(x), _p(mean), _p(rstd), _p(dgamma), _p(dbeta),
                                                    M, d, int(accumulate), _p(scratch),
                                                    scratch.numel() * scratch.element_size(), _stream()),
                 "layernorm_bwd_params")
        self.launches += 1

    def colsum(self, dy, out, scratch, accumulate=False, ld=None) -> None:
        M, N = dy.shape
        self._ck(self.lib.b200_colsum(_p(dy), _p(out), M, N, ld or N, int(accumulate), _p(scratch),
                                      scratch.numel() * scratch.element_size(), _stream()), "colsum")
        self.launches += 1

    # ---- embedding / loss -----------------------------------------------------------------------
    def embedding_fwd(self, tokens, wte, wpe, out, B, T, d, drop_p=0.0, seed=0) -> None:
        self._ck(self.lib.b200_embedding_fwd(_p(tokens), _p(wte), _p(wpe), _p(out), B, T, d, drop_p, seed,
                                             _stream()), "embedding_fwd")
        self.launches += 1

    def embedding_bwd(self, tokens, dx, dwte, dwpe, B, T, d, drop_p=0.0, seed=0, accumulate_wpe=False) -> None:
        self._ck(self.lib.b200_embedding_bwd(_p(tokens), _p(dx), _p(dwte), _p(dwpe), B, T, d, drop_p, seed,
                                             int(accumulate_wpe), _stream()), "embedding_bwd")
        self.launches += 1

    def cross_entropy(self, logits, targets, loss_rows, n_valid, loss_out, M, V, ld=None, ignore_index=-1,
                      grad_scale=1.0, write_grad=True, accumulate_loss=False, loss_scale=1.0) -> None:
        self._ck(self.lib.b200_cross_entropy(_p(logits), _p(targets), _p(loss_rows), _p(n_valid), _p(loss_out),
                                             M, V, ld or V, ignore_index, grad_scale, int(write_grad),
                                             int(accumulate_loss), loss_scale, _stream()), "cross_entropy")
        self.launches += 3

    # ---- optimizer ------------------------------------------------------------------------------
    def adamw_set_hp(self, hp, lr, b1, b2, eps, wd, ibc1, isbc2, gscale, clip) -> None:
        self._ck(self.lib.b200_adamw_set_hp(_p(hp), lr, b1, b2, eps, wd, ibc1, isbc2, gscale, clip, _stream()),
                 "adamw_set_hp")
        self.launches += 1

    def adamw_flat(self, master, m, v, grad, w_out, hp, gnorm_sq=None) -> None:
        self._ck(self.lib.b200_adamw(_p(master), _p(m), _p(v), _p(grad), _p(w_out), master.numel(), _p(hp),
                                     _p(gnorm_sq), _stream()), "adamw")
        self.launches += 1

    def sumsq(self, grad, out, scale=1.0) -> None:
        self._ck(self.lib.b200_sumsq(_p(grad), grad.numel(), scale, _p(out), _stream()), "sumsq")
        self.launches += 1

    def dropout_apply(self, x, out, M, N, drop_p, seed) -> None:
        self._ck(self.lib.b200_dropout_apply(_p(x), _p(out), M, N, drop_p, seed, _stream()), "dropout_apply")
        self.launches += 1


    def step_seed_register(self, word) -> None:
        self._seed_word = word          # keep the tensor alive for as long as the library points at it
        self._ck(self.lib.b200_step_seed_register(_p(word)), "step_seed_register")

    def step_seed_unregister(self) -> None:
        self._seed_word = None
        self._ck(self.lib.b200_step_seed_register(c_void_p(0)), "step_seed_unregister")

    def step_seed_advance(self) -> None:
        self._ck(self.lib.b200_step_seed_advance(_stream()), "step_seed_advance")
        self.launches += 1


_TIMED = ("gemm", "attention_fwd", "attention_bwd", "layernorm_fwd", "layernorm_bwd", "layernorm_bwd_fused",
          "layernorm_param_fold", "layernorm_bwd_params", "colsum", "embedding_fwd", "embedding_bwd", "cross_entropy",
          "adamw_flat", "sumsq", "dropout_apply")


def enable_op_timing(ops: "NativeOps") -> None:
    """Per-op device times (the reference has no profiling hooks).  Every front-end call is bracketed
    by CUDA events on its stream; ``op_timing_report()`` prints totals by op and shape.  For eager runs only
    (``--no-graph``): events cannot be recorded inside a graph capture, and side-stream overlap is serialised
    by the measurement itself only in the sense that each op is timed on its own stream."""
    if getattr(ops, "_timing", None) is not None:
        return
    ops._timing = []

    def wrap(name, fn):
        def timed(*a, **k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn(*a, **k)
            e1.record()
            key = name
            if name == "gemm":
                key = (f"gemm M{a[3]} N{a[4]} K{a[5]} a_mn{int(k.get('a_mn', False))} b_mn{int(k.get('b_mn', False))} "
                       f"epi{k.get('epi', 0)}{' +dbias' if k.get('dbias') is not None else ''}"
                       f"{' peer' if k.get('peer') is not None else ''}")
            ops._timing.append((key, e0, e1))
            return r
        return timed

    for n in _TIMED:
        setattr(ops, n, wrap(n, getattr(ops, n)))


def op_timing_report(ops: "NativeOps", steps: int = 1) -> str:
    torch.cuda.synchronize()
    agg = {}
    for key, e0, e1 in ops._timing:
        t = e0.elapsed_time(e1) * 1e3
        c = agg.setdefault(key, [0, 0.0])
        c[0] += 1
        c[1] += t
    ops._timing.clear()
    total = sum(v[1] for v in agg.values())
    lines = [f"| op | calls/step | us/call | us/step | share |", "|---|---|---|---|---|"]
    for key, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f"| {key} | {n / steps:.0f} | {t / n:.1f} | {t / steps:.0f} | {100 * t / total:.1f} % |")
    lines.append(f"| total | | | {total / steps:.0f} | |")
    return "\n".join(lines)


def load() -> NativeOps:
    global _ops
    if _ops is None:
        _ops = NativeOps()
        if os.environ.get("B200BENCH_OP_TIMING") == "1":
            enable_op_timing(_ops)
    return _ops


def available() -> bool:
    return os.path.exists(LIB_PATH)
