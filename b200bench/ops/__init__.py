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

EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_BIAS_RESID, EPI_DGELU, EPI_ACCUM = range(6)

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
        ull = ctypes.c_ulonglong
        L.b200_gemm_bf16_gather.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, i, vp, vp, vp, f, u, i, i, vp,
                                            vp, i, ull, ull, i, i, vp, vp, i, ull, i, i, vp, vp]
        L.b200_attention_fwd.argtypes = [vp, vp, vp, i, i, i, i, i, f, u, vp]
        L.b200_attention_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, f, u, vp]
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
             drop_p=0.0, seed=0, lda=None, ldb=None, ldd=None, bn=0, max_ctas=0, gather=None) -> None:
        lda = lda or (M if a_mn else K)
        ldb = ldb or (N if b_mn else K)
        ldd = ldd or N
        if gather is not None:
            # K14: B is a staging buffer filled by the kernel itself from the owners' partitions
            g = gather
            self._ck(self.lib.b200_gemm_bf16_gather(_p(A), _p(B), _p(D), M, N, K, lda, ldb, ldd, int(a_mn),
                                                    int(b_mn), epi, _p(bias), _p(aux), _p(D2), drop_p, seed, bn,
                                                    max_ctas, _stream(), g["table"], g["ws"], g["flat_off"],
                                                    g["shard"], g["rows"], g["cols"], _p(g["flags"]),
                                                    _p(g["epoch"]), int(g.get("self_gather", 1)),
                                                    g.get("pf_flat_off", 0), g.get("pf_rows", 0),
                                                    g.get("pf_cols", 0), _p(g.get("pf_stage")),
                                                    _p(g.get("pf_flags"))), "gemm_gather")
            self.launches += 2
            return
        self._ck(self.lib.b200_gemm_bf16(_p(A), _p(B), _p(D), M, N, K, lda, ldb, ldd, int(a_mn), int(b_mn),
                                         epi, _p(bias), _p(aux), _p(D2), drop_p, seed, bn, max_ctas,
                                         _stream()), "gemm")
        self.launches += 1

    def linear_fwd(self, x, w, out, bias=None, epi=None, aux=None, out2=None, drop_p=0.0, seed=0, gather=None) -> None:
        """out[M,N] = epi(x[M,K] @ w[N,K]^T)"""
        M, K = x.shape
        N = w.shape[0]
        if epi is None:
            epi = EPI_BIAS if bias is not None else EPI_NONE
        self.gemm(x, w, out, M, N, K, epi=epi, bias=bias, aux=aux, D2=out2, drop_p=drop_p, seed=seed, gather=gather)

    def linear_dgrad(self, dy, w, dx, epi=EPI_NONE, aux=None, gather=None) -> None:
        """dx[M,K] = epi(dy[M,N] @ w[N,K])"""
        M, N = dy.shape
        K = w.shape[1]
        self.gemm(dy, w, dx, M, K, N, b_mn=True, epi=epi, aux=aux, gather=gather)

    def linear_wgrad(self, dy, x, dw, accumulate=False) -> None:
        """dw[N,K] (+)= dy[M,N]^T @ x[M,K]"""
        M, N = dy.shape
        K = x.shape[1]
        self.gemm(dy, x, dw, N, K, M, a_mn=True, b_mn=True, epi=EPI_ACCUM if accumulate else EPI_NONE)

    # ---- attention ----------------------------------------------------------------------------
    def attention_fwd(self, qkv, out, lse, B, T, H, hd, causal=False, drop_p=0.0, seed=0) -> None:
        self._ck(self.lib.b200_attention_fwd(_p(qkv), _p(out), _p(lse), B, T, H, hd, int(causal), drop_p,
                                             seed, _stream()), "attention_fwd")
        self.launches += 1

    def attention_bwd(self, qkv, out, dout, lse, dqkv, D, dq_acc, B, T, H, hd, causal=False, drop_p=0.0,
                      seed=0) -> None:
        self._ck(self.lib.b200_attention_bwd(_p(qkv), _p(out), _p(dout), _p(lse), _p(dqkv), _p(D), _p(dq_acc),
                                             B, T, H, hd, int(causal), drop_p, seed, _stream()),
                 "attention_bwd")
        self.launches += 4         # memset + prep + main + dq convert

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

    def step_seed_advance(self) -> None:
        self._ck(self.lib.b200_step_seed_advance(_stream()), "step_seed_advance")
        self.launches += 1


def load() -> NativeOps:
    global _ops
    if _ops is None:
        _ops = NativeOps()
    return _ops


def available() -> bool:
    return os.path.exists(LIB_PATH)
