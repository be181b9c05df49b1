"""b200_colsum_static through the C ABI: the allocation-free bias-gradient column sum.

Who needs it: callers that have no way to provide a scratch buffer — C code embedded in a process
where another framework's pool already holds the GPU's memory, or a step loop that must never enter
cudaMalloc/cudaFree (driver locks, device-wide sync).  Its scratch is a fixed 512 KiB array inside the
library, shared by all calls on a GPU, so the things to check are: same numbers as the ordinary entry
point (b200_colsum with a caller buffer), correct again on a second call that finds the first call's
leftovers in the array, and a clean refusal of shapes that do not fit.

Collected last on purpose (file name), so that with ``-x`` every other test has already run.  Uses only
ctypes on the shared library; passes alone in a fresh process and after the rest of the suite.
"""
import ctypes
import math
import os
from ctypes import c_int, c_size_t, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "b200bench", "lib", "libb200bench.so")
ERR_STATIC_SCRATCH = -8


def _lib():
    lib = ctypes.CDLL(LIB)          # fails loudly if the native library was not built
    lib.b200_colsum.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p]
    lib.b200_colsum_static.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]
    lib.b200_colreduce_scratch_bytes.argtypes = [c_int, c_int, c_int]
    lib.b200_colreduce_scratch_bytes.restype = c_size_t
    lib.b200_colsum_static_max_bytes.restype = c_size_t
    return lib


def _close(a, b, M):
    a, b = a.float(), b.float()
    atol, rtol = 0.05 * math.sqrt(M / 64), 3e-2            # same tolerance as test_kernels_gpu.test_layernorm
    assert bool(((a - b).abs() <= atol + rtol * b.abs()).all()), (a - b).abs().max().item()


@pytest.mark.parametrize("M,N", [(2048, 4096), (300, 1024)])
def test_static_scratch_matches_caller_scratch_twice(M, N):
    lib = _lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.manual_seed(M + N)
    nbytes = lib.b200_colreduce_scratch_bytes(M, N, 1)
    assert nbytes <= lib.b200_colsum_static_max_bytes()
    scratch = torch.zeros(nbytes // 4 + 1, dtype=torch.float32, device=dev)
    for call in range(2):                                   # second call meets the first call's leftovers
        dy = (torch.randn(M, N, device=dev) * (1 + call)).to(torch.bfloat16)
        ref = torch.empty(N, dtype=torch.bfloat16, device=dev)
        got = torch.full((N,), float("nan"), dtype=torch.bfloat16, device=dev)
        rc = lib.b200_colsum(c_void_p(dy.data_ptr()), c_void_p(ref.data_ptr()), M, N, N, 0,
                             c_void_p(scratch.data_ptr()), c_size_t(scratch.numel() * 4), stream)
        assert rc == 0
        rc = lib.b200_colsum_static(c_void_p(dy.data_ptr()), c_void_p(got.data_ptr()), M, N, N, 0, stream)
        assert rc == 0
        torch.cuda.synchronize(dev)
        _close(got, ref, M)                                 # against the existing entry point
        _close(got, dy.float().sum(0), M)                   # and against plain PyTorch fp32
    # accumulate flag: out += column sums
    rc = lib.b200_colsum_static(c_void_p(dy.data_ptr()), c_void_p(got.data_ptr()), M, N, N, 1, stream)
    assert rc == 0
    torch.cuda.synchronize(dev)
    _close(got, 2 * dy.float().sum(0), M)


def test_static_scratch_rejects_oversize():
    lib = _lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    M, N = 64, 8 * 40000                                    # more columns than the built-in buffer serves
    dy = torch.zeros(8, dtype=torch.bfloat16, device=dev)   # never read: the call must refuse first
    rc = lib.b200_colsum_static(c_void_p(dy.data_ptr()), c_void_p(dy.data_ptr()), M, N, N, 0,
                                c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == ERR_STATIC_SCRATCH
    torch.cuda.synchronize(dev)                              # and nothing was launched
