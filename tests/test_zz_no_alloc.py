"""Allocation-free entry point (include/retr_b200.h: retr_op_small_gemm_noalloc) against the existing GEMM entry
point (retr_op_gemm), through the C ABI.  Collected LAST on purpose (file name).

The new call keeps its split-K partial tiles and tickets in file-scope static __device__ arrays that every call
shares, so it is called twice in a row with different shapes and data: the second result must not depend on what the
first left behind.  Tolerance: both paths multiply bf16 operands and accumulate in fp32; only the summation order
differs (the bar tests/test_ops_gpu.py uses for the bf16 GEMMs)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _existing(native, lib, A, W, bias, M, N, K):
    out = torch.zeros(M, N, dtype=torch.float32, device=A.device)
    a = native.GemmArgs()
    a.A, a.W, a.C = A.data_ptr(), W.data_ptr(), out.data_ptr()
    a.M, a.N, a.K, a.lda, a.ldc = M, N, K, K, N
    a.bias = bias.data_ptr()
    a.out_fp32 = 1
    native.check(lib.retr_op_gemm(native.BF16, C.byref(a), native.stream_ptr(A.device)))
    return out


def _noalloc(native, lib, A, W, bias, M, N, K, split_k):
    out = torch.full((M, N), float("nan"), dtype=torch.float32, device=A.device)
    native.check(lib.retr_op_small_gemm_noalloc(native.ptr(A), native.ptr(W), native.ptr(bias), native.ptr(out), M, N, K,
                                                split_k, native.stream_ptr(A.device)))
    return out


def test_noalloc_small_gemm_matches_existing_entry_point_twice():
    from retr_b200 import native
    lib = native.lib()
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(5)
    # the decode step's FFN second linear (M = batch, 2048 -> 256, 4 splits), then a ragged, differently split one
    for M, N, K, split_k in ((256, 256, 2048, 4), (37, 520, 1024, 2)):
        A = torch.randn(M, K, generator=g).to(dev, torch.bfloat16)
        W = (torch.randn(N, K, generator=g) / K ** 0.5).to(dev, torch.bfloat16)
        bias = torch.randn(N, generator=g).to(dev)
        want = _existing(native, lib, A, W, bias, M, N, K)
        got = _noalloc(native, lib, A, W, bias, M, N, K, split_k)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        err = (got - want).abs().max().item()
        assert err <= 2e-3 * want.abs().max().item(), (M, N, K, split_k, err)
        ref = A.float() @ W.float().t() + bias
        assert (got - ref).abs().max().item() <= 2e-3 * ref.abs().max().item()


def test_noalloc_small_gemm_rejects_oversize():
    from retr_b200 import native
    lib = native.lib()
    dev = torch.device("cuda")
    M, N, K, split_k = 2048, 1024, 1024, 4          # 8 Mi floats of partial tiles > the 1 Mi static scratch
    A = torch.zeros(M, K, dtype=torch.bfloat16, device=dev)
    W = torch.zeros(N, K, dtype=torch.bfloat16, device=dev)
    out = torch.zeros(M, N, dtype=torch.float32, device=dev)
    rc = lib.retr_op_small_gemm_noalloc(native.ptr(A), native.ptr(W), None, native.ptr(out), M, N, K, split_k,
                                        native.stream_ptr(dev))
    assert rc != 0
    assert b"exceeds the static scratch" in lib.retr_last_error()
