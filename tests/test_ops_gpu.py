"""Kernel-level parity through the C ABI (retr_op_*): each CUDA kernel against a plain PyTorch
fp32 statement of the same op on the same seeded inputs.  GPU only."""
import ctypes as C
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

FP32, BF16 = 0, 1


def _lib():
    from retr_b200 import native
    return native, native.lib()


def run_gemm(dtype, A, W, *, M, N, K, lda=None, ldc=None, scale=None, bias=None, rowbias=None, period=1,
             residual=None, relu=0, out_fp32=False, conv=None, argmax=False, want_c=True, out=None, mn_major=0, ldb=0):
    native, lib = _lib()
    dev = A.device
    out_dtype = torch.float32 if (out_fp32 or dtype == FP32) else torch.bfloat16
    ldc = ldc or N
    Cout = (out if out is not None else torch.zeros(M, ldc, dtype=out_dtype, device=dev)) if want_c else None
    keys = torch.zeros(M, dtype=torch.int64, device=dev) if argmax else None
    a = native.GemmArgs()
    a.A, a.W, a.C = A.data_ptr(), W.data_ptr(), (Cout.data_ptr() if want_c else None)
    a.M, a.N, a.K, a.lda, a.ldc = M, N, K, (lda or K), ldc
    a.scale = scale.data_ptr() if scale is not None else None
    a.bias = bias.data_ptr() if bias is not None else None
    a.rowbias = rowbias.data_ptr() if rowbias is not None else None
    a.rowbias_period = period
    if residual is not None:
        a.residual, a.ldr, a.res_fp32 = residual.data_ptr(), residual.shape[1], int(residual.dtype == torch.float32)
    a.relu, a.out_fp32 = relu, int(out_fp32)
    if conv:
        for k, v in conv.items():
            setattr(a, k, v)
    a.argmax_out = keys.data_ptr() if argmax else None
    a.mn_major, a.ldb = mn_major, ldb
    native.check(lib.retr_op_gemm(dtype, C.byref(a), native.stream_ptr(dev)))
    torch.cuda.synchronize()
    idx = None
    if argmax:
        idx = (0xFFFFFFFF - (keys & 0xFFFFFFFF))
    return Cout, idx


def ref_epilogue(acc, scale, bias, rowbias, period, residual, relu):
    M = acc.shape[0]
    if scale is not None:
        acc = acc * scale
    if bias is not None:
        acc = acc + bias
    if rowbias is not None:
        acc = acc + rowbias[torch.arange(M, device=acc.device) % period]
    if relu == 1:
        acc = acc.relu()
    if residual is not None:
        acc = acc + residual.float()
    if relu == 2:
        acc = acc.relu()
    return acc


def tol(dtype, K):
    # fp32: re-association noise; bf16: inputs are exactly representable (generated in bf16), so the
    # only error is fp32 accumulation order (+ bf16 rounding of the output where C is bf16)
    return 2e-5 if dtype == FP32 else 2e-5


@pytest.mark.parametrize("dtype", [FP32, BF16])
@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (300, 256, 256), (1000, 768, 256), (257, 2048, 256),
                                   (256, 256, 2048), (130, 72, 160), (4096, 64, 576), (64, 1000, 512)])
def test_gemm_plain(dtype, M, N, K):
    torch.manual_seed(M + N + K)
    dev = "cuda"
    tdt = torch.float32 if dtype == FP32 else torch.bfloat16
    A = torch.randn(M, K, device=dev).to(tdt)
    W = (torch.randn(N, K, device=dev) / math.sqrt(K)).to(tdt)
    bias = torch.randn(N, device=dev)
    got, _ = run_gemm(dtype, A, W, M=M, N=N, K=K, bias=bias, out_fp32=True)
    ref = A.double() @ W.double().t() + bias.double()
    err = (got.double() - ref).abs().max().item()
    assert err < 1e-4, err


@pytest.mark.parametrize("dtype", [FP32, BF16])
def test_gemm_epilogue_variants(dtype):
    torch.manual_seed(7)
    dev = "cuda"
    tdt = torch.float32 if dtype == FP32 else torch.bfloat16
    M, N, K, period = 397 * 3, 768, 256, 397
    A = torch.randn(M, K, device=dev).to(tdt)
    W = (torch.randn(N, K, device=dev) / math.sqrt(K)).to(tdt)
    scale = torch.rand(N, device=dev) + 0.5
    bias = torch.randn(N, device=dev)
    rowbias = torch.randn(period, N, device=dev)
    res32 = torch.randn(M, N, device=dev)
    acc = (A.double() @ W.double().t()).float()
    for kw in (dict(scale=scale, bias=bias, relu=1),
               dict(rowbias=rowbias, period=period),
               dict(bias=bias, residual=res32, relu=2),
               dict(scale=scale, bias=bias, residual=res32.to(tdt), relu=2),
               dict(bias=bias, residual=res32, relu=0)):
        got, _ = run_gemm(dtype, A, W, M=M, N=N, K=K, out_fp32=True, **kw)
        ref = ref_epilogue(acc, kw.get("scale"), kw.get("bias"), kw.get("rowbias"), kw.get("period", 1),
                           kw.get("residual"), kw.get("relu", 0))
        assert (got - ref).abs().max().item() < 2e-4, kw.keys()
    # native output type (bf16 in bf16 mode): rounding of the result only
    for kw in (dict(bias=bias, relu=1), dict(scale=scale, bias=bias, residual=res32.to(tdt), relu=2),
               dict(rowbias=rowbias, period=period)):
        got, _ = run_gemm(dtype, A, W, M=M, N=N, K=K, **kw)
        ref = ref_epilogue(acc, kw.get("scale"), kw.get("bias"), kw.get("rowbias"), kw.get("period", 1),
                           kw.get("residual"), kw.get("relu", 0))
        assert bool(((got.float() - ref).abs() <= ref.abs() * 2 ** -8 + 2e-4).all()), kw.keys()
    # in-place residual (x += ...), as the transformer blocks use it
    x = res32.clone()
    got, _ = run_gemm(dtype, A, W, M=M, N=N, K=K, bias=bias, residual=x, out_fp32=True, out=x)
    ref = ref_epilogue(acc, None, bias, None, 1, res32, 0)
    assert got.data_ptr() == x.data_ptr() and (x - ref).abs().max().item() < 2e-4


@pytest.mark.parametrize("dtype", [FP32, BF16])
def test_gemm_vocab_argmax_and_ragged_n(dtype):
    """N = 30522 (not a multiple of any tile), fused argmax == torch.argmax of the written logits."""
    torch.manual_seed(11)
    dev = "cuda"
    tdt = torch.float32 if dtype == FP32 else torch.bfloat16
    M, N, K = 70, 30522, 512
    A = torch.randn(M, K, device=dev).to(tdt)
    W = (torch.randn(N, K, device=dev) / math.sqrt(K)).to(tdt)
    bias = torch.randn(N, device=dev) * 0.1
    got, idx = run_gemm(dtype, A, W, M=M, N=N, K=K, bias=bias, out_fp32=True, argmax=True)
    ref = (A.double() @ W.double().t() + bias.double()).float()
    assert (got - ref).abs().max().item() < 1e-4
    assert torch.equal(idx, got.argmax(-1))
    # argmax only (C == NULL): the decode path never writes logits
    _, idx2 = run_gemm(dtype, A, W, M=M, N=N, K=K, bias=bias, argmax=True, want_c=False)
    assert torch.equal(idx2, idx)
    # ties -> first index, as torch.argmax
    A0 = torch.zeros(M, K, device=dev).to(tdt)
    _, idx3 = run_gemm(dtype, A0, W, M=M, N=N, K=K, argmax=True, want_c=False)
    assert int(idx3.max()) == 0


CONVS = [  # (Cin, Cout, k, stride, pad, dil, H)
    (64, 64, 3, 1, 1, 1, 56), (128, 128, 3, 2, 1, 1, 56), (256, 512, 1, 2, 0, 1, 56),
    (512, 512, 3, 1, 2, 2, 14), (256, 256, 3, 1, 1, 1, 14), (64, 256, 1, 1, 0, 1, 20),
    (128, 128, 3, 2, 1, 1, 15), (64, 64, 3, 1, 1, 1, 9)]    # odd sides: rows that wrap mid-tile, strided tail


@pytest.mark.parametrize("dtype", [FP32, BF16])
@pytest.mark.parametrize("out_fp32", [True, False])
@pytest.mark.parametrize("cin,cout,k,stride,pad,dil,H", CONVS)
def test_conv_implicit_gemm(dtype, out_fp32, cin, cout, k, stride, pad, dil, H):
    """NHWC implicit-GEMM conv + folded BN + residual + ReLU vs F.conv2d (NCHW fp32)."""
    torch.manual_seed(cin + cout + k + stride + H)
    dev = "cuda"
    tdt = torch.float32 if dtype == FP32 else torch.bfloat16
    n = 3
    x = torch.randn(n, cin, H, H, device=dev).to(tdt)
    w = (torch.randn(cout, cin, k, k, device=dev) / math.sqrt(cin * k * k)).to(tdt)
    scale = torch.rand(cout, device=dev) + 0.5
    shift = torch.randn(cout, device=dev) * 0.1
    ref = F.conv2d(x.double(), w.double(), stride=stride, padding=pad, dilation=dil)
    OH = ref.shape[-1]
    res = torch.randn(n, OH, OH, cout, device=dev).to(tdt)
    ref = (ref * scale.double().view(1, -1, 1, 1) + shift.double().view(1, -1, 1, 1)).permute(0, 2, 3, 1)
    ref = (ref + res.double()).relu().float().reshape(-1, cout)
    A = x.permute(0, 2, 3, 1).contiguous()                       # NHWC
    Wk = w.permute(0, 2, 3, 1).contiguous().reshape(cout, k * k * cin)   # O,(kh,kw,I)
    conv = dict(kh=k, kw=k, stride=stride, pad=pad, dil=dil, H=H, W_in=H, C_in=cin, OH=OH, OW=OH)
    got, _ = run_gemm(dtype, A, Wk, M=n * OH * OH, N=cout, K=k * k * cin, lda=cin, scale=scale, bias=shift,
                      residual=res.reshape(-1, cout), relu=2, out_fp32=out_fp32, conv=conv)
    if out_fp32 or dtype == FP32:
        assert (got - ref).abs().max().item() < 2e-4
    else:       # bf16 result: one rounding of the fp32 value (staged TMA-store epilogue, bf16 residual by TMA)
        assert bool(((got.float() - ref).abs() <= ref.abs() * 2 ** -8 + 2e-4).all())


@pytest.mark.parametrize("M,N,K", [(300, 128, 512), (1000, 64, 256), (256, 512, 2048)])
def test_gemm_relu_backward_epilogue(M, N, K):
    """relu == 3: the residual operand is the forward activation, the (scaled) result is kept where it was > 0 — the
    ReLU + FrozenBN backward (dx * (out > 0) * scale[c]) folded into the input-gradient GEMM."""
    torch.manual_seed(M + N + K)
    dev = "cuda"
    A = torch.randn(M, K, device=dev).bfloat16()
    W = (torch.randn(N, K, device=dev) / math.sqrt(K)).bfloat16()
    act = torch.randn(M, N, device=dev).relu().bfloat16()           # about half zeros, like a ReLU output
    scale = torch.rand(N, device=dev) + 0.5
    ref = (A.float() @ W.float().t()) * scale * (act.float() > 0)
    got, _ = run_gemm(BF16, A, W, M=M, N=N, K=K, scale=scale, residual=act, relu=3)
    assert bool(((got.float() - ref).abs() <= ref.abs() * 2 ** -8 + 2e-4).all())
    assert bool((got.float()[act.float() <= 0] == 0).all())


@pytest.mark.parametrize("M,N,K", [(256, 256, 4096), (200, 576, 1000), (64, 64, 128), (1000, 520, 300), (128, 2304, 25088)])
def test_gemm_mn_major(M, N, K):
    """C = A^T B with A [K,M] and B [K,N] read in place (MN-major tcgen05 operands): the weight-gradient contraction
    of engine.py:76 loss.backward() for a Linear, without transposed copies.  Ragged M, N and K."""
    torch.manual_seed(M + N + K)
    dev = "cuda"
    At = torch.randn(K, M, device=dev).bfloat16()
    Bm = torch.randn(K, N, device=dev).bfloat16()
    ref = (At.double().t() @ Bm.double()).float()
    got, _ = run_gemm(BF16, At, Bm, M=M, N=N, K=K, lda=M, out_fp32=True, mn_major=1, ldb=N)
    assert (got - ref).abs().max().item() <= 2e-5 * ref.abs().max().item() * max(1.0, math.sqrt(K) / 8)


@pytest.mark.parametrize("cin,cout,k,stride,pad,dil,H", [(64, 64, 3, 1, 1, 1, 14), (128, 128, 3, 2, 1, 1, 15),
                                                         (64, 128, 3, 1, 2, 2, 14), (64, 256, 1, 2, 0, 1, 20),
                                                         (256, 64, 3, 1, 1, 1, 9)])
def test_conv_wgrad_mn_major(cin, cout, k, stride, pad, dil, H):
    """dW[C_out, (kh,kw,C_in)] = dY^T . im2col(X) with dY [pixels, C_out] and the NHWC input read in place (TMA im2col
    as the MN-major B operand) vs the autograd weight gradient of F.conv2d."""
    torch.manual_seed(cin + cout + k + stride + H)
    dev = "cuda"
    n = 3
    x = torch.randn(n, cin, H, H, device=dev).bfloat16()
    w = torch.zeros(cout, cin, k, k, device=dev, dtype=torch.double, requires_grad=True)
    y = F.conv2d(x.double(), w, stride=stride, padding=pad, dilation=dil)
    OH = y.shape[-1]
    dy = torch.randn(n, OH, OH, cout, device=dev).bfloat16()
    y.backward(dy.double().permute(0, 3, 1, 2))
    ref = w.grad.permute(0, 2, 3, 1).reshape(cout, k * k * cin).float()
    A = x.permute(0, 2, 3, 1).contiguous()
    conv = dict(kh=k, kw=k, stride=stride, pad=pad, dil=dil, H=H, W_in=H, C_in=cin, OH=OH, OW=OH)
    got, _ = run_gemm(BF16, dy.reshape(-1, cout), A, M=cout, N=k * k * cin, K=n * OH * OH, lda=cout, out_fp32=True,
                      conv=conv, mn_major=1)
    assert (got - ref).abs().max().item() <= 1e-4 * ref.abs().max().item()


@pytest.mark.parametrize("out_bf16", [0, 1])
@pytest.mark.parametrize("eps", [1e-5, 1e-12])
def test_layernorm(out_bf16, eps):
    native, lib = _lib()
    torch.manual_seed(3)
    x = torch.randn(1000, 256, device="cuda") * 3 + 1
    g, b = torch.randn(256, device="cuda"), torch.randn(256, device="cuda")
    y = torch.empty(1000, 256, device="cuda", dtype=torch.bfloat16 if out_bf16 else torch.float32)
    native.check(lib.retr_op_layernorm(x.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(), 1000, 256, eps,
                                       out_bf16, native.stream_ptr()))
    ref = F.layer_norm(x, (256,), g, b, eps)
    if out_bf16:      # one bf16 rounding of the fp32 result
        assert bool(((y.float() - ref).abs() <= ref.abs() * 2 ** -8 + 1e-6).all())
    else:
        assert (y - ref).abs().max().item() < 2e-5


def ref_attention(q, k, v, H, key_pad, causal):
    B, Lq, d = q.shape
    Lk, dh = k.shape[1], d // H
    qh = q.double().view(B, Lq, H, dh).transpose(1, 2) / math.sqrt(dh)
    kh = k.double().view(B, Lk, H, dh).transpose(1, 2)
    vh = v.double().view(B, Lk, H, dh).transpose(1, 2)
    s = qh @ kh.transpose(-1, -2)
    if causal:
        s = s + torch.full((Lq, Lk), float("-inf"), device=q.device, dtype=s.dtype).triu(1)
    if key_pad is not None:
        s = s.masked_fill(key_pad.bool()[:, None, None, :], float("-inf"))
    p = s.softmax(-1)
    return (p @ vh).transpose(1, 2).reshape(B, Lq, d).float(), p.mean(1).float()


@pytest.mark.parametrize("dtype", [FP32, BF16])
@pytest.mark.parametrize("Lq,Lk,causal,d,H", [(397, 397, 0, 256, 8), (128, 128, 1, 256, 8), (128, 397, 0, 256, 8),
                                              (70, 70, 1, 512, 8)])
def test_attention(dtype, Lq, Lk, causal, d, H):
    native, lib = _lib()
    torch.manual_seed(Lq + Lk + d)
    dev, B = "cuda", 3
    tdt = torch.float32 if dtype == FP32 else torch.bfloat16
    q = torch.randn(B, Lq, d, device=dev).to(tdt)
    kv = torch.randn(B, Lk, 2 * d, device=dev).to(tdt)           # [k | v] rows, pitch 2d
    key_pad = torch.zeros(B, Lk, dtype=torch.uint8, device=dev)
    key_pad[0, Lk // 2:] = 1
    key_pad[2, 5:9] = 1
    if causal:
        key_pad[:, 0] = 0
        key_pad[0] = 0
        key_pad[0, Lk - 10:] = 1                                  # trailing pads only, like id masks
    out = torch.empty(B, Lq, d, device=dev, dtype=tdt)
    probs = torch.empty(B, Lq, Lk, device=dev)
    k, v = kv[:, :, :d], kv[:, :, d:]
    native.check(lib.retr_op_attention(dtype, q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), B, Lq, Lk, H,
                                       d, 2 * d, 2 * d, d, key_pad.data_ptr(), causal, probs.data_ptr(),
                                       native.stream_ptr()))
    torch.cuda.synchronize()
    ref, ref_p = ref_attention(q, k, v, H, key_pad, causal)
    valid = ~torch.isnan(ref)
    assert (out.float()[valid] - ref[valid]).abs().max().item() < (2e-5 if dtype == FP32 else 2e-2)
    pv = ~torch.isnan(ref_p)
    assert (probs[pv] - ref_p[pv]).abs().max().item() < 2e-5


# ---- fused sublayers of the decode step (cluster kernels, csrc/ffn_fused.cu) and the compacted cross-attention ----

def _bf(t):
    return t.to(torch.bfloat16)


@pytest.mark.parametrize("M", [2, 32, 33, 128, 256])
def test_decode_ffn_fused(M):
    """y + W2 relu(W1 LN(y) + b1) + b2 (models/transformer_modules.py:84-97) on the operands the kernel sees (bf16
    weights, bf16 LN output and hidden activations, fp32 accumulation).  Rows beyond a 32-row block (M = 2, 33) cover
    the tail handling; two launches must be bit-identical (fixed-order cluster reduction)."""
    native, lib = _lib()
    torch.manual_seed(M)
    d, ff = 256, 2048
    y0 = torch.randn(M, d, device="cuda") * 2
    g, b = torch.rand(d, device="cuda") + 0.5, torch.randn(d, device="cuda") * 0.1
    W1, b1 = _bf(torch.randn(ff, d, device="cuda") / 16), torch.randn(ff, device="cuda") * 0.1
    W2, b2 = _bf(torch.randn(d, ff, device="cuda") / 45), torch.randn(d, device="cuda") * 0.1
    outs = []
    for _ in range(2):
        y = y0.clone()
        native.check(lib.retr_op_decode_ffn(y.data_ptr(), g.data_ptr(), b.data_ptr(), 1e-5, W1.data_ptr(), b1.data_ptr(),
                                            W2.data_ptr(), b2.data_ptr(), M, ff, native.stream_ptr()))
        outs.append(y)
    assert torch.equal(outs[0], outs[1])
    a = _bf(F.layer_norm(y0, (d,), g, b, 1e-5)).float()
    h = _bf(torch.relu(a @ W1.float().t() + b1)).float()
    ref = y0 + h @ W2.float().t() + b2
    assert (outs[0] - ref).abs().max().item() < 2e-3 * ref.abs().max().item()


@pytest.mark.parametrize("M,with_rowbias", [(2, True), (32, False), (45, True), (128, True)])
def test_decode_out_ln_proj_fused(M, with_rowbias):
    """y <- y + Wo att + bo, q = Wq LN(y) + rowbias[*row] (models/transformer_modules.py:58-83, pre-norm)."""
    native, lib = _lib()
    torch.manual_seed(100 + M)
    d = 256
    att = _bf(torch.randn(M, d, device="cuda"))
    y0 = torch.randn(M, d, device="cuda") * 2
    Wo, bo = _bf(torch.randn(d, d, device="cuda") / 16), torch.randn(d, device="cuda") * 0.1
    Wq = _bf(torch.randn(d, d, device="cuda") / 16)
    g, b = torch.rand(d, device="cuda") + 0.5, torch.randn(d, device="cuda") * 0.1
    rb = torch.randn(7, d, device="cuda")
    row = torch.tensor([5], dtype=torch.int32, device="cuda")
    y = y0.clone()
    q = torch.zeros(M, d, device="cuda", dtype=torch.bfloat16)
    native.check(lib.retr_op_decode_out_ln_proj(att.data_ptr(), Wo.data_ptr(), bo.data_ptr(), y.data_ptr(), g.data_ptr(),
                                                b.data_ptr(), 1e-5, Wq.data_ptr(), None,
                                                rb.data_ptr() if with_rowbias else None,
                                                row.data_ptr() if with_rowbias else None, q.data_ptr(), M,
                                                native.stream_ptr()))
    y_ref = y0 + att.float() @ Wo.float().t() + bo
    assert (y - y_ref).abs().max().item() < 1e-4 * y_ref.abs().max().item()
    a = _bf(F.layer_norm(y_ref, (d,), g, b, 1e-5)).float()
    q_ref = a @ Wq.float().t() + (rb[5] if with_rowbias else 0.0)
    assert (q.float() - q_ref).abs().max().item() < 2e-2 * q_ref.abs().max().item()


def test_decode_cross_attention_compacted_equals_masked():
    """Dropping the padded keys from the cache (key_compaction) is exact up to fp32 summation order: the kernel over
    n_keys[m] dense rows against a float64 softmax over the unpadded keys of the full cache."""
    native, lib = _lib()
    torch.manual_seed(9)
    M, H, dh, S = 6, 8, 32, 397
    d = H * dh
    q = _bf(torch.randn(M, d, device="cuda"))
    K = _bf(torch.randn(M, H, S, dh, device="cuda"))
    V = _bf(torch.randn(M, H, S, dh, device="cuda"))
    pad = torch.rand(M, S, device="cuda") < 0.3
    pad[0] = False                                   # one sample without padding
    pad[1, 5:] = True                                # one with five keys
    n_keys = (~pad).sum(1).to(torch.int32)
    Kc, Vc = torch.zeros_like(K), torch.zeros_like(V)
    for m in range(M):
        idx = (~pad[m]).nonzero().squeeze(1)
        Kc[m, :, :len(idx)] = K[m, :, idx]
        Vc[m, :, :len(idx)] = V[m, :, idx]
    out = torch.zeros(M, d, device="cuda", dtype=torch.bfloat16)
    native.check(lib.retr_op_decode_cross_attention(q.data_ptr(), Kc.data_ptr(), Vc.data_ptr(), S, out.data_ptr(), M, H, d,
                                                    n_keys.data_ptr(), native.stream_ptr()))
    qh = q.double().view(M, H, 1, dh) / math.sqrt(dh)
    s = (qh @ K.double().transpose(-1, -2)).masked_fill(pad[:, None, None, :], float("-inf"))
    ref = (s.softmax(-1) @ V.double()).reshape(M, d).float()
    assert bool(((out.float() - ref).abs() <= ref.abs() * 2 ** -8 + 2e-3).all())


@pytest.mark.parametrize("Lq,Lk,causal", [(397, 397, 0), (128, 128, 1), (130, 70, 0), (1, 397, 0)])
def test_attention_tcgen05_kernel(Lq, Lk, causal, monkeypatch):
    """attention_tc.cu (tcgen05.mma + TMEM, two heads per CTA as block-diagonal MMAs, head dim 32) behind
    RETR_ATTENTION_TC=1, against the float64 reference and the default mma.sync flash kernel on the same operands."""
    native, lib = _lib()
    torch.manual_seed(Lq * 3 + Lk)
    dev, B, H, d = "cuda", 3, 8, 256
    q = torch.randn(B, Lq, 3 * d, device=dev).to(torch.bfloat16)       # q | k | v side by side, pitch 3d (encoder layout)
    kv = torch.randn(B, Lk, 3 * d, device=dev).to(torch.bfloat16)
    key_pad = torch.zeros(B, Lk, dtype=torch.uint8, device=dev)
    key_pad[0, Lk // 2:] = 1
    if Lk > 9:
        key_pad[2, 5:9] = 1
    if causal:
        key_pad[:] = 0
        key_pad[0, Lk - 10:] = 1
    outs = []
    for flag in ("1", "0"):
        monkeypatch.setenv("RETR_ATTENTION_TC", flag)
        out = torch.zeros(B, Lq, d, device=dev, dtype=torch.bfloat16)
        native.check(lib.retr_op_attention(BF16, q.data_ptr(), kv[:, :, d:].data_ptr(), kv[:, :, 2 * d:].data_ptr(),
                                           out.data_ptr(), B, Lq, Lk, H, 3 * d, 3 * d, 3 * d, d, key_pad.data_ptr(), causal,
                                           None, native.stream_ptr()))
        torch.cuda.synchronize()
        outs.append(out.float())
    ref, _ = ref_attention(q[:, :, :d], kv[:, :, d:2 * d], kv[:, :, 2 * d:], H, key_pad, causal)
    valid = ~torch.isnan(ref)
    assert (outs[0][valid] - ref[valid]).abs().max().item() < 2e-2
    assert (outs[1][valid] - ref[valid]).abs().max().item() < 2e-2
