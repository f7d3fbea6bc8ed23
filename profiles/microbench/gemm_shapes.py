"""Representative GEMM shapes of the hot path through retr_op_gemm (C ABI), timed with CUDA events back to back
(operands > L2 where the real ones are): the encoder's FF1 / FF2 / q|k|v on the packed rows and layer3's conv1 / conv3.
usage: python profiles/microbench/gemm_shapes.py [reps]"""
import ctypes as C, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from retr_b200 import native

lib = native.lib()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = torch.device("cuda")


def run(name, M, N, K, relu=0, scale=False, bias=True, res=None, out_fp32=False):
    A = torch.randn(M, K, device=dev).bfloat16()
    W = (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16()
    Cc = torch.empty(M, N, device=dev, dtype=torch.float32 if out_fp32 else torch.bfloat16)
    a = native.GemmArgs()
    a.A, a.W, a.C = A.data_ptr(), W.data_ptr(), Cc.data_ptr()
    a.M, a.N, a.K, a.lda, a.ldc = M, N, K, K, N
    keep = []
    if scale:
        s = torch.rand(N, device=dev) + 0.5; keep.append(s); a.scale = s.data_ptr()
    if bias:
        b = torch.randn(N, device=dev); keep.append(b); a.bias = b.data_ptr()
    a.rowbias_period = 1
    if res:
        r = torch.randn(M, N, device=dev, dtype=torch.float32 if res == "fp32" else torch.bfloat16)
        keep.append(r); a.residual, a.ldr, a.res_fp32 = r.data_ptr(), N, int(res == "fp32")
    a.relu, a.out_fp32 = relu, int(out_fp32)
    st = native.stream_ptr(dev)
    for _ in range(3):
        native.check(lib.retr_op_gemm(native.BF16, C.byref(a), st))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        native.check(lib.retr_op_gemm(native.BF16, C.byref(a), st))
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1000 / reps
    by = 2 * M * K + (4 if out_fp32 else 2) * M * N + ((4 if res == "fp32" else 2) * M * N if res else 0)
    print(json.dumps({"gemm": name, "M": M, "N": N, "K": K, "us": round(us, 1), "tflops": round(2 * M * N * K / us / 1e6, 1),
                      "gbs": round(by / us / 1e3)}))


run("enc_ff1", 73216, 2048, 256, relu=1)
run("enc_ff2", 73216, 256, 2048, res="fp32", out_fp32=True)
run("enc_qkv", 73216, 768, 256)
run("l3_conv3", 100352, 1024, 256, relu=2, scale=True, res="bf16")
run("l3_conv1", 100352, 256, 1024, relu=1, scale=True)
run("l4_conv3", 100352, 2048, 512, relu=2, scale=True, res="bf16")
