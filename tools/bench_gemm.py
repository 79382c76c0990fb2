"""Micro-benchmark of the MLP building blocks (CUDA events, L2 flushed by size)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
from trase_b200 import _lib
from trase_b200._deform_native import _gemm, _wgrad
lib = _lib.load()
dev = torch.device("cuda:0")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 300000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
A = torch.randn(M, 256, device=dev); B = torch.randn(256, 256, device=dev) * 0.05
D = torch.empty(M, 256, device=dev); Z = torch.randn(M, 256, device=dev); bias = torch.randn(256, device=dev)
cs = torch.zeros(256, device=dev); dW = torch.zeros(256, 256, device=dev)
mk = torch.randint(-2**31, 2**31 - 1, (M, 8), dtype=torch.int32, device=dev)
A2 = torch.randn(M, 256, device=dev)  # alternate input so consecutive launches do not hit L2

def timeit(fn, n=reps):
    for _ in range(3): fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for i in range(n): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


t = timeit(lambda i: _gemm(lib, A if i & 1 else A2, B, D, bias=bias, relu=True, mask_out=mk))
print("gemm fwd   (bias,relu)      %7.1f us  %6.0f GB/s  %6.1f TFLOP/s" % (t, 2*M*256*4/t/1e3, 2*M*256*256/t/1e6))
t = timeit(lambda i: _gemm(lib, A if i & 1 else A2, B, D, mask_in=mk, colsum=cs))
print("gemm dgrad (mask,colsum)    %7.1f us  %6.0f GB/s" % (t, 3*M*256*4/t/1e3))
t = timeit(lambda i: _wgrad(lib, A if i & 1 else A2, Z, dW))
print("wgrad 256x256               %7.1f us  %6.0f GB/s  %6.1f TFLOP/s" % (t, 2*M*256*4/t/1e3, 2*M*256*256/t/1e6))
t = timeit(lambda i: torch.matmul(A if i & 1 else A2, B.t(), out=D))
print("torch fp32 matmul           %7.1f us" % t)
torch.backends.cuda.matmul.allow_tf32 = True
t = timeit(lambda i: torch.matmul(A if i & 1 else A2, B.t(), out=D))
print("torch tf32 matmul           %7.1f us" % t)
t = timeit(lambda i: D.copy_(A if i & 1 else A2))
print("copy 2x307MB                %7.1f us  %6.0f GB/s" % (t, 2*M*256*4/t/1e3))
