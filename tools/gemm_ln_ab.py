"""Per-shape A/B of the plain and the LayerNorm-folding GEMM epilogues at the tower's shapes (M = 160 x 577 rows)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eventgpt_b200 import _lib
l = _lib.lib()
l.egpt_gemm_bf16_ln.restype = C.c_int
l.egpt_gemm_bf16_ln.argtypes = [C.c_void_p] * 4 + [C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
dev = torch.device("cuda:0")
M = 160 * 577
st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
stats = torch.rand(M, 8, device=dev) + 1.0
def run(name, N, K, epi):
    A = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
    B = (torch.randn(N, K, device=dev) * 0.02).bfloat16()
    bias = torch.randn(N, device=dev)
    colsum = torch.randn(N, device=dev)
    f32 = epi in (3, 4, 5)
    out = torch.zeros(M, N, device=dev, dtype=torch.float32 if f32 else torch.bfloat16)
    out2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    def call():
        rc = l.egpt_gemm_bf16_ln(A.data_ptr(), B.data_ptr(), bias.data_ptr(), out.data_ptr(), M, N, K, epi, out2.data_ptr(),
                                 colsum.data_ptr(), stats.data_ptr(), st)
        assert rc == 0, l.egpt_last_error()
    for _ in range(3): call()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): call()
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    print(f"{name:8s} N={N} K={K} epi={epi}: {ms*1e3:7.1f} us  {2*M*N*K/ms/1e9:6.0f} TFLOP/s")
for name, N, K, epis in (("out-proj", 1024, 1024, (3, 5)), ("fc2", 1024, 4096, (3, 5)), ("qkv", 3072, 1024, (0, 6)), ("fc1", 4096, 1024, (1, 7))):
    for e in epis: run(name, N, K, e)
