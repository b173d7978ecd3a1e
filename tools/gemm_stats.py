import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eventgpt_b200 import events as E
from eventgpt_b200._lib import lib, check
dev = torch.device("cuda:0"); torch.manual_seed(0)
fn = lib().egpt_gemm_stats; fn.restype = C.c_int
fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
for (M, N, K, epi) in [(92320, 3072, 1024, 0), (92320, 1024, 1024, 0), (92320, 1024, 1024, 3), (92320, 4096, 1024, 1), (92320, 1024, 4096, 0), (92320, 1024, 4096, 3), (92320, 4096, 4096, 0)]:
    A = torch.randn(M, K, device=dev).bfloat16(); B = torch.randn(N, K, device=dev).bfloat16(); bias = torch.randn(N, device=dev)
    out = torch.zeros(M, N, device=dev, dtype=torch.float32 if epi >= 3 else torch.bfloat16)
    st = torch.zeros(148 * 8, dtype=torch.int64, device=dev)
    for _ in range(3):
        check(fn(E.ptr(A), E.ptr(B), E.ptr(bias), E.ptr(out), M, N, K, epi, E.ptr(st), E.stream_ptr(dev)))
    torch.cuda.synchronize()
    s = st.view(148, 8).float().mean(0).tolist()
    tiles = ((M + 127) // 128) * (N // 256) / 148
    print(f"M={M} N={N} K={K} epi={epi}: MMA total {s[3]:.0f} clk/CTA ({tiles:.1f} tiles, ideal {tiles*K/64*512:.0f}); MMA wait data {s[1]:.0f} ({100*s[1]/s[3]:.1f}%), wait TMEM {s[2]:.0f} ({100*s[2]/s[3]:.1f}%); producer wait free stage {s[0]:.0f}")
