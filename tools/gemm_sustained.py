"""Sustained (power-capped) throughput of the GEMM kernel vs cuBLAS on the tower's shapes: ~3 s of back-to-back launches each,
SM clock sampled with nvidia-smi meanwhile.  The step is power limited, so this -- not the burst number -- is what matters."""
import os, subprocess, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eventgpt_b200 import events as E
from eventgpt_b200._lib import lib, check

dev = torch.device("cuda:0")
torch.manual_seed(0)


def clocks_during(fn, seconds):
    p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "100", "-i", "0"],
                         stdout=subprocess.PIPE, text=True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 0
    t0 = time.time()
    a.record()
    while time.time() - t0 < seconds:
        for _ in range(20):
            fn()
        n += 20
        torch.cuda.synchronize()
    b.record(); torch.cuda.synchronize()
    p.terminate()
    rows = [l.split(",") for l in p.stdout.read().strip().splitlines() if "," in l]
    clk = sorted(float(r[0]) for r in rows)[len(rows) // 2] if rows else 0
    pw = sorted(float(r[1]) for r in rows)[len(rows) // 2] if rows else 0
    return a.elapsed_time(b) / n, clk, pw


for (M, N, K, epi, name) in [(92320, 4096, 1024, 1, "fc1"), (92320, 1024, 4096, 3, "fc2"), (92320, 3072, 1024, 0, "qkv"), (92320, 1024, 1024, 3, "out")]:
    A = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
    B = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
    bias = torch.randn(N, device=dev)
    out = torch.zeros(M, N, device=dev, dtype=torch.float32 if epi >= 3 else torch.bfloat16)
    ours = lambda: check(lib().egpt_gemm_bf16(E.ptr(A), E.ptr(B), E.ptr(bias), E.ptr(out), M, N, K, epi, E.stream_ptr(dev)))
    ref_out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    cublas = lambda: torch.matmul(A, B.t(), out=ref_out)
    for label, fn in (("ours", ours), ("cuBLAS", cublas), ("ours", ours)):
        ms, clk, pw = clocks_during(fn, 3.0)
        print(f"{name:4s} M={M} N={N} K={K} {label:7s}: {ms:.3f} ms  {2 * M * N * K / ms / 1e9:6.0f} TFLOP/s  SM {clk:.0f} MHz  {pw:.0f} W", flush=True)
