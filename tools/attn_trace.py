"""Prints the clock64 phase trace of the attention kernel (build attention_sm100.cu with -DEGPT_ATTN_TRACE)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eventgpt_b200 import events as E
from eventgpt_b200._lib import lib, check
dev = torch.device("cuda:0"); torch.manual_seed(0)
F = 10
qkv = (torch.randn(F * 577, 3072, device=dev) * 0.5).bfloat16()
out = torch.empty(F * 577, 1024, device=dev, dtype=torch.bfloat16)
fn = lib().egpt_attention_trace; fn.restype = C.c_int; fn.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
tr = torch.zeros(1024, dtype=torch.int64, device=dev)
for _ in range(3):
    check(fn(E.ptr(qkv), F, E.ptr(out), E.ptr(tr), E.stream_ptr(dev)))
torch.cuda.synchronize()
t = tr.cpu().tolist(); t0 = min(x for x in t if x > 0)
rel = lambda x: (x - t0) if x else None
names = ["top", "s_full", "ld", "max(+o_done)", "exps+st_issue", "st_wait+arrive"]
for X in range(2):
    print(f"softmax set {X} (row group 0): start, then deltas {names[1:]}")
    for g in range(15):
        row = [rel(x) for x in t[X * 256 + g * 16: X * 256 + g * 16 + 6]]
        d = [row[i + 1] - row[i] if row[i + 1] is not None and row[i] is not None else None for i in range(5)]
        print(f"  g={g:2d} start {row[0]}  {d}")
print("S issuer: [wait s_free start, got, issued+committed]")
for g in range(15):
    print(f"  g={g:2d}", [rel(x) for x in t[512 + g * 4: 512 + g * 4 + 3]])
for X in range(2):
    print(f"PV issuer {X}: [wait p_full start, got, issued+committed]")
    for g in range(15):
        print(f"  g={g:2d}", [rel(x) for x in t[640 + X * 128 + g * 4: 640 + X * 128 + g * 4 + 3]])
