"""Prints the clock64 phase trace of the attention kernel (CTA 0, one head of its persistent loop).
Build the traced variant first:
  touch eventgpt_b200/csrc/attention_sm100.cu && make -C eventgpt_b200/csrc EXTRA="-DEGPT_ATTN_TRACE -DEGPT_ATTN_TRACE_IT=1"
(and rebuild without EXTRA afterwards).  Usage: attn_trace.py [frames]"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eventgpt_b200 import events as E
from eventgpt_b200._lib import lib, check
dev = torch.device("cuda:0"); torch.manual_seed(0)
F = int(sys.argv[1]) if len(sys.argv) > 1 else 40
qkv = (torch.randn(F * 577, 3072, device=dev) * 0.5).bfloat16()
out = torch.empty(F * 577, 1024, device=dev, dtype=torch.bfloat16)
fn = lib().egpt_attention_trace; fn.restype = C.c_int; fn.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
tr = torch.zeros(4096, dtype=torch.int64, device=dev)
for _ in range(3):
    check(fn(E.ptr(qkv), F, E.ptr(out), E.ptr(tr), E.stream_ptr(dev)))
torch.cuda.synchronize()
t = tr.cpu().tolist(); t0 = min(x for x in t if x > 0)
rel = lambda x: (x - t0) if x else None
def deltas(row):
    return [row[i + 1] - row[i] if row[i + 1] is not None and row[i] is not None else None for i in range(len(row) - 1)]
names = ["s_full wait", "ld", "max/rescale", "exps+st issue", "st_wait+arrive"]
for w in (0, 4):
    print(f"softmax warp {w} (group {w // 4}, rows 0-31) lane 0: block start, deltas {names}, block end")
    for n in range(15):
        row = [rel(x) for x in t[w * 256 + n * 16: w * 256 + n * 16 + 6]]
        print(f"  n={n:2d} start {row[0]}  {deltas(row[:6])}  end {row[5]}")
print("per softmax warp 0..7: [s_full wait, ld, max/rescale, exps+st, arrive] of blocks 3..8")
for n in range(3, 9):
    print(f"  n={n:2d}", [[t[w * 256 + n * 16 + k + 1] - t[w * 256 + n * 16 + k] for k in range(5)] for w in range(8)])
print("block start per softmax warp 0..7:")
for n in range(15):
    print(f"  n={n:2d}", [rel(t[w * 256 + n * 16 + 0]) for w in range(8)])
print("block end (P handed over) per softmax warp 0..7:")
for n in range(15):
    print(f"  n={n:2d}", [rel(t[w * 256 + n * 16 + 5]) for w in range(8)])
print("exps+st issue duration per softmax warp 0..7:")
for n in range(15):
    print(f"  n={n:2d}", [t[w * 256 + n * 16 + 4] - t[w * 256 + n * 16 + 3] for w in range(8)])
for g in range(2):
    print(f"issuer {g}: [S issue start, S committed] [pv: p_full wait start, got, committed]")
    for n in range(15):
        print(f"  n={n:2d}", [rel(x) for x in t[2048 + g * 256 + n * 8: 2048 + g * 256 + n * 8 + 5]])
print("merge warp 8 lane 0, per tile: [start (next tile's left-over dot), wait for group 0, group 0 done, O_0 read + o_free[0], group 1 done, O_1 read + o_free[1], TMA store issued]")
for n in range(16):
    row = [rel(t[3072 + n * 8 + k]) for k in (0, 1, 2, 3, 5, 6, 4)]
    if row[0] is not None: print(f"  T={n:2d}", row)
