"""Prints the clock64 phase trace of the attention kernel.
Build the traced variant first:  touch eventgpt_b200/csrc/attention_sm100.cu && make -C eventgpt_b200/csrc EXTRA=-DEGPT_ATTN_TRACE
(and rebuild without EXTRA afterwards)."""
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
tr = torch.zeros(2048, dtype=torch.int64, device=dev)
for _ in range(3):
    check(fn(E.ptr(qkv), F, E.ptr(out), E.ptr(tr), E.stream_ptr(dev)))
torch.cuda.synchronize()
t = tr.cpu().tolist(); t0 = min(x for x in t if x > 0)
rel = lambda x: (x - t0) if x else None
def deltas(row):
    return [row[i + 1] - row[i] if row[i + 1] is not None and row[i] is not None else None for i in range(len(row) - 1)]
names = ["s_full wait", "ld", "max/rescale", "exps+st issue", "st_wait+arrive"]
for X in range(2):
    nb = 15
    print(f"softmax group {X} (warp {4 * X} lane 0): block start, deltas {names}, then merge [start, tile done, O loaded + o_free, computed + staged, TMA issued]")
    for n in range(nb):
        row = [rel(x) for x in t[X * 512 + n * 16: X * 512 + n * 16 + 14]]
        print(f"  n={n:2d} start {row[0]}  {deltas(row[:6])}  end {row[5]}  merge {[row[k] for k in (10, 7, 8, 13, 9)] if row[7] else ''}")
for X in range(2):
    nb = 15
    print(f"issuer {X}: [S issue start, S committed] [pv: p_full wait start, got, committed]")
    for n in range(nb):
        print(f"  n={n:2d}", [rel(x) for x in t[1024 + X * 256 + n * 8: 1024 + X * 256 + n * 8 + 5]])
