import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eventgpt_b200 import events as E
from eventgpt_b200._lib import lib, check
dev = torch.device("cuda:0"); torch.manual_seed(0)
F = 10
qkv = (torch.randn(F * 577, 3072, device=dev) * 0.5).bfloat16()
out = torch.empty(F * 577, 1024, device=dev, dtype=torch.bfloat16)
fn = lib().egpt_attention_trace; fn.restype = C.c_int; fn.argtypes = [C.c_void_p] * 2 and [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
tr = torch.zeros(512, dtype=torch.int64, device=dev)
for _ in range(3):
    check(fn(E.ptr(qkv), F, E.ptr(out), E.ptr(tr), E.stream_ptr(dev)))
torch.cuda.synchronize()
t = tr.cpu().tolist()
t0 = min(x for x in t if x > 0)
names = ["top", "s_full", "ld_done", "max+xchg", "exps+st", "p_arrive", "o_full", "end"]
print("softmax warp 4 (cycles since first stamp; deltas)")
for g in range(12):
    row = t[g * 16: g * 16 + 8]
    rel = [(x - t0) if x else None for x in row]
    d = []
    prev = rel[0]
    for x in rel[1:]:
        d.append(None if (x is None or prev is None) else x - prev)
        if x is not None: prev = x
    print(f"g={g:2d} start {rel[0]:7d}  " + "  ".join(f"{n}:{v}" for n, v in zip(names[1:], d)))
print("MMA thread: S_issue_start, S_issued(+), pv_wait_start(+), p_full_got(+), pv_issued(+)")
for g in range(12):
    row = t[256 + g * 8: 256 + g * 8 + 5]
    rel = [(x - t0) if x else None for x in row]
    print(f"g={g:2d} ", rel)
