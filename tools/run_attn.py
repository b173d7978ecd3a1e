"""Runs the attention kernel alone (profiling target)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eventgpt_b200 import events as E
from eventgpt_b200._lib import lib, check
F = int(sys.argv[1]) if len(sys.argv) > 1 else 37
dev = torch.device("cuda:0")
torch.manual_seed(0)
qkv = (torch.randn(F * 577, 3072, device=dev) * 0.5).bfloat16()
out = torch.empty(F * 577, 1024, device=dev, dtype=torch.bfloat16)
for _ in range(3):
    check(lib().egpt_attention(E.ptr(qkv), F, E.ptr(out), E.stream_ptr(dev)))
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
