"""Times egpt_attention alone (one tower layer's worth: 160 frames x 16 heads x 577 tokens), CUDA events, for three inputs:
random-init-like scores (flat softmax, no rescale), a ramp that raises every row's maximum block after block (the lazy
rescale of the TMEM accumulator runs on EVERY block: the worst case), and sharply peaked rows (one late rescale)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eventgpt_b200 import _lib, events as E
dev = torch.device("cuda:0")
F = 160
l = _lib.lib()
def run(name, qkv):
    x = qkv.reshape(F * 577, 3072).bfloat16().to(dev)
    out = torch.empty(F * 577, 1024, device=dev, dtype=torch.bfloat16)
    call = lambda: _lib.check(l.egpt_attention(E.ptr(x), F, E.ptr(out), E.stream_ptr(dev)), "attention")
    for _ in range(3): call()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): call()
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    print(f"{name:34s} {ms*1e3:7.1f} us per layer  {F*16*4*577*577*64/ms/1e9:5.0f} TFLOP/s")
g = torch.Generator().manual_seed(0)
base = torch.randn(F, 577, 3, 16, 64, generator=g) * 0.3
run("flat (random-init scale)", base)
ramp = base.clone()
r = torch.linspace(0.0, 6.0, 577)[None, :, None, None]
sign = ramp[:, :1, 0].sign()
ramp[:, :, 1] = ramp[:, :, 1] + r * sign
ramp[:, :, 0] = ramp[:, :, 0].abs() * sign
run("rising keys (rescale every block)", ramp)
peak = base.clone()
peak[:, :, 0] = peak[:, :, 0].abs()
peak[:, 570, 1] = 8.0
run("one dominant late key", peak)
