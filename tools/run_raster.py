"""Runs the rasteriser alone on BASELINE config 2 shapes (profiling / timing target)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eventgpt_b200 import events as E, _lib
dev = torch.device("cuda:0")
rng = np.random.default_rng(2)
B, N, H, W = 32, 1_000_000, 480, 640
x = torch.from_numpy(rng.integers(0, W, B * N).astype(np.uint16)).to(dev)
y = torch.from_numpy(rng.integers(0, H, B * N).astype(np.uint16)).to(dev)
p = torch.from_numpy(rng.integers(0, 2, B * N).astype(np.uint8)).to(dev)
off = torch.arange(0, B + 1, dtype=torch.int64, device=dev) * N
for _ in range(3):
    E.raster(x, y, p, off, B, 5, H, W)
torch.cuda.synchronize()
_lib.profile_enable(True)
for _ in range(10):
    E.raster(x, y, p, off, B, 5, H, W)
pr = _lib.profile_collect()
ms = pr["raster"][0] / 10
print(f"raster (events on device, CUDA events around the whole call): {ms*1e3:.1f} us  {B*N/ms/1e3:.0f} Mev/s  "
      f"{B*(5*N+15*H*W)/ms/1e6:.0f} GB/s algorithmic")
