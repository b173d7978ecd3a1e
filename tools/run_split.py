"""Times egpt_split_time on a 32 M-event stream of 32 consecutive 50 ms windows (t int64, x/y uint16, p uint8), events
resident on the device; CUDA events around the whole call (egpt_profile_collect).  Bytes per event: t read three times
(min/max, count, scatter) + x, y, t, p moved once = 24 + 26; SURVEY 8(d)'s figure is 16 + 26 = 42 B/event."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eventgpt_b200 import events as E, _lib
dev = torch.device("cuda:0")
B, N = 32, 1_000_000
rng = np.random.default_rng(8)
t = np.sort(rng.integers(0, B * 50_000, B * N)).astype(np.int64) + 1_700_000_000_000
x = torch.from_numpy(rng.integers(0, 640, B * N).astype(np.uint16)).to(dev)
y = torch.from_numpy(rng.integers(0, 480, B * N).astype(np.uint16)).to(dev)
p = torch.from_numpy(rng.integers(0, 2, B * N).astype(np.uint8)).to(dev)
td = torch.from_numpy(t).to(dev)
for mw in (64, 4096):
    for _ in range(3):
        out = E.split_time(x, y, td, p, 50_000, mw)
    torch.cuda.synchronize()
    assert int(out[5].item()) == B and int(out[6].item()) == 0
    off = out[4][:B + 1].cpu().numpy()
    want = np.searchsorted(t, 1_700_000_000_000 + 50_000 * np.arange(B + 1))
    assert np.array_equal(off, want) and torch.equal(out[2], td) and torch.equal(out[0], x)
    _lib.profile_enable(True)
    for _ in range(10):
        E.split_time(x, y, td, p, 50_000, mw)
    ms = _lib.profile_collect()["split"][0] / 10
    _lib.profile_enable(False)
    n = B * N
    print(f"max_windows {mw}: {ms*1e3:.1f} us  {n/ms/1e3:.0f} Mev/s  {42*n/ms/1e6:.0f} GB/s at 42 B/event "
          f"({42*n/ms/1e6/6543.4:.3f} of 6543 GB/s); bytes really moved 50 B/event -> {50*n/ms/1e6:.0f} GB/s")
