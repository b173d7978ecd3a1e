"""BASELINE config 4: rasterisation-only sweep, 10^4 .. 10^8 events per window at 1280x720 (Gen4-class sensor).
Timing with CUDA events (events resident on the device); parity: bit-exact vs the vectorised CPU oracle up to
10^7 events, and for 10^8 vs the oracle on the first and last frame only (the full check takes minutes on the CPU).
Writes a markdown table to stdout."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eventgpt_b200 import events as E, _lib
from oracle import events as oev          # checker only
dev = torch.device("cuda:0")
H, W = 720, 1280
peak = 6529.4
print("| events/window | device time | Mevents/s | algorithmic GB/s (5N + 15HW bytes) | of measured HBM peak | parity |")
print("|---:|---:|---:|---:|---:|---|")
for exp in range(4, 9):
    n = 10 ** exp
    rng = np.random.default_rng(exp)
    ev = dict(x=rng.integers(0, W, n).astype(np.uint16), y=rng.integers(0, H, n).astype(np.uint16),
              p=rng.integers(0, 2, n).astype(np.uint8), t=None)
    k = n // 5
    for i in range(5):
        ev["x"][i * k], ev["y"][i * k] = W - 1, H - 1
    x, y, p = (torch.from_numpy(ev[c]).to(dev) for c in "xyp")
    off = torch.tensor([0, n], dtype=torch.int64, device=dev)
    fr, hw, _, st = E.raster(x, y, p, off, 1, 5, H, W)
    E.check_status(st, "sweep")
    frames = E.frames_to_numpy(fr, hw)
    ev["t"] = np.zeros(n, np.int8)
    if exp <= 7:
        want = oev.get_event_images_list(ev, 5)
        ok = all(np.array_equal(a, b) for a, b in zip(frames, want))
        par = "bit-exact, 5 frames" if ok else "MISMATCH"
    else:
        ok = np.array_equal(frames[0], oev.generate_event_image(ev["x"][:k], ev["y"][:k], ev["p"][:k])) and \
             np.array_equal(frames[4], oev.generate_event_image(ev["x"][4 * k:], ev["y"][4 * k:], ev["p"][4 * k:]))
        par = "bit-exact, frames 0 and 4" if ok else "MISMATCH"
    for _ in range(3):
        E.raster(x, y, p, off, 1, 5, H, W)
    _lib.profile_enable(True)
    reps = 20 if exp < 8 else 5
    for _ in range(reps):
        E.raster(x, y, p, off, 1, 5, H, W)
    ms = _lib.profile_collect()["raster"][0] / reps
    _lib.profile_enable(False)
    by = 5 * n + 15 * H * W
    print(f"| 10^{exp} | {ms*1e3:.1f} us | {n/ms/1e3:.0f} | {by/ms/1e6:.0f} | {by/ms/1e6/peak:.3f} | {par} |", flush=True)

# ---- hot-pixel variant: Zipf-distributed pixels, to stress contention on single key-plane words ---------
print()
print("| events/window (Zipf a=1.2 over pixels) | hottest pixel's share | device time | Mevents/s | vs uniform | parity |")
print("|---:|---:|---:|---:|---:|---|")
uniform = {}
for exp in (6, 7):
    n = 10 ** exp
    for kind in ("uniform", "zipf"):
        rng = np.random.default_rng(100 + exp)
        if kind == "uniform":
            lin = rng.integers(0, H * W, n)
        else:
            perm = rng.permutation(H * W)
            lin = perm[np.minimum(rng.zipf(1.2, n) - 1, H * W - 1)]
        ev = dict(x=(lin % W).astype(np.uint16), y=(lin // W).astype(np.uint16), p=rng.integers(0, 2, n).astype(np.uint8),
                  t=np.zeros(n, np.int8))
        k = n // 5
        for i in range(5):
            ev["x"][i * k], ev["y"][i * k] = W - 1, H - 1
        x, y, p = (torch.from_numpy(ev[c]).to(dev) for c in "xyp")
        off = torch.tensor([0, n], dtype=torch.int64, device=dev)
        fr, hw, _, st = E.raster(x, y, p, off, 1, 5, H, W)
        E.check_status(st, "zipf")
        frames = E.frames_to_numpy(fr, hw)
        ok = all(np.array_equal(a, b) for a, b in zip(frames, oev.get_event_images_list(ev, 5)))
        for _ in range(3):
            E.raster(x, y, p, off, 1, 5, H, W)
        _lib.profile_enable(True)
        for _ in range(20):
            E.raster(x, y, p, off, 1, 5, H, W)
        ms = _lib.profile_collect()["raster"][0] / 20
        _lib.profile_enable(False)
        if kind == "uniform":
            uniform[exp] = ms
            continue
        share = np.bincount(lin).max() / n
        print(f"| 10^{exp} | {share:.3f} | {ms*1e3:.1f} us | {n/ms/1e3:.0f} | {uniform[exp]/ms:.2f}x | {'bit-exact, 5 frames' if ok else 'MISMATCH'} |", flush=True)
