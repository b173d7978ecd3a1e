"""egpt_raster_static: the rasteriser for callers that may not allocate device memory (for
example a sensor-acquisition or plug-in thread working inside buffers the host application
gave it).  It keeps its scratch planes in a fixed array inside the library instead of a caller
workspace.  This test drives it through the C ABI and requires the frames and canvas sizes to be
byte-identical to egpt_raster's AND to the CPU oracle's, on two calls in a row whose inputs differ
(so anything the first call left in the shared array would show up in the second); it checks that
a call leaves the amount of free device memory untouched (the property the entry point exists
for), repeats on a second GPU when there is one (the array's address is looked up per device),
and checks the size limit.

Kept in a file of its own that pytest collects last."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _events(seed, n, H, W, dev):
    rng = np.random.default_rng(seed)
    x = torch.from_numpy(rng.integers(0, W, n).astype(np.uint16)).to(dev)
    y = torch.from_numpy(rng.integers(0, H, n).astype(np.uint16)).to(dev)
    p = torch.from_numpy(rng.integers(0, 2, n).astype(np.uint8)).to(dev)
    return x, y, p


def _oracle_frames(x, y, p, sizes, n_frames):
    """The same windows through the CPU restatement of common/common.py:17-37, 64-74 (checker only)."""
    from oracle import events as oev
    xs, ys, ps = x.cpu().numpy(), y.cpu().numpy(), p.cpu().numpy()
    out, s = [], 0
    for n in sizes:
        ev = dict(x=xs[s:s + n], y=ys[s:s + n], p=ps[s:s + n], t=np.zeros(n, np.int8))
        out += oev.get_event_images_list(ev, n_frames)
        s += n
    return out


def _run_both(l, x, y, p, sizes, n_frames, H, W, dev):
    """(frames, hw) from the workspace entry point and from the no-allocation entry point."""
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    B = len(sizes)
    F = B * n_frames
    off = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int64, device=dev)
    stride = l.egpt_raster_frame_stride(H, W)
    outs = []
    for static in (False, True):
        frames = torch.zeros((F, stride), dtype=torch.uint8, device=dev)
        hw = torch.zeros((F, 2), dtype=torch.int32, device=dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        args = [x.data_ptr(), y.data_ptr(), p.data_ptr(), 2, 2, 0, off.data_ptr(), B, n_frames, 0, H, W, 0,
                frames.data_ptr(), hw.data_ptr()]
        if static:
            rc = l.egpt_raster_static(*args, status.data_ptr(), stream)      # warm-up: the device address look-up
            torch.cuda.synchronize(dev)
            assert rc == 0, l.egpt_last_error()
            frames.zero_(); hw.zero_()
            torch.cuda.synchronize(dev)
            free_before = torch.cuda.mem_get_info(dev)[0]
            rc = l.egpt_raster_static(*args, status.data_ptr(), stream)
            torch.cuda.synchronize(dev)
            assert torch.cuda.mem_get_info(dev)[0] == free_before, "egpt_raster_static changed the free device memory"
        else:
            ws = torch.empty(l.egpt_raster_workspace_bytes(B, n_frames, H, W), dtype=torch.uint8, device=dev)
            rc = l.egpt_raster(*args, None, status.data_ptr(), ws.data_ptr(), ws.numel(), stream)
        assert rc == 0, l.egpt_last_error()
        torch.cuda.synchronize(dev)
        assert int(status.item()) == 0
        hw_h = hw.cpu().numpy()
        packed = [frames[f, : int(h) * int(w) * 3].cpu().numpy() for f, (h, w) in enumerate(hw_h)]
        outs.append((hw_h, packed))
    return outs


def _devices():
    return [0] + ([1] if torch.cuda.device_count() > 1 else [])


@pytest.mark.parametrize("index", _devices() if torch.cuda.is_available() else [0])
def test_static_scratch_raster_matches_workspace_raster_twice(index):
    from eventgpt_b200 import _lib
    assert torch.cuda.is_available(), "-m gpu tests need a GPU"
    dev = torch.device("cuda", index)
    torch.cuda.set_device(dev)
    l = _lib.lib()
    assert l.egpt_raster_static_max_keys() == 5 * 720 * 1280
    # call 1: three VGA windows (fills most of the built-in array); call 2: one smaller, sparser window
    calls = [(11, [250_000, 90_001, 40_000], 5, 480, 640), (12, [7_003], 5, 300, 400)]
    for seed, sizes, n_frames, H, W in calls:
        x, y, p = _events(seed, sum(sizes), H, W, dev)
        (hw_a, fr_a), (hw_b, fr_b) = _run_both(l, x, y, p, sizes, n_frames, H, W, dev)
        want = _oracle_frames(x, y, p, sizes, n_frames)
        assert np.array_equal(hw_a, hw_b)
        assert len(fr_a) == len(sizes) * n_frames == len(want)
        for a, b, w, (h_, w_) in zip(fr_a, fr_b, want, hw_b):
            assert a.size > 0 and np.array_equal(a, b)          # integer work: bit-exact
            assert (int(h_), int(w_)) == w.shape[:2] and np.array_equal(b, w.reshape(-1))      # ... and equal to the oracle
    # beyond the built-in capacity: refused with a clear message, nothing launched
    x, y, p = _events(13, 1000, 480, 640, dev)
    off = torch.tensor([0, 250, 500, 750, 1000], dtype=torch.int64, device=dev)
    dummy = torch.zeros(16, dtype=torch.int32, device=dev)
    rc = l.egpt_raster_static(x.data_ptr(), y.data_ptr(), p.data_ptr(), 2, 2, 0, off.data_ptr(), 4, 5, 0, 480, 640, 0,
                              dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), None)
    assert rc == -5 and b"built-in scratch" in l.egpt_last_error()
    # four huge ints whose product does not fit 64 bits are refused the same way
    rc = l.egpt_raster_static(x.data_ptr(), y.data_ptr(), p.data_ptr(), 2, 2, 0, off.data_ptr(), 2 ** 31 - 1, 2 ** 31 - 1, 0,
                              2 ** 31 - 1, 2 ** 31 - 1, 0, dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), None)
    assert rc == -5
    torch.cuda.set_device(0)
