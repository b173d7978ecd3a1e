"""Out-of-bounds and race evidence without compute-sanitizer (it is closed on this GPU pool; `profiles/r02_sanitizer.md`
has the refusal).  Every buffer the integer kernels write -- frames, canvas sizes, status, scratch, split outputs,
pixel values -- is carved out of a larger allocation with 4 KiB canaries on both sides; after the call the canaries must be
intact and the results must equal the CPU oracle.  Each call is repeated and must reproduce its first result bit for bit
(an unordered shared-memory or global-memory race in these kernels shows up as a differing frame)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
CANARY, FILL = 4096, 0xA5


class Guarded:
    def __init__(self, nbytes, dev):
        self.n = int(nbytes)
        self.full = torch.full((self.n + 2 * CANARY,), FILL, dtype=torch.uint8, device=dev)
        self.view = self.full[CANARY:CANARY + self.n]

    def ptr(self):
        return C.c_void_p(self.view.data_ptr())

    def intact(self):
        return bool((self.full[:CANARY] == FILL).all() and (self.full[CANARY + self.n:] == FILL).all())


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


@pytest.mark.parametrize("H,W,sizes", [(480, 640, [250_001, 33, 70_000]), (260, 346, [9_999, 5]), (720, 1280, [400_000])])
def test_raster_writes_stay_inside_their_buffers_and_repeat(H, W, sizes):
    from eventgpt_b200 import _lib
    from oracle import events as oev
    dev = torch.device("cuda:0")
    l = _lib.lib()
    rng = np.random.default_rng(H)
    n = sum(sizes)
    ev = dict(x=rng.integers(0, W, n).astype(np.uint16), y=rng.integers(0, H, n).astype(np.uint16),
              p=rng.integers(0, 2, n).astype(np.uint8))
    x, y, p = (torch.from_numpy(ev[k]).to(dev) for k in "xyp")
    off = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int64, device=dev)
    B, F = len(sizes), len(sizes) * 5
    stride = l.egpt_raster_frame_stride(H, W)
    first = None
    for rep in range(4):
        frames, hw, status = Guarded(F * stride, dev), Guarded(F * 8, dev), Guarded(4, dev)
        ws = Guarded(l.egpt_raster_workspace_bytes(B, 5, H, W), dev)
        status.view.zero_()
        rc = l.egpt_raster(x.data_ptr(), y.data_ptr(), p.data_ptr(), 2, 2, 0, off.data_ptr(), B, 5, 0, H, W, 0, frames.ptr(), hw.ptr(),
                           None, status.ptr(), ws.ptr(), ws.n, _stream(dev))
        assert rc == 0, l.egpt_last_error()
        torch.cuda.synchronize(dev)
        assert frames.intact() and hw.intact() and status.intact() and ws.intact()
        assert int(status.view.view(torch.int32).item()) == 0
        hw_h = hw.view.view(torch.int32).reshape(F, 2).cpu().numpy()
        fr = frames.view.reshape(F, stride)
        got = [fr[f, : int(h) * int(w) * 3].cpu().numpy() for f, (h, w) in enumerate(hw_h)]
        if first is None:
            first = got
            s = 0
            for b, m in enumerate(sizes):
                want = oev.get_event_images_list({k: ev[k][s:s + m] for k in "xyp"} | {"t": np.zeros(m, np.int8)}, 5)
                for i in range(5):
                    assert tuple(hw_h[b * 5 + i]) == want[i].shape[:2] and np.array_equal(got[b * 5 + i], want[i].reshape(-1))
                s += m
        else:
            assert all(np.array_equal(a, b) for a, b in zip(first, got)), f"run {rep} differs from run 0"


def test_split_and_preprocess_writes_stay_inside_their_buffers():
    from eventgpt_b200 import _lib
    from oracle import events as oev, preprocess as opp
    dev = torch.device("cuda:0")
    l = _lib.lib()
    rng = np.random.default_rng(2)
    n = 300_000
    t = np.sort(rng.integers(0, 400_000, n)).astype(np.int64)
    ev = dict(x=rng.integers(0, 640, n).astype(np.uint16), y=rng.integers(0, 480, n).astype(np.uint16),
              p=rng.integers(0, 2, n).astype(np.uint8), t=t)
    x, y, p, td = (torch.from_numpy(ev[k]).to(dev) for k in "xypt")
    ox, oy, ot, op = Guarded(2 * n, dev), Guarded(2 * n, dev), Guarded(8 * n, dev), Guarded(n, dev)
    offs, nwin, status = Guarded(8 * 65, dev), Guarded(4, dev), Guarded(4, dev)
    ws = Guarded(l.egpt_split_time_workspace_bytes(n, 64), dev)
    status.view.zero_()
    rc = l.egpt_split_time(x.data_ptr(), y.data_ptr(), td.data_ptr(), p.data_ptr(), 2, 2, 7, 0, n, 50_000, ox.ptr(), oy.ptr(), ot.ptr(),
                           op.ptr(), offs.ptr(), nwin.ptr(), 64, status.ptr(), ws.ptr(), ws.n, _stream(dev))
    assert rc == 0, l.egpt_last_error()
    torch.cuda.synchronize(dev)
    for g in (ox, oy, ot, op, offs, nwin, status, ws):
        assert g.intact()
    want = oev.split_event_by_time(ev, 50_000)
    assert int(nwin.view.view(torch.int32).item()) == len(want) == 8
    assert np.array_equal(ot.view.view(torch.int64).cpu().numpy(), np.concatenate([w["t"] for w in want]))
    assert np.array_equal(ox.view.view(torch.uint16).cpu().numpy(), np.concatenate([w["x"] for w in want]))
    # raster -> preprocess on the split windows, guarded pixel output
    F = 8 * 5
    stride = l.egpt_raster_frame_stride(480, 640)
    frames = torch.empty((F, stride), dtype=torch.uint8, device=dev)
    hw = torch.empty((F, 2), dtype=torch.int32, device=dev)
    rws = torch.empty(l.egpt_raster_workspace_bytes(8, 5, 480, 640), dtype=torch.uint8, device=dev)
    rc = l.egpt_raster(ox.ptr(), oy.ptr(), op.ptr(), 2, 2, 0, offs.ptr(), 8, 5, 0, 480, 640, 0, frames.data_ptr(), hw.data_ptr(), None,
                       status.ptr(), rws.data_ptr(), rws.numel(), _stream(dev))
    assert rc == 0, l.egpt_last_error()
    pix = Guarded(F * 3 * 336 * 336 * 2, dev)
    pws = Guarded(l.egpt_preprocess_workspace_bytes(F, 480), dev)
    rc = l.egpt_preprocess(frames.data_ptr(), stride, hw.data_ptr(), F, 480, pix.ptr(), None, status.ptr(), pws.ptr(), pws.n, _stream(dev))
    assert rc == 0, l.egpt_last_error()
    torch.cuda.synchronize(dev)
    assert pix.intact() and pws.intact() and status.intact() and int(status.view.view(torch.int32).item()) == 0
    f0 = oev.get_event_images_list(want[0], 5)[0]
    got0 = pix.view.view(torch.int16)[: 3 * 336 * 336].cpu().numpy().view(np.uint16).reshape(3, 336, 336)
    assert np.array_equal(got0, opp.pixel_values_bf16_bits(f0))
