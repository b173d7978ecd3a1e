"""Drop-in for the event half of the reference's ``common/common.py``.

Same names, arguments, return types and error behaviour as the reference
functions (cited per function); the arithmetic runs in libegpt_b200.so on the GPU.
"""
import numpy as np
import torch

from . import _lib
from . import events as E


def _capacity(x, y, sensor_size):
    """Canvas capacity.  The reference sizes the canvas from the data (y.max()+1, x.max()+1,
    common/common.py:65); on the numpy surface the same host-side max is used unless the
    caller states the sensor size."""
    if sensor_size is not None:
        return int(sensor_size[0]), int(sensor_size[1])
    return int(np.max(y)) + 1, int(np.max(x)) + 1       # ValueError on empty input, as upstream


def generate_event_image(x, y, p, sensor_size=None, device=None):
    """common/common.py:64-74: white canvas, blue where the last event at a pixel has p == 0,
    red otherwise; returns uint8[H, W, 3] with H, W = max y + 1, max x + 1."""
    dev = E._require_cuda(device)
    x, y, p = np.asarray(x), np.asarray(y), np.asarray(p)
    H, W = _capacity(x, y, sensor_size)
    n = len(x)
    xd, yd, pd = (E.to_device(a, dev) for a in (x, y, p))
    offsets = torch.tensor([0, n], dtype=torch.int64, device=dev)
    frames, hw, _, status = E.raster(xd, yd, pd, offsets, 1, 1, H, W)
    E.check_status(status, "generate_event_image")
    return E.frames_to_numpy(frames, hw)[0]


def get_event_images_list(event_npy, n, sensor_size=None, device=None):
    """common/common.py:17-37: n equal-count slices (the last takes the remainder), one frame each."""
    dev = E._require_cuda(device)
    x, y, p, t = (np.asarray(event_npy[k]) for k in ("x", "y", "p", "t"))
    total = len(t)
    if total // n == 0 and n > 1:
        # slice 0 is empty: the reference fails in y.max() (common/common.py:65)
        raise ValueError("zero-size array to reduction operation maximum which has no identity")
    H, W = _capacity(x[:total], y[:total], sensor_size)
    xd, yd, pd = (E.to_device(a[:total], dev) for a in (x, y, p))
    offsets = torch.tensor([0, total], dtype=torch.int64, device=dev)
    frames, hw, _, status = E.raster(xd, yd, pd, offsets, 1, n, H, W)
    E.check_status(status, "get_event_images_list")
    return E.frames_to_numpy(frames, hw)


def split_event_by_time(event_npy, time_interval=50000, max_windows=4096, device=None):
    """common/common.py:76-107: list of {'p','t','x','y'} dicts, one per non-empty time bin
    (t // time_interval), ascending, event order preserved inside a bin.
    ``max_windows`` bounds the SPAN of bins between the earliest and the latest event (empty bins in between count); a
    recording with a long gap that exceeds it is split again with the span it really needs, so the result is the
    reference's either way."""
    dev = E._require_cuda(device)
    p, t, x, y = (np.asarray(event_npy[k]) for k in ("p", "t", "x", "y"))
    if len(t) == 0:
        return []
    xd, yd, td, pd = (E.to_device(a, dev) for a in (x, y, t, p))
    ox, oy, ot, op, offsets, nwin, status = E.split_time(xd, yd, td, pd, time_interval, max_windows)
    if int(status.item()) & E._lib.FLAG_TOO_MANY_BINS:
        span = int(t.max()) // int(time_interval) - int(t.min()) // int(time_interval) + 1
        if span > (1 << 20):
            raise ValueError(f"split_event_by_time: the stream spans {span} time bins (limit 2^20)")
        ox, oy, ot, op, offsets, nwin, status = E.split_time(xd, yd, td, pd, time_interval, span)
    E.check_status(status, "split_event_by_time")
    nw = int(nwin.item())
    off = offsets[: nw + 1].cpu().numpy()
    hx, hy, ht, hp = (a.cpu().numpy() for a in (ox, oy, ot, op))
    out = []
    for i in range(nw):
        s, e = int(off[i]), int(off[i + 1])
        out.append({"p": hp[s:e].astype(p.dtype, copy=False), "t": ht[s:e], "x": hx[s:e], "y": hy[s:e]})
    return out


def process_event_data(event_frame_path, event_processor, device, sensor_size=None):
    """common/common.py:110-129: .npy -> 5 event frames -> CLIP preprocessing -> list of five
    bf16 [3,336,336] tensors on ``device``; returns ([H, W] of frame 0, that list).

    ``event_processor`` is accepted for signature compatibility; the resize / crop / normalise
    it would perform on the CPU (CLIP ViT-L/14-336 settings) runs on the GPU instead."""
    event_npy = np.load(event_frame_path, allow_pickle=True)
    event_npy = np.array(event_npy).item()
    dev = E._require_cuda(device)
    x, y, p, t = (np.asarray(event_npy[k]) for k in ("x", "y", "p", "t"))
    total = len(t)
    if total // E.N_FRAMES == 0:
        raise ValueError("zero-size array to reduction operation maximum which has no identity")
    H, W = _capacity(x, y, sensor_size)
    xd, yd, pd = (E.to_device(a, dev) for a in (x, y, p))
    offsets = torch.tensor([0, total], dtype=torch.int64, device=dev)
    frames, hw, _, status = E.raster(xd, yd, pd, offsets, 1, E.N_FRAMES, H, W)
    pix, _, status = E.preprocess(frames, hw, H, status=status)
    E.check_status(status, "process_event_data")
    event_image_size = [int(v) for v in hw[0].cpu().tolist()]
    return event_image_size, [pix[i] for i in range(E.N_FRAMES)]
