"""Device-side event operators (thin Python over the C ABI): time split, rasterisation,
CLIP preprocessing.  torch is used only for device memory and streams.

Reference functions replaced: common/common.py:17-37, 64-74, 76-107, 124-127.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import lib, check

CROP = 336
N_FRAMES = 5                      # hard-coded by the reference caller (common/common.py:119)
_TORCH_OF_NP = {np.dtype(np.uint8): torch.uint8, np.dtype(np.int8): torch.int8, np.dtype(np.uint16): torch.uint16,
                np.dtype(np.int16): torch.int16, np.dtype(np.uint32): torch.uint32, np.dtype(np.int32): torch.int32,
                np.dtype(np.uint64): torch.uint64, np.dtype(np.int64): torch.int64}
_CODE_OF_TORCH = {torch.uint8: 0, torch.int8: 1, torch.uint16: 2, torch.int16: 3, torch.uint32: 4, torch.int32: 5,
                  torch.uint64: 6, torch.int64: 7, torch.bool: 0}


def _require_cuda(device):
    if not torch.cuda.is_available():
        raise _lib.EgptError("eventgpt_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device(device if device is not None else "cuda")


def stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class Workspace:
    """Grow-only device scratch buffers, one set per (device, current CUDA stream, tag).  The library itself never
    allocates on the data path; this is the caller-side allocation it asks for.  The C ABI is re-entrant per
    (handle, stream, workspace) -- the reference's e2e benchmark runs prefills from two Python threads on two
    streams (pipeline/benchmark_e2e/benchmark_e2e_wallclock.py:641-649) -- so scratch is never shared between streams."""
    _bufs = {}

    @classmethod
    def get(cls, device, tag, nbytes):
        dev = torch.device(device)
        key = (dev.index or 0, int(torch.cuda.current_stream(dev).cuda_stream), tag)
        buf = cls._bufs.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = None
            cls._bufs.pop(key, None)
            buf = torch.empty(int(nbytes) + 1024, dtype=torch.uint8, device=device)
            cls._bufs[key] = buf
        return buf

    @classmethod
    def clear(cls):
        cls._bufs.clear()


def to_device(arr, device):
    """numpy integer array (any dtype the reference's .npy files use) -> device tensor, dtype preserved."""
    if isinstance(arr, torch.Tensor):
        return arr.to(device).contiguous()
    a = np.ascontiguousarray(arr)
    if a.dtype == np.bool_:
        a = a.view(np.uint8)
    if a.dtype not in _TORCH_OF_NP:
        raise TypeError(f"event arrays must be integer typed, got {a.dtype}")
    return torch.from_numpy(a).to(device, non_blocking=False)


def check_status(status, what):
    """Turns the device status word (a tensor, or its value already on the host) into the exceptions the reference raises."""
    flags = int(status.item()) if isinstance(status, torch.Tensor) else int(status)
    if flags & _lib.FLAG_EMPTY_SLICE:
        # reference: y.max() on an empty slice -> ValueError (common/common.py:65)
        raise ValueError("zero-size array to reduction operation maximum which has no identity "
                         f"({what}: a frame slice holds no events)")
    if flags & _lib.FLAG_OUT_OF_RANGE:
        raise ValueError(f"{what}: an event coordinate lies outside sensor_size (or is negative)")
    if flags & _lib.FLAG_TOO_MANY_BINS:
        raise ValueError(f"{what}: the stream spans more time bins than max_windows")
    if flags & _lib.FLAG_FILTER_TOO_WIDE:
        raise ValueError(f"{what}: frame too large for the resize filter (32 taps)")


def clip_bounds(t, first_us, interval_us, step_us, n_clips):
    """egpt_clip_bounds: (start, end) event indices int64 [n_clips, 2] of the clips [first + c*step, first + c*step + interval)
    of a time-sorted device tensor ``t`` (build_my_egpt_dsec_seq.py:137-202 with step = interval // 2)."""
    out = torch.empty((n_clips, 2), dtype=torch.int64, device=t.device)
    check(lib().egpt_clip_bounds(ptr(t), _CODE_OF_TORCH[t.dtype], t.numel(), int(first_us), int(interval_us), int(step_us),
                                 n_clips, ptr(out), stream_ptr(t.device)), "egpt_clip_bounds")
    return out


def ms_window_bounds(ms_to_idx, first_ms, window_ms, step_ms, n_windows, status=None):
    """egpt_ms_window_bounds: (start, end) int64 [n_windows, 2] from DSEC's ``ms_to_idx`` device tensor (dataset/io.py:68-87)."""
    dev = ms_to_idx.device
    out = torch.empty((n_windows, 2), dtype=torch.int64, device=dev)
    if status is None:
        status = torch.zeros(1, dtype=torch.int32, device=dev)
    check(lib().egpt_ms_window_bounds(ptr(ms_to_idx), _CODE_OF_TORCH[ms_to_idx.dtype], ms_to_idx.numel(), int(first_ms),
                                      int(window_ms), int(step_ms), n_windows, ptr(out), ptr(status), stream_ptr(dev)),
          "egpt_ms_window_bounds")
    return out, status


def raster(x, y, p, window_offsets, n_windows, n_frames, H, W, mode=_lib.RASTER_LAST_WINS, fixed_canvas=False,
           counts=False, status=None, ranges=False):
    """egpt_raster on device tensors.  Returns (frames uint8 [F, stride], hw int32 [F,2], counts|None, status).
    ``ranges=True``: ``window_offsets`` is int64 [n_windows, 2] of (start, end) pairs (egpt_raster_ranges)."""
    dev = x.device
    F = n_windows * n_frames
    stride = lib().egpt_raster_frame_stride(H, W)
    frames = torch.empty((F, stride), dtype=torch.uint8, device=dev)
    hw = torch.empty((F, 2), dtype=torch.int32, device=dev)
    cnt = torch.empty((F, 2, H, W), dtype=torch.int32, device=dev) if counts else None
    if status is None:
        status = torch.zeros(1, dtype=torch.int32, device=dev)
    nws = lib().egpt_raster_workspace_bytes(n_windows, n_frames, H, W)
    ws = Workspace.get(dev, "raster", nws)
    fn = lib().egpt_raster_ranges if ranges else lib().egpt_raster
    check(fn(ptr(x), ptr(y), ptr(p), _CODE_OF_TORCH[x.dtype], _CODE_OF_TORCH[y.dtype],
             _CODE_OF_TORCH[p.dtype], ptr(window_offsets), n_windows, n_frames, mode, H, W,
             1 if fixed_canvas else 0, ptr(frames), ptr(hw), ptr(cnt), ptr(status), ptr(ws), ws.numel(),
             stream_ptr(dev)), "egpt_raster")
    return frames, hw, cnt, status


def preprocess(frames, hw, max_h, want_pixel_values=True, want_patches=False, status=None):
    """egpt_preprocess: frames uint8 [F, stride] + hw -> bf16 [F,3,336,336] (and/or im2col patches)."""
    dev = frames.device
    F = frames.shape[0]
    pix = torch.empty((F, 3, CROP, CROP), dtype=torch.bfloat16, device=dev) if want_pixel_values else None
    patches = torch.empty((F * 576, 640), dtype=torch.bfloat16, device=dev) if want_patches else None
    if status is None:
        status = torch.zeros(1, dtype=torch.int32, device=dev)
    nws = lib().egpt_preprocess_workspace_bytes(F, max_h)
    ws = Workspace.get(dev, "preprocess", nws)
    check(lib().egpt_preprocess(ptr(frames), frames.stride(0), ptr(hw), F, max_h, ptr(pix), ptr(patches), ptr(status),
                                ptr(ws), ws.numel(), stream_ptr(dev)), "egpt_preprocess")
    return pix, patches, status


def split_time(x, y, t, p, interval, max_windows=1024, status=None):
    """egpt_split_time on device tensors.  Returns (x', y', t', p', offsets int64 [max_windows+1], n_windows int32[1], status)."""
    dev = t.device
    n = t.numel()
    ox, oy, ot, op = (torch.empty_like(a) for a in (x, y, t, p))
    offsets = torch.zeros(max_windows + 1, dtype=torch.int64, device=dev)
    nwin = torch.zeros(1, dtype=torch.int32, device=dev)
    if status is None:
        status = torch.zeros(1, dtype=torch.int32, device=dev)
    nws = lib().egpt_split_time_workspace_bytes(n, max_windows)
    ws = Workspace.get(dev, "split", nws)
    check(lib().egpt_split_time(ptr(x), ptr(y), ptr(t), ptr(p), _CODE_OF_TORCH[x.dtype], _CODE_OF_TORCH[y.dtype],
                                _CODE_OF_TORCH[t.dtype], _CODE_OF_TORCH[p.dtype], n, int(interval), ptr(ox), ptr(oy),
                                ptr(ot), ptr(op), ptr(offsets), ptr(nwin), max_windows, ptr(status), ptr(ws), ws.numel(),
                                stream_ptr(dev)), "egpt_split_time")
    return ox, oy, ot, op, offsets, nwin, status


def frames_to_numpy(frames, hw):
    """Device frame buffer -> list of uint8[H,W,3] ndarrays (what the reference returns)."""
    hw_h = hw.cpu().numpy()
    buf = frames.cpu().numpy()
    out = []
    for f, (h, w) in enumerate(hw_h):
        out.append(buf[f, : int(h) * int(w) * 3].reshape(int(h), int(w), 3).copy())
    return out
