"""Oracle (test infrastructure): numpy restatement of the reference's event
splitting and rasterisation.  See ``oracle/__init__.py`` for the usage rule.

Follows ``/root/reference/common/common.py``:
  * ``get_event_images_list``  : lines 17-37
  * ``generate_event_image``   : lines 64-74
  * ``split_event_by_time``    : lines 76-107
and the second raster mode of
``/root/reference/feasible/my_egpt_dsec_dataset/preprocess_event_images.py:58-90``.
"""
import numpy as np

WHITE = 255
RED = (255, 0, 0)    # p != 0   (common/common.py:72; -1 is painted red too)
BLUE = (0, 0, 255)   # p == 0   (common/common.py:70)


def frame_bounds(total_events, n):
    """Slice i = [i*k, (i+1)*k), k = N // n; the last slice takes the remainder
    (common/common.py:21-27)."""
    k = total_events // n
    starts = [i * k for i in range(n)]
    ends = [(i + 1) * k if i < n - 1 else total_events for i in range(n)]
    return starts, ends


def generate_event_image_loop(x, y, p):
    """Literal restatement of common/common.py:64-74 (per-event loop, last
    event at a pixel wins).  Only for small inputs."""
    height, width = int(y.max()) + 1, int(x.max()) + 1   # ValueError on empty, as upstream
    img = np.full((height, width, 3), WHITE, dtype=np.uint8)
    for x_, y_, p_ in zip(x, y, p):
        if p_ == 0:
            img[y_, x_] = BLUE
        else:
            img[y_, x_] = RED
    return img


def generate_event_image(x, y, p):
    """Vectorised last-write-wins restatement of common/common.py:64-74.

    The last occurrence of every linear pixel index is found with ``np.unique``
    on the reversed index array; proven equal to the loop form on every golden
    fixture (tests/test_oracle_golden.py)."""
    x = np.asarray(x)
    y = np.asarray(y)
    p = np.asarray(p)
    height, width = int(y.max()) + 1, int(x.max()) + 1   # ValueError on empty, as upstream
    lin = y.astype(np.int64) * width + x.astype(np.int64)
    n = lin.shape[0]
    uniq, first_rev = np.unique(lin[::-1], return_index=True)
    last = n - 1 - first_rev
    img = np.full((height * width, 3), WHITE, dtype=np.uint8)
    red = p[last] != 0
    img[uniq[red]] = RED
    img[uniq[~red]] = BLUE
    return img.reshape(height, width, 3)


def generate_event_image_negwins(x, y, p, height=None, width=None):
    """Second raster mode (preprocess_event_images.py:58-90): all p==1 events are
    painted red, then all p==0 events blue (negative wins); other p are ignored;
    optional explicit canvas size."""
    x = np.asarray(x)
    y = np.asarray(y)
    p = np.asarray(p)
    if height is None:
        height = int(y.max()) + 1
    if width is None:
        width = int(x.max()) + 1
    img = np.full((height, width, 3), WHITE, dtype=np.uint8)
    xi = x.astype(np.int32)
    yi = y.astype(np.int32)
    pi = p.astype(np.int32)
    img[yi[pi == 1], xi[pi == 1]] = RED
    img[yi[pi == 0], xi[pi == 0]] = BLUE
    return img


def get_event_images_list(event_npy, n, raster=generate_event_image):
    """common/common.py:17-37: n equal-count slices, one frame each."""
    x, y, p, t = event_npy['x'], event_npy['y'], event_npy['p'], event_npy['t']
    starts, ends = frame_bounds(len(t), n)
    return [raster(x[s:e], y[s:e], p[s:e]) for s, e in zip(starts, ends)]


def split_event_by_time(event_npy, time_interval=50000):
    """common/common.py:76-107: bin = (t // dt) * dt; unique bins ascending,
    empty bins skipped; per bin a boolean-mask select (order preserved, works
    on unsorted t)."""
    p, t, x, y = event_npy['p'], event_npy['t'], event_npy['x'], event_npy['y']
    bins = (t // time_interval) * time_interval
    out = []
    for b in np.unique(bins):
        m = bins == b
        out.append({'p': p[m], 't': t[m], 'x': x[m], 'y': y[m]})
    return out


def split_event_by_time_clips(event_npy, time_interval=50000, image_timestamps=None):
    """feasible/my_egpt_dsec_dataset/build_my_egpt_dsec_seq.py:137-225: with image timestamps, sequential clips
    of ``time_interval`` us stepping by half a clip from the first image timestamp while the clip ends no later
    than the last one; a clip with no event is dropped; each clip carries the ascending event indices
    (``np.where(mask)``) and its metadata.  Without timestamps (or when no clip holds an event): fixed bins as in
    common/common.py, each with start / end / duration."""
    p, t, x, y = event_npy['p'], event_npy['t'], event_npy['x'], event_npy['y']
    if image_timestamps is not None and len(image_timestamps) > 0:
        out = []
        seq_start, seq_end = min(image_timestamps), max(image_timestamps)
        step = time_interval // 2
        clip_start, clip_idx = seq_start, 0
        while clip_start + time_interval <= seq_end:
            clip_end = clip_start + time_interval
            idx = np.where((t >= clip_start) & (t < clip_end))[0]
            if len(idx) > 0:
                imgs = [i for i, ts in enumerate(image_timestamps) if clip_start <= ts < clip_end]
                out.append({'indices': idx, 'time_start': clip_start, 'time_end': clip_end,
                            'duration_ms': time_interval / 1000000.0 * 1000.0, 'image_timestamp': clip_end,
                            'image_idx': imgs[0] if imgs else clip_idx, 'image_indices_in_clip': imgs,
                            'num_events': len(idx)})
            clip_start += step
            clip_idx += 1
        if out:
            return out
    bins = (t // time_interval) * time_interval
    out = []
    for b in np.unique(bins):
        m = bins == b
        out.append({'p': p[m], 't': t[m], 'x': x[m], 'y': y[m], 'time_start': b, 'time_end': b + time_interval,
                    'duration_ms': time_interval / 1000.0})
    return out
