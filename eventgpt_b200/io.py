"""Upstream ingestion: the reference's DSEC slicing helpers, ``dataset/io.py:38-87``.

The reference reads DSEC ``events.h5`` files with h5py; the arithmetic is two lookups in the per-millisecond
index ``ms_to_idx`` and an ``int64`` offset add.  These mirrors take any mapping with the DSEC layout
(an open ``h5py.File`` or a dict of numpy arrays: ``events/{x,y,p,t}``, ``ms_to_idx``, ``t_offset``), so they
work without h5py on in-memory recordings; ``window_offsets_from_ms_index`` turns a whole recording into the
``window_offsets`` array that ``egpt_raster`` / ``egpt_visual_tokens`` consume, with no per-window copies."""
import math

import numpy as np


def _get(f, path):
    cur = f
    for part in path.split("/"):
        cur = cur[part]
    return cur


def _scalar(v):
    return int(np.asarray(v[()] if hasattr(v, "shape") and getattr(v, "shape", None) == () else v).reshape(-1)[0])


def _extract_by_index(f, ev_start_idx, ev_end_idx):
    """dataset/io.py:38-57"""
    ev = _get(f, "events")
    t_offset = _scalar(f["t_offset"])
    return {"p": np.asarray(ev["p"][ev_start_idx:ev_end_idx]),
            "t": np.asarray(ev["t"][ev_start_idx:ev_end_idx]).astype("int64") + t_offset,
            "x": np.asarray(ev["x"][ev_start_idx:ev_end_idx]),
            "y": np.asarray(ev["y"][ev_start_idx:ev_end_idx])}


def get_num_events(f):
    """dataset/io.py:60-62"""
    return len(_get(f, "events/t"))


def extract_from_h5_by_index(f, ev_start_idx, ev_end_idx):
    """dataset/io.py:64-66 (``f``: open file or dict instead of a path)."""
    return _extract_by_index(f, int(ev_start_idx), int(ev_end_idx))


def timewindow_to_indices(ms_to_idx, t_offset, t_last_us, t_min_us, t_max_us):
    """The index arithmetic of dataset/io.py:68-87: start = ms_to_idx[(t_min - t_offset) // 1000],
    end = ms_to_idx[floor((t_max - t_offset) / 1000)]; asserts t_max - t_offset <= t[-1] as upstream."""
    ms2idx = np.asarray(ms_to_idx, dtype="int64")
    start = int(ms2idx[(int(t_min_us) - int(t_offset)) // 1000])
    t_end = int(t_max_us) - int(t_offset)
    assert t_end <= int(t_last_us), (t_end, int(t_last_us))
    end = int(ms2idx[math.floor(t_end / 1000)])
    return start, end


def extract_from_h5_by_timewindow(f, t_min_us, t_max_us):
    """dataset/io.py:68-87 (``f``: open file or dict instead of a path)."""
    t = _get(f, "events/t")
    s, e = timewindow_to_indices(f["ms_to_idx"], _scalar(f["t_offset"]), t[-1], t_min_us, t_max_us)
    return _extract_by_index(f, s, e)


def window_offsets_from_ms_index(ms_to_idx, window_ms=50, first_ms=0, n_windows=None):
    """Offsets [n_windows + 1] of consecutive ``window_ms`` windows of a recording, straight from the
    per-millisecond index (window i = events [ms_to_idx[first + i*w], ms_to_idx[first + (i+1)*w]) )."""
    ms2idx = np.asarray(ms_to_idx, dtype="int64")
    avail = (len(ms2idx) - 1 - first_ms) // window_ms
    n = avail if n_windows is None else min(int(n_windows), avail)
    return ms2idx[first_ms + window_ms * np.arange(n + 1)]


def split_event_by_time_clips(event_npy, time_interval=50000, image_timestamps=None):
    """Mirror of ``split_event_by_time`` of the DSEC clip builder
    (feasible/my_egpt_dsec_dataset/build_my_egpt_dsec_seq.py:137-225): 50 %-overlapping clips anchored at the image
    timestamps, empty clips dropped, same metadata; fixed bins when there are no timestamps or no clip holds an event.
    Instead of one boolean mask over the whole recording per clip, the clip bounds come from two binary searches over the
    (sorted, or argsorted once) timestamps; each clip additionally carries ``start`` / ``end``, the slice of the
    time-ordered stream it covers, which is what ``clips_to_batch`` consumes."""
    t = np.asarray(event_npy["t"])
    if image_timestamps is not None and len(image_timestamps) > 0:
        ts = np.asarray(image_timestamps, dtype=np.int64)
        seq_start, seq_end = int(ts.min()), int(ts.max())
        step = time_interval // 2
        n_clips = max(0, (seq_end - time_interval - seq_start) // step + 1) if seq_end - seq_start >= time_interval else 0
        starts_us = seq_start + step * np.arange(n_clips, dtype=np.int64)
        is_sorted = bool(np.all(t[1:] >= t[:-1])) if len(t) > 1 else True
        order = None if is_sorted else np.argsort(t, kind="stable")
        ts_sorted = t if is_sorted else t[order]
        lo = np.searchsorted(ts_sorted, starts_us, side="left")
        hi = np.searchsorted(ts_sorted, starts_us + time_interval, side="left")
        ts_order = np.argsort(ts, kind="stable")
        ts_asc = ts[ts_order]
        out = []
        for c in range(n_clips):
            s, e = int(lo[c]), int(hi[c])
            if e <= s:
                continue
            cs, ce = int(starts_us[c]), int(starts_us[c]) + time_interval
            a, b = np.searchsorted(ts_asc, cs, side="left"), np.searchsorted(ts_asc, ce, side="left")
            imgs = sorted(int(i) for i in ts_order[a:b])
            idx = np.arange(s, e) if is_sorted else np.sort(order[s:e])
            out.append({"indices": idx, "start": s, "end": e, "time_start": cs, "time_end": ce,
                        "duration_ms": time_interval / 1000000.0 * 1000.0, "image_timestamp": ce,
                        "image_idx": imgs[0] if imgs else c, "image_indices_in_clip": imgs, "num_events": e - s})
        if out:
            return out
    p, x, y = np.asarray(event_npy["p"]), np.asarray(event_npy["x"]), np.asarray(event_npy["y"])
    bins = (t // time_interval) * time_interval
    out = []
    for b in np.unique(bins):
        m = bins == b
        out.append({"p": p[m], "t": t[m], "x": x[m], "y": y[m], "time_start": b, "time_end": b + time_interval,
                    "duration_ms": time_interval / 1000.0})
    return out


def clips_to_batch(event_npy, clips):
    """Overlapping clips -> one ragged batch for ``events.raster`` / ``visual_tokens``: the clips' events back to back
    (an event in the overlap of two clips appears in both) and the CSR offsets [len(clips) + 1]."""
    x, y, p = (np.asarray(event_npy[k]) for k in "xyp")
    idx = [c["indices"] for c in clips]
    offsets = np.zeros(len(clips) + 1, dtype=np.int64)
    np.cumsum([len(i) for i in idx], out=offsets[1:])
    cat = np.concatenate(idx) if idx else np.zeros(0, dtype=np.int64)
    return x[cat], y[cat], p[cat], offsets
