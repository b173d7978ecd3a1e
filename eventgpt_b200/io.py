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
