"""dataset/io.py:38-87 mirrors on an in-memory DSEC-layout recording: the same slices as a
direct restatement of the reference arithmetic, and consistency of the window-offset helper."""
import math

import numpy as np
import pytest

from eventgpt_b200 import io as eio


def make_recording(seed=0, n=200_000, dur_ms=400):
    rng = np.random.default_rng(seed)
    t = np.sort(rng.integers(0, dur_ms * 1000, n)).astype(np.uint32)
    ms_to_idx = np.searchsorted(t, np.arange(dur_ms + 1) * 1000, side="left").astype(np.uint64)   # DSEC convention
    return {"events": {"x": rng.integers(0, 640, n).astype(np.uint16), "y": rng.integers(0, 480, n).astype(np.uint16),
                       "p": rng.integers(0, 2, n).astype(np.uint8), "t": t},
            "ms_to_idx": ms_to_idx, "t_offset": np.int64(51_000_000_000)}


def reference_timewindow(f, t_min_us, t_max_us):
    """literal restatement of dataset/io.py:68-87"""
    ms2idx = np.asarray(f["ms_to_idx"], dtype="int64")
    t_offset = f["t_offset"][()]
    t = f["events"]["t"]
    s = ms2idx[(t_min_us - t_offset) // 1000]
    t_end = t_max_us - t_offset
    assert t_end <= t[-1]
    e = ms2idx[math.floor(t_end / 1000)]
    return {k: (f["events"][k][s:e].astype("int64") + t_offset if k == "t" else f["events"][k][s:e]) for k in "ptxy"}


def test_timewindow_matches_reference_arithmetic():
    f = make_recording()
    off = int(f["t_offset"])
    for (a, b) in [(0, 50_000), (50_000, 100_000), (123_456, 299_999), (0, 1_000)]:
        got = eio.extract_from_h5_by_timewindow(f, off + a, off + b)
        want = reference_timewindow(f, off + a, off + b)
        for k in "ptxy":
            assert got[k].dtype == want[k].dtype and np.array_equal(got[k], want[k])
        assert got["t"].dtype == np.int64
    with pytest.raises(AssertionError):
        eio.extract_from_h5_by_timewindow(f, off, off + 10_000_000)


def test_window_offsets_feed_the_rasteriser_layout():
    f = make_recording(1)
    offs = eio.window_offsets_from_ms_index(f["ms_to_idx"], 50)
    assert len(offs) == 9 and offs[0] == 0 and np.all(np.diff(offs) >= 0)
    t = f["events"]["t"]
    for i in range(8):
        w = t[offs[i]:offs[i + 1]]
        assert w.size == 0 or (w.min() >= i * 50_000 and w.max() < (i + 1) * 50_000)
        sl = eio.extract_from_h5_by_index(f, offs[i], offs[i + 1])
        assert len(sl["t"]) == offs[i + 1] - offs[i] and eio.get_num_events(f) == 200_000


# ---- 50 %-overlap clip splitter (build_my_egpt_dsec_seq.py:137-225) ---------------------------------------------

def _clip_recording(seed, n=50_000, dur_us=600_000, sort=True, gap=None):
    rng = np.random.default_rng(seed)
    t = rng.integers(1_000_000, 1_000_000 + dur_us, n).astype(np.int64)
    if gap is not None:                                  # a stretch without events -> dropped clips
        t = t[(t < gap[0]) | (t >= gap[1])]
    if sort:
        t = np.sort(t)
    n = len(t)
    return {"x": rng.integers(0, 640, n).astype(np.uint16), "y": rng.integers(0, 480, n).astype(np.uint16),
            "p": rng.integers(0, 2, n).astype(np.uint8), "t": t}


@pytest.mark.parametrize("sort", [True, False])
@pytest.mark.parametrize("interval", [50_000, 33_333])
def test_overlap_clips_match_oracle(sort, interval):
    from oracle import events as OE
    ev = _clip_recording(3, sort=sort, gap=(1_200_000, 1_290_000))
    stamps = [1_000_000 + 50_000 * i for i in range(12)]
    stamps[3], stamps[7] = stamps[7], stamps[3]          # the reference only takes min / max and scans the list
    want = OE.split_event_by_time_clips(ev, interval, stamps)
    got = eio.split_event_by_time_clips(ev, interval, stamps)
    assert len(got) == len(want) and len(got) > 5
    assert any(w["num_events"] != want[0]["num_events"] for w in want)
    for g, w in zip(got, want):
        np.testing.assert_array_equal(g["indices"], w["indices"])
        for k in ("time_start", "time_end", "duration_ms", "image_timestamp", "image_idx", "image_indices_in_clip", "num_events"):
            assert g[k] == w[k], k
    # consecutive clips overlap by half a clip
    assert got[1]["time_start"] - got[0]["time_start"] == interval // 2
    x, y, p, off = eio.clips_to_batch(ev, got)
    assert off[-1] == sum(w["num_events"] for w in want) and len(x) == off[-1]
    for i, w in enumerate(want):
        np.testing.assert_array_equal(x[off[i]:off[i + 1]], ev["x"][w["indices"]])
        np.testing.assert_array_equal(p[off[i]:off[i + 1]], ev["p"][w["indices"]])


def test_overlap_clips_fallbacks():
    from oracle import events as OE
    ev = _clip_recording(4, n=5_000)
    for stamps in (None, [], [1_000_000, 1_020_000], [5_000_000, 5_200_000]):   # none / too short a span / no events inside
        want = OE.split_event_by_time_clips(ev, 50_000, stamps)
        got = eio.split_event_by_time_clips(ev, 50_000, stamps)
        assert len(got) == len(want) and "indices" not in got[0]
        for g, w in zip(got, want):
            assert g["time_start"] == w["time_start"] and g["time_end"] == w["time_end"] and g["duration_ms"] == w["duration_ms"]
            for k in "xytp":
                np.testing.assert_array_equal(g[k], w[k])


def test_dsec_io_mirrors_vs_reference_golden():
    """eventgpt_b200.io against the REAL dataset/io.py functions (tests/golden/dsec_io.npz, produced by
    `python -m oracle.make_golden io`, where a stub h5py hands the reference an in-memory recording)."""
    import os
    from oracle.make_golden import IO_RANGES, IO_WINDOWS, io_recording
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "dsec_io.npz"))
    rec = io_recording()

    def summary(d):
        return [len(d["t"]), int(d["t"][0]) if len(d["t"]) else -1, int(d["t"][-1]) if len(d["t"]) else -1,
                int(d["x"].astype(np.int64).sum()), int(d["y"].astype(np.int64).sum()), int(d["p"].astype(np.int64).sum()),
                int(d["t"].dtype == np.int64)]
    assert eio.get_num_events(rec) == int(z["num_events"])
    for (a, b), want in zip(IO_WINDOWS, z["windows"].tolist()):
        assert summary(eio.extract_from_h5_by_timewindow(rec, a, b)) == want
    for (a, b), want in zip(IO_RANGES, z["ranges"].tolist()):
        assert summary(eio.extract_from_h5_by_index(rec, a, b)) == want
