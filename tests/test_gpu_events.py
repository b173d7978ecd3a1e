"""GPU parity tests (through the C ABI) for splitting, rasterisation and CLIP preprocessing:
bit-exact against the goldens made by the real reference and against the CPU oracle."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, golden_event_cases, load_golden_events

pytestmark = pytest.mark.gpu
CASES = golden_event_cases()


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "-m gpu tests need a GPU"
    return torch.device("cuda:0")


def _dev_events(ev, dev):
    from eventgpt_b200 import events as E
    return tuple(E.to_device(ev[k], dev) for k in "xyp")


@pytest.mark.parametrize("name", CASES)
def test_get_event_images_list_bit_exact(name, dev):
    """common/common.py:17-37,64-74 through eventgpt_b200.common, vs the reference's frames."""
    from eventgpt_b200 import common
    ev, frames, _ = load_golden_events(name)
    got = common.get_event_images_list(ev, 5)
    assert len(got) == 5
    for g, f in zip(got, frames):
        assert g.dtype == np.uint8 and g.shape == f.shape
        assert np.array_equal(g, f)


@pytest.mark.parametrize("name", CASES)
def test_pixel_values_bit_exact(name, dev):
    """common/common.py:124-127: raster -> Pillow-exact resize/crop -> normalise -> bf16."""
    from eventgpt_b200 import events as E
    ev, frames, pix = load_golden_events(name)
    H, W = int(ev["y"].max()) + 1, int(ev["x"].max()) + 1
    x, y, p = _dev_events(ev, dev)
    off = torch.tensor([0, len(ev["t"])], dtype=torch.int64, device=dev)
    fr, hw, _, st = E.raster(x, y, p, off, 1, 5, H, W)
    px, patches, st = E.preprocess(fr, hw, H, want_patches=True, status=st)
    E.check_status(st, name)
    assert hw.cpu().numpy().tolist() == [list(f.shape[:2]) for f in frames]
    got = px.view(torch.int16).cpu().numpy().view(np.uint16)
    assert np.array_equal(got, pix)
    # the im2col layout holds the same values: row = patch, col = c*196 + dy*14 + dx, zero padded to 640
    pt = patches.view(torch.int16).cpu().numpy().view(np.uint16).reshape(5, 24, 24, 640)
    want = pix.reshape(5, 3, 24, 14, 24, 14).transpose(0, 2, 4, 1, 3, 5).reshape(5, 24, 24, 588)
    assert np.array_equal(pt[..., :588], want) and not pt[..., 588:].any()


def test_process_event_data_matches_reference(dev, tmp_path):
    """common/common.py:110-129 end to end on the reference's own sample."""
    from eventgpt_b200 import common
    from eventgpt_b200.model import EventProcessor
    ev, frames, pix = load_golden_events("sample1")
    path = tmp_path / "sample1.npy"
    np.save(path, np.array(ev, dtype=object), allow_pickle=True)
    size, lst = common.process_event_data(str(path), EventProcessor(dev), dev)
    assert size == [480, 640] and len(lst) == 5
    for t, want in zip(lst, pix):
        assert t.dtype == torch.bfloat16 and tuple(t.shape) == (3, 336, 336) and t.is_cuda
        assert np.array_equal(t.view(torch.int16).cpu().numpy().view(np.uint16), want)


def test_event_processor_call_surface(dev):
    from eventgpt_b200.model import EventProcessor
    _, frames, pix = load_golden_events("synthetic_000003")
    out = EventProcessor(dev)(frames[2], return_tensors="pt")["pixel_values"]
    assert tuple(out.shape) == (1, 3, 336, 336) and out.dtype == torch.float32
    assert np.array_equal(out[0].bfloat16().view(torch.int16).cpu().numpy().view(np.uint16), pix[2])


def test_generate_event_image_and_errors(dev):
    from eventgpt_b200 import common
    from oracle import events as oev
    rng = np.random.default_rng(5)
    x = rng.integers(0, 50, 999).astype(np.uint16); y = rng.integers(0, 40, 999).astype(np.uint16)
    p = rng.integers(0, 2, 999).astype(np.uint8)
    assert np.array_equal(common.generate_event_image(x, y, p), oev.generate_event_image_loop(x, y, p))
    with pytest.raises(ValueError):       # empty input: y.max() raises upstream
        common.generate_event_image(x[:0], y[:0], p[:0])
    with pytest.raises(ValueError):       # fewer events than frames: an empty slice
        common.get_event_images_list({"x": x[:3], "y": y[:3], "p": p[:3], "t": np.arange(3)}, 5)
    with pytest.raises(ValueError):       # coordinate beyond the stated sensor
        common.generate_event_image(x, y, p, sensor_size=(10, 10))


def test_batched_ragged_windows_and_modes(dev):
    """One launch for several ragged windows == per-window reference results; NEG_WINS mode
    (preprocess_event_images.py:58-90); fixed canvas; per-pixel counts."""
    from eventgpt_b200 import events as E, _lib
    from oracle import events as oev
    rng = np.random.default_rng(11)
    H, W = 120, 160
    sizes = [7001, 12, 30000, 999, 5]
    evs = [dict(x=rng.integers(0, W, n).astype(np.uint16), y=rng.integers(0, H, n).astype(np.uint16),
                p=rng.integers(0, 2, n).astype(np.uint8), t=np.arange(n)) for n in sizes]
    cat = {k: np.concatenate([e[k] for e in evs]) for k in "xyp"}
    x, y, p = _dev_events(cat, dev)
    off = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int64, device=dev)
    fr, hw, cnt, st = E.raster(x, y, p, off, len(sizes), 5, H, W, counts=True)
    E.check_status(st, "batched")
    got = E.frames_to_numpy(fr, hw)
    cnt = cnt.cpu().numpy()
    for b, e in enumerate(evs):
        want = oev.get_event_images_list(e, 5)
        for i in range(5):
            assert np.array_equal(got[b * 5 + i], want[i]), (b, i)
        starts, ends = oev.frame_bounds(sizes[b], 5)
        for i, (s, t_) in enumerate(zip(starts, ends)):
            lin = e["y"][s:t_].astype(np.int64) * W + e["x"][s:t_]
            for pol in (0, 1):
                ref = np.bincount(lin[(e["p"][s:t_] != 0) == bool(pol)], minlength=H * W).reshape(H, W)
                assert np.array_equal(cnt[b * 5 + i, pol], ref)
    # negative-wins mode with an explicit canvas
    fr, hw, _, st = E.raster(x, y, p, off, len(sizes), 5, H, W, mode=_lib.RASTER_NEG_WINS, fixed_canvas=True)
    got = E.frames_to_numpy(fr, hw)
    for b, e in enumerate(evs):
        starts, ends = oev.frame_bounds(sizes[b], 5)
        for i, (s, t_) in enumerate(zip(starts, ends)):
            want = oev.generate_event_image_negwins(e["x"][s:t_], e["y"][s:t_], e["p"][s:t_], H, W)
            assert np.array_equal(got[b * 5 + i], want), (b, i)


@pytest.mark.parametrize("dtype", [np.int64, np.int32, np.int16, np.uint8])
def test_generic_dtypes_and_misaligned_views(dtype, dev):
    from eventgpt_b200 import events as E
    from oracle import events as oev
    rng = np.random.default_rng(3)
    n, H, W = 5003, 100, 120
    ev = dict(x=rng.integers(0, W, n).astype(dtype), y=rng.integers(0, H, n).astype(dtype),
              p=rng.integers(0, 2, n).astype(dtype), t=np.arange(n))
    x, y, p = _dev_events(ev, dev)
    off = torch.tensor([0, n], dtype=torch.int64, device=dev)
    fr, hw, _, st = E.raster(x, y, p, off, 1, 5, H, W)
    E.check_status(st, "dtype")
    for g, w in zip(E.frames_to_numpy(fr, hw), oev.get_event_images_list(ev, 5)):
        assert np.array_equal(g, w)
    # u16/u8 arrays at odd element offsets take the unaligned path and must agree too
    ev2 = dict(x=ev["x"].astype(np.uint16), y=ev["y"].astype(np.uint16), p=ev["p"].astype(np.uint8), t=ev["t"])
    xb, yb, pb = (E.to_device(np.concatenate([[0], ev2[k]]).astype(ev2[k].dtype), dev)[1:] for k in "xyp")
    fr, hw, _, st = E.raster(xb, yb, pb, off, 1, 5, H, W)
    for g, w in zip(E.frames_to_numpy(fr, hw), oev.get_event_images_list(ev2, 5)):
        assert np.array_equal(g, w)


@pytest.mark.parametrize("n_events", [10 ** 4, 10 ** 6, 10 ** 7])
def test_raster_sweep_720p(n_events, dev):
    """BASELINE config 4: 1280x720, bit-exact vs the vectorised oracle (itself pinned to the
    reference loop on every golden)."""
    from eventgpt_b200 import events as E
    from oracle import events as oev
    rng = np.random.default_rng(4)
    H, W = 720, 1280
    ev = dict(x=rng.integers(0, W, n_events).astype(np.uint16), y=rng.integers(0, H, n_events).astype(np.uint16),
              p=rng.integers(0, 2, n_events).astype(np.uint8), t=np.arange(n_events))
    x, y, p = _dev_events(ev, dev)
    off = torch.tensor([0, n_events], dtype=torch.int64, device=dev)
    fr, hw, _, st = E.raster(x, y, p, off, 1, 5, H, W)
    E.check_status(st, "sweep")
    for g, w in zip(E.frames_to_numpy(fr, hw), oev.get_event_images_list(ev, 5)):
        assert np.array_equal(g, w)


def test_raster_full_size_properties(dev):
    """BASELINE config 2 size (32 windows x 1M events, 640x480): size-independent properties.
    (1) idempotence: rasterising the same batch twice gives identical bytes; (2) a window's
    frames do not depend on its neighbours (window 7 alone == window 7 in the batch);
    (3) every pixel is white, red or blue and the set of painted pixels equals the set of
    pixels that received an event."""
    from eventgpt_b200 import events as E
    rng = np.random.default_rng(2)
    B, N, H, W = 32, 1_000_000, 480, 640
    xs = rng.integers(0, W, B * N).astype(np.uint16); ys = rng.integers(0, H, B * N).astype(np.uint16)
    ps = rng.integers(0, 2, B * N).astype(np.uint8)
    x, y, p = (torch.from_numpy(a).to(dev) for a in (xs, ys, ps))
    off = torch.arange(0, B + 1, dtype=torch.int64, device=dev) * N
    fr1, hw1, _, st = E.raster(x, y, p, off, B, 5, H, W)
    E.check_status(st, "full")
    a = fr1.clone()
    fr2, hw2, _, _ = E.raster(x, y, p, off, B, 5, H, W)
    assert torch.equal(a, fr2) and torch.equal(hw1, hw2)
    one = torch.tensor([0, N], dtype=torch.int64, device=dev)
    fr7, _, _, _ = E.raster(x[7 * N:8 * N], y[7 * N:8 * N], p[7 * N:8 * N], one, 1, 5, H, W)
    assert torch.equal(fr7[:, : H * W * 3], a[35:40, : H * W * 3])
    f0 = a[0, : H * W * 3].view(H, W, 3).cpu().numpy()
    k = N // 5
    painted = np.zeros(H * W, bool)
    painted[ys[:k].astype(np.int64) * W + xs[:k]] = True
    white = (f0 == 255).all(-1).reshape(-1)
    red = ((f0[..., 0] == 255) & (f0[..., 1] == 0) & (f0[..., 2] == 0)).reshape(-1)
    blue = ((f0[..., 0] == 0) & (f0[..., 1] == 0) & (f0[..., 2] == 255)).reshape(-1)
    assert (white | red | blue).all() and np.array_equal(~white, painted)


def test_split_event_by_time_matches_reference(dev):
    """common/common.py:76-107: unsorted t, empty bins skipped, order preserved."""
    from eventgpt_b200 import common
    z = np.load(os.path.join(GOLDEN, "split_stream.npz"))
    parts = common.split_event_by_time({k: z[k] for k in "xytp"}, 50000)
    assert [len(q["t"]) for q in parts] == z["counts"].tolist()
    for k in "xytp":
        cat = np.concatenate([q[k] for q in parts])
        assert cat.dtype == z["out_" + k].dtype and np.array_equal(cat, z["out_" + k])


def test_split_dsec_window_and_large_sorted_stream(dev):
    from eventgpt_b200 import common
    from oracle import events as oev
    ev, _, _ = load_golden_events("dsec_000069")       # int64 absolute microseconds
    for dt in (50000, 10000):
        got = common.split_event_by_time(ev, dt)
        want = oev.split_event_by_time(ev, dt)
        assert len(got) == len(want)
        for g, w in zip(got, want):
            for k in "xytp":
                assert np.array_equal(g[k], w[k])
    # 50 ms / 7 us = 7143 bins > max_windows: the C ABI reports EGPT_FLAG_TOO_MANY_BINS (never a silent truncation) and the
    # mirror splits again with the span the stream needs -- the reference's result either way (np.unique has no limit)
    from eventgpt_b200 import events as E, _lib
    td = E.to_device(ev["t"], dev)
    xd, yd, pd = _dev_events(ev, dev)
    out = E.split_time(xd, yd, td, pd, 7, 4096)
    assert int(out[6].item()) & _lib.FLAG_TOO_MANY_BINS and int(out[5].item()) == 0
    got = common.split_event_by_time(ev, 7, max_windows=4096)
    want = oev.split_event_by_time(ev, 7)
    assert len(got) == len(want) > 4096
    for g, w in zip(got[:50] + got[-50:], want[:50] + want[-50:]):
        for k in "xytp":
            assert np.array_equal(g[k], w[k])
    # two bursts 300 s apart (a recording with a gap): 2 windows, 6001 bins of span
    gap = {k: np.concatenate([ev[k][:1000], ev[k][-1000:]]) for k in "xytp"}
    gap["t"] = np.concatenate([ev["t"][:1000], ev["t"][-1000:] + 300_000_000])
    got, want = common.split_event_by_time(gap, 50000), oev.split_event_by_time(gap, 50000)
    assert len(got) == len(want) and all(np.array_equal(g[k], w[k]) for g, w in zip(got, want) for k in "xytp")
    rng = np.random.default_rng(9)
    n = 3_000_000
    st = dict(t=np.sort(rng.integers(0, 1_000_000, n)).astype(np.int64), x=rng.integers(0, 640, n).astype(np.uint16),
              y=rng.integers(0, 480, n).astype(np.uint16), p=rng.integers(0, 2, n).astype(np.uint8))
    got = common.split_event_by_time(st, 50000)
    want = oev.split_event_by_time(st, 50000)
    assert len(got) == len(want) == 20
    for g, w in zip(got, want):
        for k in "xytp":
            assert np.array_equal(g[k], w[k])


def test_overlap_clips_rasterise_like_reference(dev):
    """50 %-overlap clips of a recording (build_my_egpt_dsec_seq.py:137-202) as one ragged raster batch: every clip's five
    frames equal the reference rasterisation of that clip's events (events in an overlap are rasterised in both clips)."""
    from eventgpt_b200 import events as E, io as eio
    from oracle import events as oev
    rng = np.random.default_rng(21)
    n, H, W = 120_000, 120, 160
    t = np.sort(rng.integers(2_000_000, 2_400_000, n)).astype(np.int64)
    t = t[(t < 2_150_000) | (t >= 2_230_000)]            # an 80 ms hole: one clip vanishes, its neighbours are ragged
    n = len(t)
    ev = dict(x=rng.integers(0, W, n).astype(np.uint16), y=rng.integers(0, H, n).astype(np.uint16),
              p=rng.integers(0, 2, n).astype(np.uint8), t=t)
    stamps = [2_000_000 + 50_000 * i for i in range(9)]
    clips = eio.split_event_by_time_clips(ev, 50_000, stamps)
    want_clips = oev.split_event_by_time_clips(ev, 50_000, stamps)
    assert len(clips) == len(want_clips) and 10 <= len(clips) < 15
    x, y, p, off = eio.clips_to_batch(ev, clips)
    dx, dy, dp = _dev_events(dict(x=x, y=y, p=p), dev)
    fr, hw, _, st = E.raster(dx, dy, dp, torch.tensor(off, dtype=torch.int64, device=dev), len(clips), 5, H, W)
    E.check_status(st, "clips")
    got = E.frames_to_numpy(fr, hw)
    for b, c in enumerate(want_clips):
        sub = {k: ev[k][c["indices"]] for k in "xypt"}
        want = oev.get_event_images_list(sub, 5)
        for i in range(5):
            assert np.array_equal(got[b * 5 + i], want[i]), (b, i)


def test_device_clip_bounds_and_ranges_vs_reference_goldens(dev):
    """Upstream ingestion on the device.  (a) egpt_clip_bounds on the recording of tests/golden/clips.npz gives
    the clips the reference's own 50 %-overlap splitter produced (build_my_egpt_dsec_seq.py:137-202: first index and event count
    of every non-empty clip); egpt_raster_ranges rasterises them from the ONE copy of the events, equal to the oracle's frames
    of each clip.  (b) egpt_ms_window_bounds reproduces the slices the reference's extract_from_h5_by_timewindow cut
    (tests/golden/dsec_io.npz, dataset/io.py:68-87)."""
    from eventgpt_b200 import events as E
    from oracle import events as oev
    from oracle.make_golden import clips_inputs, io_recording, IO_WINDOWS
    z = np.load(os.path.join(GOLDEN, "clips.npz"))
    ev, stamps = clips_inputs()
    first, interval, step = min(stamps), 50_000, 25_000
    n_clips = (max(stamps) - interval - first) // step + 1
    td = E.to_device(ev["t"], dev)
    bounds = E.clip_bounds(td, first, interval, step, n_clips)
    b = bounds.cpu().numpy()
    keep = b[:, 1] > b[:, 0]
    assert int(keep.sum()) == int(z["a_n"])
    assert np.array_equal(b[keep, 0], z["a_first"]) and np.array_equal(b[keep, 1] - b[keep, 0], z["a_count"])
    assert np.array_equal(first + step * np.nonzero(keep)[0], z["a_time_start"])
    x, y, p = _dev_events(ev, dev)
    kept = bounds[torch.from_numpy(keep).to(dev)].contiguous()
    fr, hw, _, st = E.raster(x, y, p, kept, kept.shape[0], 5, 480, 640, ranges=True)
    E.check_status(st, "clips")
    got = E.frames_to_numpy(fr, hw)
    for c, (s_, e_) in enumerate(b[keep]):
        want = oev.get_event_images_list({k: ev[k][s_:e_] for k in "xypt"}, 5)
        for i in range(5):
            assert np.array_equal(got[c * 5 + i], want[i]), (c, i)
    # an empty clip in the batch is flagged like any empty slice (the reference drops such clips before rasterising)
    fr, hw, _, st = E.raster(x, y, p, bounds, n_clips, 5, 480, 640, ranges=True)
    with pytest.raises(ValueError):
        E.check_status(st, "clips with a hole")
    # (b) DSEC per-millisecond index
    zi = np.load(os.path.join(GOLDEN, "dsec_io.npz"))
    rec = io_recording()
    off = int(rec["t_offset"])
    ms = E.to_device(rec["ms_to_idx"], dev)
    for (t0, t1), want in zip(IO_WINDOWS, zi["windows"]):
        a_ms, b_ms = (t0 - off) // 1000, (t1 - off) // 1000
        bd, st = E.ms_window_bounds(ms, a_ms, b_ms - a_ms, 1, 1)
        assert int(st.item()) == 0
        s_, e_ = (int(v) for v in bd.cpu().numpy()[0])
        sl = {k: rec["events"][k][s_:e_] for k in "xypt"}
        assert e_ - s_ == want[0]
        if e_ > s_:
            assert int(sl["t"][0]) + off == want[1] and int(sl["t"][-1]) + off == want[2]
            assert [int(sl[k].astype(np.int64).sum()) for k in "xyp"] == list(want[3:6])
    bd, st = E.ms_window_bounds(ms, 0, 50, 50, 9)             # the ninth 50 ms window lies beyond the 400 ms index
    assert int(st.item()) & 2


@pytest.mark.parametrize("h,w", [(1100, 2000), (340, 1900), (1500, 400), (720, 1280), (483, 641), (260, 346), (337, 339), (901, 350)])
def test_preprocess_arbitrary_images_wide_and_deep_filters(h, w, dev):
    """egpt_preprocess on arbitrary uint8 images (the PNG path) at sizes that leave the common fast paths:
    (1100, 2000): rows too wide for the shared-memory staging AND more than 12 filter taps on both axes (scale 3.3);
    (340, 1900): almost no vertical scaling but a 5.6x horizontal reduction (23 taps); (1500, 400): tall, horizontal up-scale.
    And sizes that take the byte-parallel fast path (at most 12 taps per axis): (720, 1280) needs three packed words per tap
    window (9-10 taps); (483, 641) and (337, 339) have widths / heights that are no multiples of four (partial pixel quads, a last
    row quad with rows past the frame); (260, 346) is an up-scale; (901, 350) is tall with 11 vertical taps.
    Bit-exact against the CPU restatement of Pillow's fixed-point bicubic + the CLIP normalisation."""
    from eventgpt_b200 import events as E
    from oracle import preprocess as opp
    rng = np.random.default_rng(h * 10007 + w)
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    img[::7, ::5] = 255
    img[3::11, 2::3] = 0
    stride = (h * w * 3 + 255) // 256 * 256
    frames = torch.zeros((1, stride), dtype=torch.uint8, device=dev)
    frames[0, :h * w * 3] = torch.from_numpy(img.reshape(-1)).to(dev)
    hw = torch.tensor([[h, w]], dtype=torch.int32, device=dev)
    px, _, st = E.preprocess(frames, hw, h)
    E.check_status(st, "preprocess")
    got = px.view(torch.int16).cpu().numpy().view(np.uint16)[0]
    want = opp.pixel_values_bf16_bits(img)
    assert np.array_equal(got, want)


def test_preprocess_mixed_canvas_sizes_in_one_call(dev):
    """One egpt_preprocess call over frames of different data-dependent canvases inside one (720, 1280) capacity: 720x1280 (three
    packed words per tap window), 480x640 (two), 333x517 (odd sizes, row offsets at every alignment) and an empty frame.  Each
    frame must equal the oracle run on that frame alone."""
    from eventgpt_b200 import events as E
    from oracle import preprocess as opp
    rng = np.random.default_rng(77)
    H, W = 720, 1280
    sizes = [(720, 1280), (480, 640), (333, 517), (0, 0)]
    stride = H * W * 3
    frames = torch.zeros((len(sizes), stride), dtype=torch.uint8, device=dev)
    imgs = []
    for i, (h, w) in enumerate(sizes):
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        imgs.append(img)
        if h: frames[i, :h * w * 3] = torch.from_numpy(img.reshape(-1)).to(dev)
    hw = torch.tensor(sizes, dtype=torch.int32, device=dev)
    px, patches, st = E.preprocess(frames, hw, H, want_patches=True)
    E.check_status(st, "preprocess")
    got = px.view(torch.int16).cpu().numpy().view(np.uint16)
    pat = patches.view(torch.int16).cpu().numpy().view(np.uint16).reshape(len(sizes), 576, -1)
    for i, (h, w) in enumerate(sizes[:3]):
        want = opp.pixel_values_bf16_bits(imgs[i])
        assert np.array_equal(got[i], want), sizes[i]
        # im2col rows of the same pixels: patch (py, px), column c*196 + dy*14 + dx
        cols = want.reshape(3, 24, 14, 24, 14).transpose(1, 3, 0, 2, 4).reshape(576, 588)
        assert np.array_equal(pat[i][:, :588], cols) and not pat[i][:, 588:].any(), sizes[i]


@pytest.mark.parametrize("name", ["dsec_000099", "adv_signed_p"])
def test_negative_wins_mode_vs_reference_frames(name, dev):
    """EGPT_RASTER_NEG_WINS through the C ABI against frames rendered by the reference's own generate_event_image_vectorized
    (preprocess_event_images.py:58-90; tests/golden/negwins.npz): data-dependent canvas and explicit 480x640 canvas."""
    from eventgpt_b200 import events as E, _lib
    z = np.load(os.path.join(GOLDEN, "negwins.npz"))
    ev, _, _ = load_golden_events(name)
    x, y, p = _dev_events(ev, dev)
    off = torch.tensor([0, len(ev["t"])], dtype=torch.int64, device=dev)
    for tag, fixed in (("auto", False), ("vga", True)):
        fr, hw, _, st = E.raster(x, y, p, off, 1, 5, 480, 640, mode=_lib.RASTER_NEG_WINS, fixed_canvas=fixed)
        E.check_status(st, name)
        got = E.frames_to_numpy(fr, hw)
        for i in range(5):
            assert np.array_equal(got[i], z[f"{name}_{tag}_{i}"]), (tag, i)
