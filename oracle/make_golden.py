"""Oracle (test infrastructure): generates tests/golden/*.npz by running the REAL
reference code (/root/reference, build container only) on the reference's own
fixtures and on adversarial inputs.  Re-run with

    python -m oracle.make_golden

The reference has no tests or golden vectors of its own,
so these files ARE the pin: ``tests/test_oracle_golden.py`` checks the CPU
restatement in ``oracle/`` against them, and the ``-m gpu`` tests check the CUDA
path against them and against the oracle.
"""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness  # noqa: E402
from eventgpt_b200.weights import random_state_dict  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
REF = ref_harness.REF
N_FRAMES = 5


def load_npy(path):
    return np.array(np.load(path, allow_pickle=True)).item()


def adversarial_cases():
    rng = np.random.default_rng(1234)
    cases = {}
    # every event on one pixel, alternating polarity: order sensitivity
    n = 1003
    cases["adv_one_pixel"] = dict(x=np.full(n, 7, np.uint16), y=np.full(n, 3, np.uint16),
                                  t=np.arange(n, dtype=np.int64), p=(np.arange(n) % 2).astype(np.uint8))
    # remainder handling (common/common.py:27): 5, 6, 7, 9 events
    for n in (5, 6, 7, 9):
        cases[f"adv_n{n}"] = dict(x=rng.integers(0, 20, n).astype(np.uint16), y=rng.integers(0, 12, n).astype(np.uint16),
                                  t=np.sort(rng.integers(0, 50000, n)).astype(np.int64), p=rng.integers(0, 2, n).astype(np.uint8))
    # p in {-1,+1}: -1 is painted red (common/common.py:69-72)
    n = 4001
    cases["adv_signed_p"] = dict(x=rng.integers(0, 64, n).astype(np.int64), y=rng.integers(0, 48, n).astype(np.int64),
                                 t=np.sort(rng.integers(0, 50000, n)).astype(np.int64),
                                 p=(rng.integers(0, 2, n) * 2 - 1).astype(np.int64))
    # p in {0,1,2,255}: any non-zero is red
    cases["adv_p_multi"] = dict(x=rng.integers(0, 90, n).astype(np.uint16), y=rng.integers(0, 70, n).astype(np.uint16),
                                t=np.sort(rng.integers(0, 50000, n)).astype(np.uint16),
                                p=rng.choice(np.array([0, 1, 2, 255], np.uint8), n))
    # hot pixels (Zipf) on a VGA sensor: heavy collisions
    n = 60000
    z = np.minimum(rng.zipf(1.3, n) - 1, 640 * 480 - 1)
    perm = rng.permutation(640 * 480)
    lin = perm[z]
    cases["adv_zipf_vga"] = dict(x=(lin % 640).astype(np.uint16), y=(lin // 640).astype(np.uint16),
                                 t=np.sort(rng.integers(0, 50000, n)).astype(np.int64), p=rng.integers(0, 2, n).astype(np.uint8))
    # tall sensor (H > W): resize takes the other branch
    n = 30000
    cases["adv_tall"] = dict(x=rng.integers(0, 260, n).astype(np.uint16), y=rng.integers(0, 346, n).astype(np.uint16),
                             t=np.sort(rng.integers(0, 50000, n)).astype(np.int64), p=rng.integers(0, 2, n).astype(np.uint8))
    # Gen4-class 1280x720
    n = 150000
    cases["adv_720p"] = dict(x=rng.integers(0, 1280, n).astype(np.uint16), y=rng.integers(0, 720, n).astype(np.uint16),
                             t=np.sort(rng.integers(0, 50000, n)).astype(np.int64), p=rng.integers(0, 2, n).astype(np.uint8))
    # square, exactly 336 (no resize pass at all) and an upscale case
    n = 20000
    cases["adv_336sq"] = dict(x=rng.integers(0, 336, n).astype(np.uint16), y=rng.integers(0, 336, n).astype(np.uint16),
                              t=np.sort(rng.integers(0, 50000, n)).astype(np.int64), p=rng.integers(0, 2, n).astype(np.uint8))
    n = 3000
    cases["adv_small_up"] = dict(x=rng.integers(0, 120, n).astype(np.uint16), y=rng.integers(0, 100, n).astype(np.uint16),
                                 t=np.sort(rng.integers(0, 50000, n)).astype(np.int64), p=rng.integers(0, 2, n).astype(np.uint8))
    return cases


def reference_cases():
    cases = {"sample1": load_npy(f"{REF}/samples/sample1.npy")}
    d = f"{REF}/feasible/egpt_dataset/event_npy/interlaken_00_d"
    for f in ("000069", "000099"):
        cases[f"dsec_{f}"] = load_npy(f"{d}/{f}.npy")
    d = f"{REF}/feasible/my_egpt_dsec_dataset/example_event_npy"
    for i in range(10):
        cases[f"synthetic_{i:06d}"] = load_npy(f"{d}/synthetic_{i:06d}.npy")
    return cases


def main():
    os.makedirs(OUT, exist_ok=True)
    C, M = ref_harness.import_reference()
    proc = ref_harness.pil_processor()
    cases = reference_cases()
    cases.update(adversarial_cases())
    manifest = []
    for name, ev in cases.items():
        ev = {k: np.asarray(ev[k]) for k in ("x", "y", "t", "p")}
        frames = C.get_event_images_list(ev, N_FRAMES)                      # common/common.py:17-37
        hw = np.array([f.shape[:2] for f in frames], np.int32)
        pix = []
        for f in frames:                                                    # common/common.py:124-127
            pv = proc(f, return_tensors="pt")["pixel_values"][0]
            pix.append(pv.to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16))
        pix = np.stack(pix)
        packed = np.concatenate([f.reshape(-1) for f in frames])
        np.savez_compressed(os.path.join(OUT, f"events_{name}.npz"), x=ev["x"], y=ev["y"], t=ev["t"], p=ev["p"],
                            frame_hw=hw, frames=packed, pix=pix)
        manifest.append((name, len(ev["t"]), hw.tolist(), hashlib.sha256(pix.tobytes()).hexdigest()[:16]))
        print(name, len(ev["t"]), hw.tolist())

    # uint8 level -> bf16 LUT straight from the processor
    ramp = np.zeros((336, 336, 3), np.uint8)
    ramp[:] = (np.arange(336) % 256).astype(np.uint8)[None, :, None]
    pv = proc(ramp, return_tensors="pt")["pixel_values"][0]                 # [3,336,336]
    lut_f32 = pv[:, 0, :256].T.contiguous().numpy()                        # [256,3]
    lut_bf16 = pv[:, 0, :256].T.contiguous().to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    np.savez_compressed(os.path.join(OUT, "normalise_lut.npz"), f32=lut_f32, bf16=lut_bf16)

    # process_event_data end to end on the reference's own sample (common/common.py:110-129)
    size, evl = C.process_event_data(f"{REF}/samples/sample1.npy", proc, "cpu")
    ped = torch.stack(evl).view(torch.int16).numpy().view(np.uint16)
    ref_pix = np.load(os.path.join(OUT, "events_sample1.npz"))["pix"]
    assert np.array_equal(ped, ref_pix) and size == [480, 640]

    # split_event_by_time (common/common.py:76-107): unsorted t, gaps (empty bins), 3 dtypes
    rng = np.random.default_rng(77)
    n = 50000
    t = rng.integers(0, 400000, n).astype(np.int64)
    t[(t >= 100000) & (t < 200000)] += 200000            # leave bins 100000, 150000 empty
    half = n // 2
    t[:half] = np.sort(t[:half])                          # half sorted, half not
    stream = dict(x=rng.integers(0, 640, n).astype(np.uint16), y=rng.integers(0, 480, n).astype(np.uint16),
                  t=t, p=rng.integers(0, 2, n).astype(np.uint8))
    parts = C.split_event_by_time(stream, 50000)
    np.savez_compressed(os.path.join(OUT, "split_stream.npz"), **stream,
                        counts=np.array([len(q["t"]) for q in parts], np.int64),
                        out_x=np.concatenate([q["x"] for q in parts]), out_y=np.concatenate([q["y"] for q in parts]),
                        out_t=np.concatenate([q["t"] for q in parts]), out_p=np.concatenate([q["p"] for q in parts]))
    print("split bins", [len(q["t"]) for q in parts])

    # tower -> projector -> adaptor -> pooling through the reference's list branch
    # (model/EventChatModel.py:326-343), fp32 CPU, seeded weights.
    evl32 = [e.float() for e in evl]
    for layers in (2, 24):
        sd = random_state_dict(seed=0, num_layers=layers)
        C2, M2, model = ref_harness.build_reference_model(sd, num_layers=layers)
        with torch.no_grad():
            # visval_encode = tower.last_hidden_state -> projector (:194-200)
            f0 = model.visval_encode(evl32[0][None]).detach()
            lhs0 = model.get_model().visual_tower.visual_tower(evl32[0][None]).last_hidden_state
            feats = []
            for ev in evl32:
                f = model.visval_encode(ev[None]).detach()
                f = model.get_model().feature_adaptor(f)[0]
                feats.append(f)
            tokens = M2.get_spatio_temporal_features(feats)                 # [582,4096]
            # and once through prepare_inputs_labels_for_multimodal itself
            ids = torch.tensor([[1, -200, 5, 6]])
            out = model.prepare_inputs_labels_for_multimodal(ids, None, None, None, None, event_tensors=evl32)
            embeds = out[4]
            assert embeds.shape == (1, 585, 4096)
            assert torch.equal(embeds[0, 1:583], tokens)
        wsum = float(sum(v.double().abs().sum() for v in sd.values()))
        np.savez_compressed(os.path.join(OUT, f"tokens_sample1_L{layers}.npz"),
                            tokens_sub=tokens[:, ::16].numpy(), tokens_rows=tokens[[0, 4, 5, 300, 581]].numpy(),
                            lhs0_sub=lhs0[0, :, ::8].numpy(), proj0_sub=f0[0, :, ::16].numpy(),
                            absmax=np.float32(tokens.abs().max()), std=np.float32(tokens.std()),
                            weight_abs_sum=np.float64(wsum))
        print("tokens L", layers, tokens.shape, float(tokens.abs().max()), float(tokens.std()), wsum)
        # yardstick: the reference's own bf16 execution vs its fp32
        if layers == 24:
            mb = model.to(torch.bfloat16)
            with torch.no_grad():
                fb = []
                for ev in evl:
                    f = mb.visval_encode(ev[None]).detach()
                    fb.append(mb.get_model().feature_adaptor(f)[0])
                tb = M2.get_spatio_temporal_features(fb).float()
            d = (tb - tokens).abs().max().item()
            cos = torch.nn.functional.cosine_similarity(tb.flatten(), tokens.flatten(), dim=0).item()
            print("torch-bf16 vs fp32: max-abs", d, "cos", cos)
            np.savez(os.path.join(OUT, "yardstick_bf16.npz"), max_abs=np.float32(d), cosine=np.float32(cos))
    for m in manifest:
        print(m)


def synth_window(global_index, n_events=1_000_000, h=480, w=640):
    """The benchmark's window generator (bench.py synth_window), restated so that the golden does not
    depend on bench.py: dense uniform events with one event at the sensor corner in every slice."""
    rng = np.random.default_rng([1, global_index])
    x = rng.integers(0, w, n_events).astype(np.uint16)
    y = rng.integers(0, h, n_events).astype(np.uint16)
    p = rng.integers(0, 2, n_events).astype(np.uint8)
    k = n_events // N_FRAMES
    for i in range(N_FRAMES):
        x[i * k] = w - 1
        y[i * k] = h - 1
    return x, y, p


def make_full_token_goldens(only=None):
    """24-layer tokens of the REAL reference model (fp32, CPU), FULL [582, 4096] tensors stored as float16 (quantisation
    <= 2e-3 at |x| <= 4, stated in the tests), for: the reference's sample1, one DSEC window, one window of the benchmark
    workload (dense 1M events), and sample1 again under weights with outlier channels and peaked attention heads.  Next to
    each, the reference's own bf16 execution against its fp32 one (the yardstick of SURVEY F1)."""
    from eventgpt_b200.weights import outlier_state_dict
    C, M = ref_harness.import_reference()
    proc = ref_harness.pil_processor()
    dsec = load_npy(f"{REF}/feasible/egpt_dataset/event_npy/interlaken_00_d/000069.npy")
    x, y, p = synth_window(0)
    inputs = {
        "sample1": load_npy(f"{REF}/samples/sample1.npy"),
        "dsec_000069": dsec,
        "dense1m": dict(x=x, y=y, p=p, t=np.zeros(len(x), np.int8)),
    }
    pix = {}
    for name, ev in inputs.items():
        ev = {k: np.asarray(ev[k]) for k in ("x", "y", "t", "p")}
        if name == "dense1m":
            from oracle import events as oev          # the reference loop takes ~2 s per 1M events; same frames (pinned)
            frames = oev.get_event_images_list(ev, N_FRAMES)
            chk = C.get_event_images_list({k: v[:50_000] for k, v in ev.items()}, N_FRAMES)
            assert all(np.array_equal(a, b) for a, b in zip(chk, oev.get_event_images_list({k: v[:50_000] for k, v in ev.items()}, N_FRAMES)))
        else:
            frames = C.get_event_images_list(ev, N_FRAMES)
        pix[name] = [proc(f, return_tensors="pt")["pixel_values"][0].to(torch.bfloat16) for f in frames]
    jobs = [("sample1", "random", random_state_dict(0, 24)), ("dsec_000069", "random", None), ("dense1m", "random", None),
            ("sample1", "outlier", outlier_state_dict(1, 24))]
    if only:
        jobs = [j for j in jobs if j[1] == only]
    model = None
    for name, wname, sd in jobs:
        if sd is not None:
            _, M2, model = ref_harness.build_reference_model(sd, num_layers=24)
            cur_sd = sd
        evl = pix[name]
        with torch.no_grad():
            m32 = model.float()
            feats = [m32.get_model().feature_adaptor(m32.visval_encode(e.float()[None]))[0] for e in evl]
            tokens = M2.get_spatio_temporal_features(feats)
            lhs = m32.get_model().visual_tower.visual_tower(evl[0].float()[None]).last_hidden_state[0]
            mb = model.to(torch.bfloat16)
            fb = [mb.get_model().feature_adaptor(mb.visval_encode(e[None]))[0] for e in evl]
            tb = M2.get_spatio_temporal_features(fb).float()
            model.float()
        d = (tb - tokens).abs().max().item()
        cos = torch.nn.functional.cosine_similarity(tb.flatten(), tokens.flatten(), dim=0).item()
        wsum = float(sum(v.double().abs().sum() for v in cur_sd.values()))
        px = torch.stack(evl).view(torch.int16).numpy().view(np.uint16)
        out = os.path.join(OUT, f"tokens_full_{name}_{wname}.npz")
        np.savez_compressed(out, tokens_f16=tokens.numpy().astype(np.float16), pix_sha=hashlib.sha256(px.tobytes()).hexdigest(),
                            lhs_absmax_per_channel=lhs.abs().amax(0).numpy(), absmax=np.float32(tokens.abs().max()),
                            std=np.float32(tokens.std()), weight_abs_sum=np.float64(wsum),
                            bf16_max_abs=np.float32(d), bf16_cosine=np.float32(cos),
                            tokens_rows=tokens[[0, 4, 5, 300, 581]].numpy())
        print(name, wname, "tokens absmax", float(tokens.abs().max()), "std", float(tokens.std()),
              "| last_hidden_state channel absmax: median", float(lhs.abs().amax(0).median()), "max", float(lhs.abs().amax(0).max()),
              "| torch-bf16 vs fp32 max-abs", d, "cos", cos, "|", os.path.getsize(out) >> 10, "KiB")


def make_splice_golden():
    """Multimodal splice (model/EventChatModel.py:355-465) through the REAL reference: a small-vocab
    EventChatModel with a seeded bf16-representable embedding table and an identity feature_adaptor
    (the cached-features branch re-applies the adaptor, :347-353; identity makes that a no-op so the
    golden pins the splice alone)."""
    import tempfile
    from transformers import CLIPVisionConfig, CLIPVisionModel, CLIPImageProcessor
    from oracle import splice as osp
    C, M = ref_harness.import_reference()
    work = tempfile.mkdtemp()
    ck = os.path.join(work, "checkpoints", "c")
    os.makedirs(ck)
    CLIPVisionModel(CLIPVisionConfig(hidden_size=64, intermediate_size=64, num_hidden_layers=1, num_attention_heads=1,
                                     image_size=28, patch_size=14)).save_pretrained(ck)
    CLIPImageProcessor().save_pretrained(ck)
    cwd = os.getcwd()
    os.chdir(work)
    cfg = M.EventChatConfig(hidden_size=4096, num_hidden_layers=1, intermediate_size=64, num_attention_heads=32,
                            vocab_size=1024, mm_visual_tower="c", event_feature_adaptor=True)
    model = M.EventChatModel(cfg).float().eval()
    os.chdir(cwd)
    table, feats = splice_inputs()
    with torch.no_grad():
        model.get_model().embed_tokens.weight.copy_(table)
        model.get_model().feature_adaptor.weight.copy_(torch.eye(4096))
        model.get_model().feature_adaptor.bias.zero_()
    g = torch.Generator().manual_seed(321)
    L = 1700
    ids = torch.randint(1, 1000, (3, L), generator=g)
    mask = torch.ones(3, L, dtype=torch.long)
    ids[0, 5] = -200; mask[0, 40:] = 0                       # one event token, short sample
    ids[1, 3] = -200; ids[1, 1600] = -200                    # two event tokens, 1698 + 2*582 > 2048: truncated
    mask[2, 100:] = 0; mask[2, 7] = 0                        # no event token, hole in the mask
    labels = torch.where(torch.rand(3, L, generator=g) < 0.5, ids, torch.full_like(ids, -100))
    with torch.no_grad():
        out = model.prepare_inputs_labels_for_multimodal(ids, None, mask, None, labels, event_features=feats.clone())
    _, pos, am, _, emb, lab = out
    o_pos, o_am, o_emb, o_lab = osp.splice(table, ids, None, mask, labels, feats)
    assert o_pos is None and pos is None and torch.equal(o_emb, emb) and torch.equal(o_am, am) and torch.equal(o_lab, lab)
    pid = torch.arange(L)[None].expand(3, L).contiguous()
    with torch.no_grad():
        out2 = model.prepare_inputs_labels_for_multimodal(ids, pid, None, None, None, event_features=feats.clone())
    o2 = osp.splice(table, ids, pid, None, None, feats)
    assert out2[2] is None and o2[1] is None and out2[5] is None and o2[3] is None
    assert torch.equal(out2[1], o2[0]) and torch.equal(out2[4], o2[2])
    np.savez_compressed(os.path.join(OUT, "splice.npz"), ids=ids.numpy(), mask=mask.numpy(), labels=labels.numpy(),
                        emb_sub=emb[:, :, ::64].numpy(), emb_rowsum=emb.double().sum(-1).numpy(), am=am.numpy(),
                        lab=lab.numpy(), pos2=out2[1].numpy(), emb2_rowsum=out2[4].double().sum(-1).numpy())
    print("splice golden written", tuple(emb.shape))


def splice_inputs():
    """Seeded embedding table [1024,4096] and event features [4,582,4096], bf16 representable."""
    g = torch.Generator().manual_seed(123)
    table = (torch.randn(1024, 4096, generator=g) * 0.02).bfloat16().float()
    feats = torch.randn(4, 582, 4096, generator=g).bfloat16().float()
    return table, feats


def clips_inputs():
    """Seeded recording for the 50 %-overlap clip splitter: sorted t with an 80 ms hole, shuffled image timestamps."""
    rng = np.random.default_rng(31)
    t = np.sort(rng.integers(1_000_000, 1_600_000, 20_000)).astype(np.int64)
    t = t[(t < 1_200_000) | (t >= 1_280_000)]
    n = len(t)
    ev = {"x": rng.integers(0, 640, n).astype(np.uint16), "y": rng.integers(0, 480, n).astype(np.uint16),
          "p": rng.integers(0, 2, n).astype(np.uint8), "t": t}
    stamps = [1_000_000 + 50_000 * i for i in range(12)]
    stamps[3], stamps[7] = stamps[7], stamps[3]
    return ev, stamps


def make_clips_golden():
    """split_event_by_time of the DSEC clip builder (feasible/my_egpt_dsec_dataset/build_my_egpt_dsec_seq.py:137-225) through
    the REAL reference function.  The module imports h5py (absent here, unused by this function) at its top, so a stub module
    stands in for it during the import -- the same trick as the peft stub of ref_harness."""
    import contextlib, importlib.util, io, types
    had = "h5py" in sys.modules
    if not had:
        sys.modules["h5py"] = types.ModuleType("h5py")
    try:
        spec = importlib.util.spec_from_file_location(
            "ref_build_seq", os.path.join(REF, "feasible", "my_egpt_dsec_dataset", "build_my_egpt_dsec_seq.py"))
        mod = importlib.util.module_from_spec(spec)
        with contextlib.redirect_stdout(io.StringIO()):
            spec.loader.exec_module(mod)
    finally:
        if not had:
            del sys.modules["h5py"]
    ev, stamps = clips_inputs()
    out = {}
    for name, interval, st in (("a", 50_000, stamps), ("b", 33_333, stamps), ("fallback", 50_000, None)):
        with contextlib.redirect_stdout(io.StringIO()):
            clips = mod.split_event_by_time(ev, time_interval=interval, image_timestamps=st)
        out[f"{name}_n"] = np.int64(len(clips))
        out[f"{name}_time_start"] = np.array([c["time_start"] for c in clips], dtype=np.int64)
        out[f"{name}_time_end"] = np.array([c["time_end"] for c in clips], dtype=np.int64)
        if "indices" in clips[0]:
            out[f"{name}_first"] = np.array([c["indices"][0] for c in clips], dtype=np.int64)
            out[f"{name}_count"] = np.array([c["num_events"] for c in clips], dtype=np.int64)
            out[f"{name}_index_sum"] = np.array([int(c["indices"].sum()) for c in clips], dtype=np.int64)
            out[f"{name}_image_idx"] = np.array([c["image_idx"] for c in clips], dtype=np.int64)
            out[f"{name}_n_images"] = np.array([len(c["image_indices_in_clip"]) for c in clips], dtype=np.int64)
        else:
            out[f"{name}_count"] = np.array([len(c["t"]) for c in clips], dtype=np.int64)
            out[f"{name}_xsum"] = np.array([int(c["x"].astype(np.int64).sum()) for c in clips], dtype=np.int64)
    np.savez_compressed(os.path.join(OUT, "clips.npz"), **out)
    print("clips.npz:", {k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items() if k.endswith("_n")})


def io_recording(seed=0, n=200_000, dur_ms=400):
    """Seeded in-memory recording in the DSEC layout (events/{x,y,p,t}, ms_to_idx, t_offset)."""
    rng = np.random.default_rng(seed)
    t = np.sort(rng.integers(0, dur_ms * 1000, n)).astype(np.uint32)
    ms_to_idx = np.searchsorted(t, np.arange(dur_ms + 1) * 1000, side="left").astype(np.uint64)
    return {"events": {"x": rng.integers(0, 640, n).astype(np.uint16), "y": rng.integers(0, 480, n).astype(np.uint16),
                       "p": rng.integers(0, 2, n).astype(np.uint8), "t": t},
            "ms_to_idx": ms_to_idx, "t_offset": np.array(51_000_000_000, dtype=np.int64)}


IO_WINDOWS = [(51_000_050_000, 51_000_100_000), (51_000_000_000, 51_000_050_000), (51_000_123_456, 51_000_301_999),
              (51_000_399_000, 51_000_399_500)]
IO_RANGES = [(0, 10), (1234, 98_765), (199_990, 200_000)]


def make_io_golden():
    """dataset/io.py:38-87 through the REAL reference functions.  h5py is absent from the image; the reference only uses
    `h5py.File(path, 'r')` as a context manager and indexes the handle, so a stub module whose File returns the in-memory
    recording stands in for it."""
    import importlib.util, types

    class FakeFile:
        def __init__(self, rec): self.rec = rec
        def __enter__(self): return self
        def __exit__(self, *a): return False
        def __getitem__(self, key):
            node = self.rec
            for part in key.split("/"):
                node = node[part]
            return node

    rec = io_recording()
    stub = types.ModuleType("h5py")
    stub.File = lambda path, mode="r": FakeFile(rec)
    had = sys.modules.get("h5py")
    sys.modules["h5py"] = stub
    try:
        spec = importlib.util.spec_from_file_location("ref_dataset_io", os.path.join(REF, "dataset", "io.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if had is None:
            del sys.modules["h5py"]
        else:
            sys.modules["h5py"] = had
    out = {"num_events": np.int64(mod.get_num_events("recording.h5"))}
    def summary(d):
        return np.array([len(d["t"]), int(d["t"][0]) if len(d["t"]) else -1, int(d["t"][-1]) if len(d["t"]) else -1,
                         int(d["x"].astype(np.int64).sum()), int(d["y"].astype(np.int64).sum()), int(d["p"].astype(np.int64).sum()),
                         int(d["t"].dtype == np.int64)], dtype=np.int64)
    out["windows"] = np.stack([summary(mod.extract_from_h5_by_timewindow("recording.h5", a, b)) for a, b in IO_WINDOWS])
    out["ranges"] = np.stack([summary(mod.extract_from_h5_by_index("recording.h5", a, b)) for a, b in IO_RANGES])
    np.savez_compressed(os.path.join(OUT, "dsec_io.npz"), **out)
    print("dsec_io.npz:", out["windows"][:, 0].tolist(), out["ranges"][:, 0].tolist())


def make_negwins_golden():
    """Second raster mode through the REAL generate_event_image_vectorized
    (feasible/my_egpt_dsec_dataset/preprocess_event_images.py:58-90): the five slices of a DSEC window with the data-dependent
    canvas and with an explicit 480x640 canvas, plus a signed-polarity window (p = -1 draws nothing in this mode)."""
    import contextlib, importlib.util, io
    spec = importlib.util.spec_from_file_location(
        "ref_preprocess_event_images", os.path.join(REF, "feasible", "my_egpt_dsec_dataset", "preprocess_event_images.py"))
    mod = importlib.util.module_from_spec(spec)
    with contextlib.redirect_stdout(io.StringIO()):
        spec.loader.exec_module(mod)
    from oracle import events as oev
    out = {}
    for name in ("dsec_000099", "adv_signed_p"):
        z = np.load(os.path.join(OUT, f"events_{name}.npz"))
        x, y, p = z["x"], z["y"], z["p"]
        starts, ends = oev.frame_bounds(len(x), 5)
        for i, (s_, e_) in enumerate(zip(starts, ends)):
            out[f"{name}_auto_{i}"] = mod.generate_event_image_vectorized(x[s_:e_], y[s_:e_], p[s_:e_])
            out[f"{name}_vga_{i}"] = mod.generate_event_image_vectorized(x[s_:e_], y[s_:e_], p[s_:e_], 480, 640)
    np.savez_compressed(os.path.join(OUT, "negwins.npz"), **out)
    print("negwins.npz:", len(out), "frames")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "tokens_full":
    make_full_token_goldens(sys.argv[2] if len(sys.argv) > 2 else None)
    sys.exit(0)

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "negwins":
        make_negwins_golden()
    elif len(sys.argv) > 1 and sys.argv[1] == "io":
        make_io_golden()
    elif len(sys.argv) > 1 and sys.argv[1] == "splice":
        make_splice_golden()
    elif len(sys.argv) > 1 and sys.argv[1] == "clips":
        make_clips_golden()
    else:
        main()
        make_splice_golden()
        make_clips_golden()
        make_io_golden()
        make_negwins_golden()
