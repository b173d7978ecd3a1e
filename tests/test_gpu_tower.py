"""GPU parity tests (through the C ABI) for the floating-point half of the path: GEMM,
attention, tower, projector, adaptor, pooling, and the fused events -> tokens call.

Tolerance (BASELINE.json north_star): token features within max-abs 2e-2 and cosine >= 0.999
of the reference's fp32 CPU path.  The reference's own bf16 execution measures max-abs 1.8e-2
on the same weights (tests/golden/yardstick_bf16.npz), reported beside ours."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, load_golden_events

pytestmark = pytest.mark.gpu
MAX_ABS, MIN_COS = 2e-2, 0.999


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "-m gpu tests need a GPU"
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def sample_frames(dev):
    from oracle import preprocess as opp
    _, _, pix = load_golden_events("sample1")
    return [torch.from_numpy(opp.bf16_bits_to_f32(p)).to(dev).bfloat16() for p in pix]


def _cos(a, b):
    return torch.nn.functional.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


@pytest.mark.parametrize("shape", [(128, 256, 64), (100, 256, 256), (577, 1024, 1024), (2885, 3072, 1024), (1154, 1024, 4096)])
@pytest.mark.parametrize("epi", [0, 1, 2, 3, 4])
def test_gemm_vs_fp32_reference(shape, epi, dev):
    """tcgen05 GEMM + fused epilogues vs a plain torch fp32 matmul of the same bf16 operands."""
    from eventgpt_b200 import events as E
    from eventgpt_b200._lib import lib, check
    M, N, K = shape
    g = torch.Generator(device="cpu").manual_seed(M + N + K + epi)
    A = (torch.randn(M, K, generator=g) * 0.5).bfloat16().to(dev)
    B = (torch.randn(N, K, generator=g) * (K ** -0.5)).bfloat16().to(dev)
    bias = (torch.randn(N, generator=g) * 0.1).to(dev)
    ref = A.float() @ B.float().t() + bias
    if epi == 1:
        ref = ref * torch.sigmoid(1.702 * ref)
    elif epi == 2:
        ref = torch.nn.functional.gelu(ref)
    if epi == 3:
        out = torch.randn(M, N, generator=g).to(dev)
        ref = ref + out
    elif epi == 4:
        out = torch.empty(M, N, device=dev)
    else:
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    check(lib().egpt_gemm_bf16(E.ptr(A), E.ptr(B), E.ptr(bias), E.ptr(out), M, N, K, epi, E.stream_ptr(dev)), "gemm")
    torch.cuda.synchronize()
    tol = 2e-5 * K ** 0.5 if epi >= 3 else 2e-2      # fp32 outputs: accumulation order only; bf16 outputs: rounding
    assert (out.float() - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item())


def test_attention_vs_fp32_reference(dev):
    from eventgpt_b200 import events as E
    from eventgpt_b200._lib import lib, check
    g = torch.Generator(device="cpu").manual_seed(1)
    F = 3
    qkv = torch.randn(F * 577, 3072, generator=g).bfloat16().to(dev)
    out = torch.empty(F * 577, 1024, device=dev, dtype=torch.bfloat16)
    check(lib().egpt_attention(E.ptr(qkv), F, E.ptr(out), E.stream_ptr(dev)), "attention")
    q, k, v = (qkv.float().view(F, 577, 3, 16, 64)[:, :, i].transpose(1, 2) for i in range(3))
    ref = (torch.softmax(q @ k.transpose(-1, -2), -1) @ v).transpose(1, 2).reshape(F * 577, 1024)
    assert (out.float() - ref).abs().max().item() < 3e-2      # |v| <~ 4, P rounded to bf16
    assert _cos(out.float(), ref) > 0.9999


@pytest.mark.parametrize("F", [1, 2, 7, 38])
def test_attention_persistent_ranges(F, dev):
    """The attention kernel is persistent: each of (at most) 148 CTAs walks a contiguous range of (frame, head, Q tile) units.
    F = 1: fewer units than SMs, one tile per CTA; F = 2: ranges of one or two tiles that straddle heads; F = 7: 3-4 tiles, most
    CTAs start mid-head; F = 38: ~20 tiles = four heads per CTA (shared-memory and TMEM reuse across heads, the deferred merge
    of a head's last tile, double-buffered rows of the left-over key).  Heads get different score scales, and the key that is
    handled outside the MMAs (576) is made the dominant one for some rows."""
    from eventgpt_b200 import events as E
    from eventgpt_b200._lib import lib, check
    g = torch.Generator(device="cpu").manual_seed(100 + F)
    qkv = torch.randn(F, 577, 3, 16, 64, generator=g)
    qkv[:, :, 0] *= torch.linspace(0.2, 1.5, 16)[None, None, :, None]         # per-head temperature
    qkv[:, 576, 1, 3] = qkv[:, 100, 0, 3] * 3.0                                # head 3: key 576 dominates row 100 (and friends)
    x = qkv.reshape(F * 577, 3072).bfloat16().to(dev)
    out = torch.full((F * 577, 1024), float("nan"), device=dev, dtype=torch.bfloat16)
    check(lib().egpt_attention(E.ptr(x), F, E.ptr(out), E.stream_ptr(dev)), "attention")
    xf = x.float().view(F, 577, 3, 16, 64)
    q, k, v = (xf[:, :, i].transpose(1, 2) for i in range(3))
    p = torch.softmax(q @ k.transpose(-1, -2), -1)
    assert p[:, 3, 100, 576].min().item() > 0.5                                # the left-over key really matters there
    ref = (p @ v).transpose(1, 2).reshape(F * 577, 1024)
    assert torch.isfinite(out.float()).all()
    assert (out.float() - ref).abs().max().item() < 3e-2
    assert _cos(out.float(), ref) > 0.9999
    # a second launch on the same buffers gives the same bits (no state carried between launches)
    out2 = torch.empty_like(out)
    check(lib().egpt_attention(E.ptr(x), F, E.ptr(out2), E.stream_ptr(dev)), "attention")
    assert torch.equal(out, out2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_get_spatio_temporal_features(dtype, dev):
    """model/EventChatModel.py:15-38 incl. zero padding / truncation of the temporal tokens."""
    from eventgpt_b200.model import get_spatio_temporal_features
    from oracle import vit as ovit
    g = torch.Generator(device="cpu").manual_seed(2)
    f = torch.randn(5, 577, 4096, generator=g).to(dtype)
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    for nt in (None, 8, 2):
        got = get_spatio_temporal_features(list(f.to(dev)) if nt is None else f.to(dev), nt)
        ref = ovit.get_spatio_temporal_features(f.float(), nt)
        assert got.dtype == dtype and tuple(got.shape) == tuple(ref.shape)
        assert (got.float().cpu() - ref).abs().max().item() < tol
    many = torch.randn(13, 577, 256, generator=g).to(dtype)             # more than 8 frames: the two-walk kernel
    for nt in (None, 20, 4):
        got = get_spatio_temporal_features(many.to(dev), nt)
        ref = ovit.get_spatio_temporal_features(many.float(), nt)
        assert got.dtype == dtype and tuple(got.shape) == tuple(ref.shape)
        assert (got.float().cpu() - ref).abs().max().item() < tol
    small = torch.randn(3, 10, 64, generator=g).to(dev)
    ref = ovit.get_spatio_temporal_features(small.cpu())
    assert (get_spatio_temporal_features(small).cpu() - ref).abs().max().item() < 1e-5
    with pytest.raises(ValueError):
        get_spatio_temporal_features(torch.zeros(5, 7, device=dev))


@pytest.fixture(scope="module")
def frontends(dev):
    from eventgpt_b200.model import EventGPTVisualFrontEnd
    from eventgpt_b200.weights import random_state_dict
    made = {}

    def get(layers):
        if layers not in made:
            sd = random_state_dict(0, layers)
            made[layers] = (sd, EventGPTVisualFrontEnd(sd, dev, num_layers=layers))
        return made[layers]
    yield get
    for _, fe in made.values():
        fe.close()


@pytest.mark.parametrize("layers", [2, 24])
def test_tokens_vs_reference_golden(layers, frontends, sample_frames, dev):
    """List branch of prepare_inputs_labels_for_multimodal (model/EventChatModel.py:326-343)
    vs the reference model's own fp32 output (tests/golden/tokens_sample1_L*.npz)."""
    z = np.load(os.path.join(GOLDEN, f"tokens_sample1_L{layers}.npz"))
    sd, fe = frontends(layers)
    wsum = float(sum(v.double().abs().sum() for v in sd.values()))
    assert abs(wsum - float(z["weight_abs_sum"])) < 1e-6 * wsum
    # visval_encode, frame 0 (:194-200)
    pr = fe.visval_encode(sample_frames[0][None]).float().cpu()
    assert tuple(pr.shape) == (1, 577, 4096)
    assert (pr[0, :, ::16] - torch.from_numpy(z["proj0_sub"])).abs().max().item() < 2 * MAX_ABS
    # event_features through the list branch
    tok = fe.encode_events(sample_frames).float().cpu()
    assert tuple(tok.shape) == (1, 582, 4096)
    ref_sub, ref_rows = torch.from_numpy(z["tokens_sub"]), torch.from_numpy(z["tokens_rows"])
    err = max((tok[0, :, ::16] - ref_sub).abs().max().item(), (tok[0, [0, 4, 5, 300, 581]] - ref_rows).abs().max().item())
    cos = _cos(tok[0, :, ::16], ref_sub)
    y = np.load(os.path.join(GOLDEN, "yardstick_bf16.npz"))
    print(f"L={layers}: ours-bf16 vs ref-fp32 max-abs {err:.4g} cos {cos:.6f}; "
          f"torch-bf16 vs ref-fp32 (L=24) max-abs {float(y['max_abs']):.4g} cos {float(y['cosine']):.6f}")
    assert err < MAX_ABS and cos > MIN_COS
    # list of lists (batch of 2 samples) == the same sample twice
    tok2 = fe.encode_events([sample_frames, sample_frames]).float().cpu()
    assert tuple(tok2.shape) == (2, 582, 4096)
    assert torch.equal(tok2[0], tok2[1]) and torch.equal(tok2[0], tok[0])
    # tensor branch (:344-345): visval_encode only -- no adaptor, no pooling
    tb = fe.encode_events(torch.stack(sample_frames[:2]))
    assert tuple(tb.shape) == (2, 577, 4096) and torch.equal(tb[0].float().cpu(), pr[0])


def test_stagewise_vs_oracle_two_layers(frontends, sample_frames, dev):
    """tower -> projector -> adaptor -> pooling, each stage against the fp32 oracle run here."""
    from eventgpt_b200.model import get_spatio_temporal_features
    from oracle import vit as ovit
    sd, fe = frontends(2)
    x = torch.stack(sample_frames[:2])
    with torch.no_grad():
        ref_h = ovit.tower_last_hidden_state(sd, x.float().cpu(), 2)
        ref_p = ovit.visual_projector(sd, ref_h)
        ref_a = ovit.feature_adaptor(sd, ref_p)
    h = fe.tower(x).float().cpu()
    assert (h - ref_h).abs().max().item() < 5e-2 and _cos(h, ref_h) > 0.9999      # |h| ~ 4.6: bf16 output rounding
    pr = fe.visval_encode(x)
    assert (pr.float().cpu() - ref_p).abs().max().item() < MAX_ABS
    ad = fe.feature_adaptor(pr).float().cpu()
    assert (ad - ref_a).abs().max().item() < MAX_ABS and _cos(ad, ref_a) > MIN_COS
    # reference order (adaptor per frame, then pooling) through the separate calls == fused order within tolerance
    tok_ref_order = get_spatio_temporal_features([fe.feature_adaptor(f) for f in fe.visval_encode(torch.stack(sample_frames))])
    tok_fused = fe.encode_events(sample_frames)[0]
    assert (tok_ref_order.float() - tok_fused.float()).abs().max().item() < MAX_ABS


def test_fused_events_to_tokens(frontends, dev):
    """egpt_visual_tokens (events -> tokens, no host round trip) == process_event_data +
    encode_events, for a batch of ragged windows; and the golden tokens for sample1."""
    from eventgpt_b200 import events as E
    sd, fe = frontends(2)
    names = ["sample1", "dsec_000099", "adv_zipf_vga"]
    evs = [load_golden_events(n) for n in names]
    cat = {k: np.concatenate([e[0][k].astype(np.uint16 if k != "p" else np.uint8) for e in evs]) for k in "xyp"}
    sizes = [len(e[0]["t"]) for e in evs]
    x, y, p = (E.to_device(cat[k], dev) for k in "xyp")
    off = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int64, device=dev)
    tok = fe.visual_tokens(x, y, p, off, (480, 640))
    assert tuple(tok.shape) == (3, 582, 4096)
    from oracle import preprocess as opp
    for b, (_, _, pix) in enumerate(evs):
        frames = [torch.from_numpy(opp.bf16_bits_to_f32(q)).to(dev).bfloat16() for q in pix]
        want = fe.encode_events(frames)[0]
        assert torch.equal(tok[b], want), names[b]
    z = np.load(os.path.join(GOLDEN, "tokens_sample1_L2.npz"))
    assert (tok[0, :, ::16].float().cpu() - torch.from_numpy(z["tokens_sub"])).abs().max().item() < MAX_ABS
    with pytest.raises(ValueError):     # a window with fewer events than frames: empty slice
        fe.visual_tokens(x[:3], y[:3], p[:3], torch.tensor([0, 3], dtype=torch.int64, device=dev), (480, 640))


@pytest.mark.parametrize("case,weights", [("sample1", "random"), ("dsec_000069", "random"), ("dense1m", "random"),
                                          ("sample1", "outlier")])
def test_full_token_goldens_24_layers(case, weights, frontends, dev):
    """events -> tokens, 24 layers, against the FULL [582, 4096] output of the reference model itself
    (tests/golden/tokens_full_*.npz, float16-packed by oracle/make_golden.py tokens_full): the reference's own sample, a DSEC
    window, a window of the benchmark workload (dense 1M events), and sample1 under weights with massive-activation channels
    (~70x the median channel in the last hidden state) and sharply peaked attention heads.  Bound: max-abs 2e-2 and cosine
    >= 0.999 (north_star); with the outlier weights the reference's OWN bf16 run is at 6.9e-2, so the bound there is that
    yardstick.  float16 packing of the golden adds at most half an ulp of the largest token (stated per case)."""
    import hashlib
    from eventgpt_b200 import events as E
    from eventgpt_b200.model import EventGPTVisualFrontEnd
    from eventgpt_b200.weights import outlier_state_dict
    z = np.load(os.path.join(GOLDEN, f"tokens_full_{case}_{weights}.npz"))
    if case == "dense1m":
        from oracle.make_golden import synth_window          # the benchmark's generator, restated next to the golden
        x, y, p = synth_window(0)
        ev = dict(x=x, y=y, p=p)
    else:
        ev = load_golden_events(case)[0]
    n = len(ev["x"])
    xd, yd, pd = (E.to_device(np.ascontiguousarray(ev[k]), dev) for k in "xyp")
    off = torch.tensor([0, n], dtype=torch.int64, device=dev)
    # the pixel values that reach the tower are the reference's, bit for bit
    fr, hw, _, st = E.raster(xd, yd, pd, off, 1, 5, 480, 640)
    px, _, st = E.preprocess(fr, hw, 480, status=st)
    E.check_status(st, case)
    assert hashlib.sha256(px.view(torch.int16).cpu().numpy().view(np.uint16).tobytes()).hexdigest() == str(z["pix_sha"])
    if weights == "random":
        sd, fe = frontends(24)
        own = None
    else:
        sd = outlier_state_dict(1, 24)
        fe = own = EventGPTVisualFrontEnd(sd, dev, num_layers=24)
    try:
        assert abs(float(sum(v.double().abs().sum() for v in sd.values())) - float(z["weight_abs_sum"])) < 1e-6 * float(z["weight_abs_sum"])
        tok = fe.visual_tokens(xd, yd, pd, off, (480, 640))[0].float().cpu()
    finally:
        if own is not None:
            own.close()
    ref = torch.from_numpy(z["tokens_f16"].astype(np.float32))
    quant = float(z["absmax"]) * 2.0 ** -11
    err = (tok - ref).abs().max().item()
    err_rows = (tok[[0, 4, 5, 300, 581]] - torch.from_numpy(z["tokens_rows"])).abs().max().item()     # fp32 rows of the golden
    cos = _cos(tok, ref)
    bound = MAX_ABS if weights == "random" else max(MAX_ABS, float(z["bf16_max_abs"]))
    print(f"{case}/{weights}: ours-bf16 vs ref-fp32 max-abs {err:.4g} (fp32 rows {err_rows:.4g}, f16 packing <= {quant:.2g}) "
          f"cos {cos:.6f}; torch-bf16 vs ref-fp32 max-abs {float(z['bf16_max_abs']):.4g} cos {float(z['bf16_cosine']):.6f}")
    assert tuple(tok.shape) == (582, 4096)
    assert err < bound + quant and err_rows < bound and cos > MIN_COS


def test_full_batch_properties(frontends, dev):
    """BASELINE config 2 shape through the fused call at reduced depth: every window of the batch
    equals the same window processed alone (windows are independent; no cross-window state)."""
    from eventgpt_b200 import events as E
    sd, fe = frontends(2)
    rng = np.random.default_rng(2)
    B, N, H, W = 8, 200_000, 480, 640
    x = torch.from_numpy(rng.integers(0, W, B * N).astype(np.uint16)).to(dev)
    y = torch.from_numpy(rng.integers(0, H, B * N).astype(np.uint16)).to(dev)
    p = torch.from_numpy(rng.integers(0, 2, B * N).astype(np.uint8)).to(dev)
    off = torch.arange(0, B + 1, dtype=torch.int64, device=dev) * N
    tok = fe.visual_tokens(x, y, p, off, (H, W)).clone()
    assert torch.isfinite(tok.float()).all()
    for b in (0, 5):
        one = fe.visual_tokens(x[b * N:(b + 1) * N], y[b * N:(b + 1) * N], p[b * N:(b + 1) * N],
                               torch.tensor([0, N], dtype=torch.int64, device=dev), (H, W))
        assert torch.equal(one[0], tok[b])


def test_multimodal_splice_bit_exact(dev):
    """prepare_inputs_labels_for_multimodal's splice (model/EventChatModel.py:355-465) through
    egpt_splice_embeddings vs the reference's own output (tests/golden/splice.npz) and the oracle."""
    from eventgpt_b200.model import splice_event_features
    from oracle import splice as osp
    from oracle.make_golden import splice_inputs
    z = np.load(os.path.join(GOLDEN, "splice.npz"))
    table, feats = splice_inputs()
    ids, mask, labels = (torch.from_numpy(z[k]) for k in ("ids", "mask", "labels"))
    td, fd = table.to(dev).bfloat16(), feats.to(dev).bfloat16()
    pos, am, emb, lab = splice_event_features(td, ids, None, mask, labels, fd)
    assert pos is None and emb.dtype == torch.bfloat16 and tuple(emb.shape) == (3, 2048, 4096)
    e = emb.float().cpu()
    assert np.array_equal(e[:, :, ::64].numpy(), z["emb_sub"]) and np.allclose(e.double().sum(-1).numpy(), z["emb_rowsum"])
    assert np.array_equal(am.numpy(), z["am"]) and am.dtype == mask.dtype and np.array_equal(lab.numpy(), z["lab"])
    o = osp.splice(table, ids, None, mask, labels, feats)
    assert torch.equal(e, o[2])
    pid = torch.arange(ids.shape[1])[None].expand(3, -1).contiguous()
    pos2, am2, emb2, lab2 = splice_event_features(td, ids.to(dev), pid.to(dev), None, None, fd)
    assert am2 is None and lab2 is None and pos2.is_cuda and np.array_equal(pos2.cpu().numpy(), z["pos2"])
    assert np.allclose(emb2.float().cpu().double().sum(-1).numpy(), z["emb2_rowsum"])
    with pytest.raises(IndexError):            # more <event> tokens than feature entries, as upstream
        splice_event_features(td, ids, None, mask, labels, fd[:2])


def test_host_fed_stream_matches_direct_call(frontends, dev):
    """eventgpt_b200.pipeline.VisualTokenStream (pinned host events in, pinned host tokens out, copies overlapped
    with compute over two slots) returns exactly what the direct device call returns, batch after batch."""
    from eventgpt_b200 import events as E
    from eventgpt_b200.pipeline import VisualTokenStream
    sd, fe = frontends(2)
    rng = np.random.default_rng(7)
    B, N, H, W = 2, 60_000, 240, 320
    vs = VisualTokenStream(fe, B * N, B, (H, W))
    batches, slots = [], []
    for i in range(5):                                   # more batches than slots: buffers are reused
        hx = torch.from_numpy(rng.integers(0, W, B * N).astype(np.uint16)).pin_memory()
        hy = torch.from_numpy(rng.integers(0, H, B * N).astype(np.uint16)).pin_memory()
        hp = torch.from_numpy(rng.integers(0, 2, B * N).astype(np.uint8)).pin_memory()
        hoff = (torch.arange(0, B + 1, dtype=torch.int64) * N).pin_memory()
        batches.append((hx, hy, hp, hoff))
    got = []
    for b in batches:
        s = vs.submit(*b)
        got.append(vs.wait(s)[:B].clone())               # read the slot before it is reused
    vs.drain()
    for (hx, hy, hp, hoff), g in zip(batches, got):
        want = fe.visual_tokens(hx.to(dev), hy.to(dev), hp.to(dev), hoff.to(dev), (H, W)).cpu()
        assert torch.equal(g, want)
    # a window with fewer events than frames has an empty slice: the reference raises ValueError from y.max()
    # (common/common.py:65); the serving loop must deliver it for that batch -- in wait(), or at the latest when the
    # slot is reused or drained -- and carry on with the next batch
    hx, hy, hp, _ = batches[0]
    bad_off = torch.tensor([0, 3, 2 * N], dtype=torch.int64).pin_memory()
    s = vs.submit(hx, hy, hp, bad_off)
    with pytest.raises(ValueError):
        vs.wait(s)
    vs.submit(hx, hy, hp, bad_off)
    with pytest.raises(ValueError):
        vs.drain()
    s = vs.submit(*batches[1])
    assert torch.equal(vs.wait(s)[:B], got[1])
    vs.drain()


def test_stream_recording_clips_on_device(frontends, dev):
    """VisualTokenStream.submit_recording: a time-sorted recording goes to the device once; the 50 %-overlap clip bounds
    (build_my_egpt_dsec_seq.py:137-202) are found there and the clips' tokens equal those of the host-side clip splitter +
    ragged batch (eventgpt_b200.io), clip by clip."""
    from eventgpt_b200 import events as E, io as eio
    from eventgpt_b200.pipeline import VisualTokenStream
    sd, fe = frontends(2)
    rng = np.random.default_rng(33)
    n, H, W = 90_000, 240, 320
    t = np.sort(rng.integers(5_000_000, 5_300_000, n)).astype(np.int64)
    ev = dict(x=rng.integers(0, W, n).astype(np.uint16), y=rng.integers(0, H, n).astype(np.uint16),
              p=rng.integers(0, 2, n).astype(np.uint8), t=t)
    stamps = [5_000_000 + 50_000 * i for i in range(7)]
    clips = eio.split_event_by_time_clips(ev, 50_000, stamps)
    n_clips = (max(stamps) - 50_000 - min(stamps)) // 25_000 + 1
    assert len(clips) == n_clips == 11
    x, y, p, off = eio.clips_to_batch(ev, clips)
    want = fe.visual_tokens(E.to_device(x, dev), E.to_device(y, dev), E.to_device(p, dev),
                            torch.tensor(off, dtype=torch.int64, device=dev), (H, W)).cpu()
    vs = VisualTokenStream(fe, n, n_clips, (H, W))
    pin = lambda a: torch.from_numpy(a).pin_memory()
    s = vs.submit_recording(pin(ev["x"]), pin(ev["y"]), pin(ev["p"]), pin(ev["t"]), min(stamps), 50_000, 25_000, n_clips)
    got = vs.wait(s)[:n_clips].clone()
    vs.drain()
    assert torch.equal(got, want)


def test_attention_lazy_rescale_and_peaked_rows(dev):
    """The attention kernel keeps a stale row maximum unless a later key block exceeds it by more than 2^8, and then
    rescales its TMEM accumulator in place.  Drive both paths: (a) scores that grow block after block (every block
    rescales), (b) one dominant key in the LAST block (a single late rescale), (c) nearly constant scores (never)."""
    from eventgpt_b200 import events as E
    from eventgpt_b200._lib import lib, check
    g = torch.Generator(device="cpu").manual_seed(5)
    F = 2
    qkv = torch.randn(F, 577, 3, 16, 64, generator=g) * 0.3
    # (a) head 0: keys grow with their index -> every later block has a much larger maximum
    ramp = torch.linspace(0.0, 6.0, 577)[None, :, None]
    qkv[:, :, 1, 0] = qkv[:, :, 1, 0] + ramp * qkv[:, :1, 0, 0].sign()      # aligned with the sign pattern of row 0's q
    qkv[:, :, 0, 0] = qkv[:, :, 0, 0].abs() * qkv[:, :1, 0, 0].sign()       # all queries share that sign pattern
    # (b) head 1: key 570 (last block) dominates every row
    qkv[:, :, 0, 1] = qkv[:, :, 0, 1].abs()
    qkv[:, 570, 1, 1] = 8.0
    # (c) head 2: tiny q -> flat softmax
    qkv[:, :, 0, 2] *= 1e-3
    x = qkv.reshape(F * 577, 3072).bfloat16().to(dev)
    out = torch.empty(F * 577, 1024, device=dev, dtype=torch.bfloat16)
    check(lib().egpt_attention(E.ptr(x), F, E.ptr(out), E.stream_ptr(dev)), "attention")
    xf = x.float().view(F, 577, 3, 16, 64)
    q, k, v = (xf[:, :, i].transpose(1, 2) for i in range(3))
    s = q @ k.transpose(-1, -2)
    assert (s[:, 0].amax(-1) - s[:, 0, :, :192].amax(-1)).max().item() > 20.0        # (a) really needs rescaling
    ref = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(F * 577, 1024)
    err = (out.float() - ref).abs().max().item()
    assert torch.isfinite(out.float()).all() and err < 3e-2, err
    assert _cos(out.float(), ref) > 0.9999


def test_generate_front_half(frontends, dev):
    """model/EventChatModel.py:246-292 up to the hand-over to the language model: the two NotImplementedErrors, the cached
    `event_features` branch (adaptor re-applied, as upstream) and the tensors that upstream forwards to `super().generate`."""
    from eventgpt_b200.model import splice_event_features
    sd, fe = frontends(2)
    g = torch.Generator(device="cpu").manual_seed(9)
    table = (torch.randn(320, 4096, generator=g) * 0.02).bfloat16().to(dev)
    ids = torch.randint(0, 300, (2, 12), generator=g)
    ids[0, 3] = -200
    ids[1, 7] = -200                                         # EVENT_TOKEN_INDEX
    ids = ids.to(dev)
    with pytest.raises(NotImplementedError, match="please input Event"):
        fe.generate(table, ids)
    with pytest.raises(NotImplementedError, match="inputs_embeds"):
        fe.generate(table, ids, event_features=torch.zeros(2, 582, 4096), inputs_embeds=torch.zeros(1))
    feats = (torch.randn(2, 582, 4096, generator=g) * 0.1).bfloat16().to(dev)
    out = fe.generate(table, ids, event_features=feats, max_new_tokens=7)
    assert out["max_new_tokens"] == 7 and out["position_ids"] is None and out["attention_mask"] is None
    want = splice_event_features(table, ids, None, None, None, fe.feature_adaptor(feats))[2]
    assert tuple(out["inputs_embeds"].shape) == (2, 11 + 582, 4096) and torch.equal(out["inputs_embeds"], want)

    class FakeLLM:                                           # the hand-over itself
        def generate(self, **kw):
            return sorted(kw)
    assert fe.generate(table, ids, event_features=feats, llm=FakeLLM(), do_sample=False) == [
        "attention_mask", "do_sample", "inputs_embeds", "position_ids"]


def test_fused_op_order_matches_reference_order(frontends, sample_frames, dev):
    """The fused token path pools the projector's GELU output BEFORE the projector's second linear and the adaptor (both affine,
    the pooling a mean).  Cross-check against the reference's own order built from the unfused entry points: visval_encode per
    frame -> feature_adaptor per token -> get_spatio_temporal_features (model/EventChatModel.py:333-341).  Same function, one
    rounding point moved: the two must agree to bf16 rounding noise."""
    from eventgpt_b200.model import get_spatio_temporal_features
    sd, fe = frontends(2)
    fused = fe.encode_events(sample_frames)[0].float()
    per_frame = [fe.feature_adaptor(fe.visval_encode(f[None]))[0] for f in sample_frames]     # reference order, per token
    ref_order = get_spatio_temporal_features(per_frame).float()
    assert tuple(fused.shape) == tuple(ref_order.shape) == (582, 4096)
    d = (fused - ref_order).abs()
    scale = ref_order.abs().max().item()
    assert d.max().item() < 2e-2 * max(scale, 1.0) and _cos(fused, ref_order) > 0.9999, (d.max().item(), scale)


def test_config5_events_to_inputs_embeds(frontends, dev):
    """BASELINE configs[4] shape at test scale: raw events of several ragged windows (one of 2 M events, as configs[2]) ->
    visual tokens -> multimodal splice -> `inputs_embeds`, all on the device.  The tokens of every window equal that window
    processed alone; the splice is bit-exact against the oracle splice fed with those tokens; prompt rows are untouched."""
    from eventgpt_b200 import events as E
    from oracle import splice as osp
    sd, fe = frontends(2)
    rng = np.random.default_rng(55)
    H, W = 480, 640
    sizes = [2_000_000, 300_000, 1_000_000]
    n = sum(sizes)
    x = torch.from_numpy(rng.integers(0, W, n).astype(np.uint16)).to(dev)
    y = torch.from_numpy(rng.integers(0, H, n).astype(np.uint16)).to(dev)
    p = torch.from_numpy(rng.integers(0, 2, n).astype(np.uint8)).to(dev)
    off = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int64, device=dev)
    tok = fe.visual_tokens(x, y, p, off, (H, W))
    assert tuple(tok.shape) == (3, 582, 4096) and torch.isfinite(tok.float()).all()
    s, e = int(off[1]), int(off[2])
    alone = fe.visual_tokens(x[s:e].clone(), y[s:e].clone(), p[s:e].clone(), torch.tensor([0, e - s], dtype=torch.int64, device=dev), (H, W))
    assert torch.equal(alone[0], tok[1])
    # prompts: one <event> each, different lengths, right padding via the attention mask
    g = torch.Generator(device="cpu").manual_seed(7)
    table = (torch.randn(512, 4096, generator=g) * 0.02).bfloat16()
    L = 40
    ids = torch.randint(0, 500, (3, L), generator=g)
    mask = torch.ones(3, L, dtype=torch.bool)
    for b, (pos, length) in enumerate([(5, 40), (0, 23), (30, 31)]):
        ids[b, pos] = -200
        mask[b, length:] = False
    out = fe.prepare_inputs_labels_for_multimodal(table.to(dev), ids.to(dev), None, mask.to(dev), None, None, event_features=None,
                                                  event_tensors=None)
    assert out[4] is None                                            # no events given: untouched, as upstream
    from eventgpt_b200.model import splice_event_features
    pos_ids, am, emb, lab = splice_event_features(table.to(dev), ids.to(dev), None, mask.to(dev), None, tok)
    want = osp.splice(table.float(), ids, None, mask, None, tok.float().cpu())
    assert tuple(emb.shape) == tuple(want[2].shape) == (3, 39 + 582, 4096)
    assert torch.equal(emb.float().cpu(), want[2]) and torch.equal(am.cpu(), want[1])
    assert torch.equal(emb[0, :5].cpu(), table[ids[0, :5]]) and torch.equal(emb[0, 5:5 + 582], tok[0])


@pytest.mark.parametrize("n", [1, 3, 8, 11])
def test_other_frame_counts(n, frontends, dev):
    """get_event_images_list(event_npy, n) takes any n (common/common.py:17-37; the inference caller passes 5): raster,
    preprocessing and the fused token path for 1, 3, 8 and 11 frames per window -> tokens [B, n + 577, 4096]."""
    from eventgpt_b200 import events as E
    from oracle import events as oev, preprocess as opp
    sd, fe = frontends(2)
    ev, _, _ = load_golden_events("dsec_000099")
    H, W = 480, 640
    x, y, p = (E.to_device(ev[k].astype(np.uint16 if k != "p" else np.uint8), dev) for k in "xyp")
    off = torch.tensor([0, len(ev["t"])], dtype=torch.int64, device=dev)
    fr, hw, _, st = E.raster(x, y, p, off, 1, n, H, W)
    E.check_status(st, "raster")
    want_frames = oev.get_event_images_list(ev, n)
    got_frames = E.frames_to_numpy(fr, hw)
    assert len(got_frames) == n and all(np.array_equal(a, b) for a, b in zip(got_frames, want_frames))
    tok = fe.visual_tokens(x, y, p, off, (H, W), n_frames=n)
    assert tuple(tok.shape) == (1, n + 577, 4096)
    pix = [torch.from_numpy(opp.bf16_bits_to_f32(opp.pixel_values_bf16_bits(f))).to(dev).bfloat16() for f in want_frames]
    want = fe.encode_events(pix)
    assert torch.equal(tok[0], want[0])


def test_two_threads_two_streams(frontends, dev):
    """Re-entrancy per (handle, stream, workspace): two Python threads drive the same front end on two CUDA streams at once
    (the reference's e2e benchmark does this with its two prefills, pipeline/benchmark_e2e/benchmark_e2e_wallclock.py:641-649).
    Each thread's tokens equal the ones computed alone."""
    import threading
    from eventgpt_b200 import events as E
    sd, fe = frontends(2)
    rng = np.random.default_rng(77)
    H, W, N = 480, 640, 400_000
    data = []
    for k in range(2):
        x = torch.from_numpy(rng.integers(0, W, N).astype(np.uint16)).to(dev)
        y = torch.from_numpy(rng.integers(0, H, N).astype(np.uint16)).to(dev)
        p = torch.from_numpy(rng.integers(0, 2, N).astype(np.uint8)).to(dev)
        off = torch.tensor([0, N // 2, N], dtype=torch.int64, device=dev)
        data.append((x, y, p, off))
    alone = [fe.visual_tokens(*d, (H, W)).clone() for d in data]
    torch.cuda.synchronize(dev)
    out, errs = [None, None], []

    def work(k):
        try:
            s = torch.cuda.Stream(dev)
            with torch.cuda.stream(s):
                for _ in range(3):
                    out[k] = fe.visual_tokens(*data[k], (H, W))
            s.synchronize()
        except Exception as e:      # pragma: no cover
            errs.append(e)

    ts = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    for k in range(2):
        assert torch.equal(out[k], alone[k]), k


def test_torch_custom_ops(frontends, dev):
    """torch.ops.eventgpt_b200.* (eventgpt_b200/ops.py): the operators the patched reference calls
    (inference.py:49-63 -> process_event_data; model/EventChatModel.py:326-345 -> encode_events) return what the Python
    mirror returns, and pass torch.library.opcheck (schema, fake tensor, dispatch registration)."""
    import eventgpt_b200.ops  # noqa: F401
    from eventgpt_b200 import events as E
    ops = torch.ops.eventgpt_b200
    sd, fe = frontends(2)
    ev = load_golden_events("synthetic_000003")[0]
    n = len(ev["t"])
    x, y, p = (E.to_device(ev[k], dev) for k in "xyp")
    off = torch.tensor([0, n // 2, n], dtype=torch.int64, device=dev)
    H, W = int(ev["y"].max()) + 1, int(ev["x"].max()) + 1
    tok, hw, st = ops.visual_tokens(fe.handle_value, x, y, p, off, 5, H, W)
    assert int(st.item()) == 0
    assert torch.equal(tok, fe.visual_tokens(x, y, p, off, (H, W)))
    fr, hw2, st = ops.raster(x, y, p, off, 5, H, W, 0, False)
    assert torch.equal(hw, hw2)
    pix, st = ops.preprocess(fr, hw2, H)
    assert torch.equal(ops.encode_pool(fe.handle_value, pix, 5), tok)
    assert torch.equal(ops.project_pool(fe.handle_value, fe.tower(pix), 5), tok)
    feats = ops.feature_adaptor(fe.handle_value, ops.visval_encode(fe.handle_value, pix))
    pooled = ops.spatio_temporal_pool(feats.reshape(2, 5, 577, 4096), 5)
    assert (pooled.float() - tok.float()).abs().max().item() < 2 * MAX_ABS          # reference op order vs fused order
    table = torch.randn(50, 4096, device=dev).bfloat16()
    src = torch.tensor([3, -2, -3, -1, 7], dtype=torch.int32, device=dev)
    out = ops.splice(table, tok, src)
    assert torch.equal(out[0], table[3]) and torch.equal(out[1], tok[0, 0]) and torch.equal(out[2], tok[0, 1])
    assert not out[3].any() and torch.equal(out[4], table[7])
    tests = ("test_schema", "test_faketensor", "test_autograd_registration")
    torch.library.opcheck(ops.visual_tokens.default, (fe.handle_value, x, y, p, off, 5, H, W), test_utils=tests)
    torch.library.opcheck(ops.raster.default, (x, y, p, off, 5, H, W, 0, False), test_utils=tests)
    torch.library.opcheck(ops.encode_pool.default, (fe.handle_value, pix, 5), test_utils=tests)
    torch.library.opcheck(ops.splice.default, (table, tok, src), test_utils=tests)
    # cached tower outputs through the Python mirror too
    assert torch.equal(fe.project_pool(fe.tower(pix)), tok)
