"""GPU bring-up: runs each stage of the path in its own subprocess (own timeout) and prints
parity + timing.  Developer tool; the judged checks are tests/ -m gpu and bench.py.

    python tools/bringup.py [stage ...]        stages: events gemm attn misc tower fused
"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def cuda_time(fn, iters=5, warm=2):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def stage_events():
    import numpy as np
    import torch
    from conftest import golden_event_cases, load_golden_events, GOLDEN
    from eventgpt_b200 import events as E, common as Cm
    dev = torch.device("cuda:0")
    bad = 0
    for name in golden_event_cases():
        ev, frames, pix = load_golden_events(name)
        try:
            got = Cm.get_event_images_list(ev, 5)
            ok_r = all(g.shape == f.shape and np.array_equal(g, f) for g, f in zip(got, frames))
            n = len(ev["t"])
            H, W = int(ev["y"].max()) + 1, int(ev["x"].max()) + 1
            xd, yd, pd = (E.to_device(ev[k], dev) for k in "xyp")
            off = torch.tensor([0, n], dtype=torch.int64, device=dev)
            fr, hw, _, st = E.raster(xd, yd, pd, off, 1, 5, H, W)
            px, _, st = E.preprocess(fr, hw, H, status=st)
            gp = px.view(torch.int16).cpu().numpy().view(np.uint16)
            ok_p = np.array_equal(gp, pix)
            nd = int((gp != pix).sum())
            print(f"{name:24s} raster {'OK' if ok_r else 'MISMATCH'}  preprocess {'OK' if ok_p else 'MISMATCH(%d)' % nd} status {int(st.item())}")
            bad += (not ok_r) + (not ok_p)
        except Exception as ex:  # noqa
            print(f"{name:24s} EXCEPTION {type(ex).__name__}: {ex}")
            bad += 1
    z = np.load(os.path.join(GOLDEN, "split_stream.npz"))
    parts = Cm.split_event_by_time({k: z[k] for k in "xytp"}, 50000)
    ok = [len(q["t"]) for q in parts] == z["counts"].tolist()
    for k in "xytp":
        ok = ok and np.array_equal(np.concatenate([q[k] for q in parts]), z["out_" + k])
    print("split_event_by_time", "OK" if ok else "MISMATCH")
    bad += not ok
    # raster timing: config 2 shape, B=32 windows x 1M events at 640x480
    rng = np.random.default_rng(2)
    B, N, H, W = 32, 1_000_000, 480, 640
    x = torch.from_numpy(rng.integers(0, W, B * N).astype(np.uint16)).to(dev)
    y = torch.from_numpy(rng.integers(0, H, B * N).astype(np.uint16)).to(dev)
    p = torch.from_numpy(rng.integers(0, 2, B * N).astype(np.uint8)).to(dev)
    off = torch.arange(0, B + 1, dtype=torch.int64, device=dev) * N
    ms = cuda_time(lambda: E.raster(x, y, p, off, B, 5, H, W))
    by = B * (5 * N + 15 * H * W)
    print(f"raster B=32 1M ev VGA: {ms:.3f} ms  {B*N/ms/1e3:.1f} Mev/s  {by/ms/1e6:.1f} GB/s algorithmic")
    fr, hw, _, st = E.raster(x, y, p, off, B, 5, H, W)
    ms = cuda_time(lambda: E.preprocess(fr, hw, H))
    print(f"preprocess 160 frames: {ms:.3f} ms")
    print("EVENTS_BAD", bad)


GEMM_TIMING = [(92320, 3072, 1024, 0), (92320, 1024, 1024, 3), (92320, 1024, 1024, 4), (92320, 1024, 1024, 0), (92320, 4096, 1024, 1),
               (92320, 1024, 4096, 3), (92320, 1024, 4096, 4), (92320, 1024, 4096, 0), (92320, 4096, 4096, 0)]


def stage_gemm():
    import torch
    from eventgpt_b200 import events as E
    from eventgpt_b200._lib import lib, check
    dev = torch.device("cuda:0")
    torch.manual_seed(0)

    def run(M, N, K, epi):
        A = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
        B = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
        bias = torch.randn(N, device=dev) * 0.1
        ref = A.float() @ B.float().t() + bias
        if epi == 1:
            ref = ref * torch.sigmoid(1.702 * ref)
        elif epi == 2:
            ref = torch.nn.functional.gelu(ref)
        if epi == 3:
            out = torch.randn(M, N, device=dev)
            ref = ref + out
        elif epi == 4:
            out = torch.empty(M, N, device=dev)
        else:
            out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        check(lib().egpt_gemm_bf16(E.ptr(A), E.ptr(B), E.ptr(bias), E.ptr(out), M, N, K, epi, E.stream_ptr(dev)), "gemm")
        torch.cuda.synchronize()
        err = (out.float() - ref).abs().max().item()
        rel = err / ref.abs().max().item()
        return err, rel

    for (M, N, K) in [(128, 256, 64), (128, 256, 128), (100, 256, 256), (1000, 512, 1024), (4616, 3072, 1024), (2885, 1024, 4096)]:
        for epi in (4, 0, 1, 2, 3):
            try:
                err, rel = run(M, N, K, epi)
                print(f"gemm M={M} N={N} K={K} epi={epi}: max-abs {err:.4g} rel {rel:.3g} {'OK' if rel < 2e-2 else 'BAD'}")
            except Exception as ex:  # noqa
                print(f"gemm M={M} N={N} K={K} epi={epi}: EXCEPTION {ex}")
                return
    for (M, N, K, epi) in GEMM_TIMING:
        A = torch.randn(M, K, device=dev).bfloat16()
        B = torch.randn(N, K, device=dev).bfloat16()
        bias = torch.randn(N, device=dev)
        out = torch.zeros(M, N, device=dev, dtype=torch.float32 if epi >= 3 else torch.bfloat16)
        ms = cuda_time(lambda: check(lib().egpt_gemm_bf16(E.ptr(A), E.ptr(B), E.ptr(bias), E.ptr(out), M, N, K, epi, E.stream_ptr(dev))))
        ms_t = cuda_time(lambda: torch.matmul(A, B.t()))
        print(f"gemm time M={M} N={N} K={K} epi={epi}: {ms:.3f} ms  {2*M*N*K/ms/1e9:.0f} TFLOP/s   (torch.matmul {ms_t:.3f} ms {2*M*N*K/ms_t/1e9:.0f} TFLOP/s)")


def stage_attn():
    import ctypes as C
    import torch
    from eventgpt_b200 import events as E
    from eventgpt_b200._lib import lib, check
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    fns = {"tcgen05": lib().egpt_attention}
    if hasattr(lib(), "egpt_attention_legacy"):
        leg = lib().egpt_attention_legacy
        leg.restype, leg.argtypes = C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        fns["legacy"] = leg
    for F in (1, 3, 37, 160):                    # 37 / 160 frames: several heads per persistent CTA
        qkv = (torch.randn(F * 577, 3072, device=dev)).bfloat16()
        q, k, v = (qkv.float().view(F, 577, 3, 16, 64)[:, :, i].transpose(1, 2) for i in range(3))
        ref = (torch.softmax(q @ k.transpose(-1, -2), -1) @ v).transpose(1, 2).reshape(F * 577, 1024)
        for name, fn in fns.items():
            out = torch.full((F * 577, 1024), float("nan"), device=dev, dtype=torch.bfloat16)
            check(fn(E.ptr(qkv), F, E.ptr(out), E.stream_ptr(dev)), "attn")
            torch.cuda.synchronize()
            d = (out.float() - ref).abs()
            err = d.max().item()
            print(f"attention[{name}] F={F}: max-abs {err:.4g} nan {int(torch.isnan(out.float()).sum())} {'OK' if err < 3e-2 else 'BAD'}")
            if not err < 3e-2:
                dd = d.view(F, 577, 16, 64)
                print("   err by head", [round(x, 3) for x in dd.amax(dim=(0, 1, 3)).tolist()])
                print("   err by row/64", [round(x, 3) for x in dd[0].amax(dim=(1, 2)).view(-1)[:576].view(9, 64).amax(1).tolist()], round(dd[0, 576].max().item(), 3))
                print("   err by dim/8", [round(x, 3) for x in dd.amax(dim=(0, 1, 2)).view(8, 8).amax(1).tolist()])
    F = 160
    qkv = (torch.randn(F * 577, 3072, device=dev) * 0.5).bfloat16()
    out = torch.empty(F * 577, 1024, device=dev, dtype=torch.bfloat16)
    for name, fn in fns.items():
        ms = cuda_time(lambda: check(fn(E.ptr(qkv), F, E.ptr(out), E.stream_ptr(dev))))
        print(f"attention[{name}] F=160: {ms:.3f} ms  {4*577*577*64*16*F/ms/1e9:.0f} TFLOP/s")


def stage_misc():
    import torch
    from eventgpt_b200.model import get_spatio_temporal_features, _pool
    from oracle import vit as ovit
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    for dt in (torch.float32, torch.bfloat16):
        f = torch.randn(5, 577, 4096, device=dev).to(dt)
        got = get_spatio_temporal_features(list(f)).float().cpu()
        ref = ovit.get_spatio_temporal_features(f.float().cpu())
        print(f"pool {dt}: max-abs {(got-ref).abs().max().item():.4g}")
        got = get_spatio_temporal_features(f, 8).float().cpu()
        ref = ovit.get_spatio_temporal_features(f.float().cpu(), 8)
        print(f"pool nt=8 {dt}: max-abs {(got-ref).abs().max().item():.4g}")
        got = get_spatio_temporal_features(f, 2).float().cpu()
        ref = ovit.get_spatio_temporal_features(f.float().cpu(), 2)
        print(f"pool nt=2 {dt}: max-abs {(got-ref).abs().max().item():.4g}")


def stage_tower():
    import numpy as np
    import torch
    from conftest import load_golden_events, GOLDEN
    from eventgpt_b200.model import EventGPTVisualFrontEnd
    from eventgpt_b200.weights import random_state_dict
    from oracle import preprocess as opp
    dev = torch.device("cuda:0")
    _, _, pix = load_golden_events("sample1")
    frames = [torch.from_numpy(opp.bf16_bits_to_f32(p)).to(dev).bfloat16() for p in pix]
    for layers in (0, 2, 24):
        sd = random_state_dict(0, max(layers, 2) if layers else 2)
        fe = EventGPTVisualFrontEnd(sd, dev, num_layers=layers)
        if layers:
            z = np.load(os.path.join(GOLDEN, f"tokens_sample1_L{layers}.npz"))
            lhs = fe.tower(frames[0][None]).float().cpu()
            d = (lhs[0, :, ::8] - torch.from_numpy(z["lhs0_sub"])).abs().max().item()
            print(f"L={layers} last_hidden_state frame0: max-abs {d:.4g} (ref absmax {np.abs(z['lhs0_sub']).max():.3g})")
            pr = fe.visval_encode(frames[0][None]).float().cpu()
            d = (pr[0, :, ::16] - torch.from_numpy(z["proj0_sub"])).abs().max().item()
            print(f"L={layers} visval_encode frame0: max-abs {d:.4g} (ref absmax {np.abs(z['proj0_sub']).max():.3g})")
            tok = fe.encode_events(frames).float().cpu()[0]
            ref = torch.from_numpy(z["tokens_sub"])
            d = (tok[:, ::16] - ref).abs().max().item()
            cos = torch.nn.functional.cosine_similarity(tok[:, ::16].flatten(), ref.flatten(), dim=0).item()
            print(f"L={layers} tokens: max-abs {d:.4g} cosine {cos:.6f} (ref absmax {float(z['absmax']):.3g}) {'OK' if d < 2e-2 and cos > 0.999 else 'BAD'}")
        else:
            from oracle import vit as ovit
            sd0 = sd
            lhs = fe.tower(frames[0][None]).float().cpu()
            ref = ovit.tower_last_hidden_state(sd0, frames[0][None].float().cpu(), 0)
            print(f"L=0 embeddings+pre_ln: max-abs {(lhs-ref).abs().max().item():.4g}")
        fe.close()


def stage_fused():
    import numpy as np
    import torch
    from eventgpt_b200 import events as E
    from eventgpt_b200.model import EventGPTVisualFrontEnd
    from eventgpt_b200.weights import random_state_dict
    from eventgpt_b200 import _lib
    dev = torch.device("cuda:0")
    sd = random_state_dict(0, 24)
    fe = EventGPTVisualFrontEnd(sd, dev, num_layers=24)
    rng = np.random.default_rng(2)
    for B in (1, 8, 32):
        N, H, W = 1_000_000, 480, 640
        x = torch.from_numpy(rng.integers(0, W, B * N).astype(np.uint16)).to(dev)
        y = torch.from_numpy(rng.integers(0, H, B * N).astype(np.uint16)).to(dev)
        p = torch.from_numpy(rng.integers(0, 2, B * N).astype(np.uint8)).to(dev)
        off = torch.arange(0, B + 1, dtype=torch.int64, device=dev) * N
        out = torch.empty((B, 582, 4096), dtype=torch.bfloat16, device=dev)
        _lib.launch_count(True)
        fe.visual_tokens(x, y, p, off, (H, W), out=out)
        nl = _lib.launch_count(True)
        ms = cuda_time(lambda: fe.visual_tokens(x, y, p, off, (H, W), out=out, check_flags=False), iters=3, warm=1)
        print(f"fused B={B}: {ms:.2f} ms/step  {B/ms*1e3:.1f} windows/s  {B*2127.4/ms:.0f} TFLOP/s  launches/step {nl}")
    fe.close()


STAGES = {"events": stage_events, "gemm": stage_gemm, "attn": stage_attn, "misc": stage_misc, "tower": stage_tower,
          "fused": stage_fused}

if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--run":
        STAGES[sys.argv[2]]()
        sys.exit(0)
    stages = sys.argv[1:] or list(STAGES)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    for s in stages:
        t0 = time.time()
        print(f"===== stage {s} =====", flush=True)
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--run", s], timeout=420,
                               capture_output=True, text=True)
            out = (r.stdout + "\n" + r.stderr[-3000:]) if r.returncode else r.stdout
            print(out, f"[exit {r.returncode}, {time.time()-t0:.0f}s]", flush=True)
        except subprocess.TimeoutExpired as ex:
            print(f"TIMEOUT after {time.time()-t0:.0f}s\n", (ex.stdout or b"")[-3000:], flush=True)
