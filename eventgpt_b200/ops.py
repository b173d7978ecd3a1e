"""torch custom operators over the C ABI: ``torch.ops.eventgpt_b200.*``.

Each operator is one call into ``libegpt_b200.so`` on ``torch.cuda.current_stream()`` (what
``at::cuda::getCurrentCUDAStream()`` returns in C++); ctypes is only the transport.  Every operator has a
fake (meta) implementation for shape propagation, so the ops trace under FakeTensor / ``torch.export`` and
pass ``torch.library.opcheck``.  They are what the reference's call sites hit when it is patched
(INTEGRATION.md section 1):

  raster          common/common.py:17-37, 64-74      get_event_images_list / generate_event_image
  preprocess      common/common.py:124-127           event_processor(...)['pixel_values'].to(bf16)
  split_time      common/common.py:76-107            split_event_by_time
  visval_encode   model/EventChatModel.py:194-200
  feature_adaptor model/EventChatModel.py:84-85, 338
  spatio_temporal_pool  model/EventChatModel.py:15-38
  encode_pool     model/EventChatModel.py:326-343    list branch of prepare_inputs_labels_for_multimodal
  project_pool    the same from cached tower outputs
  visual_tokens   common/common.py:110-129 + model/EventChatModel.py:326-343, events -> tokens
  splice          model/EventChatModel.py:355-465    gather into inputs_embeds

A model handle (``egpt_create``) crosses the operator boundary as a Python int (the pointer value); the
front end that owns it is ``eventgpt_b200.model.EventGPTVisualFrontEnd`` (``.handle_value``).
There is no CPU implementation: a CPU tensor raises.
"""
import ctypes as C
from typing import Tuple

import torch
from torch import Tensor
from torch.library import custom_op

from . import events as E
from ._lib import check, lib

TOKENS, HID, LLM, CROP = 577, 1024, 4096, 336


def _cuda(*ts):
    for t in ts:
        if not t.is_cuda:
            raise RuntimeError("eventgpt_b200 operators run on CUDA tensors only (no CPU fallback)")


def _ws(dev, tag, nbytes):
    return E.Workspace.get(dev, tag, nbytes)


@custom_op("eventgpt_b200::raster", mutates_args=(), device_types="cuda")
def raster(x: Tensor, y: Tensor, p: Tensor, window_offsets: Tensor, n_frames: int, H: int, W: int, mode: int,
           fixed_canvas: bool) -> Tuple[Tensor, Tensor, Tensor]:
    """-> (frames uint8 [F, stride], hw int32 [F, 2], status int32 [1])"""
    _cuda(x, y, p, window_offsets)
    frames, hw, _, status = E.raster(x.contiguous(), y.contiguous(), p.contiguous(), window_offsets.contiguous(),
                                     window_offsets.numel() - 1, n_frames, H, W, mode=mode, fixed_canvas=fixed_canvas)
    return frames, hw, status


@raster.register_fake
def _(x, y, p, window_offsets, n_frames, H, W, mode, fixed_canvas):
    F = (window_offsets.numel() - 1) * n_frames
    stride = (3 * H * W + 255) // 256 * 256
    return (x.new_empty((F, stride), dtype=torch.uint8), x.new_empty((F, 2), dtype=torch.int32),
            x.new_empty((1,), dtype=torch.int32))


@custom_op("eventgpt_b200::preprocess", mutates_args=(), device_types="cuda")
def preprocess(frames: Tensor, hw: Tensor, max_h: int) -> Tuple[Tensor, Tensor]:
    """-> (pixel_values bf16 [F, 3, 336, 336], status int32 [1])"""
    _cuda(frames, hw)
    pix, _, status = E.preprocess(frames.contiguous(), hw.contiguous(), max_h)
    return pix, status


@preprocess.register_fake
def _(frames, hw, max_h):
    return (frames.new_empty((frames.shape[0], 3, CROP, CROP), dtype=torch.bfloat16), frames.new_empty((1,), dtype=torch.int32))


@custom_op("eventgpt_b200::split_time", mutates_args=(), device_types="cuda")
def split_time(x: Tensor, y: Tensor, t: Tensor, p: Tensor, interval: int, max_windows: int) \
        -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor, Tensor, Tensor]:
    """-> (x', y', t', p', offsets int64 [max_windows + 1], n_windows int32 [1], status int32 [1])"""
    _cuda(x, y, t, p)
    return E.split_time(x.contiguous(), y.contiguous(), t.contiguous(), p.contiguous(), interval, max_windows)


@split_time.register_fake
def _(x, y, t, p, interval, max_windows):
    return (torch.empty_like(x), torch.empty_like(y), torch.empty_like(t), torch.empty_like(p),
            x.new_empty((max_windows + 1,), dtype=torch.int64), x.new_empty((1,), dtype=torch.int32),
            x.new_empty((1,), dtype=torch.int32))


def _tower_ws(dev, F):
    return _ws(dev, "tower", lib().egpt_workspace_bytes(F))


@custom_op("eventgpt_b200::visval_encode", mutates_args=(), device_types="cuda")
def visval_encode(handle: int, pixel_values: Tensor) -> Tensor:
    """bf16 [F, 3, 336, 336] -> bf16 [F, 577, 4096]"""
    _cuda(pixel_values)
    x = pixel_values.to(torch.bfloat16).contiguous()
    F = x.shape[0]
    out = torch.empty((F, TOKENS, LLM), dtype=torch.bfloat16, device=x.device)
    ws = _tower_ws(x.device, F)
    check(lib().egpt_visval_encode(C.c_void_p(handle), E.ptr(x), F, E.ptr(out), E.ptr(ws), ws.numel(),
                                   E.stream_ptr(x.device)), "egpt_visval_encode")
    return out


@visval_encode.register_fake
def _(handle, pixel_values):
    return pixel_values.new_empty((pixel_values.shape[0], TOKENS, LLM), dtype=torch.bfloat16)


@custom_op("eventgpt_b200::feature_adaptor", mutates_args=(), device_types="cuda")
def feature_adaptor(handle: int, x: Tensor) -> Tensor:
    _cuda(x)
    xb = x.to(torch.bfloat16).contiguous()
    out = torch.empty_like(xb)
    check(lib().egpt_feature_adaptor(C.c_void_p(handle), E.ptr(xb), xb.numel() // LLM, E.ptr(out), E.stream_ptr(xb.device)),
          "egpt_feature_adaptor")
    return out


@feature_adaptor.register_fake
def _(handle, x):
    return x.new_empty(x.shape, dtype=torch.bfloat16)


@custom_op("eventgpt_b200::spatio_temporal_pool", mutates_args=(), device_types="cuda")
def spatio_temporal_pool(features: Tensor, n_temporal: int) -> Tensor:
    """[B, t, s, c] (bf16 or fp32) -> [B, n_temporal + s, c]"""
    _cuda(features)
    f = features.contiguous()
    if f.dtype not in (torch.bfloat16, torch.float32):
        f = f.float()
    B, t, s, c = f.shape
    out = torch.empty((B, n_temporal + s, c), dtype=f.dtype, device=f.device)
    ws = _ws(f.device, "pool", lib().egpt_pool_workspace_bytes(B, t, c))
    check(lib().egpt_spatio_temporal_pool(E.ptr(f), 1 if f.dtype == torch.float32 else 0, B, t, s, c, n_temporal, E.ptr(out),
                                          E.ptr(ws), ws.numel(), E.stream_ptr(f.device)), "egpt_spatio_temporal_pool")
    return out


@spatio_temporal_pool.register_fake
def _(features, n_temporal):
    B, t, s, c = features.shape
    dt = features.dtype if features.dtype in (torch.bfloat16, torch.float32) else torch.float32
    return features.new_empty((B, n_temporal + s, c), dtype=dt)


@custom_op("eventgpt_b200::encode_pool", mutates_args=(), device_types="cuda")
def encode_pool(handle: int, pixel_values: Tensor, n_frames: int) -> Tensor:
    """bf16 [B*n, 3, 336, 336] -> tokens bf16 [B, n + 577, 4096]"""
    _cuda(pixel_values)
    x = pixel_values.to(torch.bfloat16).contiguous()
    F = x.shape[0]
    B = F // n_frames
    out = torch.empty((B, n_frames + TOKENS, LLM), dtype=torch.bfloat16, device=x.device)
    ws = _tower_ws(x.device, F)
    check(lib().egpt_encode_pool(C.c_void_p(handle), E.ptr(x), B, n_frames, E.ptr(out), E.ptr(ws), ws.numel(),
                                 E.stream_ptr(x.device)), "egpt_encode_pool")
    return out


@encode_pool.register_fake
def _(handle, pixel_values, n_frames):
    return pixel_values.new_empty((pixel_values.shape[0] // n_frames, n_frames + TOKENS, LLM), dtype=torch.bfloat16)


@custom_op("eventgpt_b200::project_pool", mutates_args=(), device_types="cuda")
def project_pool(handle: int, tower_out: Tensor, n_frames: int) -> Tensor:
    """cached tower outputs bf16 [B*n, 577, 1024] -> tokens bf16 [B, n + 577, 4096]"""
    _cuda(tower_out)
    x = tower_out.to(torch.bfloat16).contiguous()
    F = x.shape[0]
    B = F // n_frames
    out = torch.empty((B, n_frames + TOKENS, LLM), dtype=torch.bfloat16, device=x.device)
    ws = _tower_ws(x.device, F)
    check(lib().egpt_project_pool(C.c_void_p(handle), E.ptr(x), B, n_frames, E.ptr(out), E.ptr(ws), ws.numel(),
                                  E.stream_ptr(x.device)), "egpt_project_pool")
    return out


@project_pool.register_fake
def _(handle, tower_out, n_frames):
    return tower_out.new_empty((tower_out.shape[0] // n_frames, n_frames + TOKENS, LLM), dtype=torch.bfloat16)


@custom_op("eventgpt_b200::visual_tokens", mutates_args=(), device_types="cuda")
def visual_tokens(handle: int, x: Tensor, y: Tensor, p: Tensor, window_offsets: Tensor, n_frames: int, H: int, W: int) \
        -> Tuple[Tensor, Tensor, Tensor]:
    """events -> (tokens bf16 [B, n + 577, 4096], hw int32 [B*n, 2], status int32 [1]); no host round trip"""
    _cuda(x, y, p, window_offsets)
    x, y, p, off = x.contiguous(), y.contiguous(), p.contiguous(), window_offsets.contiguous()
    dev = x.device
    B = off.numel() - 1
    out = torch.empty((B, n_frames + TOKENS, LLM), dtype=torch.bfloat16, device=dev)
    hw = torch.empty((B * n_frames, 2), dtype=torch.int32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    ws = _ws(dev, "fused", lib().egpt_visual_tokens_workspace_bytes(B, n_frames, H, W))
    check(lib().egpt_visual_tokens(C.c_void_p(handle), E.ptr(x), E.ptr(y), E.ptr(p), E._CODE_OF_TORCH[x.dtype],
                                   E._CODE_OF_TORCH[y.dtype], E._CODE_OF_TORCH[p.dtype], E.ptr(off), B, n_frames, H, W,
                                   E.ptr(out), E.ptr(hw), E.ptr(status), E.ptr(ws), ws.numel(), E.stream_ptr(dev)),
          "egpt_visual_tokens")
    return out, hw, status


@visual_tokens.register_fake
def _(handle, x, y, p, window_offsets, n_frames, H, W):
    B = window_offsets.numel() - 1
    return (x.new_empty((B, n_frames + TOKENS, LLM), dtype=torch.bfloat16), x.new_empty((B * n_frames, 2), dtype=torch.int32),
            x.new_empty((1,), dtype=torch.int32))


@custom_op("eventgpt_b200::splice", mutates_args=(), device_types="cuda")
def splice(embed_table: Tensor, event_features: Tensor, src: Tensor) -> Tensor:
    """inputs_embeds rows: embed_table[src] (src >= 0), event_features row -(src+2) (src <= -2), zero (src == -1)"""
    _cuda(embed_table, event_features, src)
    tb, fb, sb = embed_table.to(torch.bfloat16).contiguous(), event_features.to(torch.bfloat16).contiguous(), \
        src.to(torch.int32).contiguous()
    c = tb.shape[-1]
    out = torch.empty((sb.numel(), c), dtype=torch.bfloat16, device=tb.device)
    check(lib().egpt_splice_embeddings(E.ptr(tb), E.ptr(fb.reshape(-1, c)), E.ptr(sb), sb.numel(), c, E.ptr(out),
                                       E.stream_ptr(tb.device)), "egpt_splice_embeddings")
    return out


@splice.register_fake
def _(embed_table, event_features, src):
    return embed_table.new_empty((src.numel(), embed_table.shape[-1]), dtype=torch.bfloat16)


OPS = ("raster", "preprocess", "split_time", "visval_encode", "feature_adaptor", "spatio_temporal_pool", "encode_pool",
       "project_pool", "visual_tokens", "splice")
