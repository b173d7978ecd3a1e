"""Drop-in for the visual half of the reference's ``model/EventChatModel.py``:
``VisualTower`` (:45-61), ``visval_encode`` (:194-200), ``feature_adaptor`` (:84-85),
``get_spatio_temporal_features`` (:15-38) and the event branch of
``prepare_inputs_labels_for_multimodal`` (:326-345).  Same names, argument meaning and
error behaviour; the arithmetic runs in libegpt_b200.so.  No CPU fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from . import events as E
from ._lib import lib, check
from .weights import VT

TOKENS, HID, LLM = 577, 1024, 4096
IGNORE_INDEX = -100          # dataset/constants.py:7
EVENT_TOKEN_INDEX = -200     # dataset/constants.py:8
MAX_PROMPT_LEN = 2048        # model/EventChatModel.py:415


def get_spatio_temporal_features(features, num_temporal_tokens=None):
    """model/EventChatModel.py:15-38.  ``features``: list of t tensors [s, c] or a tensor [t, s, c]
    (bf16 or fp32, on the GPU) -> [num_temporal_tokens + s, c]: temporal rows (mean over s, zero
    padded / truncated to num_temporal_tokens) first, then the s spatial rows (mean over t)."""
    if isinstance(features, list):
        features = torch.stack(features)
    if features.ndim != 3:
        raise ValueError("Input features should be a 3D tensor with shape (t, s, c)")
    return _pool(features[None], num_temporal_tokens)[0]


def _pool(features, num_temporal_tokens=None):
    """[B, t, s, c] -> [B, nt + s, c] through egpt_spatio_temporal_pool."""
    if not features.is_cuda:
        raise _lib.EgptError("eventgpt_b200 has no CPU path: features must live on the GPU")
    B, t, s, c = features.shape
    nt = t if num_temporal_tokens is None else int(num_temporal_tokens)
    if features.dtype not in (torch.bfloat16, torch.float32):
        raise TypeError(f"features must be bf16 or fp32, got {features.dtype}")
    f = features.contiguous()
    out = torch.empty((B, nt + s, c), dtype=f.dtype, device=f.device)
    nws = max(lib().egpt_pool_workspace_bytes(B, t, c), B * ((s + 31) // 32) * t * c * 4)
    ws = E.Workspace.get(f.device, "pool", nws)
    check(lib().egpt_spatio_temporal_pool(E.ptr(f), 1 if f.dtype == torch.float32 else 0, B, t, s, c, nt, E.ptr(out),
                                          E.ptr(ws), ws.numel(), E.stream_ptr(f.device)), "egpt_spatio_temporal_pool")
    return out


def splice_event_features(embed_table, input_ids, position_ids, attention_mask, labels, event_features):
    """The multimodal splice of ``prepare_inputs_labels_for_multimodal`` (model/EventChatModel.py:355-465).

    ``embed_table``: the LLM's ``embed_tokens.weight`` [V, C] (bf16, GPU); ``event_features``: [NF, T, C]
    (bf16, GPU), consumed in order -- one entry per <event> token (-200) and one per sample without any,
    as upstream.  ``input_ids`` / ``attention_mask`` / ``labels`` / ``position_ids`` as the reference takes
    them (CPU or GPU).  Returns (position_ids, attention_mask, inputs_embeds, labels) = elements 1, 2, 4, 5 of
    the reference's tuple, with its ``None`` conventions.  The index bookkeeping mirrors the reference's
    Python on the host (it is O(tokens)); the embedding gather, feature copy, truncation and right padding
    are one kernel (egpt_splice_embeddings)."""
    if not (embed_table.is_cuda and event_features.is_cuda):
        raise _lib.EgptError("eventgpt_b200 has no CPU path: embed_table and event_features must live on the GPU")
    dev = embed_table.device
    ids_h = input_ids.detach().cpu()
    _labels, _position_ids, _attention_mask = labels, position_ids, attention_mask
    am_h = torch.ones_like(ids_h, dtype=torch.bool) if attention_mask is None else attention_mask.detach().cpu().bool()
    lab_h = torch.full_like(ids_h, IGNORE_INDEX) if labels is None else labels.detach().cpu()
    NF, T, C = event_features.shape
    srcs, labs, cur = [], [], 0
    for b in range(ids_h.shape[0]):
        cur_ids = ids_h[b][am_h[b]]
        cur_lab = lab_h[b][am_h[b]]
        ev_pos = torch.where(cur_ids == EVENT_TOKEN_INDEX)[0].tolist()
        if not ev_pos:                                   # :380-387 -- still consumes one feature entry
            if cur >= NF:
                raise IndexError(f"index {cur} is out of bounds for dimension 0 with size {NF}")
            srcs.append(cur_ids.to(torch.int32))
            labs.append(cur_lab)
            cur += 1
            continue
        bounds = [-1] + ev_pos + [cur_ids.shape[0]]
        s_parts, l_parts = [], []
        for i in range(len(bounds) - 1):
            s_parts.append(cur_ids[bounds[i] + 1: bounds[i + 1]].to(torch.int32))
            l_parts.append(cur_lab[bounds[i] + 1: bounds[i + 1]])
            if i < len(ev_pos):
                if cur >= NF:
                    raise IndexError(f"index {cur} is out of bounds for dimension 0 with size {NF}")
                s_parts.append(-(torch.arange(cur * T, (cur + 1) * T, dtype=torch.int32) + 2))
                l_parts.append(torch.full((T,), IGNORE_INDEX, dtype=cur_lab.dtype))
                cur += 1
        srcs.append(torch.cat(s_parts))
        labs.append(torch.cat(l_parts))
    srcs = [x[:MAX_PROMPT_LEN] for x in srcs]            # :416-418
    labs = [x[:MAX_PROMPT_LEN] for x in labs]
    B, max_len = len(srcs), max(x.shape[0] for x in srcs)
    src = torch.full((B, max_len), -1, dtype=torch.int32)
    new_labels = torch.full((B, max_len), IGNORE_INDEX, dtype=lab_h.dtype)
    new_am = torch.zeros((B, max_len), dtype=torch.bool)
    new_pos = torch.zeros((B, max_len), dtype=torch.long if position_ids is None else position_ids.dtype)
    for i, (s_, l_) in enumerate(zip(srcs, labs)):
        n = s_.shape[0]
        src[i, :n] = s_
        if n > 0:
            new_labels[i, :n] = l_
            new_am[i, :n] = True
            new_pos[i, :n] = torch.arange(0, n, dtype=new_pos.dtype)
    if int(src.max()) >= embed_table.shape[0]:
        raise IndexError("index out of range in self")        # what nn.Embedding raises upstream
    table = embed_table.to(torch.bfloat16).contiguous()
    feats = event_features.to(torch.bfloat16).contiguous()
    src_d = src.to(dev)
    out = torch.empty((B, max_len, C), dtype=torch.bfloat16, device=dev)
    check(lib().egpt_splice_embeddings(E.ptr(table), E.ptr(feats), E.ptr(src_d), B * max_len, C, E.ptr(out),
                                       E.stream_ptr(dev)), "egpt_splice_embeddings")
    odev = input_ids.device
    new_labels = None if _labels is None else new_labels.to(odev)
    new_am = None if _attention_mask is None else new_am.to(dtype=_attention_mask.dtype).to(odev)
    new_pos = None if _position_ids is None else new_pos.to(odev)
    return new_pos, new_am, out, new_labels


class EventProcessor:
    """Stands where ``CLIPImageProcessor`` does in the reference (``VisualTower.event_processor``,
    model/EventChatModel.py:55/59; called at common/common.py:125): uint8 HWC frame ->
    {'pixel_values': [1,3,336,336]}.  The values are the bf16 tensor the reference feeds the model
    (common/common.py:126), returned as float32 (an exact widening)."""
    size = {"shortest_edge": 336}
    crop_size = {"height": 336, "width": 336}
    image_mean = [0.48145466, 0.4578275, 0.40821073]
    image_std = [0.26862954, 0.26130258, 0.27577711]

    def __init__(self, device="cuda"):
        self.device = device

    def __call__(self, images, return_tensors="pt"):
        if return_tensors != "pt":
            raise ValueError(f"Unsupported tensor type: {return_tensors}")
        dev = E._require_cuda(self.device)
        imgs = images if isinstance(images, (list, tuple)) else [images]
        outs = []
        for img in imgs:
            a = np.ascontiguousarray(np.asarray(img))
            if a.ndim != 3 or a.shape[2] != 3 or a.dtype != np.uint8:
                raise ValueError("EventProcessor expects uint8 HWC RGB frames")
            h, w = a.shape[:2]
            stride = lib().egpt_raster_frame_stride(h, w)
            buf = torch.zeros((1, stride), dtype=torch.uint8, device=dev)
            buf[0, : h * w * 3] = torch.from_numpy(a.reshape(-1)).to(dev)
            hw = torch.tensor([[h, w]], dtype=torch.int32, device=dev)
            pix, _, status = E.preprocess(buf, hw, h)
            E.check_status(status, "EventProcessor")
            outs.append(pix[0].float())
        return {"pixel_values": torch.stack(outs)}

    preprocess = __call__


class VisualTower:
    """Role of model/EventChatModel.py:45-61: owns the processor; the tower weights live in the handle."""

    def __init__(self, name, device):
        self.visual_tower_name = name
        self.event_processor = EventProcessor(device)


class EventGPTVisualFrontEnd:
    """Everything between an event window and ``event_features [B, 582, 4096]``.

    ``state_dict`` uses the EventGPT checkpoint keys (see eventgpt_b200/weights.py)."""

    def __init__(self, state_dict, device="cuda", num_layers=24, visual_tower="openai/clip-vit-large-patch14-336"):
        self.device = E._require_cuda(device)
        self.num_layers = num_layers
        self.visual_tower = VisualTower(visual_tower, self.device)
        self._handle = C.c_void_p()
        with torch.cuda.device(self.device):
            self._create(state_dict)

    # ---- construction ---------------------------------------------------------------------------
    def _create(self, sd):
        dev = self.device
        keep = []

        def mat(k):
            t = sd[k].detach().to(dev, torch.bfloat16).contiguous()
            keep.append(t)
            return t.data_ptr()

        def vec(k):
            t = sd[k].detach().to(dev, torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        layers = (_lib.LayerWeights * max(1, self.num_layers))()
        for l in range(self.num_layers):
            p = f"{VT}encoder.layers.{l}."
            lw = layers[l]
            lw.ln1_w, lw.ln1_b = vec(p + "layer_norm1.weight"), vec(p + "layer_norm1.bias")
            lw.ln2_w, lw.ln2_b = vec(p + "layer_norm2.weight"), vec(p + "layer_norm2.bias")
            for n, a in (("q", "q"), ("k", "k"), ("v", "v"), ("out", "o")):
                setattr(lw, a + "_w", mat(p + f"self_attn.{n}_proj.weight"))
                setattr(lw, a + "_b", vec(p + f"self_attn.{n}_proj.bias"))
            lw.fc1_w, lw.fc1_b = mat(p + "mlp.fc1.weight"), vec(p + "mlp.fc1.bias")
            lw.fc2_w, lw.fc2_b = mat(p + "mlp.fc2.weight"), vec(p + "mlp.fc2.bias")
        w = _lib.Weights()
        w.num_layers = self.num_layers
        w.class_embedding = vec(VT + "embeddings.class_embedding")
        pw = sd[VT + "embeddings.patch_embedding.weight"].detach().reshape(HID, -1)
        t = pw.to(dev, torch.bfloat16).contiguous()
        keep.append(t)
        w.patch_weight = t.data_ptr()
        w.position_embedding = vec(VT + "embeddings.position_embedding.weight")
        w.pre_ln_w, w.pre_ln_b = vec(VT + "pre_layrnorm.weight"), vec(VT + "pre_layrnorm.bias")
        w.layers = layers
        w.proj0_w, w.proj0_b = mat("model.visual_projector.0.weight"), vec("model.visual_projector.0.bias")
        w.proj2_w, w.proj2_b = mat("model.visual_projector.2.weight"), vec("model.visual_projector.2.bias")
        w.adaptor_w, w.adaptor_b = mat("model.feature_adaptor.weight"), vec("model.feature_adaptor.bias")
        torch.cuda.synchronize(dev)
        check(lib().egpt_create(C.byref(w), dev.index or 0, C.byref(self._handle)), "egpt_create")
        del keep

    @property
    def handle_value(self):
        """The egpt_handle* as a Python int: what the ``torch.ops.eventgpt_b200.*`` operators take as ``handle``."""
        return int(self._handle.value)

    def close(self):
        if getattr(self, "_handle", None) and self._handle.value:
            lib().egpt_destroy(self._handle)
            self._handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- reference call surface -----------------------------------------------------------------
    def get_visual_tower(self):
        return self.visual_tower

    def get_model(self):
        return self

    def _ws(self, F, tag="tower"):
        return E.Workspace.get(self.device, tag, lib().egpt_workspace_bytes(F))

    def _bf16(self, x):
        if not isinstance(x, torch.Tensor) or not x.is_cuda:
            raise _lib.EgptError("eventgpt_b200 has no CPU path: event tensors must live on the GPU")
        return x.to(torch.bfloat16).contiguous()

    def tower(self, event_tensor):
        """CLIPVisionModel(x).last_hidden_state (model/EventChatModel.py:196): [B,3,336,336] -> [B,577,1024]."""
        x = self._bf16(event_tensor)
        F = x.shape[0]
        out = torch.empty((F, TOKENS, HID), dtype=torch.bfloat16, device=x.device)
        ws = self._ws(F)
        check(lib().egpt_tower(self._handle, E.ptr(x), F, E.ptr(out), E.ptr(ws), ws.numel(), E.stream_ptr(x.device)),
              "egpt_tower")
        return out

    def visval_encode(self, event_tensor):
        """model/EventChatModel.py:194-200: [B,3,336,336] -> visual_projector(tower(x)) [B,577,4096]."""
        x = self._bf16(event_tensor)
        if x.ndim != 4 or tuple(x.shape[1:]) != (3, 336, 336):
            raise ValueError(f"expected [B,3,336,336], got {tuple(x.shape)}")
        F = x.shape[0]
        out = torch.empty((F, TOKENS, LLM), dtype=torch.bfloat16, device=x.device)
        ws = self._ws(F)
        check(lib().egpt_visval_encode(self._handle, E.ptr(x), F, E.ptr(out), E.ptr(ws), ws.numel(),
                                       E.stream_ptr(x.device)), "egpt_visval_encode")
        return out

    def feature_adaptor(self, x):
        """model/EventChatModel.py:84-85: Linear(4096, 4096) on the last dimension."""
        xb = self._bf16(x)
        rows = xb.numel() // LLM
        out = torch.empty_like(xb)
        check(lib().egpt_feature_adaptor(self._handle, E.ptr(xb), rows, E.ptr(out), E.stream_ptr(xb.device)),
              "egpt_feature_adaptor")
        return out

    def encode_events(self, event_tensors):
        """Event branch of prepare_inputs_labels_for_multimodal (model/EventChatModel.py:326-345).

        list of frames / list of lists of frames ([3,336,336] each) -> event_features [B, n+577, 4096]
        (tower, projector, adaptor, pooling);  a plain tensor [B,3,336,336] -> visval_encode only
        (no adaptor, no pooling), exactly as the reference's tensor branch."""
        if not isinstance(event_tensors, list):
            return self.visval_encode(event_tensors)
        if not all(isinstance(item, list) for item in event_tensors):
            event_tensors = [event_tensors]
        n = len(event_tensors[0])
        if any(len(item) != n for item in event_tensors):
            # torch.stack at :343 would fail on ragged frame counts in the reference too
            raise RuntimeError("stack expects each sample to hold the same number of event frames")
        B = len(event_tensors)
        x = self._bf16(torch.stack([torch.stack([ev for ev in item]) for item in event_tensors]).reshape(B * n, 3, 336, 336))
        out = torch.empty((B, n + TOKENS, LLM), dtype=torch.bfloat16, device=x.device)
        ws = self._ws(B * n)
        check(lib().egpt_encode_pool(self._handle, E.ptr(x), B, n, E.ptr(out), E.ptr(ws), ws.numel(),
                                     E.stream_ptr(x.device)), "egpt_encode_pool")
        return out

    def project_pool(self, tower_out, n_frames=E.N_FRAMES):
        """Cached tower outputs [B*n, 577, 1024] (``tower``'s result) -> event_features [B, n+577, 4096]: projector, adaptor
        and pooling of the list branch (model/EventChatModel.py:199, 335-342) without re-running the tower."""
        x = self._bf16(tower_out)
        if x.ndim != 3 or tuple(x.shape[1:]) != (TOKENS, HID) or x.shape[0] % n_frames:
            raise ValueError(f"expected [B*{n_frames},577,1024], got {tuple(x.shape)}")
        B = x.shape[0] // n_frames
        out = torch.empty((B, n_frames + TOKENS, LLM), dtype=torch.bfloat16, device=x.device)
        ws = self._ws(x.shape[0])
        check(lib().egpt_project_pool(self._handle, E.ptr(x), B, n_frames, E.ptr(out), E.ptr(ws), ws.numel(),
                                      E.stream_ptr(x.device)), "egpt_project_pool")
        return out

    def prepare_inputs_labels_for_multimodal(self, embed_table, input_ids, position_ids, attention_mask, past_key_values,
                                             labels, event_tensors=None, event_image_sizes=None, event_features=None):
        """model/EventChatModel.py:309-465 with the LLM's embedding table passed in (this front end does not own
        the LLM).  Same 6-tuple as upstream.  ``event_image_sizes`` is accepted and never read, as upstream."""
        if event_tensors is None and event_features is None:
            return input_ids, position_ids, attention_mask, past_key_values, None, labels
        if input_ids.shape[1] == 1:
            return input_ids, position_ids, attention_mask, past_key_values, None, labels
        if event_features is None:
            event_features = self.encode_events(event_tensors)
        else:
            # cached branch re-applies the adaptor upstream (:347-353); kept, not "fixed"
            event_features = self.feature_adaptor(event_features)
        pos, am, emb, lab = splice_event_features(embed_table, input_ids, position_ids, attention_mask, labels, event_features)
        return None, pos, am, past_key_values, emb, lab

    def generate(self, embed_table, inputs=None, event_tensors=None, event_image_sizes=None, event_features=None, llm=None,
                 **kwargs):
        """Front half of ``EventChatModel.generate`` (model/EventChatModel.py:246-292): same argument handling and the same
        two ``NotImplementedError``s; everything up to the hand-over to the language model.  This front end does not own
        the LLM, so the embedding table is passed in and the call returns the keyword arguments upstream forwards to
        ``super().generate`` (``position_ids``, ``attention_mask``, ``inputs_embeds`` + the caller's own); with ``llm`` given
        it forwards them to ``llm.generate`` instead.  ``event_image_sizes`` is accepted and never read, as upstream."""
        position_ids = kwargs.pop("position_ids", None)
        attention_mask = kwargs.pop("attention_mask", None)
        if "inputs_embeds" in kwargs:
            raise NotImplementedError("`inputs_embeds` is not supported")
        if event_tensors is None and event_features is None:
            raise NotImplementedError("please input Event")
        _, position_ids, attention_mask, _, inputs_embeds, _ = self.prepare_inputs_labels_for_multimodal(
            embed_table, inputs, position_ids, attention_mask, None, None, event_tensors=event_tensors,
            event_image_sizes=event_image_sizes, event_features=event_features)
        out = dict(position_ids=position_ids, attention_mask=attention_mask, inputs_embeds=inputs_embeds, **kwargs)
        return llm.generate(**out) if llm is not None else out

    # ---- fused path: events -> tokens ------------------------------------------------------------
    def visual_tokens(self, x, y, p, window_offsets, sensor_size, n_frames=E.N_FRAMES, out=None, check_flags=True,
                      status=None, ranges=False):
        """process_event_data (common/common.py:110-129) + the list branch (:326-343) for a batch of
        windows, device resident end to end.  x, y, p: device integer tensors holding all windows back
        to back; window_offsets: int64 [B+1] on the device.  Returns tokens bf16 [B, n_frames+577, 4096].
        ``status`` (int32 [1] on the device, zeroed by the caller) receives the data-dependent flags when the caller
        checks them later itself (``check_flags=False``: no host synchronisation here).
        ``ranges=True``: ``window_offsets`` holds (start, end) pairs, int64 [B, 2], of possibly overlapping windows of one
        recording (``events.clip_bounds`` / ``events.ms_window_bounds``) -- egpt_visual_tokens_ranges."""
        dev = x.device
        H, W = int(sensor_size[0]), int(sensor_size[1])
        B = window_offsets.shape[0] if ranges else window_offsets.numel() - 1
        if out is None:
            out = torch.empty((B, n_frames + TOKENS, LLM), dtype=torch.bfloat16, device=dev)
        hw = torch.empty((B * n_frames, 2), dtype=torch.int32, device=dev)
        if status is None:
            status = torch.zeros(1, dtype=torch.int32, device=dev)
        nws = lib().egpt_visual_tokens_workspace_bytes(B, n_frames, H, W)
        ws = E.Workspace.get(dev, "fused", nws)
        fn = lib().egpt_visual_tokens_ranges if ranges else lib().egpt_visual_tokens
        check(fn(self._handle, E.ptr(x), E.ptr(y), E.ptr(p), E._CODE_OF_TORCH[x.dtype],
                 E._CODE_OF_TORCH[y.dtype], E._CODE_OF_TORCH[p.dtype], E.ptr(window_offsets), B,
                 n_frames, H, W, E.ptr(out), E.ptr(hw), E.ptr(status), E.ptr(ws), ws.numel(),
                 E.stream_ptr(dev)), "egpt_visual_tokens")
        if check_flags:
            E.check_status(status, "visual_tokens")
        return out
