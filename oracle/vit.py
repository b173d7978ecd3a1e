"""Oracle (test infrastructure): plain fp32 torch restatement of the reference's
tower -> projector -> adaptor -> spatio-temporal pooling stage.

Follows
  * ``model/EventChatModel.py:194-200``  visval_encode (CLIPVisionModel
    ``last_hidden_state`` -- no post_layernorm -- then ``visual_projector``)
  * ``model/EventChatModel.py:96-102``   Linear(1024,4096) . GELU(erf) . Linear(4096,4096)
  * ``model/EventChatModel.py:84-85,338`` feature_adaptor = Linear(4096,4096)
  * ``model/EventChatModel.py:15-38``    get_spatio_temporal_features
  * ``model/EventChatModel.py:326-343``  list branch of prepare_inputs_labels_for_multimodal
and, for the third-party tower (transformers 4.37.2 ``models/clip/modeling_clip.py``,
absent from /root/reference): conv 14x14/14 no bias -> 576 patches, class token,
position embedding, pre_layrnorm, 24 x {LN1, MHA 16x64 scale 0.125, +res, LN2,
fc1, QuickGELU x*sigmoid(1.702x), fc2, +res}, LN eps 1e-5.
Pinned against the real reference model via tests/golden (oracle/make_golden.py).
"""
import torch
import torch.nn.functional as F

HID = 1024
HEADS = 16
HEAD_DIM = 64
PATCH = 14
GRID = 24
TOKENS = GRID * GRID + 1     # 577
LN_EPS = 1e-5
VT = "model.visual_tower.visual_tower.vision_model."


def tower_last_hidden_state(sd, pixel_values, num_layers=24):
    """pixel_values fp32 [B,3,336,336] -> [B,577,1024] (CLIPVisionTransformer
    up to, and excluding, post_layernorm)."""
    x = pixel_values.float()
    b = x.shape[0]
    w = sd[VT + "embeddings.patch_embedding.weight"].float()
    pe = F.conv2d(x, w, stride=PATCH).flatten(2).transpose(1, 2)            # [B,576,1024]
    cls = sd[VT + "embeddings.class_embedding"].float().expand(b, 1, -1)
    h = torch.cat([cls, pe], dim=1) + sd[VT + "embeddings.position_embedding.weight"].float()[None]
    h = F.layer_norm(h, (HID,), sd[VT + "pre_layrnorm.weight"].float(),
                     sd[VT + "pre_layrnorm.bias"].float(), LN_EPS)
    for l in range(num_layers):
        p = f"{VT}encoder.layers.{l}."
        g = lambda n: sd[p + n].float()
        r = h
        y = F.layer_norm(h, (HID,), g("layer_norm1.weight"), g("layer_norm1.bias"), LN_EPS)
        q = F.linear(y, g("self_attn.q_proj.weight"), g("self_attn.q_proj.bias")) * (HEAD_DIM ** -0.5)
        k = F.linear(y, g("self_attn.k_proj.weight"), g("self_attn.k_proj.bias"))
        v = F.linear(y, g("self_attn.v_proj.weight"), g("self_attn.v_proj.bias"))
        sh = lambda t: t.view(b, TOKENS, HEADS, HEAD_DIM).transpose(1, 2)
        a = torch.softmax(sh(q) @ sh(k).transpose(-1, -2), dim=-1) @ sh(v)
        a = a.transpose(1, 2).reshape(b, TOKENS, HID)
        h = r + F.linear(a, g("self_attn.out_proj.weight"), g("self_attn.out_proj.bias"))
        r = h
        y = F.layer_norm(h, (HID,), g("layer_norm2.weight"), g("layer_norm2.bias"), LN_EPS)
        y = F.linear(y, g("mlp.fc1.weight"), g("mlp.fc1.bias"))
        y = y * torch.sigmoid(1.702 * y)
        h = r + F.linear(y, g("mlp.fc2.weight"), g("mlp.fc2.bias"))
    return h


def visual_projector(sd, x):
    """model/EventChatModel.py:96-102 (nn.GELU default = exact erf)."""
    y = F.linear(x, sd["model.visual_projector.0.weight"].float(), sd["model.visual_projector.0.bias"].float())
    y = F.gelu(y)
    return F.linear(y, sd["model.visual_projector.2.weight"].float(), sd["model.visual_projector.2.bias"].float())


def visval_encode(sd, pixel_values, num_layers=24):
    """model/EventChatModel.py:194-200."""
    return visual_projector(sd, tower_last_hidden_state(sd, pixel_values, num_layers))


def feature_adaptor(sd, x):
    """model/EventChatModel.py:84-85."""
    return F.linear(x, sd["model.feature_adaptor.weight"].float(), sd["model.feature_adaptor.bias"].float())


def get_spatio_temporal_features(features, num_temporal_tokens=None):
    """model/EventChatModel.py:15-38: temporal rows (mean over s) first, then
    spatial rows (mean over t)."""
    if isinstance(features, list):
        features = torch.stack(features)
    if features.ndim != 3:
        raise ValueError("Input features should be a 3D tensor with shape (t, s, c)")
    t = features.shape[0]
    if num_temporal_tokens is None:
        num_temporal_tokens = t
    temporal = features.mean(dim=1)
    if num_temporal_tokens > t:
        temporal = F.pad(temporal, (0, 0, 0, num_temporal_tokens - t))
    elif num_temporal_tokens < t:
        temporal = temporal[:num_temporal_tokens]
    return torch.cat([temporal, features.mean(dim=0)], dim=0)


def visual_tokens(sd, windows, num_layers=24):
    """List branch, model/EventChatModel.py:326-343: ``windows`` is a list (one
    entry per sample) of lists of [3,336,336] frames -> [B, n+577, 4096]."""
    out = []
    with torch.no_grad():
        for frames in windows:
            feats = []
            for ev in frames:
                f = visval_encode(sd, ev[None].float(), num_layers)
                feats.append(feature_adaptor(sd, f)[0])
            out.append(get_spatio_temporal_features(feats))
    return torch.stack(out)
