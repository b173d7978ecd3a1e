"""Weight containers for the visual-token stage.

The state-dict keys are the EventGPT checkpoint's own
(``model/EventChatModel.py:81-85`` in the reference):
``model.visual_tower.visual_tower.vision_model.*`` (CLIP ViT-L/14-336),
``model.visual_projector.{0,2}.*`` and ``model.feature_adaptor.*``.
There is no network on the build or GPU boxes, so benchmarks and tests use
``random_state_dict`` (seeded, bf16-representable values like the released
bf16 checkpoint); a real checkpoint's state dict can be passed instead.
"""
import torch

VT = "model.visual_tower.visual_tower.vision_model."
HID, MLP, LLM = 1024, 4096, 4096
PATCH, GRID = 14, 24
TOKENS = GRID * GRID + 1


def tower_keys(num_layers=24):
    keys = {
        VT + "embeddings.class_embedding": (HID,),
        VT + "embeddings.patch_embedding.weight": (HID, 3, PATCH, PATCH),
        VT + "embeddings.position_embedding.weight": (TOKENS, HID),
        VT + "pre_layrnorm.weight": (HID,),
        VT + "pre_layrnorm.bias": (HID,),
    }
    for l in range(num_layers):
        p = f"{VT}encoder.layers.{l}."
        for n in ("q", "k", "v", "out"):
            keys[p + f"self_attn.{n}_proj.weight"] = (HID, HID)
            keys[p + f"self_attn.{n}_proj.bias"] = (HID,)
        for n in ("layer_norm1", "layer_norm2"):
            keys[p + n + ".weight"] = (HID,)
            keys[p + n + ".bias"] = (HID,)
        keys[p + "mlp.fc1.weight"] = (MLP, HID)
        keys[p + "mlp.fc1.bias"] = (MLP,)
        keys[p + "mlp.fc2.weight"] = (HID, MLP)
        keys[p + "mlp.fc2.bias"] = (HID,)
    return keys


def head_keys():
    return {
        "model.visual_projector.0.weight": (LLM, HID),
        "model.visual_projector.0.bias": (LLM,),
        "model.visual_projector.2.weight": (LLM, LLM),
        "model.visual_projector.2.bias": (LLM,),
        "model.feature_adaptor.weight": (LLM, LLM),
        "model.feature_adaptor.bias": (LLM,),
    }


def random_state_dict(seed=0, num_layers=24):
    """Seeded random-init weights with every value bf16-representable, returned
    as fp32 CPU tensors.  Biases and LayerNorm affine terms are non-trivial on
    purpose so that a dropped bias or a swapped gamma/beta fails parity."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd = {}
    shapes = dict(tower_keys(num_layers))
    shapes.update(head_keys())
    for k, shp in shapes.items():
        r = torch.randn(shp, generator=g, dtype=torch.float32)
        if "layer_norm" in k or "layrnorm" in k:
            v = 1.0 + 0.05 * r if k.endswith("weight") else 0.02 * r
        elif k.endswith("bias"):
            v = 0.02 * r
        elif "q_proj" in k or "k_proj" in k:
            v = 0.03 * r
        elif "visual_projector" in k or "feature_adaptor" in k:
            v = r * (shp[1] ** -0.5) * 0.6
        else:
            v = 0.02 * r
        sd[k] = v.to(torch.bfloat16).to(torch.float32)
    return sd


OUTLIER_CHANNELS = (7, 300, 777, 1001)
PEAKED = {3: (0, 5), 10: (2, 9), 17: (4, 15)}          # layer -> heads with sharpened attention


def outlier_state_dict(seed=1, num_layers=24):
    """``random_state_dict`` reshaped towards what a trained CLIP-L tower looks like numerically: a few residual
    channels carry "massive activations" (every layer's fc2 bias pushes ``OUTLIER_CHANNELS`` further out, so that the
    last hidden state holds them at ~50x the median channel), and a few heads have q/k projections scaled so that
    their softmax rows are sharply peaked.  Used for a second 24-layer parity golden; every value stays
    bf16-representable."""
    sd = random_state_dict(seed, num_layers)
    pos = sd[VT + "embeddings.position_embedding.weight"]
    for i, ch in enumerate(OUTLIER_CHANNELS):
        pos[:, ch] += (8.0 if i % 2 == 0 else -6.0)
        for l in range(num_layers):
            sd[f"{VT}encoder.layers.{l}.mlp.fc2.bias"][ch] += (10.0 if i % 2 == 0 else -8.0)
    for l, heads in PEAKED.items():
        if l >= num_layers:
            continue
        for n in ("q", "k"):
            w = sd[f"{VT}encoder.layers.{l}.self_attn.{n}_proj.weight"]
            for h in heads:
                w[h * 64:(h + 1) * 64] *= 6.0
    return {k: v.to(torch.bfloat16).to(torch.float32) for k, v in sd.items()}
