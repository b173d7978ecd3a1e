"""Host-side behaviour that needs no GPU: the product path must fail loudly (never fall
back to a CPU implementation) and must mirror the reference's argument checks."""
import numpy as np
import pytest
import torch

from eventgpt_b200 import _lib, common, model
from eventgpt_b200.weights import random_state_dict, tower_keys, head_keys

needs_no_gpu = pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")


@needs_no_gpu
def test_no_cpu_fallback_anywhere():
    ev = {k: np.arange(10, dtype=np.uint16) for k in "xytp"}
    with pytest.raises(_lib.EgptError):
        common.get_event_images_list(ev, 5)
    with pytest.raises(_lib.EgptError):
        common.generate_event_image(ev["x"], ev["y"], ev["p"])
    with pytest.raises(_lib.EgptError):
        common.split_event_by_time(ev)
    with pytest.raises(_lib.EgptError):
        model.get_spatio_temporal_features(torch.zeros(5, 7, 8))
    with pytest.raises(_lib.EgptError):
        model.EventGPTVisualFrontEnd(random_state_dict(0, 1), "cuda", num_layers=1)


def test_product_does_not_import_the_oracle():
    import os, re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dirpath, _, files in os.walk(os.path.join(root, "eventgpt_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f


def test_pooling_rejects_wrong_rank_like_the_reference():
    with pytest.raises(ValueError):       # model/EventChatModel.py:19-20
        model.get_spatio_temporal_features(torch.zeros(5, 7))


def test_state_dict_keys_are_the_checkpoints():
    sd = random_state_dict(3, 2)
    assert set(sd) == set(tower_keys(2)) | set(head_keys())
    assert sd["model.visual_projector.0.weight"].shape == (4096, 1024)
    assert sd["model.feature_adaptor.weight"].shape == (4096, 4096)
    k = "model.visual_tower.visual_tower.vision_model.encoder.layers.1.mlp.fc1.weight"
    assert sd[k].shape == (4096, 1024)
    # bf16 representable, like the released bf16 checkpoint
    assert torch.equal(sd[k], sd[k].bfloat16().float())
    assert torch.equal(random_state_dict(3, 2)[k], sd[k])
