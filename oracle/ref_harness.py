"""Oracle (test infrastructure): loads the REAL reference (``/root/reference``,
read-only, present only in the build container) so that golden vectors can be
generated from its own code.  Never imported on the GPU box and never by the
product.  Recipe from SURVEY.md section 8(c).
"""
import os
import sys
import tempfile
import types

REF = "/root/reference"


def available():
    return os.path.isdir(os.path.join(REF, "common"))


def import_reference():
    """Returns (common.common, model.EventChatModel) modules of the reference."""
    if REF not in sys.path:
        sys.path.insert(0, REF)
    # peft is imported at model/EventChatModel.py:8 and never used.
    stub = types.ModuleType("peft")
    stub.get_peft_model = stub.LoraConfig = stub.TaskType = None
    had = "peft" in sys.modules
    if not had:
        sys.modules["peft"] = stub
    import model.EventChatModel as M
    if not had:
        del sys.modules["peft"]
    import common.common as C
    return C, M


def pil_processor():
    """PIL-backed CLIP processor == transformers 4.37.2 / Pillow behaviour."""
    from transformers.models.clip.image_processing_pil_clip import CLIPImageProcessorPil
    return CLIPImageProcessorPil(size={"shortest_edge": 336}, crop_size={"height": 336, "width": 336})


def build_reference_model(state_dict, num_layers=24, name="clip-vit-l14-336-rand"):
    """Constructs the reference's EventChatModel (fp32, CPU) around a random-init
    CLIP ViT-L/14-336 saved under ./checkpoints/<name>/ -- VisualTower looks there
    first (model/EventChatModel.py:52-56) -- then loads ``state_dict`` into the
    tower / projector / adaptor."""
    import torch
    from transformers import CLIPVisionConfig, CLIPVisionModel, CLIPImageProcessor
    C, M = import_reference()
    work = tempfile.mkdtemp(prefix="egpt_ref_")
    ckpt = os.path.join(work, "checkpoints", name)
    os.makedirs(ckpt)
    cfg = CLIPVisionConfig(hidden_size=1024, intermediate_size=4096, num_hidden_layers=num_layers,
                           num_attention_heads=16, image_size=336, patch_size=14,
                           hidden_act="quick_gelu", projection_dim=768)
    CLIPVisionModel(cfg).save_pretrained(ckpt)
    CLIPImageProcessor(size={"shortest_edge": 336}, crop_size={"height": 336, "width": 336}).save_pretrained(ckpt)
    cwd = os.getcwd()
    os.chdir(work)
    try:
        mcfg = M.EventChatConfig(hidden_size=4096, num_hidden_layers=1, intermediate_size=128,
                                 num_attention_heads=32, vocab_size=32000,
                                 mm_visual_tower=name, event_feature_adaptor=True)
        model = M.EventChatModel(mcfg).float().eval()
    finally:
        os.chdir(cwd)
    own = model.state_dict()
    missing = [k for k in state_dict if k not in own]
    assert not missing, missing[:5]
    with torch.no_grad():
        for k, v in state_dict.items():
            assert own[k].shape == v.shape, (k, own[k].shape, v.shape)
            own[k].copy_(v)
    model.get_model().visual_tower.event_processor = pil_processor()
    return C, M, model
