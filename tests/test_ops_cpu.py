"""torch custom operators (eventgpt_b200/ops.py): registration and shape propagation, no GPU needed.
The operators themselves have no CPU implementation (they are calls into the CUDA library)."""
import pytest
import torch


def test_ops_are_registered_with_fake_impls():
    import eventgpt_b200.ops as ops
    for name in ops.OPS:
        assert hasattr(torch.ops.eventgpt_b200, name), name
    from torch._subclasses.fake_tensor import FakeTensorMode
    with FakeTensorMode():
        x = torch.empty(1000, dtype=torch.uint16, device="cuda")
        p = torch.empty(1000, dtype=torch.uint8, device="cuda")
        t = torch.empty(1000, dtype=torch.int64, device="cuda")
        off = torch.empty(3, dtype=torch.int64, device="cuda")
        tok, hw, st = torch.ops.eventgpt_b200.visual_tokens(0, x, x, p, off, 5, 480, 640)
        assert tuple(tok.shape) == (2, 582, 4096) and tok.dtype == torch.bfloat16 and tuple(hw.shape) == (10, 2)
        fr, hw, st = torch.ops.eventgpt_b200.raster(x, x, p, off, 5, 480, 640, 0, False)
        assert tuple(fr.shape) == (10, 921600) and fr.dtype == torch.uint8
        pix, st = torch.ops.eventgpt_b200.preprocess(fr, hw, 480)
        assert tuple(pix.shape) == (10, 3, 336, 336)
        assert tuple(torch.ops.eventgpt_b200.visval_encode(0, pix).shape) == (10, 577, 4096)
        assert tuple(torch.ops.eventgpt_b200.encode_pool(0, pix, 5).shape) == (2, 582, 4096)
        hid = torch.empty((10, 577, 1024), dtype=torch.bfloat16, device="cuda")
        assert tuple(torch.ops.eventgpt_b200.project_pool(0, hid, 5).shape) == (2, 582, 4096)
        feats = torch.empty((2, 5, 577, 4096), dtype=torch.bfloat16, device="cuda")
        assert tuple(torch.ops.eventgpt_b200.spatio_temporal_pool(feats, 5).shape) == (2, 582, 4096)
        assert tuple(torch.ops.eventgpt_b200.feature_adaptor(0, feats).shape) == (2, 5, 577, 4096)
        outs = torch.ops.eventgpt_b200.split_time(x, x, t, p, 50000, 64)
        assert tuple(outs[4].shape) == (65,) and outs[2].dtype == torch.int64
        table = torch.empty((100, 4096), dtype=torch.bfloat16, device="cuda")
        src = torch.empty(40, dtype=torch.int32, device="cuda")
        assert tuple(torch.ops.eventgpt_b200.splice(table, tok, src).shape) == (40, 4096)


def test_ops_refuse_cpu_tensors():
    import eventgpt_b200.ops  # noqa: F401
    x = torch.zeros(10, dtype=torch.uint16)
    with pytest.raises((RuntimeError, NotImplementedError)):
        torch.ops.eventgpt_b200.raster(x, x, x.to(torch.uint8), torch.tensor([0, 10]), 5, 4, 4, 0, False)
