"""CPU-side checks of the boundary: the library loads, exports every symbol that
include/egpt_b200.h declares, the ctypes table matches the header, and the host-only
size queries behave.  No kernel is launched here."""
import os
import re

import pytest

from eventgpt_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "egpt_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(egpt_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    names = declared_functions()
    for must in ("egpt_raster", "egpt_split_time", "egpt_preprocess", "egpt_create", "egpt_destroy", "egpt_tower",
                 "egpt_visval_encode", "egpt_feature_adaptor", "egpt_spatio_temporal_pool", "egpt_encode_pool",
                 "egpt_visual_tokens", "egpt_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    l = _lib.lib()
    for name in declared_functions():
        assert hasattr(l, name), f"{name} declared in egpt_b200.h but not exported by libegpt_b200.so"


def test_ctypes_table_matches_header():
    assert sorted(_lib.SIGNATURES) == declared_functions()
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for name, (_, args) in _lib.SIGNATURES.items():
        m = re.search(r"\b" + name + r"\s*\(([^;]*?)\)\s*;", src, flags=re.S)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ("", "void") else params.count(",") + 1
        assert n == len(args), f"{name}: header has {n} parameters, ctypes table {len(args)}"


def test_size_queries_are_host_only_and_sane():
    l = _lib.lib()
    assert l.egpt_version() >= 100
    assert l.egpt_raster_frame_stride(480, 640) == 921600
    assert l.egpt_raster_frame_stride(3, 5) == 256
    # scratch = 2 bits per pixel per frame and event range: one range when the batch fills the machine,
    # up to 12 ranges per slice when it holds few frames
    assert l.egpt_raster_workspace_bytes(256, 5, 480, 640) <= (256 * 5 + 148) * (480 * 640 // 4 + 16) + 1024
    assert 5 * 720 * 1280 // 4 <= l.egpt_raster_workspace_bytes(1, 5, 720, 1280) <= 12 * 5 * (720 * 1280 // 4 + 16) + 1024
    assert l.egpt_raster_workspace_bytes(1, 5, 16, 20000) == 0          # a row band that wide does not fit one SM
    assert l.egpt_raster_workspace_bytes(0, 5, 720, 1280) == 0
    one = l.egpt_workspace_bytes(5)
    assert l.egpt_workspace_bytes(160) > 30 * one and l.egpt_workspace_bytes(160) < 34 * one
    assert l.egpt_visual_tokens_workspace_bytes(32, 5, 480, 640) > l.egpt_workspace_bytes(160)
    assert l.egpt_split_time_workspace_bytes(10 ** 6, 64) > 0
    assert l.egpt_preprocess_workspace_bytes(5, 480) >= 5 * 480 * 336 * 3


def test_bad_arguments_are_rejected_without_touching_the_gpu():
    l = _lib.lib()
    rc = l.egpt_raster(None, None, None, 2, 2, 0, None, 1, 5, 0, 480, 640, 0, None, None, None, None, None, 0, None)
    assert rc == -1
    assert b"null" in l.egpt_last_error()
    rc = l.egpt_gemm_bf16(None, None, None, None, 10, 256, 64, 0, None)
    assert rc == -1
    with pytest.raises(_lib.EgptError):
        _lib.check(rc, "gemm")


def test_dtype_codes():
    import numpy as np
    assert _lib.dtype_code(np.uint16) == 2 and _lib.dtype_code(np.int64) == 7 and _lib.dtype_code(np.uint8) == 0
    with pytest.raises(TypeError):
        _lib.dtype_code(np.float32)
