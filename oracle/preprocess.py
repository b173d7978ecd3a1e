"""Oracle (test infrastructure): restatement of the CLIP preprocessing that the
reference delegates to ``CLIPImageProcessor`` (``common/common.py:124-127``;
processor built at ``model/EventChatModel.py:55/59``).

Third-party arithmetic restated here (absent from /root/reference):
  * transformers 4.37.2 ``models/clip/image_processing_clip.py``: resize shortest
    edge -> 336 (``get_resize_output_image_size``, long side = int(336*long/short)),
    centre crop 336x336 (offset = (size-336)//2), rescale 1/255, normalise.
  * Pillow 10.4.0 ``src/libImaging/Resample.c``: antialiased bicubic (a=-0.5),
    two passes (horizontal first), 8-bit fixed point with PRECISION_BITS = 22,
    uint8 rounding between the passes.
Pinned against the real ``CLIPImageProcessorPil`` via tests/golden (see
``oracle/make_golden.py``).
"""
import numpy as np

CROP = 336
PRECISION_BITS = 32 - 8 - 2
CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)


def resize_geometry(h, w, size=CROP):
    """(new_h, new_w, top, left) of resize-shortest-edge + centre crop."""
    short, long_ = (w, h) if w <= h else (h, w)
    new_short, new_long = size, int(size * long_ / short)
    new_h, new_w = (new_long, new_short) if w <= h else (new_short, new_long)
    return new_h, new_w, (new_h - size) // 2, (new_w - size) // 2


def _bicubic(x):
    a = -0.5
    if x < 0.0:
        x = -x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def precompute_coeffs(in_size, out_size):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc.
    Returns (xmin[out], count[out], kk[out, ksize] int32)."""
    scale = in_size / out_size
    filterscale = max(scale, 1.0)
    support = 2.0 * filterscale
    ksize = int(np.ceil(support)) * 2 + 1
    xmins = np.zeros(out_size, np.int32)
    counts = np.zeros(out_size, np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = [_bicubic((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for v in w:
            ww += v
        for x in range(xmax):
            v = w[x] / ww if ww != 0.0 else w[x]
            if v < 0:
                kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS))
            else:
                kk[xx, x] = int(0.5 + v * (1 << PRECISION_BITS))
        xmins[xx] = xmin
        counts[xx] = xmax
    return xmins, counts, kk


def _resample_axis0(img, out_size):
    """One 8-bit Pillow pass along axis 0 of an int array [in, ...]."""
    in_size = img.shape[0]
    xmins, counts, kk = precompute_coeffs(in_size, out_size)
    src = img.astype(np.int64)
    out = np.empty((out_size,) + img.shape[1:], np.uint8)
    for xx in range(out_size):
        lo, c = int(xmins[xx]), int(counts[xx])
        k = kk[xx, :c].astype(np.int64)
        acc = (1 << (PRECISION_BITS - 1)) + np.tensordot(k, src[lo:lo + c], axes=(0, 0))
        out[xx] = np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return out


def resize_crop_u8(frame, size=CROP):
    """uint8[H,W,3] -> uint8[336,336,3]: PIL bicubic resize (horizontal pass,
    then vertical pass) followed by the centre crop."""
    h, w, _ = frame.shape
    new_h, new_w, top, left = resize_geometry(h, w, size)
    img = frame
    if new_w != w:      # Pillow skips a pass whose size is unchanged
        img = np.ascontiguousarray(np.swapaxes(_resample_axis0(np.swapaxes(img, 0, 1), new_w), 0, 1))
    if new_h != h:
        img = _resample_axis0(img, new_h)
    return np.ascontiguousarray(img[top:top + size, left:left + size])


def normalise_lut_f32():
    """float32[256,3]: value of channel c for uint8 level u after rescale (1/255,
    computed in float64 then cast to float32) and (v-mean)/std in float32."""
    u = np.arange(256, dtype=np.float64)
    v = (u * (1 / 255)).astype(np.float32)
    mean = np.array(CLIP_MEAN, dtype=np.float32)
    std = np.array(CLIP_STD, dtype=np.float32)
    return ((v[:, None] - mean[None, :]) / std[None, :]).astype(np.float32)


def f32_to_bf16_bits(a):
    """Round-to-nearest-even float32 -> bf16 bit pattern (uint16), as torch's
    ``.to(torch.bfloat16)`` does (common/common.py:126)."""
    b = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).astype(np.uint64)
    rounded = b + 0x7FFF + ((b >> 16) & 1)
    return (rounded >> 16).astype(np.uint16)


def bf16_bits_to_f32(b):
    return (np.asarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)


def pixel_values_bf16_bits(frame):
    """uint8[H,W,3] -> uint16[3,336,336] (bf16 bit patterns), the tensor
    ``process_event_data`` hands to the model (common/common.py:125-126)."""
    u8 = resize_crop_u8(frame)
    lut = f32_to_bf16_bits(normalise_lut_f32())      # [256,3]
    chw = np.stack([lut[u8[:, :, c], c] for c in range(3)], axis=0)
    return chw
