"""ORACLE (test infrastructure only — never imported by the product path).

numpy restatement of the preprocessing + batch collate the reference leaves to clients and to `torch.stack`:
  * ToTensor:   uint8 HWC -> fp32 CHW / 255           (torchvision transforms/functional.py `to_tensor`: `.div(255)`)
  * Normalize:  (x - mean) / std in fp32              (torchvision transforms/_functional_tensor.py `normalize`: sub_().div_())
    as composed in profiling/run_profiler.py:63-70 and 293-project/src/milind-code/scheduler.py:24-27
  * stack:      torch.stack(batch.inputs)             (293-project/src/scheduler.py:443)
followed by the engine's storage format: NHWC, C padded 3->4 with zeros, bf16 (round-to-nearest-even), batch padded with
zero images.  Pinned by tests/test_oracle_preprocess.py against torchvision's own transforms run in this container
(values) and against tests/golden/preprocess_golden.npz (committed fixture generated from them).

Resize / letterbox index arithmetic (integer, 10-bit fixed-point bilinear, half-pixel centres — the cv2.INTER_LINEAR
convention used by YOLOv8's LetterBox) is defined here and mirrored by the CUDA kernel; the reference has no resize code
on the serving path (its driver submits torch.randn(3,224,224), 293-project/src/test_scheduler.py:89), so that part is
parity-unpinned against the reference and pinned only between this oracle and the kernel.
"""
import numpy as np

MODE_IDENTITY, MODE_PAD, MODE_RESIZE, MODE_LETTERBOX = 0, 1, 2, 3      # 4, 5: the antialiased resizes, defined below


def f32_to_bf16_bits(x):
    """Round-to-nearest-even fp32 -> bf16 bit pattern (uint16)."""
    b = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    return ((b + 0x7FFF + ((b >> 16) & 1)) >> 16).astype(np.uint16)


def bf16_bits_to_f32(bits):
    return (np.asarray(bits, dtype=np.uint32) << 16).view(np.float32)


def normalize_lut(mean, std):
    """[3][256] bf16 bit patterns of ((u/255) - mean[c]) / std[c], each step in IEEE fp32 like torch's CPU kernels."""
    u = np.arange(256, dtype=np.float32)
    lut = np.zeros((3, 256), dtype=np.uint16)
    for c in range(3):
        x = u / np.float32(255.0)
        y = x - np.float32(mean[c])
        z = y / np.float32(std[c])
        lut[c] = f32_to_bf16_bits(z)
    return lut


def _placement(h, w, out_h, out_w, mode):
    if mode == MODE_PAD:
        return min(h, out_h), min(w, out_w), 0, 0
    if mode == MODE_LETTERBOX:
        if out_h * w <= out_w * h:
            rh, rw = out_h, max(1, (w * out_h + h // 2) // h)
        else:
            rw, rh = out_w, max(1, (h * out_w + w // 2) // w)
        rw, rh = min(rw, out_w), min(rh, out_h)
        return rh, rw, (out_h - rh) // 2, (out_w - rw) // 2
    return out_h, out_w, 0, 0


def resize_u8(img, rh, rw):
    """Bilinear resize of a uint8 HWC image, integer arithmetic identical to collate_u8_general_kernel."""
    h, w, _ = img.shape
    if (rh, rw) == (h, w):
        return img.copy()
    ys = np.arange(rh, dtype=np.int64)
    xs = np.arange(rw, dtype=np.int64)
    sy = (2 * ys + 1) * h - rh
    sx = (2 * xs + 1) * w - rw
    fy = np.where(sy < 0, 0, (sy << 10) // (2 * rh))
    fx = np.where(sx < 0, 0, (sx << 10) // (2 * rw))
    y0, wy = fy >> 10, fy & 1023
    x0, wx = fx >> 10, fx & 1023
    wy = np.where(y0 >= h - 1, 0, wy)
    y0 = np.minimum(y0, h - 1)
    wx = np.where(x0 >= w - 1, 0, wx)
    x0 = np.minimum(x0, w - 1)
    y1, x1 = np.minimum(y0 + 1, h - 1), np.minimum(x0 + 1, w - 1)
    im = img.astype(np.int64)
    wx_ = wx[None, :, None]
    wy_ = wy[:, None, None]
    top = im[y0][:, x0] * (1024 - wx_) + im[y0][:, x1] * wx_
    bot = im[y1][:, x0] * (1024 - wx_) + im[y1][:, x1] * wx_
    return ((top * (1024 - wy_) + bot * wy_ + (1 << 19)) >> 20).astype(np.uint8)


def preprocess_u8(img, out_h, out_w, mode, fill):
    """uint8 HWC (or single-channel) of any size -> uint8 [out_h, out_w, 3] canvas according to `mode`."""
    img = to_rgb(img)
    h, w, _ = img.shape
    if mode == MODE_IDENTITY:
        if (h, w) != (out_h, out_w):
            raise ValueError("identity mode needs images of the model input size")
        return img
    rh, rw, top, left = _placement(h, w, out_h, out_w, mode)
    canvas = np.full((out_h, out_w, 3), fill, dtype=np.uint8)
    if mode == MODE_PAD:
        canvas[:rh, :rw] = img[:rh, :rw]
    elif mode == MODE_RESIZE_AA:
        return resize_aa_u8(img, out_h, out_w)
    elif mode == MODE_RESIZE_SHORT_AA:
        return resize_short_side_center_crop_u8(img, out_h, out_w)
    else:
        canvas[top:top + rh, left:left + rw] = resize_u8(img, rh, rw)
    return canvas


def collate_u8(images, padded_batch, out_h, out_w, mean, std, mode=MODE_IDENTITY, fill=0):
    """List of uint8 HWC arrays -> bf16 bit patterns [padded_batch, out_h, out_w, 4]; slots >= len(images) are zero."""
    n = len(images)
    if not (0 <= n <= padded_batch):
        raise ValueError("need 0 <= n <= padded_batch")
    lut = normalize_lut(mean, std)
    out = np.zeros((padded_batch, out_h, out_w, 4), dtype=np.uint16)
    for i, img in enumerate(images):
        c = preprocess_u8(np.asarray(img), out_h, out_w, mode, fill)
        for ch in range(3):
            out[i, :, :, ch] = lut[ch][c[:, :, ch]]
    return out


def collate_f32chw(tensors, padded_batch):
    """List of fp32 CHW arrays (reference-native requests) -> bf16 bit patterns [padded_batch, H, W, 4]."""
    n = len(tensors)
    _, h, w = np.asarray(tensors[0]).shape
    out = np.zeros((padded_batch, h, w, 4), dtype=np.uint16)
    for i, t in enumerate(tensors):
        out[i, :, :, :3] = f32_to_bf16_bits(np.asarray(t, dtype=np.float32)).transpose(1, 2, 0)
    return out


# ---------------------------------------------------------------------------------------------------------------------
# The reference's own resize: transforms.Resize(224) / Resize((224, 224)) (profiling/run_profiler.py:63-70,
# 293-project/src/milind-code/scheduler.py:24-27) = bilinear WITH antialias.  Oracle of record:
# torchvision.transforms.v2.functional.resize(uint8 CHW tensor, antialias=True) on the CPU; the arithmetic lives in
# torch (ATen/native/cpu/UpSampleKernel.cpp, _compute_indices_min_size_weights_aa + the separable aa kernel), restated
# here in float64 with ONE rounding at the end.  torch's uint8 path rounds to uint8 between its horizontal and vertical
# pass and uses fixed-point weights, so the two may differ by 1 in the last bit; tests/test_oracle_preprocess.py pins this
# restatement to torchvision within that 1 LSB on the committed golden vectors (tests/golden/resize_aa_golden.npz).
# ---------------------------------------------------------------------------------------------------------------------
MODE_RESIZE_AA, MODE_RESIZE_SHORT_AA = 4, 5


def aa_weights(in_size, out_size):
    """Per output index: (first input index, normalised triangle weights), torch's antialias bilinear."""
    scale = in_size / out_size
    support = scale if scale >= 1.0 else 1.0
    invscale = 1.0 / scale if scale >= 1.0 else 1.0
    out = []
    for i in range(out_size):
        center = scale * (i + 0.5)
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size)
        j = np.arange(xmin, xmax, dtype=np.float64)
        w = np.clip(1.0 - np.abs((j - center + 0.5) * invscale), 0.0, None)
        out.append((xmin, w / w.sum()))
    return out


def resize_aa_u8(img, oh, ow):
    """Antialiased bilinear resize of a uint8 HWC image to (oh, ow): separable triangle filter, float64, one rounding."""
    h, w, c = img.shape
    src = img.astype(np.float64)
    tmp = np.empty((h, ow, c), dtype=np.float64)
    for x, (x0, wx) in enumerate(aa_weights(w, ow)):
        tmp[:, x, :] = np.tensordot(src[:, x0:x0 + len(wx), :], wx, axes=([1], [0]))
    out = np.empty((oh, ow, c), dtype=np.float64)
    for y, (y0, wy) in enumerate(aa_weights(h, oh)):
        out[y] = np.tensordot(wy, tmp[y0:y0 + len(wy)], axes=([0], [0]))
    return np.clip(np.floor(out + 0.5), 0, 255).astype(np.uint8)


def short_side_size(h, w, size):
    """torchvision _compute_resized_output_size for an int size: short side -> size, long side -> int(size * long / short)."""
    short, long_ = (w, h) if w <= h else (h, w)
    new_short, new_long = size, int(size * long_ / short)
    return (new_long, new_short) if w <= h else (new_short, new_long)      # (new_h, new_w)


def resize_short_side_center_crop_u8(img, oh, ow):
    """Resize(min(oh, ow)) (short side, antialias) followed by CenterCrop((oh, ow)) -- the crop is what lets requests of
    different aspect ratios share one batch tensor; torchvision's CenterCrop offsets: int(round((H - oh) / 2.0))."""
    h, w, _ = img.shape
    nh, nw = short_side_size(h, w, min(oh, ow))
    r = resize_aa_u8(img, nh, nw)
    top, left = int(round((nh - oh) / 2.0)), int(round((nw - ow) / 2.0))
    return r[top:top + oh, left:left + ow]


def to_rgb(img):
    """PIL's L -> RGB for a single-channel request (31 of the 2000 workload images are grayscale): the value is replicated."""
    if img.ndim == 2:
        img = img[:, :, None]
    return np.repeat(img, 3, axis=2) if img.shape[2] == 1 else img
