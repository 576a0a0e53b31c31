"""Pins oracle/preprocess.py against torchvision's own transforms (the reference's client-side preprocessing,
profiling/run_profiler.py:63-70) executed here, and against the committed fixture tests/golden/preprocess_golden.npz."""
import os

import numpy as np
import pytest
import torch

from oracle import preprocess as opre

MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
GOLD = os.path.join(os.path.dirname(__file__), "golden", "preprocess_golden.npz")


def torchvision_pipeline(img_u8_hwc):
    from torchvision import transforms
    from PIL import Image
    t = transforms.Compose([transforms.ToTensor(), transforms.Normalize(mean=list(MEAN), std=list(STD))])
    return t(Image.fromarray(img_u8_hwc))  # fp32 CHW


def test_lut_equals_torchvision_totensor_normalize_for_all_byte_values():
    # an image holding every byte value in every channel
    img = np.stack([np.arange(256, dtype=np.uint8).reshape(16, 16)] * 3, axis=-1)
    ref = torchvision_pipeline(img)                                           # [3,16,16] fp32
    ref_bits = ref.to(torch.bfloat16).view(torch.int16).numpy().astype(np.uint16)
    lut = opre.normalize_lut(MEAN, STD)
    for c in range(3):
        assert np.array_equal(lut[c], ref_bits[c].reshape(-1)), "channel %d" % c


def test_collate_identity_matches_torch_stack_of_transformed_images():
    rng = np.random.default_rng(0)
    imgs = [rng.integers(0, 256, (224, 224, 3), dtype=np.uint8) for _ in range(3)]
    want = torch.stack([torchvision_pipeline(im) for im in imgs])             # scheduler.py:443 torch.stack
    got = opre.collate_u8(imgs, 5, 224, 224, MEAN, STD)
    got_f = torch.from_numpy(opre.bf16_bits_to_f32(got[:3, :, :, :3]).copy()).permute(0, 3, 1, 2)
    assert torch.equal(got_f, want.to(torch.bfloat16).float())
    assert not got[3:].any() and not got[..., 3].any()


def test_golden_fixture():
    g = np.load(GOLD)
    got = opre.collate_u8(list(g["images"]), int(g["padded_batch"]), 224, 224, MEAN, STD)
    assert np.array_equal(got, g["expected_bits"])


def test_collate_f32_roundtrip_and_padding():
    t = [np.random.default_rng(i).standard_normal((3, 8, 6)).astype(np.float32) for i in range(2)]
    out = opre.collate_f32chw(t, 3)
    assert out.shape == (3, 8, 6, 4) and not out[2].any() and not out[..., 3].any()
    back = opre.bf16_bits_to_f32(out[0, :, :, :3]).transpose(2, 0, 1)
    assert np.array_equal(back, torch.from_numpy(t[0]).to(torch.bfloat16).float().numpy())


@pytest.mark.parametrize("hw", [(480, 640), (100, 333), (1, 1), (640, 640), (75, 3328)])
def test_letterbox_geometry(hw):
    rng = np.random.default_rng(1)
    img = rng.integers(0, 256, (*hw, 3), dtype=np.uint8)
    out = opre.preprocess_u8(img, 640, 640, opre.MODE_LETTERBOX, 114)
    rh, rw, top, left = opre._placement(hw[0], hw[1], 640, 640, opre.MODE_LETTERBOX)
    assert max(rh, rw) == 640 and abs(rh / rw - hw[0] / hw[1]) <= max(1 / rw, 1 / rh) * max(rh / rw, 1) + 1e-9
    assert (out[:top] == 114).all() and (out[:, :left] == 114).all()
    if hw == (640, 640):
        assert np.array_equal(out, img)


def test_resize_matches_float_bilinear_within_one_lsb():
    rng = np.random.default_rng(2)
    img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    got = opre.resize_u8(img, 64, 80).astype(np.int32)
    ref = torch.nn.functional.interpolate(torch.from_numpy(img).permute(2, 0, 1)[None].float(), size=(64, 80),
                                          mode="bilinear", align_corners=False, antialias=False)[0]
    ref = ref.permute(1, 2, 0).numpy()
    assert np.abs(got - ref).max() <= 1.0


def test_identity_rejects_wrong_size_and_bad_counts():
    with pytest.raises(ValueError):
        opre.preprocess_u8(np.zeros((10, 10, 3), np.uint8), 224, 224, opre.MODE_IDENTITY, 0)
    with pytest.raises(ValueError):
        opre.collate_u8([np.zeros((4, 4, 3), np.uint8)] * 3, 2, 4, 4, MEAN, STD)


# ---- the reference's antialiased Resize (profiling/run_profiler.py:63-70): oracle restatement pinned to torchvision ----
def _resize_golden():
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    import make_resize_golden as mk
    return mk, np.load(os.path.join(os.path.dirname(__file__), "golden", "resize_aa_golden.npz"))


def test_antialias_resize_oracle_within_one_lsb_of_torchvision_golden():
    mk, g = _resize_golden()
    from oracle import preprocess as op
    worst, differing, total = 0, 0, 0
    for i, (h, w) in enumerate(mk.SIZES):
        img = mk.make_image(i, h, w).permute(1, 2, 0).numpy()
        cases = [("stretch32_%d" % i, op.resize_aa_u8, 32), ("short32_%d" % i, op.resize_short_side_center_crop_u8, 32)]
        if i in mk.FULL:
            cases += [("stretch112_%d" % i, op.resize_aa_u8, 112), ("short112_%d" % i, op.resize_short_side_center_crop_u8, 112)]
        for key, fn, s in cases:
            got, want = fn(img, s, s), g[key]
            assert got.shape == want.shape == (s, s, 3), key          # sizes and crop offsets exact
            d = np.abs(got.astype(np.int16) - want.astype(np.int16))
            worst, differing, total = max(worst, int(d.max())), differing + int((d > 0).sum()), total + d.size
    print("antialias resize: worst |oracle - torchvision| = %d LSB, %d of %d values differ" % (worst, differing, total))
    assert worst <= 1                                                  # <= 1 uint8 LSB before normalisation
    gray = mk.make_image(100, 368, 304, gray=True).permute(1, 2, 0).numpy()
    got = op.resize_short_side_center_crop_u8(op.to_rgb(gray), 32, 32)
    assert np.abs(got.astype(np.int16) - g["gray_short32"].astype(np.int16)).max() <= 1


def test_short_side_size_rule_matches_torchvision():
    import torch
    import torchvision.transforms.v2.functional as TF
    from oracle import preprocess as op
    for h, w in [(375, 500), (500, 375), (333, 500), (48, 80), (2592, 3328), (203, 940), (940, 203), (224, 224), (225, 224)]:
        want = tuple(TF.resize(torch.zeros(3, h, w, dtype=torch.uint8), [224], antialias=True).shape[1:])
        assert op.short_side_size(h, w, 224) == want
    img = np.arange(6 * 8 * 3, dtype=np.uint8).reshape(6, 8, 3)
    assert np.array_equal(op.resize_aa_u8(img, 6, 8), img)             # same size: the antialias weights collapse to identity
