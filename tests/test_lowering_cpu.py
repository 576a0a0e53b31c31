"""Host-side layout rules of the lowering (no GPU): physical channel counts, ShuffleNetV2's padded halves, and the
multiply-high division the kernels use for index math (csrc/ptx.cuh `fast_div`, csrc/common.h `fast_div_mul`), restated
in Python integers and checked for exactness over the whole range the kernels rely on."""
import random

import numpy as np
import pytest

import rdb_b200  # noqa: F401
from rdb_b200 import lowering, synth


def test_physical_channel_rules():
    # multiples of 8 by default
    for c in (3, 8, 16, 40, 48, 64, 96, 160, 256, 960, 1536):
        assert lowering.phys_channels(c) == lowering.pad8(c)
    # 48-byte pixels -> 64-byte pixels; the 80-channel class branch of YOLOv8n -> 128
    assert lowering.phys_channels(24) == 32
    assert lowering.phys_channels(80) == 128
    # 128-byte lines when that costs at most 12 %: ShuffleNetV2's 58 / 116 / 232-channel halves
    assert lowering.line_channels(58) == 64
    assert lowering.line_channels(116) == 128
    assert lowering.line_channels(232) == 256
    assert lowering.line_channels(464) == 512
    # too expensive: stays a multiple of 8
    assert lowering.line_channels(48) == 48
    assert lowering.line_channels(160) == 160
    assert lowering.line_channels(80) == 80
    for c in range(1, 2049):
        p = lowering.line_channels(c)
        assert p >= c and p % 8 == 0 and p * 100 <= lowering.pad8(c) * 112


def test_shufflenet_halves_are_line_aligned(monkeypatch):
    """Every shuffle writes two halves of `line_channels(c)` channels; the chunk(2) views of the next unit start on a
    128-byte line, and every logical channel maps inside its half."""
    monkeypatch.setenv("B200_FUSE_SHUFFLE", "0")
    pb = lowering.lower("shufflenet_v2_x1_0", synth.make_model("shufflenet_v2_x1_0", n_proto=8))
    _, _, ops, _ = lowering.parse_blob(pb.to_blob())
    shuffles = [op for op in ops if op["type"] == lowering.OP_SHUFFLE]
    assert len(shuffles) == 16
    for op in shuffles:
        p = op["p"]
        c, cp, out_c = p[14], p[15], p[5]
        assert cp == lowering.line_channels(c) and out_c == 2 * cp
        assert (cp * 2) % 128 == 0          # second half starts on a 128-byte line


def test_shufflenet_shuffle_folded_into_the_last_conv_of_branch2():
    """Default lowering: no shuffle pass is left; each unit ends in a 1x1 conv that writes cat + channel_shuffle itself
    (tv models/shufflenetv2.py:24-38,96-103). Its weight rows sit in `cp` slots: channel i < c/2 in slot i, the others from
    slot cp/2 on, the remaining rows zero; the plan interpreter gives the same logits as the plan with shuffle passes."""
    import os
    import numpy as np
    import torch
    from oracle import plan_interp
    m = synth.make_model("shufflenet_v2_x1_0", n_proto=8)
    pb = lowering.lower("shufflenet_v2_x1_0", m)
    blob = pb.to_blob()
    _, bufs, ops, weights = lowering.parse_blob(blob)
    assert not [op for op in ops if op["type"] == lowering.OP_SHUFFLE]
    tails = [op for op in ops if op["type"] == lowering.OP_CONV and op["p"][18] > 0]
    assert len(tails) == 16
    wb = np.frombuffer(weights, dtype=np.uint8)
    for op in tails:
        p = op["p"]
        c, cp, cin = p[18], p[5] // 2, p[2]
        assert cp == lowering.line_channels(c) and bufs[p[3]][2] == 2 * cp and p[6] >= 0 and p[11] == lowering.ACT_RELU
        rows = wb[op["w_off"]:op["w_off"] + cp * cin * 2].view(np.uint16).reshape(cp, cin)
        used = np.zeros(cp, bool)
        used[:c // 2] = True
        used[cp // 2:cp // 2 + c // 2] = True
        assert not rows[~used].any() and rows[used].any(axis=1).all()
    x = synth.normalize_u8(synth.heldout_images(2, 224, 224), synth.IMAGENET_MEAN, synth.IMAGENET_STD)
    fused, _ = plan_interp.PlanInterpreter(blob, lowering.parse_blob).run(plan_interp.to_nhwc4(x))
    os.environ["B200_FUSE_SHUFFLE"] = "0"
    try:
        plain = lowering.lower("shufflenet_v2_x1_0", m).to_blob()
    finally:
        del os.environ["B200_FUSE_SHUFFLE"]
    ref, _ = plan_interp.PlanInterpreter(plain, lowering.parse_blob).run(plan_interp.to_nhwc4(x))
    assert torch.equal(torch.as_tensor(fused), torch.as_tensor(ref))


def _fast_div(n, d):
    """csrc/ptx.cuh fast_div with csrc/common.h fast_div_mul, in Python integers (32-bit wrap-around included)."""
    if d == 1:
        return n
    mul = ((1 << 32) // d + 1) & 0xFFFFFFFF
    q = (n * mul) >> 32
    if ((q * d) & 0xFFFFFFFF) > n:
        q -= 1
    return q


def test_fast_div_is_exact_below_2_31():
    rng = random.Random(7)
    edge = [0, 1, 2, 3, (1 << 31) - 1, (1 << 31) - 2, (1 << 30), (1 << 30) + 1]
    for d in list(range(1, 1026)) + [1536, 2048, 3136, 4096, 12544, 50176, 65535, 65536, 65537, (1 << 31) - 1]:
        ns = edge + [rng.randrange(1 << 31) for _ in range(200)]
        ns += [k * d + e for k in (1, 2, 1000, ((1 << 31) - 1) // d) for e in (-1, 0, 1)]
        for n in ns:
            if 0 <= n < (1 << 31):
                assert _fast_div(n, d) == n // d, (n, d)


@pytest.mark.parametrize("d", [1, 2, 3, 7, 28, 49, 148, 784])
def test_fast_div_exhaustive_small_range(d):
    n = np.arange(0, 1 << 20, dtype=np.uint64)
    mul = np.uint64(((1 << 32) // d + 1) & 0xFFFFFFFF)
    q = (n * mul) >> np.uint64(32)
    q = q - ((q * np.uint64(d)) > n).astype(np.uint64)
    want = n // np.uint64(d)
    if d == 1:
        q = n
    assert np.array_equal(q, want)


@pytest.mark.parametrize("kind,hw", [("resnet50", 224), ("efficientnet_v2_s", 224), ("yolov8n", 640), ("shufflenet_v2_x1_0", 224)])
def test_padded_weight_extents_do_not_change_the_plan_results(monkeypatch, kind, hw):
    """Conv weights are stored padded to the TMA box sizes (lowering.weight_k / weight_rows, recorded in p[19] / p[20]);
    the plan interpreter must produce bit-identical outputs from the padded and the exact layouts, every padded element
    must be zero, and the extents must cover the layer."""
    import numpy as np
    import torch
    from oracle import plan_interp
    m = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
    kw = {"in_hw": (hw, hw)} if kind in ("efficientnet_v2_s", "yolov8n") else {}
    padded = lowering.lower(kind, m, **kw).to_blob()
    monkeypatch.setenv("B200_PAD_WEIGHTS", "0")
    exact = lowering.lower(kind, m, **kw).to_blob()
    monkeypatch.delenv("B200_PAD_WEIGHTS")
    _, _, ops_p, w_p = lowering.parse_blob(padded)
    _, _, ops_e, _ = lowering.parse_blob(exact)
    assert len(ops_p) == len(ops_e)
    wb = np.frombuffer(w_p, dtype=np.uint8)
    n_padded = 0
    for op, oe in zip(ops_p, ops_e):
        if op["type"] != lowering.OP_CONV or op["p"][18] > 0:
            continue
        p = op["p"]
        assert oe["p"][19] == 0 and oe["p"][20] == 0
        k, cin, cout = p[8], p[2] + (p[15] if p[13] > 0 else 0), p[5]
        kpad, rows = p[19], p[20]
        assert kpad >= cin and kpad % 8 == 0 and rows >= cout and rows % 64 == 0
        w = wb[op["w_off"]:op["w_off"] + rows * k * k * kpad * 2].view(np.uint16).reshape(rows, k * k, kpad)
        assert not w[cout:].any() and not w[:, :, cin:].any()
        n_padded += int(kpad > cin or rows > cout)
    assert n_padded > 0 or kind == "resnet50"       # ResNet-50's channel counts already are multiples of the boxes
    g = torch.Generator().manual_seed(5)
    x = torch.randn(1, hw, hw, 3, generator=g)
    a, _ = plan_interp.PlanInterpreter(padded, lowering.parse_blob).run(plan_interp.to_nhwc4(x.permute(0, 3, 1, 2)))
    b, _ = plan_interp.PlanInterpreter(exact, lowering.parse_blob).run(plan_interp.to_nhwc4(x.permute(0, 3, 1, 2)))
    assert torch.equal(torch.as_tensor(a), torch.as_tensor(b))
